#!/usr/bin/env python
"""bench.py -- imagined RSSM states / second (fwd+bwd, H=15) on B200.

One "step" = one pass of the hot path named by BASELINE.json over one batch of synthetic start
states: H-step ``rollout_policy`` (actor MLP -> action sample -> RSSM
transition) + reward/value heads + lambda-returns + actor and value losses, forward, then
``actor_loss.backward()`` and ``value_loss.backward()`` (gradients into the actor and value
parameters), plus -- for N > 1 GPUs -- the NCCL all-reduce of those gradients.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload walker|...]
    python bench.py --impl reference ...   # the UNMODIFIED reference (oracle/_ref) on the host cores

Workload: N = 1 runs BASELINE.json configs[1] shapes (walker: 2 500 start states x H=15).  N > 1 runs
configs[3] -- 65 536 start states SHARDED over the ranks ("scaling": "strong") -- as the primary line,
with the weak-scaling walker line (2 500 per GPU) and the same 65 536 states on ONE GPU (rank 0, same run)
as secondary blocks.  ``--workload`` / ``--scaling`` override.

Prints ONE JSON line (rank 0).  ``value`` is timed with CUDA events, inputs resident in HBM;
``e2e`` goes through the public Python API with pinned HOST buffers, host<->device copies and the
loss read-back inside the timed region.  ``roofline`` is the dominant kernel (timed with every kernel in
plain stream order) against the measured bf16 tensor peak; ``cpu_baseline`` is the reference timed on the
box's host cores; ``reference_cuda`` is the reference's eager CUDA path on the same GPU (the north
star's ">= 10x" denominator); ``t_full`` / ``full_update`` widen the step to the observation rollout and
to the whole ``Dreamer.loss`` with the conv stack.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "imagined_rssm_states_per_sec_fwd_bwd_H15"
UNIT = "states/s"

# Workloads.  "walker" is BASELINE.json configs[1] (DMC walker-walk: A=6 tanh-normal actor, B=50 T=50 ->
# 2500 start states, H=15); "config4" is configs[3] (65 536 start states, sharded over the GPUs); "walker_large" is one
# GPU's shard of it at 8 GPUs; "cartpole" is configs[0]; "pong" is configs[2] (one-hot actor + discount head,
# N=(T-1)*B); "scaled" is configs[4] (stoch 60 / deter 600 / hidden 600, H=30).  n_states is PER GPU unless `total` is set.
WORKLOADS = {
    "walker": dict(A=6, dist="tanh_normal", n_states=2500, H=15, use_pcont=False, S=30, D=200, Hd=200),
    "walker_large": dict(A=6, dist="tanh_normal", n_states=8192, H=15, use_pcont=False, S=30, D=200, Hd=200),
    "config4": dict(A=6, dist="tanh_normal", total=65536, H=15, use_pcont=False, S=30, D=200, Hd=200),
    "cartpole": dict(A=1, dist="tanh_normal", n_states=2500, H=15, use_pcont=False, S=30, D=200, Hd=200),
    "pong": dict(A=6, dist="one_hot", n_states=2450, H=15, use_pcont=True, S=30, D=200, Hd=200),
    "scaled": dict(A=6, dist="tanh_normal", n_states=2500, H=30, use_pcont=False, S=60, D=600, Hd=600),
}


def resolve_workload(name, world, n_states=0):
    """Per-rank view of a workload: adds n_states (per GPU), n_total and the scaling mode."""
    w = dict(WORKLOADS[name])
    if "total" in w:
        if w["total"] % world:
            raise SystemExit(f"{name}: {w['total']} start states do not split over {world} GPUs")
        w["n_states"], w["n_total"], w["scaling"] = w["total"] // world, w["total"], "strong"
    else:
        if n_states:
            w["n_states"] = n_states
        w["n_total"], w["scaling"] = w["n_states"] * world, "weak"
    w["name"] = name
    return w


def workload_string(w):
    """Identical in both arms (the driver compares config.workload across them)."""
    return (f"{w['name']}: imagination fwd+bwd, A={w['A']} {w['dist']}, {w['n_total']} start states x H={w['H']}, "
            f"S/D/Hd={w['S']}/{w['D']}/{w['Hd']}, pcont={w['use_pcont']}, random-init weights")


def flops_per_state(w, ah=200, al=3, vh=200, vl=3, rh=300, rl=3, ph=200, pl=3):
    """algorithmic FLOPs per imagined state, (fwd, bwd)."""
    A, S, D, Hd = w["A"], w["S"], w["D"], w["Hd"]
    F = S + D
    a_out = 2 * A if w["dist"] == "tanh_normal" else A
    f_tr = 2 * ((A + S) * Hd + 3 * Hd * D + 3 * D * D + D * Hd + 2 * Hd * S)
    f_pol = 2 * (F * ah + (al - 1) * ah * ah + ah * a_out)
    f_val = 2 * (F * vh + (vl - 1) * vh * vh + vh)
    f_rew = 2 * (F * rh + (rl - 1) * rh * rh + rh)
    f_pc = 2 * (F * ph + (pl - 1) * ph * ph + ph) if w["use_pcont"] else 0
    fwd = f_pol + f_tr + f_rew + 2 * f_val + f_pc
    bwd = f_tr + 2 * f_pol + f_rew + f_val + 2 * f_val + f_pc
    # the reverse-time rollout kernel alone: transition data-grad + actor data-grad (weight grads are
    # separate GEMM launches), per state
    # (the actor's first Linear gets no data gradient: its input is detached, rnns.py:238)
    bwd_rollout_kernel = f_tr + f_pol - 2 * F * ah
    fwd_rollout_kernel = f_tr + f_pol
    return dict(fwd=fwd, bwd=bwd, total=fwd + bwd, fwd_rollout_kernel=fwd_rollout_kernel,
                bwd_rollout_kernel=bwd_rollout_kernel)


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(bf16_tflops=float(p["bf16_tflops"]), bf16_tflops_sustained=float(p["bf16_tflops_sustained"]),
                    hbm_gbs=float(p["hbm_gbs"]), source="measured")
    return dict(bf16_tflops=1590.0, bf16_tflops_sustained=1400.0, hbm_gbs=6650.0, source="fallback")


def ncu_traffic(workload, n_states, precision, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed
    `ncu --set full` capture of this workload (profiles/ncu_traffic.json), or None when there is none."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(path):
        return None
    with open(path) as f:
        table = json.load(f)
    return table.get(f"{workload}:{n_states}:{precision}:{kernel}")


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of one GPU while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc, self.th = index, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index),
                 "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.th = threading.Thread(target=self._read, daemon=True)
        self.th.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        if self.th is not None:
            self.th.join(timeout=2)
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 6:
                continue
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except ValueError:
                continue
            for nm, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return dict(sm_mhz=(sm[len(sm) // 2] if sm else None), sm_max_mhz=mx, reasons=sorted(reasons),
                    samples=len(sm))


# ------------------------------------------------------------------------------------------------
# this repo's arm
# ------------------------------------------------------------------------------------------------
def build_model(w, device):
    import torch

    from dreamer_pytorch_b200.dropin.agent import AtariDreamerModel

    torch.manual_seed(0)
    model = AtariDreamerModel(action_shape=(w["A"],), action_dist=w["dist"], use_pcont=w["use_pcont"],
                              stochastic_size=w["S"], deterministic_size=w["D"], hidden_size=w["Hd"])
    return model.to(device).train()


def host_start_states(w, seed, n=None):
    """Synthetic posterior start states (SURVEY.md 8(d) config 4 recipe), pinned host memory."""
    import torch

    g = torch.Generator().manual_seed(seed)
    n = n or w["n_states"]
    stoch = torch.randn(n, w["S"], generator=g)
    deter = torch.tanh(torch.randn(n, w["D"], generator=g))
    if torch.cuda.is_available():
        stoch, deter = stoch.pin_memory(), deter.pin_memory()
    return stoch, deter


class Runner:
    """One workload on this rank: model, step graph, timed regions."""

    def __init__(self, w, args, dev, rank, world, flush, collective=True):
        import torch

        from dreamer_pytorch_b200.algo import imagination_losses
        from dreamer_pytorch_b200.dropin.rnns import RSSMState

        self.w, self.args, self.dev, self.rank, self.world, self.flush = w, args, dev, rank, world, flush
        self.N, self.H = w["n_states"], w["H"]
        self.model = build_model(w, dev)
        self.actor_params = list(self.model.action_decoder.parameters())
        self.value_params = list(self.model.value_model.parameters())
        self.grad_params = self.actor_params + self.value_params
        self.h_stoch, self.h_deter = host_start_states(w, 100 + rank)
        self.d_stoch, self.d_deter = self.h_stoch.to(dev), self.h_deter.to(dev)
        self.h_losses = torch.empty(2, dtype=torch.float32).pin_memory()
        self.collective = collective and world > 1
        self.graph = None
        self._imag, self._State = imagination_losses, RSSMState
        if not args.no_graph:
            # the whole step (noise, packing, both rollouts, heads, losses, all backward kernels and -- with several
            # ranks -- the gradient all-reduce) as ONE CUDA graph
            from dreamer_pytorch_b200.graph import ImaginationGraph

            self.graph = ImaginationGraph(self.model, self.N, self.H, use_pcont=w["use_pcont"],
                                          allreduce=self.collective)

    def eager_step(self, stoch, deter):
        from dreamer_pytorch_b200.parallel import allreduce_grads

        for p in self.grad_params:
            p.grad = None
        start = self._State(stoch, stoch, stoch, deter)  # mean/std of a start state are never read
        actor_loss, value_loss, _ = self._imag(self.model, start, self.H, use_pcont=self.w["use_pcont"])
        actor_loss.backward()
        value_loss.backward()
        if self.collective:
            allreduce_grads([self.actor_params, self.value_params])
        return actor_loss, value_loss

    def step(self, stoch, deter):
        if self.graph is None:
            return self.eager_step(stoch, deter)
        losses = self.graph.step(stoch, deter)
        if self.collective and not self.graph.allreduce_in_graph:
            from dreamer_pytorch_b200.parallel import allreduce_grads

            allreduce_grads([self.actor_params, self.value_params])
        return losses

    def barrier(self):
        import torch
        import torch.distributed as dist

        if self.world > 1 and self.collective:
            dist.barrier()
        torch.cuda.synchronize()

    def warm(self, n):
        for _ in range(n):
            self.flush.zero_()
            self.step(self.d_stoch, self.d_deter)
        self.barrier()

    def timed_device(self, steps):
        """Device-resident inputs, CUDA events per step, L2 flushed between steps.  Returns total ms."""
        import torch

        if self.graph is not None:
            self.graph.stoch.copy_(self.d_stoch); self.graph.deter.copy_(self.d_deter)
            self.d_stoch, self.d_deter = self.graph.stoch, self.graph.deter  # inputs resident: the captured buffers
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        self.barrier()
        t0 = time.perf_counter()
        for a, b in ev:
            self.flush.zero_()
            a.record()
            self.step(self.d_stoch, self.d_deter)
            b.record()
        self.barrier()
        return sum(a.elapsed_time(b) for a, b in ev), time.perf_counter() - t0

    def timed_e2e(self, steps):
        """End to end through the public API with HOST buffers: H2D of the start states and D2H of the losses inside."""
        import torch

        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        self.barrier()
        for a, b in ev:
            self.flush.zero_()
            torch.cuda.synchronize()
            a.record()
            if self.graph is not None:
                al, vl = self.step(self.h_stoch, self.h_deter)  # ImaginationGraph.step copies pinned host rows into its inputs
                self.h_losses.copy_(self.graph.losses, non_blocking=True)
            else:
                s = self.h_stoch.to(self.dev, non_blocking=True)
                d = self.h_deter.to(self.dev, non_blocking=True)
                al, vl = self.step(s, d)
                self.h_losses.copy_(torch.stack([al.detach(), vl.detach()]), non_blocking=True)
            b.record()
            b.synchronize()
        self.barrier()
        return sum(a.elapsed_time(b) for a, b in ev)

    def kernel_profile(self, steps):
        """Per-kernel CUDA-event times with every kernel in PLAIN stream order (nothing runs beside anything, gated
        kernels do not wait): the dominant kernel's isolated duration, reproducible from profiles/."""
        import ctypes as C

        import torch

        from dreamer_pytorch_b200 import _lib, algo

        lib = _lib.load()
        prev = algo.set_schedule("plain")
        try:
            for _ in range(2):
                self.eager_step(self.d_stoch, self.d_deter)
            torch.cuda.synchronize()
            lib.drm_prof_enable(1)
            for _ in range(steps):
                self.flush.zero_()
                self.eager_step(self.d_stoch, self.d_deter)
            torch.cuda.synchronize()
        finally:
            algo.set_schedule(prev)
        prof = {}
        names = ["pack", "rollout_fwd", "rollout_bwd", "mlp_fwd", "mlp_bwd", "wgrad", "loss"]
        for i, nm in enumerate(names):
            tot, cnt = C.c_double(0), C.c_int(0)
            lib.drm_prof_read(i, C.byref(tot), C.byref(cnt))
            prof[nm] = dict(ms_per_step=tot.value / steps, launches_per_step=cnt.value / steps)
        lib.drm_prof_enable(0)
        if self.graph is not None:
            for p, g in zip(self.graph.params, self.graph.grads):
                p.grad = g
        return prof


def time_full_paths(w, dev, precision, reps=10):
    """t_full: the step widened to rollout_representation + KL + the RSSM / reward part of
    model_loss.backward() (conv excluded: synthetic obs_embed), and full_update: the whole Dreamer.loss with the conv
    encoder / decoder + the three backwards (BASELINE.json configs[1] "full agent step").  Eager launches, median."""
    import torch

    from dreamer_pytorch_b200.algo import dreamer_loss
    from dreamer_pytorch_b200.precision import precision as prec_ctx

    B, T, H, A = 50, 50, w["H"], w["A"]
    model = build_model(w, dev)
    g = torch.Generator().manual_seed(5)
    embed = torch.randn(T, B, 1024, generator=g).to(dev)
    obs = (torch.randint(0, 256, (T, B, 3, 64, 64), generator=g, dtype=torch.uint8).to(dev).float() / 255.0 - 0.5)
    if w["dist"] == "tanh_normal":
        act = (torch.rand(T, B, A, generator=g) * 2 - 1).to(dev)
    else:
        act = torch.nn.functional.one_hot(torch.randint(0, A, (T, B), generator=g), A).float().to(dev)
    rew = torch.randn(T, B, 1, generator=g).to(dev)
    done = (torch.rand(T, B, 1, generator=g) < 0.01).to(dev)
    out = {}
    from dreamer_pytorch_b200.precision import conv_mode, get_conv_mode

    cases = [("t_full", dict(observation=None, obs_embed=embed), get_conv_mode()),
             ("full_update", dict(observation=obs), get_conv_mode())]
    if precision == "bf16":
        cases.append(("full_update_conv_bf16", dict(observation=obs), "bf16"))
    for key, kw, cmode in cases:
        def run():
            model.zero_grad(set_to_none=True)
            ml, al, vl, _ = dreamer_loss(model, kw["observation"], act, rew, done, horizon=H, use_pcont=w["use_pcont"],
                                         obs_embed=kw.get("obs_embed"))
            ml.backward(); al.backward(); vl.backward()
        ts = []
        with prec_ctx(precision), conv_mode(cmode):
            for i in range(reps + 3):
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(); run(); b.record()
                torch.cuda.synchronize()
                if i >= 3:
                    ts.append(a.elapsed_time(b))
        ms = statistics.median(ts)
        n_imag = ((T - 1) * B if w["use_pcont"] else T * B) * H
        out[key] = dict(ms=ms, imagined_states_per_s=n_imag / (ms * 1e-3), observed_states=T * B, launch="eager",
                        conv=cmode if key != "t_full" else None,
                        what=("rollout_representation + KL + reward loss + imagination, 3 backwards; conv excluded (synthetic "
                              "obs_embed)") if key == "t_full" else
                             "whole Dreamer.loss incl. conv encoder/decoder + image log-prob, 3 backwards")
    # the whole update -- loss, 3 backwards, clip + Adam of the three groups -- as ONE graph replay (graph.UpdateGraph),
    # fed from the device-resident replay (replay.DeviceSequenceReplay): what one iteration of the training loop costs
    try:
        from collections import namedtuple

        import numpy as np

        from dreamer_pytorch_b200.graph import UpdateGraph
        from dreamer_pytorch_b200.replay import DeviceSequenceReplay

        Ex = namedtuple("Ex", ["observation", "action", "reward", "done"])
        rep = DeviceSequenceReplay(Ex(np.zeros((3, 64, 64), np.uint8), np.zeros(A, np.float32), np.float32(0),
                                      np.bool_(False)), size=20000, B=10, device=dev)
        gfill = torch.Generator().manual_seed(6)
        for _ in range(4):
            rep.append_samples(Ex(torch.randint(0, 256, (500, 10, 3, 64, 64), generator=gfill, dtype=torch.uint8),
                                  torch.rand(500, 10, A, generator=gfill) * 2 - 1, torch.randn(500, 10, generator=gfill),
                                  torch.rand(500, 10, generator=gfill) < 0.01))
        with prec_ctx(precision):
            ug = UpdateGraph(model, B, T, horizon=H, use_pcont=w["use_pcont"])
            ts, ts_s = [], []
            for i in range(reps + 3):
                torch.cuda.synchronize()
                a, m, b = (torch.cuda.Event(enable_timing=True) for _ in range(3))
                a.record()
                batch = rep.sample_batch(B, T)
                m.record()
                ug.step(batch)
                b.record()
                torch.cuda.synchronize()
                if i >= 3:
                    ts.append(a.elapsed_time(b)); ts_s.append(a.elapsed_time(m))
        out["full_update_graph"] = dict(
            ms=statistics.median(ts), replay_sample_ms=statistics.median(ts_s), launch="one CUDA graph replay",
            what="replay sample on the device (gather kernels) + whole Dreamer.loss + 3 backwards + clip_grad_norm_ + Adam "
                 "of the world-model / actor / value groups", conv=get_conv_mode())
    except Exception as e:
        out["full_update_graph"] = dict(error=f"{type(e).__name__}: {e}"[:300])
    return out


def run_b200(args):
    import torch
    import torch.distributed as dist

    from dreamer_pytorch_b200 import _lib
    from dreamer_pytorch_b200.precision import set_precision

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); use --impl reference for the CPU arm")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    _lib.ensure_init()
    set_precision(args.precision)
    name = args.workload or ("walker" if world == 1 else "config4")
    w = resolve_workload(name, world, args.n_states)
    N, H = w["n_states"], w["H"]
    flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2
    steps, warm = args.steps, max(args.warmup, 3)

    run = Runner(w, args, dev, rank, world, flush)
    run.warm(warm)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = lib.drm_launch_count()
    ms_dev, t_wall = run.timed_device(steps)
    launches = int(lib.drm_launch_count() - launches0) if run.graph is None else run.graph.launches_per_replay * steps
    # sustained block: the same step for >= 1.5 s, so that the clock record has > 10 samples under load.  Every rank must
    # replay the same number of steps (each replay holds the gradient all-reduce): agree on the slowest rank's time first.
    if world > 1:
        t0 = torch.tensor([ms_dev], dtype=torch.float64, device=dev)
        dist.all_reduce(t0, op=dist.ReduceOp.MAX)
        ms_dev = float(t0[0])
    sus_steps = max(steps, int(1.5e3 / max(ms_dev / steps, 1e-3)) + 1) if args.sustained else 0
    sus_steps = min(sus_steps, 20000)
    ms_sus = run.timed_device(sus_steps)[0] if sus_steps else 0.0
    ms_e2e = run.timed_e2e(steps)
    clocks = sampler.stop() if rank == 0 else None
    prof = run.kernel_profile(min(steps, 5))

    t = torch.tensor([ms_dev, ms_e2e, ms_sus], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_dev, ms_e2e, ms_sus = float(t[0]), float(t[1]), float(t[2])

    # ---- secondary blocks at N > 1: weak-scaling walker line, and the primary workload on ONE GPU (rank 0)
    secondary = {}
    if world > 1 and not args.no_secondary:
        if w["scaling"] == "strong":
            ww = resolve_workload("walker", world)
            r2 = Runner(ww, args, dev, rank, world, flush)
            r2.warm(warm)
            ms2, _ = r2.timed_device(steps)
            t2 = torch.tensor([ms2], dtype=torch.float64, device=dev)
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
            secondary["weak"] = dict(workload=workload_string(ww), scaling="weak", ms_per_step=float(t2[0]) / steps,
                                     value=ww["n_total"] * ww["H"] * steps / (float(t2[0]) * 1e-3), unit=UNIT)
            del r2
        torch.cuda.synchronize()
        dist.barrier()
        if rank == 0 and w["scaling"] == "strong":
            w1 = resolve_workload(name, 1)
            r1 = Runner(w1, args, dev, 0, 1, flush, collective=False)
            r1.warm(warm)
            ms1, _ = r1.timed_device(steps)
            secondary["one_gpu_same_workload"] = dict(
                workload=workload_string(w1), ms_per_step=ms1 / steps, value=w1["n_total"] * w1["H"] * steps / (ms1 * 1e-3),
                unit=UNIT, note="the whole primary workload on rank 0's GPU alone, same process, same run")
            del r1
        torch.cuda.synchronize()
        dist.barrier()
    if rank != 0:
        _finish(world, dist)
        return

    states = N * H * world * steps
    fl = flops_per_state(w)
    peaks = measured_peaks()
    value = states / (ms_dev * 1e-3)
    # dominant kernel: whichever persistent rollout kernel takes longer per step
    dom = "rollout_bwd" if prof["rollout_bwd"]["ms_per_step"] >= prof["rollout_fwd"]["ms_per_step"] else "rollout_fwd"
    dom_ms = prof[dom]["ms_per_step"] / max(prof[dom]["launches_per_step"], 1)
    dom_flops = fl["bwd_rollout_kernel" if dom == "rollout_bwd" else "fwd_rollout_kernel"] * N * H
    achieved = dom_flops / (dom_ms * 1e-3) / 1e12 if dom_ms > 0 else 0.0
    peak = peaks["bf16_tflops_sustained"]
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps,
        "warmup": warm, "ms_per_step": ms_dev / steps, "higher_is_better": True,
        "scaling": w["scaling"], "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
        "config": {"workload": workload_string(w),
                   "states_per_gpu_per_step": N * H, "global_states_per_step": N * H * world,
                   "parallelism": f"dp{world} over start states ({N} per GPU); "
                                  + ("one-shot all-reduce of the actor+value gradient bucket over NVLink peer memory (drm_allreduce)"
                                     if (run.graph is not None and getattr(run.graph, "allreduce_kind", None) == "peer_oneshot")
                                     else "NCCL all-reduce of actor+value grads")
                                  + (" inside the step's CUDA graph" if (run.graph is not None and run.graph.allreduce_in_graph) else ""),
                   "l2": "512 MiB buffer written between timed steps (L2 flush); per-step CUDA events",
                   "launch": "eager" if run.graph is None else "one CUDA graph replay per step (ImaginationGraph)",
                   "flops_per_state": fl["total"]},
        "e2e": {"value": states / (ms_e2e * 1e-3), "unit": UNIT,
                "h2d_bytes_per_step": int(run.h_stoch.numel() * 4 + run.h_deter.numel() * 4), "d2h_bytes_per_step": 8},
        "gpu_launches": launches,
        "roofline": {"bound": "tensor", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": achieved / peak, "traffic": ncu_traffic(name, N, args.precision, dom),
                     "peak_source": peaks["source"] + " bf16 sustained", "kernel_ms": dom_ms,
                     "kernel_ms_how": "CUDA events around the launch, every kernel in plain stream order "
                                      "(algo.set_schedule('plain')), L2 flushed between steps",
                     "whole_step_tflops": value * fl["total"] / 1e12,
                     "whole_step_frac": value * fl["total"] / 1e12 / peak},
        "kernels_ms_per_step": {k: round(v["ms_per_step"], 4) for k, v in prof.items()},
        "kernels_note": "isolated durations (plain stream order); in the timed step kernels run beside each other, so "
                        "these add up to more than ms_per_step",
        "wall_ms_per_step": t_wall * 1e3 / steps,
        "clocks": clocks,
    }
    if sus_steps:
        line["sustained"] = dict(steps=sus_steps, seconds=ms_sus * 1e-3, ms_per_step=ms_sus / sus_steps,
                                 value=N * H * world * sus_steps / (ms_sus * 1e-3), unit=UNIT)
    line.update(secondary)
    if world == 1:
        del run
        torch.cuda.empty_cache()
        if not args.no_full:
            try:
                line.update(time_full_paths(w, dev, args.precision))
            except Exception as e:  # a widened path must never cost the headline line
                line["t_full"] = dict(error=f"{type(e).__name__}: {e}"[:300])
        if not args.no_reference_cuda:
            line["reference_cuda"] = reference_cuda(w, dev, value, line)
        if not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(w, args.cpu_seconds)
    print(json.dumps(line), flush=True)
    _finish(world, dist)


def _finish(world, dist):
    """Multi-rank exit: meet once more, then leave without tearing NCCL down -- destroying a communicator whose collectives
    are referenced by captured CUDA graphs can block forever (seen: a 2-GPU run that had printed its line hung in
    destroy_process_group until the box's time limit)."""
    if world <= 1:
        return
    sys.stdout.flush()
    sys.stderr.flush()
    try:
        dist.barrier()
    finally:
        os._exit(0)


# ------------------------------------------------------------------------------------------------
# Reference arms.  The UNMODIFIED reference (oracle/_ref, shipped by oracle/fetch_ref.py; /root/reference in the build
# container) through its own modules: RSSMRollout.rollout_policy, AgentModel.policy, DenseModel, Dreamer.compute_return,
# FreezeParameters -- sequenced exactly as dreamer_algo.py:283-337 + :176-180 (oracle/ref_harness.imagination_step_fn).
# If the copy is missing the CPU arm falls back to the oracle restatement (kind = "port").
# ------------------------------------------------------------------------------------------------
def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def reference_step_fn(w, n_states, device="cpu"):
    """(fn, states_per_call, kind): fn() = one fwd+bwd imagination pass over n_states start states."""
    import torch

    from oracle import ref_harness as RH

    if RH.reference_available():
        RH.import_reference()
        from dreamer.models.rnns import RSSMState

        model = RH.default_reference_model(w["A"], w["dist"], w["use_pcont"], device=device, stochastic_size=w["S"],
                                           deterministic_size=w["D"], hidden_size=w["Hd"])
        algo = RH.reference_algo(model, 50, 50, w["H"], use_pcont=w["use_pcont"], device=device)
        g = torch.Generator().manual_seed(7)
        stoch = torch.randn(n_states, w["S"], generator=g).to(device)
        deter = torch.tanh(torch.randn(n_states, w["D"], generator=g)).to(device)
        fn = RH.imagination_step_fn(model, algo, RSSMState(stoch, stoch, stoch, deter))
        RH.set_as_shipped(False)
        return fn, n_states * w["H"], "reference"
    return oracle_step_fn(w, n_states) + ("port",)


def oracle_step_fn(w, n_states):
    import torch

    from oracle import dreamer_oracle as O

    shapes = O.weight_shapes(w["A"], S=w["S"], D=w["D"], Hd=w["Hd"], dist=w["dist"], use_pcont=w["use_pcont"])
    weights = O.make_weights(shapes, 0)
    g = torch.Generator().manual_seed(7)
    stoch = torch.randn(n_states, w["S"], generator=g)
    deter = torch.tanh(torch.randn(n_states, w["D"], generator=g))
    nz = O.make_noise(1, 1, n_states, w["H"], w["S"], w["A"], 8, dist=w["dist"])
    wa = {k: v.clone().requires_grad_(True) for k, v in weights.items() if k.startswith("action_decoder.")}
    wv = {k: v.clone().requires_grad_(True) for k, v in weights.items() if k.startswith("value_model.")}

    def fn():
        for p in list(wa.values()) + list(wv.values()):
            p.grad = None
        al, vl, _ = O.imagination_losses(weights, stoch, deter, nz["noise_act"], nz["eps_prior_img"], w["H"],
                                         dist=w["dist"], use_pcont=w["use_pcont"], w_actor=wa, w_value=wv)
        al.backward()
        vl.backward()
        return al.detach(), vl.detach()

    return fn, n_states * w["H"]


def _timed_calls(fn, budget_s, max_calls, warm=1, sync=None):
    for _ in range(warm):
        fn()
    ts = []
    t_all = time.perf_counter()
    while True:
        if sync:
            sync()
        t0 = time.perf_counter()
        fn()
        if sync:
            sync()
        ts.append(time.perf_counter() - t0)
        if time.perf_counter() - t_all >= budget_s or len(ts) >= max_calls:
            break
    return ts


def cpu_baseline(w, budget_s=12.0):
    """The reference on the host cores, same workload (full n_states rows): all threads (primary, anomaly off), all threads
    as shipped (anomaly + validation on), and one thread."""
    import torch

    from oracle import ref_harness as RH

    threads = host_threads()
    n = min(w["n_states"], 8192)
    torch.set_num_threads(threads)
    fn, per_call, kind = reference_step_fn(w, n)
    ts = _timed_calls(fn, budget_s, 50)
    med = statistics.median(ts)
    out = dict(value=per_call / med, unit=UNIT, cores=threads, kind=kind, cpu=cpu_model(), torch=torch.__version__,
               sample=f"median of {len(ts)} fwd+bwd passes over {n} start states x H={w['H']} "
                      f"({'unmodified reference modules (oracle/_ref)' if kind == 'reference' else 'oracle/dreamer_oracle.py restatement'}, "
                      f"torch CPU fp32, {threads} threads, anomaly mode off)")
    if kind == "reference":
        RH.set_as_shipped(True)
        try:
            ts2 = _timed_calls(fn, min(budget_s, 6.0), 10)
            out["as_shipped"] = dict(value=per_call / statistics.median(ts2), passes=len(ts2),
                                     note="autograd anomaly mode on (dreamer_algo.py:16) + distribution validation on")
        finally:
            RH.set_as_shipped(False)
    torch.set_num_threads(1)
    try:
        ts1 = _timed_calls(fn, min(budget_s, 8.0), 5, warm=1)
        out["one_thread"] = dict(value=per_call / statistics.median(ts1), passes=len(ts1), cores=1)
    finally:
        torch.set_num_threads(threads)
    return out


def reference_cuda(w, dev, our_value, line):
    """The reference's eager CUDA path on the same GPU, same workload: the north star's '>= 10x' denominator.
    Median of >= 20 synchronize-bracketed passes after 3 warm-ups, best case (anomaly + validation off) and as shipped."""
    import torch

    from oracle import ref_harness as RH

    if not RH.reference_available():
        return dict(unavailable="oracle/_ref missing: run `python -m oracle.fetch_ref` where /root/reference exists")
    n = w["n_states"]
    out = {}
    try:
        fn, per_call, _ = reference_step_fn(w, n, device=dev)
        sync = torch.cuda.synchronize
        ts = _timed_calls(fn, 60.0, 20, warm=3, sync=sync)
        med = statistics.median(ts)
        out["best_case"] = dict(value=per_call / med, ms_per_step=med * 1e3, passes=len(ts), unit=UNIT,
                                note="anomaly mode off, distribution validation off; fp32 eager, torch defaults")
        RH.set_as_shipped(True)
        try:
            ts2 = _timed_calls(fn, 30.0, 10, warm=2, sync=sync)
            med2 = statistics.median(ts2)
            out["as_shipped"] = dict(value=per_call / med2, ms_per_step=med2 * 1e3, passes=len(ts2), unit=UNIT)
        finally:
            RH.set_as_shipped(False)
        out["speedup_vs_best_case"] = dict(device_resident=our_value / out["best_case"]["value"],
                                           e2e=line["e2e"]["value"] / out["best_case"]["value"], dtype=line["dtype"])
        # the reference's whole Dreamer.loss + 3 backwards (conv included), B=50 T=50: the denominator of full_update
        if "full_update" in line and "ms" in line.get("full_update", {}):
            model = RH.default_reference_model(w["A"], w["dist"], w["use_pcont"], device=dev, stochastic_size=w["S"],
                                               deterministic_size=w["D"], hidden_size=w["Hd"])
            algo = RH.reference_algo(model, 50, 50, w["H"], use_pcont=w["use_pcont"], device=dev)
            samples = RH.synthetic_samples(w["A"], w["dist"], 50, 50, device=dev)
            full = RH.full_loss_step_fn(model, algo, samples)
            ts3 = _timed_calls(full, 60.0, 10, warm=2, sync=sync)
            med3 = statistics.median(ts3)
            out["full_update"] = dict(ms=med3 * 1e3, passes=len(ts3),
                                      speedup=med3 * 1e3 / line["full_update"]["ms"],
                                      what="reference Dreamer.loss + 3 backwards, eager CUDA, best case")
            if "ms" in line.get("full_update_graph", {}):
                full_opt = RH.full_loss_step_fn(model, algo, samples, optimize=True)
                ts4 = _timed_calls(full_opt, 60.0, 10, warm=2, sync=sync)
                med4 = statistics.median(ts4)
                out["full_update_with_optimizers"] = dict(
                    ms=med4 * 1e3, passes=len(ts4), speedup=med4 * 1e3 / line["full_update_graph"]["ms"],
                    what="reference training-loop iteration (loss, 3 backwards, 3 x clip_grad_norm_ + Adam.step, 13 .item() "
                         "reads; batch already on the device), eager CUDA, best case; speedup vs full_update_graph")
    except Exception as e:
        out["error"] = f"{type(e).__name__}: {e}"[:300]
    return out


def run_reference(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch

    name = args.workload or ("walker" if world == 1 else "config4")
    w = resolve_workload(name, world, args.n_states)
    threads = host_threads()
    torch.set_num_threads(threads)
    # a bounded sample of the workload per step: at most 8192 start states (the whole of it for walker / pong / cartpole)
    n = min(w["n_total"], 8192)
    fn, per_call, kind = reference_step_fn(w, n)
    warm = max(args.warmup, 3)
    for _ in range(warm):
        fn()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fn()
    el = time.perf_counter() - t0
    value = per_call * args.steps / el
    fl = flops_per_state(w)
    src = "unmodified reference modules (oracle/_ref)" if kind == "reference" else "oracle/dreamer_oracle.py restatement"
    sample = (f"each step = one fwd+bwd pass over {n} start states x H={w['H']} on the host cores ({src}, torch CPU fp32, "
              f"{threads} threads, anomaly mode off); one process regardless of --gpus")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": warm, "ms_per_step": el * 1e3 / args.steps, "higher_is_better": True,
        "scaling": w["scaling"], "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
        "config": {"workload": workload_string(w), "flops_per_state": fl["total"], "states_per_step": per_call},
        "cpu_baseline": dict(value=value, unit=UNIT, cores=threads, kind=kind, sample=sample, cpu=cpu_model()),
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS),
                    help="default: walker (BASELINE configs[1]) on one GPU, config4 (65 536 start states sharded) on several")
    ap.add_argument("--n-states", type=int, default=0, help="override start states per GPU (weak workloads)")
    ap.add_argument("--precision", default="bf16", choices=["fp32", "bf16"],
                    help="bf16: tcgen05 tensor-core kernels (BASELINE.json configs[1]); fp32: FFMA kernels")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-reference-cuda", action="store_true")
    ap.add_argument("--no-full", action="store_true", help="skip the t_full / full_update blocks")
    ap.add_argument("--no-secondary", action="store_true", help="N > 1: skip the weak / one-GPU secondary blocks")
    ap.add_argument("--no-sustained", dest="sustained", action="store_false")
    ap.add_argument("--no-graph", action="store_true", help="launch the step eagerly instead of replaying one CUDA graph")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
