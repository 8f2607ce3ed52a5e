#!/usr/bin/env python
"""bench.py -- imagined RSSM states / second (fwd+bwd, H=15) on B200.

One "step" = one pass of the hot path named by BASELINE.json over one batch of synthetic start
states: H-step ``rollout_policy`` (actor MLP -> action sample -> RSSM
transition) + reward/value heads + lambda-returns + actor and value losses, forward, then
``actor_loss.backward()`` and ``value_loss.backward()`` (gradients into the actor and value
parameters), plus -- for N > 1 GPUs -- the NCCL all-reduce of those gradients.  Start states are
sharded across ranks (weak scaling: every rank owns ``n_states`` start states).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload walker|walker_large|cartpole|pong]
    python bench.py --impl reference ...        # the CPU restatement (oracle/) on the host cores

Prints ONE JSON line (rank 0).  ``value`` is timed with CUDA events, inputs resident in HBM;
``e2e`` goes through the public Python API with pinned HOST buffers, host<->device copies and the
loss read-back inside the timed region.  ``roofline`` is the dominant kernel (the reverse-time
rollout kernel) against the measured bf16 tensor peak; ``cpu_baseline`` is the oracle timed on the
box's host cores on a bounded sample (rank 0, N=1 only).
"""
from __future__ import annotations

import argparse
import json
import os
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
# 2500 start states, H=15); "walker_large" is one GPU's shard of configs[3] (65 536 / 8); "cartpole" is
# configs[0]; "pong" is configs[2] (one-hot actor + discount head, N=(T-1)*B).
WORKLOADS = {
    "walker": dict(A=6, dist="tanh_normal", n_states=2500, H=15, use_pcont=False, S=30, D=200, Hd=200),
    "walker_large": dict(A=6, dist="tanh_normal", n_states=8192, H=15, use_pcont=False, S=30, D=200, Hd=200),
    "cartpole": dict(A=1, dist="tanh_normal", n_states=2500, H=15, use_pcont=False, S=30, D=200, Hd=200),
    "pong": dict(A=6, dist="one_hot", n_states=2450, H=15, use_pcont=True, S=30, D=200, Hd=200),
}


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
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
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


def host_start_states(w, seed):
    """Synthetic posterior start states (SURVEY.md 8(d) config 4 recipe), pinned host memory."""
    import torch

    g = torch.Generator().manual_seed(seed)
    n = w["n_states"]
    stoch = torch.randn(n, w["S"], generator=g)
    deter = torch.tanh(torch.randn(n, w["D"], generator=g))
    if torch.cuda.is_available():
        stoch, deter = stoch.pin_memory(), deter.pin_memory()
    return stoch, deter


def run_b200(args):
    import torch
    import torch.distributed as dist

    from dreamer_pytorch_b200 import _lib
    from dreamer_pytorch_b200.algo import imagination_losses
    from dreamer_pytorch_b200.dropin.rnns import RSSMState

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
    from dreamer_pytorch_b200.precision import set_precision

    set_precision(args.precision)
    w = dict(WORKLOADS[args.workload])
    if args.n_states:
        w["n_states"] = args.n_states
    N, H = w["n_states"], w["H"]
    model = build_model(w, dev)
    actor_params = list(model.action_decoder.parameters())
    value_params = list(model.value_model.parameters())
    grad_params = actor_params + value_params
    h_stoch, h_deter = host_start_states(w, 100 + rank)
    d_stoch, d_deter = h_stoch.to(dev), h_deter.to(dev)
    flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2
    h_losses = torch.empty(2, dtype=torch.float32).pin_memory()

    from dreamer_pytorch_b200.parallel import allreduce_grads as _allreduce

    def allreduce_grads():
        # equal shards: mean of shard means = global mean; one flat bucket per optimizer group
        _allreduce([actor_params, value_params])

    def eager_step(stoch, deter):
        for p in grad_params:
            p.grad = None
        start = RSSMState(stoch, stoch, stoch, deter)  # mean/std of a start state are never read
        actor_loss, value_loss, _ = imagination_losses(model, start, H, use_pcont=w["use_pcont"])
        actor_loss.backward()
        value_loss.backward()
        if world > 1:
            allreduce_grads()
        return actor_loss, value_loss

    graph = None
    if not args.no_graph:
        # the whole step (noise, packing, both rollouts, heads, losses, all backward kernels) as ONE CUDA graph
        from dreamer_pytorch_b200.graph import ImaginationGraph

        graph = ImaginationGraph(model, N, H, use_pcont=w["use_pcont"])

    def step(stoch, deter):
        if graph is None:
            return eager_step(stoch, deter)
        losses = graph.step(stoch, deter)
        if world > 1:
            allreduce_grads()
        return losses

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up
    for _ in range(max(args.warmup, 3)):
        flush.zero_()
        step(d_stoch, d_deter)
    barrier()

    # ---- timed region 1: device-resident inputs, CUDA events per step, L2 flushed between steps
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    if graph is not None:
        graph.stoch.copy_(d_stoch); graph.deter.copy_(d_deter)
        d_stoch, d_deter = graph.stoch, graph.deter  # inputs resident in HBM: the captured buffers themselves
    else:
        lib.drm_prof_enable(1)
    launches0 = lib.drm_launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    t_wall0 = time.perf_counter()
    for a, b in ev:
        flush.zero_()
        a.record()
        step(d_stoch, d_deter)
        b.record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = int(lib.drm_launch_count() - launches0) if graph is None else graph.launches_per_replay * args.steps
    ms_dev = sum(a.elapsed_time(b) for a, b in ev)
    import ctypes as C

    prof_steps = args.steps
    if graph is not None:
        # per-kernel CUDA-event times (roofline.kernel_ms): the same step run eagerly -- events cannot be read inside
        # a replayed graph -- with the same L2 flush between steps
        prof_steps = min(args.steps, 5)
        lib.drm_prof_enable(1)
        for _ in range(prof_steps):
            flush.zero_()
            eager_step(d_stoch, d_deter)
        torch.cuda.synchronize()
        for p, g in zip(graph.params, graph.grads):
            p.grad = g

    prof = {}
    names = ["pack", "rollout_fwd", "rollout_bwd", "mlp_fwd", "mlp_bwd", "wgrad", "loss"]
    for i, nm in enumerate(names):
        tot, cnt = C.c_double(0), C.c_int(0)
        lib.drm_prof_read(i, C.byref(tot), C.byref(cnt))
        prof[nm] = dict(ms_per_step=tot.value / prof_steps, launches_per_step=cnt.value / prof_steps)
    lib.drm_prof_enable(0)

    # ---- timed region 2: end to end through the public API with HOST buffers
    barrier()
    ev2 = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t0 = time.perf_counter()
    for a, b in ev2:
        flush.zero_()
        torch.cuda.synchronize()
        a.record()
        if graph is not None:
            al, vl = step(h_stoch, h_deter)  # ImaginationGraph.step copies pinned host rows straight into its input buffers
        else:
            s = h_stoch.to(dev, non_blocking=True)
            d = h_deter.to(dev, non_blocking=True)
            al, vl = step(s, d)
        h_losses.copy_(graph.losses if graph is not None else torch.stack([al.detach(), vl.detach()]), non_blocking=True)
        b.record()
        b.synchronize()
    barrier()
    ms_e2e = sum(a.elapsed_time(b) for a, b in ev2)
    clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([ms_dev, ms_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_dev, ms_e2e = float(t[0]), float(t[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    states = N * H * world * args.steps
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
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
        "config": {"workload": f"{args.workload}: imagination fwd+bwd, A={w['A']} {w['dist']}, "
                               f"{N} start states per GPU x H={H}, S/D/Hd={w['S']}/{w['D']}/{w['Hd']}, "
                               f"pcont={w['use_pcont']}, random-init weights",
                   "states_per_gpu_per_step": N * H, "global_states_per_step": N * H * world,
                   "parallelism": f"dp{world} over start states; NCCL all-reduce of actor+value grads",
                   "l2": "512 MiB buffer written between timed steps (L2 flush); per-step CUDA events",
                   "launch": "eager" if graph is None else "one CUDA graph replay per step (ImaginationGraph); kernels_ms_per_step "
                                                           "from an eager pass of the same step",
                   "flops_per_state": fl["total"]},
        "e2e": {"value": states / (ms_e2e * 1e-3), "unit": UNIT,
                "h2d_bytes_per_step": int(h_stoch.numel() * 4 + h_deter.numel() * 4), "d2h_bytes_per_step": 8},
        "gpu_launches": launches,
        "roofline": {"bound": "tensor", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": achieved / peak, "traffic": ncu_traffic(args.workload, N, args.precision, dom), "peak_source": peaks["source"] + " bf16 sustained",
                     "kernel_ms": dom_ms, "whole_step_tflops": value * fl["total"] / 1e12},
        "kernels_ms_per_step": {k: round(v["ms_per_step"], 4) for k, v in prof.items()},
        "kernels_note": "CUDA-event durations of the launches; kernels run beside each other (gated ones include their "
                        "waiting), so they add up to more than ms_per_step",
        "wall_ms_per_step": t_wall * 1e3 / args.steps,
        "clocks": clocks,
    }
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(w, args.cpu_seconds)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle (a PyTorch-CPU restatement of the reference's algorithm, oracle/dreamer_oracle.py)
# timed on the host cores.  /root/reference does not exist on the GPU box, so kind = "port".
# ------------------------------------------------------------------------------------------------
def oracle_step_fn(w, n_states, threads):
    """Returns (fn, states_per_call): fn() runs one fwd+bwd imagination pass of the oracle on n_states rows."""
    import torch

    from oracle import dreamer_oracle as O

    torch.set_num_threads(threads)
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
        return float(al.detach()), float(vl.detach())

    return fn, n_states * w["H"]


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_baseline(w, budget_s=15.0):
    """Oracle on the host cores for about ``budget_s`` seconds on the same workload (full n_states rows)."""
    threads = host_threads()
    fn, per_call = oracle_step_fn(w, w["n_states"], threads)
    fn()  # warm-up
    t0 = time.perf_counter()
    calls = 0
    while True:
        fn()
        calls += 1
        el = time.perf_counter() - t0
        if el >= budget_s or calls >= 50:
            break
    return dict(value=per_call * calls / el, unit=UNIT, cores=threads, kind="port",
                sample=f"{calls} fwd+bwd passes of {w['n_states']} start states x H={w['H']} "
                       f"(oracle/dreamer_oracle.py, torch CPU fp32, {threads} threads, anomaly mode off)")


def run_reference(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w = dict(WORKLOADS[args.workload])
    if args.n_states:
        w["n_states"] = args.n_states
    threads = host_threads()
    fn, per_call = oracle_step_fn(w, w["n_states"], threads)
    for _ in range(max(1, min(args.warmup, 2))):
        fn()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fn()
    el = time.perf_counter() - t0
    value = per_call * args.steps / el
    fl = flops_per_state(w)
    sample = (f"each step = one fwd+bwd pass over {w['n_states']} start states x H={w['H']} on the host cores "
              f"(oracle/dreamer_oracle.py restatement of the reference, torch CPU fp32, {threads} threads); "
              f"one process regardless of --gpus")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(1, min(args.warmup, 2)), "ms_per_step": el * 1e3 / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: imagination fwd+bwd, A={w['A']} {w['dist']}, "
                               f"{w['n_states']} start states x H={w['H']}, S/D/Hd={w['S']}/{w['D']}/{w['Hd']}, "
                               f"pcont={w['use_pcont']}, random-init weights",
                   "flops_per_state": fl["total"]},
        "cpu_baseline": dict(value=value, unit=UNIT, cores=threads, kind="port", sample=sample),
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
    ap.add_argument("--workload", default="walker", choices=sorted(WORKLOADS))
    ap.add_argument("--n-states", type=int, default=0, help="override start states per GPU")
    ap.add_argument("--precision", default="bf16", choices=["fp32", "bf16"],
                    help="bf16: tcgen05 tensor-core kernels (BASELINE.json configs[1]); fp32: FFMA kernels")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch the step eagerly instead of replaying one CUDA graph")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
