"""Per-kernel device time of ONE whole update (Dreamer.loss + 3 backwards) through algo.dreamer_loss, grouped by kernel
name (kineto/CUPTI through torch.profiler).  Answers "what share of a full update is the imagination step now".

    python tools/full_trace.py [--precision bf16] [--rssm-only] > gpurun_out/full_trace.txt
"""
import argparse
import collections
import json
import os
import sys
import tempfile

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--rssm-only", action="store_true", help="synthetic obs_embed: no conv (t_full of SURVEY 8(d))")
    ap.add_argument("--workload", default="walker")
    a = ap.parse_args()
    from bench import WORKLOADS, build_model
    from dreamer_pytorch_b200.algo import dreamer_loss
    from dreamer_pytorch_b200.precision import set_precision

    dev = torch.device("cuda:0")
    set_precision(a.precision)
    w = dict(WORKLOADS[a.workload])
    model = build_model(w, dev)
    B, T, H, A = 50, 50, w["H"], w["A"]
    g = torch.Generator().manual_seed(5)
    embed = torch.randn(T, B, 1024, generator=g).to(dev)
    obs = torch.randint(0, 256, (T, B, 3, 64, 64), generator=g, dtype=torch.uint8).to(dev).float() / 255.0 - 0.5
    if w["dist"] == "tanh_normal":
        act = (torch.rand(T, B, A, generator=g) * 2 - 1).to(dev)
    else:
        act = torch.nn.functional.one_hot(torch.randint(0, A, (T, B), generator=g), A).float().to(dev)
    rew = torch.randn(T, B, 1, generator=g).to(dev)
    done = (torch.rand(T, B, 1, generator=g) < 0.01).to(dev)

    def run():
        model.zero_grad(set_to_none=True)
        ml, al, vl, _ = dreamer_loss(model, None if a.rssm_only else obs, act, rew, done, horizon=H,
                                     use_pcont=w["use_pcont"], obs_embed=embed if a.rssm_only else None)
        ml.backward(); al.backward(); vl.backward()

    for _ in range(4):
        run()
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    reps = 3
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(reps):
            run()
            torch.cuda.synchronize()
    path = os.path.join(tempfile.mkdtemp(), "t.json")
    prof.export_chrome_trace(path)
    ev = json.load(open(path))["traceEvents"]
    dev_ev = [e for e in ev if e.get("ph") == "X" and e.get("cat") in ("kernel", "gpu_memset", "gpu_memcpy")]
    tot = collections.defaultdict(lambda: [0.0, 0])
    for e in dev_ev:
        k = e["name"][:100]
        tot[k][0] += e["dur"] / reps
        tot[k][1] += 1
    span = (max(e["ts"] + e["dur"] for e in dev_ev) - min(e["ts"] for e in dev_ev)) / reps
    allk = sum(v[0] for v in tot.values())
    print(f"# {a.workload} {a.precision} rssm_only={a.rssm_only}: {len(dev_ev) // reps} device activities per update, "
          f"sum of durations {allk:.1f} us, wall span per update ~{span:.1f} us")
    print(f"{'us/update':>10} {'n':>5} {'share':>6}  name")
    for k, (us, n) in sorted(tot.items(), key=lambda kv: -kv[1][0])[:40]:
        print(f"{us:10.1f} {n // reps:5d} {100 * us / allk:5.1f}%  {k}")


if __name__ == "__main__":
    main()
