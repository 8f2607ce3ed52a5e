"""Device timeline of ONE replayed imagination step (kineto/CUPTI through torch.profiler): every kernel / memset /
memcpy with its stream, start offset and duration, so that the critical path of the captured graph can be read.

    python tools/step_trace.py [--n 2500] [--h 15] [--precision bf16] [--eager]  > gpurun_out/step_trace.txt
"""
import argparse
import json
import os
import sys
import tempfile

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=2500)
    ap.add_argument("--h", type=int, default=15)
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--eager", action="store_true")
    a = ap.parse_args()
    from bench import WORKLOADS, build_model  # the walker model of BASELINE.json configs[1]
    from dreamer_pytorch_b200.precision import set_precision
    from dreamer_pytorch_b200.graph import ImaginationGraph

    dev = torch.device("cuda:0")
    set_precision(a.precision)
    model = build_model(dict(WORKLOADS["walker"]), dev)
    g = ImaginationGraph(model, a.n, a.h)
    stoch = torch.randn(a.n, model.stochastic_size, device=dev)
    deter = torch.randn(a.n, model.deterministic_size, device=dev)
    run = (lambda: g._eager()) if a.eager else (lambda: g.step(stoch, deter))
    for _ in range(5):
        run()
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(3):
            run()
            torch.cuda.synchronize()
    path = os.path.join(tempfile.mkdtemp(), "t.json")
    prof.export_chrome_trace(path)
    ev = json.load(open(path))["traceEvents"]
    dev_ev = [e for e in ev if e.get("ph") == "X" and e.get("cat") in ("kernel", "gpu_memset", "gpu_memcpy")]
    dev_ev.sort(key=lambda e: e["ts"])
    # the last step starts at its noise kernel (one generator launch opens every step)
    firsts = [i for i, e in enumerate(dev_ev) if "distribution_elementwise" in e["name"]]
    lo = firsts[-1] if firsts else 0
    last = dev_ev[lo:]
    t0 = last[0]["ts"]
    print(f"# {len(last)} device activities in the last step; span {last[-1]['ts'] + last[-1]['dur'] - t0:.1f} us")
    print(f"{'start':>9} {'dur':>8} {'end':>9} strm  name")
    for e in last:
        s = e["ts"] - t0
        print(f"{s:9.1f} {e['dur']:8.1f} {s + e['dur']:9.1f} {e['args'].get('stream', '?'):>4}  {e['name'][:90]}")


if __name__ == "__main__":
    main()
