"""drm_allreduce (mean over NVLink peer memory) against NCCL, 2 and 4 ranks on one box.  Needs that many GPUs:
skipped on the single-GPU test box (bench.py --gpus N exercises it at N = 2 .. 8; PeerBucket.create runs the same
comparison as a self-test on every start-up and falls back to NCCL on all ranks if it fails)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.environ["REPO"])
rank = int(os.environ["RANK"]); torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
from dreamer_pytorch_b200.parallel import PeerBucket
dev = torch.device("cuda", rank)
groups = [[torch.empty(200, 230, device=dev), torch.empty(200, device=dev), torch.empty(12, 200, device=dev), torch.empty(12, device=dev)],
          [torch.empty(200, 230, device=dev), torch.empty(200, device=dev), torch.empty(1, 200, device=dev), torch.empty(1, device=dev)]]
b = PeerBucket.create(groups, algo=2 if dist.get_world_size() >= 4 else 1)  # 4 ranks: the two-shot kernel
assert b is not None, "peer mapping unavailable"
for it in range(5):
    g = torch.Generator(device=dev).manual_seed(it * 10 + rank)
    b.flat.copy_(torch.randn(b.flat.numel(), generator=g, device=dev))
    want = b.flat.clone(); dist.all_reduce(want, op=dist.ReduceOp.AVG)
    b.allreduce(); torch.cuda.synchronize()
    assert float((b.flat - want).abs().max()) <= 1e-6 * float(want.abs().max()), it
    other = b.flat.clone(); dist.broadcast(other, 0)
    assert torch.equal(other, b.flat)  # bit-identical on every rank
assert b.provide(groups[0]).data_ptr() == b.flat.data_ptr() and b.provide([torch.empty(3, device=dev)]) is None
g = torch.cuda.CUDAGraph()
s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    with torch.cuda.graph(g, stream=s):
        b.allreduce()
torch.cuda.current_stream().wait_stream(s)
for it in range(3):
    b.flat.fill_(float(rank + it)); torch.cuda.synchronize(); dist.barrier()
    g.replay(); torch.cuda.synchronize()
    assert torch.allclose(b.flat, torch.full_like(b.flat, it + (dist.get_world_size() - 1) / 2))
print("PEER_OK", rank, flush=True)
dist.barrier(); os._exit(0)
"""


@pytest.mark.parametrize("ranks", [2, 4])  # 2: one-shot kernel; 4: reduce-scatter + all-gather form
def test_peer_allreduce_matches_nccl(tmp_path, ranks):
    if torch.cuda.device_count() < ranks:
        pytest.skip(f"needs {ranks} GPUs")
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    env = dict(os.environ, REPO=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={ranks}",
                          "--master-addr", "127.0.0.1", "--master-port", str(29531 + ranks), str(script)],
                         env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and out.stdout.count("PEER_OK") == ranks, out.stdout[-2000:] + out.stderr[-3000:]
