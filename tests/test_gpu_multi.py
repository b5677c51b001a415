"""Two or more GPUs on one box: the temporal-segment stitch over NCCL and as ONE peer-memory kernel per rank
(``svr2_segment_stitch_peer`` through ``segments.PeerStitcher``) against the sequential parent loop of the reference
(inference_cli.py:1163-1198), bit-exact.  One process per GPU under torchrun, as bench.py runs.  Skips on a one-GPU box (the
driver's GPU test box has one; `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu` runs it)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on one box")
def test_segment_stitch_nccl_and_peer_memory_match_parent_loop(svr2):
    n = min(torch.cuda.device_count(), 4)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "scripts", "nccl_segments_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    print(r.stdout[-4000:])
    assert r.returncode == 0, r.stderr[-3000:]
    assert "MISMATCH" not in r.stdout
    assert "== sequential parent loop (bit-exact)" in r.stdout
    assert "one peer-memory kernel per rank" in r.stdout and "== NCCL halo exchange + gather (bit-exact)" in r.stdout
