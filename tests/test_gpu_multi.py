"""Multi-GPU parity: the hash-repartitioned join over 2 GPUs (both exchange modes) against
the generator-defined answer.  Skipped on a single-GPU box."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("bits,nprobe,extra", [(16, 300_000, ""), (23, 40_000_000, ""), (23, 20_000_000, "skew")])
def test_two_gpu_partitioned_join_digest(bits, nprobe, extra):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    # bits=23: the local table (2 x 64 MB) is large enough for the radix fan-out, so the p2p mode
    # takes the exchange fused with the partitioning; bits=16 takes the plain scatter
    env = dict(os.environ, IGL_PART_MB="4")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", os.path.join(ROOT, "tests", "multi_gpu_worker.py"), str(bits), str(nprobe)] + ([extra] if extra else []),
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "mode=p2p" in r.stdout and "mode=nccl" in r.stdout
    if bits >= 23:
        assert "mode=p2p fused=True" in r.stdout, r.stdout[-500:]
