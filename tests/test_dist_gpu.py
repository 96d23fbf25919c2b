"""Multi-GPU parity on real GPUs (2 ranks, NCCL): the h2o queries over row shards through aqgpu_groupby_run_dist,
the global result assembled by the product (ShardPlan.gather_ordered) and compared with the CPU oracle by
scripts/dist_check.py.  Skipped on boxes with fewer than 2 GPUs (run it with `gpurun --gpus 2`)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.parametrize("n,k", [(2_000_000, 10), (3_000_001, 100)])
def test_sharded_h2o_two_ranks(n, k):
    if _ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "scripts", "dist_check.py"), str(n), str(k)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0 and "[dist_check] ALL OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_sharded_windowed_trade_two_ranks():
    """rows repartitioned by symbol owner, windowed aggregates local to the owner"""
    if _ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29534", os.path.join(ROOT, "scripts", "dist_check_trade.py"), "400000", "3000"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0 and "[dist_check_trade] ALL OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
