"""Sharded runs on real GPUs: two ranks under torch.distributed.run, every solver through the public
drivers against the fp64 oracle -- W, H, objective history and stop iteration on every rank, the AO-ADMM inner-break
pattern, the l2n prox and the Kullback-Leibler branch sharded, and the sharded upload against the full walk, plane
for plane.  The exchange runs over peer memory (csrc/peer.cu).  Skipped on a box with a single GPU."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_ranks_match_the_oracle_on_every_path():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tools", "mgpu_check.py")]
    run = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1500)
    lines = [ln for ln in run.stdout.splitlines() if "sharded" in ln or "MGPU" in ln or "pattern" in ln
             or "stop iteration" in ln]
    print("\n".join(lines))
    assert run.returncode == 0, run.stdout[-4000:]
    assert "MGPU CHECK PASS" in run.stdout and " BAD" not in run.stdout
