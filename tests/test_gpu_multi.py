"""The fused multi-GPU path under pytest: whenever the suite runs on a box with two or more GPUs, two
ranks (one process per GPU, torchrun) shard C1-C4 clusters in contiguous node blocks, run every op through
hs_run_batch — the persistent decide kernel exchanging the packed arg-best key through peer-mapped mailboxes over
NVLink — and rank 0 compares every decision with the single-scheduler CPU oracle (tools/shard_parity.py): node, global
position, score bits, device rows, annotation strings, FailedNodes.  Skipped on a single-GPU box."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _run(world, env_extra=None):
    env = dict(os.environ, **(env_extra or {}))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tools", "shard_parity.py")]
    return subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=1500)


def _gpus():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.skipif(_gpus() < 2, reason="needs two GPUs on one box")
def test_fused_mailbox_path_matches_the_oracle_on_two_ranks():
    r = _run(2)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("shard parity")]
    assert len(lines) == 4 and all(" 0 mismatches" in ln for ln in lines), lines


@pytest.mark.skipif(_gpus() < 2, reason="needs two GPUs on one box")
def test_nccl_per_decision_baseline_matches_the_oracle_on_two_ranks():
    r = _run(2, {"HS_SHARD_NCCL": "1"})
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("shard parity")]
    assert len(lines) == 4 and all(" 0 mismatches" in ln for ln in lines), lines
