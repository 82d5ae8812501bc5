"""The scan probe tool (hs_scan_probe, the HBM-roofline evidence of bench.py's scan_hbm leg) runs and reports on both of its
inventories.  Its choice against the oracle is test_gpu_parity.py::test_sweep_probe_matches_oracle_choice; this one keeps
the tool itself alive when include/hami_sched.h grows."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize("flags", [[], ["--realistic"]], ids=["idle", "realistic"])
def test_scan_probe_reports(flags):
    exe = os.path.join(ROOT, "k8s-vgpu-scheduler_b200", "hs_scan_probe")
    proc = subprocess.run([exe] + flags + ["65536"], capture_output=True, text=True, timeout=300)
    assert proc.returncode == 0, proc.stderr[-400:]
    rows = [json.loads(x) for x in proc.stdout.strip().splitlines()]
    assert len(rows) >= 3
    for r in rows:
        assert r["fits"] is True and r["ms"] > 0 and r["algorithmic_gbps"] > 0, r
        assert abs(r["inventory_mb"] - 65536 * 256 / 1e6) < 0.01
