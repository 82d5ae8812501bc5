"""Namespace quota on the GPU path through the C ABI, against the oracle: streams whose pods live in
namespaces with a ResourceQuota on NVIDIA memory and/or cores (and one on MLU memory with its x256 factor).  Every
decision, every FailedNodes reason ("… ResourceQuotaNotFit") and the Used counters must match after every batch.
"""
import ctypes as C
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import oracle_abi as oa  # noqa: E402
import parity  # noqa: E402
from oracle_abi import abi  # noqa: E402
from test_quota import _lib, quota_get  # noqa: E402

pytestmark = pytest.mark.gpu
MEM, CORE = 0, 1


def run_both(cfg, nodes, ops, quotas, chunk):
    """quotas: [(ns, vendor, which, limit)].  Returns (mismatches, number of quota no-fits)."""
    import k8s_vgpu_scheduler_b200.engine as eng
    L = _lib()
    o = oa.Oracle(cfg.get("node_policy"), cfg.get("gpu_policy"))
    e = eng.Engine(cfg.get("node_policy"), cfg.get("gpu_policy"))
    for n in nodes:
        o.node_upsert(n)
        e.node_upsert(n)
    for ns, vendor, which, limit in quotas:
        assert L.oracle_quota_set(o.h, ns.encode(), vendor.encode(), which, limit) == 0
        e.quota_set(ns, vendor, which, limit)
    bad, quota_nofit = [], 0
    for s in range(0, len(ops), chunk):
        part = ops[s:s + chunk]
        ref = []
        for kind, arg in part:
            if kind == "filter":
                ref.append(o.filter(arg))
            else:
                rc = o.pod_delete(arg)
                ref.append({"status": rc, "rc": rc})
        e.run_batch(eng.PackedOps(part))
        got = [e.batch_get(i) for i in range(len(part))]
        bad += parity.compare(part, ref, got)
        quota_nofit += sum(1 for r in ref if r.get("status") == abi.HS_NOFIT and any("ResourceQuotaNotFit" in v for v in r["failed"].values()))
        for ns, vendor, which, _ in quotas:
            assert e.quota_get(ns, vendor, which) == quota_get(L, o, ns, vendor, which), (s, ns, vendor, which)
    snap_o = {n["name"]: o.snapshot(n["name"]) for n in nodes}
    snap_e = {n["name"]: e.snapshot(n["name"]) for n in nodes}
    assert snap_o == snap_e
    e.close()
    o.close()
    return bad, quota_nofit


def with_namespaces(ops, names):
    out = []
    for i, (kind, arg) in enumerate(ops):
        out.append((kind, dict(arg, ns=names[i % len(names)]) if kind == "filter" else arg))
    return out


def test_quota_on_mixed_feature_stream():
    """multi-container pods, init containers, base-type padding (nodes whose device Type is exactly "NVIDIA")"""
    import k8s_vgpu_scheduler_b200.gen as gen
    cfg, nodes, ops = gen.mixed_features(n_nodes=24, n_pods=600)
    ops = with_namespaces(ops, ["team-a", "team-b", "free"])
    quotas = [("team-a", "NVIDIA", MEM, 600000), ("team-b", "NVIDIA", CORE, 900), ("team-b", "NVIDIA", MEM, 2000000)]
    bad, qn = run_both(cfg, nodes, ops, quotas, chunk=97)
    assert bad == []
    assert qn > 10


def test_quota_on_heterogeneous_cluster_with_mlu_factor():
    import k8s_vgpu_scheduler_b200.gen as gen
    cfg, nodes, ops = gen.config4(n_nodes=200, n_pods=1200)
    ops = with_namespaces(ops, ["q1", "q2"])
    quotas = [("q1", "NVIDIA", MEM, 900000), ("q1", "MLU", MEM, 400), ("q2", "MLU", CORE, 300)]   # MLU memory limit counts x256
    bad, qn = run_both(cfg, nodes, ops, quotas, chunk=400)
    assert bad == []
    assert qn > 10


def test_quota_set_later_sees_earlier_usage_and_zero_blocks():
    import k8s_vgpu_scheduler_b200.engine as eng
    import k8s_vgpu_scheduler_b200.gen as gen
    cfg, nodes, ops = gen.config2(n_nodes=50, n_pods=300)
    fil = [(k, dict(a, ns="late")) for k, a in ops if k == "filter"]
    e = eng.Engine(cfg.get("node_policy"), cfg.get("gpu_policy"))
    for n in nodes:
        e.node_upsert(n)
    e.run_batch(eng.PackedOps(fil[:100]))
    used = e.quota_get("late", "NVIDIA", MEM)
    assert used is not None and used[1] > 0 and used[2] is False                # tracked although no limit is configured
    e.quota_set("late", "NVIDIA", MEM, used[1])                                  # the limit is already reached
    e.run_batch(eng.PackedOps(fil[100:110]))
    res = [e.batch_get(i) for i in range(10)]
    assert all(r["status"] == abi.HS_NOFIT for r in res)
    assert all("ResourceQuotaNotFit" in v for v in res[0]["failed"].values())
    e.quota_del("late", "NVIDIA", MEM)
    e.run_batch(eng.PackedOps(fil[110:120]))
    assert any(e.batch_get(i)["status"] == abi.HS_FIT for i in range(10))
    e.quota_set("late", "NVIDIA", MEM, 0)                                        # explicit zero: hard block (quota.go:30-36)
    e.run_batch(eng.PackedOps(fil[120:125]))
    assert all(e.batch_get(i)["status"] == abi.HS_NOFIT for i in range(5))
    e.close()
