"""hs_run_batch_async: a long batch runs as a pipeline (compile | device | usage-cache update) while the caller already
reads the results of the chunks that are done.  Whatever the interleaving, every decision, the final usage cache and
the pod cache must equal the oracle's — and the synchronous call's."""
import pytest

import parity
from oracle_abi import abi

pytestmark = pytest.mark.gpu


def _run_async(cfg, nodes, ops, read):
    import k8s_vgpu_scheduler_b200.engine as eng
    e = eng.Engine(cfg.get("node_policy"), cfg.get("gpu_policy"))
    for n in nodes:
        e.node_upsert(n)
    packed = eng.PackedOps(ops)
    e.run_batch_async(packed)
    if read == "in_order":
        got = [e.batch_get(i) for i in range(len(ops))]
    elif read == "backwards":                       # the last op first: waits for the whole batch
        got = [e.batch_get(i) for i in reversed(range(len(ops)))][::-1]
    else:                                           # a summary walks the pipeline, then the results are read again
        e.batch_summary()
        got = [e.batch_get(i) for i in range(len(ops))]
    snap = {n["name"]: e.snapshot(n["name"]) for n in nodes}     # any other call lets the pipeline finish first
    pods = {a["uid"]: e.pod_get(a["uid"]) for k, a in ops if k == "filter"}
    e.close()
    return got, snap, pods


@pytest.mark.parametrize("read", ["in_order", "backwards", "summary"])
def test_async_batch_equals_oracle(read, monkeypatch):
    import k8s_vgpu_scheduler_b200.gen as gen
    monkeypatch.setenv("HS_PIPE_CHUNK", "64")
    monkeypatch.setenv("HS_FAIL_CAP", "4")
    for cfg, nodes, ops in (gen.config3(n_nodes=150, n_pods=900), gen.mixed_features(n_nodes=24, n_pods=700),
                            gen.config1(n_pods=400, departures=0.3)):
        ref, rsnap = parity.run_oracle(cfg, nodes, ops)
        got, gsnap, _ = _run_async(cfg, nodes, ops, read)
        assert parity.compare(ops, ref, got, rsnap, gsnap) == []


def test_async_and_sync_leave_the_same_pod_cache(monkeypatch):
    import k8s_vgpu_scheduler_b200.engine as eng
    import k8s_vgpu_scheduler_b200.gen as gen
    monkeypatch.setenv("HS_PIPE_CHUNK", "100")
    cfg, nodes, ops = gen.config2(n_nodes=120, n_pods=1500)
    _, snap_a, pods_a = _run_async(cfg, nodes, ops, "summary")
    e = eng.Engine(cfg.get("node_policy"), cfg.get("gpu_policy"))
    for n in nodes:
        e.node_upsert(n)
    e.run_batch(eng.PackedOps(ops))
    snap_s = {n["name"]: e.snapshot(n["name"]) for n in nodes}
    pods_s = {a["uid"]: e.pod_get(a["uid"]) for k, a in ops if k == "filter"}
    e.close()
    assert snap_a == snap_s and pods_a == pods_s
    assert sum(1 for v in pods_s.values() if v) > 100


def test_windows_decide_like_the_one_op_path(monkeypatch):
    """HS_WINDOW=1: runs of common-shape ops are decided 16 at a time from one sweep (hs_window.cuh) — same decisions."""
    import k8s_vgpu_scheduler_b200.gen as gen
    monkeypatch.setenv("HS_WINDOW", "1")
    for cfg, nodes, ops in (gen.config3(n_nodes=300, n_pods=1500), gen.config2(n_nodes=200, n_pods=1500),
                            gen.config1(n_pods=300, departures=0.3), gen.config4(n_nodes=300, n_pods=1500),
                            gen.mixed_features(n_nodes=24, n_pods=600)):
        ref, rsnap = parity.run_oracle(cfg, nodes, ops)
        got, gsnap = parity.run_engine(cfg, nodes, ops)
        assert parity.compare(ops, ref, got, rsnap, gsnap) == []
