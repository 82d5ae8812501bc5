"""A scheduler's life, not just a pod stream: between batches of Filter/Delete ops the cluster changes under the engine —
nodes leave and come back with different devices, pods arrive through the informer path (usage_apply), a namespace
gets a quota and loses it again, single hs_filter calls and dry runs are mixed in — and after every step the engine
must agree with the oracle on every decision and on the usage of every node.  Everything goes through the C ABI.
"""
import ctypes as C
import os
import random
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import oracle_abi as oa  # noqa: E402
import parity  # noqa: E402
from oracle_abi import abi  # noqa: E402
from test_quota import _lib, quota_get  # noqa: E402

pytestmark = pytest.mark.gpu


def _decision(r):
    if r["status"] == abi.HS_FIT:
        return (0, r["node"], parity.f32bits(r["score"]), r["devices"], r["annotations"])
    return (r["status"], r.get("failed"))


def test_small_reason_buffer_drains_and_resumes(monkeypatch):
    """With room for only 5 no-fit decisions per launch the engine must stop, hand the reasons to the host and resume at
    the right op, many times per batch, without changing a single decision."""
    import k8s_vgpu_scheduler_b200.gen as gen
    monkeypatch.setenv("HS_FAIL_CAP", "5")
    for cfg, nodes, ops in (gen.config1(n_pods=400, departures=0.2), gen.mixed_features(n_nodes=16, n_pods=500)):
        ref, rsnap = parity.run_oracle(cfg, nodes, ops)
        got, gsnap = parity.run_engine(cfg, nodes, ops, chunk=170)
        assert sum(1 for r in ref if r.get("status") == abi.HS_NOFIT) > 40
        assert parity.compare(ops, ref, got, rsnap, gsnap) == []


def test_long_batches_compiled_in_chunks_under_the_device(monkeypatch):
    """hs_run_batch compiles a long batch in chunks while the device runs the previous chunk.  With a chunk of 50 ops
    (and a reason buffer of 3) a 700-op batch crosses many chunk borders: releases that refer to decisions of earlier
    chunks, re-filtered pods, drained reason buffers in the middle of a chunk — no decision may change."""
    import k8s_vgpu_scheduler_b200.gen as gen
    monkeypatch.setenv("HS_PIPE_CHUNK", "50")
    monkeypatch.setenv("HS_FAIL_CAP", "3")
    for cfg, nodes, ops in (gen.mixed_features(n_nodes=24, n_pods=700), gen.config3(n_nodes=150, n_pods=900),
                            gen.config1(n_pods=500, departures=0.3)):
        ref, rsnap = parity.run_oracle(cfg, nodes, ops)
        got, gsnap = parity.run_engine(cfg, nodes, ops)          # one batch
        assert parity.compare(ops, ref, got, rsnap, gsnap) == []
    monkeypatch.delenv("HS_FAIL_CAP")
    cfg, nodes, ops = gen.config4(n_nodes=300, n_pods=2500)
    ref, rsnap = parity.run_oracle(cfg, nodes, ops)
    got, gsnap = parity.run_engine(cfg, nodes, ops)
    assert parity.compare(ops, ref, got, rsnap, gsnap) == []


def test_node_reregisters_with_reordered_and_replaced_devices():
    """A node re-registers while pods are bound to it: same UUIDs in another order keep their usage and the pods release
    against the right devices; a replaced UUID starts empty and the share of the pods that sat on the old one is gone."""
    import k8s_vgpu_scheduler_b200.engine as eng
    import k8s_vgpu_scheduler_b200.gen as gen
    cfg, nodes, ops = gen.config2(n_nodes=6, n_pods=400)
    o = oa.Oracle(cfg["node_policy"], cfg["gpu_policy"])
    e = eng.Engine(cfg["node_policy"], cfg["gpu_policy"])
    for n in nodes:
        o.node_upsert(n); e.node_upsert(n)

    def run(part):
        ref = []
        for kind, arg in part:
            ref.append(o.filter(arg) if kind == "filter" else {"status": o.pod_delete(arg)})
        e.run_batch(eng.PackedOps(part))
        got = [e.batch_get(i) for i in range(len(part))]
        for r in ref:
            r.setdefault("rc", r["status"])
        assert parity.compare(part, ref, got) == []
        for n in nodes:
            assert e.snapshot(n["name"]) == o.snapshot(n["name"]), n["name"]

    run(ops[:150])
    n0 = nodes[0]
    n0["devices"] = [dict(d, index=i) for i, d in enumerate(reversed(n0["devices"]))]           # same UUIDs, other order
    o.node_upsert(n0); e.node_upsert(n0)
    assert e.snapshot(n0["name"]) == o.snapshot(n0["name"])
    run(ops[150:300])
    n1 = nodes[1]
    n1["devices"] = [dict(d, id=d["id"] + "-new") if i == 0 else d for i, d in enumerate(n1["devices"])]   # one card swapped
    o.node_upsert(n1); e.node_upsert(n1)
    assert e.snapshot(n1["name"]) == o.snapshot(n1["name"])
    run(ops[300:350])
    # the swapped card's original UUID is registered again: the pods that sat on it still hold it (the reference joins
    # pods and devices by UUID on every rebuild), so its usage is back — no over-commit
    old0 = n1["devices"][0]["id"][:-len("-new")]
    n1["devices"] = [dict(d, id=old0) if i == 0 else d for i, d in enumerate(n1["devices"])]
    o.node_upsert(n1); e.node_upsert(n1)
    assert e.snapshot(n1["name"]) == o.snapshot(n1["name"])
    run(ops[350:])
    e.close()
    o.close()


def test_cluster_changes_between_batches():
    import k8s_vgpu_scheduler_b200.engine as eng
    import k8s_vgpu_scheduler_b200.gen as gen
    L = _lib()
    L.oracle_node_remove.argtypes = [C.c_void_p, C.c_char_p]
    rng = random.Random(7)
    cfg, nodes, ops = gen.mixed_features(n_nodes=40, n_pods=1400)
    _, spare, _ = gen.config2(n_nodes=12, n_pods=1)
    for i, n in enumerate(spare):
        n["name"] = "spare-%02d" % i
    o = oa.Oracle(cfg["node_policy"], cfg["gpu_policy"])
    e = eng.Engine(cfg["node_policy"], cfg["gpu_policy"])
    live_nodes = {}
    for n in nodes:
        o.node_upsert(n); e.node_upsert(n)
        live_nodes[n["name"]] = n
    removed = []
    ext = 0
    ops = [(k, dict(a, ns=("q" if i % 3 == 0 else "default")) if k == "filter" else a) for i, (k, a) in enumerate(ops)]
    step = 100
    for s in range(0, len(ops), step):
        part = ops[s:s + step]
        # ---- a batch through hs_run_batch, or the same ops one call at a time
        ref = []
        for kind, arg in part:
            if kind == "filter":
                ref.append(o.filter(arg))
            else:
                rc = o.pod_delete(arg)
                ref.append({"status": rc, "rc": rc})
        if (s // step) % 4 == 3:
            got = []
            for kind, arg in part:
                if kind == "filter":
                    got.append(e.filter(arg))
                else:
                    rc = e.pod_delete(arg)
                    got.append({"status": rc, "rc": rc})
        else:
            e.run_batch(eng.PackedOps(part))
            got = [e.batch_get(i) for i in range(len(part))]
        assert parity.compare(part, ref, got) == [], "step %d" % s
        # ---- a dry run changes nothing on either side
        probe = dict(next(a for k, a in part if k == "filter"), uid="probe-%d" % s)
        want = o.filter(probe)
        o.pod_delete(probe["uid"])
        assert _decision(e.filter(probe, dry=True)) == _decision(want), "dry %d" % s
        # ---- the cluster changes
        kind = (s // step) % 5
        if kind == 0 and len(live_nodes) > 30:             # a node leaves
            name = rng.choice(sorted(live_nodes))
            assert L.oracle_node_remove(o.h, name.encode()) == 0 and e.node_remove(name) == 0
            removed.append(live_nodes.pop(name))
        elif kind == 1 and removed:                        # … and comes back; its pods count again (join by UUID)
            n = removed.pop()
            o.node_upsert(n); e.node_upsert(n)
            live_nodes[n["name"]] = n
        elif kind == 2:                                    # a pod someone else bound shows up through the informer
            n = live_nodes[rng.choice(sorted(live_nodes))]
            d = rng.choice(n["devices"])
            uid = "ext-%d" % ext
            ext += 1
            a = o.usage_apply(uid, n["name"], d["id"], 1, 1024, 10)
            b = e.usage_apply(uid, n["name"], d["id"], 1, 1024, 10)
            assert a == b == 0
        elif kind == 3:                                    # the namespace gets a quota just above what it uses now
            used = e.quota_get("q", "NVIDIA", 0)
            limit = (used[1] if used else 0) + 200000
            e.quota_set("q", "NVIDIA", 0, limit)
            assert L.oracle_quota_set(o.h, b"q", b"NVIDIA", 0, limit) == 0
        else:                                              # … and loses it again
            e.quota_del("q", "NVIDIA", 0)
            L.oracle_quota_del(o.h, b"q", b"NVIDIA", 0)
        assert e.quota_get("q", "NVIDIA", 0) == quota_get(L, o, "q", "NVIDIA", 0), "quota %d" % s
        for name in live_nodes:
            assert e.snapshot(name) == o.snapshot(name), (s, name)
    e.close()
    o.close()
