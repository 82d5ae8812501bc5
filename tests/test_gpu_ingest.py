"""Ingest entry points that change the HBM inventory, through the C ABI on the GPU:

  hs_node_register   the node's devices arrive as register-annotation STRINGS (JSON for NVIDIA/Ascend, the legacy
                     CSV for DCU, hami.io/node-nvidia-score for pair scores) — decisions must equal the oracle's,
                     which is fed the same devices as structs;
  hs_pod_ingest      restart replay: a fresh engine rebuilds its usage from the pods' "*-devices-allocated"
                     annotations (the text Filter itself produced) — usage and every later decision must match;
  hs_pod_shrink      init containers done → usage drops to the app containers' share (expected usage computed
                     with the Python ingest oracle).
"""
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import parity  # noqa: E402
from oracle_abi import abi  # noqa: E402

pytestmark = pytest.mark.gpu


def registrations(node):
    """The node's devices as the annotations the device plugins would write (MarshalNodeDevices' omitempty JSON,
    EncodeNodeDevices' CSV).  MLU devices do not come from an annotation in the reference (cambricon/device.go:209-232)."""
    import ingest_oracle as O
    by_vendor = {}
    for d in node["devices"]:
        by_vendor.setdefault(d["vendor"], []).append(d)
    regs = []
    for vendor, devs in by_vendor.items():
        if vendor == "DCU":
            text = O.encode_node_devices([{"ID": d["id"], "Count": d["count"], "Devmem": d["devmem"], "Devcore": d["devcore"],
                                           "Type": d["type"], "Numa": d["numa"], "Health": d["health"], "Index": d["index"],
                                           "Mode": d["mode"]} for d in devs])
            regs.append((vendor, abi.HS_REG_CSV, text, None))
            continue
        items = []
        for d in devs:
            o = {}
            for k in ("id", "index", "count", "devmem", "devcore", "type", "numa", "mode", "health"):
                if d[k]:                                  # omitempty
                    o[k] = d[k]
            if d.get("network_id", -1) >= 0:
                o["custominfo"] = {"NetworkID": d["network_id"]}
            items.append(o)
        pair = None
        if vendor == "NVIDIA" and node.get("pair_scores") is not None:
            n = len(node["devices"])
            ids = [x["id"] for x in node["devices"]]
            ps = node["pair_scores"]
            at = (lambda i, j: ps[i][j]) if isinstance(ps[0], (list, tuple)) else (lambda i, j: ps[i * n + j])
            pair = json.dumps([{"uuid": ids[i], "score": {ids[j]: int(at(i, j)) for j in range(n) if j != i}} for i in range(n)])
        regs.append((vendor, abi.HS_REG_JSON, json.dumps(items), pair))
    return regs


def engine_with_registered_nodes(cfg, nodes):
    import k8s_vgpu_scheduler_b200.engine as eng
    e = eng.Engine(cfg.get("node_policy"), cfg.get("gpu_policy"))
    for n in nodes:
        if any(d["vendor"] == "MLU" for d in n["devices"]):
            e.node_upsert(n)
            continue
        got = e.node_register(n["name"], registrations(n), n.get("device_cordon"), n.get("hami_vnpu_core", -1))
        assert got == len(n["devices"]), e.error()
    return e, eng


def run_stream(e, eng, ops):
    e.run_batch(eng.PackedOps(ops))
    return [e.batch_get(i) for i in range(len(ops))]


@pytest.mark.parametrize("name,kw", [("config4", dict(n_nodes=300, n_pods=1500)), ("config3", dict(n_nodes=200, n_pods=1200)),
                                     ("mixed_features", dict(n_nodes=24, n_pods=600))])
def test_registered_from_annotations_decides_like_the_oracle(name, kw):
    import k8s_vgpu_scheduler_b200.gen as gen
    cfg, nodes, ops = getattr(gen, name)(**kw)
    ref, ref_snap = parity.run_oracle(cfg, nodes, ops)
    e, eng = engine_with_registered_nodes(cfg, nodes)
    got = run_stream(e, eng, ops)
    snap = {n["name"]: e.snapshot(n["name"]) for n in nodes}
    e.close()
    assert parity.compare(ops, ref, got, ref_snap, snap) == []


def test_register_skips_bad_vendors_and_removes_empty_nodes():
    import k8s_vgpu_scheduler_b200.engine as eng
    e = eng.Engine("binpack", "spread")
    good = '[{"id":"GPU-a","count":10,"devmem":8192,"devcore":100,"type":"NVIDIA-V100","health":true}]'
    assert e.node_register("n1", [("NVIDIA", abi.HS_REG_JSON, good, None), ("DCU", abi.HS_REG_CSV, "garbage", None)]) == 1
    assert e.node_count() == 1 and "DCU" in e.error()
    assert e.node_register("n1", [("NVIDIA", abi.HS_REG_JSON, "[]", None)]) == 0       # vendor reports zero devices
    assert e.node_count() == 0
    assert e.node_register("n2", [("NVIDIA", abi.HS_REG_JSON, "not-valid-json", None)]) == 0
    assert e.node_count() == 0
    # vastai dies (custominfo.DeviceStrategy = "die"): the card a die sits on (AIC) is registered with it; a node that mixes
    # dies and whole cards is skipped loudly
    die = '[{"id":"va-0","count":1,"type":"Vastai","health":true,"custominfo":{"DeviceStrategy":"die","AIC":"1"}},' \
          '{"id":"va-1","count":1,"type":"Vastai","health":true,"custominfo":{"DeviceStrategy":"die","AIC":"0"}}]'
    mixed = '[{"id":"va-0","count":1,"type":"Vastai","health":true,"custominfo":{"DeviceStrategy":"die","AIC":"1"}},' \
            '{"id":"va-1","count":1,"type":"Vastai","health":true}]'
    card = '[{"id":"va-0","count":1,"type":"Vastai","health":true,"custominfo":{"DeviceStrategy":"card"}}]'
    assert e.node_register("n3", [("Vastai", abi.HS_REG_JSON, die, None), ("NVIDIA", abi.HS_REG_JSON, good, None)]) == 3
    assert e.node_register("n5", [("Vastai", abi.HS_REG_JSON, mixed, None), ("NVIDIA", abi.HS_REG_JSON, good.replace("GPU-a", "GPU-b"), None)]) == 1
    assert "die" in e.error() and e.node_count() == 2
    assert e.node_register("n4", [("Vastai", abi.HS_REG_JSON, card, None)]) == 1
    e.close()


def test_restart_replay_from_pod_annotations():
    import k8s_vgpu_scheduler_b200.gen as gen
    cfg, nodes, ops = gen.mixed_features(n_nodes=24, n_pods=600)
    cut = 400
    ref, _ = parity.run_oracle(cfg, nodes, ops, snapshot=False)
    # what survives at the cut: pods that fit and were not deleted since
    alive = {}
    for (kind, arg), r in zip(ops[:cut], ref[:cut]):
        if kind == "filter":
            alive.pop(arg["uid"], None)
            if r["status"] == abi.HS_FIT:
                alive[arg["uid"]] = (arg, r)
        else:
            alive.pop(arg, None)
    assert len(alive) > 50
    e, eng = engine_with_registered_nodes(cfg, nodes)
    for uid, (pod, r) in alive.items():
        assert e.pod_ingest(uid, r["annotations"]["hami.io/vgpu-node"], r["annotations"], pod.get("n_init", 0)) == 0, e.error()
        assert e.pod_ingest(uid, r["annotations"]["hami.io/vgpu-node"], r["annotations"], pod.get("n_init", 0)) == 0   # informer resync: idempotent
    # the rebuilt usage equals the usage of an oracle that lived through the first `cut` ops
    _, snap_cut = parity.run_oracle(cfg, nodes, ops[:cut])
    assert {n["name"]: e.snapshot(n["name"]) for n in nodes} == snap_cut
    # and the stream continues bit for bit
    got = run_stream(e, eng, ops[cut:])
    e.close()
    assert parity.compare(ops[cut:], ref[cut:], got) == []


def test_shrink_after_init_containers():
    import ingest_oracle as O
    import k8s_vgpu_scheduler_b200.gen as gen
    cfg, nodes, ops = gen.mixed_features(n_nodes=24, n_pods=600)
    ref, _ = parity.run_oracle(cfg, nodes, ops, snapshot=False)
    alive = {}
    for (kind, arg), r in zip(ops, ref):
        if kind == "filter":
            alive.pop(arg["uid"], None)
            if r["status"] == abi.HS_FIT:
                alive[arg["uid"]] = (arg, r)
        else:
            alive.pop(arg, None)
    with_init = [u for u, (p, r) in alive.items() if p.get("n_init", 0) > 0]
    assert len(with_init) >= 5
    e, eng = engine_with_registered_nodes(cfg, nodes)
    checklist = {"NVIDIA": "hami.io/vgpu-devices-allocated"}
    want = {n["name"]: {d["id"]: [0, 0, 0] for d in n["devices"]} for n in nodes}
    for uid, (pod, r) in alive.items():
        node = r["annotations"]["hami.io/vgpu-node"]
        n_init = pod.get("n_init", 0)
        assert e.pod_ingest(uid, node, r["annotations"], n_init) == 0
        raw = O.decode_pod_devices(checklist, r["annotations"])
        if uid in with_init:
            assert e.pod_shrink(uid, r["annotations"], n_init) == 0
            assert e.pod_shrink(uid, r["annotations"], n_init) == 0                     # only once
            assert e.pod_ingest(uid, node, r["annotations"], n_init) == 0               # a resync must not re-inflate it (pods.go:79-86)
            eff = O.app_containers_only_device_usage(n_init, raw)
        else:
            eff = O.collapse_init_container_usage(n_init, raw)
        for devs in eff.values():
            for d in devs[0]:
                if d["UUID"] in want[node]:
                    u = want[node][d["UUID"]]
                    u[0] += max(d["Slots"], 1); u[1] += d["Usedmem"]; u[2] += d["Usedcores"]
    got = {n["name"]: e.snapshot(n["name"]) for n in nodes}
    e.close()
    assert got == {n["name"]: [tuple(want[n["name"]][d["id"]]) for d in n["devices"]] for n in nodes}
