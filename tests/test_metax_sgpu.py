"""metax sGPU ("Metax-SGPU"): the oracle against the reference's own tables of pkg/device/metax/sdevice_test.go
(extracted into tests/golden/metax_sgpu_*.json by tools/extract_golden.py), then GPU parity."""
import ctypes as C
import json
import random

import pytest

import oracle_abi as oa
from oracle_abi import abi
from test_oracle_golden import go_device, ids, load, run_fit

SGPU = "Metax-SGPU"
CLASS, QOS, TOPO = "metax-tech.com/sgpu-app-class", "metax-tech.com/sgpu-qos-policy", "metax-tech.com/sgpu-topology-aware"


@pytest.fixture(scope="module")
def orc():
    o = oa.Oracle()
    yield o
    o.close()


@pytest.mark.parametrize("case", load("metax_sgpu_fit"), ids=ids(load("metax_sgpu_fit")))
def test_fit_table(orc, case):
    # metax/sdevice_test.go:739-1273 (TestMetaxSDevices_Fit)
    annos = case["annos"] if isinstance(case["annos"], dict) else {}
    fit, got, reason = run_fit(orc, SGPU, case["devices"], case["request"], annos)
    assert fit == case["wantFit"]
    if fit:
        assert len(got) == case["wantLen"] and [g[0] for g in got] == case["wantDevIDs"]
    assert reason == case["wantReason"]


def _dev(d, **kw):
    base = {"Type": SGPU, "Health": True, "Count": 100, "Totalmem": 1000, "Totalcore": 100}
    base.update(go_device(d))
    base.update(kw)
    return base


@pytest.mark.parametrize("case", load("metax_sgpu_check_qos"), ids=ids(load("metax_sgpu_check_qos")))
def test_check_device_qos_table(orc, case):
    # sdevice_test.go:592-678 (TestCheckDeviceQos) through Fit: the one gate that can fail here is the QoS one
    dev = _dev(case["usage"], Usedcores=0)
    fit, _, reason = run_fit(orc, SGPU, [dev], {"Nums": 1, "Type": SGPU, "Memreq": 1, "Coresreq": case["request"]["Coresreq"]},
                             {QOS: case["reqQos"]})
    if case["request"]["Coresreq"] == 100 and dev.get("Used", 0) > 0:
        assert reason == "1/1 ExclusiveDeviceAllocateConflict"          # passed the QoS gate, stopped at the exclusivity one
    else:
        assert fit == case["expected"] and (fit or reason == "1/1 CardQosPolicyMismatch")


@pytest.mark.parametrize("case", load("metax_sgpu_check_app_class"), ids=ids(load("metax_sgpu_check_app_class")))
def test_check_app_class_table(orc, case):
    # sdevice_test.go:2103-2272 (TestCheckAppClass) through Fit
    dev = _dev(case["usage"])
    annos = {CLASS: case["reqClass"]} if case["reqClass"] else {}
    fit, _, reason = run_fit(orc, SGPU, [dev], {"Nums": 1, "Type": SGPU, "Memreq": 1, "Coresreq": 0}, annos)
    assert fit == case["expected"] and (fit or reason == "1/1 CardAppClassMismatch")


@pytest.mark.parametrize("case", load("metax_sgpu_prioritize_exclusive"), ids=ids(load("metax_sgpu_prioritize_exclusive")))
def test_prioritize_exclusive_table(orc, case):
    # sdevice_test.go:1326-1607: Fit walks the list backwards, so the candidate list of the table is handed over reversed
    devs = [_dev(d, Index=i) for i, d in enumerate(case["candidateDevices"])][::-1]
    fit, got, _ = run_fit(orc, SGPU, devs, {"Nums": case["require"], "Type": SGPU, "Memreq": 1, "Coresreq": 100}, {})
    assert fit and [g[0] for g in got] == [d["ID"] for d in case["expectedDevices"]]


@pytest.mark.parametrize("case", load("metax_sgpu_prioritize_online"), ids=ids(load("metax_sgpu_prioritize_online")))
def test_prioritize_online_table(orc, case):
    # sdevice_test.go:2274-2773 (require = 1)
    devs = [_dev(d, Index=i, Used=len(go_device(d).get("PodInfos") or [])) for i, d in enumerate(case["candidateDevices"])][::-1]
    fit, got, _ = run_fit(orc, SGPU, devs, {"Nums": 1, "Type": SGPU, "Memreq": 1, "Coresreq": 0}, {CLASS: "online"})
    assert fit and [g[0] for g in got] == [d["ID"] for d in case["expectedDevices"]]
    assert got[0][3] == 0          # an online pod's pick records no cores (sdevice.go:459-462)


def _score(orc, pod_devices, previous, annos):
    picks, zones, has_zone = [], [], False
    for c, ctr in enumerate(pod_devices):
        for d in ctr:
            ci = d.get("CustomInfo") if isinstance(d.get("CustomInfo"), dict) else None
            has_zone = has_zone or ci is not None
            zones.append(int((ci or {}).get("LinkZone", 0)))
            picks.append(abi.hs_alloc_entry(c, b"Metax-SGPU", abi._b(d.get("UUID", "")), 0, 0, int(d.get("Usedcores", 0)), None, 0, 0))
    arr = (abi.hs_alloc_entry * max(1, len(picks)))(*picks)
    z = (C.c_int32 * max(1, len(zones)))(*zones)
    prev = oa.devs_array([oa.dev_from_go(go_device(d)) for d in previous])
    kv = (abi.hs_kv * max(1, len(annos)))(*[abi.hs_kv(abi._b(k), abi._b(v)) for k, v in annos.items()])
    L = oa.lib()
    L.oracle_sgpu_score.restype = C.c_float
    L.oracle_sgpu_score.argtypes = [C.c_void_p, C.POINTER(abi.hs_alloc_entry), C.c_uint32, C.POINTER(C.c_int32), C.POINTER(oa.oracle_dev),
                                    C.c_uint32, C.POINTER(abi.hs_kv), C.c_uint32]
    return L.oracle_sgpu_score(orc.h, arr, len(picks), z if has_zone else None, prev, len(previous), kv, len(annos))


@pytest.mark.parametrize("case", load("metax_sgpu_score_exclusive"), ids=ids(load("metax_sgpu_score_exclusive")))
def test_score_exclusive_table(orc, case):
    # sdevice_test.go:1706-2101 (TestScoreExclusiveDevices): topology awareness switched on by the pod's annotation
    assert _score(orc, case["podDevices"], case["previous"], {TOPO: "true"}) == case["expectedScore"]
    assert _score(orc, case["podDevices"], case["previous"], {}) == 0      # MetaxConfig.TopologyAware defaults to false


@pytest.mark.parametrize("case", load("metax_sgpu_score_online"), ids=ids(load("metax_sgpu_score_online")))
def test_score_online_table(orc, case):
    # sdevice_test.go:2775-2945 (TestScoreOnlineDevices)
    assert _score(orc, case["pickDevs"], case["devs"], {CLASS: "online"}) == case["expectedScore"]


# ---- GPU parity ------------------------------------------------------------------------------------------------------
def sgpu_node(rng, name, cards=8):
    devs = []
    for i in range(cards):
        devs.append({"id": "%s-sgpu-%d" % (name, i), "index": i, "count": rng.choice([4, 8, 16]), "devmem": 65536, "devcore": 100,
                     "type": SGPU, "numa": i // 4, "mode": "", "health": rng.random() > 0.04, "vendor": SGPU,
                     "sgpu": {"qos_policy": rng.choice(["", "best-effort", "fixed-share", "burst-share", "best-effort"]),
                              "link_zone": rng.choice([0, 1, 1, 2, 2, 3])}})
    return {"name": name, "devices": devs}


def sgpu_pod(rng, uid):
    annos = {}
    r = rng.random()
    if r < 0.2:
        annos[CLASS] = "online"
    elif r < 0.35:
        annos[CLASS] = "offline"
    if annos.get(CLASS) != "online":
        annos[QOS] = rng.choice(["best-effort", "best-effort", "fixed-share", "burst-share"])
    if rng.random() < 0.5:
        annos[TOPO] = rng.choice(["true", "true", "false", "bogus"])
    if rng.random() < 0.08:
        annos["hami.io/gpu-scheduler-policy"] = "mutex"
    if rng.random() < 0.1:
        annos["hami.io/node-scheduler-policy"] = rng.choice(["spread", "binpack"])
    classed = CLASS in annos
    nums = 1 if classed else rng.choice([1, 1, 2, 2, 4])
    whole = (not classed) and rng.random() < 0.45
    req = {"key": SGPU, "type": SGPU, "nums": nums, "memreq": 0 if whole else rng.choice([0, 4096, 16384]),
           "mempct": 100, "coresreq": 100 if whole else rng.choice([0, 10, 30, 50])}
    if req["memreq"]:
        req["mempct"] = 0
    ctrs = [[req]]
    if not classed and rng.random() < 0.2:
        ctrs.append([{"key": SGPU, "type": SGPU, "nums": 1, "memreq": 2048, "mempct": 0, "coresreq": rng.choice([10, 100])}])
    pod = {"uid": uid, "ns": "default", "name": uid, "annotations": annos, "containers": ctrs, "n_init": 0}
    if len(ctrs) > 1 and rng.random() < 0.3:
        pod["n_init"] = 1
    return pod


def sgpu_cluster(seed, n_nodes=24, n_pods=900, topo_default=False):
    rng = random.Random(seed)
    nodes = [sgpu_node(rng, "mx-%02d" % i, cards=rng.choice([2, 4, 8, 8])) for i in range(n_nodes)]
    ops, live = [], []
    for p in range(n_pods):
        if live and rng.random() < 0.3:
            ops.append(("delete", live.pop(rng.randrange(len(live)))))
            continue
        uid = "pod-%04d" % p
        ops.append(("filter", sgpu_pod(rng, uid)))
        live.append(uid)
    return {"node_policy": rng.choice(["binpack", "spread"]), "gpu_policy": rng.choice(["binpack", "spread"])}, nodes, ops


def _vendors(topo_default):
    v = [list(x) for x in abi.DEFAULT_VENDORS]
    for x in v:
        if x[0] == SGPU:
            while len(x) < 9:
                x.append(None)
            x[8] = 1 if topo_default else 0
    return [tuple(x) for x in v]


@pytest.mark.gpu
@pytest.mark.parametrize("seed,topo_default", [(4, False), (7, True)])
def test_metax_sgpu_gpu_parity(seed, topo_default):
    import parity
    import k8s_vgpu_scheduler_b200.engine as eng
    cfg, nodes, ops = sgpu_cluster(seed)
    o = oa.Oracle(cfg["node_policy"], cfg["gpu_policy"], vendors=_vendors(topo_default))
    e = eng.Engine(cfg["node_policy"], cfg["gpu_policy"], vendors=_vendors(topo_default))
    for n in nodes:
        o.node_upsert(n)
        e.node_upsert(n)
    ref, got = [], []
    rng = random.Random(seed + 100)
    for s in range(0, len(ops), 150):
        part = ops[s:s + 150]
        for kind, arg in part:
            ref.append(o.filter(arg) if kind == "filter" else {"status": o.pod_delete(arg)})
        e.run_batch(eng.PackedOps(part))
        got += [e.batch_get(i) for i in range(len(part))]
        assert parity.compare(ops[:len(ref)], ref, got) == [], "batch at %d" % s
        for n in nodes:
            assert e.snapshot(n["name"]) == o.snapshot(n["name"])
            assert e.snapshot_node_sgpu(n["name"]) == o.snapshot_node_sgpu(n["name"]), (s, n["name"])
            dev_rows, mirror_rows = e.device_rows(n["name"])
            assert dev_rows == mirror_rows, (s, n["name"], [(hex(a[3] & 0xffffffff), hex(b[3] & 0xffffffff)) for a, b in zip(dev_rows, mirror_rows) if a != b])
        # the device plugin catches up with some of the QoS switches: those nodes register again with the new policy
        # (JitteryQosCache.Sync retires the switch), others with the old one (the switch stays in flight)
        for n in rng.sample(nodes, 5):
            for d in n["devices"]:
                if rng.random() < 0.5:
                    d["sgpu"]["qos_policy"] = rng.choice(["", "best-effort", "fixed-share", "burst-share"])
            o.node_upsert(n)
            e.node_upsert(n)
    fits = [r for (k, a), r in zip(ops, ref) if k == "filter" and r["status"] == abi.HS_FIT]
    nofit = [r for (k, a), r in zip(ops, ref) if k == "filter" and r["status"] == abi.HS_NOFIT]
    assert len(fits) > 250 and len(nofit) > 20
    reasons = " ".join(v for r in nofit for v in r["failed"].values())
    assert "CardQosPolicyMismatch" in reasons and "CardAppClassMismatch" in reasons
    assert len({round(r["score"], 1) for r in fits}) > 20
    assert any(abs(r["score"]) > 5000 for r in fits)                 # the +-10000 weight was in play
    assert all("metax-tech.com/gpu-devices-allocated" in r["annotations"] for r in fits)
    e.close()
    o.close()


# ---- registration ----------------------------------------------------------------------------------------------------
def test_get_node_devices_from_the_node_annotation():
    # sdevice_test.go:507-590 (TestGetMetaxSDevices) + convertMetaxSDeviceToHAMIDevice (protocol.go:96-116)
    import k8s_vgpu_scheduler_b200.engine as eng
    anno = json.dumps([{"uuid": "GPU-123", "bdf": "0000:1a:00.0", "model": "MXC500", "totalDevCount": 16, "totalCompute": 100,
                        "totalVRam": 32768, "availableDevCount": 16, "availableCompute": 100, "availableVRam": 32768,
                        "numa": 1, "healthy": True, "qosPolicy": "fixed-share"},
                       {"uuid": "GPU-456", "totalDevCount": 8, "totalCompute": 100, "totalVRam": 65536, "numa": 0, "healthy": False,
                        "qosPolicy": "", "linkZone": 2}])
    devs, pair = eng.decode_node_devices(SGPU, abi.HS_REG_METAX_SGPU, anno, None, "node-a")
    assert pair is None
    assert [(d["ID"], d["Index"], d["Count"], d["Devmem"], d["Devcore"], d["Type"], d["Numa"], d["Mode"], d["Health"], d["DeviceVendor"])
            for d in devs] == [("GPU-123", 0, 16, 32768, 100, SGPU, 1, "", True, SGPU), ("GPU-456", 1, 8, 65536, 100, SGPU, 0, "", False, SGPU)]
    assert [d["Sgpu"] for d in devs] == [{"qos_policy": "fixed-share", "link_zone": 0}, {"qos_policy": "", "link_zone": 2}]
    for bad, msg in ((None, "annos not found metax-tech.com/node-gpu-devices"), ("", "unexpected end of JSON input"),
                     ("[]", "no sdevice found on node"), ('{"uuid":"x"}', "cannot unmarshal object"),
                     ('[{"uuid":"x","totalVRam":"big"}]', "MetaxSDeviceInfo.totalVRam of type int32")):
        with pytest.raises(eng.IngestError) as ei:
            eng.decode_node_devices(SGPU, abi.HS_REG_METAX_SGPU, bad, None, "n")
        assert msg in str(ei.value), (bad, str(ei.value))


@pytest.mark.gpu
def test_sgpu_nodes_registered_from_the_annotation_decide_identically():
    import parity
    import k8s_vgpu_scheduler_b200.engine as eng
    cfg, nodes, ops = sgpu_cluster(11, n_nodes=16, n_pods=400)
    o = oa.Oracle(cfg["node_policy"], cfg["gpu_policy"], vendors=_vendors(False))
    e = eng.Engine(cfg["node_policy"], cfg["gpu_policy"], vendors=_vendors(False))
    for n in nodes:
        o.node_upsert(n)
        anno = json.dumps([{"uuid": d["id"], "totalDevCount": d["count"], "totalCompute": d["devcore"], "totalVRam": d["devmem"],
                            "numa": d["numa"], "healthy": d["health"], "qosPolicy": d["sgpu"]["qos_policy"],
                            "linkZone": d["sgpu"]["link_zone"]} for d in n["devices"]])
        assert e.node_register(n["name"], [(SGPU, abi.HS_REG_METAX_SGPU, anno, None)]) == len(n["devices"]), e.error()
    ref = [o.filter(arg) if kind == "filter" else {"status": o.pod_delete(arg)} for kind, arg in ops]
    e.run_batch(eng.PackedOps(ops))
    got = [e.batch_get(i) for i in range(len(ops))]
    assert parity.compare(ops, ref, got) == []
    for n in nodes:
        assert e.snapshot_node_sgpu(n["name"]) == o.snapshot_node_sgpu(n["name"])
    assert sum(r["status"] == abi.HS_FIT for r in ref) > 100
    e.close()
    o.close()


# ---- the in-flight QoS switch cache (JitteryQosCache), replayed through the oracle's Filter / register flow -----------
def _qos_node(policies):
    return {"name": "mx-q", "devices": [
        {"id": uid, "index": i, "count": 16, "devmem": 32768, "devcore": 100, "type": SGPU, "numa": 0, "mode": "", "health": True,
         "vendor": SGPU, "sgpu": {"qos_policy": pol, "link_zone": 0}} for i, (uid, pol) in enumerate(policies)]}


def _qos_pod(uid, qos, card, cores):
    return {"uid": uid, "ns": "default", "name": uid, "annotations": {QOS: qos, "metax-tech.com/use-gpuuuid": card},
            "containers": [[{"key": SGPU, "type": SGPU, "nums": card.count(",") + 1, "memreq": 1024, "mempct": 0, "coresreq": cores}]]}


def _pending(o):
    return {uid: p for uid, (_, _, p) in zip(("GPU-123", "GPU-456", "GPU-789"), o.snapshot_node_sgpu("mx-q")) if p is not None}


def test_add_jittery_qos_vector():
    # sdevice_test.go:680-737 (TestAddJitteryQos): a best-effort pod on GPU-123 (best-effort) and GPU-456 (burst-share) at 50
    # cores, and on GPU-789 (best-effort) at 100 cores -> {GPU-456: best-effort, GPU-789: ""}; the device that already runs the
    # pod's policy gets no entry, a whole card is expected to drop its policy
    o = oa.Oracle()
    o.node_upsert(_qos_node([("GPU-123", "best-effort"), ("GPU-456", "burst-share"), ("GPU-789", "best-effort")]))
    assert o.filter(_qos_pod("p1", "best-effort", "GPU-123,GPU-456", 50))["status"] == abi.HS_FIT
    assert _pending(o) == {"GPU-456": "best-effort"}
    assert o.filter(_qos_pod("p2", "best-effort", "GPU-789", 100))["status"] == abi.HS_FIT
    assert _pending(o) == {"GPU-456": "best-effort", "GPU-789": ""}
    o.close()


@pytest.mark.parametrize("first,placed,want", [
    # qos_test.go:24-120 (TestJitteryQosCacheSync): every device reports best-effort at the sync
    (("burst-share", "best-effort"), ("fixed-share", "burst-share"), {"GPU-123": "fixed-share", "GPU-789": "burst-share"}),   # "no sync"
    (("fixed-share", "best-effort"), ("best-effort", "burst-share"), {"GPU-789": "burst-share"}),                             # one retired
    (("fixed-share", "burst-share"), ("best-effort", "best-effort"), {}),                                                    # both retired
], ids=["no-sync", "sync-one", "sync-both"])
def test_jittery_qos_cache_sync_vectors(first, placed, want):
    o = oa.Oracle()
    o.node_upsert(_qos_node([("GPU-123", first[0]), ("GPU-456", "best-effort"), ("GPU-789", first[1])]))
    assert o.filter(_qos_pod("a", placed[0], "GPU-123", 50))["status"] == abi.HS_FIT
    assert o.filter(_qos_pod("b", placed[1], "GPU-789", 50))["status"] == abi.HS_FIT
    assert _pending(o) == {"GPU-123": placed[0], "GPU-789": placed[1]}          # the currentCache of the reference's case
    o.node_upsert(_qos_node([("GPU-123", "best-effort"), ("GPU-456", "best-effort"), ("GPU-789", "best-effort")]))   # Sync(devices)
    assert _pending(o) == want
    o.close()


def test_check_device_qos_reads_the_pending_switch():
    # sdevice.go:528-556 with a cache entry (JitteryQosCache.Get, qos_test.go:122-160): while GPU-123 is switching to fixed-share a
    # second fixed-share pod shares it, a best-effort pod — the policy the device still REPORTS — is refused
    o = oa.Oracle()
    o.node_upsert(_qos_node([("GPU-123", "best-effort"), ("GPU-456", "best-effort"), ("GPU-789", "best-effort")]))
    assert o.filter(_qos_pod("a", "fixed-share", "GPU-123", 30))["status"] == abi.HS_FIT
    assert _pending(o) == {"GPU-123": "fixed-share"}
    assert o.filter(_qos_pod("b", "fixed-share", "GPU-123", 30))["status"] == abi.HS_FIT
    r = o.filter(_qos_pod("c", "best-effort", "GPU-123", 30))
    assert r["status"] == abi.HS_NOFIT and "CardQosPolicyMismatch" in " ".join(r["failed"].values())
    o.close()


@pytest.mark.parametrize("node_policy", ["binpack", "spread"])
def test_score_node_is_policy_neutral_through_filter(node_policy):
    # sdevice_test.go:2947-3035 (TestScoreNodePolicyIndependence) at the level where it matters: the node that keeps its MetaLink
    # zones whole wins the topology-aware pod under BOTH node policies (a policy-signed weight would flip the choice under one)
    def node(name, zones):
        return {"name": name, "devices": [
            {"id": "%s-%d" % (name, i), "index": i, "count": 16, "devmem": 32768, "devcore": 100, "type": SGPU, "numa": 0, "mode": "",
             "health": True, "vendor": SGPU, "sgpu": {"qos_policy": "", "link_zone": z}} for i, z in enumerate(zones)]}
    pod = {"uid": "t", "ns": "default", "name": "t", "annotations": {TOPO: "true"},
           "containers": [[{"key": SGPU, "type": SGPU, "nums": 2, "memreq": 0, "mempct": 100, "coresreq": 100}]]}
    scores = {}
    for order in (("whole", "split"), ("split", "whole")):          # the candidate order must not decide it either
        o = oa.Oracle(node_policy, "binpack")
        o.node_upsert(node("whole", [1, 1, 2, 2]))                  # taking one zone's two cards breaks no pair elsewhere
        o.node_upsert(node("split", [1, 1, 1, 1]))                  # taking two of four breaks five of the six pairs
        r = o.filter(pod, node_names=list(order))
        assert r["status"] == abi.HS_FIT and r["node"] == "whole", (node_policy, order, r["node"], r["score"])
        scores[order] = r["score"]
        o.close()
    assert len(set(scores.values())) == 1


def test_allocated_annotation_is_metaxs_own_json():
    # PatchAnnotations (sdevice.go:142-171) + convertHAMIPodDeviceToMetaxPodDevice (protocol.go:129-145; protocol_test.go:118-198):
    # per container a list of {"uuid", "compute", "vRam"} with Go's omitempty on the two numbers, beside the two hami keys
    o = oa.Oracle()
    o.node_upsert(_qos_node([("GPU-123", ""), ("GPU-456", ""), ("GPU-789", "")]))
    pod = {"uid": "j", "ns": "default", "name": "j", "annotations": {},
           "containers": [[{"key": SGPU, "type": SGPU, "nums": 2, "memreq": 1024, "mempct": 0, "coresreq": 0}],
                          [{"key": SGPU, "type": SGPU, "nums": 1, "memreq": 10240, "mempct": 0, "coresreq": 20}]]}
    r = o.filter(pod)
    assert r["status"] == abi.HS_FIT
    ctrs = r["devices"][SGPU]
    want = [[{"uuid": d[0]} | ({"compute": d[3]} if d[3] else {}) | ({"vRam": d[2]} if d[2] else {}) for d in c] for c in ctrs]
    got = r["annotations"]["metax-tech.com/gpu-devices-allocated"]
    assert got == json.dumps(want, separators=(",", ":"))
    assert [len(c) for c in json.loads(got)] == [2, 1] and "compute" not in json.loads(got)[0][0] and json.loads(got)[1][0]["compute"] == 20
    assert r["annotations"]["metax-tech.com/predicate-time"] == ""               # the caller stamps it
    assert r["annotations"]["hami.io/metax-sgpu-devices-to-allocate"] == r["annotations"]["hami.io/metax-sgpu-devices-allocated"]
    assert QOS not in r["annotations"]        # the "best-effort" default is the webhook's (MutateAdmission, :89-129), not Filter's
    o.close()


@pytest.mark.parametrize("node_policy,want", [("binpack", 600000.0), ("spread", -600000.0)])
def test_override_score_weights_the_policy_neutral_term(node_policy, want):
    # policy/node_policy_test.go:303-407 (TestOverrideScore, the two MetaX-SGPU cases): four idle cards of one MetaLink zone, a
    # topology-aware pod takes two whole cards -> raw score 60 (scoreExclusiveDevices), weighted by +10000 under binpack and
    # -10000 under spread on top of a default node score of 0
    o = oa.Oracle(node_policy, "binpack")
    o.node_upsert({"name": "node1", "devices": [
        {"id": "GPU-%d" % (i + 1), "index": i, "count": 16, "devmem": 32768, "devcore": 100, "type": SGPU, "numa": 0, "mode": "",
         "health": True, "vendor": SGPU, "sgpu": {"qos_policy": "", "link_zone": 1}} for i in range(4)]})
    r = o.filter({"uid": "t", "ns": "default", "name": "t", "annotations": {TOPO: "true"},
                  "containers": [[{"key": SGPU, "type": SGPU, "nums": 2, "memreq": 0, "mempct": 100, "coresreq": 100}]]})
    assert r["status"] == abi.HS_FIT and r["score"] == want, r["score"]
    o.close()
