"""kunlun whole XPUs: the oracle against the reference's own kunlun tests — the Test_graphSelect and
Test_ScoreNode tables (extracted into tests/golden by tools/extract_golden.py) and the non-table tests of
pkg/device/kunlun/device_test.go and topo_test.go, transcribed with their file:line."""
import ctypes as C

import pytest

import oracle_abi as oa
from oracle_abi import abi
from test_oracle_golden import ids, load, run_fit


@pytest.fixture(scope="module")
def orc():
    o = oa.Oracle()
    yield o
    o.close()


def xpus(used=(), unhealthy=(), n=8):
    # topo_test.go:31-52 (newDevices) / device_test.go:521-531
    return [{"ID": "xpu-%d" % i, "Index": i, "Used": 1 if i in used else 0, "Count": 1, "Totalmem": 98304, "Totalcore": 100,
             "Type": "kunlun", "Health": i not in unhealthy} for i in range(n)]


def fit(orc, devices, nums, annos=None):
    ok, got, reason = run_fit(orc, "kunlun", devices, {"Nums": nums, "Type": "kunlun"}, annos or {})
    return ok, [g[1] for g in got], got, reason


@pytest.mark.parametrize("case", load("kunlun_graph_select"), ids=ids(load("kunlun_graph_select")))
def test_graph_select_table(orc, case):
    # kunlun/device_test.go:83-279: graghSelect with FitXPU through Fit (the picks are the returned card indices in order)
    devs = [dict(d, ID="xpu-%d" % d["Index"], Type="kunlun", Totalmem=98304, Totalcore=100) for d in case["args"]["d"]]
    ok, idx, got, _ = fit(orc, devs, case["args"]["c"]["Nums"])
    assert idx == list(case["want1"] or [])
    assert ok == bool(case["want1"])
    for g in got:
        assert g[2:4] == (98304, 100)      # the pick takes the whole card (device.go:215-221)


@pytest.mark.parametrize("case", load("kunlun_score_node"), ids=ids(load("kunlun_score_node")))
def test_score_node_table(case):
    # kunlun/device_test.go:419-517
    prev = [d["Index"] for d in case["args"]["usage"] if d.get("Used", 0) > 0]
    cur = [d["Idx"] for ctr in case["args"]["podDevices"] for d in ctr]
    assert oa.kunlun_calc_score(prev, cur) == case["want"]


def test_topo_unit_vectors():
    # topo_test.go:125-192: getvalue / countbubble / calcscore
    cs = oa.kunlun_calc_score
    assert cs([], []) == 2000                       # 1 - 1
    assert cs([0, 1, 2, 3], [0, 1, 2, 3, 4, 5, 6, 7]) == 3000      # prev 0+1=1 … cur 0+0=0: -1
    assert cs([], [0]) == 0                         # 1 → 2+1=3: +2
    assert cs([0], [0, 1]) == 3000                  # 3 → 2: -1
    assert cs([0, 1, 2, 3], [4]) == 0               # prev 0+1 = 1, cur 1+2 = 3: +2
    assert cs([0, 4], [0, 1, 4, 5]) == 1000         # 4 → 2: -2 falls through to the default


def test_graph_select_topo_vectors(orc):
    # topo_test.go:259-366
    assert fit(orc, xpus(), 8)[1] == [0, 1, 2, 3, 4, 5, 6, 7]                 # TestGraghSelect_AllEightAvailable
    assert fit(orc, xpus(used=[3]), 8)[0] is False                           # …EightNotFullyAvailable
    ok, idx, _, _ = fit(orc, xpus(used=[0, 1, 2]), 1)                         # …SingleDevice: the wing with exactly one free card
    assert ok and idx == [3]
    ok, idx, _, _ = fit(orc, xpus(used=[4, 5, 6, 7]), 4)                      # …FourFromLeftWing
    assert ok and idx == [0, 1, 2, 3]
    ok, idx, _, _ = fit(orc, xpus(used=[1, 2, 3, 5, 6, 7]), 2)                # TestInterconnect_TwoDevices: 0-4 is an inter-group pair
    assert ok and sorted(idx) == [0, 4]
    ok, idx, _, _ = fit(orc, xpus(used=[2, 3, 6, 7]), 4)                      # TestInterconnect_FourDevices: 0-1-4-5
    assert ok and idx == [0, 1, 4, 5]
    assert fit(orc, xpus(), 3)[0] is False                                    # TestInterconnect_UnsupportedCount
    assert fit(orc, xpus(), 3)[3] == "1/8 NumaNotFit"


def test_fit_reasons_and_uuid_annotations(orc):
    # device_test.go:70-81 (TestKunlunDevices_Fit_NumaNotFit)
    devs = [{"Index": i, "Used": 1, "ID": "x%d" % i, "Type": "kunlun"} for i in range(8)]
    ok, _, _, reason = fit(orc, devs, 1)
    assert not ok and reason == "1/8 NumaNotFit"
    # device_test.go:519-569 (TestKunlunDevices_Fit_UseUUID): the legacy pair, the standard pair, the literal key
    for key, want in (("baidu.com/use-gpuuuid", "xpu-5"), ("hami.io/use-xpu-uuid", "xpu-3"), ("hami.io/use-xpu-uuid", "xpu-6")):
        ok, _, got, _ = fit(orc, xpus(), 1, {key: want})
        assert ok and [g[0] for g in got] == [want]
    ok, _, _, reason = fit(orc, xpus(), 1, {"hami.io/use-xpu-uuid": "xpu-99"})
    assert not ok and "CardUuidMismatch" in reason and reason == "8/8 CardUuidMismatch"
    # device_test.go:571-607 (TestKunlunDevices_Fit_NoUseUUID)
    every = ",".join("xpu-%d" % i for i in range(8))
    for key in ("baidu.com/nouse-gpuuuid", "hami.io/no-use-xpu-uuid"):
        ok, _, _, reason = fit(orc, xpus(), 1, {key: every})
        assert not ok and "CardUuidMismatch" in reason
    # both pairs at once: a card must pass both
    ok, _, got, _ = fit(orc, xpus(), 1, {"hami.io/use-xpu-uuid": "xpu-1,xpu-2", "baidu.com/use-gpuuuid": "xpu-2,xpu-3"})
    assert ok and [g[0] for g in got] == ["xpu-2"]


# ---- the shared "XPU" type (kunlun/vdevice.go) ------------------------------------------------------------------------
def fit_vxpu(orc, devices, request, annos=None):
    ok, got, reason = run_fit(orc, "XPU", devices, dict(request, Type="XPU"), annos or {})
    return ok, [g[1] for g in got], got, reason


@pytest.mark.parametrize("case", load("kunlun_graph_select_vxpu"), ids=ids(load("kunlun_graph_select_vxpu")))
def test_graph_select_vxpu_table(orc, case):
    # kunlun/device_test.go:281-417: graghSelect with FitVXPU
    devs = [dict(d, ID="xpu-%d" % d["Index"], Type="XPU", Totalcore=100) for d in case["args"]["d"]]
    ok, idx, got, _ = fit_vxpu(orc, devs, case["args"]["c"])
    assert idx == list(case["want1"] or []) and ok == bool(case["want1"])
    for g in got:
        assert g[2] == case["args"]["c"].get("Memreq", 0)       # the pick records the request's memory (vdevice.go:282)


@pytest.mark.parametrize("case", load("kunlun_fit_vxpu_direct"), ids=ids(load("kunlun_fit_vxpu_direct")))
def test_fit_vxpu_direct(orc, case):
    # kunlun/vdevice_test.go:498-541: one card, one device asked
    dev = dict(case["usage"], ID="xpu-0", Index=0, Type="XPU")
    ok, _, _, _ = fit_vxpu(orc, [dev], dict(case["request"], Nums=1))
    assert ok == case["want"]


def test_vxpu_fit_reasons(orc):
    # device_test.go:32-55 (TestKunlunVDevices_Fit_Mutex): shareable cards, all in use
    devs = [{"ID": "x%d" % i, "Index": i, "Used": 1, "Usedmem": 24576, "Totalmem": 98304, "Health": True, "Type": "XPU"} for i in range(8)]
    assert fit_vxpu(orc, devs, {"Nums": 1, "Memreq": 24576})[0] is True
    ok, _, _, reason = fit_vxpu(orc, devs, {"Nums": 1, "Memreq": 24576}, {"hami.io/gpu-scheduler-policy": "mutex"})
    assert not ok and reason == "8/8 ExclusiveDeviceAllocateConflict"
    # device_test.go:57-68 (TestKunlunVDevices_Fit_NumaNotFit)
    devs = [{"ID": "x%d" % i, "Index": i, "Used": 0, "Usedmem": 0, "Totalmem": 1024, "Type": "XPU"} for i in range(8)]
    ok, _, _, reason = fit_vxpu(orc, devs, {"Nums": 1, "Memreq": 2048})
    assert not ok and reason == "1/8 NumaNotFit"
    # device_test.go:609-651 (TestKunlunVDevices_Fit_UUIDAnnotations)
    devs = [{"ID": "xpu-%d" % i, "Index": i, "Count": 10, "Totalmem": 98304, "Totalcore": 100, "Health": True, "Type": "XPU"} for i in range(8)]
    req = {"Nums": 1, "Memreq": 98304}
    for key, want in (("hami.io/use-xpu-uuid", "xpu-3"), ("baidu.com/use-gpuuuid", "xpu-2")):
        ok, _, got, _ = fit_vxpu(orc, devs, req, {key: want})
        assert ok and got[0][0] == want
    every = ",".join("xpu-%d" % i for i in range(8))
    for key in ("hami.io/no-use-xpu-uuid", "baidu.com/nouse-gpuuuid"):
        ok, _, _, reason = fit_vxpu(orc, devs, req, {key: every})
        assert not ok and "CardUuidMismatch" in reason


# ---- GPU parity ----------------------------------------------------------------------------------------------------
def _kunlun_node(name, cards=8, extra=None):
    devs = [{"id": "%s-kunlun-%d" % (name, i), "index": i, "count": 100, "devmem": 98304, "devcore": 100, "type": "kunlun",
             "numa": 1 if i >= cards // 2 else 0, "mode": "", "health": True, "vendor": "kunlun"} for i in range(cards)]
    return {"name": name, "devices": devs + (extra or [])}


def _kunlun_cluster(seed, n_nodes=40, n_pods=900):
    import random
    rng = random.Random(seed)
    nodes = []
    for i in range(n_nodes):
        name = "kl-%02d" % i
        r = rng.random()
        if r < 0.7:
            n = _kunlun_node(name)
            if rng.random() < 0.15:
                n["devices"][rng.randrange(8)]["health"] = False
        elif r < 0.8:
            n = _kunlun_node(name, cards=4)                 # a half-populated box: one wing only
        elif r < 0.9:                                        # kunlun beside NVIDIA: ScoreNode reads 0 on such a node
            n = _kunlun_node(name, extra=[])
            n["devices"] = n["devices"][:4] + [
                {"id": "%s-gpu-%d" % (name, j), "index": j, "count": 10, "devmem": 81920, "devcore": 100, "type": "NVIDIA-H100",
                 "numa": 0, "mode": "hami-core", "health": True, "vendor": "NVIDIA"} for j in range(4)]
        else:
            n = {"name": name, "devices": [
                {"id": "%s-gpu-%d" % (name, j), "index": j, "count": 10, "devmem": 81920, "devcore": 100, "type": "NVIDIA-H100",
                 "numa": j // 4, "mode": "hami-core", "health": True, "vendor": "NVIDIA"} for j in range(8)]}
        nodes.append(n)
    for i in range(n_nodes // 3):                            # shared XPUs (registered from hami.io/node-register-xpu in the reference)
        name = "vx-%02d" % i
        nodes.append({"name": name, "devices": [
            {"id": "%s-xpu-%d" % (name, j), "index": j, "count": 10, "devmem": 98304, "devcore": 100, "type": "XPU-P800",
             "numa": j // 4, "mode": "", "health": rng.random() > 0.05, "vendor": "XPU"} for j in range(8)]})
    kl_nodes = [n for n in nodes if any(d["vendor"] == "kunlun" for d in n["devices"])]
    ops, live = [], []
    for p in range(n_pods):
        if live and rng.random() < 0.33:
            ops.append(("delete", live.pop(rng.randrange(len(live)))))
            continue
        uid = "pod-%04d" % p
        annos = {}
        r = rng.random()
        if r < 0.1:
            annos["hami.io/node-scheduler-policy"] = rng.choice(["spread", "binpack"])
        if rng.random() < 0.1:
            annos["hami.io/gpu-scheduler-policy"] = rng.choice(["spread", "binpack"])
        if rng.random() < 0.12:                              # uuid lists, either pair or both
            n = rng.choice(kl_nodes)
            ids_ = [d["id"] for d in n["devices"] if d["vendor"] == "kunlun"]
            pick = rng.sample(ids_, rng.randint(1, len(ids_)))
            key = rng.choice(["hami.io/use-xpu-uuid", "baidu.com/use-gpuuuid", "hami.io/no-use-xpu-uuid", "baidu.com/nouse-gpuuuid"])
            annos[key] = ",".join(pick)
            if rng.random() < 0.4:
                key2 = rng.choice(["hami.io/use-xpu-uuid", "baidu.com/use-gpuuuid", "hami.io/no-use-xpu-uuid", "baidu.com/nouse-gpuuuid"])
                annos[key2] = ",".join(rng.sample(ids_, rng.randint(1, len(ids_))))
        kl = lambda n: {"key": "kunlun", "type": "kunlun", "nums": n, "memreq": 0, "mempct": 100, "coresreq": 0}  # noqa: E731
        nv = lambda n: {"key": "NVIDIA", "type": "NVIDIA", "nums": n, "memreq": 8000, "mempct": 101, "coresreq": 10}  # noqa: E731
        vx = lambda n, m: {"key": "XPU", "type": "XPU", "nums": n, "memreq": m, "mempct": 0 if m != 98304 else 100,  # noqa: E731
                           "coresreq": m * 100 // 98304}
        r = rng.random()
        if r < 0.3:
            ctrs = [[vx(rng.choice([1, 1, 2, 4, 8, 3]), rng.choice([24576, 24576, 49152, 98304]))]]
            if rng.random() < 0.2:
                ctrs.append([vx(rng.choice([1, 2]), rng.choice([24576, 49152]))])
            if rng.random() < 0.2:
                annos["hami.io/gpu-scheduler-policy"] = "mutex"
            if rng.random() < 0.1:
                annos[rng.choice(["hami.io/use-xpu-uuid", "baidu.com/nouse-gpuuuid"])] = ",".join(
                    "vx-%02d-xpu-%d" % (rng.randrange(n_nodes // 3), rng.randrange(8)) for _ in range(rng.randint(1, 5)))
        elif r < 0.38:
            ctrs = [[nv(rng.choice([1, 2]))]]
        elif r < 0.45:
            ctrs = [[kl(rng.choice([1, 2])), nv(1)]]          # both vendors in one container: only the mixed nodes can serve it
        else:
            ctrs = [[kl(rng.choice([1, 1, 2, 2, 4, 4, 8, 3]))]]
            if rng.random() < 0.25:
                ctrs.append([kl(rng.choice([1, 2, 4]))])
        pod = {"uid": uid, "ns": "default", "name": uid, "annotations": annos, "containers": ctrs, "n_init": 0}
        if len(ctrs) > 1 and rng.random() < 0.4:
            pod["n_init"] = 1
        ops.append(("filter", pod))
        live.append(uid)
    return {"node_policy": "binpack", "gpu_policy": rng.choice(["binpack", "spread"])}, nodes, ops


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [3, 8])
def test_kunlun_gpu_parity(seed):
    import parity
    import k8s_vgpu_scheduler_b200.engine as eng
    cfg, nodes, ops = _kunlun_cluster(seed)
    ref, ref_snap = parity.run_oracle(cfg, nodes, ops)
    got, snap = parity.run_engine(cfg, nodes, ops, chunk=200)
    assert parity.compare(ops, ref, got, ref_snap, snap) == []
    fits = [r for (k, a), r in zip(ops, ref) if k == "filter" and r["status"] == abi.HS_FIT and any(r["devices"].get("kunlun", []))]
    nofit = [r for (k, a), r in zip(ops, ref) if k == "filter" and r["status"] == abi.HS_NOFIT]
    assert len(fits) > 100 and len(nofit) > 20
    scores = {round(r["score"]) // 1000 for r in fits}
    assert len(scores) >= 3                                  # the wing score moved the node choice around (0 / 1000 / 2000 / 3000 + usage)
    assert any("CardUuidMismatch" in v for r in nofit for v in r["failed"].values())
    assert any("NumaNotFit" in v for r in nofit for v in r["failed"].values())
    assert all("hami.io/kunlun-allocated" in r["annotations"] and "BAIDU_COM_DEVICE_IDX" in r["annotations"] for r in fits)
    vfits = [r for (k, a), r in zip(ops, ref) if k == "filter" and r["status"] == abi.HS_FIT and any(r["devices"].get("XPU", []))]
    assert len(vfits) > 60 and all("hami.io/xpu-devices-allocated" in r["annotations"] and "hami.io/xpu-devices-to-allocate" in r["annotations"] for r in vfits)
    assert any("ExclusiveDeviceAllocateConflict" in v for r in nofit for v in r["failed"].values())
    # the same all-kunlun nodes registered from capacity (kunlun/device.go:64-91) decide identically
    e = eng.Engine(cfg.get("node_policy"), cfg.get("gpu_policy"))
    n_cap = 0
    for n in nodes:
        full = all(d["vendor"] == "kunlun" and d["health"] for d in n["devices"])
        if full:
            nd = len(n["devices"])
            assert e.node_register(n["name"], [("kunlun", abi.HS_REG_CAPACITY_KUNLUN, None, None, None, nd, 0)]) == nd, e.error()
            n_cap += 1
        else:
            e.node_upsert(n)
    assert n_cap > 10
    e.run_batch(eng.PackedOps(ops))
    got2 = [e.batch_get(i) for i in range(len(ops))]
    e.close()
    assert parity.compare(ops, ref, got2) == []


@pytest.mark.gpu
def test_general_path_in_streaming_mode(monkeypatch):
    """HS_NO_RESIDENT=1 makes the decide loop take its streaming mode on a small inventory: there the general
    evaluator's node list of a group is dealt to its 8 lanes (one thread evaluates AND commits a node).  Same stream,
    same answers."""
    import parity
    monkeypatch.setenv("HS_NO_RESIDENT", "1")
    cfg, nodes, ops = _kunlun_cluster(11, n_nodes=30, n_pods=500)
    ref, ref_snap = parity.run_oracle(cfg, nodes, ops)
    got, snap = parity.run_engine(cfg, nodes, ops, chunk=125)
    assert parity.compare(ops, ref, got, ref_snap, snap) == []


def test_vxpu_get_node_devices_vectors():
    # vdevice_test.go:215-277 (Test_GetNodeDevices): missing annotation, invalid JSON, empty list, and a list whose devices
    # take the vendor word; ascend words the empty list the same way (ascend/device.go:212), the others say "gpu"
    import k8s_vgpu_scheduler_b200.engine as eng
    for anno, msg in ((None, "annos not found"), ("not-json", "invalid"), ("[]", "no device found on node")):
        with pytest.raises(eng.IngestError) as ei:
            eng.decode_node_devices("XPU", abi.HS_REG_JSON, anno, None, "n")
        assert msg in str(ei.value)
    devs, _ = eng.decode_node_devices("XPU", abi.HS_REG_JSON,
                                      '[{"id":"xpu-0","index":0,"count":1,"devmem":98304,"devcore":100,"type":"XPU","health":true}]', None, "n")
    assert [(d["ID"], d["DeviceVendor"], d["Type"], d["Devmem"], d["Health"]) for d in devs] == [("xpu-0", "XPU", "XPU", 98304, True)]
    for vendor, msg in (("Ascend910B3", "no device found on node"), ("NVIDIA", "no gpu found on node"), ("Iluvatar", "no gpu found on node")):
        with pytest.raises(eng.IngestError) as ei:
            eng.decode_node_devices(vendor, abi.HS_REG_JSON, "[]", None, "n")
        assert msg in str(ei.value)
