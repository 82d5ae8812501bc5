"""enflame shared vGCU ("Enflame"): the oracle against the reference's own tests (pkg/device/enflame/
device_test.go — not table-driven, so the vectors are transcribed here with their file:line), then GPU parity."""
import json
import random

import pytest

import oracle_abi as oa
from oracle_abi import abi
from test_oracle_golden import run_fit

PROFILES = {"1g.6gb": "0", "3g.20gb": "1", "6g.40gb": "2"}


@pytest.fixture(scope="module")
def orc():
    o = oa.Oracle()
    yield o
    o.close()


def drs(used=0, count=6, health=True, profiles=None, totalcore=0, usedmem=0, idx=0, max_slice=0):
    # device_test.go:162-181
    ci = {"minor": str(idx), "index": str(idx), "profiles": dict(PROFILES if profiles is None else profiles)}
    if max_slice:
        ci["maxSlice"] = max_slice
    return {"ID": "node-a-enflame-drs-%d" % idx, "Index": idx, "Count": count, "Used": used, "Totalmem": 40960, "Usedmem": usedmem,
            "Totalcore": totalcore, "Type": "Enflame", "Health": health, "CustomInfo": ci}


def fit(orc, devices, request, annos=None):
    return run_fit(orc, "Enflame", devices, dict(request, Type="Enflame"), annos or {})


def test_fit_select_profile_by_request(orc):
    # device_test.go:160-195 (TestFit_SelectProfileByRequest): a direct request for 3 slices
    ok, got, reason = fit(orc, [drs()], {"Nums": 1, "Memreq": 3})
    assert ok and reason == "" and len(got) == 1
    assert got[0][2:] == (20480, 50, "3g.20gb", 3, 0)


def test_fit_select_profile_by_memory_core_request(orc):
    # device_test.go:197-234: 20480 MiB -> 20 GB, 40 % cores: the 3g.20gb profile (50 %)
    ok, got, reason = fit(orc, [drs()], {"Nums": 1, "Memreq": 20480, "Coresreq": 40, "MemPercentagereq": -1})
    assert ok and reason == "" and got[0][2:] == (20480, 50, "3g.20gb", 3, 0)
    # a value no larger than twice the biggest profile is read as GB (normalizeMemoryRequestToGB, device.go:645-657)
    ok, got, _ = fit(orc, [drs()], {"Nums": 1, "Memreq": 7, "MemPercentagereq": -1})
    assert ok and got[0][4] == "3g.20gb"
    ok, got, _ = fit(orc, [drs()], {"Nums": 1, "Memreq": 81, "MemPercentagereq": -1})      # 81 > 80: MiB, rounds up to 1 GB
    assert ok and got[0][4] == "1g.6gb"
    ok, _, reason = fit(orc, [drs()], {"Nums": 1, "Memreq": 0, "Coresreq": 101, "MemPercentagereq": -1})
    assert not ok and reason == "1/1 ModeNotFit"


def test_fit_profile_not_found(orc):
    # device_test.go:255-282
    ok, _, reason = fit(orc, [drs(health=False, profiles={"1g.6gb": "0"})], {"Nums": 1, "Memreq": 3})
    assert not ok and "ModeNotFit" in reason


def test_fit_mutex_rejects_used_device(orc):
    # device_test.go:284-315
    ok, _, reason = fit(orc, [drs(used=1, count=2)], {"Nums": 1, "Memreq": 3}, {"hami.io/gpu-scheduler-policy": "mutex"})
    assert not ok and reason == "1/1 ExclusiveDeviceAllocateConflict"


def test_fit_unhealthy_device_rejected(orc):
    # device_test.go:317-347
    ok, _, reason = fit(orc, [drs(health=False)], {"Nums": 1, "Memreq": 3})
    assert not ok and "CardNotHealth" in reason


def test_fit_oversized_profile_rejected(orc):
    # device_test.go:459-: a profile whose memory does not fit int32 MiB is ModeNotFit
    ok, _, reason = fit(orc, [drs(profiles={"3g.9999999gb": "1"})], {"Nums": 1, "Memreq": 3})
    assert not ok and "ModeNotFit" in reason


def test_core_share_and_gates(orc):
    # ceil(size * 100 / maxSlice) (device.go:594); the core gate only applies when Totalcore > 0 (:436)
    ok, got, _ = fit(orc, [drs(totalcore=100)], {"Nums": 1, "Memreq": 1})
    assert ok and got[0][2:4] == (6144, 17)
    ok, got, _ = fit(orc, [drs(totalcore=100, max_slice=12)], {"Nums": 1, "Memreq": 6})
    assert ok and got[0][3] == 50
    d = drs(totalcore=100)
    d["Usedcores"] = 60
    ok, _, reason = fit(orc, [d], {"Nums": 1, "Memreq": 3})
    assert not ok and reason == "1/1 CardInsufficientCore"
    ok, _, reason = fit(orc, [drs(usedmem=30000)], {"Nums": 1, "Memreq": 3})
    assert not ok and reason == "1/1 CardInsufficientMemory"
    ok, _, reason = fit(orc, [drs(used=6)], {"Nums": 1, "Memreq": 3})
    assert not ok and reason == "1/1 CardTimeSlicingExhausted"
    ok, _, reason = fit(orc, [drs()], {"Nums": 1, "Memreq": 3}, {"enflame.com/use-gpuuuid": "other"})
    assert not ok and reason == "1/1 CardUuidMismatch"


def vgcu_node(name, cards=4, capacity=6, profiles=None, unhealthy=()):
    profiles = PROFILES if profiles is None else profiles
    devs = []
    for i in range(cards):
        devs.append({"id": "%s-enflame-drs-%d" % (name, i), "index": i, "count": capacity, "devmem": 40960, "devcore": 100, "type": "Enflame",
                     "numa": 0, "mode": "", "health": i not in unhealthy, "vendor": "Enflame",
                     "enflame": {"minor": str(i), "index": str(i), "max_slice": 6, "profiles": profiles}})
    return {"name": name, "devices": devs}


def vgcu_pod(uid, mode, value, cores=0, annos=None, names=None, extra=None):
    req = {"key": "Enflame", "type": "Enflame", "nums": 1, "memreq": value, "mempct": 0 if mode == "direct" else -1, "coresreq": cores}
    ctrs = [[req]] + (extra or [])
    return {"uid": uid, "ns": "default", "name": uid, "annotations": annos or {}, "containers": ctrs, "n_init": 0,
            "container_names": names or ["main"] + ["side-%d" % i for i in range(len(ctrs) - 1)]}


def test_filter_annotations_through_the_oracle():
    # PatchAnnotations, device.go:254-300 (TestPatchAnnotations_DRSFields, device_test.go:349-396)
    o = oa.Oracle()
    o.node_upsert(vgcu_node("node-a", cards=1))
    r = o.filter(vgcu_pod("p1", "direct", 3, names=["pod-gcu-example1"]))
    assert r["status"] == 0
    a = r["annotations"]
    assert a["hami.io/enflame-vgpu-devices-allocated"] == "node-a-enflame-drs-0,Enflame,20480,50:;"
    assert a["enflame.com/gcu-assigned"] == "false" and a["enflame.com/gcu-assigned-time"] == ""
    assert a["enflame.com/gcu-assigned-index"] == "0" and a["enflame.com/gcu-assigned-id"] == "0" and a["enflame.com/gcu-assigned-minor"] == "0"
    assert a["enflame.com/gcu-request-size"] == "3"
    assert json.loads(a["assigned-containers"]) == {"pod-gcu-example1": {"allocated": False, "request": 3, "profileID": "1", "profileName": "3g.20gb"}}
    assert "hami.io/enflame-vgpu-devices-to-allocate" not in a
    # AddResourceUsage counts the slices inside a Filter (:372-380), the cached pod counts one slot per entry (scheduler.go:782)
    assert o.snapshot("node-a") == [(1, 20480, 50)]
    o.close()


def _cluster(seed, n_nodes=30, n_pods=700):
    rng = random.Random(seed)
    nodes = []
    for i in range(n_nodes):
        r = rng.random()
        if r < 0.6:
            nodes.append(vgcu_node("gcu-%02d" % i, cards=rng.choice([1, 2, 4, 8]), capacity=rng.choice([4, 6, 8]),
                                   unhealthy=[0] if rng.random() < 0.1 else ()))
        elif r < 0.8:         # another profile table (equal slice counts with different memory, a name that does not parse)
            nodes.append(vgcu_node("gcu-%02d" % i, cards=rng.choice([2, 4]),
                                   profiles={"1g.6gb": "0", "2g.12gb": "3", "2g.16gb": "4", "6g.40gb": "2", "whole": "9"}))
        else:
            nodes.append({"name": "gpu-%02d" % i, "devices": [
                {"id": "gpu-%02d-%d" % (i, j), "index": j, "count": 10, "devmem": 81920, "devcore": 100, "type": "NVIDIA-H100",
                 "numa": 0, "mode": "hami-core", "health": True, "vendor": "NVIDIA"} for j in range(4)]})
    ops, live = [], []
    for p in range(n_pods):
        if live and rng.random() < 0.3:
            ops.append(("delete", live.pop(rng.randrange(len(live)))))
            continue
        uid = "pod-%04d" % p
        annos = {}
        if rng.random() < 0.1:
            annos["hami.io/gpu-scheduler-policy"] = "mutex"
        if rng.random() < 0.08:
            annos["enflame.com/use-gpuuuid"] = ",".join("gcu-%02d-enflame-drs-%d" % (rng.randrange(n_nodes), rng.randrange(4)) for _ in range(6))
        r = rng.random()
        if r < 0.12:
            pod = {"uid": uid, "ns": "default", "name": uid, "annotations": annos, "n_init": 0,
                   "containers": [[{"key": "NVIDIA", "type": "NVIDIA", "nums": 1, "memreq": 8000, "mempct": 101, "coresreq": 10}]]}
        elif r < 0.55:
            pod = vgcu_pod(uid, "direct", rng.choice([1, 2, 3, 3, 6, 5]), annos=annos)
        else:
            extra = []
            if rng.random() < 0.25:
                extra = [[{"key": "Enflame", "type": "Enflame", "nums": 1, "memreq": rng.choice([1, 3]), "mempct": 0, "coresreq": 0}]]
            pod = vgcu_pod(uid, "spec", rng.choice([6, 12, 6144, 16000, 20480, 40960, 50000, 0]), cores=rng.choice([0, 0, 20, 40, 100]),
                           annos=annos, extra=extra)
            if extra and rng.random() < 0.4:
                pod["n_init"] = 1
        ops.append(("filter", pod))
        live.append(uid)
    return {"node_policy": rng.choice(["binpack", "spread"]), "gpu_policy": rng.choice(["binpack", "spread"])}, nodes, ops


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [5, 9])
def test_enflame_vgcu_gpu_parity(seed):
    import parity
    cfg, nodes, ops = _cluster(seed)
    ref, ref_snap = parity.run_oracle(cfg, nodes, ops)
    got, snap = parity.run_engine(cfg, nodes, ops, chunk=175)
    assert parity.compare(ops, ref, got, ref_snap, snap) == []
    fits = [r for (k, a), r in zip(ops, ref) if k == "filter" and r["status"] == abi.HS_FIT and any(r["devices"].get("Enflame", []))]
    nofit = [r for (k, a), r in zip(ops, ref) if k == "filter" and r["status"] == abi.HS_NOFIT]
    assert len(fits) > 150 and len(nofit) > 15
    assert len({d[4] for r in fits for c in r["devices"]["Enflame"] for d in c}) >= 4          # several profiles were handed out
    assert any("ModeNotFit" in v for r in nofit for v in r["failed"].values())
    assert all("assigned-containers" in r["annotations"] for r in fits)


def test_get_node_devices_from_the_drs_annotation():
    # device_test.go:33-52 (TestGetNodeDevices_DRSAnnotation) + the error returns of device.go:198-252
    import k8s_vgpu_scheduler_b200.engine as eng
    anno = '{"devices":[{"index":"0","minor":"0","capacity":6}],"profiles":{"1g.6gb":"0","3g.20gb":"1","6g.40gb":"2"}}'
    devs, _ = eng.decode_node_devices("Enflame", abi.HS_REG_ENFLAME_DRS, anno, None, "node-a")
    assert len(devs) == 1
    d = devs[0]
    assert (d["ID"], d["Type"], d["Devmem"], d["Count"], d["Devcore"], d["Health"]) == ("node-a-enflame-drs-0", "Enflame", 40960, 6, 100, True)
    assert d["Enflame"] == {"minor": "0", "index": "0", "max_slice": 6, "profiles": PROFILES}
    # index that is not a number falls back to the position, a blank minor to the index, capacity may be a string
    anno = '{"devices":[{"index":"x","minor":" ","capacity":"4"},{"index":"7","minor":"m7","capacity":2.0}],"profiles":{"2g.16gb":"a"}}'
    devs, _ = eng.decode_node_devices("Enflame", abi.HS_REG_ENFLAME_DRS, anno, None, "n")
    assert [(x["ID"], x["Index"], x["Count"], x["Devmem"], x["Enflame"]["minor"]) for x in devs] == [
        ("n-enflame-drs-0", 0, 4, 16384, "0"), ("n-enflame-drs-7", 7, 2, 16384, "m7")]
    for bad, msg in (('{"devices":[{"index":"0","capacity":6}],"profiles":{"whole":"0"}}', "no valid drs profiles found on node n"),
                     ('{"devices":[{"index":"0","capacity":0}],"profiles":{"1g.6gb":"0"}}', "invalid drs capacity on node n"),
                     ('{"devices":[],"profiles":{"1g.6gb":"0"}}', "no drs devices found on node n"),
                     ('{"devices":', "failed to parse enflame.com/gcu-drs-capacity")):
        with pytest.raises(eng.IngestError) as ei:
            eng.decode_node_devices("Enflame", abi.HS_REG_ENFLAME_DRS, bad, None, "n")
        assert msg in str(ei.value)


@pytest.mark.gpu
def test_enflame_nodes_registered_from_the_annotation_decide_identically():
    import parity
    import k8s_vgpu_scheduler_b200.engine as eng
    cfg, nodes, ops = _cluster(13, n_nodes=20, n_pods=400)
    for n in nodes:          # the annotation carries no health: make the upsert twin healthy too
        for d in n["devices"]:
            d["health"] = True
    ref, _ = parity.run_oracle(cfg, nodes, ops)
    e = eng.Engine(cfg.get("node_policy"), cfg.get("gpu_policy"))
    n_reg = 0
    for n in nodes:
        if n["devices"][0]["vendor"] != "Enflame":
            e.node_upsert(n)
            continue
        profiles = n["devices"][0]["enflame"]["profiles"]
        anno = json.dumps({"devices": [{"index": str(d["index"]), "minor": str(d["index"]), "capacity": d["count"]} for d in n["devices"]],
                           "profiles": profiles})
        assert e.node_register(n["name"], [("Enflame", abi.HS_REG_ENFLAME_DRS, anno, None)]) == len(n["devices"]), e.error()
        n_reg += 1
    assert n_reg > 8
    e.run_batch(eng.PackedOps(ops))
    got = [e.batch_get(i) for i in range(len(ops))]
    e.close()
    assert parity.compare(ops, ref, got) == []


def test_drs_capacity_parsing_vectors():
    # device_test.go:420-437 (TestParseDRSCapacity_RejectsNaNAndInf) + parseDRSCapacity / clampToInt32 (device.go:643-675) as
    # they show through GetNodeDevices: a JSON number is a float64 (refused above MaxInt32, truncated toward zero below), a
    # string goes through Atoi after TrimSpace and is clamped, anything else — and a capacity that is not positive — is refused
    import k8s_vgpu_scheduler_b200.engine as eng

    def count(cap):
        anno = '{"devices":[{"index":"0","minor":"0","capacity":%s}],"profiles":{"1g.6gb":"0"}}' % cap
        try:
            return eng.decode_node_devices("Enflame", abi.HS_REG_ENFLAME_DRS, anno, None, "n")[0][0]["Count"]
        except eng.IngestError as e:
            assert "invalid drs capacity on node n" in str(e)
            return None
    assert [count(c) for c in ("42", "42.0", "42.9", "2147483647.0", '"2147483648"', '" 7 "')] == [42, 42, 42, 2 ** 31 - 1, 2 ** 31 - 1, 7]
    assert [count(c) for c in ("2147483648", "1e300", "3e9", '"7.5"', "-3", '"-9999999999"', "true", "null", "0")] == [None] * 9
