"""Pin the CPU oracle against the reference's own known-answer tests.

The fixtures under tests/golden/ are the DATA of the reference's table-driven
tests (extracted by tools/extract_golden.py; each file records its source
file:line).  Every case is replayed against oracle/liboracle.so with exact
equality — float32 scores included.  No GPU needed.
"""
import ctypes as C
import json
import os
import struct

import pytest

import oracle_abi as oa
from oracle_abi import abi

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    with open(os.path.join(GOLDEN, name + ".json")) as f:
        return json.load(f)["cases"]


def f32(x):
    return struct.unpack("f", struct.pack("f", float(x)))[0]


def ids(cases):
    return [c.get("name", str(i)) for i, c in enumerate(cases)]


def as_list(x):
    return x if isinstance(x, list) else ([] if not x else x)


def make_device(args):
    # makeDevice(id, numa, Type, used, count, totalmem, usedmem, usedcores, totalcore) — score_test.go:3829-3842
    i, numa, typ, used, count, totalmem, usedmem, usedcores, totalcore = args
    return {"ID": i, "Numa": numa, "Type": typ, "Used": used, "Count": count, "Totalmem": totalmem,
            "Usedmem": usedmem, "Usedcores": usedcores, "Totalcore": totalcore, "Health": True}


def go_device(d):
    if isinstance(d, dict) and d.get("_call") == "makeDevice":
        return make_device(d["args"])
    return d


def go_devices(device_lists):
    out = []
    for dl in device_lists:
        out.append((go_device(dl["Device"]), dl.get("Score", 0)))
    return out


def pod_from(annos=None, reqs=None, n_init=0, uid="pod"):
    annos = annos if isinstance(annos, dict) else {}
    ctrs = []
    for m in reqs or []:
        m = m if isinstance(m, dict) else {}
        ctrs.append([oa.req_from_go(k, v) for k, v in m.items()])
    return {"uid": uid, "name": uid, "annotations": annos, "containers": ctrs, "n_init": n_init}


def node_info_from_go(ni, name="node"):
    """Go device.NodeInfo literal -> hs_node_info dict (pair scores, node annotations)."""
    if not isinstance(ni, dict):
        return None
    devs = []
    for vendor in sorted((ni.get("Devices") or {}).keys()):
        for d in ni["Devices"][vendor]:
            dd = {"id": d.get("ID", ""), "index": d.get("Index", 0), "count": d.get("Count", 0),
                  "devmem": d.get("Devmem", 0), "devcore": d.get("Devcore", 0), "type": d.get("Type", ""),
                  "numa": d.get("Numa", 0), "mode": d.get("Mode", ""), "health": d.get("Health", True),
                  "vendor": vendor, "_scores": (d.get("DevicePairScore") or {}).get("Scores") or {}}
            ci = d.get("CustomInfo")
            if isinstance(ci, dict) and "NetworkID" in ci:
                dd["network_id"] = int(ci["NetworkID"])
            devs.append(dd)
    n = len(devs)
    ps = [[int(devs[i]["_scores"].get(devs[j]["id"], 0)) for j in range(n)] for i in range(n)]
    annos = (((ni.get("Node") or {}).get("ObjectMeta") or {}).get("Annotations")) or {}
    out = {"name": ni.get("ID", name), "devices": devs, "pair_scores": ps if n else None}
    if "hami.io/device-cordon" in annos:
        out["device_cordon"] = annos["hami.io/device-cordon"]
    if "hami-vnpu-core" in annos:
        out["hami_vnpu_core"] = 1 if annos["hami-vnpu-core"] == "true" else 0
    return out


def parse_reason(s):
    # convertReasonToMap, score_test.go:3844-3870
    out = {}
    for r in s.split(", "):
        parts = r.split(" ", 1)
        if len(parts) != 2:
            continue
        cp = parts[0].split("/", 1)
        if len(cp) != 2:
            continue
        try:
            out[parts[1]] = int(cp[0])
        except ValueError:
            continue
    return out


@pytest.fixture(scope="module")
def orc():
    o = oa.Oracle()
    yield o
    o.close()


def run_fit(orc, vendor, devices, request, annos=None, node_info=None, type_override=None):
    devs = [oa.dev_from_go(go_device(d)) for d in devices]
    arr = oa.devs_array(devs)
    r = oa.req_from_go("k", request)
    if type_override is not None:
        r["type"] = type_override
    req = abi.hs_container_request(abi._b(r["key"]), abi._b(r["type"]), r["nums"], r["memreq"], r["mempct"], r["coresreq"])
    p, keep = abi.build_pod(pod_from(annos))
    nip = None
    if node_info is not None:
        ni, keep2 = abi.build_node(node_info)
        nip = C.byref(ni)
    res = abi.hs_filter_result()
    rc = orc.L.oracle_fit(orc.h, abi._b(vendor), arr, len(devs), C.byref(req), C.byref(p), nip, C.byref(res))
    assert rc in (0, 1), orc.L.oracle_last_error(orc.h)
    out = abi.read_result(res)
    got = []
    for lst in out["devices"].values():
        for c in lst:
            got += c
    return rc == 1, got, out["failed"].get("", "")


# ----------------------------------------------------------------------------- scores
@pytest.mark.parametrize("case", load("compute_score"), ids=ids(load("compute_score")))
def test_compute_score(case):
    # policy/gpu_policy_test.go:505-656
    d = oa.dev_from_go(case["device"] or {}) if case.get("device") is not None else None
    if d is None:
        pytest.skip("nil Device is a Go-pointer case; the oracle holds devices by value")
    reqs = [oa.req_from_go(k, v) for k, v in (case.get("requests") or {}).items()]
    arr = (abi.hs_container_request * max(1, len(reqs)))()
    for i, r in enumerate(reqs):
        arr[i] = abi.hs_container_request(abi._b(r["key"]), abi._b(r["type"]), r["nums"], r["memreq"], r["mempct"], r["coresreq"])
    got = oa.lib().oracle_compute_score(C.byref(d), arr, len(reqs), 1, 1, 1)
    assert got == f32(case["expectedScore"])


def test_compute_score_with_resource_weights():
    # policy/gpu_policy_test.go:658-742 (TestComputeScoreWithResourceWeights): gpu-a(1,90,10)/gpu-b(7,10,20) of
    # (10,100,100), request 40 % memory; default -> 16/15, memory-heavy (1,1,3) -> 26/27, core-heavy (1,3,1) -> 34/17.
    L = oa.lib()
    a = oa.dev_from_go({"ID": "gpu-a", "Type": "NVIDIA", "Count": 10, "Used": 1, "Totalcore": 100, "Usedcores": 90,
                        "Totalmem": 100, "Usedmem": 10})
    b = oa.dev_from_go({"ID": "gpu-b", "Type": "NVIDIA", "Count": 10, "Used": 7, "Totalcore": 100, "Usedcores": 10,
                        "Totalmem": 100, "Usedmem": 20})
    req = abi.hs_container_request(b"NVIDIA", b"NVIDIA", 1, 0, 40, 0)
    for w, (wa, wb) in {(1, 1, 1): (16, 15), (1, 1, 3): (26, 27), (1, 3, 1): (34, 17)}.items():
        assert L.oracle_compute_score(C.byref(a), C.byref(req), 1, *w) == f32(wa)
        assert L.oracle_compute_score(C.byref(b), C.byref(req), 1, *w) == f32(wb)


@pytest.mark.parametrize("case", load("device_less"), ids=ids(load("device_less")))
def test_device_less(case):
    # policy/gpu_policy_test.go:125-399
    dl = case["deviceLists"]
    a = oa.dev_from_go(dl[0]["Device"], dl[0].get("Score", 0))
    b = oa.dev_from_go(dl[1]["Device"], dl[1].get("Score", 0))
    got = oa.lib().oracle_less(abi._b(case.get("policy", "")), 1 if case.get("numaBind") else 0, C.byref(a), C.byref(b))
    assert bool(got) == bool(case["expectedLess"])


@pytest.mark.parametrize("case", load("node_default_score"), ids=ids(load("node_default_score")))
def test_node_default_score(case):
    # policy/node_policy_test.go:416-514; device1/device2 are the locals at :417-443
    local = {
        "device1": {"ID": "device1", "Index": 1, "Used": 50, "Count": 100, "Usedmem": 100, "Totalmem": 100,
                    "Totalcore": 100, "Usedcores": 100, "Numa": 0, "Type": "GPU", "Health": True},
        "device2": {"ID": "device2", "Index": 2, "Used": 75, "Count": 150, "Usedmem": 200, "Totalmem": 200,
                    "Totalcore": 200, "Usedcores": 200, "Numa": 1, "Type": "CPU", "Health": False},
    }
    devs = []
    for dl in (case["devices"].get("DeviceLists") or []) if isinstance(case["devices"], dict) else []:
        d = dl["Device"]
        if isinstance(d, dict) and "_id" in d:
            d = local[d["_id"]]
        devs.append(oa.dev_from_go(d))
    got = oa.lib().oracle_node_default_score(oa.devs_array(devs), len(devs))
    assert got == f32(case["wantScore"])


# ----------------------------------------------------------------------------- selectors / parsers
@pytest.mark.parametrize("case", load("check_uuid"), ids=ids(load("check_uuid")))
def test_check_uuid(case):
    # device/devices_test.go:1370-1475
    annos = case.get("annos") if isinstance(case.get("annos"), dict) else {}
    kv = (abi.hs_kv * max(1, len(annos)))(*[abi.hs_kv(abi._b(k), abi._b(v)) for k, v in annos.items()])
    got = oa.lib().oracle_check_uuid(kv, len(annos), abi._b(case["id"]), b"nvidia.com/use-gpuuuid", b"nvidia.com/nouse-gpuuuid")
    assert bool(got) == bool(case["want"])


@pytest.mark.parametrize("case", load("check_type"), ids=ids(load("check_type")))
def test_check_type(case):
    # device/devices_test.go:1477-1560
    annos = case.get("annos") if isinstance(case.get("annos"), dict) else {}
    kv = (abi.hs_kv * max(1, len(annos)))(*[abi.hs_kv(abi._b(k), abi._b(v)) for k, v in annos.items()])
    got = oa.lib().oracle_check_type(kv, len(annos), abi._b(case["cardType"]), b"useKey", b"noUseKey")
    assert bool(got) == bool(case["want"])


@pytest.mark.parametrize("case", load("weights_parse"), ids=ids(load("weights_parse")))
def test_weights_parse(case):
    # util/device_scoring_weights_test.go:65-103
    out = (C.c_int64 * 3)()
    rc = oa.lib().oracle_parse_weights(abi._b(case["value"]), out)
    if case.get("wantErr"):
        assert rc == abi.HS_ERR_WEIGHTS
    else:
        assert rc == 0
        w = case["want"]
        assert (out[0], out[1], out[2]) == (w.get("Slot", 0), w.get("Core", 0), w.get("Memory", 0))


# ----------------------------------------------------------------------------- vendor Fit
def _has_mig(devices):
    return any(go_device(d).get("Mode") == "mig" or go_device(d).get("MigProfiles") for d in devices)


def _fit_table(orc, case, vendor, type_fix=None, annos_fix=None):
    devices = [dict(go_device(d)) for d in case["devices"]]
    annos = dict(case["annos"]) if isinstance(case.get("annos"), dict) else {}
    request = dict(case["request"])
    if type_fix:
        type_fix(case, devices, request, annos)
    node_info = None
    if annos_fix:
        node_info = annos_fix(case, devices)
    fit, got, reason = run_fit(orc, vendor, devices, request, annos, node_info)
    assert fit == bool(case["wantFit"])
    if case["wantFit"]:
        assert len(got) == case["wantLen"]
        for i, want_id in enumerate(case.get("wantDevIDs") or []):
            assert got[i][0] == want_id
    assert reason == case["wantReason"]


@pytest.mark.parametrize("case", load("nvidia_fit"), ids=ids(load("nvidia_fit")))
def test_nvidia_fit(orc, case):
    # nvidia/device_test.go:753-1224
    _fit_table(orc, case, "NVIDIA")


@pytest.mark.parametrize("case", load("hygon_fit"), ids=ids(load("hygon_fit")))
def test_hygon_fit(orc, case):
    # hygon/device_test.go:813-1313
    _fit_table(orc, case, "DCU")


@pytest.mark.parametrize("case", load("cambricon_fit"), ids=ids(load("cambricon_fit")))
def test_cambricon_fit(orc, case):
    # cambricon/device_test.go:805-1315
    _fit_table(orc, case, "MLU")


def _ascend_fix(word):
    def fix(case, devices, request, annos):
        # the reference's loop (ascend/device_test.go:2446-2459) stamps the vendor word in
        if "type mismatch" not in case["name"]:
            request["Type"] = word
        if "user assign use uuid mismatch" in case["name"]:
            annos["hami.io/use-%s-uuid" % word] = "dev-0"
        if "user assign no use uuid match" in case["name"]:
            annos["hami.io/no-use-%s-uuid" % word] = "dev-0"
        for d in devices:
            d["Type"] = word
    return fix


def _ascend_node(word, nets):
    def mk(case, devices):
        annos = case.get("nodeAnnotation") if isinstance(case.get("nodeAnnotation"), dict) else {}
        out = {"name": "node1", "devices": [
            {"id": "dev-%d" % i, "index": 0, "vendor": word, "type": word, "network_id": n} for i, n in enumerate(nets)]}
        if "hami-vnpu-core" in annos:
            out["hami_vnpu_core"] = 1 if annos["hami-vnpu-core"] == "true" else 0
        return out
    return mk


@pytest.mark.parametrize("word", ["Ascend910B3", "Ascend310P", "Ascend910A"])
@pytest.mark.parametrize("case", load("ascend_fit"), ids=ids(load("ascend_fit")))
def test_ascend_fit(orc, case, word):
    # ascend/device_test.go:1706-2520; nodeInfo NetworkIDs 0,1,1 at :2475-2497
    _fit_table(orc, case, word, _ascend_fix(word), _ascend_node(word, [0, 1, 1]))


@pytest.mark.parametrize("case", load("ascend_fit_910c"), ids=ids(load("ascend_fit_910c")))
def test_ascend_fit_910c(orc, case):
    # ascend/device_test.go:2522-2672
    _fit_table(orc, case, "Ascend910C", _ascend_fix("Ascend910C"), _ascend_node("Ascend910C", [0, 0, 0]))


# ----------------------------------------------------------------------------- score.go tables
@pytest.mark.parametrize("case", load("fit_in_certain_device"), ids=ids(load("fit_in_certain_device")))
def test_fit_in_certain_device(orc, case):
    # score_test.go:3291-3829: nvidia Fit over getNodeResources(node, "NVIDIA") with reason histograms
    devices = [go_device(dl["Device"]) for dl in case["args"]["node"]["Devices"]["DeviceLists"]]
    devices = [d for d in devices if "nvidia" in d.get("Type", "").lower()]  # getNodeResources, score.go:43-52
    pod = case["args"].get("pod") or {}
    annos = ((pod.get("ObjectMeta") or {}).get("Annotations")) if isinstance(pod, dict) else {}
    fit, got, reason = run_fit(orc, "NVIDIA", devices, case["args"]["request"], annos or {})
    assert fit == bool(case["want1"])
    want2 = case.get("want2") if isinstance(case.get("want2"), dict) else {}
    want_list = [(d["UUID"], d.get("Usedmem", 0), d.get("Usedcores", 0)) for d in (want2.get("NVIDIA") or [])]
    assert [(g[0], g[2], g[3]) for g in got] == want_list
    want3 = case.get("want3") if isinstance(case.get("want3"), dict) else {}
    assert parse_reason(reason) == want3


@pytest.mark.parametrize("case", load("nvidia_gpu_topology"), ids=ids(load("nvidia_gpu_topology")))
def test_nvidia_gpu_topology(orc, case):
    # score_test.go:4211-4360
    devices = [go_device(dl["Device"]) for dl in case["args"]["node"]["Devices"]["DeviceLists"]]
    pod = case["args"].get("pod") or {}
    annos = ((pod.get("ObjectMeta") or {}).get("Annotations")) or {}
    ni = node_info_from_go(case["args"]["nodeInfo"], "node-1")
    fit, got, reason = run_fit(orc, "NVIDIA", devices, case["args"]["request"], annos, ni)
    assert fit == bool(case["want1"])
    want = [(d["UUID"], d.get("Usedmem", 0), d.get("Usedcores", 0)) for d in case["want2"]["NVIDIA"]]
    assert [(g[0], g[2], g[3]) for g in got] == want
    assert reason == case["want3"]


@pytest.mark.parametrize("case", load("fit_in_devices"), ids=ids(load("fit_in_devices")))
def test_fit_in_devices(orc, case):
    # score_test.go:3872-4046
    devs = [oa.dev_from_go(go_device(dl["Device"]), dl.get("Score", 0)) for dl in case["args"]["node"]["Devices"]["DeviceLists"]]
    reqs = [oa.req_from_go(k, v) for k, v in case["args"]["requests"].items()]
    rarr = (abi.hs_container_request * max(1, len(reqs)))()
    for i, r in enumerate(reqs):
        rarr[i] = abi.hs_container_request(abi._b(r["key"]), abi._b(r["type"]), r["nums"], r["memreq"], r["mempct"], r["coresreq"])
    ctr = abi.hs_container(rarr, len(reqs))
    annos = case["args"].get("annos") if isinstance(case["args"].get("annos"), dict) else {}
    p, keep = abi.build_pod(pod_from(annos))
    res = abi.hs_filter_result()
    policy = case["args"]["node"]["Devices"].get("Policy")
    rc = orc.L.oracle_fit_in_devices(orc.h, oa.devs_array(devs), len(devs), abi._b(policy) if policy else None, -1,
                                     C.byref(ctr), C.byref(p), None, C.byref(res))
    assert rc in (0, 1)
    out = abi.read_result(res)
    assert (rc == 1) == bool(case["want1"])
    reason = list(out["failed"].values())[0] if out["failed"] else ""
    assert reason == case["want2"]


@pytest.mark.parametrize("case", load("calc_score"), ids=ids(load("calc_score")))
def test_calc_score(orc, case):
    # score_test.go:76-3289: per node, expected devices per container AND the exact float32 node score
    args = case["args"]
    task = args.get("task") or {}
    annos = ((task.get("ObjectMeta") or {}).get("Annotations")) or {}
    n_init = len(((task.get("Spec") or {}).get("InitContainers")) or [])
    pod = pod_from(annos, args["nums"], n_init, uid="task")
    want_nodes = {n["NodeID"]: n for n in ((case["wants"].get("want") or {}).get("NodeList") or [])}
    want_failed = case["wants"].get("failedNodes") if isinstance(case["wants"].get("failedNodes"), dict) else {}
    failed = {}
    got_nodes = {}
    for node_id, nu in args["nodes"].items():
        dl = nu["Devices"]
        devs = [oa.dev_from_go(go_device(x["Device"]), x.get("Score", 0)) for x in dl.get("DeviceLists") or []]
        p, keep = abi.build_pod(pod)
        res = abi.hs_filter_result()
        rc = orc.L.oracle_score_node(orc.h, abi._b(node_id), oa.devs_array(devs), len(devs),
                                     abi._b(dl.get("Policy", "")), 1 if dl.get("NumaBind") else 0,
                                     C.byref(p), None, C.byref(res))
        assert rc in (abi.HS_FIT, abi.HS_NOFIT), orc.L.oracle_last_error(orc.h)
        out = abi.read_result(res)
        if rc == abi.HS_NOFIT:
            failed[node_id] = out["failed"][node_id]
        else:
            got_nodes[node_id] = out
    if failed:
        # the reference only compares failedNodes when any node failed (score_test.go:3275-3278)
        assert failed == want_failed
        return
    for node_id, out in got_nodes.items():
        want = want_nodes[node_id]
        want_devs = {}
        for typ, ctrs in (want.get("Devices") or {}).items():
            want_devs[typ] = [[(d["UUID"], d.get("Idx", 0), d.get("Usedmem", 0), d.get("Usedcores", 0)) for d in (c or [])]
                              for c in ctrs]
        assert out["devices"] == want_devs
        assert out["score"] == f32(want.get("Score", 0))
