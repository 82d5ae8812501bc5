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
