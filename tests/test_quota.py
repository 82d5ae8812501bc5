"""Namespace ResourceQuota on the Fit path: device.QuotaManager (pkg/device/quota.go) and the vendors'
fitQuota (nvidia/device.go:666-692, cambricon/device.go:351-376).

CPU part: the oracle against the reference's vectors — tests/golden/nvidia_fit_quota.json (TestFitQuota of
nvidia/device_test.go, 11 cases, extracted) and the scalar cases of pkg/device/quota_test.go (TestFitQuota,
TestFitQuotaExplicitZeroBlocks, TestAddUsageAndRmUsage; transcribed below with their line numbers).
GPU part (test_gpu_quota.py): the engine against the oracle on streams in quota'd namespaces.
"""
import ctypes as C
import json
import os

import pytest

import oracle_abi as oa
from oracle_abi import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MEM, CORE = 0, 1


def _lib():
    L = oa.lib()
    L.oracle_quota_set.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32, C.c_int64]
    L.oracle_quota_del.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32]
    L.oracle_quota_get.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32, C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                   C.POINTER(C.c_int32)]
    L.oracle_quota_add_usage.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32, C.c_int64]
    L.oracle_fit_quota_scalar.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int64, C.c_int32, C.c_int64]
    L.oracle_fit_quota.argtypes = [C.c_void_p, C.c_char_p, C.c_uint32, C.POINTER(abi.hs_alloc_entry), C.c_uint32,
                                   C.POINTER(abi.hs_alloc_entry), C.c_uint32, C.c_int32, C.c_char_p, C.c_char_p, C.c_int64, C.c_int64]
    return L


def quota_get(L, o, ns, vendor, which):
    lim, used, st = C.c_int64(), C.c_int64(), C.c_int32()
    rc = L.oracle_quota_get(o.h, ns.encode(), vendor.encode(), which, C.byref(lim), C.byref(used), C.byref(st))
    return None if rc != 0 else (lim.value, used.value, bool(st.value))


def _entries(rows):
    arr = (abi.hs_alloc_entry * max(len(rows), 1))()
    for i, (ctr, key, d) in enumerate(rows):
        arr[i] = abi.hs_alloc_entry(ctr, key.encode(), d["UUID"].encode(), 0, d["Usedmem"], d["Usedcores"])
    return arr, len(rows)


def test_fit_quota_golden_nvidia():
    L = _lib()
    doc = json.load(open(os.path.join(ROOT, "tests", "golden", "nvidia_fit_quota.json")))
    assert len(doc["cases"]) == 11
    o = oa.Oracle("binpack", "spread")
    assert L.oracle_quota_set(o.h, b"default", b"NVIDIA", MEM, 2048) == 0      # limits.nvidia.com/gpumem: 2048 (device_test.go:1324-1339)
    for c in doc["cases"]:
        n_init = c["pod"]["args"][0]
        tmp = [(0, "NVIDIA", d) for d in (c["tmpDevs"].get("NVIDIA", []) if isinstance(c["tmpDevs"], dict) else [])]
        alloc = []
        if c["allocated"]:
            for key, ctrs in c["allocated"].items():
                key = "NVIDIA" if key == "NvidiaGPUDevice" else key
                for ci, ctr in enumerate(ctrs):
                    alloc += [(ci, key, d) for d in ctr]
        ta, nt = _entries(tmp)
        aa, na = _entries(alloc)
        got = L.oracle_fit_quota(o.h, c["ns"].encode(), n_init, ta, nt, aa, na, 1 if c["allocated"] is not None else 0, b"NVIDIA",
                                 c["devUUID"].encode(), c["memreq"], c["coresreq"])
        assert bool(got) == c["expectedResult"], c["name"]
    o.close()


def test_fit_quota_scalar_cases():
    """pkg/device/quota_test.go:121-162 (TestFitQuota) and :164-182 (TestFitQuotaExplicitZeroBlocks)."""
    L = _lib()
    o = oa.Oracle("binpack", "spread")
    ns = b"testns"
    L.oracle_quota_set(o.h, ns, b"NVIDIA", MEM, 2000)
    L.oracle_quota_set(o.h, ns, b"NVIDIA", CORE, 400)
    L.oracle_quota_add_usage(o.h, ns, b"NVIDIA", MEM, 1000)
    L.oracle_quota_add_usage(o.h, ns, b"NVIDIA", CORE, 200)
    fit = lambda n, mem, factor, core, dev=b"NVIDIA": bool(L.oracle_fit_quota_scalar(o.h, n, dev, mem, factor, core))   # noqa: E731
    assert fit(ns, 500, 1, 100)
    assert not fit(ns, 1500, 1, 100)
    assert not fit(ns, 500, 1, 300)
    assert fit(ns, 1000, 2, 100)
    assert not fit(ns, 5000, 2, 100)
    assert fit(b"otherns", 1500, 1, 100)
    assert fit(ns, 1000, 1, 100, b"unknown-device")
    L.oracle_quota_set(o.h, b"team-zero", b"NVIDIA", MEM, 0)                     # an explicit zero blocks
    assert not fit(b"team-zero", 500, 1, 0)
    L.oracle_quota_del(o.h, b"team-zero", b"NVIDIA", MEM)                        # DelQuota: limit gone, entry stays (quota.go:271-289)
    assert fit(b"team-zero", 500, 1, 0)
    assert quota_get(L, o, "team-zero", "NVIDIA", MEM) == (0, 0, False)
    o.close()


def test_add_and_rm_usage_clamps_at_zero():
    """pkg/device/quota_test.go:184-231 (TestAddUsageAndRmUsage)."""
    L = _lib()
    o = oa.Oracle("binpack", "spread")
    ns = b"testns"
    L.oracle_quota_add_usage(o.h, ns, b"NVIDIA", MEM, 1000)
    L.oracle_quota_add_usage(o.h, ns, b"NVIDIA", CORE, 100)
    assert quota_get(L, o, "testns", "NVIDIA", MEM) == (0, 1000, False)          # auto-created by usage tracking: no limit set
    assert quota_get(L, o, "testns", "NVIDIA", CORE) == (0, 100, False)
    L.oracle_quota_add_usage(o.h, ns, b"NVIDIA", MEM, -1000)
    L.oracle_quota_add_usage(o.h, ns, b"NVIDIA", CORE, -100)
    L.oracle_quota_add_usage(o.h, ns, b"NVIDIA", MEM, -1000)                     # a second removal must not go negative
    assert quota_get(L, o, "testns", "NVIDIA", MEM) == (0, 0, False)
    assert quota_get(L, o, "testns", "NVIDIA", CORE) == (0, 0, False)
    o.close()


def test_filter_tracks_quota_usage_in_the_oracle():
    import k8s_vgpu_scheduler_b200.gen as gen
    L = _lib()
    cfg, nodes, ops = gen.mixed_features(n_nodes=12, n_pods=120)
    o = oa.Oracle(cfg["node_policy"], cfg["gpu_policy"])
    for n in nodes:
        o.node_upsert(n)
    L.oracle_quota_set(o.h, b"default", b"NVIDIA", MEM, 400000)
    fits = nofit_quota = 0
    for kind, arg in ops:
        if kind == "filter":
            r = o.filter(arg)
            fits += r["status"] == abi.HS_FIT
            nofit_quota += r["status"] == abi.HS_NOFIT and any("ResourceQuotaNotFit" in v for v in r["failed"].values())
        else:
            o.pod_delete(arg)
        lim, used, st = quota_get(L, o, "default", "NVIDIA", MEM)
        assert 0 <= used <= 400000 and st
    assert fits > 5 and nofit_quota > 5
    o.close()


@pytest.mark.parametrize("nums,cores,mem,q_cores,q_mem,want", [
    # pkg/scheduler/scheduler_test.go:2066-2370 (Test_ResourceQuota): node1 holds four cards of 2000 / 8000 / 4000 / 6000 MiB
    (2, 100, 2000, 200, 4000, None),                                                             # "multi device Resourcequota pass"
    (2, 60, 3000, 200, 4000, "1/4 AllocatedCardsInsufficientRequest, 3/4 ResourceQuotaNotFit"),  # the quota gate sits before the memory gate
    (1, 0, 0, 100, 2000, None),                                                                  # unspecified -> a whole card; only the 2000 MiB one fits the quota
    (1, 0, 0, 100, 1500, "4/4 ResourceQuotaNotFit"),
], ids=["multi-pass", "multi-deny", "whole-card-pass", "whole-card-deny"])
def test_resource_quota_filter_vectors(nums, cores, mem, q_cores, q_mem, want):
    L = _lib()
    o = oa.Oracle("binpack", "spread")           # NewScheduler()'s defaults: node policy binpack, GPU policy spread
    o.node_upsert({"name": "node1", "devices": [
        {"id": "device%d" % (i + 1), "index": i % 2, "count": 10, "devmem": m, "devcore": 100, "type": "NVIDIA", "numa": 0, "mode": "hami",
         "health": True, "vendor": "NVIDIA"} for i, m in enumerate((2000, 8000, 4000, 6000))]})
    L.oracle_quota_set(o.h, b"default", b"NVIDIA", CORE, q_cores)
    L.oracle_quota_set(o.h, b"default", b"NVIDIA", MEM, q_mem)
    r = o.filter({"uid": "t", "ns": "default", "name": "t", "annotations": {}, "containers": [[
        {"key": "NVIDIA", "type": "NVIDIA", "nums": nums, "memreq": mem, "mempct": 100 if mem == 0 else 101, "coresreq": cores}]]})
    if want is None:
        assert r["status"] == abi.HS_FIT and r["node"] == "node1", r.get("failed")
        if nums == 1:
            assert [d[0] for d in r["devices"]["NVIDIA"][0]] == ["device1"]
    else:
        assert r["status"] == abi.HS_NOFIT and r["failed"] == {"node1": want}
    o.close()


def test_cambricon_whole_card_request_meets_the_memory_quota():
    # cambricon/device_test.go:1317-1378 (TestDevices_Fit_ResourceQuotaWholeCardRequest): a request that names no memory is an
    # implicit whole-card ask; its memory is resolved from the card BEFORE fitQuota, so a 40960 MiB card cannot pass a quota
    # of limits.cambricon.com/mlu.smlu.vmemory: 50 (x MemoryFactor 256 = 12800 MiB)
    L = _lib()
    o = oa.Oracle("binpack", "spread")
    o.node_upsert({"name": "node", "devices": [
        {"id": "dev-0", "index": 0, "count": 100, "devmem": 40960, "devcore": 100, "type": "MLU", "numa": 0, "mode": "", "health": True, "vendor": "MLU"}]})
    L.oracle_quota_set(o.h, b"cambricon-quota-test-ns", b"MLU", MEM, 50)
    pod = {"uid": "w", "ns": "cambricon-quota-test-ns", "name": "w", "annotations": {}, "containers": [[
        {"key": "MLU", "type": "MLU", "nums": 1, "memreq": 0, "mempct": 100, "coresreq": 10}]]}
    r = o.filter(pod)
    assert r["status"] == abi.HS_NOFIT and r["failed"] == {"node": "1/1 ResourceQuotaNotFit"}, r
    r = o.filter(dict(pod, ns="another-namespace"))           # the quota is the namespace's
    assert r["status"] == abi.HS_FIT
    o.close()
