"""Dynamic MIG: the oracle against the reference's own MIG tests.

The reference's MIG tests are not table-driven, so their vectors are transcribed by hand here, each with the Go
file:line it comes from: pkg/device/nvidia/mig_topology_test.go, mig_capability_test.go,
pkg/scheduler/mig_allocation_test.go (the table case "mode is mig", score_test.go:3613-3669, is in
tests/golden/fit_in_certain_device.json and runs in test_oracle_golden.py).
"""
import ctypes as C

import pytest

import oracle_abi as oa
from oracle_abi import abi
from test_oracle_golden import pod_from, run_fit


def a100(with_meta=True):
    # mig_capability_test.go:27-33 (a100MigProfiles) / mig_topology_test.go:26-30
    p = [
        {"Name": "1g.5gb", "MemoryMB": 5120, "Core": 14, "SliceCount": 1, "Placements": [{"Start": i, "Size": 1} for i in range(7)]},
        {"Name": "2g.10gb", "MemoryMB": 10240, "Core": 28, "SliceCount": 2, "Placements": [{"Start": 0, "Size": 2}, {"Start": 2, "Size": 2}, {"Start": 4, "Size": 2}]},
        {"Name": "3g.20gb", "MemoryMB": 20480, "Core": 42, "SliceCount": 3, "Placements": [{"Start": 0, "Size": 4}, {"Start": 4, "Size": 4}]},
    ]
    if not with_meta:
        for q in p:
            q["MemoryMB"] = q["Core"] = q["SliceCount"] = 0
    return p


def profiles_c(go_profiles, keep):
    return abi.build_mig_profiles(oa.mig_profiles_from_go(go_profiles), keep)


def can_place(go_profiles, occupied, requested):
    keep = []
    arr, n = profiles_c(go_profiles, keep)
    occ = oa.placements_array(occupied)
    req = (C.c_char_p * max(1, len(requested)))(*[r.encode() for r in requested])
    return oa.lib().oracle_mig_can_place(arr, n, occ, len(occupied), req, len(requested)) == 1


def select_placement(go_profiles, occupied, name):
    keep = []
    arr, n = profiles_c(go_profiles, keep)
    out = abi.hs_mig_placement()
    ok = oa.lib().oracle_mig_select_placement(arr, n, oa.placements_array(occupied), len(occupied), name.encode(), C.byref(out))
    return (out.start, out.size) if ok else None


def select_candidate(go_profiles, occupied, memory):
    keep = []
    arr, n = profiles_c(go_profiles, keep)
    out = abi.hs_mig_placement()
    idx = C.c_int32(-1)
    ok = oa.lib().oracle_mig_select_candidate(arr, n, oa.placements_array(occupied), len(occupied), memory, C.byref(idx), C.byref(out))
    return (go_profiles[idx.value]["Name"], (out.start, out.size)) if ok else None


def test_can_place_mig_profiles_a100():
    # mig_topology_test.go:25-44
    allowed = a100(with_meta=False)
    assert can_place(allowed, [], ["3g.20gb", "2g.10gb", "1g.5gb", "1g.5gb"])
    assert can_place(allowed, [(4, 4)], ["2g.10gb", "1g.5gb", "1g.5gb"])
    assert not can_place(allowed, [(0, 4), (4, 4)], ["1g.5gb"])
    assert not can_place(allowed, [(0, 2), (2, 2), (4, 2)], ["2g.10gb"])


def test_select_mig_placement_uses_balanced_packing():
    # mig_topology_test.go:46-69
    allowed = a100(with_meta=False)
    occupied = []
    for profile, start in (("3g.20gb", 4), ("2g.10gb", 0), ("1g.5gb", 3), ("1g.5gb", 2)):
        got = select_placement(allowed, occupied, profile)
        assert got is not None and got[0] == start
        occupied.append(got)


def test_add_resource_usage_uses_reported_profile_and_placement(orc):
    # mig_capability_test.go:35-51: Usedmem 10000 -> profile 2g.10gb (10240 MB, 28 cores) at {0, 2}; through
    # fitInDevices so that AddResourceUsage itself runs
    assert select_candidate(a100(), [], 10000) == ("2g.10gb", (0, 2))
    dev = {"ID": "GPU-a", "Type": "NVIDIA", "Mode": "mig", "Health": True, "Count": 7, "Totalmem": 40960, "Totalcore": 100,
           "MigProfiles": a100()}
    fit, devices, _ = fit_in_devices(orc, [dev], {"Nums": 1, "Type": "NVIDIA", "Memreq": 10000})
    assert fit and devices["NVIDIA"] == [[("GPU-a", 0, 10240, 28, "2g.10gb", 0, 2)]]


def test_custom_filter_uses_reported_placement_capacity():
    # mig_capability_test.go:53-69
    occ = [(0, 2), (2, 2), (4, 2)]
    assert select_candidate(a100(), occ, 10000) is None
    assert select_candidate(a100(), occ, 5000) == ("1g.5gb", (6, 1))


def test_custom_filter_ignores_queued_allocations_for_other_gpus(orc):
    # mig_capability_test.go:71-89: a pick queued on GPU-b does not consume GPU-a's last placement, one on GPU-a does.
    # Driven through Fit: two GPUs whose six 2g slices are taken; a 2-device request of 5000 MB finds one placement
    # on each GPU (the second visit of a GPU sees the first pick queued on it and fails)
    used = [{"Profile": "2g.10gb", "Placement": {"Start": s, "Size": 2}} for s in (0, 2, 4)]
    gpu = lambda u: {"ID": u, "Type": "NVIDIA", "Mode": "mig", "Health": True, "Count": 7, "Used": 3, "Totalmem": 40960,  # noqa: E731
                     "Usedmem": 30720, "Totalcore": 100, "MigProfiles": a100(), "MigAllocationsInUse": used}
    fit, got, reason = run_fit(orc, "NVIDIA", [gpu("GPU-a"), gpu("GPU-b")], {"Nums": 2, "Type": "NVIDIA", "Memreq": 5000})
    assert fit and [g[0] for g in got] == ["GPU-b", "GPU-a"]
    fit, got, reason = run_fit(orc, "NVIDIA", [gpu("GPU-a")], {"Nums": 2, "Type": "NVIDIA", "Memreq": 5000})
    assert not fit and "1/1 CardMigTopologyInfeasible" in reason


def test_fit_uses_effective_percentage_memory_for_mig_placement(orc):
    # mig_capability_test.go:91-110: 25 % of 40960 = 10240 MB needs a 2g slice, only a 1g placement remains
    dev = {"ID": "GPU-a", "Type": "NVIDIA", "Mode": "mig", "Health": True, "Count": 7, "Used": 3, "Totalmem": 40960,
           "Usedmem": 30720, "Totalcore": 100, "MigProfiles": a100(),
           "MigAllocationsInUse": [{"Profile": "2g.10gb", "Placement": {"Start": s, "Size": 2}} for s in (0, 2, 4)]}
    fit, _, _ = run_fit(orc, "NVIDIA", [dev], {"Nums": 1, "Type": "NVIDIA", "MemPercentagereq": 25})
    assert not fit


# ---- pkg/scheduler/mig_allocation_test.go ------------------------------------------------------------------
ONE_1G = [{"name": "1g.5gb", "memory_mb": 5120, "core": 14, "placements": [(6, 1)]}]


def mig_node(profiles):
    return {"name": "node1", "devices": [{"id": "GPU-a", "index": 0, "count": 7, "devmem": 40960, "devcore": 100, "type": "NVIDIA",
                                          "mode": "mig", "health": True, "vendor": "NVIDIA", "mig_profiles": profiles}]}


def test_get_nodes_usage_restores_mig_allocation_by_profile_and_placement():
    # mig_allocation_test.go:33-58
    o = oa.Oracle()
    o.node_upsert(mig_node(ONE_1G))
    o.usage_apply("pod-1", "node1", "GPU-a", 1, 5120, 14)
    assert o.pod_set_mig("pod-1", [("GPU-a", "1g.5gb", 6, 1)]) == 0
    assert o.snapshot_node_mig("node1") == [(1 << 6, 1)]


def test_get_nodes_usage_fails_closed_without_mig_allocation():
    # mig_allocation_test.go:60-82
    o = oa.Oracle()
    o.node_upsert(mig_node(ONE_1G))
    o.usage_apply("pod-1", "node1", "GPU-a", 1, 5120, 14)
    assert o.snapshot_node_mig("node1") == [(0, 0)]
    o.pod_delete("pod-1")
    assert o.snapshot_node_mig("node1") == [(0, 1)]


def test_get_nodes_usage_restores_unconsumed_mig_reservations():
    # mig_allocation_test.go:84-117: one cached device entry, two reservations: the second is restored from the annotation
    o = oa.Oracle()
    o.node_upsert(mig_node([{"name": "1g.5gb", "memory_mb": 5120, "core": 14, "placements": [(5, 1), (6, 1)]}]))
    o.usage_apply("pod-1", "node1", "GPU-a", 1, 5120, 14)
    o.pod_set_mig("pod-1", [("GPU-a", "1g.5gb", 6, 1), ("GPU-a", "1g.5gb", 5, 1)])
    assert o.snapshot_node_mig("node1") == [((1 << 6) | (1 << 5), 1)]


def test_reservation_on_a_non_mig_or_unknown_device_fails_closed():
    # scheduler.go:788-792,810-841
    o = oa.Oracle()
    node = mig_node(ONE_1G)
    node["devices"].append({"id": "GPU-b", "index": 1, "count": 10, "devmem": 40960, "devcore": 100, "type": "NVIDIA",
                            "mode": "hami-core", "health": True, "vendor": "NVIDIA"})
    o.node_upsert(node)
    o.usage_apply("pod-1", "node1", "GPU-b", 1, 1000, 10)
    o.pod_set_mig("pod-1", [("GPU-b", "1g.5gb", 6, 1)])
    assert o.snapshot_node_mig("node1") == [(0, 1), (0, 0)]        # a reservation on a hami-core GPU
    o.pod_set_mig("pod-1", [("GPU-zz", "1g.5gb", 6, 1)])
    assert o.snapshot_node_mig("node1") == [(0, 0), (0, 1)]        # unknown GPU: every MIG-mode device of the node
    o.pod_set_mig("pod-1", [])
    assert o.snapshot_node_mig("node1") == [(0, 1), (0, 1)]


def fit_in_devices(orc, devices, request, annos=None):
    devs = [oa.dev_from_go(d) for d in devices]
    r = oa.req_from_go("NVIDIA", request)
    rarr = (abi.hs_container_request * 1)(abi.hs_container_request(abi._b(r["key"]), abi._b(r["type"]), r["nums"], r["memreq"], r["mempct"], r["coresreq"]))
    ctr = abi.hs_container(rarr, 1)
    p, keep = abi.build_pod(pod_from(annos or {}))
    res = abi.hs_filter_result()
    rc = orc.L.oracle_fit_in_devices(orc.h, oa.devs_array(devs), len(devs), None, -1, C.byref(ctr), C.byref(p), None, C.byref(res))
    out = abi.read_result(res)
    return rc == 1, out["devices"], (list(out["failed"].values())[0] if out["failed"] else "")


def single_gpu(mode, placements):
    # mig_allocation_test.go:147-164 (singleNvidiaGPUNode)
    d = {"ID": "GPU-a", "Count": 7, "Totalmem": 40960, "Totalcore": 100, "Mode": mode, "Type": "NVIDIA", "Health": True}
    if mode == "mig":
        d["MigProfiles"] = [{"Name": "1g.5gb", "MemoryMB": 5120, "Core": 14, "Placements": [{"Start": s, "Size": z} for s, z in placements]}]
    return d


TWO = {"Nums": 2, "Type": "NVIDIA", "Memreq": 4000}    # twoNvidiaDeviceRequest, mig_allocation_test.go:119-145


def test_fit_in_devices_allows_multiple_mig_placements_on_single_gpu(orc):
    # mig_allocation_test.go:207-235
    dev = single_gpu("mig", [(i, 1) for i in range(7)])
    fit, got, reason = run_fit(orc, "NVIDIA", [dev], TWO, {"nvidia.com/vgpu-mode": "mig"})
    assert fit and len(got) == 2
    fit, devices, reason = fit_in_devices(orc, [dev], TWO, {"nvidia.com/vgpu-mode": "mig"})
    assert fit, reason
    got = devices["NVIDIA"]
    assert len(got) == 1 and len(got[0]) == 2 and got[0][0][0] == "GPU-a" and got[0][1][0] == "GPU-a"
    assert got[0][0][4:] == ("1g.5gb", 6, 1) and got[0][1][4:] == ("1g.5gb", 5, 1)      # 1g placements are taken from the top
    assert got[0][0][2:4] == (5120, 14)


def test_fit_in_devices_rejects_insufficient_mig_placements_on_single_gpu(orc):
    # mig_allocation_test.go:237-256
    dev = single_gpu("mig", [(6, 1)])
    fit, _, direct = run_fit(orc, "NVIDIA", [dev], TWO, {"nvidia.com/vgpu-mode": "mig"})
    assert not fit and "1/1 CardMigTopologyInfeasible" in direct
    fit, _, reason = fit_in_devices(orc, [dev], TWO, {"nvidia.com/vgpu-mode": "mig"})
    assert not fit and reason == direct


def test_fit_in_devices_retains_physical_count_check_for_non_mig_gpu(orc):
    # mig_allocation_test.go:258-268
    fit, _, reason = fit_in_devices(orc, [single_gpu("hami-core", None)], TWO, {"nvidia.com/vgpu-mode": "hami-core"})
    assert not fit and reason == "NodeInsufficientDevice"


def test_is_mig_request(orc):
    # mig_allocation_test.go:166-205 through fitInDevices' gate (score.go:99): with a MIG-mode device a request for more
    # devices than the node has GPUs reaches Fit unless the pod pins another mode
    dev = single_gpu("mig", [(6, 1)])
    _, _, reason = fit_in_devices(orc, [dev], TWO, {"nvidia.com/vgpu-mode": "hami-core"})
    assert reason == "NodeInsufficientDevice"            # explicit HAMi-core mode
    _, _, reason = fit_in_devices(orc, [dev], TWO, {})
    assert "CardMigTopologyInfeasible" in reason         # MIG mode without annotation
    _, _, reason = fit_in_devices(orc, [dev], TWO, {"nvidia.com/vgpu-mode": "mig"})
    assert "CardMigTopologyInfeasible" in reason         # MIG mode annotation


@pytest.fixture(scope="module")
def orc():
    o = oa.Oracle()
    yield o
    o.close()


def test_filter_commits_fail_closed_until_the_informer_delivers_the_reservation():
    # scheduler.go:1132-1139 caches args.Pod as it came in; getNodesUsage (:799-802) then marks the GPU unhealthy
    o = oa.Oracle(node_policy="binpack", gpu_policy="binpack")
    o.node_upsert(mig_node([{"name": "1g.5gb", "memory_mb": 5120, "core": 14, "slice_count": 1, "placements": [(i, 1) for i in range(7)]},
                            {"name": "2g.10gb", "memory_mb": 10240, "core": 28, "slice_count": 2, "placements": [(0, 2), (2, 2), (4, 2)]}]))
    pod = lambda uid, mem, n=1: {"uid": uid, "name": uid, "containers": [[{"type": "NVIDIA", "nums": n, "memreq": mem, "mempct": 101}]]}  # noqa: E731
    r = o.filter(pod("p1", 4000, 2))
    assert r["status"] == 0
    assert r["devices"]["NVIDIA"] == [[("GPU-a", 0, 5120, 14, "1g.5gb", 6, 1), ("GPU-a", 0, 5120, 14, "1g.5gb", 5, 1)]]
    assert r["annotations"]["hami.io/vgpu-mig-allocations"] == (
        '[{"containerIndex":0,"deviceIndex":0,"gpuUUID":"GPU-a","profile":"1g.5gb","placement":{"start":6,"size":1}},'
        '{"containerIndex":0,"deviceIndex":1,"gpuUUID":"GPU-a","profile":"1g.5gb","placement":{"start":5,"size":1}}]')
    assert o.snapshot_node_mig("node1") == [(0, 0)]
    r2 = o.filter(pod("p2", 4000))
    assert r2["status"] == 1 and r2["failed"]["node1"] == "1/1 CardNotHealth"
    o.pod_set_mig("p1", [("GPU-a", "1g.5gb", 6, 1), ("GPU-a", "1g.5gb", 5, 1)])
    assert o.snapshot_node_mig("node1") == [((1 << 6) | (1 << 5), 1)]
    r3 = o.filter(pod("p3", 8000))
    assert r3["status"] == 0 and r3["devices"]["NVIDIA"] == [[("GPU-a", 0, 10240, 28, "2g.10gb", 0, 2)]]
