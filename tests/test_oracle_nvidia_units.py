"""The small NVIDIA unit tests of pkg/device/nvidia/device_test.go that are not table-driven (cordon, whole-card booking of
a 0 % request, the gputype / vgpu-mode / numa-bind annotations), replayed through the oracle's Fit hook."""
import pytest

import oracle_abi as oa
from test_oracle_golden import run_fit

NV = "NVIDIA"


@pytest.fixture(scope="module")
def orc():
    o = oa.Oracle()
    yield o
    o.close()


def two():
    return [{"ID": "dev-0", "Count": 100, "Totalmem": 128, "Totalcore": 100, "Type": NV, "Health": True},
            {"ID": "dev-1", "Count": 100, "Totalmem": 128, "Totalcore": 100, "Type": NV, "Health": True}]


def cordon(uuids):
    devs = [{"id": d["ID"], "index": i, "count": 100, "devmem": 128, "devcore": 100, "type": NV, "numa": 0, "mode": "", "health": True,
             "vendor": NV} for i, d in enumerate(two())]
    return {"name": "node", "devices": devs, "device_cordon": uuids}


REQ = {"Nums": 1, "Memreq": 64, "Coresreq": 50, "Type": NV}


def test_fit_device_cordon(orc):
    # device_test.go:1226-1298 (TestFit_DeviceCordon)
    fit, got, reason = run_fit(orc, NV, two(), REQ, None, cordon("dev-1, "))          # trailing separator and blank are tolerated
    assert fit and [d[0] for d in got] == ["dev-0"], reason
    fit, got, reason = run_fit(orc, NV, two(), REQ, None, cordon("dev-0,dev-1"))
    assert not fit and reason == "2/2 CardCordoned"
    devs = two()
    devs[1].update(Used=1, Usedcores=50, Usedmem=64)                                  # running pods on a cordoned device stay
    fit, got, _ = run_fit(orc, NV, devs, REQ, None, cordon("dev-1"))
    assert fit and [d[0] for d in got] == ["dev-0"]
    node = cordon("")
    del node["device_cordon"]
    assert run_fit(orc, NV, two(), REQ, None, node)[0] and run_fit(orc, NV, two(), REQ, None, None)[0]   # nothing cordoned


def test_zero_memory_percentage_books_the_whole_card(orc):
    # device_test.go:2055-2093 (TestZeroMemoryPercentageIsAccountedAsWholeCard), Fit half: the request the generator makes of
    # gpumem-percentage 0 is {Memreq 0, MemPercentagereq 100} (tests/test_requests.py pins that half)
    card = lambda usedmem: [{"ID": "dev-0", "Index": 0, "Used": 0, "Count": 10, "Usedmem": usedmem, "Totalmem": 11264,   # noqa: E731
                             "Totalcore": 100, "Usedcores": 0, "Type": NV, "Health": True}]
    req = {"Nums": 1, "Memreq": 0, "MemPercentagereq": 100, "Coresreq": 0, "Type": NV}
    fit, got, reason = run_fit(orc, NV, card(0), req)
    assert fit and got[0][2] == 11264, (reason, got)
    fit, _, reason = run_fit(orc, NV, card(11260), req)
    assert not fit and "CardInsufficientMemory" in reason


def test_mig_device_without_profiles_does_not_fit(orc):
    # device_test.go:2644-2669 (TestFit_MigPercentageRequestRejectsUndersizedTemplate): the deprecated MigTemplate is not read
    # by dynamic MIG, a MIG-mode GPU that reports no profiles offers nothing
    devs = [{"ID": "dev-0", "Index": 0, "Used": 0, "Count": 1, "Totalmem": 8192, "Totalcore": 100, "Type": NV, "Health": True, "Mode": "mig"}]
    fit, _, _ = run_fit(orc, NV, devs, {"Nums": 1, "MemPercentagereq": 50, "Coresreq": 10, "Type": NV})
    assert not fit


@pytest.mark.parametrize("annos,card,want", [
    # TestCheckGPUtype_NoUse / _EmptyAnnotation / _InUseMismatch (device_test.go:2420-2436, :2582-2585)
    ({"nvidia.com/nouse-gputype": "A100"}, "NVIDIA-A100", False), ({"nvidia.com/nouse-gputype": "A100"}, "NVIDIA-V100", True),
    ({"nvidia.com/use-gputype": ""}, "NVIDIA-A100", True), ({"nvidia.com/use-gputype": "   "}, "NVIDIA-A100", True),
    ({"nvidia.com/nouse-gputype": ""}, "NVIDIA-A100", True), ({"nvidia.com/nouse-gputype": "   "}, "NVIDIA-A100", True),
    ({"nvidia.com/use-gputype": "A100"}, "NVIDIA-V100", False),
    # TestCheckType_AllocateMode (:2438-2452)
    ({"nvidia.com/vgpu-mode": "mps"}, "NVIDIA-A100", False), ({"nvidia.com/vgpu-mode": "hami-core"}, "NVIDIA-A100", True),
])
def test_gputype_and_vgpu_mode_annotations(orc, annos, card, want):
    devs = [{"ID": "dev-0", "Count": 10, "Totalmem": 8192, "Totalcore": 100, "Type": card, "Mode": "hami-core", "Health": True}]
    fit, _, reason = run_fit(orc, NV, devs, {"Nums": 1, "Memreq": 1024, "Coresreq": 10, "Type": NV}, annos)
    assert fit == want, (annos, card, reason)
    if not want:
        assert reason == "1/1 CardTypeMismatch"


@pytest.mark.parametrize("value,enforced", [("true", True), ("false", False), ("notabool", False), (None, False)])
def test_assert_numa(orc, value, enforced):
    # device_test.go:1573-1605 (TestAssertNuma) as it shows in Fit: two idle cards on different NUMA nodes, a pod asking for both
    devs = [{"ID": "dev-0", "Count": 10, "Totalmem": 8192, "Totalcore": 100, "Type": NV, "Numa": 0, "Health": True},
            {"ID": "dev-1", "Count": 10, "Totalmem": 8192, "Totalcore": 100, "Type": NV, "Numa": 1, "Health": True}]
    annos = {} if value is None else {"nvidia.com/numa-bind": value}
    fit, _, reason = run_fit(orc, NV, devs, {"Nums": 2, "Memreq": 1024, "Coresreq": 10, "Type": NV}, annos)
    assert fit == (not enforced), (value, reason)
    if enforced:
        assert "NumaNotFit" in reason


def test_one_container_two_device_types_keep_their_own_devices():
    # score_test.go:4403-4446 (Test_fitInDevices_MultiTypePartition): one container asks for one device of each of two types on
    # a node that holds both; each type's entry must list only its own device
    from oracle_abi import abi
    o = oa.Oracle()
    o.node_upsert({"name": "node-1", "devices": [
        {"id": "uuid-a", "index": 0, "count": 4, "devmem": 8192, "devcore": 100, "type": NV, "numa": 0, "mode": "hami-core", "health": True, "vendor": NV},
        {"id": "uuid-b", "index": 1, "count": 4, "devmem": 8192, "devcore": 100, "type": "DCU", "numa": 0, "mode": "", "health": True, "vendor": "DCU"}]})
    r = o.filter({"uid": "p", "ns": "default", "name": "p", "annotations": {}, "containers": [[
        {"key": NV, "type": NV, "nums": 1, "memreq": 1024, "mempct": 101, "coresreq": 1},
        {"key": "DCU", "type": "DCU", "nums": 1, "memreq": 1024, "mempct": 101, "coresreq": 1}]]})
    assert r["status"] == abi.HS_FIT
    assert {k: [[d[0] for d in c] for c in v] for k, v in r["devices"].items()} == {NV: [["uuid-a"]], "DCU": [["uuid-b"]]}
    o.close()


def test_stale_pod_device_allocations_touch_nothing():
    # scheduler_test.go:175-395 (Test_getNodesUsage_StalePodDeviceAllocation): a cached pod that names a UUID the node no
    # longer has adds no usage anywhere and leaves every real device healthy and schedulable; a valid one beside it still counts
    from oracle_abi import abi

    def fresh(devs):
        o = oa.Oracle()
        o.node_upsert({"name": "node1", "devices": [
            {"id": u, "index": i + 1, "count": 10, "devmem": 1024, "devcore": 100, "type": NV, "numa": 1, "mode": "hami", "health": True,
             "vendor": NV} for i, u in enumerate(devs)]})
        return o
    o = fresh(["GPU-B"])                                                    # StaleOnly
    o.usage_apply("1111", "node1", "GPU-A", 1, 100, 10)
    assert o.snapshot("node1") == [(0, 0, 0)]
    o.close()
    o = fresh(["GPU-B"])                                                    # ValidOnly
    o.usage_apply("2222", "node1", "GPU-B", 1, 200, 20)
    assert o.snapshot("node1") == [(1, 200, 20)]
    o.close()
    o = fresh(["GPU-B"])                                                    # MixedStaleAndValid
    o.usage_apply("1111", "node1", "GPU-A", 1, 100, 10)
    o.usage_apply("2222", "node1", "GPU-B", 1, 200, 20)
    assert o.snapshot("node1") == [(1, 200, 20)]
    o.close()
    o = fresh(["GPU-B", "GPU-C"])                                           # MultipleDevices: a ghost UUID beside two real cards
    o.usage_apply("3333", "node1", "GPU-GHOST", 1, 512, 50)
    assert o.snapshot("node1") == [(0, 0, 0), (0, 0, 0)]
    r = o.filter({"uid": "p", "ns": "default", "name": "p", "annotations": {}, "containers": [[
        {"key": NV, "type": NV, "nums": 2, "memreq": 512, "mempct": 101, "coresreq": 10}]]})
    assert r["status"] == abi.HS_FIT and sorted(d[0] for d in r["devices"][NV][0]) == ["GPU-B", "GPU-C"]     # both still healthy
    o.close()
