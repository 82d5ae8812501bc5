"""ascend PatchAnnotations (ascend/device.go:217-258) against the strings of ascend/device_test.go:262-459 (Test_PatchAnnotations,
…_VNPUCoreMode, …_ZeroCore), through the oracle's Filter: the runtime annotation huawei.com/<word> is []RuntimeInfo with Go's
omitempty — the template name unless the pod runs in hami-core mode, then memory and core."""
import json

import oracle_abi as oa
from oracle_abi import abi


def one(word, uuid, memreq, cores, annos=None, vnpu_core=None):
    o = oa.Oracle()
    node = {"name": "node", "devices": [{"id": uuid, "index": 0, "count": 8, "devmem": 32768 if word == "Ascend910A" else 65536, "devcore": 32,
                                          "type": word, "numa": 0, "mode": "", "health": True, "vendor": word, "network_id": 0}]}
    if vnpu_core is not None:
        node["hami_vnpu_core"] = vnpu_core
    o.node_upsert(node)
    r = o.filter({"uid": "p", "ns": "default", "name": "p", "annotations": annos or {}, "containers": [[
        {"key": word, "type": word, "nums": 1, "memreq": memreq, "mempct": 0, "coresreq": cores}]]})
    o.close()
    assert r["status"] == abi.HS_FIT, r
    return r["annotations"]


def test_template_mode_names_the_template():
    # :262-346 "exist device": 8738 MiB on an Ascend910A is template vir08
    a = one("Ascend910A", "device-0", 8738, 1)
    assert a["huawei.com/Ascend910A"] == '[{"UUID":"device-0","temp":"vir08","memory":8738,"core":1}]'
    assert a["hami.io/Ascend910A-devices-to-allocate"] == a["hami.io/Ascend910A-devices-allocated"] == "device-0,Ascend910A,8738,1:;"
    assert a["predicate-time"] == ""


def test_hami_core_mode_carries_cores_and_memory():
    # :348-402 Test_PatchAnnotations_VNPUCoreMode (the test compares the decoded objects; the text follows the struct's field order)
    a = one("Ascend910B3", "ascend-uuid-1", 8192, 5, {"huawei.com/vnpu-mode": "hami-core"}, vnpu_core=1)
    assert json.loads(a["huawei.com/Ascend910B3"]) == [{"UUID": "ascend-uuid-1", "core": 5, "memory": 8192}]
    assert a["huawei.com/Ascend910B3"] == '[{"UUID":"ascend-uuid-1","memory":8192,"core":5}]'


def test_zero_core_is_omitted():
    # :404-459 Test_PatchAnnotations_ZeroCore: omitempty drops "core": 0
    a = one("Ascend910A", "device-zero-core", 8738, 0)
    assert a["huawei.com/Ascend910A"] == '[{"UUID":"device-zero-core","temp":"vir08","memory":8738}]'
