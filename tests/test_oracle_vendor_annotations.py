"""PatchAnnotations of cambricon and mthreads (cambricon/device_test.go:330-430 Test_PatchAnnotations, mthreads/device.go:155-180)
through the oracle's Filter: the vendor's own keys beside the two hami ones."""
import oracle_abi as oa
from oracle_abi import abi


def run(node, pod):
    o = oa.Oracle()
    o.node_upsert(node)
    r = o.filter(pod)
    o.close()
    assert r["status"] == abi.HS_FIT, r
    return r["annotations"]


def mlu_node():
    return {"name": "node", "devices": [{"id": "device-0", "index": 0, "count": 100, "devmem": 512000, "devcore": 100, "type": "MLU", "numa": 0,
                                          "mode": "", "health": True, "vendor": "MLU"}]}


def test_cambricon_dsmlu_keys():
    req = [{"key": "MLU", "type": "MLU", "nums": 1, "memreq": 256000, "mempct": 0, "coresreq": 1}]
    a = run(mlu_node(), {"uid": "p", "ns": "default", "name": "p", "annotations": {}, "containers": [req]})
    want = {"CAMBRICON_DSMLU_ASSIGNED": "false", "CAMBRICON_DSMLU_PROFILE": "0_1_1000",
            "hami.io/cambricon-mlu-devices-to-allocate": "device-0,MLU,256000,1:;", "hami.io/cambricon-mlu-devices-allocated": "device-0,MLU,256000,1:;"}
    assert {k: a[k] for k in want} == want
    # a container without a request in front keeps its (empty) slot in the encoded list (:383-410)
    a = run(mlu_node(), {"uid": "p", "ns": "default", "name": "p", "annotations": {}, "containers": [[], req]})
    assert a["hami.io/cambricon-mlu-devices-allocated"] == ";device-0,MLU,256000,1:;" and a["CAMBRICON_DSMLU_PROFILE"] == "0_1_1000"


def test_mthreads_gpu_index_and_node_keys():
    node = {"name": "mt-node", "devices": [{"id": "mt-node-mthreads-%d" % i, "index": i, "count": 100, "devmem": 32768, "devcore": 16, "type": "Mthreads",
                                             "numa": 0, "mode": "", "health": True, "vendor": "Mthreads"} for i in range(2)]}
    a = run(node, {"uid": "p", "ns": "default", "name": "p", "annotations": {}, "containers": [[
        {"key": "Mthreads", "type": "Mthreads", "nums": 1, "memreq": 1024, "mempct": 0, "coresreq": 2}]]})
    assert a["mthreads.com/gpu-index"] in ("0", "1") and a["mthreads.com/predicate-time"] == ""
    assert "hami.io/mthreads-vgpu-devices-allocated" in a
