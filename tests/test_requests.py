"""Request extraction: hs_generate_requests — device.Resourcereqs for one container over the
vendors' GenerateResourceRequests — against the reference's own tables:

  tests/golden/nvidia_generate_requests.json        nvidia/device_test.go TestGenerateResourceRequests (15 cases)
  tests/golden/hygon_generate_requests*.json        hygon/device_test.go Test_GenerateResourceRequests (+ OutOfRangeValues)
  tests/golden/cambricon_generate_requests.json     cambricon/device_test.go Test_GenerateResourceRequests
  tests/golden/ascend_generate_requests*.json       ascend/device_test.go Test_GenerateResourceRequests (+ OutOfRangeValues)
plus the MemoryFactor / DefaultMemory cases of nvidia/device_test.go:1995-2045 and a seeded differential run against
the Python restatement below (test infrastructure, written from the reference independently of the C++).
Host code only: no GPU needed.
"""
import ctypes as C
import json
import os
import random
import re
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import k8s_vgpu_scheduler_b200.engine as eng  # noqa: E402
from k8s_vgpu_scheduler_b200 import abi  # noqa: E402

MAXI32 = 2 ** 31 - 1
SUFFIX = {"": 1, "k": 10 ** 3, "M": 10 ** 6, "G": 10 ** 9, "T": 10 ** 12, "Ki": 2 ** 10, "Mi": 2 ** 20, "Gi": 2 ** 30, "Ti": 2 ** 40}

NVIDIA = dict(word="NVIDIA", kind=abi.HS_REQ_NVIDIA, count="nvidia.com/gpu", mem="nvidia.com/gpumem",
              pct="nvidia.com/gpumem-percentage", cores="nvidia.com/gpucores", default_mem=0, default_cores=0, factor=1)
DCU = dict(word="DCU", kind=abi.HS_REQ_DCU, count="hygon.com/dcunum", mem="hygon.com/dcumem", pct=None, cores="hygon.com/dcucores",
           default_mem=0, default_cores=0, factor=1)
MLU = dict(word="MLU", kind=abi.HS_REQ_MLU, count="cambricon.com/mlu", mem="cambricon.com/mlu.smlu.vmemory", pct=None,
           cores="cambricon.com/mlu.smlu.vcore", default_mem=0, default_cores=0, factor=256)
ASCEND_910A = dict(word="Ascend910A", kind=abi.HS_REQ_ASCEND, count="huawei.com/Ascend910A", mem="huawei.com/Ascend910A-memory",
                   pct=None, cores=None, default_mem=0, default_cores=0, factor=1)
ASCEND_CORE = dict(word="Ascend910B3", kind=abi.HS_REQ_ASCEND, count="huawei.com/Ascend910B3", mem="huawei.com/Ascend910B3-memory",
                   pct=None, cores="huawei.com/Ascend910B3-core", default_mem=0, default_cores=0, factor=1)


def quantity(q):
    """resource.Quantity as the views the reference reads: (AsInt64 value, ok, Value() rounded up, Sign(), String() holds a letter)."""
    if isinstance(q, int):
        return q, True, q, (q > 0) - (q < 0), False
    s = q["_quantity"]
    m = re.match(r"^([+-]?)(\d+)(?:\.(\d+))?(m|k|M|G|T|Ki|Mi|Gi|Ti)?$", s)
    assert m, s
    sign = -1 if m.group(1) == "-" else 1
    whole, frac, suf = int(m.group(2)), m.group(3), m.group(4) or ""
    if suf == "m":
        milli = sign * (whole * 1000 + int((frac or "0").ljust(3, "0")[:3])) if frac else sign * whole
        exact = milli % 1000 == 0
        val = -((-milli) // 1000) if milli < 0 else -(-milli // 1000)       # round away from zero like Value()
        return (milli // 1000 if exact else 0), exact, val, (milli > 0) - (milli < 0)
    if frac is not None:          # decimal-form literals are held as inf.Dec: AsInt64 reports !ok (device_test "16.0Gi")
        v = sign * int(round(float("%d.%s" % (whole, frac)) * SUFFIX[suf]))
        return 0, False, v, (v > 0) - (v < 0)
    v = sign * whole * SUFFIX[suf]
    # the canonical String(): a binary-format quantity keeps a binary suffix from 1024 up, a decimal one gains k/M/G per factor of 1000
    lettered = (abs(v) >= 1024 and v % 1024 == 0) if suf.endswith("i") else (v != 0 and v % 1000 == 0)
    return v, True, v, (v > 0) - (v < 0), lettered


def oracle_generate(v, limits, requests):
    """GenerateResourceRequests restated from nvidia/device.go:520-600, hygon/device.go:168-239,
    cambricon/device.go:250-305, ascend/device.go:303-378.  Returns the request dict or None (zero value)."""
    def get(name):
        if not name:
            return None
        return limits.get(name, requests.get(name))
    if v["kind"] == abi.HS_REQ_ENFLAME:          # enflame/device.go:322-366: direct slices, else memory / core by spec (Quantity.Value())
        wrap = lambda x: ((x + 2 ** 31) % 2 ** 32) - 2 ** 31   # noqa: E731  int32(x)
        q = get(v["count"])
        if q is not None and q[1] and q[0] > 0:
            return {"Nums": 1, "Type": v["word"], "Memreq": q[0], "MemPercentagereq": 0, "Coresreq": 0} if q[0] <= MAXI32 else None
        mem, core = get(v["mem"]), get(v["cores"])
        if mem is None and core is None:
            return None
        if (mem is not None and mem[2] > MAXI32) or (core is not None and core[2] > MAXI32):
            return None
        return {"Nums": 1, "Type": v["word"], "Memreq": wrap(mem[2]) if mem is not None else 0, "MemPercentagereq": -1,
                "Coresreq": wrap(core[2]) if core is not None else 0}
    if v["kind"] == abi.HS_REQ_METAX_SGPU:       # metax/sdevice.go:222-293: Limits only, vmemory in Gi unless it carries a unit
        cnt, mem, core = limits.get(v["count"]), limits.get(v["mem"]), limits.get(v["cores"])
        if cnt is None or not cnt[1] or cnt[0] <= 0 or cnt[0] > MAXI32:
            return None
        cores, memnum = 100, 0
        if core is not None and core[1]:
            if core[0] < 0 or core[0] > MAXI32:
                return None
            cores = core[0]
        if mem is not None and mem[1]:
            if mem[4]:
                memnum = mem[0] // 1024 // 1024 if mem[0] >= 0 else -(-mem[0] // 1024 // 1024)
            else:
                if mem[0] < 0 or mem[0] > MAXI32 // 1024:
                    return None
                memnum = mem[0] * 1024
            if memnum < 0 or memnum > MAXI32:
                return None
        return {"Nums": cnt[0], "Type": v["word"], "Memreq": memnum, "MemPercentagereq": 100 if memnum == 0 else 0, "Coresreq": cores}
    if v["kind"] == abi.HS_REQ_NEURON:           # awsneuron/device.go:287-334 (+ :115-154)
        def valid(q, maxv):
            return q[1] and 0 < q[0] <= maxv
        q = get(v["count"])
        if q is not None:
            return {"Nums": q[0], "Type": v["word"], "Memreq": 0, "MemPercentagereq": 0, "Coresreq": v["default_cores"]} if valid(q, MAXI32) else None
        q = get(v["cores"])
        if q is None or not valid(q, MAXI32 * 2):
            return None
        per = 2 if v["default_cores"] <= 0 else min(v["default_cores"], 2)
        if q[0] <= per:
            nums, cores = 1, q[0]
        elif q[0] % per or q[0] // per > MAXI32:
            return None
        else:
            nums, cores = q[0] // per, per
        return {"Nums": nums, "Type": v["word"], "Memreq": 0, "MemPercentagereq": 0, "Coresreq": cores}
    cnt = get(v["count"])
    if cnt is None or not cnt[1]:
        return None
    n = cnt[0]
    factor = max(v["factor"], 1)
    mem, core = get(v["mem"]), get(v["cores"])
    memnum = 0
    wrap = lambda x: ((x + 2 ** 31) % 2 ** 32) - 2 ** 31   # noqa: E731  int32(x)
    if v["kind"] == abi.HS_REQ_NVIDIA:
        if mem is not None:
            if not mem[1] or mem[0] < 0 or mem[0] > MAXI32 // factor:
                return None
            memnum = mem[0] * factor
        pctn = 101
        pct = get(v["pct"])
        if pct is not None and pct[1]:
            p = min(pct[0], 100)
            pctn = p if p > 0 else 101
        if pctn == 101 and memnum == 0:
            if v["default_mem"]:
                memnum = v["default_mem"]
            else:
                pctn = 100
        cores = v["default_cores"]
        if core is not None and core[1]:
            cores = wrap(core[0])
    elif v["kind"] == abi.HS_REQ_DCU:
        if n <= 0 or n > MAXI32:
            return None
        if mem is not None and mem[1]:
            if mem[0] < 0 or mem[0] > MAXI32:
                return None
            memnum = mem[0] * (v["factor"] if v["factor"] > 1 else 1)
            if memnum > MAXI32:
                return None
        cores = 100
        if core is not None and core[1]:
            if core[0] < 0 or core[0] > MAXI32:
                return None
            cores = core[0]
        pctn = 100 if memnum == 0 else 0
    elif v["kind"] == abi.HS_REQ_MLU:
        if mem is not None:
            if not mem[1] or mem[0] < 0 or mem[0] > MAXI32 // factor:
                return None
            memnum = mem[0] * factor
        cores = 100
        if core is not None and core[1]:
            cores = wrap(core[0])
        pctn = 100 if memnum == 0 else 0
    elif v["kind"] == abi.HS_REQ_ILUVATAR:       # iluvatar/device.go:201-254
        if mem is not None:
            if not mem[1] or mem[0] < 0 or mem[0] > MAXI32 // factor:
                return None
            memnum = mem[0] * factor
        cores = wrap(core[0]) if core is not None and core[1] else 0
        pctn = 100 if memnum == 0 else 0
    elif v["kind"] == abi.HS_REQ_COUNT:          # biren/device.go:141-165, vastai/device.go:143-167
        memnum, pctn, cores = 0, 100, 0
    elif v["kind"] == abi.HS_REQ_ENFLAME:        # enflame/device.go:322-366
        pass                                     # handled before the count lookup below
    elif v["kind"] == abi.HS_REQ_KUNLUN_VXPU:    # kunlun/vdevice.go:173-213 (trimMemory :62-70)
        if mem is not None and mem[1]:
            memnum = 24576 if mem[0] <= 24576 else (49152 if mem[0] <= 49152 else 98304)
        cores = memnum * 100 // 98304
        pctn = 0
        if memnum == 0:
            memnum, pctn, cores = 98304, 100, 100
    elif v["kind"] == abi.HS_REQ_WHOLE_CARD:     # enflame/gcu.go:91-111
        memnum, pctn, cores = 100, 100, 100
    elif v["kind"] == abi.HS_REQ_METAX:          # metax/device.go:134-154
        memnum, pctn, cores = 0, 100, 100
    elif v["kind"] == abi.HS_REQ_AMD:            # amd/device.go:217-266: Limits only
        cnt, mem, core = limits.get(v["count"]), limits.get(v["mem"]), limits.get(v["cores"])
        if cnt is None or not cnt[1] or cnt[0] <= 0 or cnt[0] > MAXI32:
            return None
        n = cnt[0]
        if mem is not None:
            if not mem[1] or mem[0] < 0 or mem[0] > MAXI32:
                return None
            memnum = mem[0]
        cores = 100
        if core is not None:
            if not core[1] or core[0] < 1 or core[0] > 100:
                return None
            cores = core[0]
        pctn = 0
    elif v["kind"] == abi.HS_REQ_MTHREADS:       # mthreads/device.go:202-264
        if n <= 0 or n > MAXI32:
            return None
        if mem is not None:
            if not mem[1] or mem[0] < 0 or mem[0] > MAXI32 // factor:
                return None
            memnum = mem[0] * factor
        cores = wrap(core[0]) if core is not None and core[1] else 0
        pctn = 100 if memnum == 0 else 0
        godiv = lambda a, b: abs(a) // abs(b) * (1 if (a < 0) == (b < 0) else -1)   # noqa: E731
        memnum, cores = godiv(memnum, n), godiv(cores, n)
    else:
        if n <= 0 or n > MAXI32:
            return None
        if mem is not None:
            if mem[3] < 0:
                return None
            if mem[1]:
                if mem[0] > MAXI32:
                    return None
                memnum = mem[0] * (v["factor"] if v["factor"] > 1 else 1)
                if memnum > MAXI32:
                    return None
        cores = 0
        if core is not None:
            if core[2] < 0 or core[2] > MAXI32:
                return None
            cores = core[2]
        pctn = 100 if memnum == 0 else 0
    return {"Nums": wrap(n), "Type": v["word"], "Memreq": wrap(memnum), "MemPercentagereq": pctn, "Coresreq": cores}


def product_generate(vendors, limits, requests):
    L = eng.lib()
    L.hs_generate_requests.argtypes = [C.POINTER(abi.hs_vendor_resources), C.c_uint32, C.POINTER(abi.hs_container_resources),
                                       C.POINTER(abi.hs_container_request), C.c_uint32]
    va = (abi.hs_vendor_resources * len(vendors))(*[
        abi.hs_vendor_resources(abi._b(v["word"]), v["kind"], abi._b(v["count"]), abi._b(v["mem"]), abi._b(v["pct"]), abi._b(v["cores"]),
                                v["default_mem"], v["default_cores"], v["factor"]) for v in vendors])

    def arr(d):
        a = (abi.hs_quantity * max(len(d), 1))()
        for i, (k, q) in enumerate(d.items()):
            a[i] = abi.hs_quantity(abi._b(k), q[0], 1 if q[1] else 0, q[2], q[3], 1 if len(q) > 4 and q[4] else 0)
        return a
    la, ra = arr(limits), arr(requests)
    ctr = abi.hs_container_resources(la, len(limits), ra, len(requests))
    out = (abi.hs_container_request * 32)()
    n = L.hs_generate_requests(va, len(vendors), C.byref(ctr), out, 32)
    assert n >= 0
    return [{"Nums": out[i].nums, "Type": out[i].type.decode(), "Memreq": out[i].memreq, "MemPercentagereq": out[i].mempct,
             "Coresreq": out[i].coresreq, "key": out[i].key.decode()} for i in range(n)]


def resources(case):
    ctr = case.get("ctr") or case.get("args") or case.get("container") or {}
    res = ctr.get("Resources", {}) if isinstance(ctr, dict) else {}
    conv = lambda d: {k: quantity(q) for k, q in d.items()} if isinstance(d, dict) else {}   # noqa: E731
    return conv(res.get("Limits", {})), conv(res.get("Requests", {}))


def want_of(case):
    w = case["want"] if "want" in case else case["expected"]
    if not isinstance(w, dict) or w.get("Nums", 0) <= 0:
        return None
    return {k: w.get(k, 0) for k in ("Nums", "Memreq", "MemPercentagereq", "Coresreq")} | {"Type": w["Type"]}


def strip(r):
    return None if r is None else {k: r[k] for k in ("Nums", "Type", "Memreq", "MemPercentagereq", "Coresreq")}


def _plain(word, kind, count, mem=None, cores=None, factor=1):
    return dict(word=word, kind=kind, count=count, mem=mem, pct=None, cores=cores, default_mem=0, default_cores=0, factor=factor)


ILUVATAR = _plain("MR-V100", abi.HS_REQ_ILUVATAR, "iluvatar.ai/MR-V100-vgpu", "iluvatar.ai/MR-V100.vMem", "iluvatar.ai/MR-V100.vCore", 256)
BIREN = _plain("Biren", abi.HS_REQ_COUNT, "birentech.com/gpu")
VASTAI = _plain("Vastai", abi.HS_REQ_COUNT, "vastaitech.com/va")
MTHREADS = _plain("Mthreads", abi.HS_REQ_MTHREADS, "mthreads.com/vgpu", "mthreads.com/sgpu-memory", "mthreads.com/sgpu-core", 512)
GCU = _plain("GCU", abi.HS_REQ_WHOLE_CARD, "enflame.com/gcu")
AMD = _plain("AMD", abi.HS_REQ_AMD, "amd.com/gpu", "amd.com/gpu-mem", "amd.com/gpu-core-pct")
METAX = _plain("Metax-GPU", abi.HS_REQ_METAX, "metax-tech.com/gpu")
# the golden table keeps the Go identifiers of the two resource names as its map keys (vdevice_test.go:430,444-445)
XPU = _plain("XPU", abi.HS_REQ_KUNLUN_VXPU, "KunlunResourceVCount", "KunlunResourceVMemory")
ENFLAME = _plain("Enflame", abi.HS_REQ_ENFLAME, "enflame.com/drs-gcu", "enflame.com/gcu-memory", "enflame.com/gcu-core")
SGPU = _plain("Metax-SGPU", abi.HS_REQ_METAX_SGPU, "metax-tech.com/sgpu", "metax-tech.com/vmemory", "metax-tech.com/vcore")
NEURON = dict(_plain("AWSNeuron", abi.HS_REQ_NEURON, "aws.amazon.com/neuron", None, "aws.amazon.com/neuroncore"), default_cores=2)

TABLES = [("nvidia_generate_requests", NVIDIA, 15), ("hygon_generate_requests", DCU, 5), ("hygon_generate_requests_range", DCU, 2),
          ("cambricon_generate_requests", MLU, 6), ("ascend_generate_requests", ASCEND_910A, 3),
          ("ascend_generate_requests_range", ASCEND_CORE, 7),
          ("iluvatar_generate_requests", ILUVATAR, 5), ("biren_generate_requests", BIREN, 4), ("vastai_generate_requests", VASTAI, 4),
          ("mthreads_generate_requests", MTHREADS, 8), ("gcu_generate_requests", GCU, 3), ("metax_generate_requests", METAX, 2),
          ("awsneuron_generate_requests", NEURON, 8), ("kunlun_vxpu_generate_requests", XPU, 3),
          ("metax_sgpu_generate_requests", SGPU, 10)]


@pytest.mark.parametrize("name,vendor,count", TABLES, ids=[t[0] for t in TABLES])
def test_generate_requests_golden(name, vendor, count):
    cases = json.load(open(os.path.join(ROOT, "tests", "golden", name + ".json")))["cases"]
    assert len(cases) == count
    for c in cases:
        limits, requests = resources(c)
        want = want_of(c)
        ref = oracle_generate(vendor, limits, requests)
        ref = ref if ref and ref["Nums"] > 0 else None
        assert strip(ref) == want, ("oracle", c["name"])
        got = product_generate([vendor], limits, requests)
        assert (strip(got[0]) if got else None) == want, ("product", c["name"])
        if got:
            assert got[0]["key"] == vendor["word"]


def test_enflame_generate_requests():
    # enflame/device_test.go:112-158 (TestGenerateResourceRequests, …_ByMemoryCore) + the fall-through rules of device.go:322-366
    q = lambda n: (n, True, n, 1)   # noqa: E731
    v = _plain("Enflame", abi.HS_REQ_ENFLAME, "enflame.com/drs-gcu", "enflame.com/gcu-memory", "enflame.com/gcu-core")
    got = product_generate([v], {"enflame.com/drs-gcu": q(3)}, {})
    assert strip(got[0]) == {"Nums": 1, "Type": "Enflame", "Memreq": 3, "MemPercentagereq": 0, "Coresreq": 0}
    got = product_generate([v], {"enflame.com/gcu-memory": q(20480), "enflame.com/gcu-core": q(40)}, {})
    assert strip(got[0]) == {"Nums": 1, "Type": "Enflame", "Memreq": 20480, "MemPercentagereq": -1, "Coresreq": 40}
    got = product_generate([v], {}, {"enflame.com/drs-gcu": q(2)})                       # Requests when Limits has none
    assert strip(got[0])["Memreq"] == 2 and strip(got[0])["MemPercentagereq"] == 0
    got = product_generate([v], {"enflame.com/drs-gcu": (0, True, 0, 0), "enflame.com/gcu-core": q(30)}, {})   # count 0: by spec
    assert strip(got[0]) == {"Nums": 1, "Type": "Enflame", "Memreq": 0, "MemPercentagereq": -1, "Coresreq": 30}
    assert product_generate([v], {"enflame.com/drs-gcu": q(2 ** 31)}, {}) == []          # too large: the zero request
    assert product_generate([v], {"enflame.com/gcu-memory": q(2 ** 31)}, {}) == []
    assert product_generate([v], {}, {}) == []


def test_nvidia_memory_factor_and_default_memory():
    # nvidia/device_test.go:1995-2018 (MemoryFactor 1000: 4 -> 4000) and :2020-2045 (DefaultMemory applies when neither is set)
    q = lambda n: (n, True, n, 1)   # noqa: E731
    v = dict(NVIDIA, factor=1000)
    got = product_generate([v], {"nvidia.com/gpu": q(1), "nvidia.com/gpumem": q(4)}, {})
    assert strip(got[0]) == strip(oracle_generate(v, {"nvidia.com/gpu": q(1), "nvidia.com/gpumem": q(4)}, {}))
    assert got[0]["Memreq"] == 4000 and got[0]["MemPercentagereq"] == 101
    v = dict(NVIDIA, default_mem=3000, default_cores=30)
    got = product_generate([v], {"nvidia.com/gpu": q(2)}, {})
    assert (got[0]["Memreq"], got[0]["MemPercentagereq"], got[0]["Coresreq"], got[0]["Nums"]) == (3000, 101, 30, 2)


def test_resourcereqs_keeps_only_positive_counts_in_vendor_order():
    q = lambda n: (n, True, n, (n > 0) - (n < 0))   # noqa: E731
    vendors = [NVIDIA, DCU, MLU, ASCEND_CORE]
    limits = {"nvidia.com/gpu": q(2), "hygon.com/dcunum": q(0), "cambricon.com/mlu": q(1), "cambricon.com/mlu.smlu.vmemory": q(10),
              "huawei.com/Ascend910B3": q(-1)}
    got = product_generate(vendors, limits, {})
    assert [g["Type"] for g in got] == ["NVIDIA", "MLU"] and got[1]["Memreq"] == 2560


def test_amd_generate_requests_golden():
    """amd/device_test.go Test_GenerateResourceRequests (t.Run blocks + loop table → ingest_funcs.json): Limits only."""
    cases = json.load(open(os.path.join(ROOT, "tests", "golden", "ingest_funcs.json")))["amd_generate_requests"]
    assert len(cases) == 9
    q = lambda d: {k: quantity(v) for k, v in d.items()}   # noqa: E731
    for c in cases:
        limits, requests = q(c["limits"]), q(c["requests"])
        ref = oracle_generate(AMD, limits, requests)
        got = product_generate([AMD], limits, requests)
        got = strip(got[0]) if got else None
        assert strip(ref) == got, c["name"]
        if "want_partial" in c:
            assert all(got[k] == v for k, v in c["want_partial"].items()), c["name"]
        else:
            assert got == c["want"], c["name"]


def test_product_matches_restatement_on_random_containers():
    rng = random.Random(42)
    vendors = [NVIDIA, dict(NVIDIA, word="NV2", factor=4, default_mem=777, default_cores=5), DCU, dict(DCU, word="DCU2", factor=3),
               MLU, ASCEND_910A, ASCEND_CORE, dict(ASCEND_CORE, word="AscF", factor=2),
               ILUVATAR, BIREN, MTHREADS, GCU, AMD, METAX, XPU, ENFLAME, NEURON, dict(NEURON, word="Neuron1", default_cores=1),
               dict(NEURON, word="Neuron0", default_cores=0)]
    names = sorted({v[k] for v in vendors for k in ("count", "mem", "pct", "cores") if v[k]})
    pool = [0, 1, 2, 8, -1, -100, 100, 101, 150, 4096, 2 ** 31 - 1, 2 ** 31, 2200000000, 2 ** 32 + 50, 8388607, 8388608,
            {"_quantity": "16Gi"}, {"_quantity": "16.0Gi"}, {"_quantity": "-1m"}, {"_quantity": "1500m"}, {"_quantity": "2000m"}]
    kept = 0
    for _ in range(4000):
        limits = {n: quantity(rng.choice(pool)) for n in names if rng.random() < 0.45}
        requests = {n: quantity(rng.choice(pool)) for n in names if rng.random() < 0.25}
        want = []
        for v in vendors:
            r = oracle_generate(v, limits, requests)
            if r and r["Nums"] > 0:
                want.append(strip(r))
        got = [strip(g) for g in product_generate(vendors, limits, requests)]
        assert got == want, (limits, requests)
        kept += len(want)
    assert kept > 2000


def test_metax_sgpu_vmemory_unit_rule():
    # metax/sdevice.go:259-281: "has a unit" is read off Quantity.String(), whose canonical form turns 16000 into "16k"
    sg = lambda **kw: {"metax-tech.com/sgpu": quantity(1)} | {"metax-tech.com/" + k: quantity({"_quantity": v}) for k, v in kw.items()}   # noqa: E731
    for lim, want in ((sg(vmemory="16000"), (0, 100)), (sg(vmemory="16001"), (16001 * 1024, 0)), (sg(vmemory="2Gi"), (2048, 0)),
                      (sg(vmemory="512Ki"), (0, 100)), (sg(vmemory="3M"), (2, 0)), (sg(vmemory="0"), (0, 100))):
        ref = oracle_generate(SGPU, lim, {})
        got = product_generate([SGPU], lim, {})
        assert (ref["Memreq"], ref["MemPercentagereq"]) == want, lim
        assert strip(got[0]) == strip(ref), lim
    # Requests are not read at all (Limits only), and a lettered negative quantity lands below zero: the zero request
    assert product_generate([SGPU], {}, {"metax-tech.com/sgpu": quantity(1)}) == []
    assert product_generate([SGPU], sg(vmemory="-4Gi"), {}) == []


def test_resourcereqs_per_container_sequences():
    # pkg/scheduler/config/config_test.go:540-800 (Test_Resourcereqs) with device.Resourcereqs' loop (devices.go:639-690) on top of
    # hs_generate_requests: custom NVIDIA resource names, a GPU container first / in the middle / last among plain ones
    nv = dict(NVIDIA, count="hami.io/gpu", mem="hami.io/gpumem", pct="hami.io/gpumem-percentage", cores="hami.io/gpucores",
              default_mem=0, default_cores=0)
    gpu = {"hami.io/gpu": quantity(1), "hami.io/gpucores": quantity(30), "hami.io/gpumem": quantity(1000)}
    cpu, mem = {"cpu": quantity(1)}, {"memory": quantity(2000)}
    want = {"Nums": 1, "Type": "NVIDIA", "Memreq": 1000, "MemPercentagereq": 101, "Coresreq": 30}

    def reqs(containers):           # PodDeviceRequests: one map per container, keyed by the vendor word
        return [{r["key"]: strip(r) for r in product_generate([nv], limits, {})} for limits in containers]
    assert reqs([cpu]) == [{}]
    assert reqs([gpu]) == [{"NVIDIA": want}]
    assert reqs([gpu, cpu]) == [{"NVIDIA": want}, {}]
    assert reqs([gpu, cpu, mem]) == [{"NVIDIA": want}, {}, {}]
    assert reqs([cpu, gpu, mem]) == [{}, {"NVIDIA": want}, {}]
    assert reqs([cpu, mem, gpu]) == [{}, {}, {"NVIDIA": want}]
    # several vendors configured at once: each container's map holds exactly the vendors it asks for
    both = dict(gpu) | {"metax-tech.com/sgpu": quantity(1)}
    got = reqs([both])                      # only NVIDIA is in the list handed in: the sGPU resource is not seen
    assert got == [{"NVIDIA": want}]
    got = [{r["key"]: strip(r) for r in product_generate([nv, SGPU], both, {})}]
    assert set(got[0]) == {"NVIDIA", "Metax-SGPU"} and got[0]["NVIDIA"] == want and got[0]["Metax-SGPU"]["MemPercentagereq"] == 100


def test_nvidia_memory_factor_default_memory_and_zero_percentage():
    # nvidia/device_test.go:1995-2053 (TestGenerateResourceRequests_MemoryFactor, …_DefaultMemory) and :2055-2093
    # (TestZeroMemoryPercentageIsAccountedAsWholeCard, the request half; the Fit half is in test_oracle_manual_golden.py)
    q = quantity
    gen = lambda v, lim: strip((product_generate([v], lim, {}) or [None])[0])                       # noqa: E731
    both = lambda v, lim: (gen(v, lim), strip(oracle_generate(v, lim, {})))                         # noqa: E731
    f2 = dict(NVIDIA, factor=2)
    got, ref = both(f2, {"nvidia.com/gpu": q(1), "nvidia.com/gpumem": q(1024)})
    assert got == ref and got["Memreq"] == 2048
    got, ref = both(f2, {"nvidia.com/gpu": q(1), "nvidia.com/gpumem": q({"_quantity": "1Gi"})})
    assert got is None and ref is None                      # 2^30 x 2 leaves int32: the zero request
    d512 = dict(NVIDIA, default_mem=512)
    base, ref = both(d512, {"nvidia.com/gpu": q(1)})
    assert base == ref and (base["Memreq"], base["MemPercentagereq"]) == (512, 101)
    for extra in ({"nvidia.com/gpumem-percentage": (0, True, 0, 0, False)}, {"nvidia.com/gpumem": (0, True, 0, 0, False)}):
        got, ref = both(d512, {"nvidia.com/gpu": q(1)} | extra)     # a 0 is "unset": it lands on the default like the absent key
        assert got == ref == base, extra
    got, ref = both(NVIDIA, {"nvidia.com/gpu": q(1), "nvidia.com/gpumem-percentage": (0, True, 0, 0, False)})
    assert got == ref and (got["Memreq"], got["MemPercentagereq"]) == (0, 100)       # no default either: the whole card


def test_memory_factor_overflow_gives_the_zero_request():
    # hygon/device_test.go:662-682 and ascend/device_test.go:1567-1624 (Test_GenerateResourceRequests_MemoryFactorOverflow): a value
    # that fits int32 but leaves it once scaled by MemoryFactor yields the zero request, on ascend's soft-partition and trim paths alike
    q = quantity
    for v, lim in ((dict(DCU, factor=1024), {"hygon.com/dcunum": q(1), "hygon.com/dcumem": q(3000000)}),
                   (dict(ASCEND_CORE, factor=10), {"huawei.com/Ascend910B3": q(1), "huawei.com/Ascend910B3-core": q(10),
                                                   "huawei.com/Ascend910B3-memory": q(300000000)}),
                   (dict(ASCEND_910A, factor=10), {"huawei.com/Ascend910A": q(1), "huawei.com/Ascend910A-memory": q(300000000)})):
        assert product_generate([v], lim, {}) == [], v["word"]
        ref = oracle_generate(v, lim, {})
        assert ref is None or ref["Nums"] <= 0, v["word"]
    # and just below the edge they still go through
    got = product_generate([dict(DCU, factor=1024)], {"hygon.com/dcunum": q(1), "hygon.com/dcumem": q(2097151)}, {})
    assert got and got[0]["Memreq"] == 2097151 * 1024
