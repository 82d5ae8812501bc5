"""Ingest codecs: node register annotations, pair scores, pod device annotations, init-container
collapse.  Three layers, all on the CPU (the decoders are host code inside libhami_sched.so and need no GPU):

  1. the Python oracle (oracle/ingest_oracle.py) against the reference's own test vectors (tests/golden/*.json,
     extracted by tools/extract_golden.py and tools/extract_golden_ingest.py);
  2. the product's C-ABI decoders against the same vectors;
  3. product vs oracle on seeded random inputs, well-formed and mangled (error TEXT compared where the reference
     builds it itself, error-ness only where it comes out of encoding/json).
The entry points that change the HBM inventory (hs_node_register, hs_pod_ingest, hs_pod_shrink) are in
test_gpu_ingest.py.
"""
import json
import os
import random
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, ROOT)

import ingest_oracle as O  # noqa: E402
import k8s_vgpu_scheduler_b200.engine as eng  # noqa: E402
from k8s_vgpu_scheduler_b200 import abi  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
FIELDS = ("ID", "Index", "Count", "Devmem", "Devcore", "Type", "Numa", "Mode", "Health")


def gold(name):
    return json.load(open(os.path.join(GOLD, name + ".json")))


def norm_dev(d, mode_default=""):
    return tuple(d.get(f, {"Mode": mode_default, "Health": False, "ID": "", "Type": ""}.get(f, 0)) for f in FIELDS)


def err_text(e):
    return None if e is None else e["args"][0]


# ---- the two implementations behind one face -----------------------------------------------------------
class OracleImpl:
    name = "oracle"

    @staticmethod
    def csv(s):
        try:
            return O.decode_node_devices(s), None
        except O.DecodeError as e:
            return None, str(e)

    @staticmethod
    def js(s, pair=None):
        try:
            devs = O.unmarshal_node_devices(s)
            return devs, (O.pair_score_matrix(pair, devs) if pair is not None else None), None
        except O.DecodeError as e:
            return None, None, str(e)

    @staticmethod
    def pod(checklist, annos):
        try:
            return O.decode_pod_devices(checklist, annos), None
        except O.DecodeError as e:
            return None, str(e)

    @staticmethod
    def collapse(mode, n_init, raw):
        return (O.collapse_init_container_usage if mode == 0 else O.app_containers_only_device_usage)(n_init, raw)


class ProductImpl:
    name = "product"

    @staticmethod
    def csv(s):
        try:
            return eng.decode_node_devices("DCU", abi.HS_REG_CSV, s)[0], None
        except eng.IngestError as e:
            return ([] if e.rc == abi.HS_ERR_NOTFOUND else None), (None if e.rc == abi.HS_ERR_NOTFOUND else str(e))

    @staticmethod
    def js(s, pair=None):
        try:
            devs, m = eng.decode_node_devices("NVIDIA", abi.HS_REG_JSON, s, pair)
            return devs, m, None
        except eng.IngestError as e:
            if e.rc == abi.HS_ERR_NOTFOUND:
                return [], ([] if pair is not None else None), None
            return None, None, str(e)

    @staticmethod
    def pod(checklist, annos):
        try:
            return eng.decode_pod_devices(checklist, annos), None
        except eng.IngestError as e:
            return None, str(e)

    @staticmethod
    def collapse(mode, n_init, raw):
        return eng.collapse_pod_devices(mode, n_init, raw)


IMPLS = [OracleImpl, ProductImpl]


def strip_slots(pd):
    return {k: [[{f: d[f] for f in ("UUID", "Type", "Usedmem", "Usedcores")} for d in (c or [])] for c in v] for k, v in pd.items()}


def with_slots(pd):
    return {k: [[{f: d.get(f, 0) for f in ("UUID", "Type", "Usedmem", "Usedcores", "Slots")} for d in (c or [])] for c in v]
            for k, v in pd.items()}


# ---- 1 + 2: golden vectors ---------------------------------------------------------------------------------
@pytest.mark.parametrize("impl", IMPLS, ids=lambda i: i.name)
def test_decode_node_devices_golden(impl):
    cases = gold("decode_node_devices")["cases"]
    assert len(cases) == 17
    for c in cases:
        got, err = impl.csv(c["args"])
        assert err == err_text(c["want"]["err"]), c["name"]
        if err is None:
            want = c["want"]["di"] or []
            assert [norm_dev(d) for d in got] == [norm_dev(d) for d in want], c["name"]


@pytest.mark.parametrize("impl", IMPLS, ids=lambda i: i.name)
def test_unmarshal_and_marshal_node_devices_golden(impl):
    for c in gold("unmarshal_node_devices")["cases"]:
        got, _, err = impl.js(c["args"]["str"])
        assert (err is not None) == c["wantErr"], c["name"]
        assert [norm_dev(d) for d in got] == [norm_dev(d) for d in c["want"]], c["name"]
    # MarshalNodeDevices' output (the text the device plugin writes) decodes back to its input
    for c in gold("marshal_node_devices")["cases"]:
        got, _, err = impl.js(c["want"])
        assert err is None
        assert [norm_dev(d) for d in got] == [norm_dev(d) for d in c["args"]["dlist"]], c["name"]
    # EncodeNodeDevices' output decodes back through DecodeNodeDevices
    for c in gold("encode_node_devices")["cases"]:
        got, err = impl.csv(c["want"])
        assert err is None
        assert [norm_dev(d) for d in got] == [norm_dev(d) for d in c["args"]], c["name"]


@pytest.mark.parametrize("impl", IMPLS, ids=lambda i: i.name)
def test_decode_pod_devices_golden(impl):
    names = {"InRequestDevices": "hami.io/vgpu-devices-to-allocate", "SupportDevices": "hami.io/vgpu-devices-allocated"}
    for c in gold("decode_pod_devices")["cases"]:
        ck = c["args"]["checklist"]
        checklist = {"NVIDIA": names[ck["_id"]]} if isinstance(ck, dict) and "_id" in ck else {}
        annos = {}
        for k, v in c["args"]["annos"].items():
            annos[names[json.loads(k)["x"]["_id"]] if k.startswith("{") else k] = v
        got, err = impl.pod(checklist, annos)
        assert err is None
        want = c["want"] if isinstance(c["want"], dict) else {}
        assert strip_slots(got) == strip_slots(want), c["name"]
    # TestDecodePodDevices_BadAnnotation (devices_test.go:297-302)
    got, err = impl.pod({"NVIDIA": "hami.io/vgpu-devices-to-allocate"}, {"hami.io/vgpu-devices-to-allocate": "uuid,type,100:;"})
    assert got is None and "missing fields" in err


@pytest.mark.parametrize("impl", IMPLS, ids=lambda i: i.name)
def test_init_container_collapse_golden(impl):
    cases = gold("ingest_funcs")["init_container"]
    assert len(cases) == 15
    for c in cases:
        mode = 0 if c["func"] == "CollapseInitContainerUsage" else 1
        got = impl.collapse(mode, c["n_init"], c["raw"])
        if "expected" in c:
            assert with_slots(got) == with_slots(c["expected"]), c["name"]
        else:
            assert got[c["expect_slots"]["key"]][0][0]["Slots"] == c["expect_slots"]["slots"], c["name"]


@pytest.mark.parametrize("impl", IMPLS, ids=lambda i: i.name)
def test_nvidia_get_node_devices_golden(impl):
    for c in gold("ingest_funcs")["nvidia_get_node_devices"]:
        devs, m, err = impl.js(c["register"], c.get("pair_scores"))
        assert (err is not None) == c["want_err"], c["name"]
        if err is None:
            assert len(devs) == c["want_len"]
        if "want_score" in c:
            w = c["want_score"]
            j = [d["ID"] for d in devs].index(w["peer"])
            assert m[w["device"] * len(devs) + j] == w["score"]


_CAP_FMT = {"MLU": abi.HS_REG_CAPACITY_MLU, "Mthreads": abi.HS_REG_CAPACITY_MTHREADS, "Metax-GPU": abi.HS_REG_CAPACITY_METAX}


def test_neuron_capacity_node_devices_product_vs_oracle():
    """awsneuron GetNodeDevices (awsneuron/device.go:156-198): device count, total cores, the sticky cores-per-device and the
    instance-type label in; IDs, Count, the core bit mask and the ring flag out."""
    rng = random.Random(5)
    keys = ("Index", "ID", "Count", "Devmem", "Devcore", "Type", "Numa", "Health", "DeviceVendor", "NetworkID")
    for _ in range(400):
        counts = rng.choice([0, 1, 2, 4, 8, 16])
        total = rng.choice([0, counts, counts * 2, counts * 2 + 1, 32, 7])
        per = rng.choice([0, 0, 1, 2, 4, 8])
        itype = rng.choice(["inf2.48xlarge", "trn1.32xlarge", "Inf1", "", None, "m5.large"])
        try:
            want = O.neuron_node_devices("n1", counts, total, per, itype)
        except ValueError as e:
            with pytest.raises(eng.IngestError) as ei:
                eng.decode_node_devices("AWSNeuron", abi.HS_REG_CAPACITY_NEURON, itype, None, "n1", counts, total, per)
            assert str(e) in str(ei.value)
            continue
        devs, _ = eng.decode_node_devices("AWSNeuron", abi.HS_REG_CAPACITY_NEURON, itype, None, "n1", counts, total, per)
        assert [{k: d[k] for k in keys} for d in devs] == want, (counts, total, per, itype)
    # awsneuron/device_test.go:244-293 (Test_GetNodeDevices): node "test", label inf2, neuron 1, neuroncore 2
    one = O.neuron_node_devices("test", 1, 2, 0, "inf2")
    assert (one[0]["ID"], one[0]["Count"], one[0]["Devmem"], one[0]["Devcore"], one[0]["NetworkID"]) == ("test-AWSNeuron-0", 2, 0, 3, 1)


def test_metax_capacity_node_devices_golden():
    """metax/device_test.go TestGetNodeDevices compares Index, ID, Devcore and Devmem (its Count literal differs from the code's)."""
    cases = gold("metax_get_node_devices")["cases"]
    assert len(cases) == 1
    for c in cases:
        count = list(c["node"]["Status"]["Capacity"].values())[0]
        want = [{k: d[k] for k in ("Index", "ID", "Devcore", "Devmem")} for d in c["expected"]]
        ref = O.capacity_node_devices("Metax-GPU", c["node"]["ObjectMeta"]["Name"], count, 0)
        got, err = _capacity_product("Metax-GPU", c["node"]["ObjectMeta"]["Name"], count, 0)
        assert err is None and got == ref
        assert [{k: d[k] for k in ("Index", "ID", "Devcore", "Devmem")} for d in got] == want
    got, err = _capacity_product("Metax-GPU", "n", 0, 0)
    assert got is None and "device not found metax-tech.com/gpu" in err
_CAP_KEYS = ("Index", "ID", "Count", "Devmem", "Devcore", "Type", "Numa", "Health", "DeviceVendor")


def _capacity_product(vendor, node, cores, memory):
    try:
        devs, _ = eng.decode_node_devices(vendor, _CAP_FMT[vendor], None, None, node, cores, memory)
        return [{k: d[k] for k in _CAP_KEYS} for d in devs], None
    except eng.IngestError as e:
        return None, str(e)


def test_capacity_node_devices_golden():
    """cambricon / mthreads Test_GetNodeDevices: oracle and product against the reference's own expectations."""
    cases = gold("ingest_funcs")["capacity_node_devices"]
    assert {c["vendor"] for c in cases} == {"MLU", "Mthreads"}
    for c in cases:
        assert c["want"], c["name"]
        assert O.capacity_node_devices(c["vendor"], c["node"], c["cores"], c["memory"]) == c["want"], c["source"]
        got, err = _capacity_product(c["vendor"], c["node"], c["cores"], c["memory"])
        assert err is None and got == c["want"], c["source"]


def test_capacity_node_devices_product_vs_oracle():
    rng = random.Random(77)
    for vendor in ("MLU", "Mthreads"):
        for _ in range(300):
            cores = rng.choice([0, 1, 15, 16, 17, 99, 100, 101, 800, 1600, 128, -5, rng.randrange(1, 3000)])
            memory = rng.choice([0, 1, 1536, 96, 2 ** 31, 2 ** 52, -7, rng.randrange(0, 1 << 20)])
            node = "n%d" % rng.randrange(100)
            got, err = _capacity_product(vendor, node, cores, memory)
            try:
                want = O.capacity_node_devices(vendor, node, cores, memory)
            except ValueError as e:
                assert got is None and str(e) in err
                continue
            assert got == want, (vendor, cores, memory)


# ---- 3: product vs oracle, seeded random inputs -------------------------------------------------------
def _rand_csv(rng):
    n = rng.randrange(0, 5)
    segs = []
    for i in range(n):
        f = ["GPU-%04x" % rng.randrange(1 << 16), str(rng.choice([1, 10, 100, -3, 2 ** 31 - 1])), str(rng.randrange(0, 200000)),
             str(rng.choice([100, 0, 50])), rng.choice(["NVIDIA-Tesla P4", "DCU-Z100", ""]), str(rng.randrange(0, 4)),
             rng.choice(["true", "false", "T", "0", "1", "True"])]
        if rng.random() < 0.5:
            f += [str(rng.randrange(0, 16)), rng.choice(["hami-core", "mig", "mps", ""])]
        segs.append(",".join(f))
    s = ":".join(segs) + (":" if rng.random() < 0.9 else "")
    r = rng.random()
    if r < 0.35 and s:                                 # mangle one character / field
        i = rng.randrange(len(s))
        s = s[:i] + rng.choice([",", ":", "x", "-", "+", "2147483648", " ", ""]) + s[i + 1:]
    elif r < 0.45:
        s = s.replace("true", rng.choice(["yes", "TRUE", "tRUE"]), 1)
    return s


def test_csv_decoder_matches_oracle_on_random_input():
    rng = random.Random(20260101)
    errs = 0
    for _ in range(3000):
        s = _rand_csv(rng)
        a, ea = OracleImpl.csv(s)
        b, eb = ProductImpl.csv(s)
        if ea is None and any(d["Numa"] != (d["Numa"] & 0x7FFFFFFF) and d["Numa"] >= 0 or d["Numa"] < -2 ** 31 or d["Index"] > 255 for d in a):
            assert eb is not None and "engine's" in eb      # Go's 64-bit int vs the 32-byte HBM record: refused loudly, never truncated
            errs += 1
            continue
        assert ea == eb, repr(s)
        if ea is None:
            assert [norm_dev(d) for d in a] == [norm_dev(d) for d in (b or [])], repr(s)
        else:
            errs += 1
    assert 300 < errs < 2700          # both the accept and the reject side are exercised


def _rand_json_devices(rng):
    devs = []
    for i in range(rng.randrange(0, 5)):
        d = {}
        def put(k, v):
            if rng.random() < 0.9:
                d[rng.choice([k, k.upper(), k.capitalize()]) if rng.random() < 0.2 else k] = v
        put("id", "GPU-%d" % i)
        put("index", rng.choice([i, 0, 7]))
        put("count", rng.choice([10, 1, 2 ** 31 - 1]))
        put("devmem", rng.randrange(0, 200000))
        put("devcore", 100)
        put("type", rng.choice(["NVIDIA A100-SXM4-80GB", "Ascend910B3", "xé\"q\\"]))
        put("numa", rng.randrange(0, 2))
        put("mode", rng.choice(["hami-core", "mps", ""]))
        put("health", rng.random() < 0.8)
        if rng.random() < 0.3:
            d["custominfo"] = {"NetworkID": rng.choice([0, 1, 3.0, 2.7]), "other": [1, {"a": None}]}
        if rng.random() < 0.2:
            d["unknownField"] = {"nested": [1, 2, 3]}
        if rng.random() < 0.15:                     # type confusion: encoding/json rejects these
            k = rng.choice(["count", "devmem", "index", "health", "id", "numa"])
            d[k] = rng.choice(["12", 1.5, -1 if k == "index" else 2 ** 31, None, [1], True if k != "health" else "true"])
        devs.append(d)
    text = json.dumps(devs, ensure_ascii=rng.random() < 0.5)
    r = rng.random()
    if r < 0.08:
        text = text[:rng.randrange(len(text) + 1)]      # truncated document
    elif r < 0.12:
        text = rng.choice(["null", "{}", "[null]", "[1]", "", "[]x", " [ ] "])
    return text, devs


def test_json_decoder_matches_oracle_on_random_input():
    rng = random.Random(77)
    errs = 0
    for _ in range(3000):
        text, devs = _rand_json_devices(rng)
        pair = None
        if rng.random() < 0.5:
            ids = [d.get("id", "") for d in devs if isinstance(d.get("id", ""), str)]
            pair = json.dumps([{"uuid": u, "score": {v: rng.randrange(0, 300) for v in ids if v != u and rng.random() < 0.8}}
                               for u in ids if rng.random() < 0.9])
        a, ma, ea = OracleImpl.js(text, pair)
        b, mb, eb = ProductImpl.js(text, pair)
        if ea is None and any(not -2 ** 31 <= d["Numa"] < 2 ** 31 or d["Index"] > 255 for d in a):
            assert eb is not None and "engine's" in eb      # wider than the HBM record: refused loudly, never truncated
            errs += 1
            continue
        assert (ea is None) == (eb is None), (text, pair, ea, eb)
        if ea is None:
            assert [norm_dev(d) for d in a] == [norm_dev(d) for d in b], text
            assert [d["NetworkID"] for d in a] == [d["NetworkID"] for d in b], text
            if a:                                      # an empty list has no matrix on either side
                assert ma == mb, (text, pair)
        else:
            errs += 1
    assert 200 < errs < 2500


def test_pod_decoder_and_collapse_match_oracle_on_random_input():
    rng = random.Random(5)
    keys = {"NVIDIA": "hami.io/vgpu-devices-allocated", "DCU": "hami.io/dcu-devices-allocated", "Ascend910B3": "huawei.com/x"}
    errs = 0
    for _ in range(3000):
        annos = {}
        for k, anno in keys.items():
            if rng.random() < 0.6:
                ctrs = []
                for c in range(rng.randrange(0, 5)):
                    devs = ["%s,%s,%d,%d" % (rng.choice(["u0", "u1", "u2", "u3"]), k, rng.randrange(0, 50000), rng.randrange(0, 101))
                            for _ in range(rng.randrange(0, 4))]
                    ctrs.append(":".join(devs) + (":" if devs else ""))
                s = ";".join(ctrs) + (";" if rng.random() < 0.8 else "")
                if rng.random() < 0.1 and s:
                    i = rng.randrange(len(s))
                    s = s[:i] + rng.choice(["x", ",", ":", ";", "99999999999"]) + s[i + 1:]
                annos[anno] = s
        if rng.random() < 0.3:
            annos["unrelated"] = "v"
        a, ea = OracleImpl.pod(keys, annos)
        b, eb = ProductImpl.pod(keys, annos)
        assert ea == eb, annos
        if ea is not None:
            errs += 1
            continue
        assert strip_slots(a) == strip_slots(b), annos
        n_init = rng.randrange(0, 3)
        for mode in (0, 1):
            assert with_slots(OracleImpl.collapse(mode, n_init, a)) == with_slots(ProductImpl.collapse(mode, n_init, a)), (mode, n_init, annos)
    assert 50 < errs < 1500


# ---- dynamic MIG: what the register annotation and the pod annotation carry ------------------------------------
def test_mig_profiles_from_the_register_annotation():
    # nvidia/device_test.go:2466-2482 (TestGetNodeDevices_MigProfilesFromNode): name and placements survive the decode
    c = [x for x in gold("ingest_funcs")["nvidia_get_node_devices"] if x["name"] == "TestGetNodeDevices_MigProfilesFromNode"][0]
    devs, _ = eng.decode_node_devices("NVIDIA", abi.HS_REG_JSON, c["register"])
    assert len(devs) == 1 and devs[0]["Mode"] == "mig"
    assert devs[0]["MigProfiles"] == [{"name": "1g.5gb", "memory_mb": 5120, "core": 14, "slice_count": 1, "placements": [(6, 1)]}]


def test_decode_mig_allocations_golden():
    # nvidia/mig_profiles_test.go:36-88
    one = '[{"containerIndex":0,"deviceIndex":0,"gpuUUID":"GPU-a","profile":"1g.5gb","placement":{"start":6,"size":1}}]'
    assert eng.decode_mig_allocations(one) == [("GPU-a", "1g.5gb", 6, 1)]                       # TestEncodeDecodeMigAllocations
    assert eng.decode_mig_allocations("") == []
    dup = ('[{"containerIndex":0,"deviceIndex":0,"gpuUUID":"GPU-a","profile":"1g.5gb","placement":{"start":6,"size":1}},'
           '{"containerIndex":0,"deviceIndex":0,"gpuUUID":"GPU-a","profile":"1g.5gb","placement":{"start":5,"size":1}}]')
    with pytest.raises(eng.IngestError, match="duplicate MIG allocation for container 0 device 0"):  # …RejectsDuplicateDeviceIndex
        eng.decode_mig_allocations(dup)
    full = ('[{"containerIndex":0,"deviceIndex":0,"gpuUUID":"GPU-a","profile":"2g.10gb","placement":{"start":0,"size":2},'
            '"migUUID":"MIG-a","gpuInstanceID":4,"computeInstanceID":0}]')
    assert eng.decode_mig_allocations(full) == [("GPU-a", "2g.10gb", 0, 2)]                      # …PreservesZeroRuntimeInstanceIDs
    part = '[{"containerIndex":0,"deviceIndex":0,"gpuUUID":"GPU-a","profile":"2g.10gb","placement":{"start":0,"size":2},"migUUID":"MIG-a"}]'
    with pytest.raises(eng.IngestError, match="partial runtime identity"):                       # …RejectsPartialRuntimeIdentity
        eng.decode_mig_allocations(part)
    with pytest.raises(eng.IngestError, match="MIG allocation 0 is incomplete"):
        eng.decode_mig_allocations('[{"containerIndex":0,"deviceIndex":0,"gpuUUID":"GPU-a","profile":"1g.5gb","placement":{"start":6,"size":0}}]')
    with pytest.raises(eng.IngestError):
        eng.decode_mig_allocations("not json")
    # the scheduler's own test annotation, two reservations on one GPU (scheduler/mig_allocation_test.go:96-99)
    two = '''[
        {"containerIndex":0,"deviceIndex":0,"gpuUUID":"GPU-a","profile":"1g.5gb","placement":{"start":6,"size":1}},
        {"containerIndex":0,"deviceIndex":1,"gpuUUID":"GPU-a","profile":"1g.5gb","placement":{"start":5,"size":1}}
    ]'''
    assert eng.decode_mig_allocations(two) == [("GPU-a", "1g.5gb", 6, 1), ("GPU-a", "1g.5gb", 5, 1)]


def test_hygon_and_ascend_get_node_devices_vectors():
    # hygon/device_test.go:242-330 and ascend/device_test.go:168-262 (Test_GetNodeDevices): the rows those tests compare, the
    # error returns with the texts of the code they reach (devices.go:317-384 DecodeNodeDevices, ascend/device.go:197-215)
    def run(vendor, fmt, anno):
        devs, _ = eng.decode_node_devices(vendor, fmt, anno, None, "n")
        return [(d["ID"], d["Index"], d["Count"], d["Devmem"], d["Devcore"], d["Type"], d["Numa"], d["Health"], d["Mode"], d["DeviceVendor"]) for d in devs]
    assert run("DCU", abi.HS_REG_CSV, "test-0,1,1024,100,DCU,0,true,0,test:") == [("test-0", 0, 1, 1024, 100, "DCU", 0, True, "test", "DCU")]
    for anno, msg in ((None, "annos not found"), (":", "no gpu found on node"), ("", "node annotation missing device separator"),
                      ("test-0,1,1024,100,DCU:", "unexpected field count 5")):
        with pytest.raises(eng.IngestError) as ei:
            run("DCU", abi.HS_REG_CSV, anno)
        assert msg in str(ei.value), (anno, str(ei.value))
    anno = '[{"ID":"GPU-0","Count":4,"Devmem":8738,"Devcore":8,"Type":"huawei.com/Ascend910","Numa":0,"Health":true}]'
    assert run("Ascend910A", abi.HS_REG_JSON, anno) == [("GPU-0", 0, 4, 8738, 8, "huawei.com/Ascend910", 0, True, "", "Ascend910A")]
    for anno, msg in (("[]", "no device found on node"), (None, "annos not found"), ("", "unexpected end of JSON input")):
        with pytest.raises(eng.IngestError) as ei:
            run("Ascend910A", abi.HS_REG_JSON, anno)
        assert msg in str(ei.value), (anno, str(ei.value))


def test_plain_vendor_get_node_devices_vectors():
    # Test_GetNodeDevices of iluvatar (device_test.go:35-133), amd (:149-170), biren (:83-160), vastai (:83-160) and
    # TestGCUDevices_GetNodeDevices (enflame/gcu_test.go:194-300), through the product's decoder
    def run(vendor, fmt, anno, **kw):
        devs, _ = eng.decode_node_devices(vendor, fmt, anno, None, kw.pop("node", "n"), **kw)
        return [(d["ID"], d["Index"], d["Count"], d["Devmem"], d["Devcore"], d["Type"], d["Numa"], d["Health"], d["Mode"], d["DeviceVendor"]) for d in devs]
    assert run("MR-V100", abi.HS_REG_CSV, "GPU-bad51c5a-ed4c-591d-91bf-c04a12e19eae,10,8192,100,MR-V100,0,true:") == [
        ("GPU-bad51c5a-ed4c-591d-91bf-c04a12e19eae", 0, 10, 8192, 100, "MR-V100", 0, True, "hami-core", "MR-V100")]
    assert run("AMD", abi.HS_REG_JSON, '[{"id":"GPU-0","index":0,"count":1,"devmem":8192,"devcore":100,"type":"amd","numa":0,"health":true}]') == [
        ("GPU-0", 0, 1, 8192, 100, "amd", 0, True, "", "AMD")]
    # biren and vastai overwrite Devcore with 100 after the decode (biren/device.go:84, vastai/device.go:86): capacity_cores asks for it
    for word in ("Biren", "Vastai"):
        anno = '[{"id":"7-0","count":1,"devmem":32768,"type":"%s","health":true,"devicepairscore":{}}]' % word
        assert run(word, abi.HS_REG_JSON, anno, cores=100) == [("7-0", 0, 1, 32768, 100, word, 0, True, "", word)]
        assert run(word, abi.HS_REG_JSON, anno)[0][4] == 0                     # what the annotation itself says
        for bad, msg in ((None, "annos not found"), (":", "invalid character"), ("", "unexpected end of JSON input"), ("[]", "no gpu found on node")):
            with pytest.raises(eng.IngestError) as ei:
                run(word, abi.HS_REG_JSON, bad, cores=100)
            assert msg in str(ei.value), (bad, str(ei.value))
    # whole-card GCUs come from the node's capacity
    assert run("GCU", abi.HS_REG_CAPACITY_GCU, None, node="testnode", cores=2) == [
        ("testnode-GCU-0", 0, 1, 100, 100, "GCU", 0, True, "", "GCU"), ("testnode-GCU-1", 1, 1, 100, 100, "GCU", 0, True, "", "GCU")]
    with pytest.raises(eng.IngestError) as ei:
        run("GCU", abi.HS_REG_CAPACITY_GCU, None, node="testnode", cores=0)
    assert "device not found enflame.com/gcu" in str(ei.value)
