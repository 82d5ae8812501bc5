"""Vendors with a plain Fit: iluvatar (pkg/device/iluvatar/device.go:267-371 — hygon's gate list,
one device type per configured CommonWord) and biren (pkg/device/biren/device.go:190-248 — count-only gates).

CPU: the oracle against the reference's own tables (tests/golden/iluvatar_fit.json = Test_Fit,
tests/golden/biren_fit.json = TestDevices_Fit) and TestFit_MutexRejectsUsedDevice (iluvatar/device_test.go:421-442).
GPU: a mixed cluster (NVIDIA + iluvatar + biren nodes) through hs_run_batch against the oracle, bit for bit.
"""
import json
import os
import random

import pytest

import oracle_abi as oa
import parity
from oracle_abi import abi
from test_oracle_golden import run_fit

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load(name):
    return json.load(open(os.path.join(ROOT, "tests", "golden", name + ".json")))["cases"]


@pytest.fixture(scope="module")
def orc():
    o = oa.Oracle("binpack", "spread")
    yield o
    o.close()


@pytest.mark.parametrize("case", load("iluvatar_fit"), ids=[c["name"] for c in load("iluvatar_fit")])
def test_iluvatar_fit(orc, case):
    annos = case["annos"] if isinstance(case["annos"], dict) else {}
    fit, got, reason = run_fit(orc, "MR-V100", case["devices"], case["request"], annos)
    assert fit == case["wantOK"]
    if fit:
        assert [g[0] for g in got] == case["wantDevIDs"] and len(got) == case["wantLen"]
    assert reason == case.get("wantReason", "")


def test_iluvatar_mutex_rejects_used_device(orc):
    dev = {"ID": "dev-0", "Index": 0, "Used": 1, "Count": 100, "Totalmem": 128, "Totalcore": 100, "Numa": 0, "Type": "MR-V100", "Health": True}
    fit, got, reason = run_fit(orc, "MR-V100", [dev], {"Nums": 1, "Type": "MR-V100", "Memreq": 64, "Coresreq": 50},
                               {"hami.io/gpu-scheduler-policy": "mutex"})
    assert not fit and reason == "1/1 ExclusiveDeviceAllocateConflict"


@pytest.mark.parametrize("case", load("biren_fit"), ids=[c["name"] for c in load("biren_fit")])
def test_biren_fit(orc, case):
    annos = case["annos"] if isinstance(case["annos"], dict) else {}
    fit, got, reason = run_fit(orc, "Biren", case["devices"], case["request"], annos)
    assert fit == case["wantFit"]
    if fit:
        assert [g[0] for g in got] == case["wantDevIDs"] and len(got) == case["wantLen"]
        assert all(g[2] == 0 for g in got)                     # Usedmem stays 0 (biren/device.go:231-237)
    assert reason == case["wantReason"]


@pytest.mark.parametrize("case", load("mthreads_fit"), ids=[c["name"] for c in load("mthreads_fit")])
def test_mthreads_fit(orc, case):
    # mthreads/device_test.go:686-1186
    annos = case["annos"] if isinstance(case["annos"], dict) else {}
    fit, got, reason = run_fit(orc, "Mthreads", case["devices"], case["request"], annos)
    assert fit == case["wantFit"]
    if fit:
        assert [g[0] for g in got] == case["wantDevIDs"] and len(got) == case["wantLen"]
    assert reason == case["wantReason"]


@pytest.mark.parametrize("case", load("vastai_fit"), ids=[c["name"] for c in load("vastai_fit")])
def test_vastai_fit_card_mode(orc, case):
    # vastai/device_test.go:542-901.  Card mode (no CustomInfo DeviceStrategy "die") has biren's gate list; the two die-mode
    # cases exercise the AIC exclusivity under mutex (the card a die sits on rides in NetworkID).
    annos = case.get("annos") or {}
    fit, got, reason = run_fit(orc, "Vastai", case["devices"], case["request"], annos)
    assert fit == case["wantFit"]
    if fit:
        assert [g[0] for g in got] == case["wantDevIDs"] and len(got) == case["wantLen"]
    assert reason == case["wantReason"]


@pytest.mark.parametrize("case", load("gcu_fit"), ids=[c["name"] for c in load("gcu_fit")])
def test_enflame_gcu_fit(orc, case):
    # enflame/gcu_test.go:433-677: whole cards — a card in use at all is a conflict
    fit, got, reason = run_fit(orc, "GCU", case["devices"], case["request"], {})
    assert fit == case["wantFit"]
    if fit:
        assert [g[0] for g in got] == case["wantDevIDs"] and len(got) == case["wantLen"]
        assert all((g[2], g[3]) == (case["request"]["Memreq"], case["request"]["Coresreq"]) for g in got)
    assert reason == case["wantReason"]


def _mthreads_node(name, n):
    return {"name": name, "devices": [{"id": "%s-mthreads-%d" % (name, d), "index": d, "count": 100, "devmem": 49152, "devcore": 16,
                                       "type": "Mthreads", "numa": 0, "mode": "", "health": True, "vendor": "Mthreads"} for d in range(n)]}


def test_mthreads_containers_of_a_pod_share_one_card():
    # customFilterRule (mthreads/device.go:266-276; device_test.go:613-684): the second container may only land on the card
    # the first one got, and the annotation set is mthreads' own (device.go:155-177)
    o = oa.Oracle("binpack", "spread")
    o.node_upsert(_mthreads_node("mt-0", 3))
    req = {"key": "Mthreads", "type": "Mthreads", "nums": 1, "memreq": 1024, "mempct": 0, "coresreq": 4}
    r = o.filter({"uid": "p1", "ns": "default", "name": "p1", "annotations": {}, "n_init": 0, "containers": [[req], [req], [req]]})
    assert r["status"] == abi.HS_FIT
    cards = [c[0][0] for c in r["devices"]["Mthreads"]]
    assert len(set(cards)) == 1
    a = r["annotations"]
    assert "hami.io/mthreads-vgpu-devices-allocated" in a and "hami.io/mthreads-vgpu-devices-to-allocate" not in a
    assert a["mthreads.com/predicate-node"] == "mt-0" and a["mthreads.com/predicate-time"] == "" and a["mthreads.com/gpu-index"] in "012"
    # two cards in the first container: no single card can satisfy the rule for the second one
    two = dict(req, nums=2)
    r = o.filter({"uid": "p2", "ns": "default", "name": "p2", "annotations": {}, "n_init": 0, "containers": [[two], [req]]})
    assert r["status"] == abi.HS_NOFIT and "CardNotFoundCustomFilterRule" in r["failed"]["mt-0"]
    o.close()


AMD_CASES = json.load(open(os.path.join(ROOT, "tests", "golden", "ingest_funcs.json")))["amd_fit"]


@pytest.mark.parametrize("case", AMD_CASES, ids=[c["name"][:40] for c in AMD_CASES])
def test_amd_fit(orc, case):
    # amd/device_test.go:300-478 (TestDevices_Fit, t.Run blocks extracted by tools/extract_golden_ingest.py)
    fit, got, reason = run_fit(orc, "AMD", case["devices"], case["request"], case["annos"] or {})
    assert fit == case["wantFit"]
    if case["wantUUIDs"]:
        assert [g[0] for g in got] == case["wantUUIDs"]
    if "wantLen" in case:
        assert len(got) == case["wantLen"]
    if "wantUsedmem" in case:
        assert got[0][2] == case["wantUsedmem"]
    if "wantUsedcores" in case:
        assert got[0][3] == case["wantUsedcores"]
    if "wantReason" in case:
        assert reason == case["wantReason"]
    if "wantReasonContains" in case:
        assert case["wantReasonContains"] in reason


def test_amd_allocation_keeps_the_product_type():
    # amd/device_test.go:354-370: the annotation names the registered product type, the PodDevices key stays "AMD"
    o = oa.Oracle("binpack", "spread")
    o.node_upsert({"name": "amd-0", "devices": [{"id": "dev-0", "index": 0, "count": 2, "devmem": 1000, "devcore": 100,
                                                 "type": "AMD_Instinct_MI300X_VF", "numa": 0, "mode": "", "health": True, "vendor": "AMD"}]})
    r = o.filter({"uid": "p", "ns": "default", "name": "p", "annotations": {}, "n_init": 0,
                  "containers": [[{"key": "AMD", "type": "AMD", "nums": 1, "memreq": 100, "mempct": 0, "coresreq": 10}]]})
    o.close()
    assert r["status"] == abi.HS_FIT
    assert r["annotations"]["hami.io/amd-devices-allocated"] == "dev-0,AMD_Instinct_MI300X_VF,100,10:;"
    assert [d[0] for d in r["devices"]["AMD"][0]] == ["dev-0"]
    assert r["devices"].get("AMD_Instinct_MI300X_VF") == [[]]        # base-type padding (score.go:278-282), no devices under it


def test_vastai_best_combination_by_card(orc):
    # vastai/device_test.go:954-1060 (TestComputeBestCombination) through Fit on a node of idle dies: all dies fit, so the
    # choice is computeBestCombination's — cards by die count, dies inside a card in walk order
    def dies(spec):
        return [{"ID": "dev%d" % i, "Index": i, "Count": 1, "Used": 0, "Type": "Vastai", "Health": True,
                 "CustomInfo": {"AIC": aic, "DeviceStrategy": "die"}} for i, aic in spec]
    # walk order is the list from the END: ids are numbered so that the expected picks read like the reference's table
    fit, got, _ = run_fit(orc, "Vastai", dies([(3, "1"), (2, "1"), (1, "1")]), {"Nums": 2, "Type": "Vastai"}, {})
    assert fit and [g[0] for g in got] == ["dev1", "dev2"]
    six = [(6, "3"), (5, "2"), (4, "2"), (3, "1"), (2, "1"), (1, "1")]
    fit, got, _ = run_fit(orc, "Vastai", dies(six), {"Nums": 3, "Type": "Vastai"}, {})
    assert fit and [g[0] for g in got] == ["dev1", "dev2", "dev3"]
    fit, got, _ = run_fit(orc, "Vastai", dies(six), {"Nums": 4, "Type": "Vastai"}, {})
    assert fit and [g[0] for g in got] == ["dev1", "dev2", "dev3", "dev4"]
    fit, got, reason = run_fit(orc, "Vastai", dies([(3, "2"), (2, "1"), (1, "1")]), {"Nums": 5, "Type": "Vastai"}, {})
    assert not fit and reason == "3/3 AllocatedCardsInsufficientRequest"
    fit, got, _ = run_fit(orc, "Vastai", dies(six), {"Nums": 1, "Type": "Vastai"}, {})
    assert fit and [g[0] for g in got] == ["dev1"]                                  # nums == 1: the first die of the walk


# ---- GPU parity on a generated cluster ---------------------------------------------------------------------
def _cluster(n_nodes, n_pods, seed=11):
    rng = random.Random(seed)
    nodes, kinds = [], []
    for i in range(n_nodes):
        kind = ("NVIDIA", "MR-V100", "BI-V150", "Biren", "AMD", "Mthreads", "Vastai", "GCU")[i % 8]
        nd = rng.randrange(2, 9)
        devs = []
        for d in range(nd):
            if kind == "NVIDIA":
                devs.append({"id": "GPU-%03d-%d" % (i, d), "index": d, "count": 10, "devmem": 81920, "devcore": 100,
                             "type": "NVIDIA A100-SXM4-80GB", "numa": d // 4, "mode": "hami-core", "health": True, "vendor": "NVIDIA"})
            elif kind == "GCU":
                devs.append({"id": "GCU-%03d-%d" % (i, d), "index": d, "count": 1, "devmem": 49152, "devcore": 100,
                             "type": "GCU", "numa": 0, "mode": "", "health": rng.random() > 0.05, "vendor": "GCU"})
            elif kind in ("Biren", "Vastai"):
                devs.append({"id": "%s-%03d-%d" % (kind[:2].upper(), i, d), "index": d, "count": rng.choice([1, 1, 2]), "devmem": 0,
                             "devcore": 0, "type": kind, "numa": 0, "mode": "", "health": rng.random() > 0.05, "vendor": kind})
                if kind == "Vastai" and (i // 8) % 2 == 0:      # every other vastai node is a node of DIES: d // 2 or d // 3 is the card
                    devs[-1]["network_id"] = d // (2 + (i // 16) % 2)
            elif kind == "Mthreads":
                devs.append({"id": "pv-%03d-mthreads-%d" % (i, d), "index": d, "count": 100, "devmem": 49152, "devcore": 16,
                             "type": "Mthreads", "numa": 0, "mode": "", "health": True, "vendor": "Mthreads"})
            elif kind == "AMD":
                devs.append({"id": "AMD-%03d-%d" % (i, d), "index": d, "count": 4, "devmem": 196608, "devcore": rng.choice([64, 100, 304]),
                             "type": rng.choice(["AMD_Instinct_MI300X_VF", "AMD_Instinct_MI250"]), "numa": 0, "mode": "",
                             "health": rng.random() > 0.05, "vendor": "AMD"})
            else:
                devs.append({"id": "ILU-%03d-%d" % (i, d), "index": d, "count": 4, "devmem": 32768, "devcore": 100,
                             "type": kind, "numa": 0, "mode": "", "health": rng.random() > 0.05, "vendor": kind})
        nodes.append({"name": "pv-%03d" % i, "devices": devs})
        kinds.append(kind)
    ops, live = [], []
    for p in range(n_pods):
        kind = rng.choice(["NVIDIA", "MR-V100", "BI-V150", "Biren", "Biren", "Vastai", "GCU", "AMD", "AMD", "Mthreads", "Mthreads", "MR-V50"])   # MR-V50: registered, no node has it
        annos = {}
        if rng.random() < 0.3:
            annos["hami.io/gpu-scheduler-policy"] = rng.choice(["binpack", "spread", "mutex"])
        if rng.random() < 0.3:
            annos["hami.io/node-scheduler-policy"] = rng.choice(["binpack", "spread"])
        if kind == "GCU":
            req = {"key": kind, "type": kind, "nums": rng.choice([1, 1, 2, 4]), "memreq": rng.choice([0, 100, 49152]), "mempct": 100,
                   "coresreq": rng.choice([0, 100, 150])}
        elif kind in ("Biren", "Vastai"):
            req = {"key": kind, "type": kind, "nums": rng.choice([1, 1, 2, 3]), "memreq": rng.choice([0, 512]), "mempct": 0,
                   "coresreq": rng.choice([0, 20, 150])}
            if rng.random() < 0.15:
                n = rng.choice([x for x, k in zip(nodes, kinds) if k == kind])
                dom = "birentech.com" if kind == "Biren" else "vastaitech.com"
                annos[rng.choice([dom + "/use-gpuuuid", dom + "/nouse-gpuuuid"])] = n["devices"][0]["id"]
        elif kind == "Mthreads":
            req = {"key": kind, "type": kind, "nums": rng.choice([1, 1, 1, 2]), "memreq": rng.choice([0, 1024, 8192]),
                   "mempct": rng.choice([100, 25]), "coresreq": rng.choice([0, 2, 4, 16])}
            if req["memreq"]:
                req["mempct"] = 0
            if rng.random() < 0.15:
                n = rng.choice([x for x, k in zip(nodes, kinds) if k == "Mthreads"])
                annos[rng.choice(["mthreads.ai/use-gpuuuid", "mthreads.ai/nouse-gpuuuid"])] = n["devices"][0]["id"]
        elif kind == "AMD":
            req = {"key": kind, "type": kind, "nums": rng.choice([1, 1, 2, 4]), "memreq": rng.choice([0, 0, 65536, 98304]), "mempct": 0,
                   "coresreq": rng.choice([0, 1, 25, 50, 101])}
            r = rng.random()
            if r < 0.15:
                annos[rng.choice(["amd.com/use-gputype", "amd.com/nouse-gputype"])] = rng.choice(["MI300X", "mi250, MI210", "Instinct"])
            elif r < 0.25:
                n = rng.choice([x for x, k in zip(nodes, kinds) if k == "AMD"])
                annos[rng.choice(["amd.com/use-gpu-uuid", "amd.com/nouse-gpu-uuid"])] = n["devices"][0]["id"]
        else:
            req = {"key": kind, "type": kind, "nums": rng.choice([1, 1, 2, 4]), "memreq": rng.choice([0, 4096, 16384]),
                   "mempct": rng.choice([101, 101, 50]), "coresreq": rng.choice([0, 30, 100])}
            if req["memreq"] == 0 and req["mempct"] == 101:
                req["mempct"] = 100
            if kind != "NVIDIA" and rng.random() < 0.15:
                n = rng.choice([x for x, k in zip(nodes, kinds) if k == kind] or nodes)
                annos["hami.io/%s-%s-uuid" % (rng.choice(["use", "no-use"]), kind)] = n["devices"][0]["id"]
        uid = "pv-%05d" % p
        ctrs = [[req]]
        if kind == "Mthreads" and rng.random() < 0.5:      # several containers: customFilterRule ties them to one card
            ctrs = [[dict(req, nums=rng.choice([1, 1, 2]))] for _ in range(rng.choice([2, 3]))]
        elif rng.random() < 0.15:
            ctrs.append([dict(req, nums=1)])
        ops.append(("filter", {"uid": uid, "ns": "default", "name": uid, "annotations": annos, "containers": ctrs, "n_init": 0}))
        live.append(uid)
        if live and rng.random() < 0.35:
            ops.append(("delete", live.pop(rng.randrange(len(live)))))
    return {"node_policy": "binpack", "gpu_policy": "spread"}, nodes, ops


@pytest.mark.gpu
def test_plain_vendors_gpu_parity():
    cfg, nodes, ops = _cluster(60, 900)
    ref, ref_snap = parity.run_oracle(cfg, nodes, ops)
    got, snap = parity.run_engine(cfg, nodes, ops, chunk=300)
    assert parity.compare(ops, ref, got, ref_snap, snap) == []
    fits = {}
    for (kind, arg), r in zip(ops, ref):
        if kind == "filter" and r["status"] == abi.HS_FIT:
            k = arg["containers"][0][0]["type"]
            fits[k] = fits.get(k, 0) + 1
    assert fits.get("Biren", 0) > 20 and fits.get("MR-V100", 0) > 20 and fits.get("BI-V150", 0) > 20 and "MR-V50" not in fits
    assert fits.get("AMD", 0) > 20 and fits.get("Mthreads", 0) > 20 and fits.get("Vastai", 0) > 10 and fits.get("GCU", 0) > 10
    assert any(r["status"] == abi.HS_NOFIT and any("CardNotFoundCustomFilterRule" in v for v in r["failed"].values())
               for (kind, arg), r in zip(ops, ref) if kind == "filter")
    amd = next(r for (kind, arg), r in zip(ops, ref) if kind == "filter" and r["status"] == abi.HS_FIT and arg["containers"][0][0]["type"] == "AMD")
    assert ",AMD_Instinct_" in amd["annotations"]["hami.io/amd-devices-allocated"]      # the device's type, not the key
    sample = next(r for (kind, arg), r in zip(ops, ref) if kind == "filter" and r["status"] == abi.HS_FIT and arg["containers"][0][0]["type"] == "Biren")
    assert "hami.io/Biren-devices-allocated" in sample["annotations"] and "hami.io/Biren-devices-to-allocate" in sample["annotations"]


@pytest.mark.gpu
def test_capacity_registered_nodes_gpu_parity():
    """mthreads and cambricon nodes have no register annotation: their devices come from node capacity
    (mthreads/device.go:130-153, cambricon/device.go:209-232).  Register those through HS_REG_CAPACITY_* and the
    stream must decide exactly like the oracle fed with the derived device lists."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ingest_oracle as O
    import k8s_vgpu_scheduler_b200.engine as eng
    cfg, nodes, ops = _cluster(60, 900, seed=23)
    mlu = []
    for i, cards in enumerate((2, 8, 5)):                       # cambricon nodes: cores = 100 per card
        name = "mlu-%02d" % i
        devs = O.capacity_node_devices("MLU", name, cards * 100, cards * 96)
        assert len(devs) == cards and devs[0]["Devmem"] == 24576
        mlu.append((name, cards, {"name": name, "devices": [
            {"id": d["ID"], "index": d["Index"], "count": d["Count"], "devmem": d["Devmem"], "devcore": d["Devcore"], "type": d["Type"],
             "numa": 0, "mode": "", "health": True, "vendor": "MLU"} for d in devs]}))
    nodes = nodes + [m[2] for m in mlu]
    rng = random.Random(5)
    for p in range(60):
        uid = "mlu-pod-%03d" % p
        req = {"key": "MLU", "type": "MLU", "nums": rng.choice([1, 1, 2]), "memreq": rng.choice([0, 4096, 24576]), "mempct": 0,
               "coresreq": rng.choice([0, 25, 100])}
        if req["memreq"] == 0:
            req["mempct"] = 100
        ops.insert(rng.randrange(len(ops)), ("filter", {"uid": uid, "ns": "default", "name": uid, "annotations": {}, "containers": [[req]], "n_init": 0}))
    ref, ref_snap = parity.run_oracle(cfg, nodes, ops)
    e = eng.Engine(cfg.get("node_policy"), cfg.get("gpu_policy"))
    n_cap = 0
    for n in nodes:
        vendor = n["devices"][0]["vendor"]
        nd = len(n["devices"])
        if vendor == "Mthreads":
            assert e.node_register(n["name"], [("Mthreads", abi.HS_REG_CAPACITY_MTHREADS, None, None, None, 16 * nd, 96 * nd)]) == nd, e.error()
            n_cap += 1
        elif vendor == "MLU":
            assert e.node_register(n["name"], [("MLU", abi.HS_REG_CAPACITY_MLU, None, None, None, 100 * nd, 96 * nd)]) == nd, e.error()
            n_cap += 1
        else:
            e.node_upsert(n)
    assert n_cap >= 10
    e.run_batch(eng.PackedOps(ops))
    got = [e.batch_get(i) for i in range(len(ops))]
    snap = {n["name"]: e.snapshot(n["name"]) for n in nodes}
    assert e.node_register("mlu-zero", [("MLU", abi.HS_REG_CAPACITY_MLU, None, None, None, 0, 64)]) == 0 and "device not found" in e.error()
    e.close()
    assert parity.compare(ops, ref, got, ref_snap, snap) == []
    fit_mlu = sum(1 for (k, a), r in zip(ops, ref) if k == "filter" and a["uid"].startswith("mlu-pod") and r["status"] == abi.HS_FIT)
    assert fit_mlu > 20


# ---- metax whole GPUs ("Metax-GPU", pkg/device/metax/device.go) -------------------------------------------
@pytest.mark.parametrize("case", load("metax_fit"), ids=[c["name"] for c in load("metax_fit")])
def test_metax_fit(orc, case):
    # metax/device_test.go:485-883 (TestMetaxDevices_Fit)
    annos = case["annos"] if isinstance(case["annos"], dict) else {}
    fit, got, reason = run_fit(orc, "Metax-GPU", case["devices"], case["request"], annos)
    assert fit == case["wantFit"]
    if fit:
        assert [g[0] for g in got] == case["wantDevIDs"] and len(got) == case["wantLen"]
    assert reason == case["wantReason"]


def _score_node(orc, n, policy, losses=None, scores=None):
    import ctypes as C
    L = oa.lib()
    L.oracle_vendor_score_node.restype = C.c_float
    L.oracle_vendor_score_node.argtypes = [C.c_void_p, C.c_char_p, C.c_uint32, C.c_char_p, C.c_char_p, C.c_char_p]
    b = lambda s: None if s is None else s.encode()   # noqa: E731
    return L.oracle_vendor_score_node(orc.h, b"Metax-GPU", n, policy.encode(), b(losses), b(scores))


def test_metax_score_node_and_parse_annos(orc):
    # metax/device_test.go:400-483 (Test_ScoreNode) and :101-156 (TestParseMetaxAnnos: value = table[index])
    for c in load("metax_score_node"):
        an = c["args"]["node"]["ObjectMeta"]["Annotations"]
        n = sum(len(ctr) for ctr in c["args"]["podDevices"])
        got = _score_node(orc, n, c["args"]["policy"], an.get("metax-tech.com/gpu.topology.losses"), an.get("metax-tech.com/gpu.topology.scores"))
        assert got == c["want"], c["name"]
    cases = load("metax_parse_annos")
    assert len(cases) == 8
    for c in cases:
        assert 2000 - _score_node(orc, c["index"], "binpack", losses=c["name"]) == c["value"]
    # every error reads as 0; the policy picks the annotation; neither policy nor annotation: 0
    for bad in ('{"1":1.5}', '{"a":1}', '[1]', '{"1":1', '{"1":"2"}', ''):
        assert _score_node(orc, 1, "binpack", losses=bad) == 2000.0
    assert _score_node(orc, 1, "binpack", losses='{"1":null,"01":7}') == 1993.0
    assert _score_node(orc, 1, "spread", losses='{"1":5}') == 0.0 and _score_node(orc, 1, "other", losses='{"1":5}', scores='{"1":5}') == 0.0
    assert _score_node(orc, 2, "spread", scores=' {"2": 300} ') == 1700.0


def _metax_cluster(n_nodes=48, n_pods=700, seed=41):
    import sys
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    rng = random.Random(seed)
    nodes = []
    for i in range(n_nodes):
        name = "mx-%03d" % i
        if i % 4 == 3:          # NVIDIA nodes in between: a Metax pod must never land there
            devs = [{"id": "GPU-m%03d-%d" % (i, d), "index": d, "count": 10, "devmem": 81920, "devcore": 100, "type": "NVIDIA A100-SXM4-80GB",
                     "numa": 0, "mode": "hami-core", "health": True, "vendor": "NVIDIA"} for d in range(4)]
            nodes.append({"name": name, "devices": devs})
            continue
        nd = rng.choice([2, 4, 8, 8])
        devs = [{"id": "%s-metax-%d" % (name, d), "index": d, "count": 1, "devmem": 65536, "devcore": 100, "type": "Metax-GPU", "numa": 0,
                 "mode": "", "health": True, "vendor": "Metax-GPU"} for d in range(nd)]
        node = {"name": name, "devices": devs}
        table = lambda: json.dumps({str(k): rng.choice([0, 10, 110, 270, 540, 930, 1240]) for k in range(1, nd + 1) if rng.random() < 0.85})  # noqa: E731
        r = rng.random()
        if r < 0.75:
            node["metax_losses"] = table()
        if 0.15 < r < 0.9:
            node["metax_scores"] = table()
        if i % 11 == 5:
            node["metax_losses"] = '{"1": 1.5}'                # a table that does not parse: every entry reads 0
        nodes.append(node)
    ops, live = [], []
    for p in range(n_pods):
        uid = "mxp-%05d" % p
        annos = {}
        if rng.random() < 0.7:
            annos["hami.io/node-scheduler-policy"] = rng.choice(["binpack", "spread", "spread", "weird"])
        if rng.random() < 0.2:
            annos["hami.io/gpu-scheduler-policy"] = rng.choice(["binpack", "spread", "mutex"])
        if rng.random() < 0.12:
            req = {"key": "NVIDIA", "type": "NVIDIA", "nums": 1, "memreq": 4096, "mempct": 101, "coresreq": 10}
        else:
            req = {"key": "Metax-GPU", "type": "Metax-GPU", "nums": rng.choice([1, 1, 2, 2, 4, 8]), "memreq": 0, "mempct": 100, "coresreq": 100}
        ctrs = [[req]]
        n_init = 0
        if rng.random() < 0.16:
            ctrs.append([dict(req, nums=1)])                   # a second app container: customFilterRule + Count 1 = never fits on metax
            if rng.random() < 0.5:
                n_init = 1                                     # … but an init container runs on its own copy, and its devices count in ScoreNode's sum
        ops.append(("filter", {"uid": uid, "ns": "default", "name": uid, "annotations": annos, "containers": ctrs, "n_init": n_init}))
        live.append(uid)
        while len(live) > 40:
            ops.append(("delete", live.pop(rng.randrange(len(live)))))
    return {"node_policy": "binpack", "gpu_policy": "spread"}, nodes, ops


def test_metax_cluster_oracle_prefers_low_loss_nodes():
    cfg, nodes, ops = _metax_cluster()
    ref, _ = parity.run_oracle(cfg, nodes, ops, snapshot=False)
    fits = [(a, r) for (k, a), r in zip(ops, ref) if k == "filter" and r["status"] == abi.HS_FIT and a["containers"][0][0]["type"] == "Metax-GPU"]
    assert len(fits) > 150
    assert any(r["score"] > 1000 for _, r in fits) and any(r["score"] < 100 for _, r in fits)   # with and without the ScoreNode term
    by_name = {n["name"]: n for n in nodes}
    assert all(by_name[r["node"]]["devices"][0]["vendor"] == "Metax-GPU" for _, r in fits)
    sample = fits[0][1]
    assert "hami.io/metax-gpu-devices-allocated" in sample["annotations"] and "hami.io/metax-gpu-devices-to-allocate" in sample["annotations"]


@pytest.mark.gpu
def test_metax_gpu_parity():
    cfg, nodes, ops = _metax_cluster()
    ref, ref_snap = parity.run_oracle(cfg, nodes, ops)
    got, snap = parity.run_engine(cfg, nodes, ops, chunk=250)
    assert parity.compare(ops, ref, got, ref_snap, snap) == []
    # the same inventory registered from capacity + the two annotations decides identically
    import k8s_vgpu_scheduler_b200.engine as eng
    e = eng.Engine(cfg.get("node_policy"), cfg.get("gpu_policy"))
    for n in nodes:
        if n["devices"][0]["vendor"] != "Metax-GPU":
            e.node_upsert(n)
            continue
        nd = len(n["devices"])
        assert e.node_register(n["name"], [("Metax-GPU", abi.HS_REG_CAPACITY_METAX, n.get("metax_losses"), n.get("metax_scores"), None, nd, 0)]) == nd, e.error()
    e.run_batch(eng.PackedOps(ops))
    got2 = [e.batch_get(i) for i in range(len(ops))]
    e.close()
    assert parity.compare(ops, ref, got2) == []


# ---- AWS Neuron (pkg/device/awsneuron/device.go) -----------------------------------------------------------
def _neuron_gold():
    return json.load(open(os.path.join(ROOT, "tests", "golden", "ingest_funcs.json")))["awsneuron"]


@pytest.mark.parametrize("case", _neuron_gold()["fit"], ids=[c["name"] for c in _neuron_gold()["fit"]])
def test_awsneuron_fit(orc, case):
    # awsneuron/device_test.go:876-1110 (TestDevices_Fit)
    fit, got, reason = run_fit(orc, "AWSNeuron", case["devices"], case["request"], case["annos"])
    assert fit == case["wantFit"]
    assert len(got) == case["wantLen"]
    if fit:
        assert [g[0] for g in got] == case["wantDevIDs"]
    assert reason == case["wantReason"]


@pytest.mark.parametrize("case", _neuron_gold()["graph_select"], ids=[c["name"] for c in _neuron_gold()["graph_select"]])
def test_awsneuron_graph_select(orc, case):
    # awsneuron/device_test.go:814-874 (Test_graphSelect): 16 devices, the node type on device 0 only; through Fit with
    # Nums = count the picks are the devices whose Index equals the selected positions
    devs = []
    for i in range(16):
        d = {"ID": "d%d" % i, "Index": i, "Used": case["used"].get(str(i), 0), "Count": 2, "Totalcore": 3, "Type": "AWSNeuron", "Health": case["health"]}
        if i == 0:
            d["CustomInfo"] = {"AWSNodeType": case["node_type"]}
        devs.append(d)
    fit, got, reason = run_fit(orc, "AWSNeuron", devs, {"Nums": case["count"], "Type": "AWSNeuron", "Coresreq": 2}, {})
    assert fit == bool(case["want"])
    if fit:
        assert [g[0] for g in got] == ["d%d" % i for i in case["want"]]
        assert all(g[3] == 3 for g in got)                      # the whole card: Usedcores = Totalcore
    else:
        assert reason == "1/16 NumaNotFit"


def _neuron_cluster(n_nodes=40, n_pods=900, seed=77):
    import sys
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ingest_oracle as O
    rng = random.Random(seed)
    nodes, caps = [], {}
    for i in range(n_nodes):
        name = "nr-%03d" % i
        if i % 5 == 4:
            devs = [{"id": "GPU-n%03d-%d" % (i, d), "index": d, "count": 10, "devmem": 81920, "devcore": 100, "type": "NVIDIA A100-SXM4-80GB",
                     "numa": 0, "mode": "hami-core", "health": True, "vendor": "NVIDIA"} for d in range(4)]
            nodes.append({"name": name, "devices": devs})
            continue
        counts = rng.choice([2, 4, 8, 8])
        itype = rng.choice(["inf2.48xlarge", "trn1.32xlarge", "trn1.32xlarge", "Inf1.6xlarge", ""])
        caps[name] = (counts, counts * 2, itype)
        devs = O.neuron_node_devices(name, counts, counts * 2, 0, itype)
        nodes.append({"name": name, "devices": [
            {"id": d["ID"], "index": d["Index"], "count": d["Count"], "devmem": d["Devmem"], "devcore": d["Devcore"], "type": d["Type"], "numa": 0,
             "mode": "", "health": rng.random() > 0.06, "vendor": "AWSNeuron", "network_id": d["NetworkID"]} for d in devs]})
    ops, live = [], []
    for p in range(n_pods):
        uid = "nrp-%05d" % p
        annos = {}
        if rng.random() < 0.35:
            annos["hami.io/gpu-scheduler-policy"] = rng.choice(["binpack", "spread", "mutex", "binpack,numa"])
        if rng.random() < 0.4:
            annos["hami.io/node-scheduler-policy"] = rng.choice(["binpack", "spread"])
        def one():
            r = rng.random()
            if r < 0.45:       # by core: one device, 1 or 2 cores
                return {"key": "AWSNeuron", "type": "AWSNeuron", "nums": 1, "memreq": 0, "mempct": 0, "coresreq": rng.choice([1, 1, 2])}
            if r < 0.9:        # by device or by many cores: whole devices
                return {"key": "AWSNeuron", "type": "AWSNeuron", "nums": rng.choice([2, 2, 3, 4, 4, 8]), "memreq": 0, "mempct": 0, "coresreq": rng.choice([2, 2, 1])}
            return {"key": "NVIDIA", "type": "NVIDIA", "nums": 1, "memreq": 4096, "mempct": 101, "coresreq": 10}
        ctrs = [[one()]]
        n_init = 0
        if rng.random() < 0.3:
            ctrs.append([one()])
            if rng.random() < 0.3:
                ctrs.append([one()])
            if rng.random() < 0.3:
                n_init = 1
        if rng.random() < 0.1:
            tgt = rng.choice([n for n in nodes if n["devices"][0]["vendor"] == "AWSNeuron"])
            annos[rng.choice(["aws.amazon.com/use-neuron-uuid", "aws.amazon.com/nouse-neuron-uuid"])] = tgt["devices"][0]["id"]
        ops.append(("filter", {"uid": uid, "ns": "default", "name": uid, "annotations": annos, "containers": ctrs, "n_init": n_init}))
        live.append(uid)
        while len(live) > 45:
            ops.append(("delete", live.pop(rng.randrange(len(live)))))
    return {"node_policy": "binpack", "gpu_policy": "spread"}, nodes, ops, caps


def test_awsneuron_cluster_oracle_shape():
    cfg, nodes, ops, _ = _neuron_cluster()
    ref, _ = parity.run_oracle(cfg, nodes, ops, snapshot=False)
    fits = [(a, r) for (k, a), r in zip(ops, ref) if k == "filter" and r["status"] == abi.HS_FIT and a["containers"][0][0]["type"] == "AWSNeuron"]
    multi = [r for a, r in fits if a["containers"][0][0]["nums"] > 1]
    single = [r for a, r in fits if a["containers"][0][0]["nums"] == 1 and len(a["containers"]) == 1]
    assert len(multi) > 40 and len(single) > 100
    cores = {c[0][3] for r in single for c in r["devices"]["AWSNeuron"] if c}
    assert cores <= {0, 1, 2, 3} and {1, 3} <= cores                 # core masks, not percentages
    assert any("NumaNotFit" in v for (k, a), r in zip(ops, ref) if k == "filter" and r["status"] == abi.HS_NOFIT for v in r["failed"].values())
    assert set(fits[0][1]["annotations"]) >= {"hami.io/aws-neuron-devices-allocated"} and "hami.io/AWSNeuron-devices-to-allocate" not in fits[0][1]["annotations"]


@pytest.mark.gpu
def test_awsneuron_gpu_parity():
    cfg, nodes, ops, caps = _neuron_cluster()
    ref, ref_snap = parity.run_oracle(cfg, nodes, ops)
    got, snap = parity.run_engine(cfg, nodes, ops, chunk=300)
    assert parity.compare(ops, ref, got, ref_snap, snap) == []
    # registered from capacity + the instance-type label: the same devices (health aside), so the same first decisions
    import k8s_vgpu_scheduler_b200.engine as eng
    for name, (counts, total, itype) in list(caps.items())[:6]:
        devs, _ = eng.decode_node_devices("AWSNeuron", abi.HS_REG_CAPACITY_NEURON, itype, None, name, counts, total, 0)
        want = next(n for n in nodes if n["name"] == name)["devices"]
        assert [(d["ID"], d["Index"], d["Count"], d["Devcore"], d["NetworkID"]) for d in devs] == \
               [(d["id"], d["index"], d["count"], d["devcore"], d["network_id"]) for d in want]


def test_metax_table_parser_product_vs_oracle(orc):
    """The product's json.Unmarshal-into-map[int]int (hs_ingest.cc: decode_int_map) against the oracle's parseMetaxAnnos
    on documents a node annotation could hold, mangled ones included: any error must read as 0 on both sides."""
    import ctypes as C
    import k8s_vgpu_scheduler_b200.engine as eng
    L = eng.lib()
    L.hs_debug_metax_table.restype = C.c_int64
    L.hs_debug_metax_table.argtypes = [C.c_char_p, C.c_int64]
    rng = random.Random(99)
    docs = ['{"1":0,"2":110,"3":270,"4":540,"5":580,"6":730,"7":930,"8":1240}', 'null', '{}', ' { "1" : 5 , "2" : null } ', '{"1":1.0}', '{"1":1e2}',
            '{"+1":3}', '{"01":4,"1":9}', '{"1":9,"01":4}', '{"-1":3,"1":-7}', '{"a":1,"1":2}', '{"1":"2"}', '[1,2]', '{"1":2,}', '{"1":2', '{"1":2}x',
            '{"1":01}', '{"1":-0}', '{"1":9223372036854775807}', '{"1":9223372036854775808}', '{"9223372036854775808":1}', '', '   ', '{"1":true}',
            '{"1":{"2":3}}', '{"1" 2}', '{1:2}']
    for _ in range(300):
        n = rng.randrange(0, 6)
        body = ",".join('"%s":%s' % (rng.choice(["0", "1", "2", "3", "8", "31", "40"]), rng.choice(["0", "10", "110", "-5", "null", "32767", "1.5", '"x"']))
                        for _ in range(n))
        s = "{" + body + "}"
        if rng.random() < 0.3 and s:
            i = rng.randrange(len(s))
            s = s[:i] + rng.choice(["", ",", "}", "{", " ", "x", '"']) + s[i + 1:]
        docs.append(s)
    checked = 0
    for doc in docs:
        for index in (0, 1, 2, 3, 8, 31, 40, -1):
            want = 2000.0 - _score_node(orc, index, "binpack", losses=doc) if index >= 0 else None
            got = L.hs_debug_metax_table(doc.encode(), index)
            if want is not None:
                if abs(got) <= 1 << 22:
                    assert float(got) == want, (doc, index, got, want)
                else:                                   # beyond float32's exact integers: both sides saw the big value
                    assert abs(want) > 1e6, (doc, index, got, want)
                checked += 1
    assert checked > 2000


def test_override_score_metax_vectors(orc):
    # policy/node_policy_test.go:217-302 (TestOverrideScore, the first two cases): two picks per vendor, DCU's ScoreNode adds 0,
    # Metax-GPU's reads the node's topology losses ({"1":123,"2":321} -> 2000 - 321 under binpack); an unrelated annotation -> 0
    assert _score_node(orc, 2, "binpack", losses='{"1":123,"2":321}') == 1679.0
    assert _score_node(orc, 2, "binpack") == 0.0
