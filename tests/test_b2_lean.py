"""B2, the lean CPU baseline (oracle/b2_lean.cc: SoA, no strings, persistent thread pool), must decide like the oracle on
the request shape it handles — it is only worth timing if it is right.  Node, score bits and device picks per Filter."""
import random

import pytest

import oracle_abi as oa
import parity
from oracle_abi import abi


def _run_b2(cfg, nodes, ops, threads):
    b2 = oa.B2(nodes, threads)
    compiled, index_of, skipped = [], {}, 0
    for kind, arg in ops:
        if kind == "filter":
            op = oa.b2_compile(arg, cfg)
            assert op is not None, arg
            index_of[arg["uid"]] = len(compiled)
            compiled.append(op)
        else:
            op = oa.b2_op()
            op.kind, op.rel_op = 1, index_of[arg]
            compiled.append(op)
    sec, out = b2.run(compiled)
    res = [(out[i].node, out[i].score, [out[i].picks[k] for k in range(out[i].n_picks)], [out[i].memreq[k] for k in range(out[i].n_picks)], out[i].cores)
           for i in range(len(compiled))]
    b2.close()
    return res


def _check(cfg, nodes, ops, threads):
    ref, _ = parity.run_oracle(cfg, nodes, ops, snapshot=False)
    got = _run_b2(cfg, nodes, ops, threads)
    idx = {n["name"]: i for i, n in enumerate(nodes)}
    slot = {d["id"]: k for n in nodes for k, d in enumerate(n["devices"])}
    fits = 0
    for i, ((kind, arg), r, g) in enumerate(zip(ops, ref, got)):
        if kind != "filter":
            continue
        if r["status"] == abi.HS_NOFIT:
            assert g[0] == -1, (i, g)
            continue
        assert r["status"] == abi.HS_FIT
        fits += 1
        devs = r["devices"]["NVIDIA"][0]            # [(uuid, idx?, mem, cores)] rows of the one container
        assert g[0] == idx[r["node"]], (i, r["node"], g)
        assert parity.f32bits(g[1]) == parity.f32bits(r["score"]), (i, r["score"], g[1])
        assert [slot[d[0]] for d in devs] == g[2], (i, devs, g)
        assert [d[2] for d in devs] == g[3] and all(d[3] == g[4] for d in devs), (i, devs, g)
    return fits


@pytest.mark.parametrize("threads", [1, 4])
def test_b2_matches_the_oracle_on_the_baseline_configs(threads):
    import k8s_vgpu_scheduler_b200.gen as gen
    total = 0
    for name, kw in (("c1", dict(n_pods=300, departures=0.3)), ("c2", dict(n_nodes=60, n_pods=600)), ("c3", dict(n_nodes=80, n_pods=600))):
        cfg, nodes, ops = gen.CONFIGS[name](**kw)
        total += _check(cfg, nodes, ops, threads)
    assert total > 600


def test_b2_matches_the_oracle_on_every_gpu_policy():
    rng = random.Random(22)
    nodes = []
    for i in range(40):
        nd = rng.choice([2, 4, 8, 8])
        typ = "NVIDIA" if i % 4 == 0 else "NVIDIA A100-SXM4-80GB"        # exact type: the request term enters ComputeScore
        devs = [{"id": "GPU-%03d-%d" % (i, d), "index": d, "count": rng.choice([4, 10]), "devmem": rng.choice([40960, 81920]), "devcore": 100,
                 "type": typ, "numa": rng.randrange(0, 3), "mode": "hami-core", "health": rng.random() > 0.07, "vendor": "NVIDIA"} for d in range(nd)]
        node = {"name": "b2-%03d" % i, "devices": devs}
        if i % 3 == 0:
            node["pair_scores"] = [[0 if a == b else rng.choice([10, 20, 40, 60]) for b in range(nd)] for a in range(nd)]
            for a in range(nd):
                for b in range(a):
                    node["pair_scores"][a][b] = node["pair_scores"][b][a]
        nodes.append(node)
    policies = ["binpack", "spread", "mutex", "topology-aware", "binpack,numa", "numa", "spread,numa", "numa,binpack", "binpack,spread",
                "topology-aware,spread"]
    ops, live = [], []
    for p in range(1500):
        uid = "b2p-%05d" % p
        annos = {"hami.io/gpu-scheduler-policy": rng.choice(policies)}
        if rng.random() < 0.5:
            annos["hami.io/node-scheduler-policy"] = rng.choice(["binpack", "spread"])
        if rng.random() < 0.3:
            annos["nvidia.com/numa-bind"] = rng.choice(["true", "1", "false"])
        mem = rng.choice([0, 2000, 9000, 30000])
        req = {"key": "NVIDIA", "type": "NVIDIA", "nums": rng.choice([1, 1, 2, 2, 3, 4]), "memreq": mem,
               "mempct": 101 if mem else rng.choice([10, 50, 100]), "coresreq": rng.choice([0, 10, 30, 100, 120])}
        ops.append(("filter", {"uid": uid, "ns": "default", "name": uid, "annotations": annos, "containers": [[req]], "n_init": 0}))
        live.append(uid)
        while len(live) > 70:
            ops.append(("delete", live.pop(rng.randrange(len(live)))))
    assert _check({"node_policy": "binpack", "gpu_policy": "spread"}, nodes, ops, 3) > 500
