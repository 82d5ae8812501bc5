"""Every vendor of DEFAULT_VENDORS in ONE cluster and ONE stream: the per-vendor streams of the other suites (their own node
and pod generators) interleaved, so each pod meets every other vendor's nodes (type-substring filters, "Metax-GPU" beside
"Metax-SGPU", "XPU" beside whole kunlun cards, the shared geometry table of MIG and enflame profiles, sGPU class counts
beside plain usage rows) and every node sees every other vendor's pods.  Oracle vs engine, every field, plus the usage
snapshot of every node at the end."""
import random

import pytest

import test_enflame as t_enflame
import test_gpu_mig as t_mig
import test_kunlun as t_kunlun
import test_metax_sgpu as t_sgpu
import test_plain_vendors as t_plain
from oracle_abi import abi


def _mig_stream(seed, n_nodes, n_pods):
    rng = random.Random(seed)
    nodes = t_mig.make_cluster(rng, n_nodes)
    ops, live = [], []
    for p in range(n_pods):
        if live and rng.random() < 0.3:
            ops.append(("delete", live.pop(rng.randrange(len(live)))))
            continue
        pod = t_mig.make_pod(rng, "mig-%04d" % p)
        pod.setdefault("ns", "default")
        ops.append(("filter", pod))
        live.append(pod["uid"])
    return nodes, ops


def mixed_cluster(seed, scale=1):
    streams = {
        "kl": t_kunlun._kunlun_cluster(seed, 10 * scale, 160 * scale)[1:3],
        "ef": t_enflame._cluster(seed + 1, 8 * scale, 140 * scale)[1:3],
        "sg": t_sgpu.sgpu_cluster(seed + 2, 8 * scale, 180 * scale)[1:3],
        "pv": t_plain._cluster(14 * scale, 220 * scale, seed + 3)[1:3],
        "mx": t_plain._metax_cluster(8 * scale, 120 * scale, seed + 4)[1:3],
        "nr": t_plain._neuron_cluster(8 * scale, 140 * scale, seed + 5)[1:3],
        "mg": _mig_stream(seed + 6, 6 * scale, 120 * scale),
    }
    nodes, queues = [], {}
    for tag, (ns, ops) in streams.items():
        for n in ns:
            n = dict(n)
            n["name"] = tag + "." + n["name"]
            nodes.append(n)
        q = []
        for kind, arg in ops:
            if kind == "filter":
                arg = dict(arg)
                arg["uid"] = tag + "." + arg["uid"]
                arg["name"] = arg["uid"]
                q.append((kind, arg))
            else:
                q.append((kind, tag + "." + arg))
        queues[tag] = q
    rng = random.Random(seed * 7 + 1)
    rng.shuffle(nodes)
    ops = []
    while queues:
        tag = rng.choices(list(queues), weights=[len(q) for q in queues.values()])[0]
        ops.append(queues[tag].pop(0))
        if not queues[tag]:
            del queues[tag]
    ids = [d["id"] for n in nodes for d in n["devices"]]
    assert len(ids) == len(set(ids)), "the generators' device ids collide"
    return {"node_policy": rng.choice(["binpack", "spread"]), "gpu_policy": rng.choice(["binpack", "spread"])}, nodes, ops


def tag_counts(ops, results):
    out = {}
    for (kind, arg), r in zip(ops, results):
        if kind != "filter":
            continue
        c = out.setdefault(arg["uid"].split(".")[0], [0, 0])
        c[0 if r["status"] == abi.HS_FIT else 1] += 1
    return out


def test_mixed_stream_exercises_every_vendor_through_the_oracle():
    import parity
    cfg, nodes, ops = mixed_cluster(5)
    ref, _ = parity.run_oracle(cfg, nodes, ops)
    counts = tag_counts(ops, ref)
    assert set(counts) == {"kl", "ef", "sg", "pv", "mx", "nr", "mg"}
    assert all(c[0] >= 20 for c in counts.values()), counts              # every vendor's pods land somewhere …
    assert sum(c[1] for c in counts.values()) >= 40, counts              # … and the stream has its share of refusals
    # the streams really mix: the NVIDIA pods the vendor suites carry land on other streams' NVIDIA nodes too
    assert any(r["node"].split(".")[0] != arg["uid"].split(".")[0]
               for (kind, arg), r in zip(ops, ref) if kind == "filter" and r["status"] == abi.HS_FIT)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [5, 23])
def test_all_vendors_in_one_cluster_gpu_parity(seed):
    import parity
    cfg, nodes, ops = mixed_cluster(seed)
    ref, ref_snap = parity.run_oracle(cfg, nodes, ops)
    got, got_snap = parity.run_engine(cfg, nodes, ops)
    assert parity.compare(ops, ref, got, ref_snap, got_snap) == []
    counts = tag_counts(ops, ref)
    assert len(counts) == 7 and all(c[0] >= 15 for c in counts.values()), counts


@pytest.mark.gpu
def test_all_vendors_in_one_cluster_streaming_mode(monkeypatch):
    import parity
    monkeypatch.setenv("HS_NO_RESIDENT", "1")
    cfg, nodes, ops = mixed_cluster(31)
    ref, ref_snap = parity.run_oracle(cfg, nodes, ops)
    got, got_snap = parity.run_engine(cfg, nodes, ops)
    assert parity.compare(ops, ref, got, ref_snap, got_snap) == []
