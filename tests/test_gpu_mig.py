"""Dynamic MIG on the GPU against the oracle: the reference's own MIG scenarios through the C ABI,
then a seeded stream over a cluster that mixes GPUs in MIG mode with hami-core GPUs — Filters in batches, the
informer delivering (or withholding) each pod's hami.io/vgpu-mig-allocations, deletes — compared decision by
decision (node, score bits, devices with profile + placement, annotation text, FailedNodes) and state by state
(occupied slices and health of every device after every step)."""
import json
import random

import pytest

import oracle_abi as oa
import parity

pytestmark = pytest.mark.gpu

A100 = [
    {"name": "1g.5gb", "memory_mb": 5120, "core": 14, "slice_count": 1, "placements": [(i, 1) for i in range(7)]},
    {"name": "2g.10gb", "memory_mb": 10240, "core": 28, "slice_count": 2, "placements": [(0, 2), (2, 2), (4, 2)]},
    {"name": "3g.20gb", "memory_mb": 20480, "core": 42, "slice_count": 3, "placements": [(0, 4), (4, 4)]},
    {"name": "7g.40gb", "memory_mb": 40960, "core": 100, "slice_count": 7, "placements": [(0, 8)]},
]
# a second geometry: duplicate profile names (the first one's placements count), equal memory with different slice counts
ODD = [
    {"name": "2g.10gb", "memory_mb": 10240, "core": 30, "slice_count": 3, "placements": [(4, 2), (0, 2)]},
    {"name": "1g.10gb", "memory_mb": 10240, "core": 14, "slice_count": 1, "placements": [(1, 1), (3, 1), (5, 1)]},
    {"name": "2g.10gb", "memory_mb": 10240, "core": 28, "slice_count": 2, "placements": [(2, 2)]},
    {"name": "4g.20gb", "memory_mb": 20480, "core": 57, "slice_count": 4, "placements": [(0, 4)]},
]


def gpu(uuid, index, mode, profiles=None, numa=0, health=True):
    d = {"id": uuid, "index": index, "count": 7 if mode == "mig" else 10, "devmem": 40960, "devcore": 100,
         "type": "NVIDIA-A100", "numa": numa, "mode": mode, "health": health, "vendor": "NVIDIA"}
    if profiles is not None:
        d["mig_profiles"] = profiles
    return d


def both(node_policy="binpack", gpu_policy="binpack"):
    import k8s_vgpu_scheduler_b200.engine as eng
    return eng.Engine(node_policy, gpu_policy), oa.Oracle(node_policy, gpu_policy)


def pod(uid, mem, n=1, annos=None, cores=0, mempct=101):
    return {"uid": uid, "name": uid, "annotations": annos or {},
            "containers": [[{"type": "NVIDIA", "nums": n, "memreq": mem, "mempct": mempct, "coresreq": cores}]]}


def reservations_of(result):
    """What the informer would decode from the patched pod: [(gpu uuid, profile, start, size)] in annotation order."""
    raw = result["annotations"].get("hami.io/vgpu-mig-allocations")
    if not raw:
        return []
    return [(a["gpuUUID"], a["profile"], a["placement"]["start"], a["placement"]["size"]) for a in json.loads(raw)]


def same_state(e, o, nodes):
    for n in nodes:
        assert e.snapshot_node_mig(n["name"]) == o.snapshot_node_mig(n["name"]), n["name"]
        assert e.snapshot(n["name"]) == o.snapshot(n["name"]), n["name"]


def test_reference_scenarios_through_the_abi():
    # mig_allocation_test.go:207-256 + scheduler.go:799-802,1132-1139: two instances on one GPU, fail closed until the
    # informer delivers the reservation, then the next profile goes beside them
    e, o = both()
    node = {"name": "node1", "devices": [gpu("GPU-a", 0, "mig", A100)]}
    e.node_upsert(node)
    o.node_upsert(node)
    for step, (p, want_status) in enumerate([(pod("p1", 4000, 2), 0), (pod("p2", 4000), 1)]):
        r, g = o.filter(p), e.filter(p)
        assert parity.compare([("filter", p)], [r], [g]) == []
        assert g["status"] == want_status
        same_state(e, o, [node])
    r1 = o.filter(pod("p1b", 1, 1))     # still unhealthy
    assert r1["failed"]["node1"] == "1/1 CardNotHealth"
    assert e.filter(pod("p1b", 1, 1))["failed"] == r1["failed"]
    res = [("GPU-a", "1g.5gb", 6, 1), ("GPU-a", "1g.5gb", 5, 1)]
    assert e.pod_set_mig("p1", res) == 0 and o.pod_set_mig("p1", res) == 0
    same_state(e, o, [node])
    assert e.snapshot_node_mig("node1") == [((1 << 6) | (1 << 5), 1)]
    p3 = pod("p3", 8000)
    r, g = o.filter(p3), e.filter(p3)
    assert parity.compare([("filter", p3)], [r], [g]) == []
    assert g["devices"]["NVIDIA"] == [[("GPU-a", 0, 10240, 28, "2g.10gb", 0, 2)]]
    # one placement left for a request of two (mig_allocation_test.go:237-256): CardMigTopologyInfeasible
    e2, o2 = both()
    one = {"name": "n", "devices": [gpu("GPU-a", 0, "mig", [{"name": "1g.5gb", "memory_mb": 5120, "core": 14, "placements": [(6, 1)]}])]}
    e2.node_upsert(one)
    o2.node_upsert(one)
    p = pod("q", 4000, 2, {"nvidia.com/vgpu-mode": "mig"})
    r, g = o2.filter(p), e2.filter(p)
    assert r["status"] == 1 and "1/1 CardMigTopologyInfeasible" in r["failed"]["n"]
    assert parity.compare([("filter", p)], [r], [g]) == []
    # a hami-core pod asking for two devices on the one-GPU node keeps the physical count check (:258-268)
    p = pod("q2", 4000, 2, {"nvidia.com/vgpu-mode": "hami-core"})
    r, g = o2.filter(p), e2.filter(p)
    assert r["failed"]["n"] == "NodeInsufficientDevice" and g["failed"] == r["failed"]
    for x in (e, o, e2, o2):
        x.close()


def test_informer_rules_match_the_oracle():
    # scheduler.go:788-841: reservation on a hami-core GPU, on an unknown GPU, left over beyond the cached entries
    e, o = both()
    node = {"name": "node1", "devices": [gpu("GPU-a", 0, "mig", A100), gpu("GPU-b", 1, "hami-core"), gpu("GPU-c", 2, "mig", ODD)]}
    for x in (e, o):
        x.node_upsert(node)
        x.usage_apply("pod-1", "node1", "GPU-b", 1, 1000, 10)
        x.usage_apply("pod-2", "node1", "GPU-a", 1, 5120, 14)
    same_state(e, o, [node])
    steps = [("pod-1", [("GPU-b", "1g.5gb", 6, 1)]), ("pod-1", [("GPU-zz", "1g.5gb", 6, 1)]), ("pod-1", []),
             ("pod-2", [("GPU-a", "1g.5gb", 6, 1), ("GPU-a", "2g.10gb", 0, 2), ("GPU-c", "4g.20gb", 0, 4)]),
             ("pod-2", [("GPU-c", "4g.20gb", 0, 4)]), ("pod-2", [("GPU-a", "3g.20gb", 4, 4)])]
    for uid, res in steps:
        assert e.pod_set_mig(uid, res) == o.pod_set_mig(uid, res) == 0
        same_state(e, o, [node])
        p = pod("probe", 9000)
        r, g = o.filter(p), e.filter(p)
        assert parity.compare([("filter", p)], [r], [g]) == [], (uid, res)
        if g["status"] == 0:
            e.pod_delete("probe")
            o.pod_delete("probe")
        same_state(e, o, [node])
    e.close()
    o.close()


def make_cluster(rng, n_nodes):
    nodes = []
    for i in range(n_nodes):
        kind = rng.random()
        devs = []
        nd = rng.choice([1, 2, 4, 8])
        for d in range(nd):
            if kind < 0.45 or (kind < 0.8 and rng.random() < 0.5):
                devs.append(gpu("GPU-%d-%d" % (i, d), d, "mig", rng.choice([A100, A100, ODD]), numa=d // 4, health=rng.random() > 0.03))
            else:
                devs.append(gpu("GPU-%d-%d" % (i, d), d, rng.choice(["hami-core", "hami-core", "mps", ""]), numa=d // 4))
        nodes.append({"name": "node-%d" % i, "devices": devs})
    return nodes


def make_pod(rng, uid):
    r = rng.random()
    annos = {}
    if r < 0.25:
        annos["nvidia.com/vgpu-mode"] = "mig"
    elif r < 0.35:
        annos["nvidia.com/vgpu-mode"] = "hami-core"
    if rng.random() < 0.15:
        annos["hami.io/gpu-scheduler-policy"] = rng.choice(["spread", "binpack"])
    if rng.random() < 0.1:
        annos["nvidia.com/numa-bind"] = "true"
    n = rng.choice([1, 1, 1, 2, 2, 3, 4])
    if rng.random() < 0.2:
        p = pod(uid, 0, n, annos, mempct=rng.choice([10, 25, 50, 100]))
    else:
        p = pod(uid, rng.choice([1000, 4000, 5120, 6000, 10240, 12000, 20000, 30000]), n, annos, cores=rng.choice([0, 0, 10, 30]))
    if rng.random() < 0.2:        # a second container, sometimes an init container in front
        p["containers"].append([{"type": "NVIDIA", "nums": rng.choice([1, 2]), "memreq": rng.choice([3000, 9000]), "mempct": 101, "coresreq": 0}])
        if rng.random() < 0.4:
            p["n_init"] = 1
    return p


@pytest.mark.parametrize("seed,node_policy,gpu_policy", [(1, "binpack", "binpack"), (2, "spread", "spread"), (3, "binpack", "spread")])
def test_mig_stream_parity(seed, node_policy, gpu_policy):
    import k8s_vgpu_scheduler_b200.engine as eng
    rng = random.Random(seed)
    nodes = make_cluster(rng, 22)
    e, o = both(node_policy, gpu_policy)
    for n in nodes:
        e.node_upsert(n)
        o.node_upsert(n)
    live, serial, mismatches, n_mig_picks, n_fit, n_nofit = [], 0, [], 0, 0, 0
    for step in range(16):
        ops = []
        for _ in range(rng.randint(20, 60)):
            if live and rng.random() < 0.22:
                ops.append(("delete", live.pop(rng.randrange(len(live)))))
            else:
                serial += 1
                ops.append(("filter", make_pod(rng, "pod-%d" % serial)))
        ref = []
        for kind, arg in ops:
            ref.append(o.filter(arg) if kind == "filter" else {"status": o.pod_delete(arg)})
        packed = eng.PackedOps(ops)
        e.run_batch(packed)
        got = [e.batch_get(i) for i in range(len(ops))]
        bad = parity.compare(ops, ref, got)
        mismatches += ["step %d: %s" % (step, b) for b in bad]
        assert not mismatches, mismatches[:3]
        same_state(e, o, nodes)
        deleted = {arg for kind, arg in ops if kind == "delete"}
        for (kind, arg), r in zip(ops, ref):
            if kind != "filter":
                continue
            n_fit += r["status"] == 0
            n_nofit += r["status"] == 1
            if r["status"] != 0 or arg["uid"] in deleted:
                continue
            live.append(arg["uid"])
            res = reservations_of(r)
            n_mig_picks += len(res)
            # the informer: most pods get their reservation delivered, some stay fail-closed for a while, a few get a
            # truncated or foreign one
            roll = rng.random()
            if roll < 0.75:
                pass
            elif roll < 0.9:
                continue
            elif roll < 0.95:
                res = res[:-1]
            else:
                res = res + [("GPU-nowhere", "1g.5gb", 0, 1)]
            assert e.pod_set_mig(arg["uid"], res) == o.pod_set_mig(arg["uid"], res) == 0
        same_state(e, o, nodes)
    assert n_mig_picks > 40 and n_fit > 100 and n_nofit > 5, (n_mig_picks, n_fit, n_nofit)
    e.close()
    o.close()


def test_register_annotation_and_pod_annotations_carry_mig_state():
    # the reference-facing ingest path: hs_node_register decodes DeviceInfo.migProfiles from hami.io/node-nvidia-register
    # (devices.go:127-144), a Filter's own annotations come back through hs_pod_ingest the way onAddPod sees the patched
    # pod (scheduler.go:137-171) — against an oracle fed the same facts directly
    import k8s_vgpu_scheduler_b200.engine as eng
    from oracle_abi import abi
    e, o = both()
    reg = json.dumps([
        {"id": "GPU-0", "count": 7, "devmem": 40960, "devcore": 100, "type": "NVIDIA-A100", "numa": 0, "health": True, "mode": "mig",
         "migProfiles": [{"name": p["name"], "memoryMB": p["memory_mb"], "core": p["core"], "sliceCount": p["slice_count"],
                          "placements": [{"start": s, "size": z} for s, z in p["placements"]]} for p in A100]},
        {"id": "GPU-1", "index": 1, "count": 10, "devmem": 40960, "devcore": 100, "type": "NVIDIA-A100", "numa": 0, "health": True, "mode": "hami-core"},
    ])
    assert e.node_register("node1", [("NVIDIA", abi.HS_REG_JSON, reg, None)]) == 2
    node = {"name": "node1", "devices": [gpu("GPU-0", 0, "mig", A100), gpu("GPU-1", 1, "hami-core")]}
    o.node_upsert(node)
    seq = [pod("a", 9000, 2, {"nvidia.com/vgpu-mode": "mig"}), pod("b", 3000, 1, {"nvidia.com/vgpu-mode": "mig"}),
           pod("c", 18000, 1, {"nvidia.com/vgpu-mode": "mig"}), pod("d", 18000, 1, {"nvidia.com/vgpu-mode": "mig"})]
    for p in seq:
        r, g = o.filter(p), e.filter(p)
        assert parity.compare([("filter", p)], [r], [g]) == [], p["uid"]
        same_state(e, o, [node])
        if g["status"] != 0:
            continue
        # the informer: the patched pod comes back with the annotations Filter returned
        annos = dict(g["annotations"])
        annos["hami.io/vgpu-node"] = "node1"
        assert e.pod_ingest(p["uid"], "node1", annos) == 0
        assert o.pod_set_mig(p["uid"], reservations_of(r)) == 0
        same_state(e, o, [node])
    occ, healthy = e.snapshot_node_mig("node1")[0]
    assert healthy == 1 and occ != 0
    # a mangled annotation decodes to no reservation: the GPU fails closed again (scheduler.go:760-766,799-802)
    annos = {"hami.io/vgpu-devices-allocated": "GPU-0,NVIDIA,10240,28:;", "hami.io/vgpu-mig-allocations": "[{broken"}
    assert e.pod_ingest("a", "node1", annos) == 0
    assert e.snapshot_node_mig("node1")[0][1] == 0
    e.close()
    o.close()
