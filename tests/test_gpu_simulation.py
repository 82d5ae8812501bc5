"""Scheduler.filterSimulation (pkg/scheduler/scheduler.go:1088-1090,1156-1199) through the C ABI:

  hs_filter_dry       decides like Filter but leaves the inventory and the pod cache untouched;
  hs_filter_simulate  the candidate nodes come with the request as register annotations, carry no usage
                      (buildTransientNodeInfo) and are scored in a transient inventory.
The oracle side is the same Filter: for the dry call it decides on a copy-equivalent state (decide, then undo), for
the simulation it is a fresh oracle that holds only the transient nodes.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import oracle_abi as oa  # noqa: E402
import parity  # noqa: E402
from oracle_abi import abi  # noqa: E402
from test_gpu_ingest import registrations  # noqa: E402

pytestmark = pytest.mark.gpu


def same_decision(ref, got):
    if ref["status"] != got["status"]:
        return False
    if ref["status"] == abi.HS_FIT:
        return (ref["node"], parity.f32bits(ref["score"]), ref["devices"], ref["annotations"]) == \
               (got["node"], parity.f32bits(got["score"]), got["devices"], got["annotations"])
    return ref["failed"] == got["failed"]


def test_dry_filter_decides_without_committing():
    import k8s_vgpu_scheduler_b200.engine as eng
    import k8s_vgpu_scheduler_b200.gen as gen
    cfg, nodes, ops = gen.mixed_features(n_nodes=24, n_pods=600)
    warm, probes = ops[:300], [a for k, a in ops[300:] if k == "filter"][:120]
    o = oa.Oracle(cfg.get("node_policy"), cfg.get("gpu_policy"))
    e = eng.Engine(cfg.get("node_policy"), cfg.get("gpu_policy"))
    for n in nodes:
        o.node_upsert(n)
        e.node_upsert(n)
    for kind, arg in warm:
        if kind == "filter":
            o.filter(arg)
        else:
            o.pod_delete(arg)
    e.run_batch(eng.PackedOps(warm))
    before = {n["name"]: e.snapshot(n["name"]) for n in nodes}
    fits = 0
    for pod in probes:
        pod = dict(pod, uid="dry-" + pod["uid"])
        ref = o.filter(pod)
        o.pod_delete(pod["uid"])                                  # undo: the dry call must equal a decision on untouched state
        got = e.filter(pod, dry=True)
        assert same_decision(ref, got), pod["uid"]
        fits += ref["status"] == abi.HS_FIT
    assert fits > 20
    assert {n["name"]: e.snapshot(n["name"]) for n in nodes} == before       # nothing was committed
    assert {n["name"]: o.snapshot(n["name"]) for n in nodes} == before
    # a dry call on a pod the engine already holds does not release it either (simulation runs before TakeAndDeletePod)
    held = next(a for k, a in warm if k == "filter")
    e.filter(held, dry=True)
    assert {n["name"]: e.snapshot(n["name"]) for n in nodes} == before
    e.close()
    o.close()


def test_simulation_on_transient_nodes():
    import k8s_vgpu_scheduler_b200.engine as eng
    import k8s_vgpu_scheduler_b200.gen as gen
    cfg, nodes, ops = gen.config3(n_nodes=40, n_pods=200)
    cfg4, nodes4, ops4 = gen.config4(n_nodes=60, n_pods=200)
    e = eng.Engine(cfg.get("node_policy"), cfg.get("gpu_policy"))
    for n in nodes[:20]:                                           # the live inventory: must stay out of the simulation
        e.node_upsert(n)
    e.run_batch(eng.PackedOps(ops[:100]))
    live_before = {n["name"]: e.snapshot(n["name"]) for n in nodes[:20]}
    pods = [a for k, a in ops if k == "filter"][:40] + [a for k, a in ops4 if k == "filter"][:60]
    for i, pod in enumerate(pods):
        # a different handful of template nodes per request, some of them without any device annotation
        pool = (nodes[20:] if i < 40 else [n for n in nodes4 if not any(d["vendor"] == "MLU" for d in n["devices"])])
        chosen = [pool[(i * 7 + j * 3) % len(pool)] for j in range(1 + i % 6)]
        chosen = list({n["name"]: n for n in chosen}.values())
        sim = [(n["name"], registrations(n), n.get("device_cordon"), n.get("hami_vnpu_core", -1)) for n in chosen]
        if i % 5 == 0:
            sim.insert(len(sim) // 2, ("empty-template-%d" % i, [], None, -1))
        o = oa.Oracle(cfg.get("node_policy"), cfg.get("gpu_policy"))
        for n in chosen:
            o.node_upsert(n)
        ref = o.filter(pod, [s[0] for s in sim])
        o.close()
        got = e.filter_simulate(pod, sim)
        assert same_decision(ref, got), (i, ref, got)
        if i % 5 == 0 and ref["status"] == abi.HS_NOFIT:
            assert got["failed"]["empty-template-%d" % i] == "node unregistered"
    assert {n["name"]: e.snapshot(n["name"]) for n in nodes[:20]} == live_before
    e.close()


def test_simulation_keeps_metax_and_neuron_fit_classes():
    """The transient inventory must judge a vendor with the parent's own Fit class: metax nodes keep their ScoreNode
    term (2000 - loss / score tables) and customFilterRule, AWS Neuron nodes their core masks and graphSelect."""
    import k8s_vgpu_scheduler_b200.engine as eng
    from test_plain_vendors import _metax_cluster, _neuron_cluster
    checked = 0
    for make in (_metax_cluster, _neuron_cluster):
        made = make()
        cfg, nodes, ops = made[:3]
        caps = made[3] if len(made) > 3 else {}
        e = eng.Engine(cfg.get("node_policy"), cfg.get("gpu_policy"))
        pods = [a for k, a in ops if k == "filter"][:80]
        for i, pod in enumerate(pods):
            chosen = [nodes[(i * 5 + j * 3) % len(nodes)] for j in range(2 + i % 5)]
            chosen = list({n["name"]: n for n in chosen}.values())
            sim, twins = [], []
            for n in chosen:
                if n["devices"][0]["vendor"] == "Metax-GPU":
                    regs = [("Metax-GPU", abi.HS_REG_CAPACITY_METAX, n.get("metax_losses"), n.get("metax_scores"), None, len(n["devices"]), 0)]
                elif n["name"] in caps:                      # registered from capacity + instance type: every device healthy
                    counts, total, itype = caps[n["name"]]
                    regs = [("AWSNeuron", abi.HS_REG_CAPACITY_NEURON, itype, None, n["name"], counts, total, 0)]
                    n = dict(n, devices=[dict(d, health=True) for d in n["devices"]])
                else:
                    regs = registrations(n)
                sim.append((n["name"], regs, None, -1))
                twins.append(n)
            o = oa.Oracle(cfg.get("node_policy"), cfg.get("gpu_policy"))
            for n in twins:
                o.node_upsert(n)
            ref = o.filter(pod, [s[0] for s in sim])
            o.close()
            got = e.filter_simulate(pod, sim)
            assert same_decision(ref, got), (make.__name__, i, ref, got)
            checked += ref["status"] == abi.HS_FIT
        e.close()
    assert checked > 40


def test_simulation_honours_the_namespace_quota():
    """filterSimulation goes through the same Fit, fitQuota included: a namespace that has used up its quota is
    refused on transient nodes too, with ResourceQuotaNotFit."""
    import k8s_vgpu_scheduler_b200.engine as eng
    import k8s_vgpu_scheduler_b200.gen as gen
    cfg, nodes, ops = gen.config2(n_nodes=8, n_pods=40)
    e = eng.Engine(cfg.get("node_policy"), cfg.get("gpu_policy"))
    for n in nodes[:4]:
        e.node_upsert(n)
    pod = dict(next(a for k, a in ops if k == "filter"), ns="tight", uid="q-pod")
    pod["containers"] = [[{"key": "NVIDIA", "type": "NVIDIA", "nums": 1, "memreq": 4000, "mempct": 101, "coresreq": 10}]]
    pod["annotations"] = {}
    sim = [(n["name"], registrations(n), None, -1) for n in nodes[4:]]
    assert e.filter_simulate(pod, sim)["status"] == abi.HS_FIT
    e.quota_set("tight", "NVIDIA", 0, 3000)             # less memory than the pod asks for
    got = e.filter_simulate(pod, sim)
    assert got["status"] == abi.HS_NOFIT
    assert all("ResourceQuotaNotFit" in r for r in got["failed"].values()), got["failed"]
    e.quota_del("tight", "NVIDIA", 0)
    assert e.filter_simulate(pod, sim)["status"] == abi.HS_FIT
    e.close()
