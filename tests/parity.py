"""Shared parity harness: run one op stream through the CPU oracle and through the CUDA
engine (C ABI) and compare every decision bit for bit."""
import struct

import oracle_abi as oa
from oracle_abi import abi


def f32bits(x):
    return struct.unpack("I", struct.pack("f", float(x)))[0]


def run_oracle(cfg, nodes, ops, faithful=False, threads=1, snapshot=True):
    o = oa.Oracle(cfg.get("node_policy"), cfg.get("gpu_policy"), faithful=faithful, threads=threads)
    for n in nodes:
        o.node_upsert(n)
    results = []
    for kind, arg in ops:
        if kind == "filter":
            results.append(o.filter(arg))
        else:
            rc = o.pod_delete(arg)
            results.append({"status": rc, "rc": rc})
    snap = {n["name"]: o.snapshot(n["name"]) for n in nodes} if snapshot else None
    o.close()
    return results, snap


def run_engine(cfg, nodes, ops, batch=True, snapshot=True, chunk=None):
    import k8s_vgpu_scheduler_b200.engine as eng
    e = eng.Engine(cfg.get("node_policy"), cfg.get("gpu_policy"))
    for n in nodes:
        e.node_upsert(n)
    results = []
    if batch:
        step = chunk or len(ops) or 1
        for s in range(0, len(ops), step):
            part = ops[s:s + step]
            packed = eng.PackedOps(part)
            e.run_batch(packed)
            for i in range(len(part)):
                results.append(e.batch_get(i))
    else:
        for kind, arg in ops:
            if kind == "filter":
                results.append(e.filter(arg))
            else:
                rc = e.pod_delete(arg)
                results.append({"status": rc, "rc": rc})
    snap = {n["name"]: e.snapshot(n["name"]) for n in nodes} if snapshot else None
    e.close()
    return results, snap


def compare(ops, ref, got, ref_snap=None, got_snap=None, max_report=5):
    """Returns a list of mismatch descriptions (empty == bit-exact parity)."""
    bad = []
    assert len(ref) == len(got) == len(ops)
    for i, ((kind, arg), r, g) in enumerate(zip(ops, ref, got)):
        if kind != "filter":
            if (r["status"] == 0) != (g["status"] == 0):
                bad.append("op %d delete %s: oracle %s engine %s" % (i, arg, r["status"], g["status"]))
            continue
        if r["status"] != g["status"]:
            bad.append("op %d (%s) status: oracle %s engine %s" % (i, arg.get("uid"), r["status"], g["status"]))
        elif r["status"] == abi.HS_FIT:
            if r["node"] != g["node"] or r["node_index"] != g["node_index"]:
                bad.append("op %d node: oracle %s[%s] engine %s[%s]" % (i, r["node"], r["node_index"], g["node"], g["node_index"]))
            elif f32bits(r["score"]) != f32bits(g["score"]):
                bad.append("op %d score bits: oracle %r engine %r" % (i, r["score"], g["score"]))
            elif r["devices"] != g["devices"]:
                bad.append("op %d devices: oracle %r engine %r" % (i, r["devices"], g["devices"]))
            elif r["annotations"] != g["annotations"]:
                bad.append("op %d annotations: oracle %r engine %r" % (i, r["annotations"], g["annotations"]))
        elif r["status"] == abi.HS_NOFIT:
            if r["failed"] != g["failed"]:
                diff = {k: (r["failed"].get(k), g["failed"].get(k)) for k in set(r["failed"]) | set(g["failed"])
                        if r["failed"].get(k) != g["failed"].get(k)}
                bad.append("op %d failedNodes differ (%d nodes), e.g. %r" % (i, len(diff), list(diff.items())[:2]))
        if len(bad) >= max_report:
            break
    if ref_snap is not None and got_snap is not None and ref_snap != got_snap:
        diff = [k for k in ref_snap if ref_snap[k] != got_snap.get(k)]
        bad.append("final usage differs on %d nodes, e.g. %s: %r vs %r" % (len(diff), diff[0], ref_snap[diff[0]], got_snap.get(diff[0])))
    return bad
