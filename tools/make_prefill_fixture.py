#!/usr/bin/env python3
"""Writes tests/golden/c3_prefill_pods.json.gz: the pod cache (podManager content) of the headline C3 cluster after the
bench's prefill phase, computed by the CPU ORACLE alone.  `bench.py --impl reference` starts from it, so the reference
arm is timed on the same cluster state as the GPU arm (same 16,000 prefill filters) without spending minutes of host
time on the prefill in every run.  The GPU engine reaches the identical state on its own (bench.py's parity digest).

  python tools/make_prefill_fixture.py [--workload c3] [--threads N]
"""
import argparse
import gzip
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import bench as B  # noqa: E402
import oracle_abi as oa  # noqa: E402
import k8s_vgpu_scheduler_b200.engine as eng  # noqa: E402  (PackedOps only marshals)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--threads", type=int, default=B.ncpu())
    args = ap.parse_args()
    B.WORKLOAD = args.workload
    wl = B.WORKLOADS[args.workload]
    cfg, nodes, fill, _ = B.make_stream(1, 3, wl["nodes"], wl["fill"], wl["batch"])
    o = oa.Oracle(cfg["node_policy"], cfg["gpu_policy"], faithful=False, threads=args.threads)
    for n in nodes:
        o.node_upsert(n)
    t0 = time.time()
    step = 1000
    for s in range(0, len(fill), step):
        o.run_ops(eng.PackedOps(fill[s:s + step]))
        print("prefill %d / %d ops, %.0f s" % (min(s + step, len(fill)), len(fill), time.time() - t0), flush=True)
    pods = []
    for uid in B.live_uids([fill]):
        got = o.pod_get(uid)
        if got is not None and got[1]:
            pods.append([uid, got[0], [list(r) for r in got[1]]])
    out = {"workload": args.workload, "nodes": wl["nodes"], "prefill_filters": wl["fill"], "fill_ops": len(fill),
           "seed": "gen.SEED", "pods": pods}
    path = os.path.join(ROOT, "tests", "golden", "%s_prefill_pods.json.gz" % args.workload)
    with gzip.open(path, "wt", compresslevel=9) as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote %s: %d pods, %d bytes" % (path, len(pods), os.path.getsize(path)))


if __name__ == "__main__":
    main()
