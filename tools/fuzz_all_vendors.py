#!/usr/bin/env python
"""Differential soak of the all-vendors stream (tests/test_gpu_all_vendors.py) over many seeds, in both sweep modes:
oracle vs engine, every field of every decision plus the final usage snapshot.  Needs a GPU.
    python tools/fuzz_all_vendors.py --seeds 100:140 --scale 2 [--streaming]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seeds", default="100:120")
    ap.add_argument("--scale", type=int, default=2)
    ap.add_argument("--streaming", action="store_true", help="HS_NO_RESIDENT=1: the streaming sweep on every launch")
    a = ap.parse_args()
    if a.streaming:
        os.environ["HS_NO_RESIDENT"] = "1"
    import parity
    import test_gpu_all_vendors as t
    lo, hi = (int(x) for x in a.seeds.split(":"))
    bad, total_ops, t0 = [], 0, time.time()
    for seed in range(lo, hi):
        cfg, nodes, ops = t.mixed_cluster(seed, a.scale)
        ref, ref_snap = parity.run_oracle(cfg, nodes, ops)
        got, got_snap = parity.run_engine(cfg, nodes, ops)
        diff = parity.compare(ops, ref, got, ref_snap, got_snap)
        total_ops += len(ops)
        if diff:
            bad.append({"seed": seed, "first": [str(d)[:400] for d in diff[:3]]})
    print(json.dumps({"seeds": [lo, hi], "scale": a.scale, "streaming": a.streaming, "ops": total_ops, "mismatching_seeds": bad,
                      "seconds": round(time.time() - t0, 1)}))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
