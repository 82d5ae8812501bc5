#!/usr/bin/env python3
"""Diagnostic: where does one decision's time go inside hs_decide_kernel?  Runs a stream with
HS_TIMING=1 and prints block 0's cumulative SM clocks per phase (not a bench number)."""
import ctypes as C
import os
import sys

os.environ["HS_TIMING"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import k8s_vgpu_scheduler_b200.engine as eng  # noqa: E402
import k8s_vgpu_scheduler_b200.gen as gen  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "c2"
n_nodes = int(sys.argv[2]) if len(sys.argv) > 2 else 5000
n_pods = int(sys.argv[3]) if len(sys.argv) > 3 else 6000
cfg, nodes, ops = gen.CONFIGS[workload](n_nodes=n_nodes, n_pods=n_pods)
e = eng.Engine(cfg["node_policy"], cfg["gpu_policy"])
for n in nodes:
    e.node_upsert(n)
half = len(ops) // 2
e.run_batch(eng.PackedOps(ops[:half]))
L = eng.lib()
L.hs_debug_timing.argtypes = [C.c_void_p, C.POINTER(C.c_ulonglong)]
t0 = (C.c_ulonglong * 8)()
L.hs_debug_timing(e.h, t0)
part = ops[half:]
e.run_batch(eng.PackedOps(part))
ms = e.last_device_ms()
nf, nn, _ = e.batch_summary()
t1 = (C.c_ulonglong * 8)()
L.hs_debug_timing(e.h, t1)
names = ["stage", "release", "sweep", "reduce", "barrier", "winread", "commit/explain"]
d = [t1[i] - t0[i] for i in range(7)]
tot = sum(d)
nfilt = sum(1 for k, _ in part if k == "filter")
print("ops %d filters %d fits %d nofits %d device_ms %.3f  -> %.2f us/op, %.2f us/decision" %
      (len(part), nfilt, nf, nn, ms, 1e3 * ms / len(part), 1e3 * ms / nfilt))
for nm, v in zip(names, d):
    print("  %-15s %12d clk  %5.1f%%  %.2f us/decision @1.965GHz" % (nm, v, 100.0 * v / tot, v / 1965.0 / nfilt))
# windows (hs_window.cuh): the resolver's clocks and the window counters
L.hs_debug_window_stats.argtypes = [C.c_void_p, C.POINTER(C.c_ulonglong), C.POINTER(C.c_uint32)]
t16, st = (C.c_ulonglong * 16)(), (C.c_uint32 * 8)()
L.hs_debug_window_stats(e.h, t16, st)
if st[0]:
    print("windows %d (both batches), ops decided in windows %d (%.1f per window), stops: no-fit %d, tracker full %d; winners loaded late %d"
          % (st[0], st[1], st[1] / st[0], st[2], st[3], st[4]))
    for nm, i in (("whole windows (any CTA 0 view)", 7), ("stage+sweep", 8), ("wait arrivals", 9), ("pre-track", 10), ("ops in order: rest", 11),
                   ("  fit of the tracked nodes", 12), ("  scan + block max", 13), ("  commit + node score", 14), ("  release ops", 15)):
        print("  %-32s %12d clk  %.2f us/window @1.965GHz" % (nm, t16[i], t16[i] / 1965.0 / st[0]))
L.hs_debug_host_seconds.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
hs = (C.c_double * 4)()
L.hs_debug_host_seconds(e.h, hs)
print('host seconds (cumulative, both batches): compile %.4f device+copies %.4f absorb %.4f summary %.4f  over %d ops' % (hs[0], hs[1], hs[2], hs[3], len(ops)))
e.close()
