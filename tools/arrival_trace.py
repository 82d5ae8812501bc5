#!/usr/bin/env python3
"""Diagnostic: who is the long pole of a decision?  HS_TRACE=1 makes every CTA stamp its arrival at each op's barrier;
this prints, per Filter, how long after the previous barrier completed the CTAs arrived (median / last), which CTA was
last, and whether that CTA had just committed the previous decision."""
import ctypes as C
import os
import statistics
import sys

os.environ["HS_TRACE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import k8s_vgpu_scheduler_b200.engine as eng  # noqa: E402
import k8s_vgpu_scheduler_b200.gen as gen  # noqa: E402

cfg, nodes, ops = gen.config3(n_nodes=5000, n_pods=24000)
e = eng.Engine(cfg["node_policy"], cfg["gpu_policy"])
for n in nodes:
    e.node_upsert(n)
half = len(ops) // 2
e.run_batch(eng.PackedOps(ops[:half]))
part = ops[half:half + 500]                 # one launch (below the pipeline's chunking)
e.run_batch(eng.PackedOps(part))
L = eng.lib()
N = 512 * 320
buf = (C.c_uint64 * N)()
o, st = C.c_uint32(), C.c_uint32()
L.hs_debug_trace.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
assert L.hs_debug_trace(e.h, buf, N, C.byref(o), C.byref(st)) == 0
stride = st.value
rows = []
for r in range(len(part)):
    row = buf[r * stride:(r + 1) * stride]
    arr = [(t, c) for c, t in enumerate(row[:stride - 2]) if t]
    if not arr:
        continue
    rows.append((r, arr, row[stride - 2], row[stride - 1]))
prev_done, prev_win = None, None
gaps_med, gaps_last, last_is_prev_winner, period, tail = [], [], 0, [], []
for r, arr, win, done in rows:
    if prev_done:
        ts = sorted(t for t, _ in arr)
        t_last, c_last = max(arr)
        gaps_med.append((ts[len(ts) // 2] - prev_done) / 1e3)
        gaps_last.append((t_last - prev_done) / 1e3)
        tail.append((done - t_last) / 1e3)
        period.append((done - prev_done) / 1e3)
        if prev_win and c_last == prev_win - 1:
            last_is_prev_winner += 1
    prev_done, prev_win = done, win
n = len(period)
print("filters traced: %d (CTAs %d)" % (n, len(rows[0][1])))
print("period between barrier completions (CTA 0's view): median %.2f us, mean %.2f us" % (statistics.median(period), statistics.mean(period)))
print("after the previous completion: median CTA arrives after %.2f us, LAST CTA after %.2f us (medians over ops)" %
      (statistics.median(gaps_med), statistics.median(gaps_last)))
print("last arrival -> CTA 0 sees completion: median %.2f us" % statistics.median(tail))
print("the last CTA to arrive was the previous winner's CTA in %d of %d ops (%.0f %%)" % (last_is_prev_winner, n, 100.0 * last_is_prev_winner / n))
e.close()
