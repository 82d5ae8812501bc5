#!/usr/bin/env python3
"""bench.py — scheduling decisions/sec of the B200 Filter/Score engine on BASELINE.json's headline config.

Workload (config.workload): C3 = 5,000 nodes x 8 B200-class 180 GB devices with NVLink pair scores
(20 % PCIe-only nodes), topology-aware 2/4/8-GPU pods, 25 % numa-bind, departures holding ~70 %
occupancy.  A "step" is one batch of the pod stream (filters + deletes) run through
the persistent decide kernel; a decision is one Filter outcome including its commit.

  value  = decisions/s with the batch's request descriptors already resident in HBM (CUDA events around
           the kernel on the engine's stream, hs_last_device_ms);
  e2e    = decisions/s through the C ABI calls a scheduler makes (hs_run_batch_async with host buffers: request
           compile, H2D, kernel, D2H of the decisions — a pipeline over 512-op chunks) plus reading every
           result back as the annotation map / FailedNodes (hs_batch_summary, which follows the pipeline);
  roofline = algorithmic inventory bytes per launch / launch duration against the measured HBM copy peak;
  cpu_baseline = the CPU oracle (C++ restatement of the reference) timed on this box's host cores on a
           bounded sample of the same stream, from the same cluster state, with a digest parity check.

`--impl reference` times the reference's CPU algorithm (oracle in its faithful shape: usage rebuilt from
the whole pod cache on every Filter, one thread per node chunk) instead; the Go binary itself cannot be
built here or on the box (no Go toolchain).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

N_NODES = 5000
BATCH_OPS = 4096
FILL_FILTERS = 16000
WORKLOAD = "c3"
METRIC = "sched decisions/sec at 5k nodes x 8 GPUs"
UNIT = "decisions/s"

# BASELINE.json's configs at their stated sizes: nodes, prefill filters (enough to reach the generator's steady
# occupancy), ops per step, and what the line's config.workload says
WORKLOADS = {
    "c1": dict(nodes=4, fill=200, batch=1000, gen=dict(departures=0.3),
               desc="C1: 4 nodes x 8 x 80 GB, identical 3000 MiB / 30-core pods, binpack/binpack, 30% departures"),
    "c2": dict(nodes=1000, fill=8000, batch=4096, gen={},
               desc="C2: 1,000 nodes x 8 x 80 GB, mixed 1-8 GPU pods (gpumem / gpumem-percentage), spread/spread, ~70% occupancy"),
    "c3": dict(nodes=5000, fill=16000, batch=4096, gen={},
               desc="C3: 5000 nodes x 8 x 180 GB, NVLink pair scores (20% PCIe nodes), topology-aware 2/4/8-GPU pods, 25% numa-bind, ~70% occupancy"),
    "c4": dict(nodes=20000, fill=120000, batch=4096, gen={},
               desc="C4: 20,000 heterogeneous nodes (NVIDIA x4 models / Ascend910B3 / MLU / DCU), type and UUID use/nouse selectors, 2% unhealthy, ~60% occupancy"),
}


def ncpu():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    def __init__(self, index=0):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for i, nm in enumerate(names):
                if f[4 + i].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def make_stream(steps, warmup, n_nodes, fill_filters, batch_ops):
    import k8s_vgpu_scheduler_b200.gen as gen
    # enough filters for the fill phase plus (warmup + steps) batches; deletes ride along in the stream
    n_pods = fill_filters + (warmup + steps + 1) * batch_ops
    kw = dict(WORKLOADS.get(WORKLOAD, {}).get("gen", {}))
    if WORKLOAD != "c1":
        kw["n_nodes"] = n_nodes
    cfg, nodes, ops = gen.CONFIGS[WORKLOAD](n_pods=n_pods, **kw)
    # split: fill = prefix holding `fill_filters` filter ops
    cnt, cut = 0, 0
    for i, (k, _) in enumerate(ops):
        if k == "filter":
            cnt += 1
            if cnt == fill_filters:
                cut = i + 1
                break
    fill, rest = ops[:cut], ops[cut:]
    batches = [rest[i * batch_ops:(i + 1) * batch_ops] for i in range(warmup + steps)]
    return cfg, nodes, fill, batches


def live_uids(op_lists):
    """uids filtered and not deleted since, in stream order (whether each one holds devices is the pod cache's business)"""
    live = {}
    for ops in op_lists:
        for kind, arg in ops:
            if kind == "filter":
                live[arg["uid"]] = True
            else:
                live.pop(arg, None)
    return list(live)


def pod_cache(engine, uids):
    """podManager's content for `uids` (hs_pod_get): [(uid, node, [(uuid, slots, mem, cores)])]"""
    out = []
    for u in uids:
        got = engine.pod_get(u)
        if got is not None and got[1]:
            out.append((u, got[0], got[1]))
    return out


def inject_pods(target, pods, kind):
    """Recreate a pod cache (and with it the usage of every device) in a fresh oracle / engine, pod by pod, under the
    pods' own uids — so the deletes of the measured stream release exactly what they release in the timed engine."""
    for uid, node, rows in pods:
        for uuid, slots, mem, cores in rows:
            rc = target.usage_apply(uid, node, uuid, slots, mem, cores)
            assert rc == 0, (kind, uid, rc)


def algorithmic_bytes(ops, n_nodes):
    """per decision N_candidates x D x 32 B (+ N x 128 B pair tile when topology-aware)
    + nums x 16 B commit."""
    total = 0
    for kind, pod in ops:
        if kind != "filter":
            continue
        total += n_nodes * 8 * 32
        if "topology-aware" in (pod.get("annotations") or {}).get("hami.io/gpu-scheduler-policy", ""):
            total += n_nodes * 128
        total += 16 * sum(r["nums"] for c in pod["containers"] for r in c)
    return total


def decide_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the decide kernel, from the committed ncu --set full
    capture (profiles/r02_decide_fast_ncu_summary.json); None when the summary is missing."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_decide_fast_ncu_summary.json")) as f:
            return int(json.load(f)["dram_bytes_per_launch"])
    except (OSError, KeyError, ValueError):
        return None


def pct(xs, q):
    xs = sorted(xs)
    if not xs:
        return None
    return xs[min(len(xs) - 1, int(q * (len(xs) - 1) + 0.5))]


def latency_leg(e, eng, ops, n_filters):
    """p50 / p99 of ONE Scheduler.Filter (scheduler.go:1067-1154) as an extender sees it: hs_filter per pod, host
    submit -> result visible (perf_counter around the C-ABI call; the pod struct is marshalled beforehand, as the cgo
    shim would have it).  Deletes of the stream are applied in between (untimed) so the occupancy holds."""
    import ctypes as C
    from k8s_vgpu_scheduler_b200 import abi
    L = e.L
    res = abi.hs_filter_result()
    lat, fits = [], 0
    for kind, arg in ops:
        if len(lat) >= n_filters:
            break
        if kind != "filter":
            e.pod_delete(arg)
            continue
        p, keep = abi.build_pod(arg)
        t0 = time.perf_counter()
        rc = L.hs_filter(e.h, C.byref(p), None, 0, C.byref(res))
        lat.append(time.perf_counter() - t0)
        fits += rc == 0
    us = [1e6 * x for x in lat]
    return {"calls": len(us), "fits": fits, "p50_us": round(pct(us, 0.50), 1), "p99_us": round(pct(us, 0.99), 1),
            "max_us": round(max(us), 1), "mean_us": round(statistics.mean(us), 1),
            "what": "single hs_filter calls (one Scheduler.Filter each: compile, launch, commit, result assembly), host "
                    "perf_counter submit -> result visible; deletes of the stream applied between calls"}


def run_ours(args):
    import torch  # device memory / streams / distributed are torch's job on the multi-GPU path
    import k8s_vgpu_scheduler_b200.engine as eng
    import oracle_abi as oa

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 or args.gpus > 1 or args.workload == "c5":
        import bench_multi
        return bench_multi.run(args, world, rank, local_rank)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local_rank)

    extra = 3 if not args.no_latency else 0          # batches after the timed ones feed the single-call latency leg
    cfg, nodes, fill, batches = make_stream(args.steps + extra, args.warmup, args.nodes, args.fill, args.batch)
    e = eng.Engine(cfg["node_policy"], cfg["gpu_policy"], cuda_device=local_rank)
    for n in nodes:
        e.node_upsert(n)
    fill_packed = eng.PackedOps(fill)
    packed = [eng.PackedOps(b) for b in batches[:args.warmup + args.steps]]
    e.run_batch(fill_packed)
    fill_fit, fill_nofit, _ = e.batch_summary()

    for i in range(args.warmup):
        e.run_batch(packed[i])
        e.batch_summary()
    # podManager's content at the start of the measured stream, for the CPU baseline / parity sample
    pods0 = pod_cache(e, live_uids([fill] + batches[:args.warmup])) if not args.no_cpu else None

    sampler = ClockSampler(local_rank)
    sampler.start()
    host0 = e.host_seconds()
    launches0 = e.launch_count()
    bytes0 = e.copy_bytes()
    dev_ms, e2e_s, fits, nofits, decisions, bytes_alg = [], [], 0, 0, 0, 0
    gaps_us = []
    torch.cuda.synchronize()
    for i in range(args.warmup, args.warmup + args.steps):
        t0 = time.perf_counter()
        e.run_batch_async(packed[i])                # returns once the pipeline runs; results are read chunk by chunk
        nf, nn, _ = e.batch_summary(digest=False)   # materialise every result: node, rows, annotation map / FailedNodes
        t1 = time.perf_counter()
        e2e_s.append(t1 - t0)
        dev_ms.append(e.last_device_ms())
        fits += nf
        nofits += nn
        nfilt = sum(1 for k, _ in batches[i] if k == "filter")
        decisions += nfilt
        bytes_alg += algorithmic_bytes(batches[i], args.nodes)
        ts = sorted(t for t in e.decision_ns(len(batches[i])) if t)
        gaps_us += [(b - a) / 1e3 for a, b in zip(ts, ts[1:])]
    torch.cuda.synchronize()
    clocks = sampler.stop()
    launches = e.launch_count() - launches0
    host1 = e.host_seconds()
    bytes1 = e.copy_bytes()
    host_ms = [round(1e3 * (b - a) / args.steps, 3) for a, b in zip(host0, host1)]

    total_dev_s = sum(dev_ms) / 1e3
    value = decisions / total_dev_s
    e2e_val = decisions / sum(e2e_s)
    peak, peak_src = measured_peak()
    achieved = bytes_alg / total_dev_s / 1e9
    wl = WORKLOADS[WORKLOAD]
    out = {
        "metric": METRIC if WORKLOAD == "c3" else "sched decisions/sec on BASELINE config %s (%d nodes x 8)" % (WORKLOAD.upper(), args.nodes),
        "value": round(value, 1), "unit": UNIT, "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(statistics.mean(dev_ms), 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int32+f32", "data": "synthetic",
        "config": {"workload": wl["desc"] if args.nodes == wl["nodes"] else "%s stream on %d nodes (scaled run)" % (WORKLOAD, args.nodes),
                   "ops_per_step": args.batch, "decisions_timed": decisions, "fits": fits, "no_fits": nofits,
                   "prefill_filters": args.fill, "prefill_fits": fill_fit,
                   "l2": "inventory (%.2f MB) is L2/register resident by design; no flush: each decision depends on the "
                         "previous commit" % (args.nodes * 256 / 1e6)},
        "e2e": {"value": round(e2e_val, 1), "unit": UNIT, "h2d_bytes_per_step": int((bytes1[0] - bytes0[0]) // args.steps),
                "d2h_bytes_per_step": int((bytes1[1] - bytes0[1]) // args.steps), "bytes": "measured at the copies (hs_debug_copy_bytes)",
                "ms_per_step": round(1e3 * statistics.mean(e2e_s), 3),
                "host_ms_per_step": {"compile": host_ms[0], "device_and_copies": host_ms[1], "absorb": host_ms[2],
                                     "materialise": host_ms[3]}},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(achieved / peak, 4), "traffic": decide_traffic(), "peak_source": peak_src,
                     "kernel": "hs_decide_fast_kernel" if WORKLOAD in ("c1", "c2", "c3") else
                               "hs_decide_fast_kernel, and hs_decide_kernel for the ranges that hold general-path pods",
                     "note": "algorithmic bytes = decisions x (N x 8 x 32 B + N x 128 B topology tile + nums x 16 B); "
                             "the inventory is L2/register resident, HBM evidence is the scan probe (profiles/)"},
    }
    if gaps_us:
        out["latency"] = {"in_batch_completion_gap_us": {"p50": round(pct(gaps_us, 0.5), 2), "p99": round(pct(gaps_us, 0.99), 2),
                                                         "max": round(max(gaps_us), 2),
                                                         "what": "time between consecutive decisions becoming final on the device inside hs_run_batch (%globaltimer)"}}
    if not args.no_latency:
        lat_ops = [op for b in batches[args.warmup + args.steps:] for op in b]
        out.setdefault("latency", {})["filter"] = latency_leg(e, eng, lat_ops, args.latency_calls)

    if not args.no_scan:
        out["scan_hbm"] = scan_probe(args.scan_nodes, peak)
    if not args.no_cpu:
        out["cpu_baseline"] = cpu_baseline(cfg, nodes, pods0, batches[args.warmup], eng, oa, args, local_rank)
        if "latency" in out and "filter" in out["latency"] and out["cpu_baseline"].get("filter_ms"):
            out["latency"]["cpu_oracle_filter_ms"] = out["cpu_baseline"].pop("filter_ms")
    e.close()
    print(json.dumps(out))


def scan_probe(n_nodes, peak):
    """The inventory scan in isolation on an inventory larger than L2: hs_sweep_kernel timed with CUDA
    events by the hs_scan_probe tool (same library), on the idle homogeneous inventory AND on a realistic one (C4's
    vendor / type mix, ~70 % of the devices in use, 2 % unhealthy; requests with a type selector, a UUID list, numa-bind).
    algorithmic bytes = nodes x 8 x 32 B."""
    exe = os.path.join(ROOT, "k8s-vgpu-scheduler_b200", "hs_scan_probe")
    tma = os.environ.get("HS_SWEEP_TMA", "1")[:1] != "0"      # the library's default: inventory staged by cp.async.bulk
    out = {"kernel": "hs_sweep_tma_kernel" if tma else "hs_sweep_kernel", "nodes": n_nodes, "peak": peak, "unit": "GB/s"}
    try:
        for key, flag in (("idle", []), ("realistic", ["--realistic"])):
            proc = subprocess.run([exe] + flag + [str(n_nodes)], capture_output=True, text=True, timeout=600)
            if proc.returncode != 0 or not proc.stdout.strip():
                raise RuntimeError("hs_scan_probe %s exited %d: %s" % (" ".join(flag), proc.returncode, proc.stderr.strip()[-200:]))
            lines = proc.stdout.strip().splitlines()
            rows = [json.loads(x) for x in lines]
            out[key] = [{"request": r["request"], "ms": r["ms"], "achieved": r["algorithmic_gbps"],
                         "frac": round(r["algorithmic_gbps"] / peak, 4), "fits": r["fits"]} for r in rows]
            out["inventory_mb"] = rows[0]["inventory_mb"]
        # the headline scan figure is the REALISTIC inventory's plain request; the idle one stands beside it
        out["achieved"] = out["realistic"][0]["achieved"]
        out["frac"] = out["realistic"][0]["frac"]
        out["ms"] = out["realistic"][0]["ms"]
        out["request"] = out["realistic"][0]["request"]
    except Exception as ex:  # the probe is an extra; the bench line must still be printed
        out["error"] = str(ex)
    return out


def b2_baseline(cfg, nodes, pods0, probe_ops, sample_ops, e2, oa, threads):
    """B2 of BASELINE.md §3: the lean CPU restatement (oracle/b2_lean.cc — SoA inventory, no strings, no copies, cached node
    scores, order-free verdict, a persistent thread pool) on the SAME sample from the same pod cache; its decisions (node,
    score bits, device picks) are compared with the engine's results for the sample.  None when the stream holds request
    shapes B2 does not handle (type / UUID selectors, other vendors, several containers)."""
    from k8s_vgpu_scheduler_b200 import abi
    import struct
    known = {uid: (node, rows) for uid, node, rows in pods0}

    def compile_ops(ops, index_of):
        out = []
        for kind, arg in ops:
            if kind == "filter":
                op = oa.b2_compile(arg, cfg)
                if op is None:
                    return None
                index_of[arg["uid"]] = len(out)
            else:
                op = oa.b2_op()
                if arg in index_of:                      # a pod filtered earlier in this stream
                    op.kind, op.rel_op = 1, index_of.pop(arg)
                elif arg in known:                       # a pod of the cache the sample started from
                    node, rows = known.pop(arg)
                    op.kind, op.rel_node = 2, b2.node_index[node]
                    for uuid, slots, mem, cores in rows:
                        d = b2.dev_slot[uuid][1]
                        op.rel_rows[d][0], op.rel_rows[d][1], op.rel_rows[d][2] = max(slots, 1), mem, cores
                else:                                    # never held anything
                    op.kind, op.rel_node = 2, -1
            out.append(op)
        return out

    b2 = oa.B2(nodes, threads)
    try:
        for uid, node, rows in pods0:
            for uuid, slots, mem, cores in rows:
                b2.usage_add(uuid, slots, mem, cores)
        # the probe ops ran before the sample on the oracle and on the engine: same here (one stream, so in-run releases resolve)
        stream = compile_ops(list(probe_ops) + list(sample_ops), {})
        if stream is None:
            return None
        n_probe = len(probe_ops)
        sec_all, out = b2.run(stream)           # the whole stream once: the decisions that are checked
        rebased = None
        timings = {}
        for T in sorted({1, threads}):         # timing: the sample alone, from the state after the probe, at 1 thread and at all cores
            b2s = oa.B2(nodes, T)
            for uid, node, rows in pods0:
                for uuid, slots, mem, cores in rows:
                    b2s.usage_add(uuid, slots, mem, cores)
            b2s.run(stream[:n_probe])
            if rebased is None:
                rebased = []
                for op in stream[n_probe:]:
                    if op.kind == 1 and op.rel_op < n_probe:      # releases a probe-phase decision: an explicit (node, rows) release
                        src = out[op.rel_op]
                        q = oa.b2_op()
                        q.kind, q.rel_node = 2, src.node
                        for k in range(src.n_picks):
                            d = src.picks[k]
                            q.rel_rows[d][0] += 1; q.rel_rows[d][1] += src.memreq[k]; q.rel_rows[d][2] += src.cores
                        rebased.append(q)
                    elif op.kind == 1:
                        q = oa.b2_op()
                        q.kind, q.rel_op = 1, op.rel_op - n_probe
                        rebased.append(q)
                    else:
                        rebased.append(op)
            timings[T], _ = b2s.run(rebased)
            b2s.close()
        best_T = min(timings, key=timings.get)
        sec = timings[best_T]
        # decisions of the sample against the engine's
        slot = b2.dev_slot
        bad, nfil = 0, 0
        for i, (kind, arg) in enumerate(sample_ops):
            if kind != "filter":
                continue
            nfil += 1
            g = e2.batch_get(i)
            d = out[n_probe + i]
            if g["status"] == abi.HS_NOFIT:
                bad += d.node != -1
                continue
            rows = next(iter(g["devices"].values()))[0]
            same = (d.node == b2.node_index[g["node"]] and
                    struct.pack("f", d.score) == struct.pack("f", g["score"]) and
                    [slot[r[0]][1] for r in rows] == [d.picks[k] for k in range(d.n_picks)])
            bad += not same
        return {"value": round(nfil / sec, 1), "unit": UNIT, "cores": best_T, "kind": "port (lean: SoA, no strings, persistent thread pool)",
                "seconds": round(sec, 3), "filters": nfil, "mismatches_vs_engine": int(bad),
                "by_threads": {str(T): round(nfil / t, 1) for T, t in sorted(timings.items())}}
    finally:
        b2.close()


def cpu_baseline(cfg, nodes, pods0, batch, eng, oa, args, local_rank):
    """The CPU oracle on a bounded, CONTIGUOUS sample of the first timed batch — filters and deletes — from the very
    pod cache the timed engine had at that point, and the same sample through a second engine handle: the FNV digest
    over every result (status, node, score bits, annotation strings, FailedNodes; delete statuses) must be equal."""
    threads = ncpu()
    o = oa.Oracle(cfg["node_policy"], cfg["gpu_policy"], faithful=False, threads=threads)
    for n in nodes:
        o.node_upsert(n)
    inject_pods(o, pods0, "oracle")
    # size the sample for roughly args.cpu_seconds of CPU work: time a few filters first
    k = 0
    nfilt = 0
    while k < len(batch) and nfilt < 8:
        nfilt += batch[k][0] == "filter"
        k += 1
    probe_ops = batch[:k]
    _, _, d_probe_o, sec = o.run_ops(eng.PackedOps(probe_ops))
    per = max(sec / max(nfilt, 1), 1e-6)
    want = int(max(16, args.cpu_seconds / per))
    end, nf2 = k, 0
    while end < len(batch) and nf2 < want:
        nf2 += batch[end][0] == "filter"
        end += 1
    sample_ops = batch[k:end]
    sample = eng.PackedOps(sample_ops)
    nf, nn, digest_o, sec = o.run_ops(sample)
    o.close()
    # the same ops, from the same pod cache, through a second engine handle: digest parity
    e2 = eng.Engine(cfg["node_policy"], cfg["gpu_policy"], cuda_device=local_rank)
    for n in nodes:
        e2.node_upsert(n)
    inject_pods(e2, pods0, "engine")
    e2.run_batch(eng.PackedOps(probe_ops))
    _, _, d_probe_e = e2.batch_summary()
    e2.run_batch(sample)
    gf, gn, digest_e = e2.batch_summary()
    b2 = b2_baseline(cfg, nodes, pods0, probe_ops, sample_ops, e2, oa, threads)
    e2.close()
    n_del = sum(1 for kk, _ in sample_ops if kk != "filter")
    return {"value": round(nf2 / sec, 2), "unit": UNIT, "cores": threads, "kind": "port",
            "sample": "%d contiguous ops of the first timed batch (%d filters, %d deletes) from the same pod cache, oracle in "
                      "its lean shape (incremental usage, %d threads over node chunks)" % (len(sample_ops), nf2, n_del, threads),
            "seconds": round(sec, 2), "filter_ms": round(1e3 * per, 3), "b2_lean": b2,
            "parity": {"digest_match": digest_o == digest_e and d_probe_o == d_probe_e, "ops": len(sample_ops) + len(probe_ops),
                       "deletes": n_del, "fits": [nf, gf], "no_fits": [nn, gn]}}


def run_reference(args):
    """The reference's CPU algorithm for the path (oracle, faithful shape) on the box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import k8s_vgpu_scheduler_b200.engine as eng  # PackedOps only marshals; no engine handle is created
    import oracle_abi as oa
    threads = ncpu()
    step_ops = args.ref_ops
    prefill = args.ref_prefill
    multi = args.gpus > 1 or int(os.environ.get("WORLD_SIZE", "1")) > 1
    if multi:
        # the arm's config at N > 1 is C5 (bench_multi.py): the same 100,000-node inventory and fill, a smaller sample
        import bench_multi as BM
        import k8s_vgpu_scheduler_b200.gen as gen
        step_ops = 4            # one Filter over 100,000 nodes takes the host a few hundred ms: keep the sample bounded
        prefill = 128           # (the GPU arm fills with BM.FILL_FILTERS; the CPU arm could not do that within minutes)
        n_nodes = BM.N_NODES
        cfg, nodes, ops = gen.config5(n_nodes=n_nodes, n_pods=prefill + (args.warmup + args.steps + 1) * step_ops)
        cnt, cut = 0, 0
        for i, (k, _) in enumerate(ops):
            if k == "filter":
                cnt += 1
                if cnt == prefill:
                    cut = i + 1
                    break
        fill, rest = ops[:cut], ops[cut:]
        batches = [rest[i * step_ops:(i + 1) * step_ops] for i in range(args.warmup + args.steps)]
    else:
        # the GPU arm's own stream: same prefill, same step boundaries; each timed step is a bounded sample (the first
        # `step_ops` ops) of that step's 4,096-op batch
        n_nodes = args.nodes
        prefill = args.fill
        cfg, nodes, fill, full = make_stream(args.steps, args.warmup, args.nodes, args.fill, args.batch)
        batches = [b[:step_ops] for b in full]
    o = oa.Oracle(cfg["node_policy"], cfg["gpu_policy"], faithful=False, threads=threads)
    for n in nodes:
        o.node_upsert(n)
    fixture = os.path.join(ROOT, "tests", "golden", "%s_prefill_pods.json.gz" % WORKLOAD)
    state_from = "oracle prefill (lean shape, untimed)"
    loaded = False
    if not multi and os.path.exists(fixture):
        import gzip
        with gzip.open(fixture, "rt") as f:
            fx = json.load(f)
        if fx.get("nodes") == n_nodes and fx.get("prefill_filters") == prefill and fx.get("fill_ops") == len(fill):
            inject_pods(o, [(u, nd, [tuple(r) for r in rows]) for u, nd, rows in fx["pods"]], "oracle")
            state_from = "pod cache after the same %d prefill filters (tests/golden/%s, written by the oracle itself: tools/make_prefill_fixture.py)" % (prefill, os.path.basename(fixture))
            loaded = True
    if not loaded:
        o.run_ops(eng.PackedOps(fill))       # untimed prefill in the lean shape (it is only state set-up)
    o.L.oracle_set_mode(o.h, 1, threads)     # timed: rebuild usage from the pod cache on every Filter
    packed = [eng.PackedOps(b) for b in batches]
    for i in range(args.warmup):
        o.run_ops(packed[i])
    secs, decisions = [], 0
    for i in range(args.warmup, args.warmup + args.steps):
        _, _, _, sec = o.run_ops(packed[i])
        secs.append(sec)
        decisions += sum(1 for k, _ in batches[i] if k == "filter")
    o.close()
    value = decisions / sum(secs)
    sample = ("the first %d ops of each %d-op step of the %s stream, cluster state: %s; oracle (C++ restatement of the "
              "reference) in its faithful shape: usage rebuilt from the whole pod cache per Filter, %d threads over node chunks"
              % (step_ops, step_ops if multi else args.batch, "C5" if multi else WORKLOAD.upper(), state_from, threads))
    out = {"impl": "reference",
           "metric": METRIC if (multi or WORKLOAD == "c3") else "sched decisions/sec on BASELINE config %s (%d nodes x 8)" % (WORKLOAD.upper(), n_nodes),
           "value": round(value, 2), "unit": UNIT, "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * statistics.mean(secs), 3),
           "higher_is_better": True, "scaling": "strong" if multi else "weak", "vs_baseline": None, "dtype": "int32+f32",
           "data": "synthetic",
           "config": {"workload": ("C5: %d-node inventory (C2 device/request mix), one host, no sharding (bounded sample)" % n_nodes) if multi
                      else WORKLOADS[WORKLOAD]["desc"],
                      "ops_per_step": step_ops if multi else args.batch, "prefill_filters": prefill,
                      "sample_ops_per_step": step_ops},
           "cpu_baseline": {"value": round(value, 2), "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
           "e2e": {"value": round(value, 2), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0,
           "note": "Go reference Filter/Score: unavailable on this box (no Go toolchain / module cache)"}
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--nodes", type=int, default=None)
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--fill", type=int, default=None)
    ap.add_argument("--no-latency", action="store_true", help="skip the single-call Filter latency leg")
    ap.add_argument("--latency-calls", type=int, default=10000)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-scan", action="store_true", help="skip the HBM scan probe")
    ap.add_argument("--scan-nodes", type=int, default=2000000)
    ap.add_argument("--ref-ops", type=int, default=96, help="ops per step of the --impl reference arm")
    ap.add_argument("--ref-prefill", type=int, default=1500)
    ap.add_argument("--workload", default="c3", choices=["c1", "c2", "c3", "c4", "c5"],
                    help="diagnostic: another config's stream (c5 = the multi-GPU workload, runnable on 1 GPU for the scaling base)")
    args = ap.parse_args()
    global WORKLOAD
    WORKLOAD = args.workload
    wl = WORKLOADS.get(WORKLOAD if WORKLOAD != "c5" else "c3")
    if args.nodes is None:
        args.nodes = wl["nodes"] if WORKLOAD != "c5" else N_NODES
    if args.batch is None:
        args.batch = wl["batch"]
    if args.fill is None:
        args.fill = wl["fill"]
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
