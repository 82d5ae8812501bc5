"""bench.py --gpus N (N > 1): the sharded path of SURVEY §8e on ONE node, one process per GPU (torchrun).

Workload: C5 — a 100,000-node synthetic inventory (C2's device and request mix) partitioned in contiguous
node blocks across the ranks.  Every rank sees the same pod stream.  Default (fused): ONE persistent decide
kernel per GPU runs the whole batch; per decision the CTAs of a GPU combine through the local L2 slot, the
rank's leader CTA stores the packed (score, global position) key into every peer's mailbox over NVLink
(peer-mapped memory) and polls its own, and the rank that owns the winning node commits — no launch and no
library call per decision.  HS_SHARD_NCCL=1 selects the baseline: per decision a sweep kernel, an NCCL
all-reduce(MAX, uint64) and an owner-commit kernel.  Time = CUDA events, MAX over ranks.
scaling: "strong" (the inventory
and the stream are fixed as N grows).
"""
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_NODES = 100000
BATCH_OPS = 1024
FILL_FILTERS = 4096


def partition(n_nodes, world):
    """contiguous blocks, block not cyclic: 'highest index wins' == highest rank, then highest local index"""
    per = (n_nodes + world - 1) // world
    return [(min(n_nodes, r * per), min(n_nodes, (r + 1) * per)) for r in range(world)]


METRIC = "sched decisions/sec on the 100k-node inventory sharded over N GPUs"


def sharded_mismatches(ops, ref, gathered):
    """Compare a sharded run with the single-scheduler oracle.  gathered: per rank (pos_base, {op: full result of a fit
    this rank owns}, {op: FailedNodes of this shard}, {op: (global position, score bits) seen as HS_FIT_REMOTE})."""
    import parity
    from k8s_vgpu_scheduler_b200 import abi
    bad = []
    for i, ((kind, arg), r) in enumerate(zip(ops, ref)):
        if kind != "filter":
            continue
        owners = [(glo, m[i]) for glo, m, _, _ in gathered if i in m]
        if r["status"] == abi.HS_FIT:
            if len(owners) != 1:
                bad.append("op %d: %d owners" % (i, len(owners)))
                continue
            glo, g = owners[0]
            if (g["node"], glo + g["node_index"], parity.f32bits(g["score"]), g["devices"], g["annotations"]) != \
               (r["node"], r["node_index"], parity.f32bits(r["score"]), r["devices"], r["annotations"]):
                bad.append("op %d differs: %r vs %r" % (i, g["node"], r["node"]))
            for _, _, _, rem in gathered:
                if i in rem and rem[i] != (r["node_index"], parity.f32bits(r["score"])):
                    bad.append("op %d remote view differs" % i)
        elif r["status"] == abi.HS_NOFIT:
            merged = {}
            for _, _, nf, _ in gathered:
                merged.update(nf.get(i, {}))
            if owners or merged != r["failed"]:
                bad.append("op %d no-fit differs" % i)
    return bad


def shard_results(e, n_ops):
    import parity
    from k8s_vgpu_scheduler_b200 import abi
    got = [e.batch_get(i) for i in range(n_ops)]
    mine = {i: g for i, g in enumerate(got) if g["status"] == abi.HS_FIT}
    nofit = {i: g["failed"] for i, g in enumerate(got) if g["status"] == abi.HS_NOFIT}
    remote = {i: (g["node_index"], parity.f32bits(g["score"])) for i, g in enumerate(got) if g["status"] == abi.HS_FIT_REMOTE}
    return mine, nofit, remote


def parity_leg(e, eng, dist, cfg, nodes, lo, history, extra_batch, rank, world, seconds):
    """After the timed steps: one more batch of the stream through the SAME sharded engines (fused path), the first ops
    of it compared in full — node, global position, score bits, device rows, annotation strings, FailedNodes — with
    the CPU oracle started from the ranks' gathered pod caches.  Untimed."""
    import bench as B
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_abi as oa
    pods_here = B.pod_cache(e, B.live_uids(history))
    gathered_pods = [None] * world
    dist.all_gather_object(gathered_pods, pods_here)
    # bound the sample by the oracle's cost: ~30 us per node per thread
    per = 30e-6 * len(nodes) / max(B.ncpu(), 1)
    want = int(max(8, min(256, seconds / per)))
    end, nf = 0, 0
    while end < len(extra_batch) and nf < want:
        nf += extra_batch[end][0] == "filter"
        end += 1
    sample = extra_batch[:end]
    e.run_batch(eng.PackedOps(sample))
    mine, nofit, remote = shard_results(e, len(sample))
    gathered = [None] * world
    dist.all_gather_object(gathered, (lo, mine, nofit, remote))
    out = None
    if rank == 0:
        o = oa.Oracle(cfg["node_policy"], cfg["gpu_policy"], faithful=False, threads=B.ncpu())
        for n in nodes:
            o.node_upsert(n)
        for part in gathered_pods:
            B.inject_pods(o, part, "oracle")
        ref = []
        for kind, arg in sample:
            if kind == "filter":
                ref.append(o.filter(arg))
            else:
                rc = o.pod_delete(arg)
                ref.append({"status": rc, "rc": rc})
        o.close()
        bad = sharded_mismatches(sample, ref, gathered)
        out = {"ops": len(sample), "filters": nf, "deletes": len(sample) - nf, "mismatches": len(bad), "digest_match": not bad,
               "what": "the batch after the timed steps, same sharded engines and path, every field of every decision "
                       "against the CPU oracle started from the ranks' gathered pod caches", "first": bad[:3]}
    return out


def run(args, world, rank, local_rank):
    import torch
    import torch.distributed as dist
    import k8s_vgpu_scheduler_b200.engine as eng
    import k8s_vgpu_scheduler_b200.gen as gen
    import bench as B

    if world != args.gpus:
        raise SystemExit("launch with torchrun --nproc-per-node %d" % args.gpus)
    if world == 1:   # plain `python bench.py --workload c5`: a one-rank group
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29533")
        os.environ.setdefault("RANK", "0")
        os.environ.setdefault("WORLD_SIZE", "1")
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n_nodes = args.nodes if args.nodes != B.N_NODES else N_NODES
    batch = min(args.batch, BATCH_OPS)
    n_pods = FILL_FILTERS + (args.warmup + args.steps + 2) * batch
    cfg, nodes, ops = gen.config5(n_nodes=n_nodes, n_pods=n_pods)
    lo, hi = partition(n_nodes, world)[rank]

    e = eng.Engine(cfg["node_policy"], cfg["gpu_policy"], cuda_device=local_rank, reserve_nodes=hi - lo)
    for n in nodes[lo:hi]:
        e.node_upsert(n)
    uid = [eng.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    e.shard_init(rank, world, uid[0], lo)
    if world > 1 and not os.environ.get("HS_SHARD_NCCL"):   # fused path: exchange the peer-mailbox IPC handles
        boxes = [None] * world
        dist.all_gather_object(boxes, e.shard_mailbox())
        e.shard_connect(boxes)

    cnt, cut = 0, 0
    for i, (k, _) in enumerate(ops):
        if k == "filter":
            cnt += 1
            if cnt == FILL_FILTERS:
                cut = i + 1
                break
    fill, rest = ops[:cut], ops[cut:]
    batches = [rest[i * batch:(i + 1) * batch] for i in range(args.warmup + args.steps + 1)]
    packed = [eng.PackedOps(b) for b in batches[:args.warmup + args.steps]]
    e.run_batch(eng.PackedOps(fill))
    for i in range(args.warmup):
        e.run_batch(packed[i])

    sampler = B.ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = e.launch_count()
    dev_ms, e2e_s, decisions, fits, nofits, remote = [], [], 0, 0, 0, 0
    dist.barrier()
    torch.cuda.synchronize()
    for i in range(args.warmup, args.warmup + args.steps):
        t0 = time.perf_counter()
        e.run_batch(packed[i])
        nf, nn, _ = e.batch_summary(digest=False)
        t1 = time.perf_counter()
        e2e_s.append(t1 - t0)
        dev_ms.append(e.last_device_ms())
        decisions += sum(1 for k, _ in batches[i] if k == "filter")
        fits += nf
        nofits += nn
    torch.cuda.synchronize()
    dist.barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = e.launch_count() - launches0
    # MAX over ranks of the device time and of the end-to-end time; fits are owned by exactly one rank each
    t = torch.tensor([sum(dev_ms), sum(e2e_s)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    c = torch.tensor([fits, nofits], dtype=torch.int64, device="cuda")
    dist.all_reduce(c, op=dist.ReduceOp.SUM)
    total_dev_s, total_e2e_s = float(t[0]) / 1e3, float(t[1])
    fused = world > 1 and not os.environ.get("HS_SHARD_NCCL")
    mode_desc = ("one persistent decide kernel per GPU; per decision the ranks' leader CTAs exchange the packed arg-best key "
                 "through peer-mapped mailboxes over NVLink (fused, no launch, no NCCL call)") if fused else \
                "per decision sweep -> NCCL allreduce(MAX,uint64) -> owner commit (HS_SHARD_NCCL=1 baseline)"
    mode_note = ("latency-bound: local slot barrier + one NVLink mailbox round trip per decision" if fused
                 else "latency-bound: 3 kernel launches + one 8-byte NCCL allreduce per decision")
    parity_out = None
    if not getattr(args, "no_cpu", False):
        parity_out = parity_leg(e, eng, dist, cfg, nodes, lo, [fill] + batches[:args.warmup + args.steps],
                                batches[args.warmup + args.steps], rank, world, getattr(args, "cpu_seconds", 12.0))
    # the scaling base: the SAME inventory and the same batches on ONE GPU (rank 0, unsharded engine), so the line
    # carries its own 1-GPU number for this workload (bench.py --gpus 1 runs C3, the metric's config)
    single = None
    if world > 1 and rank == 0 and not os.environ.get("HS_NO_SINGLE"):
        e1 = eng.Engine(cfg["node_policy"], cfg["gpu_policy"], cuda_device=local_rank, reserve_nodes=n_nodes)
        for n in nodes:
            e1.node_upsert(n)
        e1.run_batch(eng.PackedOps(fill))
        for i in range(args.warmup):
            e1.run_batch(packed[i])
        ms1 = 0.0
        for i in range(args.warmup, args.warmup + args.steps):
            e1.run_batch(packed[i])
            ms1 += e1.last_device_ms()
        single = {"value": round(decisions / (ms1 / 1e3), 1), "unit": B.UNIT, "n_gpus": 1,
                  "ms_per_step": round(ms1 / args.steps, 4), "what": "same inventory and batches, unsharded, on rank 0's GPU"}
        e1.close()
    if world > 1:
        dist.barrier()
    if rank == 0:
        peak, peak_src = B.measured_peak()
        bytes_alg = decisions * (n_nodes * 8 * 32)
        out = {
            "metric": METRIC if world > 1 else B.METRIC, "value": round(decisions / total_dev_s, 1), "unit": B.UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * total_dev_s / args.steps, 4),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int32+f32", "data": "synthetic",
            "config": {"workload": "C5: %d-node inventory (C2 device/request mix) sharded in contiguous blocks over %d GPUs; "
                                   "%s" % (n_nodes, world, mode_desc),
                       "ops_per_step": batch, "decisions_timed": decisions, "fits": int(c[0]),
                       "no_fits": int(c[1]) // world, "l2": "each rank's shard (%.1f MB) is L2 resident" % ((hi - lo) * 256 / 1e6)},
            "e2e": {"value": round(decisions / total_e2e_s, 1), "unit": B.UNIT,
                    "h2d_bytes_per_step": batch * 160, "d2h_bytes_per_step": batch * 160},
            "gpu_launches": int(launches), "clocks": clocks, "single_gpu_same_workload": single, "parity": parity_out,
            "roofline": {"bound": "hbm", "achieved": round(bytes_alg / total_dev_s / 1e9, 1), "peak": peak * world,
                         "unit": "GB/s", "frac": round(bytes_alg / total_dev_s / 1e9 / (peak * world), 4), "traffic": None,
                         "peak_source": peak_src + " x n_gpus", "kernel": "hs_decide_kernel" if fused else "hs_sweep_kernel",
                         "note": mode_note},
        }
        print(json.dumps(out))
    e.close()
    dist.destroy_process_group()
