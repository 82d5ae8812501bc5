"""The sharded protocol on CPU: world_size 2 over gloo, each rank holds a contiguous block of
nodes in its own oracle, keys are merged with all_reduce(MAX); the decisions must equal the single
scheduler's.  Covers the host logic of bench_multi.py (partition, key order, owner-only commit)."""
import os
import struct
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def f32_ord(x, spread):
    u = struct.unpack("I", struct.pack("f", float(x) + 0.0))[0]
    u = (~u & 0xFFFFFFFF) if u & 0x80000000 else (u | 0x80000000)
    return (~u & 0xFFFFFFFF) if spread else u


def make_key(score, spread, pos):
    # same layout as make_key in csrc/hs_eval.cuh: score(32) | pos+1 (24) | epoch(8)
    return (f32_ord(score, spread) << 32) | (((pos + 1) & 0xFFFFFF) << 8)


def _worker(rank, world, port, q):
    import bench_multi
    import oracle_abi as oa
    import k8s_vgpu_scheduler_b200.gen as gen
    from oracle_abi import abi
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg, nodes, ops = gen.config2(n_nodes=37, n_pods=260)
    lo, hi = bench_multi.partition(len(nodes), world)[rank]
    o = oa.Oracle(cfg["node_policy"], cfg["gpu_policy"])
    for n in nodes[lo:hi]:
        o.node_upsert(n)
    spread = cfg["node_policy"] == "spread"
    SIGN = 1 << 63
    out = []
    for kind, arg in ops:
        if kind != "filter":
            o.pod_delete(arg)          # only the owner knows the pod; NOTFOUND elsewhere
            out.append(None)
            continue
        r = o.filter(arg)              # local decision (commits locally)
        key = make_key(r["score"], spread, lo + r["node_index"]) if r["status"] == abi.HS_FIT else 0
        v = key ^ SIGN
        t = torch.tensor([v - (1 << 64) if v >= SIGN else v], dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)     # signed order of key^SIGN == unsigned order of key
        win = (int(t[0]) & 0xFFFFFFFFFFFFFFFF) ^ SIGN
        if r["status"] == abi.HS_FIT and win != key:
            o.pod_delete(arg["uid"])   # another rank's node won: undo the local commit
        out.append(((win >> 8) & 0xFFFFFF) - 1 if (win >> 8) & 0xFFFFFF else -1)
    if rank == 0:
        q.put(out)
    dist.barrier()
    o.close()
    dist.destroy_process_group()


def test_sharded_protocol_matches_single_scheduler():
    import parity
    import k8s_vgpu_scheduler_b200.gen as gen
    from oracle_abi import abi
    cfg, nodes, ops = gen.config2(n_nodes=37, n_pods=260)
    ref, _ = parity.run_oracle(cfg, nodes, ops, snapshot=False)
    want = [None if k != "filter" else (r["node_index"] if r["status"] == abi.HS_FIT else -1)
            for (k, _), r in zip(ops, ref)]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got == want
    assert any(w is not None and w >= 0 for w in want)


def test_partition_is_contiguous_and_covers():
    import bench_multi
    for n, w in [(100000, 8), (37, 2), (5, 4), (8, 8)]:
        parts = bench_multi.partition(n, w)
        assert parts[0][0] == 0 and parts[-1][1] == n
        assert all(parts[i][1] == parts[i + 1][0] for i in range(w - 1))
        assert all(a <= b for a, b in parts)
