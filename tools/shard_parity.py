#!/usr/bin/env python3
"""Multi-GPU parity check (run under torchrun, one rank per GPU): a small cluster sharded over the ranks,
every decision compared with the single-scheduler oracle — node, score bits, device rows, annotations.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/shard_parity.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench_multi  # noqa: E402
import parity  # noqa: E402
import k8s_vgpu_scheduler_b200.engine as eng  # noqa: E402
import k8s_vgpu_scheduler_b200.gen as gen  # noqa: E402
from k8s_vgpu_scheduler_b200 import abi  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    bad_total = 0
    for name, kw in [("c2", dict(n_nodes=333, n_pods=1500)), ("c3", dict(n_nodes=200, n_pods=1200)),
                     ("c4", dict(n_nodes=300, n_pods=1500)), ("c1", dict(n_pods=300, departures=0.3))]:
        cfg, nodes, ops = gen.CONFIGS[name](**kw)
        lo, hi = bench_multi.partition(len(nodes), world)[rank]
        e = eng.Engine(cfg["node_policy"], cfg["gpu_policy"], cuda_device=local)
        for n in nodes[lo:hi]:
            e.node_upsert(n)
        uid = [eng.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        e.shard_init(rank, world, uid[0], lo)
        if world > 1 and not os.environ.get("HS_SHARD_NCCL"):   # fused path: exchange the peer-mailbox IPC handles
            boxes = [None] * world
            dist.all_gather_object(boxes, e.shard_mailbox())
            e.shard_connect(boxes)
        got = []
        for s in range(0, len(ops), 500):
            part = ops[s:s + 500]
            e.run_batch(eng.PackedOps(part))
            got += [e.batch_get(i) for i in range(len(part))]
        # each decision is owned by exactly one rank: gather the owners' full results on rank 0
        mine = {i: g for i, g in enumerate(got) if g["status"] == abi.HS_FIT}
        nofit = {i: g["failed"] for i, g in enumerate(got) if g["status"] == abi.HS_NOFIT}
        remote = {i: (g["node_index"], parity.f32bits(g["score"])) for i, g in enumerate(got) if g["status"] == abi.HS_FIT_REMOTE}
        gathered = [None] * world
        dist.all_gather_object(gathered, (lo, mine, nofit, remote))
        if rank == 0:
            ref, _ = parity.run_oracle(cfg, nodes, ops, snapshot=False)
            bad = bench_multi.sharded_mismatches(ops, ref, gathered)
            print("shard parity %-3s world %d: %d ops, %d mismatches %s" % (name, world, len(ops), len(bad), bad[:3]))
            bad_total += len(bad)
        e.close()
        dist.barrier()
    dist.destroy_process_group()
    if rank == 0 and bad_total:
        sys.exit(1)


if __name__ == "__main__":
    main()
