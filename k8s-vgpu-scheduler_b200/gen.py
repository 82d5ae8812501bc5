"""Seeded synthetic clusters and pod streams for the BASELINE.json configs.

Pure data generation (numpy PCG64, seed 20260601): both the CPU oracle and the CUDA
engine are fed the very same dicts.  Every config can be scaled down (n_nodes / n_pods)
so the parity tests finish in seconds; bench.py uses the full sizes.
"""
import numpy as np

SEED = 20260601

NVIDIA_MODELS = [
    ("NVIDIA A100-SXM4-80GB", 81920),
    ("NVIDIA H100 80GB HBM3", 81559),
    ("NVIDIA A100-PCIE-40GB", 40960),
    ("NVIDIA L40S", 46068),
]


def _nv_dev(node_i, d, typ, mem, numa, health=True, mode="hami-core"):
    return {"id": "GPU-%06d-%d" % (node_i, d), "index": d, "count": 10, "devmem": mem, "devcore": 100,
            "type": typ, "numa": numa, "mode": mode, "health": health, "vendor": "NVIDIA"}


def _nv_pod(uid, nums, memreq=0, mempct=101, cores=0, annos=None, extra_ctrs=None, n_init=0):
    if mempct == 101 and memreq == 0:
        mempct = 100  # GenerateResourceRequests default (nvidia/device.go:572-578)
    ctrs = [[{"key": "NVIDIA", "type": "NVIDIA", "nums": nums, "memreq": memreq, "mempct": mempct, "coresreq": cores}]]
    if extra_ctrs:
        ctrs = extra_ctrs + ctrs
    return {"uid": uid, "ns": "default", "name": uid, "annotations": annos or {}, "containers": ctrs, "n_init": n_init}


def config1(n_nodes=4, n_pods=1000, departures=0.0, seed=SEED):
    """C1: 4 nodes x 8 x 80 GB, identical 3000 MiB / 30-core pods, binpack/binpack (saturates after 96 fits, F10)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    nodes = [{"name": "node-%04d" % i,
              "devices": [_nv_dev(i, d, "NVIDIA A100-SXM4-80GB", 81920, d // 4) for d in range(8)]}
             for i in range(n_nodes)]
    ops, live = [], []
    for p in range(n_pods):
        uid = "c1-%06d" % p
        ops.append(("filter", _nv_pod(uid, 1, memreq=3000, cores=30)))
        live.append(uid)
        if departures > 0 and rng.random() < departures and live:
            ops.append(("delete", live.pop(int(rng.integers(len(live))))))
    return {"node_policy": "binpack", "gpu_policy": "binpack"}, nodes, ops


def config2(n_nodes=1000, n_pods=50000, seed=SEED, occupancy=0.7):
    """C2: N nodes x 8 x 80 GB, mixed 1-8 GPU pods (gpumem and gpumem-percentage), spread/spread, departures."""
    rng = np.random.Generator(np.random.PCG64(seed + 2))
    nodes = [{"name": "node-%05d" % i,
              "devices": [_nv_dev(i, d, "NVIDIA A100-SXM4-80GB", 81920, d // 4) for d in range(8)]}
             for i in range(n_nodes)]
    # mean request ~ 2.35 GPUs at ~35 % of a card: keep about `occupancy` of the slots' memory busy
    target_live = max(4, int(occupancy * n_nodes * 8 / (2.35 * 0.45)))
    ops, live = [], []
    for p in range(n_pods):
        uid = "c2-%06d" % p
        nums = int(rng.choice([1, 2, 4, 8], p=[0.55, 0.20, 0.15, 0.10]))
        cores = int(rng.choice([0, 10, 25, 50, 100]))
        if rng.random() < 0.5:
            pod = _nv_pod(uid, nums, memreq=int(rng.integers(1000, 40001)), cores=cores)
        else:
            pod = _nv_pod(uid, nums, mempct=int(rng.integers(5, 101)), cores=cores)
        ops.append(("filter", pod))
        live.append(uid)
        while len(live) > target_live:
            ops.append(("delete", live.pop(int(rng.integers(len(live))))))
    return {"node_policy": "spread", "gpu_policy": "spread"}, nodes, ops


def _pcie_pairs(rng, n=8):
    # calculateGPUPairScore's PCIe levels are 10..60 (calculate_score.go:230-283); symmetric, mixed per pair
    m = [[0] * n for _ in range(n)]
    for i in range(n):
        for j in range(i + 1, n):
            v = int(rng.choice([10, 20, 30, 40, 50, 60]))
            m[i][j] = m[j][i] = v
    return m


def config3(n_nodes=5000, n_pods=20000, seed=SEED, occupancy=0.7):
    """C3: N nodes x 8 B200-class 180 GB devices, NUMA 4+4, NVSwitch pair score 1800 (20 % PCIe-only nodes
    with mixed 10-60 scores), topology-aware 2/4/8-GPU pods, 25 % numa-bind."""
    rng = np.random.Generator(np.random.PCG64(seed + 3))
    nodes = []
    for i in range(n_nodes):
        devs = [_nv_dev(i, d, "NVIDIA B200", 184320, d // 4) for d in range(8)]
        if rng.random() < 0.2:
            pairs = _pcie_pairs(rng)
        else:
            pairs = [[0 if a == b else 1800 for b in range(8)] for a in range(8)]
        nodes.append({"name": "node-%05d" % i, "devices": devs, "pair_scores": pairs})
    target_live = max(4, int(occupancy * n_nodes * 8 / (4.2 * 0.5)))
    ops, live = [], []
    for p in range(n_pods):
        uid = "c3-%06d" % p
        nums = int(rng.choice([2, 4, 8], p=[0.5, 0.3, 0.2]))
        annos = {"hami.io/gpu-scheduler-policy": "topology-aware"}
        if rng.random() < 0.25:
            annos["nvidia.com/numa-bind"] = "true"
        cores = int(rng.choice([0, 25, 50, 100]))
        if rng.random() < 0.5:
            pod = _nv_pod(uid, nums, memreq=int(rng.integers(8000, 120001)), cores=cores, annos=annos)
        else:
            pod = _nv_pod(uid, nums, mempct=int(rng.integers(10, 101)), cores=cores, annos=annos)
        ops.append(("filter", pod))
        live.append(uid)
        while len(live) > target_live:
            ops.append(("delete", live.pop(int(rng.integers(len(live))))))
    return {"node_policy": "binpack", "gpu_policy": "spread"}, nodes, ops


def config4(n_nodes=20000, n_pods=200000, seed=SEED, occupancy=0.6):
    """C4: heterogeneous NVIDIA (4 models) / Ascend910B3 / MLU / DCU nodes, type and UUID use/nouse selectors,
    2 % unhealthy devices."""
    rng = np.random.Generator(np.random.PCG64(seed + 4))
    nodes, nv_uuids = [], []
    kinds = rng.choice(4, size=n_nodes, p=[0.5, 0.2, 0.15, 0.15])
    for i in range(n_nodes):
        k = int(kinds[i])
        devs = []
        for d in range(8):
            health = bool(rng.random() >= 0.02)
            if k == 0:
                typ, mem = NVIDIA_MODELS[int(rng.integers(4))] if d == 0 else (devs[0]["type"], devs[0]["devmem"])
                devs.append(_nv_dev(i, d, typ, mem, d // 4, health))
                nv_uuids.append(devs[-1]["id"])
            elif k == 1:
                devs.append({"id": "NPU-%06d-%d" % (i, d), "index": d, "count": 20, "devmem": 65536, "devcore": 20,
                             "type": "Ascend910B3", "numa": d // 4, "mode": "", "health": health,
                             "vendor": "Ascend910B3", "network_id": d // 4})
            elif k == 2:
                devs.append({"id": "node-%06d-cambricon-mlu-%d" % (i, d), "index": d, "count": 100, "devmem": 49152,
                             "devcore": 100, "type": "MLU", "numa": 0, "mode": "", "health": health, "vendor": "MLU"})
            else:
                devs.append({"id": "DCU-%06d-%d" % (i, d), "index": d, "count": 4, "devmem": 65536, "devcore": 100,
                             "type": "DCU-K100" if i % 2 else "DCU-Z100L", "numa": d // 4, "mode": "", "health": health,
                             "vendor": "DCU"})
        node = {"name": "node-%05d" % i, "devices": devs}
        if k == 1:
            node["hami_vnpu_core"] = 1 if i % 3 == 0 else -1
        nodes.append(node)
    target_live = max(4, int(occupancy * n_nodes * 8 / (1.8 * 0.5)))
    ops, live = [], []
    for p in range(n_pods):
        uid = "c4-%06d" % p
        v = int(rng.choice(4, p=[0.5, 0.2, 0.15, 0.15]))
        nums = int(rng.choice([1, 2, 4], p=[0.6, 0.25, 0.15]))
        annos = {}
        if v == 0:
            r = rng.random()
            if r < 0.30:
                annos["nvidia.com/use-gputype"] = str(rng.choice(["A100", "H100", "a100-sxm4,L40S", "A100-PCIE"]))
            elif r < 0.40:
                annos["nvidia.com/nouse-gputype"] = str(rng.choice(["A100-PCIE", "L40S", "h100"]))
            elif r < 0.45 and nv_uuids:
                k = int(rng.integers(1, 5))
                annos["nvidia.com/use-gpuuuid"] = ",".join(str(x) for x in rng.choice(nv_uuids, size=k))
            elif r < 0.50 and nv_uuids:
                k = int(rng.integers(1, 5))
                annos["nvidia.com/nouse-gpuuuid"] = ",".join(str(x) for x in rng.choice(nv_uuids, size=k))
            cores = int(rng.choice([0, 10, 25, 50, 100]))
            if rng.random() < 0.5:
                pod = _nv_pod(uid, nums, memreq=int(rng.integers(1000, 40001)), cores=cores, annos=annos)
            else:
                pod = _nv_pod(uid, nums, mempct=int(rng.integers(5, 101)), cores=cores, annos=annos)
        elif v == 1:
            mem = int(rng.choice([0, 16384, 32768]))
            if rng.random() < 0.3:
                annos["huawei.com/vnpu-mode"] = "hami-core"
            pod = {"uid": uid, "ns": "default", "name": uid, "annotations": annos, "n_init": 0,
                   "containers": [[{"key": "Ascend910B3", "type": "Ascend910B3", "nums": nums, "memreq": mem,
                                    "mempct": 100 if mem == 0 else 0, "coresreq": int(rng.choice([0, 5, 10]))}]]}
        elif v == 2:
            mem = int(rng.choice([0, 16, 32, 64])) * 256
            pod = {"uid": uid, "ns": "default", "name": uid, "annotations": annos, "n_init": 0,
                   "containers": [[{"key": "MLU", "type": "MLU", "nums": nums, "memreq": mem,
                                    "mempct": 100 if mem == 0 else 0, "coresreq": int(rng.choice([10, 25, 50, 100]))}]]}
        else:
            mem = int(rng.choice([0, 8192, 16384, 32768]))
            if rng.random() < 0.2:
                annos["hygon.com/use-dcutype"] = "K100"
            pod = {"uid": uid, "ns": "default", "name": uid, "annotations": annos, "n_init": 0,
                   "containers": [[{"key": "DCU", "type": "DCU", "nums": nums, "memreq": mem,
                                    "mempct": 100 if mem == 0 else 0, "coresreq": int(rng.choice([10, 25, 50, 100]))}]]}
        ops.append(("filter", pod))
        live.append(uid)
        while len(live) > target_live:
            ops.append(("delete", live.pop(int(rng.integers(len(live))))))
    return {"node_policy": "binpack", "gpu_policy": "spread"}, nodes, ops


def config5(n_nodes=100000, n_pods=100000, seed=SEED):
    """C5: 100 000-node inventory with C2's device and request mix (sharded across GPUs by bench.py)."""
    return config2(n_nodes=n_nodes, n_pods=n_pods, seed=seed + 5)


def mixed_features(n_nodes=24, n_pods=600, seed=SEED):
    """Not a BASELINE config: a small stream that exercises every request feature the path has
    (multi-container, init containers, mutex, chains, weights, numa-bind, cordon, type-equality scoring)."""
    rng = np.random.Generator(np.random.PCG64(seed + 9))
    nodes = []
    for i in range(n_nodes):
        typ = "NVIDIA" if i % 3 == 0 else "NVIDIA A100-SXM4-80GB"
        devs = [_nv_dev(i, d, typ, 40960 if i % 5 == 0 else 81920, (d // 2) % 2, health=bool(rng.random() > 0.05))
                for d in range(int(rng.integers(2, 9)))]
        node = {"name": "mx-%03d" % i, "devices": devs}
        if i % 4 == 0:
            node["pair_scores"] = _pcie_pairs(rng, len(devs))
        if i % 7 == 0:
            node["device_cordon"] = " %s , %s" % (devs[0]["id"], "GPU-unknown")
        nodes.append(node)
    policies = ["binpack", "spread", "mutex", "topology-aware", "binpack,numa", "numa", "spread,numa",
                "mutex,topology-aware", "numa,binpack", "binpack,spread"]
    ops, live = [], []
    for p in range(n_pods):
        uid = "mx-%05d" % p
        annos = {}
        if rng.random() < 0.7:
            annos["hami.io/gpu-scheduler-policy"] = str(rng.choice(policies))
        if rng.random() < 0.5:
            annos["hami.io/node-scheduler-policy"] = str(rng.choice(["binpack", "spread"]))
        if rng.random() < 0.25:
            annos["nvidia.com/numa-bind"] = str(rng.choice(["true", "1", "false", "yes"]))
        if rng.random() < 0.2:
            annos["hami.io/device-scoring-weights"] = str(rng.choice(["slot=1,core=1,memory=3", "slot=0,core=3,memory=1",
                                                                      "memory=2,slot=5,core=0"]))
        if rng.random() < 0.1:
            annos["nvidia.com/vgpu-mode"] = str(rng.choice(["hami-core", "mps", "hami-core,mig"]))
        n_ctr = int(rng.choice([1, 1, 1, 2, 3]))
        n_init = int(rng.choice([0, 0, 1])) if n_ctr > 1 else 0
        ctrs = []
        for c in range(n_ctr):
            if n_ctr > 1 and rng.random() < 0.2:
                ctrs.append([])
                continue
            nums = int(rng.choice([1, 1, 2, 3, 4]))
            cores = int(rng.choice([0, 10, 30, 50, 100, 120]))
            if rng.random() < 0.5:
                r = {"key": "NVIDIA", "type": "NVIDIA", "nums": nums, "memreq": int(rng.integers(500, 30000)),
                     "mempct": 101, "coresreq": cores}
            else:
                r = {"key": "NVIDIA", "type": "NVIDIA", "nums": nums, "memreq": 0,
                     "mempct": int(rng.integers(1, 101)), "coresreq": cores}
            ctrs.append([r])
        if not any(ctrs):
            ctrs[-1] = [{"key": "NVIDIA", "type": "NVIDIA", "nums": 1, "memreq": 1000, "mempct": 101, "coresreq": 10}]
        ops.append(("filter", {"uid": uid, "ns": "default", "name": uid, "annotations": annos, "containers": ctrs,
                               "n_init": n_init}))
        live.append(uid)
        while len(live) > 60:
            ops.append(("delete", live.pop(int(rng.integers(len(live))))))
        if rng.random() < 0.02 and live:
            # the same pod filtered again: TakeAndDeletePod first (scheduler.go:1092-1094)
            again = dict(ops[-1][1]) if ops[-1][0] == "filter" else None
            if again:
                ops.append(("filter", again))
    return {"node_policy": "binpack", "gpu_policy": "spread"}, nodes, ops


CONFIGS = {"c1": config1, "c2": config2, "c3": config3, "c4": config4, "c5": config5, "mixed": mixed_features}
