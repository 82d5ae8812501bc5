"""ctypes loader for the CPU oracle (oracle/liboracle.so) — test infrastructure.

Builds the library on first use (make -C oracle).  The structs are the product's
own ABI structs (k8s-vgpu-scheduler_b200/abi.py), so the same Python dicts feed
the oracle and the CUDA engine.
"""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import k8s_vgpu_scheduler_b200.abi as abi  # noqa: E402


class oracle_dev(C.Structure):
    _fields_ = [("id", C.c_char_p), ("index", C.c_uint32), ("used", C.c_int32), ("count", C.c_int32),
                ("usedmem", C.c_int32), ("totalmem", C.c_int32), ("totalcore", C.c_int32), ("usedcores", C.c_int32),
                ("mode", C.c_char_p), ("numa", C.c_int32), ("type", C.c_char_p), ("health", C.c_uint8),
                ("network_id", C.c_int32), ("score", C.c_float),
                ("mig_profiles", C.POINTER(abi.hs_mig_profile)), ("n_mig_profiles", C.c_uint32),
                ("mig_in_use", C.POINTER(abi.hs_mig_placement)), ("n_mig_in_use", C.c_uint32),
                ("enflame", C.POINTER(abi.hs_enflame_info)), ("sgpu", C.POINTER(abi.hs_metax_sgpu_info)),
                ("online_pods", C.c_int32), ("offline_pods", C.c_int32)]


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    so = os.path.join(ROOT, "oracle", "liboracle.so")
    srcs = [os.path.join(ROOT, "oracle", f) for f in ("oracle.cc", "oracle_capi.cc", "b2_lean.cc", "oracle.h", "oracle_capi.h")]
    srcs.append(os.path.join(ROOT, "include", "hami_sched.h"))
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs if os.path.exists(s)):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-s"])
    L = C.CDLL(so)
    P = C.POINTER
    L.oracle_create.restype = C.c_void_p
    L.oracle_create.argtypes = [P(abi.hs_config)]
    L.oracle_destroy.argtypes = [C.c_void_p]
    L.oracle_set_mode.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.oracle_node_upsert.argtypes = [C.c_void_p, P(abi.hs_node_info)]
    L.oracle_node_remove.argtypes = [C.c_void_p, C.c_char_p]
    L.oracle_usage_apply.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int32, C.c_int32, C.c_int32]
    L.oracle_pod_delete.argtypes = [C.c_void_p, C.c_char_p]
    L.oracle_pod_get.argtypes = [C.c_void_p, C.c_char_p, P(C.c_char_p), P(abi.hs_pod_usage_row), C.c_uint32]
    L.oracle_filter.argtypes = [C.c_void_p, P(abi.hs_pod), P(C.c_char_p), C.c_uint32, P(abi.hs_filter_result)]
    L.oracle_snapshot_node.argtypes = [C.c_void_p, C.c_char_p, P(abi.hs_device_usage_row), C.c_uint32]
    L.oracle_run_ops.argtypes = [C.c_void_p, P(abi.hs_op), C.c_uint32, P(C.c_uint32), P(C.c_uint32), P(C.c_uint64),
                                 P(C.c_double)]
    L.oracle_last_error.restype = C.c_char_p
    L.oracle_last_error.argtypes = [C.c_void_p]
    L.oracle_score_node.argtypes = [C.c_void_p, C.c_char_p, P(oracle_dev), C.c_uint32, C.c_char_p, C.c_int32,
                                    P(abi.hs_pod), P(abi.hs_node_info), P(abi.hs_filter_result)]
    L.oracle_fit_in_devices.argtypes = [C.c_void_p, P(oracle_dev), C.c_uint32, C.c_char_p, C.c_int32,
                                        P(abi.hs_container), P(abi.hs_pod), P(abi.hs_node_info),
                                        P(abi.hs_filter_result)]
    L.oracle_fit.argtypes = [C.c_void_p, C.c_char_p, P(oracle_dev), C.c_uint32, P(abi.hs_container_request),
                             P(abi.hs_pod), P(abi.hs_node_info), P(abi.hs_filter_result)]
    L.oracle_compute_score.restype = C.c_float
    L.oracle_compute_score.argtypes = [P(oracle_dev), P(abi.hs_container_request), C.c_uint32, C.c_int64, C.c_int64,
                                       C.c_int64]
    L.oracle_less.argtypes = [C.c_char_p, C.c_int32, P(oracle_dev), P(oracle_dev)]
    L.oracle_sort.argtypes = [C.c_char_p, C.c_int32, P(oracle_dev), C.c_uint32, P(C.c_int32)]
    L.oracle_node_default_score.restype = C.c_float
    L.oracle_node_default_score.argtypes = [P(oracle_dev), C.c_uint32]
    L.oracle_check_uuid.argtypes = [P(abi.hs_kv), C.c_uint32, C.c_char_p, C.c_char_p, C.c_char_p]
    L.oracle_check_type.argtypes = [P(abi.hs_kv), C.c_uint32, C.c_char_p, C.c_char_p, C.c_char_p]
    L.oracle_parse_weights.argtypes = [C.c_char_p, P(C.c_int64)]
    L.oracle_gen_reason.argtypes = [P(C.c_int32), C.c_int32, C.c_char_p, C.c_size_t]
    L.oracle_collapse.argtypes = [P(abi.hs_alloc_entry), C.c_uint32, C.c_uint32, P(abi.hs_alloc_entry), C.c_uint32]
    L.oracle_pod_set_mig.argtypes = [C.c_void_p, C.c_char_p, P(abi.hs_mig_reservation), C.c_uint32]
    L.oracle_snapshot_node_mig.argtypes = [C.c_void_p, C.c_char_p, P(C.c_uint32), P(C.c_uint8), C.c_uint32]
    L.oracle_mig_select_placement.argtypes = [P(abi.hs_mig_profile), C.c_uint32, P(abi.hs_mig_placement), C.c_uint32,
                                              C.c_char_p, P(abi.hs_mig_placement)]
    L.oracle_mig_can_place.argtypes = [P(abi.hs_mig_profile), C.c_uint32, P(abi.hs_mig_placement), C.c_uint32,
                                       P(C.c_char_p), C.c_uint32]
    L.oracle_mig_select_candidate.argtypes = [P(abi.hs_mig_profile), C.c_uint32, P(abi.hs_mig_placement), C.c_uint32,
                                              C.c_int32, P(C.c_int32), P(abi.hs_mig_placement)]
    _lib = L
    return L


_mig_keep = []   # ctypes arrays the oracle_dev rows point into (fixtures are tiny; kept for the life of the process)


def mig_profiles_from_go(profiles):
    """Go []device.MigProfile literal -> the dicts abi.build_mig_profiles takes"""
    return [{"name": p.get("Name", ""), "memory_mb": int(p.get("MemoryMB", 0)), "core": int(p.get("Core", 0)),
             "slice_count": int(p.get("SliceCount", 0)),
             "placements": [(int(q.get("Start", 0)), int(q.get("Size", 0))) for q in (p.get("Placements") or [])]}
            for p in (profiles or [])]


def placements_array(pl):
    arr = (abi.hs_mig_placement * max(1, len(pl)))()
    for i, (st, sz) in enumerate(pl):
        arr[i] = abi.hs_mig_placement(st, sz)
    return arr


def sgpu_from_go(ci):
    """Go CustomInfo of a metax sGPU device ({"QosPolicy", "LinkZone"}) -> hs_metax_sgpu_info*"""
    if not isinstance(ci, dict) or not ("QosPolicy" in ci or "LinkZone" in ci):
        return None
    return abi.build_sgpu_info({"qos_policy": ci.get("QosPolicy", ""), "link_zone": int(ci.get("LinkZone", 0))}, _mig_keep)


def pod_classes_from_go(d):
    """DeviceUsage.PodInfos -> (pods with metax-tech.com/sgpu-app-class online, … offline)"""
    on = off = 0
    for pi in d.get("PodInfos") or []:
        pod = pi.get("Pod") if isinstance(pi, dict) else None
        annos = (((pod or {}).get("ObjectMeta") or {}).get("Annotations")) or {}
        cls = annos.get("metax-tech.com/sgpu-app-class")
        on += cls == "online"
        off += cls == "offline"
    return on, off


def enflame_from_go(ci):
    """Go CustomInfo of an enflame DRS device ({"minor","index","profiles": {name: id},"maxSlice"}) -> hs_enflame_info*"""
    if not isinstance(ci, dict) or "profiles" not in ci:
        return None
    return abi.build_enflame_info({"minor": ci.get("minor", ""), "index": ci.get("index", ""),
                                   "max_slice": int(ci.get("maxSlice", 0)), "profiles": ci.get("profiles") or {}}, _mig_keep)


def dev_from_go(d, score=0.0):
    """Go DeviceUsage literal (fixture dict with Go field names) -> oracle_dev"""
    mp, nmp = abi.build_mig_profiles(mig_profiles_from_go(d.get("MigProfiles")), _mig_keep)
    inuse = [(int(a["Placement"].get("Start", 0)), int(a["Placement"].get("Size", 0))) for a in (d.get("MigAllocationsInUse") or [])]
    iu = placements_array(inuse) if inuse else None
    _mig_keep.append(iu)
    net = -1
    ci = d.get("CustomInfo")
    if isinstance(ci, dict) and "NetworkID" in ci:
        net = int(ci["NetworkID"])
    if isinstance(ci, dict) and ci.get("DeviceStrategy") == "die" and "AIC" in ci:   # vastai die: the card id (digit strings in the fixtures) rides in NetworkID
        net = int(ci["AIC"])
    if isinstance(ci, dict) and "AWSNodeType" in ci:      # carried as 1 (contains "inf"/"Inf": ring) / 0 (awsneuron/device.go:414)
        net = 1 if ("inf" in ci["AWSNodeType"] or "Inf" in ci["AWSNodeType"]) else 0
    return oracle_dev(abi._b(d.get("ID", "")), int(d.get("Index", 0)), int(d.get("Used", 0)), int(d.get("Count", 0)),
                      int(d.get("Usedmem", 0)), int(d.get("Totalmem", 0)), int(d.get("Totalcore", 0)),
                      int(d.get("Usedcores", 0)), abi._b(d.get("Mode", "")), int(d.get("Numa", 0)),
                      abi._b(d.get("Type", "")), 1 if d.get("Health", False) else 0, net, float(score),
                      mp, nmp, iu, len(inuse), enflame_from_go(ci), sgpu_from_go(ci), *pod_classes_from_go(d))


def devs_array(devs):
    arr = (oracle_dev * max(1, len(devs)))()
    for i, d in enumerate(devs):
        arr[i] = d
    return arr


def kunlun_calc_score(prev, cur):
    """calcscore (kunlun/topo.go:74-95) on the card indices in use before / picked now"""
    L = lib()
    L.oracle_kunlun_calc_score.restype = C.c_float
    L.oracle_kunlun_calc_score.argtypes = [C.POINTER(C.c_int32), C.c_uint32, C.POINTER(C.c_int32), C.c_uint32]
    a = (C.c_int32 * max(1, len(prev)))(*prev)
    b = (C.c_int32 * max(1, len(cur)))(*cur)
    return L.oracle_kunlun_calc_score(a, len(prev), b, len(cur))


def req_from_go(key, r):
    return {"key": key, "type": r.get("Type", ""), "nums": int(r.get("Nums", 0)), "memreq": int(r.get("Memreq", 0)),
            "mempct": int(r.get("MemPercentagereq", 0)), "coresreq": int(r.get("Coresreq", 0))}


class Oracle:
    """Thin OO wrapper: the scheduler-level oracle (Filter over an inventory)."""

    def __init__(self, node_policy=None, gpu_policy=None, faithful=False, threads=1, **kw):
        self.L = lib()
        cfg, self._keep = abi.build_config(node_policy, gpu_policy, **kw)
        self.h = self.L.oracle_create(C.byref(cfg))
        self.L.oracle_set_mode(self.h, 1 if faithful else 0, threads)

    def close(self):
        if self.h:
            self.L.oracle_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def node_upsert(self, node):
        ni, keep = abi.build_node(node)
        rc = self.L.oracle_node_upsert(self.h, C.byref(ni))
        if rc != 0:
            raise RuntimeError(self.L.oracle_last_error(self.h).decode())

    def usage_apply(self, uid, node, uuid, slots, mem, cores):
        return self.L.oracle_usage_apply(self.h, abi._b(uid), abi._b(node), abi._b(uuid), slots, mem, cores)

    def pod_delete(self, uid):
        return self.L.oracle_pod_delete(self.h, abi._b(uid))

    def pod_set_mig(self, uid, reservations):
        keep = []
        arr, n = abi.build_mig_reservations(reservations, keep)
        return self.L.oracle_pod_set_mig(self.h, abi._b(uid), arr, n)

    def snapshot_node_mig(self, node, cap=16):
        occ = (C.c_uint32 * cap)()
        hl = (C.c_uint8 * cap)()
        n = self.L.oracle_snapshot_node_mig(self.h, abi._b(node), occ, hl, cap)
        return [(occ[i], hl[i]) for i in range(max(0, n))]

    def snapshot_node_sgpu(self, node, cap=16):
        on, off, pend = (C.c_int32 * cap)(), (C.c_int32 * cap)(), (C.c_char_p * cap)()
        self.L.oracle_snapshot_node_sgpu.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_char_p), C.c_uint32]
        n = self.L.oracle_snapshot_node_sgpu(self.h, abi._b(node), on, off, pend, cap)
        return [(on[i], off[i], pend[i].decode() if pend[i] is not None else None) for i in range(max(0, n))]

    def pod_get(self, uid):
        rows = (abi.hs_pod_usage_row * 16)()
        node = C.c_char_p()
        n = self.L.oracle_pod_get(self.h, abi._b(uid), C.byref(node), rows, 16)
        if n < 0:
            return None
        return node.value.decode(), [(rows[i].uuid.decode(), rows[i].slots, rows[i].usedmem, rows[i].usedcores) for i in range(n)]

    def filter(self, pod, node_names=None):
        p, keep = abi.build_pod(pod)
        res = abi.hs_filter_result()
        if node_names is None:
            names, n = None, 0
        else:
            names = (C.c_char_p * max(1, len(node_names)))(*[abi._b(x) for x in node_names])
            n = len(node_names)
        rc = self.L.oracle_filter(self.h, C.byref(p), names, n, C.byref(res))
        out = abi.read_result(res)
        out["rc"] = rc
        if rc < 0:
            out["error"] = self.L.oracle_last_error(self.h).decode()
        return out

    def run_ops(self, packed):
        """packed: engine.PackedOps.  Returns (n_fit, n_nofit, digest, seconds)."""
        nf, nn, dg, sec = C.c_uint32(), C.c_uint32(), C.c_uint64(), C.c_double()
        rc = self.L.oracle_run_ops(self.h, packed.arr, packed.n, C.byref(nf), C.byref(nn), C.byref(dg), C.byref(sec))
        if rc != 0:
            raise RuntimeError(self.L.oracle_last_error(self.h).decode())
        return nf.value, nn.value, dg.value, sec.value

    def snapshot(self, node, cap=16):
        rows = (abi.hs_device_usage_row * cap)()
        n = self.L.oracle_snapshot_node(self.h, abi._b(node), rows, cap)
        return [(rows[i].used, rows[i].usedmem, rows[i].usedcores) for i in range(min(n, cap))]


# ---- B2: the lean CPU baseline (oracle/b2_lean.cc) -----------------------------------------------------------
class b2_op(C.Structure):
    _fields_ = [("kind", C.c_int32), ("nums", C.c_int32), ("memreq", C.c_int32), ("mempct", C.c_int32), ("coresreq", C.c_int32),
                ("node_spread", C.c_uint8), ("mutex_gate", C.c_uint8), ("topology", C.c_uint8), ("numa_bind", C.c_uint8),
                ("sort_mode", C.c_uint8), ("chain", C.c_uint8 * 3), ("w_slot", C.c_float), ("w_core", C.c_float), ("w_mem", C.c_float),
                ("rel_op", C.c_int32), ("rel_node", C.c_int32), ("rel_rows", (C.c_int32 * 3) * 8)]


class b2_decision(C.Structure):
    _fields_ = [("node", C.c_int32), ("score", C.c_float), ("n_picks", C.c_int32), ("picks", C.c_int8 * 8),
                ("memreq", C.c_int32 * 8), ("cores", C.c_int32)]


def _policy_contains(policy, name):
    return any(p.strip() == name for p in policy.split(","))


def b2_compile(pod, cfg):
    """The part of Filter that looks at the pod only (util.go:277-298, scheduler.go:641-652, gpu_policy.go:77-167), for the
    request shape B2 handles; None when the pod is outside it."""
    an = pod.get("annotations") or {}
    ctrs = pod["containers"]
    if len(ctrs) != 1 or len(ctrs[0]) != 1 or pod.get("n_init", 0) or ctrs[0][0]["type"] != "NVIDIA":
        return None
    if any(k in an for k in ("hami.io/device-scoring-weights", "nvidia.com/use-gputype", "nvidia.com/nouse-gputype",
                             "nvidia.com/use-gpuuuid", "nvidia.com/nouse-gpuuuid", "nvidia.com/vgpu-mode")):
        return None
    r = ctrs[0][0]
    gpu = an.get("hami.io/gpu-scheduler-policy", cfg.get("gpu_policy") or "spread")
    node = an.get("hami.io/node-scheduler-policy", cfg.get("node_policy") or "binpack")
    numa = an.get("nvidia.com/numa-bind") in ("1", "t", "T", "TRUE", "true", "True")
    op = b2_op()
    op.kind, op.nums, op.memreq, op.mempct, op.coresreq = 0, r["nums"], r["memreq"], r["mempct"], r["coresreq"]
    op.node_spread = 1 if node == "spread" else 0
    op.mutex_gate = 1 if _policy_contains(gpu, "mutex") else 0
    op.topology = 1 if _policy_contains(gpu, "topology-aware") else 0
    op.numa_bind = 1 if numa else 0
    op.w_slot = op.w_core = op.w_mem = 1.0
    if "," in gpu or gpu == "numa":
        op.sort_mode = 5
        chain, seen = [], set()
        for t in gpu.split(","):
            k = {"binpack": 1, "spread": 2, "numa": 3}.get(t.strip(), 0)
            if k and k not in seen:
                chain.append(k)
                seen.add(k)
        if not chain:
            chain = [2]
        if numa and chain[0] != 3:
            chain = [3] + [k for k in chain if k != 3]
        for i in range(3):
            op.chain[i] = chain[i] if i < len(chain) else 0
    elif gpu == "mutex":
        op.sort_mode = 2
    elif numa:
        op.sort_mode = 3 if gpu == "binpack" else 4
    else:
        op.sort_mode = 0 if gpu == "binpack" else 1
    return op


class B2:
    """Thin wrapper over oracle/b2_lean.cc: SoA inventory, persistent thread pool."""

    def __init__(self, nodes, threads):
        self.L = lib()
        self.L.b2_create.restype = C.c_void_p
        self.L.b2_create.argtypes = [C.c_int32, C.c_int32]
        self.L.b2_destroy.argtypes = [C.c_void_p]
        self.L.b2_node_set.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int16)]
        self.L.b2_usage_add.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32]
        self.L.b2_run.restype = C.c_double
        self.L.b2_run.argtypes = [C.c_void_p, C.POINTER(b2_op), C.c_int32, C.POINTER(b2_decision)]
        self.h = self.L.b2_create(len(nodes), threads)
        self.node_index = {n["name"]: i for i, n in enumerate(nodes)}
        self.dev_slot = {}
        for i, n in enumerate(nodes):
            devs = n["devices"]
            arr = (C.c_int32 * (6 * len(devs)))()
            for d, dv in enumerate(devs):
                arr[d * 6:d * 6 + 6] = [dv["count"], dv["devmem"], dv["devcore"], dv.get("numa", 0), 1 if dv.get("health", True) else 0,
                                        1 if dv["type"] == "NVIDIA" else 0]
                self.dev_slot[dv["id"]] = (i, d)
            pair = None
            if n.get("pair_scores"):
                pair = (C.c_int16 * 64)()
                for a, row in enumerate(n["pair_scores"]):
                    for b, v in enumerate(row):
                        pair[a * 8 + b] = v
            self.L.b2_node_set(self.h, i, len(devs), arr, pair)

    def usage_add(self, uuid, slots, mem, cores):
        i, d = self.dev_slot[uuid]
        self.L.b2_usage_add(self.h, i, d, max(slots, 1), mem, cores)

    def run(self, ops):
        arr = (b2_op * max(len(ops), 1))(*ops)
        out = (b2_decision * max(len(ops), 1))()
        sec = self.L.b2_run(self.h, arr, len(ops), out)
        return sec, out

    def close(self):
        if self.h:
            self.L.b2_destroy(self.h)
            self.h = None
