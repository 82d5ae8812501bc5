"""ctypes binding of libhami_sched.so — the CUDA engine behind include/hami_sched.h.

This is the Python twin of the cgo shim in INTEGRATION.md: it only marshals.
Loading fails loudly when the library has not been built (run
`python -c "import __graft_entry__ as g; g.build()"` or `make -C k8s-vgpu-scheduler_b200/csrc`);
creating an Engine fails loudly when no CUDA device is usable.  No CPU fallback.
"""
import ctypes as C
import os

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhami_sched.so")
_lib = None

EXPORTS = [
    "hs_create", "hs_destroy", "hs_last_error", "hs_node_upsert", "hs_node_remove", "hs_node_count",
    "hs_usage_apply", "hs_pod_delete", "hs_filter", "hs_run_batch", "hs_run_batch_async", "hs_batch_wait", "hs_batch_get", "hs_rollback",
    "hs_snapshot_node", "hs_pod_get", "hs_pod_set_mig", "hs_snapshot_node_mig", "hs_snapshot_node_sgpu", "hs_debug_device_rows", "hs_decode_mig_allocations", "hs_last_device_ms", "hs_launch_count", "hs_batch_summary", "hs_sweep_probe",
    "hs_shard_unique_id", "hs_shard_init", "hs_sweep_static", "hs_shard_mailbox", "hs_shard_connect",
    "hs_decode_node_devices", "hs_decode_pod_devices", "hs_collapse_pod_devices", "hs_node_register", "hs_pod_ingest", "hs_pod_shrink", "hs_filter_dry", "hs_filter_simulate", "hs_quota_set", "hs_quota_del", "hs_quota_get", "hs_quota_memory_factor", "hs_generate_requests", "hs_debug_timing", "hs_debug_window_stats", "hs_debug_decision_ns", "hs_debug_copy_bytes", "hs_debug_host_seconds", "hs_debug_divcheck", "hs_debug_metax_table",
]


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("libhami_sched.so is not built (%s): the CUDA engine is required, there is no CPU fallback"
                           % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    P = C.POINTER
    L.hs_create.restype = C.c_void_p
    L.hs_create.argtypes = [P(abi.hs_config)]
    L.hs_destroy.argtypes = [C.c_void_p]
    L.hs_last_error.restype = C.c_char_p
    L.hs_last_error.argtypes = [C.c_void_p]
    L.hs_node_upsert.argtypes = [C.c_void_p, P(abi.hs_node_info)]
    L.hs_node_remove.argtypes = [C.c_void_p, C.c_char_p]
    L.hs_node_count.argtypes = [C.c_void_p]
    L.hs_usage_apply.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int32, C.c_int32, C.c_int32]
    L.hs_pod_delete.argtypes = [C.c_void_p, C.c_char_p]
    L.hs_filter.argtypes = [C.c_void_p, P(abi.hs_pod), P(C.c_char_p), C.c_uint32, P(abi.hs_filter_result)]
    L.hs_filter_dry.argtypes = [C.c_void_p, P(abi.hs_pod), P(C.c_char_p), C.c_uint32, P(abi.hs_filter_result)]
    L.hs_filter_simulate.argtypes = [C.c_void_p, P(abi.hs_pod), P(abi.hs_sim_node), C.c_uint32, P(abi.hs_filter_result)]
    L.hs_run_batch.argtypes = [C.c_void_p, P(abi.hs_op), C.c_uint32]
    L.hs_run_batch_async.argtypes = [C.c_void_p, P(abi.hs_op), C.c_uint32]
    L.hs_batch_wait.argtypes = [C.c_void_p, C.c_uint32]
    L.hs_batch_get.argtypes = [C.c_void_p, C.c_uint32, P(abi.hs_filter_result)]
    L.hs_batch_summary.argtypes = [C.c_void_p, P(C.c_uint32), P(C.c_uint32), P(C.c_uint64)]
    L.hs_shard_unique_id.argtypes = [C.c_void_p]
    L.hs_shard_init.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64]
    L.hs_rollback.argtypes = [C.c_void_p, C.c_char_p]
    L.hs_snapshot_node.argtypes = [C.c_void_p, C.c_char_p, P(abi.hs_device_usage_row), C.c_uint32]
    L.hs_pod_get.argtypes = [C.c_void_p, C.c_char_p, P(C.c_char_p), P(abi.hs_pod_usage_row), C.c_uint32]
    L.hs_pod_set_mig.argtypes = [C.c_void_p, C.c_char_p, P(abi.hs_mig_reservation), C.c_uint32]
    L.hs_snapshot_node_mig.argtypes = [C.c_void_p, C.c_char_p, P(C.c_uint32), P(C.c_uint8), C.c_uint32]
    L.hs_decode_node_devices.argtypes = [P(abi.hs_vendor_registration), P(P(abi.hs_device_info)), P(C.c_uint32), P(P(C.c_int32))]
    L.hs_decode_pod_devices.argtypes = [P(abi.hs_kv), C.c_uint32, P(abi.hs_kv), C.c_uint32, P(P(abi.hs_pod_device)), P(C.c_uint32)]
    L.hs_collapse_pod_devices.argtypes = [C.c_int32, C.c_uint32, P(abi.hs_pod_device), C.c_uint32, P(P(abi.hs_pod_device)), P(C.c_uint32)]
    L.hs_node_register.argtypes = [C.c_void_p, C.c_char_p, P(abi.hs_vendor_registration), C.c_uint32, C.c_char_p, C.c_int8]
    L.hs_pod_ingest.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, P(abi.hs_kv), C.c_uint32, C.c_uint32]
    L.hs_quota_set.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32, C.c_int64]
    L.hs_quota_del.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32]
    L.hs_quota_get.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32, P(C.c_int64), P(C.c_int64), P(C.c_int32)]
    L.hs_quota_memory_factor.argtypes = [C.c_void_p, C.c_char_p, C.c_int32]
    L.hs_pod_shrink.argtypes = [C.c_void_p, C.c_char_p, P(abi.hs_kv), C.c_uint32, C.c_uint32]
    L.hs_last_device_ms.restype = C.c_double
    L.hs_last_device_ms.argtypes = [C.c_void_p]
    L.hs_launch_count.restype = C.c_uint64
    L.hs_launch_count.argtypes = [C.c_void_p]
    _lib = L
    return L


def unique_id():
    """128-byte NCCL unique id (rank 0 creates it, the others receive it)."""
    buf = C.create_string_buffer(128)
    rc = lib().hs_shard_unique_id(buf)
    if rc != 0:
        raise RuntimeError("hs_shard_unique_id failed: %s" % lib().hs_last_error(None).decode())
    return buf.raw


class IngestError(Exception):
    """a decoder refused its input; .rc is the status, str() the reference's error text"""

    def __init__(self, rc, msg):
        super().__init__(msg)
        self.rc = rc


def _kvs(d):
    items = list(d.items()) if isinstance(d, dict) else list(d)
    arr = (abi.hs_kv * max(len(items), 1))()
    for i, (k, v) in enumerate(items):
        arr[i].key, arr[i].value = abi._b(k), abi._b(v)
    return arr, len(items)


def _registration(vendor, fmt, annotation, pair_scores=None, node_name=None, cores=0, memory=0, aux=0):
    """annotation formats: (vendor, fmt, annotation[, pair_scores]); capacity formats: (vendor, fmt, None, None, node, cores, memory)."""
    return abi.hs_vendor_registration(abi._b(vendor), fmt, abi._b(annotation), abi._b(pair_scores), abi._b(node_name), cores, memory, aux)


def decode_node_devices(vendor, fmt, annotation, pair_scores=None, node_name=None, cores=0, memory=0, aux=0):
    """hs_decode_node_devices (host only, no GPU): (list of DeviceInfo dicts, pair-score matrix or None)."""
    L = lib()
    reg = _registration(vendor, fmt, annotation, pair_scores, node_name, cores, memory, aux)
    devs, n, pair = C.POINTER(abi.hs_device_info)(), C.c_uint32(), C.POINTER(C.c_int32)()
    rc = L.hs_decode_node_devices(C.byref(reg), C.byref(devs), C.byref(n), C.byref(pair))
    if rc != 0:
        raise IngestError(rc, L.hs_last_error(None).decode())
    out = [{"ID": devs[i].id.decode(), "Index": devs[i].index, "Count": devs[i].count, "Devmem": devs[i].devmem,
            "Devcore": devs[i].devcore, "Type": devs[i].type.decode(), "Numa": devs[i].numa, "Mode": devs[i].mode.decode(),
            "Health": bool(devs[i].health), "DeviceVendor": devs[i].vendor.decode(), "NetworkID": devs[i].network_id}
           for i in range(n.value)]
    for i in range(n.value):      # enflame DRS CustomInfo
        if devs[i].enflame:
            ef = devs[i].enflame.contents
            out[i]["Enflame"] = {"minor": ef.minor.decode(), "index": ef.index.decode(), "max_slice": ef.max_slice,
                                 "profiles": {ef.profiles[k].key.decode(): ef.profiles[k].value.decode() for k in range(ef.n_profiles)}}
    for i in range(n.value):      # metax sGPU CustomInfo (QosPolicy, LinkZone)
        if devs[i].sgpu:
            sg = devs[i].sgpu.contents
            out[i]["Sgpu"] = {"qos_policy": (sg.qos_policy or b"").decode(), "link_zone": sg.link_zone}
    for i in range(n.value):      # DeviceInfo.MIGProfiles, present for GPUs that report any
        if devs[i].n_mig_profiles:
            out[i]["MigProfiles"] = [
                {"name": p.name.decode(), "memory_mb": p.memory_mb, "core": p.core, "slice_count": p.slice_count,
                 "placements": [(p.placements[j].start, p.placements[j].size) for j in range(p.n_placements)]}
                for p in (devs[i].mig_profiles[k] for k in range(devs[i].n_mig_profiles))]
    return out, ([pair[i] for i in range(n.value * n.value)] if pair else None)


def decode_mig_allocations(raw):
    """hs_decode_mig_allocations (host only): [(gpu_uuid, profile, start, size)]; IngestError where the reference's
    DecodeMigAllocations returns an error."""
    L = lib()
    out, n = C.POINTER(abi.hs_mig_reservation)(), C.c_uint32()
    L.hs_decode_mig_allocations.argtypes = [C.c_char_p, C.POINTER(C.POINTER(abi.hs_mig_reservation)), C.POINTER(C.c_uint32)]
    rc = L.hs_decode_mig_allocations(abi._b(raw), C.byref(out), C.byref(n))
    if rc != 0:
        raise IngestError(rc, L.hs_last_error(None).decode())
    return [(out[i].gpu_uuid.decode(), out[i].profile.decode(), out[i].start, out[i].size) for i in range(n.value)]


def _pod_devices_out(devs, n):
    out = {}
    for i in range(n):
        d = devs[i]
        ctrs = out.setdefault(d.key.decode(), [])
        while len(ctrs) <= d.container:
            ctrs.append([])
        if d.uuid is not None:
            ctrs[d.container].append({"UUID": d.uuid.decode(), "Type": d.type.decode(), "Usedmem": d.usedmem,
                                      "Usedcores": d.usedcores, "Slots": d.slots})
    return out


def decode_pod_devices(checklist, annos):
    """hs_decode_pod_devices: {key: [[device dict, …] per container]}."""
    L = lib()
    ck, nck = _kvs(checklist)
    an, nan = _kvs(annos)
    devs, n = C.POINTER(abi.hs_pod_device)(), C.c_uint32()
    rc = L.hs_decode_pod_devices(ck, nck, an, nan, C.byref(devs), C.byref(n))
    if rc != 0:
        raise IngestError(rc, L.hs_last_error(None).decode())
    return _pod_devices_out(devs, n.value)


def collapse_pod_devices(mode, n_init, raw):
    """hs_collapse_pod_devices; raw in the shape decode_pod_devices returns."""
    L = lib()
    flat = []
    for key, ctrs in raw.items():
        for c, devs in enumerate(ctrs):
            if not devs:
                flat.append(abi.hs_pod_device(abi._b(key), c, None, None, 0, 0, 0))
            for d in devs or []:
                flat.append(abi.hs_pod_device(abi._b(key), c, abi._b(d["UUID"]), abi._b(d["Type"]), d["Usedmem"], d["Usedcores"], 0))
    arr = (abi.hs_pod_device * max(len(flat), 1))(*flat)
    devs, n = C.POINTER(abi.hs_pod_device)(), C.c_uint32()
    rc = L.hs_collapse_pod_devices(mode, n_init, arr, len(flat), C.byref(devs), C.byref(n))
    if rc != 0:
        raise IngestError(rc, L.hs_last_error(None).decode())
    return _pod_devices_out(devs, n.value)


class PackedOps:
    """A pod stream marshalled once into the C structs hs_run_batch takes."""

    def __init__(self, ops):
        """ops: list of ("filter", pod_dict) | ("delete", uid)"""
        self.keep = []
        self.n = len(ops)
        self.arr = (abi.hs_op * max(1, self.n))()
        self.pods = (abi.hs_pod * max(1, self.n))()
        for i, (kind, arg) in enumerate(ops):
            if kind == "filter":
                p, k = abi.build_pod(arg)
                self.arr[i].kind = abi.HS_OP_FILTER
            else:
                p, k = abi.build_pod({"uid": arg})
                self.arr[i].kind = abi.HS_OP_DELETE
            self.pods[i] = p
            self.keep.append(k)
            self.arr[i].pod = C.pointer(self.pods[i])


class Engine:
    def __init__(self, node_policy=None, gpu_policy=None, **kw):
        self.L = lib()
        cfg, self._keep = abi.build_config(node_policy, gpu_policy, **kw)
        self.h = self.L.hs_create(C.byref(cfg))
        if not self.h:
            raise RuntimeError(self.L.hs_last_error(None).decode())

    def close(self):
        if getattr(self, "h", None):
            self.L.hs_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def error(self):
        return self.L.hs_last_error(self.h).decode()

    def _check(self, rc):
        if rc < 0:
            raise RuntimeError("hami_sched error %d: %s" % (rc, self.error()))
        return rc

    def node_upsert(self, node):
        ni, keep = abi.build_node(node)
        return self._check(self.L.hs_node_upsert(self.h, C.byref(ni)))

    def node_count(self):
        return self.L.hs_node_count(self.h)

    def node_remove(self, name):
        return self.L.hs_node_remove(self.h, abi._b(name))

    def usage_apply(self, uid, node, uuid, slots, mem, cores):
        return self.L.hs_usage_apply(self.h, abi._b(uid), abi._b(node), abi._b(uuid), slots, mem, cores)

    def pod_delete(self, uid):
        return self.L.hs_pod_delete(self.h, abi._b(uid))

    def filter_simulate(self, pod, sim_nodes):
        """sim_nodes: [(name, regs, device_cordon, hami_vnpu_core)] with regs as in node_register."""
        p, keep = abi.build_pod(pod)
        arr = (abi.hs_sim_node * max(len(sim_nodes), 1))()
        hold = []
        for i, (name, regs, cordon, vnpu) in enumerate(sim_nodes):
            ra = (abi.hs_vendor_registration * max(len(regs), 1))(*[_registration(*r) for r in regs])
            hold.append(ra)
            arr[i].name, arr[i].regs, arr[i].n_regs = abi._b(name), ra, len(regs)
            arr[i].device_cordon, arr[i].hami_vnpu_core = abi._b(cordon), vnpu
        res = abi.hs_filter_result()
        rc = self.L.hs_filter_simulate(self.h, C.byref(p), arr, len(sim_nodes), C.byref(res))
        out = abi.read_result(res)
        out["rc"] = rc
        if rc < 0:
            out["error"] = self.error()
        return out

    def filter(self, pod, node_names=None, dry=False):
        p, keep = abi.build_pod(pod)
        res = abi.hs_filter_result()
        if node_names is None:
            names, n = None, 0
        else:
            names = (C.c_char_p * max(1, len(node_names)))(*[abi._b(x) for x in node_names])
            n = len(node_names)
        rc = (self.L.hs_filter_dry if dry else self.L.hs_filter)(self.h, C.byref(p), names, n, C.byref(res))
        out = abi.read_result(res)
        out["rc"] = rc
        if rc < 0:
            out["error"] = self.error()
        return out

    def run_batch(self, packed):
        """packed: PackedOps. Returns number of ops run; results via batch_get(i)."""
        return self._check(self.L.hs_run_batch(self.h, packed.arr, packed.n))

    def run_batch_async(self, packed):
        """Start the batch and return: results are read with batch_get / batch_summary while later chunks still run.
        `packed` must stay alive until the batch has finished (batch_wait(), a complete batch_summary, or any other call)."""
        self._async_keep = packed
        return self._check(self.L.hs_run_batch_async(self.h, packed.arr, packed.n))

    def batch_wait(self, i=0xFFFFFFFF):
        return self._check(self.L.hs_batch_wait(self.h, i))

    def batch_get(self, i):
        res = abi.hs_filter_result()
        rc = self.L.hs_batch_get(self.h, i, C.byref(res))
        out = abi.read_result(res)
        out["rc"] = rc
        return out

    def batch_get_raw(self, i, res):
        return self.L.hs_batch_get(self.h, i, C.byref(res))

    def batch_summary(self, digest=True):
        """(n_fit, n_nofit, digest) over every result of the last run_batch; digest=False materialises every
        result (node, rows, annotation map / FailedNodes) without hashing it."""
        nf, nn, dg = C.c_uint32(), C.c_uint32(), C.c_uint64()
        self._check(self.L.hs_batch_summary(self.h, C.byref(nf), C.byref(nn), C.byref(dg) if digest else None))
        return nf.value, nn.value, dg.value

    def shard_init(self, rank, world, unique_id, pos_base):
        """unique_id: 128 bytes from unique_id() on rank 0 (None when world == 1)."""
        buf = C.create_string_buffer(bytes(unique_id), 128) if unique_id is not None else None
        return self._check(self.L.hs_shard_init(self.h, rank, world, buf, pos_base))

    def shard_mailbox(self):
        buf = C.create_string_buffer(64)
        self.L.hs_shard_mailbox.argtypes = [C.c_void_p, C.c_void_p]
        self._check(self.L.hs_shard_mailbox(self.h, buf))
        return buf.raw

    def shard_connect(self, handles):
        """handles: list of every rank's 64-byte mailbox handle, in rank order."""
        blob = C.create_string_buffer(b"".join(handles), 64 * len(handles))
        self.L.hs_shard_connect.argtypes = [C.c_void_p, C.c_void_p, C.c_int32]
        return self._check(self.L.hs_shard_connect(self.h, blob, len(handles)))

    def node_register(self, node, regs, device_cordon=None, hami_vnpu_core=-1):
        """regs: [(vendor, HS_REG_*, annotation, pair_scores or None)]; returns the number of devices registered."""
        arr = (abi.hs_vendor_registration * max(len(regs), 1))(*[_registration(*r) for r in regs])
        return self._check(self.L.hs_node_register(self.h, abi._b(node), arr, len(regs), abi._b(device_cordon), hami_vnpu_core))

    def pod_ingest(self, uid, node, annos, n_init=0, ns="default"):
        an, n = _kvs(annos)
        return self.L.hs_pod_ingest(self.h, abi._b(uid), abi._b(ns), abi._b(node), an, n, n_init)

    def quota_set(self, ns, vendor, which, limit):
        return self._check(self.L.hs_quota_set(self.h, abi._b(ns), abi._b(vendor), which, limit))

    def quota_del(self, ns, vendor, which):
        return self._check(self.L.hs_quota_del(self.h, abi._b(ns), abi._b(vendor), which))

    def quota_get(self, ns, vendor, which):
        lim, used, st = C.c_int64(), C.c_int64(), C.c_int32()
        rc = self.L.hs_quota_get(self.h, abi._b(ns), abi._b(vendor), which, C.byref(lim), C.byref(used), C.byref(st))
        return None if rc != 0 else (lim.value, used.value, bool(st.value))

    def pod_shrink(self, uid, annos, n_init=0):
        an, n = _kvs(annos)
        return self.L.hs_pod_shrink(self.h, abi._b(uid), an, n, n_init)

    def snapshot(self, node, cap=16):
        rows = (abi.hs_device_usage_row * cap)()
        n = self.L.hs_snapshot_node(self.h, abi._b(node), rows, cap)
        if n < 0:
            raise RuntimeError("snapshot failed: %d %s" % (n, self.error()))
        return [(rows[i].used, rows[i].usedmem, rows[i].usedcores) for i in range(min(n, cap))]

    def pod_get(self, uid):
        """podManager.GetPod: (node name, [(uuid, slots, usedmem, usedcores)]) or None when the cache does not hold the pod."""
        rows = (abi.hs_pod_usage_row * 16)()
        node = C.c_char_p()
        n = self.L.hs_pod_get(self.h, abi._b(uid), C.byref(node), rows, 16)
        if n < 0:
            return None
        return node.value.decode(), [(rows[i].uuid.decode(), rows[i].slots, rows[i].usedmem, rows[i].usedcores) for i in range(n)]

    def pod_set_mig(self, uid, reservations):
        """The informer delivers the pod's decoded hami.io/vgpu-mig-allocations: [(gpu_uuid, profile, start, size)]."""
        keep = []
        arr, n = abi.build_mig_reservations(reservations, keep)
        return self.L.hs_pod_set_mig(self.h, abi._b(uid), arr, n)

    def snapshot_node_mig(self, node, cap=16):
        """[(occupied slice mask, healthy)] per device of the node, as getNodesUsage would hand them to Fit."""
        occ = (C.c_uint32 * cap)()
        hl = (C.c_uint8 * cap)()
        n = self.L.hs_snapshot_node_mig(self.h, abi._b(node), occ, hl, cap)
        return [(occ[i], hl[i]) for i in range(max(0, min(n, cap)))]

    def device_rows(self, node):
        """(device rows, host mirror rows) of one node: 8 x (used, usedmem, usedcores, dflags) each (test hook)."""
        dev, mir = (C.c_int32 * 32)(), (C.c_int32 * 32)()
        self.L.hs_debug_device_rows.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        self._check(self.L.hs_debug_device_rows(self.h, abi._b(node), dev, mir))
        return [tuple(dev[i * 4:i * 4 + 4]) for i in range(8)], [tuple(mir[i * 4:i * 4 + 4]) for i in range(8)]

    def snapshot_node_sgpu(self, node, cap=16):
        """[(online pods, offline pods, pending QoS or None)] per device of the node (metax sGPU bookkeeping)."""
        on, off, pend = (C.c_int32 * cap)(), (C.c_int32 * cap)(), (C.c_char_p * cap)()
        self.L.hs_snapshot_node_sgpu.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_char_p), C.c_uint32]
        n = self.L.hs_snapshot_node_sgpu(self.h, abi._b(node), on, off, pend, cap)
        return [(on[i], off[i], pend[i].decode() if pend[i] is not None else None) for i in range(max(0, min(n, cap)))]

    def decision_ns(self, n):
        """%globaltimer ns at which each Filter decision of the last run_batch became final (0 = not a Filter)."""
        out = (C.c_uint64 * max(n, 1))()
        self.L.hs_debug_decision_ns.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.c_uint32]
        k = self.L.hs_debug_decision_ns(self.h, out, n)
        return [out[i] for i in range(max(k, 0))]

    def copy_bytes(self):
        """cumulative (host->device, device->host) bytes of the batch path, counted at the copies"""
        out = (C.c_uint64 * 2)()
        self.L.hs_debug_copy_bytes.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
        self.L.hs_debug_copy_bytes(self.h, out)
        return out[0], out[1]

    def host_seconds(self):
        """cumulative host seconds: (request compile, device run + copies, absorb, materialise)"""
        out = (C.c_double * 4)()
        self.L.hs_debug_host_seconds.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        self.L.hs_debug_host_seconds(self.h, out)
        return [out[i] for i in range(4)]

    def last_device_ms(self):
        return self.L.hs_last_device_ms(self.h)

    def launch_count(self):
        return self.L.hs_launch_count(self.h)
