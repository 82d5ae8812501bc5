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
    "hs_usage_apply", "hs_pod_delete", "hs_filter", "hs_run_batch", "hs_batch_get", "hs_rollback",
    "hs_snapshot_node", "hs_last_device_ms", "hs_launch_count", "hs_batch_summary", "hs_sweep_probe",
    "hs_shard_unique_id", "hs_shard_init", "hs_sweep_static", "hs_shard_mailbox", "hs_shard_connect", "hs_debug_timing", "hs_debug_host_seconds", "hs_debug_divcheck",
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
    L.hs_run_batch.argtypes = [C.c_void_p, P(abi.hs_op), C.c_uint32]
    L.hs_batch_get.argtypes = [C.c_void_p, C.c_uint32, P(abi.hs_filter_result)]
    L.hs_batch_summary.argtypes = [C.c_void_p, P(C.c_uint32), P(C.c_uint32), P(C.c_uint64)]
    L.hs_shard_unique_id.argtypes = [C.c_void_p]
    L.hs_shard_init.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64]
    L.hs_rollback.argtypes = [C.c_void_p, C.c_char_p]
    L.hs_snapshot_node.argtypes = [C.c_void_p, C.c_char_p, P(abi.hs_device_usage_row), C.c_uint32]
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

    def node_remove(self, name):
        return self.L.hs_node_remove(self.h, abi._b(name))

    def usage_apply(self, uid, node, uuid, slots, mem, cores):
        return self.L.hs_usage_apply(self.h, abi._b(uid), abi._b(node), abi._b(uuid), slots, mem, cores)

    def pod_delete(self, uid):
        return self.L.hs_pod_delete(self.h, abi._b(uid))

    def filter(self, pod, node_names=None):
        p, keep = abi.build_pod(pod)
        res = abi.hs_filter_result()
        if node_names is None:
            names, n = None, 0
        else:
            names = (C.c_char_p * max(1, len(node_names)))(*[abi._b(x) for x in node_names])
            n = len(node_names)
        rc = self.L.hs_filter(self.h, C.byref(p), names, n, C.byref(res))
        out = abi.read_result(res)
        out["rc"] = rc
        if rc < 0:
            out["error"] = self.error()
        return out

    def run_batch(self, packed):
        """packed: PackedOps. Returns number of ops run; results via batch_get(i)."""
        return self._check(self.L.hs_run_batch(self.h, packed.arr, packed.n))

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

    def snapshot(self, node, cap=16):
        rows = (abi.hs_device_usage_row * cap)()
        n = self.L.hs_snapshot_node(self.h, abi._b(node), rows, cap)
        if n < 0:
            raise RuntimeError("snapshot failed: %d %s" % (n, self.error()))
        return [(rows[i].used, rows[i].usedmem, rows[i].usedcores) for i in range(min(n, cap))]

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
