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
                ("network_id", C.c_int32), ("score", C.c_float)]


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    so = os.path.join(ROOT, "oracle", "liboracle.so")
    srcs = [os.path.join(ROOT, "oracle", f) for f in ("oracle.cc", "oracle_capi.cc", "oracle.h", "oracle_capi.h")]
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
    _lib = L
    return L


def dev_from_go(d, score=0.0):
    """Go DeviceUsage literal (fixture dict with Go field names) -> oracle_dev"""
    net = -1
    ci = d.get("CustomInfo")
    if isinstance(ci, dict) and "NetworkID" in ci:
        net = int(ci["NetworkID"])
    if isinstance(ci, dict) and "AWSNodeType" in ci:      # carried as 1 (contains "inf"/"Inf": ring) / 0 (awsneuron/device.go:414)
        net = 1 if ("inf" in ci["AWSNodeType"] or "Inf" in ci["AWSNodeType"]) else 0
    return oracle_dev(abi._b(d.get("ID", "")), int(d.get("Index", 0)), int(d.get("Used", 0)), int(d.get("Count", 0)),
                      int(d.get("Usedmem", 0)), int(d.get("Totalmem", 0)), int(d.get("Totalcore", 0)),
                      int(d.get("Usedcores", 0)), abi._b(d.get("Mode", "")), int(d.get("Numa", 0)),
                      abi._b(d.get("Type", "")), 1 if d.get("Health", False) else 0, net, float(score))


def devs_array(devs):
    arr = (oracle_dev * max(1, len(devs)))()
    for i, d in enumerate(devs):
        arr[i] = d
    return arr


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
