"""hs_sweep_static — the entry point for callers that must not allocate device memory (a scheduler plugin
living in a fixed, pre-carved device-memory pool, or a latency-critical path that must stay out of the
driver's allocator).  Its scratch is a file-scope static device array, so every call shares it.

The test drives it through the C ABI and checks the winner it reports against the allocating entry point
(hs_sweep_probe) bit for bit — the same exactness the other GPU tests use — on two calls in a row (the
second call finds the first call's data in the shared array) and for a second request shape.  The output
buffer is allocated by the TEST (a caller-owned device buffer); the entry point itself allocates nothing.
Collected last on purpose (file name), and it must also pass alone in a fresh process.
"""
import ctypes as C
import struct

import pytest

from oracle_abi import abi

pytestmark = pytest.mark.gpu


def _setup():
    import k8s_vgpu_scheduler_b200.engine as eng
    import k8s_vgpu_scheduler_b200.gen as gen
    cfg, nodes, ops = gen.config3(n_nodes=700, n_pods=1500)
    e = eng.Engine(cfg["node_policy"], cfg["gpu_policy"])
    for n in nodes:
        e.node_upsert(n)
    e.run_batch(eng.PackedOps(ops[:1200]))          # a half-full cluster
    L = eng.lib()
    L.hs_sweep_static.argtypes = [C.c_void_p, C.POINTER(abi.hs_pod), C.c_void_p, C.c_void_p]
    L.hs_sweep_probe.argtypes = [C.c_void_p, C.POINTER(abi.hs_pod), C.c_uint32, C.POINTER(C.c_double),
                                 C.POINTER(C.c_int32), C.POINTER(C.c_float)]
    pods = [arg for kind, arg in ops[1200:] if kind == "filter"]
    return e, L, pods


def _probe(e, L, pod):
    p, keep = abi.build_pod(dict(pod, uid="static-check"))
    ms, idx, sc = C.c_double(), C.c_int32(), C.c_float()
    rc = L.hs_sweep_probe(e.h, C.byref(p), 1, C.byref(ms), C.byref(idx), C.byref(sc))
    return rc, idx.value, struct.unpack("I", struct.pack("f", sc.value))[0]


def _static(e, L, pod, out):
    p, keep = abi.build_pod(dict(pod, uid="static-check"))
    rc = L.hs_sweep_static(e.h, C.byref(p), C.c_void_p(out.data_ptr()), None)
    assert rc == 0, e.error()
    key = int(out[0].item()) & 0xFFFFFFFFFFFFFFFF
    pos = ((key >> 8) & 0xFFFFFF) - 1
    u = key >> 32
    if pod.get("annotations", {}).get("hami.io/node-scheduler-policy", "binpack") == "spread":
        u = ~u & 0xFFFFFFFF
    bits = (u & 0x7FFFFFFF) if u & 0x80000000 else (~u & 0xFFFFFFFF)
    return (abi.HS_FIT if pos >= 0 else abi.HS_NOFIT), pos, (bits if pos >= 0 else 0), int(out[1].item())


def test_static_sweep_matches_allocating_sweep_twice_in_a_row():
    import torch
    e, L, pods = _setup()
    out = torch.zeros(2, dtype=torch.int64, device="cuda")      # the CALLER's output buffer
    plain = next(p for p in pods if "nvidia.com/numa-bind" not in p["annotations"])
    numa = next(p for p in pods if "nvidia.com/numa-bind" in p["annotations"])
    for pod in (plain, plain, numa, plain):                      # consecutive calls share the static array
        want = _probe(e, L, pod)
        rc, pos, bits, ctas = _static(e, L, pod, out)
        assert (rc, pos) == want[:2]
        if rc == abi.HS_FIT:
            assert bits == want[2]
        assert 1 <= ctas <= 1024
    e.close()


def test_static_sweep_rejects_what_it_cannot_hold():
    import torch
    e, L, pods = _setup()
    out = torch.zeros(2, dtype=torch.int64, device="cuda")
    two = dict(pods[0], containers=pods[0]["containers"] * 2)   # two request entries: not the static shape
    p, keep = abi.build_pod(two)
    assert L.hs_sweep_static(e.h, C.byref(p), C.c_void_p(out.data_ptr()), None) == abi.HS_ERR_UNSUPPORTED
    assert "hs_sweep_static" in e.error()
    e.close()
