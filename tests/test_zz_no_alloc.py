"""hs_sweep_static — the entry point for callers that must not allocate device memory (a scheduler plugin
living in a fixed, pre-carved device-memory pool, or a latency-critical path that must stay out of the
driver's allocator).  Its scratch is a file-scope static device array, so every call shares it.

The test drives it through the C ABI and checks the winner it reports against the allocating entry point
(hs_sweep_probe) bit for bit — the same exactness the other GPU tests use — on two calls in a row (the
second call finds the first call's data in the shared array) and for a second request shape, and against
the CPU oracle's own choice for the same pod on the same cluster state.  The output buffer is allocated by
the TEST (a caller-owned device buffer); the entry point itself allocates nothing, which the test checks by
reading the driver's free-memory figure around the calls.
Collected last on purpose (file name), and it must also pass alone in a fresh process.
"""
import ctypes as C
import struct

import pytest

from oracle_abi import abi

pytestmark = pytest.mark.gpu


def _setup(with_oracle=False):
    import k8s_vgpu_scheduler_b200.engine as eng
    import k8s_vgpu_scheduler_b200.gen as gen
    cfg, nodes, ops = gen.config3(n_nodes=700, n_pods=1500)
    e = eng.Engine(cfg["node_policy"], cfg["gpu_policy"])
    for n in nodes:
        e.node_upsert(n)
    e.run_batch(eng.PackedOps(ops[:1200]))          # a half-full cluster
    if with_oracle:                                  # the same cluster state in the CPU oracle
        import os
        import oracle_abi as oa
        o = oa.Oracle(cfg["node_policy"], cfg["gpu_policy"], threads=len(os.sched_getaffinity(0)))
        for n in nodes:
            o.node_upsert(n)
        o.run_ops(eng.PackedOps(ops[:1200]))
        e.oracle = o
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


def _oracle_choice(o, pod):
    """The oracle's Filter for the same pod, undone straight away (the scan commits nothing)."""
    r = o.filter(dict(pod, uid="static-check"))
    if r["status"] == abi.HS_FIT:
        o.pod_delete("static-check")
        return abi.HS_FIT, r["node_index"], struct.unpack("I", struct.pack("f", r["score"]))[0]
    return r["status"], -1, 0


def test_static_sweep_matches_allocating_sweep_twice_in_a_row():
    import torch
    e, L, pods = _setup(with_oracle=True)
    out = torch.zeros(2, dtype=torch.int64, device="cuda")      # the CALLER's output buffer
    plain = next(p for p in pods if "nvidia.com/numa-bind" not in p["annotations"])
    numa = next(p for p in pods if "nvidia.com/numa-bind" in p["annotations"])
    free_after_first = None
    for pod in (plain, plain, numa, plain):                      # consecutive calls share the static array
        want = _probe(e, L, pod)
        rc, pos, bits, ctas = _static(e, L, pod, out)
        assert (rc, pos) == want[:2]
        if rc == abi.HS_FIT:
            assert bits == want[2]
        assert (rc, pos, bits) == _oracle_choice(e.oracle, pod)  # and the CPU oracle agrees, directly
        assert 1 <= ctas <= 1024
        # the headline property: after the first call (which resolves the array's address) the entry point leaves
        # the device's free memory exactly where it was
        torch.cuda.synchronize()
        if free_after_first is None:
            free_after_first = torch.cuda.mem_get_info()[0]
            for _ in range(3):
                _static(e, L, pod, out)
            torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_after_first
    e.oracle.close()
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
