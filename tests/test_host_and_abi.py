"""CPU-side checks: the C-ABI library loads and exports every symbol include/hami_sched.h declares,
refuses to run without a GPU (no CPU fallback), the generator is deterministic, and the oracle's
two shapes (faithful rebuild / lean incremental, 1 / T threads) agree."""
import ctypes as C
import os
import re

import pytest

import parity
from oracle_abi import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _built():
    import k8s_vgpu_scheduler_b200.engine as eng
    if not os.path.exists(eng.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return eng


def test_abi_exports_every_declared_symbol():
    eng = _built()
    L = C.CDLL(eng.LIB_PATH)
    hdr = open(os.path.join(ROOT, "include", "hami_sched.h")).read()
    declared = sorted(set(re.findall(r"\b(hs_[a-z_]+)\s*\(", hdr)))
    assert declared, "no entry points found in the header"
    for name in declared:
        assert hasattr(L, name), "libhami_sched.so does not export %s" % name
    assert set(eng.EXPORTS) <= set(declared)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    eng = _built()
    with pytest.raises(RuntimeError, match="no usable CUDA device|CUDA"):
        eng.Engine()


def test_struct_sizes_match_header():
    # guards the ctypes mirror against drifting from include/hami_sched.h
    assert C.sizeof(abi.hs_container_request) == 32
    assert C.sizeof(abi.hs_alloc_entry) == 56      # + mig_profile / mig_start / mig_size
    assert C.sizeof(abi.hs_mig_profile) == 40 and C.sizeof(abi.hs_mig_placement) == 8
    assert C.sizeof(abi.hs_device_info) == 104 and C.sizeof(abi.hs_container) == 24 and C.sizeof(abi.hs_enflame_info) == 40
    assert C.sizeof(abi.hs_kv) == 16
    assert C.sizeof(abi.hs_device_usage_row) == 12


def test_generator_is_deterministic():
    import k8s_vgpu_scheduler_b200.gen as gen
    a = gen.config3(n_nodes=50, n_pods=80)
    b = gen.config3(n_nodes=50, n_pods=80)
    assert a == b
    cfg, nodes, ops = gen.config4(n_nodes=80, n_pods=100)
    assert len(nodes) == 80 and sum(1 for k, _ in ops if k == "filter") == 100


@pytest.mark.parametrize("name,kw", [("c1", dict(n_pods=200, departures=0.3)), ("c2", dict(n_nodes=40, n_pods=400)),
                                     ("c3", dict(n_nodes=40, n_pods=300)), ("c4", dict(n_nodes=60, n_pods=400)),
                                     ("mixed", dict())])
def test_oracle_faithful_lean_and_threads_agree(name, kw):
    import k8s_vgpu_scheduler_b200.gen as gen
    cfg, nodes, ops = gen.CONFIGS[name](**kw)
    lean, s1 = parity.run_oracle(cfg, nodes, ops)
    faithful, s2 = parity.run_oracle(cfg, nodes, ops, faithful=True)
    threaded, s3 = parity.run_oracle(cfg, nodes, ops, threads=4)
    assert not parity.compare(ops, lean, faithful, s1, s2)
    assert not parity.compare(ops, lean, threaded, s1, s3)
    assert any(r["status"] == abi.HS_FIT for r in lean)


def test_c1_saturates_after_96_fits():
    # F10: 3000 MiB / 30-core pods on 4 x 8 x (80 GB, 100 cores, 10 slots): 3 per GPU (cores bound) = 96 fits
    import k8s_vgpu_scheduler_b200.gen as gen
    cfg, nodes, ops = gen.config1()
    res, _ = parity.run_oracle(cfg, nodes, ops, snapshot=False)
    assert sum(1 for r in res if r["status"] == abi.HS_FIT) == 96
    nofit = [r for r in res if r["status"] == abi.HS_NOFIT]
    assert len(nofit) == 904
    assert set(nofit[0]["failed"].values()) == {"8/8 CardInsufficientCore"}


def test_header_enumerators_match_the_python_mirror():
    """Every `HS_NAME = value` enumerator of include/hami_sched.h that abi.py mirrors must carry the same value, and the
    vendor-facing families (fit kinds, register formats, request kinds) must be mirrored completely."""
    import re
    from k8s_vgpu_scheduler_b200 import abi as A
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "include", "hami_sched.h")).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    enums = {m.group(1): int(m.group(2)) for m in re.finditer(r"\b(HS_[A-Z0-9_]+)\s*=\s*(-?\d+)", text)}
    assert len(enums) > 30
    mirrored = 0
    for name, value in enums.items():
        if hasattr(A, name):
            assert getattr(A, name) == value, name
            mirrored += 1
        else:
            assert not name.startswith(("HS_FIT_", "HS_REG_", "HS_REQ_")), "abi.py lacks " + name
    assert mirrored > 25
