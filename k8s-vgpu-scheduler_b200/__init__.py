"""B200-native HAMi Filter/Score engine — Python side of the C ABI.

Holds only what the hot path needs: csrc/ (CUDA kernels + the C ABI +
the C++ host mirror of pkg/scheduler for this path) and thin ctypes bindings.
The engine itself is libhami_sched.so; importing `engine` fails loudly when it
has not been built — there is no CPU fallback.
"""
from . import abi  # noqa: F401
