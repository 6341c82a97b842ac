"""ctypes loader for ``libdtf_kernels.so`` (the hand-written sm_100a kernels).

The library is built in-tree by :mod:`distributed_tensorflow_b200.utils.build`.  On a machine
with a GPU a missing or unloadable library is a hard error -- there is deliberately no eager /
PyTorch fallback for the GPU ops, so a broken build can never pass silently.
"""
from __future__ import annotations

import ctypes
from pathlib import Path

_LIB = None
_PATH = Path(__file__).resolve().parents[1] / "_lib" / "libdtf_kernels.so"


class KernelError(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    global _LIB
    if _LIB is None:
        from ..utils import build as _build
        # (re)build when the library is missing or was built from other sources than the tree holds
        # (content hash of csrc/kernels + include, stamped next to the .so and compiled into it)
        if not _PATH.exists() or _build.built_hash(_PATH) != _build.source_hash("kernels"):
            _build.build_kernels()
        L = ctypes.CDLL(str(_PATH))
        L.dtf_last_error.restype = ctypes.c_char_p
        L.dtf_launch_count.restype = ctypes.c_longlong
        L.dtf_reset_launch_count.restype = None
        _LIB = L
    return _LIB


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise KernelError("%s failed (%d): %s" % (what, rc, lib().dtf_last_error().decode()))


def launch_count() -> int:
    return int(lib().dtf_launch_count())


def reset_launch_count() -> None:
    lib().dtf_reset_launch_count()


def ptr(t) -> ctypes.c_void_p:
    """Device pointer of a torch tensor (or None)."""
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def stream_ptr(stream=None) -> ctypes.c_void_p:
    import torch
    s = torch.cuda.current_stream() if stream is None else stream
    return ctypes.c_void_p(s.cuda_stream)
