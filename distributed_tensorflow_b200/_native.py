"""ctypes bindings for ``libdtf_native.so`` (C ABI in ``include/dtf_native.h``).

The library is the C++ CPU runtime: crc32c, LevelDB-format tables, TF checkpoint-V2 tensor bundles,
TFRecord framing, optimizer apply kernels and the TCP parameter-server transport.  It is built
in-tree on first use (``g++``, a few seconds).
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

_PATH = Path(__file__).resolve().parent / "_lib" / "libdtf_native.so"
_LIB = None


class NativeError(RuntimeError):
    pass


def _sig(L, name, restype, *argtypes):
    fn = getattr(L, name)
    fn.restype = restype
    fn.argtypes = list(argtypes)


def lib() -> C.CDLL:
    global _LIB
    if _LIB is not None:
        return _LIB
    from .utils.build import build_native
    build_native()  # no-op when up to date
    L = C.CDLL(str(_PATH))
    vp, cp, sz, i, ll, f = C.c_void_p, C.c_char_p, C.c_size_t, C.c_int, C.c_longlong, C.c_float
    u32 = C.c_uint32
    _sig(L, "dtfn_last_error", cp)
    _sig(L, "dtfn_crc32c", u32, vp, sz)
    _sig(L, "dtfn_crc32c_extend", u32, u32, vp, sz)
    _sig(L, "dtfn_crc32c_mask", u32, u32)
    _sig(L, "dtfn_crc32c_unmask", u32, u32)
    _sig(L, "dtfn_table_builder_new", vp, cp, i)
    _sig(L, "dtfn_table_builder_add", i, vp, vp, sz, vp, sz)
    _sig(L, "dtfn_table_builder_finish", i, vp)
    _sig(L, "dtfn_table_builder_abandon", None, vp)
    _sig(L, "dtfn_table_open", vp, cp)
    _sig(L, "dtfn_table_count", ll, vp)
    _sig(L, "dtfn_table_entry", i, vp, ll, C.POINTER(vp), C.POINTER(sz), C.POINTER(vp), C.POINTER(sz))
    _sig(L, "dtfn_table_find", i, vp, vp, sz, C.POINTER(vp), C.POINTER(sz))
    _sig(L, "dtfn_table_close", None, vp)
    _sig(L, "dtfn_bundle_writer_new", vp, cp)
    _sig(L, "dtfn_bundle_writer_add", i, vp, cp, i, i, C.POINTER(ll), vp, sz)
    _sig(L, "dtfn_bundle_writer_finish", i, vp)
    _sig(L, "dtfn_bundle_writer_abandon", None, vp)
    _sig(L, "dtfn_bundle_merge", i, C.POINTER(cp), i, cp)
    _sig(L, "dtfn_bundle_reader_open", vp, cp)
    _sig(L, "dtfn_bundle_reader_num_shards", i, vp)
    _sig(L, "dtfn_bundle_reader_count", ll, vp)
    _sig(L, "dtfn_bundle_reader_entry", i, vp, ll, C.POINTER(vp), C.POINTER(sz), C.POINTER(i), C.POINTER(i),
         C.POINTER(ll), C.POINTER(i), C.POINTER(ll), C.POINTER(ll), C.POINTER(u32))
    _sig(L, "dtfn_bundle_reader_lookup", i, vp, cp)
    _sig(L, "dtfn_bundle_reader_read", i, vp, ll, vp, sz)
    _sig(L, "dtfn_bundle_reader_num_slices", i, vp, ll)
    _sig(L, "dtfn_bundle_reader_close", None, vp)
    _sig(L, "dtfn_record_writer_open", vp, cp)
    _sig(L, "dtfn_record_writer_write", i, vp, vp, sz)
    _sig(L, "dtfn_record_writer_flush", i, vp)
    _sig(L, "dtfn_record_writer_close", None, vp)
    fp = C.POINTER(f)
    _sig(L, "dtfn_apply_sgd", None, fp, fp, sz, f)
    _sig(L, "dtfn_apply_adagrad", None, fp, fp, fp, sz, f)
    _sig(L, "dtfn_apply_adam", None, fp, fp, fp, fp, sz, f, f, f, f, f, f)
    _sig(L, "dtfn_apply_adadelta", None, fp, fp, fp, fp, sz, f, f, f)
    _sig(L, "dtfn_apply_ftrl", None, fp, fp, fp, fp, sz, f, f, f, f)
    _sig(L, "dtfn_apply_rmsprop", None, fp, fp, fp, fp, sz, f, f, f, f)
    _sig(L, "dtfn_server_start", vp, cp, i)
    _sig(L, "dtfn_server_port", i, vp)
    _sig(L, "dtfn_server_join", None, vp)
    _sig(L, "dtfn_server_stop", None, vp)
    _sig(L, "dtfn_server_inject_drop", None, vp, i)
    _sig(L, "dtfn_server_num_variables", ll, vp)
    _sig(L, "dtfn_client_connect", vp, cp, i, i)
    _sig(L, "dtfn_client_close", None, vp)
    _sig(L, "dtfn_client_set_timeout", None, vp, i)
    _sig(L, "dtfn_client_ping", i, vp)
    _sig(L, "dtfn_client_init_var", i, vp, cp, i, i, C.POINTER(ll), vp, sz, i)
    _sig(L, "dtfn_client_pull", ll, vp, cp, vp, sz)
    _sig(L, "dtfn_client_list", ll, vp, C.c_char_p, sz)
    _sig(L, "dtfn_client_var_info", i, vp, cp, C.POINTER(i), C.POINTER(ll), C.POINTER(ll))
    _sig(L, "dtfn_client_push_apply", i, vp, cp, i, f, fp, sz)
    _sig(L, "dtfn_client_push_pull", i, vp, i, C.POINTER(cp), i, f, C.POINTER(fp), C.POINTER(sz), cp, ll,
         C.POINTER(fp), C.POINTER(ll))
    _sig(L, "dtfn_client_assign_add_i64", i, vp, cp, ll, C.POINTER(ll))
    _sig(L, "dtfn_client_queue_create", i, vp, cp, i)
    _sig(L, "dtfn_client_enqueue", i, vp, cp, i)
    _sig(L, "dtfn_client_dequeue", i, vp, cp, C.POINTER(i), i)
    _sig(L, "dtfn_client_save", i, vp, cp, cp)
    _sig(L, "dtfn_client_shutdown_server", i, vp)
    _LIB = L
    return L


def last_error() -> str:
    return lib().dtfn_last_error().decode(errors="replace")


def check(rc: int, what: str) -> int:
    if rc < 0:
        raise NativeError("%s failed (%d): %s" % (what, rc, last_error()))
    return rc


def crc32c(data: bytes) -> int:
    return int(lib().dtfn_crc32c(data, len(data)))


def masked_crc32c(data: bytes) -> int:
    L = lib()
    return int(L.dtfn_crc32c_mask(L.dtfn_crc32c(data, len(data))))
