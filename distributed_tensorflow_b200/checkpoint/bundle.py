"""Python face of the C++ tensor-bundle code (``csrc/native/bundle.cc``): TF checkpoint-V2 files.

``<prefix>.index`` + ``<prefix>.data-SSSSS-of-NNNNN`` as written by ``tf.train.Saver`` in the
reference (/root/reference/trainer/task.py:143; sharded variant trainer/task_supervisor.py:145).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Iterable, List, Sequence, Tuple

import numpy as np

from .. import _native

# tensorflow DataType enum values (types.proto)
DTYPE_TO_TF = {
    np.dtype(np.float32): 1, np.dtype(np.float64): 2, np.dtype(np.int32): 3, np.dtype(np.uint8): 4,
    np.dtype(np.int16): 5, np.dtype(np.int8): 6, np.dtype(np.int64): 9, np.dtype(np.bool_): 10,
    np.dtype(np.float16): 19,
}
TF_TO_DTYPE = {v: k for k, v in DTYPE_TO_TF.items()}
DT_BFLOAT16 = 14  # carried as raw uint16 on the NumPy side


class BundleWriter:
    """Writes a single-shard bundle.  Tensor names may be added in any order."""

    def __init__(self, prefix: str):
        self._L = _native.lib()
        self._h = self._L.dtfn_bundle_writer_new(prefix.encode())
        if not self._h:
            raise _native.NativeError("BundleWriter(%r): %s" % (prefix, _native.last_error()))
        self.prefix = prefix

    def add(self, name: str, array, tf_dtype: int = None) -> None:
        a = np.asarray(array)
        if not a.flags.c_contiguous:
            a = a.copy(order="C")
        if tf_dtype is None:
            if a.dtype not in DTYPE_TO_TF:
                raise TypeError("unsupported dtype %s for tensor %r" % (a.dtype, name))
            tf_dtype = DTYPE_TO_TF[a.dtype]
        dims = (C.c_longlong * max(1, a.ndim))(*a.shape)
        rc = self._L.dtfn_bundle_writer_add(self._h, name.encode(), tf_dtype, a.ndim, dims,
                                            a.ctypes.data_as(C.c_void_p), a.nbytes)
        _native.check(rc, "bundle add %r" % name)

    def finish(self) -> None:
        h, self._h = self._h, None
        _native.check(self._L.dtfn_bundle_writer_finish(h), "bundle finish")

    def __enter__(self):
        return self

    def __exit__(self, et, ev, tb):
        if self._h:
            if et is None:
                self.finish()
            else:
                self._L.dtfn_bundle_writer_abandon(self._h)
                self._h = None


class BundleReader:
    def __init__(self, prefix: str):
        self._L = _native.lib()
        self._h = self._L.dtfn_bundle_reader_open(prefix.encode())
        if not self._h:
            raise _native.NativeError("BundleReader(%r): %s" % (prefix, _native.last_error()))
        self.prefix = prefix
        self.num_shards = int(self._L.dtfn_bundle_reader_num_shards(self._h))
        self._entries: Dict[str, Tuple[int, int, Tuple[int, ...], int, int, int, int]] = {}
        self._slice_entries: Dict[bytes, Tuple[int, int, Tuple[int, ...], int, int, int, int]] = {}
        n = int(self._L.dtfn_bundle_reader_count(self._h))
        name_p, name_n = C.c_void_p(), C.c_size_t()
        dtype, ndims, shard = C.c_int(), C.c_int(), C.c_int()
        dims = (C.c_longlong * 8)()
        off, size, crc = C.c_longlong(), C.c_longlong(), C.c_uint32()
        for i in range(n):
            rc = self._L.dtfn_bundle_reader_entry(self._h, i, C.byref(name_p), C.byref(name_n), C.byref(dtype),
                                                  C.byref(ndims), dims, C.byref(shard), C.byref(off), C.byref(size),
                                                  C.byref(crc))
            _native.check(rc, "bundle entry")
            raw = C.string_at(name_p.value, name_n.value)
            rec = (i, dtype.value, tuple(dims[d] for d in range(ndims.value)), shard.value, off.value, size.value, crc.value)
            if raw[:1] == b"\x00":
                # slice of a partitioned variable, stored under an OrderedCode key (TF's EncodeTensorNameSlice);
                # reached through the full tensor's entry, not by name
                self._slice_entries[raw] = rec
                continue
            self._entries[raw.decode()] = rec

    def keys(self) -> List[str]:
        return list(self._entries)

    def has_tensor(self, name: str) -> bool:
        return name in self._entries

    def entry(self, name: str) -> dict:
        i, dt, shape, shard, off, size, crc = self._entries[name]
        return {"dtype": dt, "shape": shape, "shard_id": shard, "offset": off, "size": size, "crc32c": crc}

    def num_slices(self, name: str) -> int:
        """> 0 when ``name`` is a partitioned variable stored as slices (``get_tensor`` assembles them)."""
        return int(self._L.dtfn_bundle_reader_num_slices(self._h, self._entries[name][0]))

    def get_variable_to_shape_map(self) -> Dict[str, Tuple[int, ...]]:
        return {k: v[2] for k, v in self._entries.items()}

    def get_tensor(self, name: str) -> np.ndarray:
        i, dt, shape, shard, off, size, crc = self._entries[name]
        np_dtype = TF_TO_DTYPE.get(dt, np.dtype(np.uint16) if dt == DT_BFLOAT16 else None)
        if np_dtype is None:
            raise TypeError("tensor %r has unsupported TF dtype %d" % (name, dt))
        out = np.empty(shape, dtype=np_dtype)
        if out.nbytes != size:
            raise _native.NativeError("tensor %r: index says %d bytes, shape/dtype give %d" % (name, size, out.nbytes))
        rc = self._L.dtfn_bundle_reader_read(self._h, i, out.ctypes.data_as(C.c_void_p), out.nbytes)
        _native.check(rc, "bundle read %r" % name)
        return out

    def close(self) -> None:
        if self._h:
            self._L.dtfn_bundle_reader_close(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def merge_bundles(prefixes: Sequence[str], merged_prefix: str) -> None:
    """MergeV2Checkpoints: n single-shard bundles -> one n-shard bundle (sources are consumed)."""
    arr = (C.c_char_p * len(prefixes))(*[p.encode() for p in prefixes])
    _native.check(_native.lib().dtfn_bundle_merge(arr, len(prefixes), merged_prefix.encode()), "bundle merge")


def write_bundle(prefix: str, tensors: Dict[str, np.ndarray]) -> None:
    with BundleWriter(prefix) as w:
        for k in sorted(tensors):
            w.add(k, tensors[k])


def read_bundle(prefix: str, names: Iterable[str] = None) -> Dict[str, np.ndarray]:
    with BundleReader(prefix) as r:
        return {k: r.get_tensor(k) for k in (names if names is not None else r.keys())}
