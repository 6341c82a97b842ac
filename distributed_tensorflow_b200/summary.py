"""TensorBoard event files: ``FileWriter(logdir).add_summary(summary, step)``.

The reference writes scalar summaries ``loss`` (and ``training/hptuning/metric`` in standalone mode)
through ``tf.summary.FileWriter`` (/root/reference/trainer/task.py:72-74, 80, 95, 98).  Files are
named ``events.out.tfevents.<unix ts>.<hostname>``, framed as TFRecords by the C++ record writer
(``csrc/native/bundle.cc``) and start with an ``Event{file_version: "brain.Event:2"}`` record.
Event/Summary messages are hand-encoded protobuf (only four fields are needed), so there is no
dependency on TensorFlow or TensorBoard for writing.
"""
from __future__ import annotations

import os
import socket
import struct
import threading
import time
from typing import Dict, Iterable, List, Optional, Tuple, Union

from . import _native


# ---- tiny protobuf encoder -------------------------------------------------------------------
def _varint(v: int) -> bytes:
    out = bytearray()
    v &= (1 << 64) - 1
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _key(field: int, wire: int) -> bytes:
    return _varint((field << 3) | wire)


def _bytes_field(field: int, b: bytes) -> bytes:
    return _key(field, 2) + _varint(len(b)) + b


class Summary:
    """``Summary{ repeated Value{tag=1, simple_value=2} value = 1 }``."""

    def __init__(self, values: Iterable[Tuple[str, float]] = ()):
        self.values: List[Tuple[str, float]] = [(str(t), float(v)) for t, v in values]

    def add_scalar(self, tag: str, value: float) -> "Summary":
        self.values.append((str(tag), float(value)))
        return self

    def SerializeToString(self) -> bytes:
        out = b""
        for tag, v in self.values:
            val = _bytes_field(1, tag.encode()) + _key(2, 5) + struct.pack("<f", v)
            out += _bytes_field(1, val)
        return out


class SummaryRegistry:
    """``tf.summary.scalar`` + ``merge_all``: named scalar slots evaluated into one Summary."""

    def __init__(self):
        self.tags: List[str] = []

    def scalar(self, tag: str) -> str:
        self.tags.append(tag)
        return tag

    def merge_all(self):
        tags = list(self.tags)

        def evaluate(values: Dict[str, float]) -> Summary:
            return Summary((t, values[t]) for t in tags)
        return evaluate


def _event(wall_time: float, step: int = 0, file_version: str = None, summary: bytes = None,
           graph_def: bytes = None) -> bytes:
    # Event{ double wall_time = 1; int64 step = 2; string file_version = 3; bytes graph_def = 4; Summary summary = 5 }
    out = _key(1, 1) + struct.pack("<d", wall_time)
    if step:
        out += _key(2, 0) + _varint(step)
    if file_version is not None:
        out += _bytes_field(3, file_version.encode())
    if graph_def is not None:
        out += _bytes_field(4, graph_def)
    if summary is not None:
        out += _bytes_field(5, summary)
    return out


class FileWriter:
    _lock = threading.Lock()
    _last_ts = 0

    def __init__(self, logdir: str, graph=None, filename_suffix: str = ""):
        os.makedirs(logdir, exist_ok=True)
        with FileWriter._lock:  # two writers in one process/second must not share a file name
            ts = max(int(time.time()), FileWriter._last_ts + 1)
            FileWriter._last_ts = ts
        self.logdir = logdir
        self.path = os.path.join(logdir, "events.out.tfevents.%010d.%s%s" % (ts, socket.gethostname(), filename_suffix))
        self._L = _native.lib()
        self._h = self._L.dtfn_record_writer_open(self.path.encode())
        if not self._h:
            raise _native.NativeError("FileWriter: %s" % _native.last_error())
        self._write(_event(time.time(), file_version="brain.Event:2"))
        if graph is not None:
            self.add_graph(graph)
        self.flush()

    def _write(self, record: bytes) -> None:
        if self._h is None:
            raise ValueError("FileWriter is closed")
        _native.check(self._L.dtfn_record_writer_write(self._h, record, len(record)), "event write")

    def add_graph(self, graph) -> None:
        """``graph`` is a serialized GraphDef (bytes) or an object with ``as_graph_def_bytes()``."""
        b = graph if isinstance(graph, (bytes, bytearray)) else graph.as_graph_def_bytes()
        self._write(_event(time.time(), graph_def=bytes(b)))

    def add_summary(self, summary: Union[Summary, bytes], global_step: Optional[int] = None) -> None:
        b = summary if isinstance(summary, (bytes, bytearray)) else summary.SerializeToString()
        self._write(_event(time.time(), step=int(global_step or 0), summary=bytes(b)))

    def flush(self) -> None:
        if self._h is not None:
            self._L.dtfn_record_writer_flush(self._h)

    def close(self) -> None:
        if self._h is not None:
            self._L.dtfn_record_writer_flush(self._h)
            self._L.dtfn_record_writer_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def read_events(path: str) -> List[dict]:
    """Decode an event file written by :class:`FileWriter` (verifies both CRCs of every record)."""
    out = []
    with open(path, "rb") as f:
        data = f.read()
    L = _native.lib()
    pos = 0
    while pos + 12 <= len(data):
        (n,) = struct.unpack_from("<Q", data, pos)
        (hcrc,) = struct.unpack_from("<I", data, pos + 8)
        if L.dtfn_crc32c_mask(L.dtfn_crc32c(data[pos:pos + 8], 8)) != hcrc:
            raise ValueError("bad length crc at %d" % pos)
        rec = data[pos + 12:pos + 12 + n]
        (dcrc,) = struct.unpack_from("<I", data, pos + 12 + n)
        if L.dtfn_crc32c_mask(L.dtfn_crc32c(rec, len(rec))) != dcrc:
            raise ValueError("bad data crc at %d" % pos)
        out.append(_decode_event(rec))
        pos += 12 + n + 4
    return out


def _decode_fields(b: bytes):
    i = 0
    while i < len(b):
        key = 0
        shift = 0
        while True:
            c = b[i]
            i += 1
            key |= (c & 0x7F) << shift
            shift += 7
            if not c & 0x80:
                break
        field, wire = key >> 3, key & 7
        if wire == 0:
            v = 0
            shift = 0
            while True:
                c = b[i]
                i += 1
                v |= (c & 0x7F) << shift
                shift += 7
                if not c & 0x80:
                    break
            yield field, wire, v
        elif wire == 1:
            yield field, wire, b[i:i + 8]
            i += 8
        elif wire == 5:
            yield field, wire, b[i:i + 4]
            i += 4
        elif wire == 2:
            n = 0
            shift = 0
            while True:
                c = b[i]
                i += 1
                n |= (c & 0x7F) << shift
                shift += 7
                if not c & 0x80:
                    break
            yield field, wire, b[i:i + n]
            i += n
        else:
            raise ValueError("unsupported wire type %d" % wire)


def _decode_event(rec: bytes) -> dict:
    ev = {"step": 0}
    for field, wire, v in _decode_fields(rec):
        if field == 1:
            ev["wall_time"] = struct.unpack("<d", v)[0]
        elif field == 2:
            ev["step"] = v
        elif field == 3:
            ev["file_version"] = v.decode()
        elif field == 4:
            ev["graph_def"] = v
        elif field == 5:
            vals = {}
            for f2, _, val in _decode_fields(v):
                if f2 == 1:
                    tag, sv = None, None
                    for f3, _, x in _decode_fields(val):
                        if f3 == 1:
                            tag = x.decode()
                        elif f3 == 2:
                            sv = struct.unpack("<f", x)[0]
                    vals[tag] = sv
            ev["summary"] = vals
    return ev
