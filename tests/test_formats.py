"""On-disk formats: crc32c, LevelDB-format table, TF checkpoint-V2 bundle, checkpoint state file,
TFRecord event files, SavedModel.  TensorFlow is not installable in this environment, so golden
bytes are hand-derived from the format specification and files are cross-checked
with TensorBoard's own readers where they exist."""
import os
import struct

import numpy as np
import pytest

from distributed_tensorflow_b200 import _native
from distributed_tensorflow_b200.checkpoint import bundle as B
from distributed_tensorflow_b200.checkpoint import saver as S
from distributed_tensorflow_b200.summary import FileWriter, Summary, read_events
import ctypes as C


def test_crc32c_known_answers():
    assert _native.crc32c(b"") == 0
    assert _native.crc32c(b"123456789") == 0xE3069283
    assert _native.crc32c(b"\x00" * 32) == 0x8A9136AA
    assert _native.crc32c(b"\xff" * 32) == 0x62A8AB43
    L = _native.lib()
    c = L.dtfn_crc32c(b"hello ", 6)
    assert L.dtfn_crc32c_extend(c, b"world", 5) == _native.crc32c(b"hello world")
    m = L.dtfn_crc32c_mask(0xE3069283)
    assert m == ((0xE3069283 >> 15 | 0xE3069283 << 17) + 0xA282EAD8) & 0xFFFFFFFF
    assert L.dtfn_crc32c_unmask(m) == 0xE3069283


def test_crc32c_matches_tensorboard_helper():
    tb = pytest.importorskip("tensorboard.compat.tensorflow_stub.pywrap_tensorflow")
    data = os.urandom(1000)
    assert _native.masked_crc32c(data) == tb.masked_crc32c(data)


def _table_roundtrip(tmp_path, items, block_size=0):
    L = _native.lib()
    path = str(tmp_path / "t.tbl").encode()
    b = L.dtfn_table_builder_new(path, block_size)
    for k, v in items:
        assert L.dtfn_table_builder_add(b, k, len(k), v, len(v)) == 0
    assert L.dtfn_table_builder_finish(b) == 0
    t = L.dtfn_table_open(path)
    assert t, _native.last_error()
    out = []
    kp, kl, vp, vl = C.c_void_p(), C.c_size_t(), C.c_void_p(), C.c_size_t()
    for i in range(L.dtfn_table_count(t)):
        L.dtfn_table_entry(t, i, C.byref(kp), C.byref(kl), C.byref(vp), C.byref(vl))
        out.append((C.string_at(kp.value, kl.value), C.string_at(vp.value, vl.value)))
    L.dtfn_table_close(t)
    return out, path.decode()


def test_table_roundtrip_many_blocks_and_footer(tmp_path):
    items = [(b"key%06d" % i, os.urandom(i % 50)) for i in range(2000)]
    out, path = _table_roundtrip(tmp_path, items, block_size=512)   # forces hundreds of data blocks
    assert out == items
    raw = open(path, "rb").read()
    assert struct.unpack("<Q", raw[-8:])[0] == 0xDB4775248B80FB57
    assert len(raw) > 48


def test_table_rejects_unsorted_keys_and_corruption(tmp_path):
    L = _native.lib()
    b = L.dtfn_table_builder_new(str(tmp_path / "u.tbl").encode(), 0)
    assert L.dtfn_table_builder_add(b, b"b", 1, b"", 0) == 0
    assert L.dtfn_table_builder_add(b, b"a", 1, b"", 0) < 0
    L.dtfn_table_builder_abandon(b)
    _, path = _table_roundtrip(tmp_path, [(b"a", b"1"), (b"b", b"2")])
    raw = bytearray(open(path, "rb").read())
    raw[3] ^= 0xFF
    open(path, "wb").write(raw)
    assert not L.dtfn_table_open(path.encode())
    assert "checksum" in _native.last_error()


def test_bundle_golden_bytes_single_scalar(tmp_path):
    """One f32 scalar 'w' = 1.0: every byte of .index up to the first block trailer is derivable."""
    prefix = str(tmp_path / "g")
    B.write_bundle(prefix, {"w": np.array(1.0, np.float32)})
    data = open(prefix + ".data-00000-of-00001", "rb").read()
    assert data == struct.pack("<f", 1.0)
    crc = _native.masked_crc32c(data)
    header = bytes([0x08, 0x01, 0x1A, 0x02, 0x08, 0x01])                      # num_shards=1, version{producer=1}
    entry = bytes([0x08, 0x01, 0x12, 0x00, 0x28, 0x04, 0x35]) + struct.pack("<I", crc)  # dtype, shape{}, size=4, crc
    block = (bytes([0, 0, len(header)]) + header + bytes([0, 1, len(entry)]) + b"w" + entry
             + struct.pack("<II", 0, 1))                                       # one restart at 0
    raw = open(prefix + ".index", "rb").read()
    assert raw[:len(block)] == block
    assert raw[len(block)] == 0                                               # no compression
    assert struct.unpack("<I", raw[len(block) + 1:len(block) + 5])[0] == _native.masked_crc32c(block + b"\0")


def test_bundle_roundtrip_dtypes_and_crc_check(tmp_path):
    prefix = str(tmp_path / "m")
    t = {"weight": np.array(2.0, np.float32), "bias": np.array(10.0, np.float32), "global_step": np.array(7, np.int32),
         "emb": np.random.randn(50, 8).astype(np.float32), "ids": np.arange(5, dtype=np.int64),
         "weight/Adagrad": np.array(0.1, np.float32)}
    B.write_bundle(prefix, t)
    with B.BundleReader(prefix) as r:
        assert r.keys() == sorted(t) and r.num_shards == 1
        assert r.get_variable_to_shape_map()["emb"] == (50, 8)
        assert r.entry("global_step")["dtype"] == 3 and r.entry("weight")["dtype"] == 1 and r.entry("ids")["dtype"] == 9
        for k, v in t.items():
            got = r.get_tensor(k)
            assert got.dtype == v.dtype and got.shape == v.shape and np.array_equal(got, v)
    raw = bytearray(open(prefix + ".data-00000-of-00001", "rb").read())
    raw[0] ^= 1
    open(prefix + ".data-00000-of-00001", "wb").write(raw)
    with B.BundleReader(prefix) as r, pytest.raises(_native.NativeError, match="checksum"):
        r.get_tensor(sorted(t)[0])


def test_bundle_merge_shards(tmp_path):
    p0, p1, m = str(tmp_path / "p0"), str(tmp_path / "p1"), str(tmp_path / "model.ckpt-5")
    B.write_bundle(p0, {"weight": np.array(1.0, np.float32), "global_step": np.array(5, np.int32)})
    B.write_bundle(p1, {"bias": np.array(2.0, np.float32)})
    B.merge_bundles([p0, p1], m)
    files = sorted(os.listdir(tmp_path))
    assert files == ["model.ckpt-5.data-00000-of-00002", "model.ckpt-5.data-00001-of-00002", "model.ckpt-5.index"]
    with B.BundleReader(m) as r:
        assert r.num_shards == 2
        assert r.entry("bias")["shard_id"] == 1 and r.entry("weight")["shard_id"] == 0
        assert float(r.get_tensor("bias")) == 2.0 and int(r.get_tensor("global_step")) == 5
    B.write_bundle(p0, {"dup": np.zeros(1, np.float32)})
    B.write_bundle(p1, {"dup": np.zeros(1, np.float32)})
    with pytest.raises(_native.NativeError, match="more than one shard"):
        B.merge_bundles([p0, p1], str(tmp_path / "x"))


class _Store:
    def __init__(self, **kw):
        self.t = {k: np.array(v) for k, v in kw.items()}

    def snapshot(self):
        return dict(self.t)

    def assign(self, k, v):
        self.t[k] = v


def test_saver_state_file_max_to_keep_and_restore(tmp_path):
    d = str(tmp_path / "ckpt")
    st = _Store(weight=np.float32(1), bias=np.float32(2), global_step=np.int32(0))
    sv = S.Saver(max_to_keep=5)
    for step in range(7):
        st.t["global_step"] = np.array(step, np.int32)
        p = sv.save(st, os.path.join(d, "model.ckpt"), global_step=step)
        assert p.endswith("model.ckpt-%d" % step)
    kept = sorted(f for f in os.listdir(d) if f.endswith(".index"))
    assert kept == ["model.ckpt-%d.index" % s for s in range(2, 7)]
    text = open(os.path.join(d, "checkpoint")).read().splitlines()
    assert text[0] == 'model_checkpoint_path: "model.ckpt-6"'
    assert text[1:] == ['all_model_checkpoint_paths: "model.ckpt-%d"' % s for s in range(2, 7)]
    assert S.latest_checkpoint(d) == os.path.join(d, "model.ckpt-6")
    assert os.path.exists(os.path.join(d, "model.ckpt-6.meta"))
    st2 = _Store()
    assert sorted(S.Saver().restore(st2, S.latest_checkpoint(d))) == ["bias", "global_step", "weight"]
    assert int(st2.t["global_step"]) == 6 and st2.t["global_step"].dtype == np.int32
    assert S.latest_checkpoint(str(tmp_path / "nope")) is None
    with pytest.raises(FileNotFoundError):
        S.Saver().restore(st2, os.path.join(d, "model.ckpt-0"))


def test_meta_file_parses_as_metagraphdef(tmp_path):
    mg_pb2 = pytest.importorskip("tensorboard.compat.proto.meta_graph_pb2")
    d = str(tmp_path)
    S.Saver(sharded=False).save(_Store(w=np.float32(1)), os.path.join(d, "model.ckpt"), global_step=3)
    mg = mg_pb2.MetaGraphDef()
    mg.ParseFromString(open(os.path.join(d, "model.ckpt-3.meta"), "rb").read())
    assert mg.saver_def.max_to_keep == 5 and mg.saver_def.version == 2
    assert [n.name for n in mg.graph_def.node] == ["w"]


def test_event_file_framing_and_tensorboard_loader(tmp_path):
    w = FileWriter(str(tmp_path), graph=b"")
    w.add_summary(Summary([("loss", 1.5), ("training/hptuning/metric", 1.5)]), 100)
    w.add_summary(Summary([("loss", 0.25)]), 200)
    w.close()
    assert os.path.basename(w.path).startswith("events.out.tfevents.")
    ev = read_events(w.path)
    assert ev[0]["file_version"] == "brain.Event:2"
    sums = [e for e in ev if "summary" in e]
    assert [(e["step"], e["summary"]) for e in sums] == [
        (100, {"loss": 1.5, "training/hptuning/metric": 1.5}), (200, {"loss": 0.25})]
    loader_mod = pytest.importorskip("tensorboard.backend.event_processing.event_file_loader")
    event_pb2 = pytest.importorskip("tensorboard.compat.proto.event_pb2")
    events = [event_pb2.Event.FromString(rec) for rec in loader_mod.RawEventFileLoader(w.path).Load()]
    assert events[0].file_version == "brain.Event:2" and events[1].HasField("graph_def")
    got = [(e.step, v.tag, v.simple_value) for e in events for v in e.summary.value]
    assert got == [(100, "loss", 1.5), (100, "training/hptuning/metric", 1.5), (200, "loss", 0.25)]


def test_merged_partitions_also_appear_as_one_tf_style_partitioned_variable(tmp_path):
    """``<var>/part_<r>`` entries written by the ranks are, after the merge, ALSO reachable the way TensorFlow stores a
    PartitionedVariable: an entry for the full tensor that lists its slices (BundleEntryProto.slices) plus one entry
    per slice under an OrderedCode key (EncodeTensorNameSlice), pointing at the same bytes -- no data is duplicated.
    Golden bytes of the keys are derived by hand from ordered_code.cc (no TensorFlow on this box to produce them)."""
    rng = np.random.RandomState(3)
    parts = [rng.randn(n, 4).astype(np.float32) for n in (3, 2, 2)]
    accs = [np.full((p.shape[0], 4), 0.1 + r, np.float32) for r, p in enumerate(parts)]
    dense = [np.arange(5, dtype=np.float32) + 10 * r for r in range(3)]
    prefixes = []
    for r in range(3):
        p = str(tmp_path / ("part-%05d" % r))
        t = {"embedding/part_%d" % r: parts[r], "embedding/part_%d/Adagrad" % r: accs[r], "dense_flat/part_%d" % r: dense[r]}
        if r == 0:
            t["global_step"] = np.array(7, np.int64)
        B.write_bundle(p, t)
        prefixes.append(p)
    m = str(tmp_path / "model.ckpt-7")
    size_before = sum(os.path.getsize(p + ".data-00000-of-00001") for p in prefixes)
    B.merge_bundles(prefixes, m)
    assert sum(os.path.getsize("%s.data-%05d-of-00003" % (m, r)) for r in range(3)) == size_before
    with B.BundleReader(m) as r:
        keys = r.keys()
        assert {"embedding", "embedding/Adagrad", "dense_flat", "embedding/part_0", "embedding/part_2/Adagrad", "global_step"} <= set(keys)
        assert r.num_slices("embedding") == 3 and r.num_slices("embedding/part_1") == 0 and r.num_slices("global_step") == 0
        assert r.get_variable_to_shape_map()["embedding"] == (7, 4) and r.get_variable_to_shape_map()["dense_flat"] == (15,)
        assert np.array_equal(r.get_tensor("embedding"), np.concatenate(parts))          # assembled from the slices
        assert np.array_equal(r.get_tensor("embedding/Adagrad"), np.concatenate(accs))
        assert np.array_equal(r.get_tensor("dense_flat"), np.concatenate(dense))
        assert np.array_equal(r.get_tensor("embedding/part_1"), parts[1])                 # the per-rank names still work
        # slice keys: 00 | "embedding" 00 01 | 01 02 (two dims) | start, length of dim 0 | 0x80 (start 0), 0x7f (length -1)
        want = [b"\x00embedding\x00\x01\x01\x02" + bytes([0x80 + s, 0x80 + n]) + b"\x80\x7f" for s, n in ((0, 3), (3, 2), (5, 2))]
        assert sorted(r._slice_entries) == sorted(want + [k.replace(b"embedding", b"embedding/Adagrad") for k in want]
                                                  + [b"\x00dense_flat\x00\x01\x01\x01" + bytes([0x80 + 5 * i, 0x85]) for i in range(3)])
        e = r._slice_entries[want[1]]
        assert (e[3], e[4], e[5]) == tuple(r.entry("embedding/part_1")[k] for k in ("shard_id", "offset", "size"))
    # a single partition is simply a second name for the same bytes
    p = str(tmp_path / "solo")
    B.write_bundle(p, {"embedding/part_0": parts[0]})
    B.merge_bundles([p], str(tmp_path / "solo.ckpt"))
    with B.BundleReader(str(tmp_path / "solo.ckpt")) as r:
        assert r.num_slices("embedding") == 0 and np.array_equal(r.get_tensor("embedding"), parts[0])


def test_ordered_code_signed_numbers_longer_than_one_byte(tmp_path):
    """Row offsets of real shards need the multi-byte form of WriteSignedNumIncreasing: 1 250 000 = 0x1312d0 has 21
    significant bits -> 4 bytes, header 0xf0 XORed onto the big-endian value: f0 13 12 d0; 64 -> c0 40; 8127 -> df bf."""
    parts = [np.zeros((1250000, 1), np.int8), np.zeros((64, 1), np.int8), np.zeros((8191 - 64, 1), np.int8), np.ones((5, 1), np.int8)]
    prefixes = []
    for r, a in enumerate(parts):
        p = str(tmp_path / ("p%d" % r))
        B.write_bundle(p, {"t/part_%d" % r: a})
        prefixes.append(p)
    B.merge_bundles(prefixes, str(tmp_path / "m"))
    with B.BundleReader(str(tmp_path / "m")) as r:
        keys = sorted(r._slice_entries)
        head = b"\x00t\x00\x01\x01\x02"
        assert head + b"\x80" + b"\xf0\x13\x12\xd0" + b"\x80\x7f" in keys                      # start 0, length 1 250 000
        assert head + b"\xf0\x13\x12\xd0" + b"\xc0\x40" + b"\x80\x7f" in keys                  # start 1 250 000, length 64
        assert head + b"\xf0\x13\x13\x10" + b"\xdf\xbf" + b"\x80\x7f" in keys                  # start 1 250 064, length 8127 = 0x1fbf
        assert int(r.get_tensor("t").sum()) == 5 and r.get_tensor("t").shape == (1250000 + 8191 + 5, 1)


def test_meta_file_next_to_a_checkpoint_is_an_importable_meta_graph(tmp_path):
    """``model.ckpt-N.meta`` (reference trainer/task.py:143,215-223: Saver handed to the Supervisor): graph with the
    saver sub-graph its saver_def names, and the variables / trainable_variables collections of VariableDef protos."""
    from tensorboard.compat.proto import meta_graph_pb2
    from distributed_tensorflow_b200.checkpoint.saver import Saver
    from distributed_tensorflow_b200.models.linear import LinearModel
    from distributed_tensorflow_b200.optim import get_optimizer
    from distributed_tensorflow_b200.session import LocalSession
    sess = LocalSession(LinearModel(), get_optimizer("adam", 0.01))
    sess.run_init()
    sess.train_step(np.array([[1.0]]), np.array([[12.0]]))
    prefix = Saver().save(sess, str(tmp_path / "model.ckpt"), global_step=1)
    mg = meta_graph_pb2.MetaGraphDef()
    mg.ParseFromString(open(prefix + ".meta", "rb").read())
    nodes = {n.name: n for n in mg.graph_def.node}
    sd = mg.saver_def
    assert sd.filename_tensor_name.split(":")[0] in nodes and sd.save_tensor_name.split(":")[0] in nodes and sd.restore_op_name in nodes
    for n in mg.graph_def.node:
        for i in n.input:
            assert i.lstrip("^").split(":")[0] in nodes
    saved = [s.decode() for s in nodes["save/SaveV2/tensor_names"].attr["value"].tensor.string_val]
    assert saved == ["beta1_power", "beta2_power", "bias", "bias/Adam", "bias/Adam_1", "global_step", "weight", "weight/Adam", "weight/Adam_1"]
    vdefs = mg.collection_def["variables"].bytes_list.value
    assert len(vdefs) == 9
    w = [v for v in vdefs if v.startswith(b"\n\x08weight:0")][0]
    assert w == b"\n\x08weight:0\x12\rweight/Assign\x1a\rweight/read:02\x16weight/initial_value:08\x01"
    tv = mg.collection_def["trainable_variables"].bytes_list.value
    assert sorted(v[2:2 + v[1]] for v in tv) == [b"bias:0", b"weight:0"]
