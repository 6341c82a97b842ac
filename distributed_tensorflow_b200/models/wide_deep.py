"""Wide-and-deep on the B200-native parameter server (north-star configs 3 and 4).

Model::

    deep : concat(emb[id_1] .. emb[id_F], dense[0:dense_in] zero-padded to dense_pad)
           -> Linear(K0, h1) ReLU -> Linear(h1, h2) ReLU -> Linear(h2, h3) ReLU -> Linear(h3, 1)
    wide : sum_f wide_w[id_f] + wide_b
    loss : mean sigmoid cross-entropy of (deep + wide) against the label

Every rank is worker and parameter-server shard at once: embedding / wide row ``r`` lives in the HBM
of rank ``r % world``, the flat dense parameter vector is range-partitioned.  The training step is
executed by the C++ plan (``csrc/kernels/wd_plan.cu``): one ctypes call launches route, fused pull
(peer-HBM gather + tcgen05 GEMM-1), the MLP, loss, backward, fused push (dgrad GEMM scattering rows
into the owners' mailboxes) and the ps-side optimizer applies.  This is the scale-up of the
reference's ps/worker training loop (/root/reference/trainer/task.py:123-143, 232-246) -- async
(Hogwild) like the reference, plus the sync-replica mode of the north star.

:class:`TorchReference` restates the same step in plain PyTorch fp32 and is the numerics oracle.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, asdict
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from ..ops import _lib

OPT_IDS = {"sgd": 0, "adadelta": 1, "adagrad": 2, "adam": 3, "ftrl": 4, "rmsprop": 5}   # same ids as dtf_native.h


@dataclass
class WideDeepConfig:
    emb_rows: int = 10_000_000
    emb_dim: int = 128
    slots: int = 26
    dense_in: int = 13
    dense_pad: int = 64
    h1: int = 1024
    h2: int = 512
    h3: int = 256
    batch: int = 8192            # per rank
    optimizer: str = "adagrad"
    lr: float = 0.05
    sync: bool = False           # False = async / Hogwild (the reference's semantics)
    emb_init_scale: float = 0.05
    seed: int = 1234

    @property
    def k0(self) -> int:
        return self.slots * self.emb_dim + self.dense_pad


class _CConfig(C.Structure):
    _fields_ = [("world", C.c_int), ("rank", C.c_int), ("batch", C.c_int), ("slots", C.c_int), ("emb_dim", C.c_int),
                ("dense_in", C.c_int), ("dense_pad", C.c_int), ("h1", C.c_int), ("h2", C.c_int), ("h3", C.c_int),
                ("emb_rows", C.c_longlong), ("optimizer", C.c_int), ("sync", C.c_int), ("no_device_barrier", C.c_int),
                ("lr", C.c_float), ("emb_init_scale", C.c_float), ("overlap_apply_blocks", C.c_int)]


class _CInfo(C.Structure):
    _fields_ = [("K0", C.c_int)] + [(n, C.c_longlong) for n in (
        "rows_local", "P", "P_pad", "dense_lo", "dense_hi", "off_w1", "off_w2", "off_w3", "off_b1", "off_b2", "off_b3",
        "off_w4", "off_b4", "off_wide_b", "small_begin", "slab_bytes", "step")]


class _DevArray:
    """Zero-copy view of a raw device pointer for ``torch.as_tensor``."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (ptr, False), "version": 2}


_TYPESTR = {torch.float32: ("<f4", 4), torch.int32: ("<i4", 4), torch.uint8: ("|u1", 1), torch.int16: ("<i2", 2),
            torch.int64: ("<i8", 8)}


def _bind(L):
    if getattr(L, "_wd_bound", False):
        return
    vp, i = C.c_void_p, C.c_int
    L.dtf_wd_create.restype = vp
    L.dtf_wd_create.argtypes = [C.POINTER(_CConfig)]
    L.dtf_wd_destroy.restype = None
    L.dtf_wd_destroy.argtypes = [vp]
    L.dtf_wd_info.argtypes = [vp, C.POINTER(_CInfo)]
    L.dtf_wd_buffer.argtypes = [vp, C.c_char_p, C.POINTER(vp), C.POINTER(C.c_size_t)]
    L.dtf_wd_slab_handle.argtypes = [vp, vp]
    L.dtf_wd_slab_ptr.restype = vp
    L.dtf_wd_slab_ptr.argtypes = [vp]
    L.dtf_wd_attach_ipc.argtypes = [vp, vp]
    L.dtf_wd_attach_ptrs.argtypes = [vp, C.POINTER(vp)]
    L.dtf_wd_init_params.argtypes = [vp, C.c_ulonglong, vp]
    L.dtf_wd_broadcast_dense.argtypes = [vp, vp]
    L.dtf_wd_refresh_embedding.argtypes = [vp, vp]
    for fn in (L.dtf_wd_step_compute, L.dtf_wd_step):
        fn.argtypes = [vp, vp, vp, vp, vp]
    L.dtf_wd_barrier.argtypes = [vp, i, vp]
    L.dtf_wd_step_apply.argtypes = [vp, vp]
    L.dtf_wd_predict.argtypes = [vp, vp, vp, i, vp]
    L.dtf_wd_read_loss.argtypes = [vp, C.POINTER(C.c_float), vp]
    L.dtf_wd_loss_ptr.restype = vp
    L.dtf_wd_loss_ptr.argtypes = [vp]
    L.dtf_wd_last_parity.argtypes = [vp]
    L.dtf_wd_step_overlapped.argtypes = [vp, vp, vp, vp, vp, vp]
    L.dtf_wd_join.argtypes = [vp, vp]
    L.dtf_wd_set_next_batch.argtypes = [vp, vp, vp]
    L.dtf_wd_prefetched_steps.restype = C.c_longlong
    L.dtf_wd_prefetched_steps.argtypes = [vp]
    L.dtf_wd_overlap_reset.restype = None
    L.dtf_wd_overlap_reset.argtypes = [vp]
    L.dtf_symm_handle_size.restype = i
    L._wd_bound = True


class WideDeepTrainer:
    """One rank of the wide-and-deep job (worker + ps shard)."""

    name = "wide_deep"

    def __init__(self, cfg: WideDeepConfig, rank: int = 0, world: int = 1, device: Optional[torch.device] = None,
                 no_device_barrier: bool = False):
        if not torch.cuda.is_available():
            raise RuntimeError("wide_deep needs a CUDA device (B200); the toy linear model is the CPU path")
        self.cfg, self.rank, self.world = cfg, rank, world
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self._L = _lib.lib()
        _bind(self._L)
        cc = _CConfig(world=world, rank=rank, batch=cfg.batch, slots=cfg.slots, emb_dim=cfg.emb_dim,
                      dense_in=cfg.dense_in, dense_pad=cfg.dense_pad, h1=cfg.h1, h2=cfg.h2, h3=cfg.h3,
                      emb_rows=cfg.emb_rows, optimizer=OPT_IDS[cfg.optimizer], sync=int(cfg.sync),
                      no_device_barrier=int(no_device_barrier), lr=cfg.lr, emb_init_scale=cfg.emb_init_scale,
                      overlap_apply_blocks=int(os.environ.get("DTF_APPLY_BLOCKS", "0")))
        with torch.cuda.device(self.device):
            self._h = self._L.dtf_wd_create(C.byref(cc))
        if not self._h:
            raise _lib.KernelError("dtf_wd_create: " + self._L.dtf_last_error().decode())
        self.info = _CInfo()
        self._L.dtf_wd_info(self._h, C.byref(self.info))
        # The partitioning rule has one statement in Python (placement.RowShard / RangeShard: what checkpoints,
        # export, serving and the tests use to address shards) and one in the C++ executor that lays the
        # buffers out; a disagreement would silently mis-assemble every checkpoint, so it is fatal here.
        from ..placement import RangeShard, RowShard
        self.row_shard = RowShard(rows=cfg.emb_rows, world=world)
        self.dense_shard = RangeShard(n=int(self.info.P_pad), world=world, align=256)
        if (self.row_shard.local_rows(rank) != self.info.rows_local
                or self.dense_shard.bounds(rank) != (self.info.dense_lo, self.info.dense_hi)):
            raise _lib.KernelError("placement mismatch between placement.py and the C++ plan: rows %d vs %d, dense %s vs %s"
                                   % (self.row_shard.local_rows(rank), self.info.rows_local, self.dense_shard.bounds(rank),
                                      (self.info.dense_lo, self.info.dense_hi)))
        self._attached = world == 1
        self._views: Dict[str, torch.Tensor] = {}
        # loss accumulators live after cnt[16] + counter[4] in the per-step scratch (one per parity)
        self._loss_views = [self.buffer("step_scratch", torch.float32)[20:21]]
        if not cfg.sync:
            self._loss_views.append(self.buffer("step_scratch1", torch.float32)[20:21])
        self.loss_view = self._loss_views[0]
        self._apply_stream: Optional[torch.cuda.Stream] = None

    # ------------------------------------------------------------------ set-up
    def slab_handle(self) -> bytes:
        n = self._L.dtf_symm_handle_size()
        buf = C.create_string_buffer(n)
        _lib.check(self._L.dtf_wd_slab_handle(self._h, buf), "dtf_wd_slab_handle")
        return buf.raw

    def slab_ptr(self) -> int:
        return int(self._L.dtf_wd_slab_ptr(self._h))

    def attach_ipc(self, handles: List[bytes]) -> None:
        blob = b"".join(handles)
        with torch.cuda.device(self.device):
            _lib.check(self._L.dtf_wd_attach_ipc(self._h, blob), "dtf_wd_attach_ipc")
        self._attached = True

    def attach_peers(self, dist_module=None) -> None:
        """Map every peer's slab (collective over the default process group).  Slabs are VMM
        allocations shared as file descriptors over Unix sockets; CUDA-IPC handles are only the
        fallback (their mappings serve fine-grained remote reads about 3x slower)."""
        from ..parallel import symm
        if self.world == 1:
            return
        self._L.dtf_wd_slab_is_vmm.argtypes = [C.c_void_p]
        self._L.dtf_wd_slab_fd.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        self._L.dtf_wd_attach_fds.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        import torch.distributed as dist
        d = dist_module or dist
        flags = [None] * self.world
        d.all_gather_object(flags, bool(self._L.dtf_wd_slab_is_vmm(self._h)))
        if all(flags):
            fd = C.c_int(-1)
            _lib.check(self._L.dtf_wd_slab_fd(self._h, C.byref(fd)), "dtf_wd_slab_fd")
            fds = symm.exchange_fds(fd.value, self.rank, self.world, d)
            arr = (C.c_int * self.world)(*[f if f is not None else -1 for f in fds])
            with torch.cuda.device(self.device):
                _lib.check(self._L.dtf_wd_attach_fds(self._h, arr), "dtf_wd_attach_fds")
            symm.close_fds([fd.value] + [f for f in fds if f is not None])
            self._attached = True
            # measured dense apply, NVLS vs peer loads/stores: 44 vs 33 us on 2 GPUs, 35 vs 45 us on 8
            if os.environ.get("DTF_MULTICAST", "1" if self.world >= 4 else "0") != "0":
                self._attach_multicast(d)
        else:
            handles = [None] * self.world
            d.all_gather_object(handles, self.slab_handle())
            self.attach_ipc(handles)
        d.barrier()

    multicast = False

    def _attach_multicast(self, d) -> None:
        """Collective, best effort: one multicast object (NVLS) over all ranks' slabs, so that the dense
        apply sums the replicas' gradients inside NVSwitch (``multimem.ld_reduce``) and broadcasts the
        new weights with one ``multimem.st``.  Any failure on any rank leaves every rank on the plain
        peer load / store path."""
        from ..parallel import symm
        L = self._L
        L.dtf_mc_usable_bytes.restype = C.c_longlong
        L.dtf_mc_usable_bytes.argtypes = [C.c_int, C.c_size_t]
        L.dtf_mc_create.argtypes = [C.c_int, C.c_size_t, C.POINTER(C.c_longlong), C.POINTER(C.c_int)]
        L.dtf_mc_import.argtypes = [C.c_int, C.c_size_t, C.POINTER(C.c_longlong)]
        L.dtf_mc_add_device.argtypes = [C.c_longlong]
        L.dtf_mc_bind_map.argtypes = [C.c_longlong, C.c_void_p, C.POINTER(C.c_void_p)]
        L.dtf_wd_attach_multicast.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong]

        def agree(ok: bool) -> bool:
            flags = [None] * self.world
            d.all_gather_object(flags, bool(ok))
            return all(flags)

        with torch.cuda.device(self.device):
            L.dtf_symm_mapped_size.restype = C.c_longlong
            L.dtf_symm_mapped_size.argtypes = [C.c_void_p]
            mapped = int(L.dtf_symm_mapped_size(C.c_void_p(self.slab_ptr())))
            nbytes = int(L.dtf_mc_usable_bytes(self.world, mapped)) if (L.dtf_mc_supported() and mapped > 0) else 0
            sizes = [None] * self.world
            d.all_gather_object(sizes, nbytes)
            nbytes = min(sizes)
            if nbytes <= 0:
                return
            mc_id, fd = C.c_longlong(0), C.c_int(-1)
            ok = True
            if self.rank == 0:
                ok = L.dtf_mc_create(self.world, nbytes, C.byref(mc_id), C.byref(fd)) == 0
            if not agree(ok):
                return
            my_fd = fd.value if self.rank == 0 else os.open(os.devnull, os.O_RDONLY)
            fds = symm.exchange_fds(my_fd, self.rank, self.world, d)
            if self.rank != 0:
                ok = L.dtf_mc_import(fds[0], nbytes, C.byref(mc_id)) == 0
            symm.close_fds([my_fd] + [f for f in fds if f is not None])
            if not agree(ok):
                return
            if not agree(L.dtf_mc_add_device(mc_id) == 0):
                return
            ptr = C.c_void_p()
            ok = L.dtf_mc_bind_map(mc_id, C.c_void_p(self.slab_ptr()), C.byref(ptr)) == 0
            ok = ok and L.dtf_wd_attach_multicast(self._h, ptr, nbytes) == 0
            if not agree(ok):
                L.dtf_wd_attach_multicast(self._h, None, 0)
                return
            self.multicast = True
            self._mc_id = int(mc_id.value)

    def attach_local(self, peers: List["WideDeepTrainer"]) -> None:
        """Peers living in this process (several ranks emulated on one device, or one process
        driving several GPUs)."""
        arr = (C.c_void_p * self.world)(*[p.slab_ptr() for p in peers])
        _lib.check(self._L.dtf_wd_attach_ptrs(self._h, arr), "dtf_wd_attach_ptrs")
        self._attached = True

    def init_params(self, seed: Optional[int] = None) -> None:
        assert self._attached, "attach peers before initialising (owners broadcast their dense shard)"
        with torch.cuda.device(self.device):
            _lib.check(self._L.dtf_wd_init_params(self._h, seed if seed is not None else self.cfg.seed, self._stream()),
                       "dtf_wd_init_params")

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # ------------------------------------------------------------------ buffers
    def buffer(self, name: str, dtype=torch.float32, shape=None) -> torch.Tensor:
        """Zero-copy torch view of a named plan buffer (bf16 buffers: pass torch.bfloat16)."""
        key = "%s/%s/%s" % (name, dtype, shape)
        if key in self._views:
            return self._views[key]
        ptr, nbytes = C.c_void_p(), C.c_size_t()
        _lib.check(self._L.dtf_wd_buffer(self._h, name.encode(), C.byref(ptr), C.byref(nbytes)), "dtf_wd_buffer")
        base_dtype = torch.int16 if dtype == torch.bfloat16 else dtype
        typestr, isz = _TYPESTR[base_dtype]
        n = nbytes.value // isz
        with torch.cuda.device(self.device):
            t = torch.as_tensor(_DevArray(ptr.value, (n,), typestr), device=self.device)
        if dtype == torch.bfloat16:
            t = t.view(torch.bfloat16)
        if shape is not None:
            t = t[: int(np.prod(shape))].view(*shape)
        self._views[key] = t
        return t

    def dense_views(self) -> Dict[str, torch.Tensor]:
        """fp32 views into this rank's *gradient* buffer and (world == 1 only) master weights, by layer."""
        c, inf = self.cfg, self.info
        out = {}
        for src, prefix in (("dense_grad", "grad_"), ("dense_master", "")):
            if src == "dense_master" and self.world != 1:
                continue
            flat = self.buffer(src, torch.float32)
            out[prefix + "w1"] = flat[inf.off_w1: inf.off_w1 + c.h1 * inf.K0].view(c.h1, inf.K0)
            out[prefix + "w2"] = flat[inf.off_w2: inf.off_w2 + c.h2 * c.h1].view(c.h2, c.h1)
            out[prefix + "w3"] = flat[inf.off_w3: inf.off_w3 + c.h3 * c.h2].view(c.h3, c.h2)
            out[prefix + "b1"] = flat[inf.off_b1: inf.off_b1 + c.h1]
            out[prefix + "b2"] = flat[inf.off_b2: inf.off_b2 + c.h2]
            out[prefix + "b3"] = flat[inf.off_b3: inf.off_b3 + c.h3]
            out[prefix + "w4"] = flat[inf.off_w4: inf.off_w4 + c.h3]
            out[prefix + "b4"] = flat[inf.off_b4: inf.off_b4 + 1]
            out[prefix + "wide_b"] = flat[inf.off_wide_b: inf.off_wide_b + 1]
        return out

    # ------------------------------------------------------------------ stepping
    def _ptrs(self, ids, dense, labels=None):
        assert ids.dtype == torch.int32 and ids.is_cuda and ids.is_contiguous() and ids.numel() == self.cfg.batch * self.cfg.slots
        assert dense.dtype == torch.float32 and dense.is_contiguous() and dense.numel() == self.cfg.batch * self.cfg.dense_in
        if labels is not None:
            assert labels.dtype == torch.float32 and labels.numel() == self.cfg.batch
        return C.c_void_p(ids.data_ptr()), C.c_void_p(dense.data_ptr()), C.c_void_p(labels.data_ptr() if labels is not None else 0)

    def step(self, ids: torch.Tensor, dense: torch.Tensor, labels: torch.Tensor, overlap: Optional[bool] = None,
             next_batch: Optional[Tuple[torch.Tensor, torch.Tensor]] = None) -> None:
        """One full training step on device-resident inputs (asynchronous on the current stream).

        In async mode (``cfg.sync == False``) the parameter-server side (barrier + applies) runs on an
        internal second stream and overlaps the next step's compute phase, unless ``overlap=False``.

        ``next_batch=(ids, dense)`` (async mode) announces the NEXT step's inputs: device tensors that are
        already filled on the current stream and stay untouched until that step has run.  Their embedding rows
        are then gathered on a low-priority stream while this step computes, and the next ``step`` -- called
        with that same ``ids`` tensor -- runs layer 1 as a plain GEMM on them instead of the fused pull (a
        parameter prefetch: rows one more step stale, as a TF worker that reads its variables early)."""
        a, b, c = self._ptrs(ids, dense, labels)
        if overlap is None:
            overlap = not self.cfg.sync
        if next_batch is not None and not self.cfg.sync:
            na, nb, _ = self._ptrs(next_batch[0], next_batch[1])
            self._L.dtf_wd_set_next_batch(self._h, na, nb)
        if overlap:
            if self._apply_stream is None:
                self._apply_stream = torch.cuda.Stream(device=self.device)
            _lib.check(self._L.dtf_wd_step_overlapped(self._h, a, b, c, self._stream(),
                                                      C.c_void_p(self._apply_stream.cuda_stream)), "dtf_wd_step_overlapped")
        else:
            _lib.check(self._L.dtf_wd_step(self._h, a, b, c, self._stream()), "dtf_wd_step")

    def prefetched_steps(self) -> int:
        """How many steps ran layer 1 on prefetched rows (see ``step(next_batch=...)``)."""
        return int(self._L.dtf_wd_prefetched_steps(self._h))

    def stamps_count(self) -> int:
        """Marks issued so far (``DTF_STAMPS=1`` diagnostics; -1 when off)."""
        self._L.dtf_wd_stamps_count.restype = C.c_longlong
        self._L.dtf_wd_stamps_count.argtypes = [C.c_void_p]
        return int(self._L.dtf_wd_stamps_count(self._h))

    def stamps_read(self, start: int = 0, stop: Optional[int] = None, clear: bool = False) -> str:
        """``index name microseconds`` per phase mark in [start, stop): the device-side timeline of the
        overlapped streams.  Reads on a private stream, so it answers while the training streams are stuck."""
        buf = C.create_string_buffer(1 << 20)
        self._L.dtf_wd_stamps_read.argtypes = [C.c_void_p, C.c_longlong, C.c_longlong, C.c_int, C.c_char_p, C.c_size_t]
        stop = self.stamps_count() if stop is None else stop
        if int(self._L.dtf_wd_stamps_read(self._h, start, stop, int(clear), buf, len(buf))) != 0:
            return ""
        return buf.value.decode()

    def set_barrier_timeout_ms(self, ms: int) -> None:
        """How long a device barrier waits for a peer before it records an error (0 = forever)."""
        self._L.dtf_wd_set_barrier_timeout_ms.restype = None
        self._L.dtf_wd_set_barrier_timeout_ms.argtypes = [C.c_longlong]
        self._L.dtf_wd_set_barrier_timeout_ms(int(ms))

    def check_error(self) -> None:
        """Raise ``ConnectionError`` if a device barrier of this plan gave up on a peer (dead or stalled
        rank).  Cheap: one 4-byte read on a private stream, the training streams are not synchronised."""
        self._L.dtf_wd_check_error.argtypes = [C.c_void_p]
        rc = int(self._L.dtf_wd_check_error(self._h))
        if rc > 0:
            raise ConnectionError("rank %d: device barrier timed out waiting for rank %d (peer dead or stalled)"
                                  % (self.rank, rc - 1))
        if rc < 0:
            raise RuntimeError("dtf_wd_check_error: " + _lib.lib().dtf_last_error().decode())

    def join(self) -> None:
        """Make the current stream wait for every ps-side apply issued so far (call before reading
        shards, predicting, checkpointing, or ending a CUDA-graph capture of overlapped steps)."""
        _lib.check(self._L.dtf_wd_join(self._h, self._stream()), "dtf_wd_join")

    def overlap_reset(self) -> None:
        """Forget cross-step event dependencies (after a full device synchronisation)."""
        self._L.dtf_wd_overlap_reset(self._h)

    def step_compute(self, ids, dense, labels) -> None:
        a, b, c = self._ptrs(ids, dense, labels)
        _lib.check(self._L.dtf_wd_step_compute(self._h, a, b, c, self._stream()), "dtf_wd_step_compute")

    def barrier(self, publish_counts: bool) -> None:
        _lib.check(self._L.dtf_wd_barrier(self._h, int(publish_counts), self._stream()), "dtf_wd_barrier")

    def step_apply(self) -> None:
        _lib.check(self._L.dtf_wd_step_apply(self._h, self._stream()), "dtf_wd_step_apply")

    def predict(self, ids: torch.Tensor, dense: torch.Tensor, n: Optional[int] = None) -> torch.Tensor:
        """Probabilities for the first ``n`` examples (default: a full batch).  ``ids`` / ``dense`` may
        hold just ``n`` rows; the result is a view of the plan's ``prob`` buffer."""
        n = int(n if n is not None else ids.shape[0])
        self.join()
        assert ids.dtype == torch.int32 and ids.is_cuda and ids.is_contiguous() and ids.numel() >= n * self.cfg.slots
        assert dense.dtype == torch.float32 and dense.is_contiguous() and dense.numel() >= n * self.cfg.dense_in
        _lib.check(self._L.dtf_wd_predict(self._h, C.c_void_p(ids.data_ptr()), C.c_void_p(dense.data_ptr()), n,
                                          self._stream()), "dtf_wd_predict")
        return self.buffer("prob", torch.float32)[:n]

    # ------------------------------------------------------------------ parameter-server shard state
    # TF slot names per optimizer, in the order slot1, slot2 of the plan
    _SLOT_NAMES = {"adagrad": ("Adagrad",), "adam": ("Adam", "Adam_1"), "sgd": (), "adadelta": ("Adadelta", "Adadelta_1"),
                   "ftrl": ("Ftrl", "Ftrl_1"), "rmsprop": ("RMSProp", "RMSProp_1")}

    def named_shards(self) -> Dict[str, torch.Tensor]:
        self.join()
        return self._named_shards()

    def _named_shards(self) -> Dict[str, torch.Tensor]:
        """This rank's ps shard as named device tensors (fp32 masters + optimizer slots), using the
        ``<variable>/part_<rank>`` convention for partitioned variables and TF slot names."""
        c, r = self.cfg, self.rank
        n = self.info.rows_local
        out = {"embedding/part_%d" % r: self.buffer("emb_master", torch.float32)[: n * c.emb_dim].view(n, c.emb_dim),
               "wide_weights/part_%d" % r: self.buffer("wide", torch.float32)[:n],
               "dense_flat/part_%d" % r: self.buffer("dense_master", torch.float32)}
        for i, slot in enumerate(self._SLOT_NAMES[c.optimizer]):
            k = "slot%d" % (i + 1)
            out["embedding/part_%d/%s" % (r, slot)] = self.buffer("emb_" + k, torch.float32)[: n * c.emb_dim].view(n, c.emb_dim)
            out["wide_weights/part_%d/%s" % (r, slot)] = self.buffer("wide_" + k, torch.float32)[:n]
            out["dense_flat/part_%d/%s" % (r, slot)] = self.buffer("dense_" + k, torch.float32)
        return out

    def load_shards(self, tensors: Dict[str, "np.ndarray"], step: Optional[int] = None) -> None:
        """Inverse of :meth:`named_shards` (restore): copies host arrays into the shard, refreshes the
        bf16 working copies and re-broadcasts this rank's dense range to its peers."""
        with torch.cuda.device(self.device):
            for name, dst in self.named_shards().items():
                if name not in tensors:
                    raise KeyError("checkpoint has no tensor %r (was it written with world=%d?)" % (name, self.world))
                src = torch.from_numpy(np.ascontiguousarray(tensors[name])).to(self.device)
                dst.copy_(src.view_as(dst))
            _lib.check(self._L.dtf_wd_refresh_embedding(self._h, self._stream()), "dtf_wd_refresh_embedding")
            _lib.check(self._L.dtf_wd_broadcast_dense(self._h, self._stream()), "dtf_wd_broadcast_dense")
            torch.cuda.synchronize(self.device)
        if step is not None:
            self.set_global_step(int(step))

    step_offset = 0   # kept for callers that add it; the device counter itself is restored now

    def export_config(self) -> dict:
        d = asdict(self.cfg)
        d["world"] = self.world
        return d

    def profile(self, on: bool) -> None:
        """Per-kernel tracing: CUDA events between the launches of every executor call."""
        self._L.dtf_wd_profile_enable.restype = None
        self._L.dtf_wd_profile_enable.argtypes = [C.c_void_p, C.c_int]
        self._L.dtf_wd_profile_enable(self._h, int(on))

    def profile_read(self) -> List[tuple]:
        """[(kernel group, microseconds), ...] since the last read (synchronises the device)."""
        self._L.dtf_wd_profile_read.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
        buf = C.create_string_buffer(1 << 16)
        self._L.dtf_wd_profile_read(self._h, buf, len(buf))
        out = []
        for line in buf.value.decode().splitlines():
            k, _, v = line.rpartition("=")
            out.append((k, float(v)))
        return out

    def last_loss_sum(self) -> torch.Tensor:
        """Device scalar: sum over the micro-batch of the per-example loss of the last step."""
        return self._loss_views[int(self._L.dtf_wd_last_parity(self._h)) if not self.cfg.sync else 0]

    def global_step(self) -> int:
        """Number of applies this shard has performed -- read from the device-resident step state, so
        it also counts steps executed by CUDA-graph replays (synchronises the device)."""
        torch.cuda.synchronize(self.device)
        return int(self.buffer("step_state", torch.int64)[0].item())

    def set_global_step(self, step: int) -> None:
        """Restore path: also the ``t`` of Adam's bias correction."""
        torch.cuda.synchronize(self.device)
        self.buffer("step_state", torch.int64)[0:1].fill_(int(step))
        torch.cuda.synchronize(self.device)

    def close(self) -> None:
        if self._h:
            self._views.clear()
            if getattr(self, "_mc_id", None) is not None and os.environ.get("DTF_MC_DESTROY", "0") == "1":
                # explicit teardown of the NVLS object (unmap + release, once); by default it is left to process exit
                self._L.dtf_wd_attach_multicast(self._h, None, 0)
                self._L.dtf_mc_destroy.argtypes = [C.c_longlong]
                self._L.dtf_mc_destroy(self._mc_id)
                self._mc_id = None
            self._L.dtf_wd_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class TorchReference:
    """The same step in plain PyTorch (fp32 math on bf16-rounded operands) for ``world`` replicas that
    share one parameter set -- the oracle for the CUDA path."""

    def __init__(self, cfg: WideDeepConfig, emb: torch.Tensor, wide: torch.Tensor, dense: Dict[str, torch.Tensor]):
        self.cfg = cfg
        self.emb, self.wide = emb.clone().float(), wide.clone().float()
        self.p = {k: v.clone().float() for k, v in dense.items()}
        init = {"adagrad": 0.1, "ftrl": 0.1, "rmsprop": 1.0}.get(cfg.optimizer, 0.0)
        self.s1 = {k: torch.full_like(v, init) for k, v in self._all().items()}
        self.s2 = {k: torch.zeros_like(v) for k, v in self._all().items()}
        self.t = 0

    def _all(self):
        d = dict(self.p)
        d["emb"], d["wide"] = self.emb, self.wide
        return d

    @staticmethod
    def _bf(x):
        return x.to(torch.bfloat16).float()

    def forward(self, ids, dense):
        c = self.cfg
        B = ids.shape[0]
        e = self._bf(self.emb)[ids.long()].reshape(B, c.slots * c.emb_dim)
        x = torch.zeros(B, c.dense_pad, device=ids.device)
        x[:, :c.dense_in] = dense
        a = torch.cat([e, self._bf(x)], dim=1)
        p = self.p
        z1 = a @ self._bf(p["w1"]).t() + p["b1"]
        h1 = self._bf(torch.relu(z1))
        z2 = h1 @ self._bf(p["w2"]).t() + p["b2"]
        h2 = self._bf(torch.relu(z2))
        z3 = h2 @ self._bf(p["w3"]).t() + p["b3"]
        h3 = self._bf(torch.relu(z3))
        logit = h3 @ p["w4"] + p["b4"] + self.wide[ids.long()].sum(1) + p["wide_b"]
        return a, h1, h2, h3, logit

    def grads(self, ids, dense, labels, denom):
        """Gradients of sum(loss)/denom; returns (loss_sum, dense grads, emb row grads [B,F,128], dlogit)."""
        c = self.cfg
        a, h1, h2, h3, logit = self.forward(ids, dense)
        loss = torch.nn.functional.binary_cross_entropy_with_logits(logit, labels, reduction="sum")
        dl = (torch.sigmoid(logit) - labels) / denom
        p = self.p
        g = {"w4": h3.t() @ dl, "b4": dl.sum().view(1), "wide_b": dl.sum().view(1)}
        dz3 = self._bf((dl[:, None] * p["w4"][None, :]) * (h3 > 0))
        g["w3"], g["b3"] = dz3.t() @ h2, dz3.sum(0)
        dz2 = self._bf((dz3 @ self._bf(p["w3"])) * (h2 > 0))
        g["w2"], g["b2"] = dz2.t() @ h1, dz2.sum(0)
        dz1 = self._bf((dz2 @ self._bf(p["w2"])) * (h1 > 0))
        g["w1"], g["b1"] = dz1.t() @ a, dz1.sum(0)
        da = self._bf(dz1 @ self._bf(p["w1"])[:, : c.slots * c.emb_dim]).view(-1, c.slots, c.emb_dim)
        return loss, g, da, dl

    def _apply(self, name, w, g, rows=None):
        c = self.cfg
        s1, s2 = self.s1[name], self.s2[name]
        if rows is not None:
            w_, s1_, s2_ = w[rows], s1[rows], s2[rows]
        else:
            w_, s1_, s2_ = w, s1, s2
        if c.optimizer == "sgd":
            w_ = w_ - c.lr * g
        elif c.optimizer == "adagrad":
            s1_ = s1_ + g * g
            w_ = w_ - c.lr * g / s1_.sqrt()
        elif c.optimizer == "adam":
            lr_t = c.lr * (1 - 0.999 ** self.t) ** 0.5 / (1 - 0.9 ** self.t)
            s1_ = 0.9 * s1_ + 0.1 * g
            s2_ = 0.999 * s2_ + 0.001 * g * g
            w_ = w_ - lr_t * s1_ / (s2_.sqrt() + 1e-8)
        elif c.optimizer == "adadelta":
            s1_ = 0.95 * s1_ + 0.05 * g * g
            u = (s2_ + 1e-8).sqrt() / (s1_ + 1e-8).sqrt() * g
            s2_ = 0.95 * s2_ + 0.05 * u * u
            w_ = w_ - c.lr * u
        elif c.optimizer == "ftrl":
            n_new = s1_ + g * g
            sigma = (n_new.sqrt() - s1_.sqrt()) / c.lr
            s2_ = s2_ + g - sigma * w_
            w_ = -s2_ * c.lr / n_new.sqrt()
            s1_ = n_new
        elif c.optimizer == "rmsprop":
            s1_ = s1_ + (g * g - s1_) * 0.1
            s2_ = c.lr * g / (s1_ + 1e-10).sqrt()
            w_ = w_ - s2_
        else:
            raise KeyError(c.optimizer)
        if rows is not None:
            w[rows], s1[rows], s2[rows] = w_, s1_, s2_
        else:
            w.copy_(w_), s1.copy_(s1_), s2.copy_(s2_)

    def sync_step(self, batches):
        """Sync-replica: gradients of all replicas' micro-batches are summed, one apply."""
        c = self.cfg
        self.t += 1
        denom = float(sum(b[0].shape[0] for b in batches))
        total = 0.0
        gsum, rows_all, da_all, dl_all = None, [], [], []
        for ids, dense, labels in batches:
            loss, g, da, dl = self.grads(ids, dense, labels, denom)
            total += float(loss)
            gsum = g if gsum is None else {k: gsum[k] + g[k] for k in g}
            rows_all.append(ids.long().reshape(-1))
            da_all.append(da.reshape(-1, c.emb_dim))
            dl_all.append(dl[:, None].expand(-1, c.slots).reshape(-1))
        for k, v in gsum.items():
            self._apply(k, self.p[k], v)
        rows = torch.cat(rows_all)
        uniq, inv = torch.unique(rows, return_inverse=True)
        ge = torch.zeros(uniq.numel(), c.emb_dim, device=rows.device).index_add_(0, inv, torch.cat(da_all))
        gw = torch.zeros(uniq.numel(), device=rows.device).index_add_(0, inv, torch.cat(dl_all))
        self._apply("emb", self.emb, ge, rows=uniq)
        self._apply("wide", self.wide, gw, rows=uniq)
        return total
