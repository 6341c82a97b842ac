"""``python -m trainer.task --model=wide_deep --device=cuda``: the reference's training lifecycle
(/root/reference/trainer/task.py:215-258 -- Supervisor-managed session, chief duties, periodic
checkpoints, master-only export) driving the B200 wide-and-deep model instead of the toy graph.

One process per GPU (``torchrun`` sets RANK / LOCAL_RANK / WORLD_SIZE); rank 0 is the chief
("master").  Checkpointing is the GPU flavour of ``Saver(sharded=True)``: every rank copies its ps
shard device -> pinned host on a side stream and writes its own bundle shard, the chief merges the
indices into one TF checkpoint-V2 (``model.ckpt-<step>.data-0000r-of-0000R``) -- SURVEY.md row K9.
"""
from __future__ import annotations

import datetime
import os
import threading
import time
from typing import Dict, Optional

import numpy as np
import torch

from ..checkpoint import bundle as _bundle
from ..checkpoint.saver import Saver, checkpoint_exists, latest_checkpoint, update_checkpoint_state
from ..summary import FileWriter, Summary
from .wide_deep import WideDeepConfig, WideDeepTrainer


class _Dist:
    """Rank, world and the host control plane of this process.

    Two ways in, one interface out (``barrier`` / ``all_gather_object`` / ``broadcast_object`` and
    ``pg`` with torch.distributed's call shapes):

    * ``TF_CONFIG`` (the reference's way, /root/reference/trainer/task.py:110-127): cluster and role come
      from the environment, ``master`` is rank 0 and the chief, ``worker`` i is rank 1 + i, every task
      starts its own :class:`Server` on the address the cluster lists for it, and all host-side
      coordination goes through the C++ transport (``parallel/control.py``); no torchrun, no gloo.
    * ``torchrun`` (RANK / WORLD_SIZE set): gloo for the collectives; rank 0 additionally hosts a control
      store so that the checkpoint service below works the same way.
    """

    def __init__(self, timeout_ms: int = 30000, connect_retry_secs: float = 60.0):
        from ..cluster import ClusterSpec, Server, parse_tf_config
        from ..parallel.control import TransportGroup, rank_layout
        self.pg = None
        self.server = None
        self.cluster = None
        self.task_type, self.task_index = "master", 0
        self.timeout_ms = int(timeout_ms)
        tf_cfg = parse_tf_config() if "RANK" not in os.environ else None
        if tf_cfg is not None:
            self.cluster, self.task_type, self.task_index = tf_cfg.cluster, tf_cfg.task_type, tf_cfg.task_index
            rank, self.world, self.store_addr = rank_layout(tf_cfg.cluster, tf_cfg.task_type, tf_cfg.task_index)
            if rank is None:
                raise ValueError("task type %r does not train (only master / worker do)" % tf_cfg.task_type)
            self.rank = rank
            self.server = Server(tf_cfg.cluster, job_name=tf_cfg.task_type, task_index=tf_cfg.task_index)
            ndev = max(1, torch.cuda.device_count())
            self.local_rank = self.rank % ndev       # CUDA_VISIBLE_DEVICES per task (one device each) also works
            self.pg = TransportGroup(self.store_addr, self.rank, self.world, timeout_s=max(1.0, timeout_ms / 1000.0),
                                     connect_retry_secs=connect_retry_secs)
        else:
            self.world = int(os.environ.get("WORLD_SIZE", "1"))
            self.rank = int(os.environ.get("RANK", "0"))
            self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
            if self.world > 1:
                import torch.distributed as dist
                os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
                if not dist.is_initialized():
                    dist.init_process_group("gloo", rank=self.rank, world_size=self.world)
                self.pg = dist
            if self.rank == 0:
                self.server = Server(ClusterSpec({"master": ["127.0.0.1:0"]}), "master", 0)
            self.store_addr = self.broadcast_object("127.0.0.1:%d" % self.server.port if self.rank == 0 else None)
        # names on the control store are scoped by a per-incarnation nonce chosen by the chief
        self.nonce = self.broadcast_object("%x" % int(time.time() * 1000) if self.rank == 0 else None)

    @property
    def is_chief(self) -> bool:
        return self.rank == 0

    def control_client(self):
        """A fresh connection to the control store (one per thread that blocks on queues)."""
        from ..session import PSClient
        return PSClient(self.store_addr, timeout_ms=self.timeout_ms, connect_retry_secs=10.0)

    def name(self, what: str) -> str:
        return "wd/%s/%s" % (self.nonce, what)

    def barrier(self):
        if self.pg:
            self.pg.barrier()

    def all_gather_object(self, obj):
        if not self.pg:
            return [obj]
        out = [None] * self.world
        self.pg.all_gather_object(out, obj)
        return out

    def broadcast_object(self, obj, src=0):
        if not self.pg:
            return obj
        box = [obj]
        self.pg.broadcast_object_list(box, src=src)
        return box[0]

    def close(self) -> None:
        if self.pg is not None and hasattr(self.pg, "close"):
            self.pg.close()
        if self.server is not None:
            self.server.stop()
            self.server = None


def make_trainer(cfg: WideDeepConfig, d: _Dist) -> WideDeepTrainer:
    torch.cuda.set_device(d.local_rank)
    tr = WideDeepTrainer(cfg, rank=d.rank, world=d.world, device=torch.device("cuda", d.local_rank))
    if d.world > 1:
        tr.attach_peers(d.pg)
    return tr


class WideDeepCheckpointer:
    """Save / restore of the sharded model in TF checkpoint-V2 layout: every rank writes
    the bundle shard of its own ps shard, the chief merges the indices.

    Two ways to save.  :meth:`save` is the plain collective (every rank calls it at the same step).
    The *service* is what the Supervisor's timer thread uses (reference trainer/task.py:215-223,
    ``save_model_secs=60``): :meth:`start_service` starts one thread per non-chief rank that waits on a
    queue of the control store; the chief's :meth:`request_save` posts one request per rank, writes its
    own shard, collects the acknowledgements and merges.  Shards are copied device -> pinned host on a
    side stream without synchronising the training streams, so a checkpoint taken during Hogwild
    training is not an atomic cut -- exactly like TF's (the ps snapshots while workers keep applying).
    """

    def __init__(self, trainer: WideDeepTrainer, d: _Dist, logdir: str, max_to_keep: int = 5):
        self.tr, self.d, self.logdir = trainer, d, logdir
        self._saver = Saver(max_to_keep=max_to_keep)      # bookkeeping of retained checkpoints (chief)
        self._cuda = getattr(trainer, "device", torch.device("cpu")).type == "cuda"
        self._side = torch.cuda.Stream(device=trainer.device) if self._cuda else None
        self._pinned: Dict[str, torch.Tensor] = {}
        self._lock = threading.Lock()                       # one snapshot at a time per rank
        self._svc: Optional[threading.Thread] = None
        self._svc_stop = threading.Event()
        self._svc_error: Optional[BaseException] = None
        self._ctl = None
        if d.rank == 0:
            os.makedirs(logdir, exist_ok=True)
            self._saver.recover_last_checkpoints(logdir)

    @property
    def last_checkpoints(self):
        return list(self._saver._last_checkpoints)

    def _snapshot_to_host(self, wait_for_training: bool = True) -> Dict[str, np.ndarray]:
        """Host views of this rank's shard (pinned staging buffers, reused: callers hold ``_lock`` until
        they are done with the arrays)."""
        if not self._cuda:
            return {k: t.detach().clone().numpy() for k, t in self.tr.named_shards().items()}
        if wait_for_training:
            shards = self.tr.named_shards()                 # current stream waits for the pending applies
            self._side.wait_stream(torch.cuda.current_stream(self.tr.device))
        else:
            shards = self.tr._named_shards()                # service thread: no dependency on the training streams
        with torch.cuda.device(self.tr.device), torch.cuda.stream(self._side):
            for k, t in shards.items():
                if k not in self._pinned:
                    self._pinned[k] = torch.empty(t.shape, dtype=t.dtype).pin_memory()
                self._pinned[k].copy_(t, non_blocking=True)
        self._side.synchronize()
        return {k: v.numpy() for k, v in self._pinned.items()}

    def _part(self, prefix: str, rank: int) -> str:
        return "%s_temp/part-%05d" % (prefix, rank)

    def _write_part(self, prefix: str, step: int, wait_for_training: bool) -> None:
        with self._lock:                                    # one snapshot + write at a time per rank
            host = self._snapshot_to_host(wait_for_training)
            if self.d.rank == 0:
                host["global_step"] = np.array(step, dtype=np.int64)
            os.makedirs(os.path.dirname(self._part(prefix, self.d.rank)), exist_ok=True)
            _bundle.write_bundle(self._part(prefix, self.d.rank), host)

    def _merge(self, prefix: str) -> None:
        _bundle.merge_bundles([self._part(prefix, r) for r in range(self.d.world)], prefix)
        try:
            os.rmdir(prefix + "_temp")
        except OSError:
            pass
        s = self._saver
        try:
            # <prefix>.meta like tf.train.Saver: graph (placeholders, one VariableV2 per saved tensor, saver sub-graph),
            # saver_def, variable collections -- shapes only, no tensor is loaded for it
            from ..checkpoint.saver import _write_meta
            with _bundle.BundleReader(prefix) as r:
                dtypes = {k: _bundle.TF_TO_DTYPE.get(r.entry(k)["dtype"], np.dtype(np.float32)) for k in r.keys()}
                shapes = {k: np.broadcast_to(np.zeros((), dtypes[k]), shp) for k, shp in r.get_variable_to_shape_map().items()}
            _write_meta(prefix + ".meta", sorted(shapes), True, s.max_to_keep, WideDeepServingSpec(self.tr.export_config()), shapes)
        except Exception as e:  # noqa: BLE001  (the .meta file is optional; restore never reads it)
            print("checkpoint %s: no .meta written (%s)" % (prefix, e))
        s._last_checkpoints.append(prefix)
        while len(s._last_checkpoints) > s.max_to_keep:
            s._delete(s._last_checkpoints.pop(0))
        update_checkpoint_state(self.logdir, prefix, s._last_checkpoints)

    def save(self, step: int) -> str:
        """Collective: every rank must call this at the same step."""
        prefix = os.path.join(self.logdir, "model.ckpt-%d" % step)
        if self.d.rank == 0:
            os.makedirs(prefix + "_temp", exist_ok=True)
        self.d.barrier()
        self._write_part(prefix, step, wait_for_training=True)
        self.d.barrier()
        if self.d.rank == 0:
            self._merge(prefix)
        self.d.barrier()
        return prefix

    # ------------------------------------------------------------------ checkpoint service
    def _req_queue(self, rank: int) -> str:
        return self.d.name("ckpt_req/%d" % rank)

    def start_service(self) -> None:
        """Non-chief ranks: serve the chief's save requests from a background thread until
        :meth:`stop_service`.  (The chief needs no thread; its Supervisor timer calls request_save.)"""
        if self.d.rank == 0 or self._svc is not None:
            return
        self._ctl = self.d.control_client()
        self._svc_stop.clear()

        def loop():
            try:
                while not self._svc_stop.is_set():
                    step = self._ctl.dequeue(self._req_queue(self.d.rank), timeout_ms=200)
                    if step is None:
                        continue
                    prefix = os.path.join(self.logdir, "model.ckpt-%d" % step)
                    self._write_part(prefix, step, wait_for_training=False)
                    self._ctl.enqueue(self.d.name("ckpt_done"), self.d.rank)
            except BaseException as e:  # noqa: BLE001  (surfaced by stop_service)
                self._svc_error = e

        self._svc = threading.Thread(target=loop, name="wd-checkpoint-service", daemon=True)
        self._svc.start()

    def stop_service(self) -> None:
        if self._svc is not None:
            self._svc_stop.set()
            self._svc.join(30)
            self._svc = None
        if self._ctl is not None:
            self._ctl.close()
            self._ctl = None
        if self._svc_error is not None:
            e, self._svc_error = self._svc_error, None
            raise e

    def request_save(self, step: int, timeout_s: float = 600.0, abort=None) -> Optional[str]:
        """Chief only, any thread: checkpoint every rank's shard as of "about now" without stopping training.
        ``abort()`` turning true (the Supervisor is stopping) gives the save up and returns None."""
        assert self.d.rank == 0
        prefix = os.path.join(self.logdir, "model.ckpt-%d" % step)
        os.makedirs(prefix + "_temp", exist_ok=True)
        ctl = self.d.control_client()
        try:
            for r in range(1, self.d.world):
                ctl.enqueue(self._req_queue(r), int(step))
            self._write_part(prefix, step, wait_for_training=False)
            deadline = time.time() + timeout_s
            seen = set()
            while len(seen) < self.d.world - 1:
                if abort is not None and abort():
                    return None
                if time.time() > deadline:
                    raise ConnectionError("checkpoint %s: ranks %s did not write their shard within %.0f s"
                                          % (prefix, sorted(set(range(1, self.d.world)) - seen), timeout_s))
                r = ctl.dequeue(self.d.name("ckpt_done"), timeout_ms=500)
                if r is not None:
                    seen.add(int(r))
        finally:
            ctl.close()
        self._merge(prefix)
        return prefix

    def restore_from(self, path: str) -> int:
        """This rank's shard (and the step) out of a merged checkpoint."""
        with _bundle.BundleReader(path) as r:
            if r.num_shards != self.d.world:
                raise RuntimeError("checkpoint %s was written by %d ranks, this job has %d" % (path, r.num_shards, self.d.world))
            names = [k for k in r.keys() if ("/part_%d" % self.d.rank) in k + "/"]
            tensors = {k: r.get_tensor(k) for k in names}
            step = int(r.get_tensor("global_step"))
        self.tr.load_shards(tensors, step=step)
        return step

    def restore_latest(self) -> Optional[int]:
        """Collective.  Chief semantics of Supervisor.prepare_session: restore if a checkpoint exists, else None."""
        path = self.d.broadcast_object(latest_checkpoint(self.logdir) if self.d.rank == 0 else None)
        if path is None:
            return None
        step = self.restore_from(path)
        self.d.barrier()
        return step


def export_wide_deep(tr: WideDeepTrainer, d: _Dist, saved_model_path: str, model_version: int) -> Optional[str]:
    """SavedModel-compatible export; all ranks contribute their shards, the chief writes the files."""
    from ..export.saved_model import SavedModelBuilder, SERVING, DEFAULT_SERVING_SIGNATURE_DEF_KEY, build_signature_def
    export_path = os.path.join(saved_model_path, str(model_version))
    host = {k: v.detach().cpu().numpy() for k, v in tr.named_shards().items() if k.count("/") == 1}   # no slots
    shards = d.all_gather_object(host) if d.world <= 1 else None
    if d.world > 1:
        tmp = export_path + "_shards"
        if d.rank == 0:
            os.makedirs(tmp, exist_ok=True)
        d.barrier()
        _bundle.write_bundle(os.path.join(tmp, "part-%05d" % d.rank), host)
        d.barrier()
    if d.rank != 0:
        d.barrier()
        return None

    class _Store:
        model = WideDeepServingSpec(tr.export_config())

        def snapshot(self_inner):
            if d.world == 1:
                return dict(shards[0])
            out = {}
            for r in range(d.world):
                out.update(_bundle.read_bundle(os.path.join(export_path + "_shards", "part-%05d" % r)))
            return out

    builder = SavedModelBuilder(export_path)
    store = _Store()
    builder.add_meta_graph_and_variables(store, [SERVING], signature_def_map={
        DEFAULT_SERVING_SIGNATURE_DEF_KEY: build_signature_def(store.model)}, model=store.model)
    builder.save()
    if d.world > 1:
        import shutil
        shutil.rmtree(export_path + "_shards", ignore_errors=True)
        d.barrier()
    return export_path


class WideDeepServingSpec:
    """Signature of the exported wide-and-deep model: ``keys`` pass through next to ``prediction``;
    ``features`` carries, per example, the F sparse ids followed by the dense features."""
    name = "wide_deep"

    def __init__(self, config: dict):
        self._config = dict(config)
        width = config["slots"] + config["dense_in"]
        self.signature_inputs = {"keys": ("keys:0", np.int32, (-1, 1)), "features": ("features:0", np.float32, (-1, width))}
        self.signature_outputs = {"keys": ("keys_out:0", np.int32, (-1, 1)), "prediction": ("prediction:0", np.float32, (-1, 1))}
        self.legacy_inputs = {"keys": "keys", "X": "features"}
        self.legacy_outputs = {"keys": "keys", "predict": "prediction"}

    def export_config(self) -> dict:
        return self._config


def synthetic_batch(cfg: WideDeepConfig, gen: torch.Generator, device):
    ids = torch.randint(0, cfg.emb_rows, (cfg.batch, cfg.slots), generator=gen, dtype=torch.int32)
    dense = torch.randn(cfg.batch, cfg.dense_in, generator=gen)
    # a learnable target: depends on the dense features and on the parity of the first id
    logit = dense[:, 0] - 0.5 * dense[:, 1] + (ids[:, 0] % 2).float() - 0.5
    labels = (torch.rand(cfg.batch, generator=gen) < torch.sigmoid(logit)).float()
    return ids.to(device, non_blocking=True), dense.to(device, non_blocking=True), labels.to(device, non_blocking=True)


class WideDeepSession:
    """What :class:`Supervisor` manages for the GPU job: readiness, init / restore, the step counter.

    The chief decides "restore checkpoint X" or "initialise" (Supervisor.prepare_or_wait_for_session, the
    restore-or-init of reference trainer/task.py:215-226) and publishes the decision on the control
    store; every other rank polls ``is_ready()`` like a TF non-chief, and when the decision is there
    loads its own shard of X (or initialises its shard) and joins the barrier that ends start-up.
    """

    def __init__(self, tr: WideDeepTrainer, d: _Dist, ckpt: WideDeepCheckpointer):
        self.tr, self.d, self.ckpt = tr, d, ckpt
        self.model = WideDeepServingSpec(tr.export_config())
        self.steps = 0                     # steps of this rank; restored from the checkpoint
        self._ready = False
        self._ctl = d.control_client()

    # -- chief ---------------------------------------------------------------------------------
    def _publish(self, decision) -> None:
        from ..parallel.control import put_object
        put_object(self._ctl, self.d.name("start"), decision)

    def _finish_startup(self) -> None:
        if self.tr.device.type == "cuda":
            torch.cuda.synchronize(self.tr.device)
        self.d.barrier()                   # every shard is in place before anyone pulls from a peer
        self._ready = True

    def run_init(self) -> None:
        self._publish({"restore": None})
        self.tr.init_params()
        self._finish_startup()

    def restore(self, path: str) -> None:
        self._publish({"restore": path})
        self.steps = self.ckpt.restore_from(path) // max(1, self.d.world)

    def end_restore(self) -> None:
        self._finish_startup()

    # -- every rank ------------------------------------------------------------------------------
    def is_ready(self) -> bool:
        if self._ready or self.d.rank == 0:
            return self._ready
        from ..parallel.control import get_object
        decision = get_object(self._ctl, self.d.name("start"))
        if decision is None:
            return False
        if decision["restore"] is None:
            self.tr.init_params()
        else:
            self.steps = self.ckpt.restore_from(decision["restore"]) // max(1, self.d.world)
        self._finish_startup()
        return True

    def global_step(self) -> int:
        """Like the reference, every worker's step bumps the shared counter (task.py:138)."""
        return self.steps * self.d.world

    def snapshot(self) -> Dict[str, np.ndarray]:
        """Shapes only (zero-stride views), for graph.pbtxt; checkpoints go through the checkpointer."""
        return {k: np.broadcast_to(np.float32(0), tuple(t.shape)) for k, t in self.tr._named_shards().items()}

    def close(self) -> None:
        self._ctl.close()


class WideDeepSaver:
    """The ``saver`` the Supervisor's timer thread calls (``saver.save(sess, save_path, global_step)``):
    triggers the checkpoint service, so periodic checkpoints never synchronise the training streams."""

    def __init__(self, ckpt: WideDeepCheckpointer, timeout_s: float = 600.0):
        self.ckpt, self.timeout_s = ckpt, timeout_s
        self.abort = None                   # set to Supervisor.should_stop once the supervisor exists

    @property
    def last_checkpoints(self):
        return self.ckpt.last_checkpoints

    def recover_last_checkpoints(self, logdir: str) -> None:
        pass                                # done by the checkpointer on the chief

    def restore(self, sess: WideDeepSession, path: str):
        sess.restore(path)
        return [path]

    def save(self, sess: WideDeepSession, save_path: str, global_step=None) -> str:
        return self.ckpt.request_save(int(global_step if global_step is not None else sess.global_step()),
                                      timeout_s=self.timeout_s, abort=self.abort)


def run_wide_deep(flags, auto_stop: bool = False, trainer_factory=None) -> None:
    """The reference's distributed main() (trainer/task.py:109-261) for the B200 wide-and-deep job:
    roles from TF_CONFIG (or torchrun), one Server per task, Supervisor-managed session with the 60 s
    checkpoint timer and chief duties, master-only export, optional ps auto-stop.

    ``trainer_factory(cfg, d)`` replaces the GPU trainer in the CPU tests of this lifecycle
    (tests/test_wide_deep_lifecycle.py); product runs always take the CUDA path."""
    from ..supervisor import Supervisor
    from .wide_deep import OPT_IDS
    if trainer_factory is None and (flags.device != "cuda" or not torch.cuda.is_available()):
        raise SystemExit("--model=wide_deep needs --device=cuda and a visible GPU")
    start = datetime.datetime.now()
    timeout_ms = int(getattr(flags, "ps_timeout_ms", 30000))
    d = _Dist(timeout_ms=timeout_ms, connect_retry_secs=float(getattr(flags, "connect_retry_secs", 60)))
    opt = flags.optimizer if flags.optimizer in OPT_IDS else None
    print("Use the optimizer: {}".format(flags.optimizer))
    if opt is None:
        print("Unknow optimizer: {}, exit now".format(flags.optimizer))
        raise SystemExit(1)
    cfg = WideDeepConfig(emb_rows=flags.embedding_rows, emb_dim=flags.embedding_dim, batch=flags.batch_size,
                         optimizer=opt, lr=flags.learning_rate, sync=bool(flags.sync_replicas))
    tr = (trainer_factory or make_trainer)(cfg, d)
    on_gpu = tr.device.type == "cuda"
    tr.set_barrier_timeout_ms(timeout_ms)
    is_chief = d.is_chief
    logdir = getattr(flags, "checkpoint_path", None) or getattr(flags, "checkpoint_dir")
    out_dir = getattr(flags, "output_path", None) or getattr(flags, "tensorboard_dir")
    ckpt = WideDeepCheckpointer(tr, d, logdir)
    session = WideDeepSession(tr, d, ckpt)
    saver = WideDeepSaver(ckpt, timeout_s=max(5.0, 4.0 * timeout_ms / 1000.0))
    sv = Supervisor(is_chief=is_chief, logdir=logdir, summary_op=None, saver=saver,
                    global_step="global_step", save_model_secs=float(getattr(flags, "save_model_secs", 60)),
                    recovery_wait_secs=min(1.0, float(getattr(flags, "recovery_wait_secs", 30))),
                    session_factory=lambda target: session)
    # a checkpoint in flight when training ends normally is completed (the other ranks' services stay up until the
    # chief has stopped); it is only given up when the job is stopping because of an error
    saver.abort = sv.coord.has_exception
    failed = None
    try:
        ckpt.start_service()               # non-chief ranks answer the chief's checkpoint requests from now on
        with sv.managed_session(d.server.target if d.server is not None else "") as sess:
            print("Save tensorboard files into: {}".format(out_dir))
            writer = FileWriter(out_dir, filename_suffix=".rank%d" % d.rank)
            print("Run training with epoch number: {}".format(flags.max_epochs))
            gen = torch.Generator().manual_seed(100 + d.rank + 7919 * sess.steps)
            steps_per_epoch = max(1, flags.train_steps // max(1, flags.max_epochs))
            sess_local_steps = [0]             # steps taken by this process (sess.steps also counts restored ones)
            if on_gpu:
                torch.cuda.set_stream(torch.cuda.Stream(device=tr.device, priority=-1))   # compute outranks the ps-side apply
            # Async mode on the GPU: the input pipeline runs one batch ahead, and every step announces the next
            # batch so that its embedding rows are gathered while the step computes (WideDeepTrainer.step).
            prefetch = on_gpu and not cfg.sync and os.environ.get("DTF_PREFETCH", "1") != "0"

            def draw():
                batch = synthetic_batch(cfg, gen, torch.device("cpu"))
                if d.world > 1:
                    # copy-engine transfers are kept away from the device-barrier kernels of the overlapped apply
                    # (they stalled intermittently when mixed inside CUDA graphs); the host-side
                    # batch generation above is far longer than the apply, so this wait is normally already over
                    tr.join()
                return tuple(t.to(tr.device) for t in batch)

            total = flags.max_epochs * steps_per_epoch
            cur = draw()
            for epoch in range(flags.max_epochs):
                for _ in range(steps_per_epoch):
                    nxt = draw() if (prefetch and sess_local_steps[0] + 1 < total) else None
                    if nxt is not None:
                        tr.step(*cur, next_batch=nxt[:2])
                    else:
                        tr.step(*cur)
                    sess_local_steps[0] += 1
                    cur = nxt if nxt is not None else (draw() if sess_local_steps[0] < total else None)
                    sess.steps += 1
                tr.check_error()           # a peer that stopped answering the device barriers: ConnectionError
                if sv.should_stop():
                    break
                if epoch % flags.checkpoint_period == 0:
                    loss = float(tr.last_loss_sum()) / cfg.batch
                    print("Epoch: {}, loss: {}".format(epoch, loss))
                    if is_chief:
                        writer.add_summary(Summary([("loss", loss)]), sess.global_step())
            writer.close()
            if on_gpu:
                torch.cuda.synchronize()
            tr.check_error()
            print("[{}] End of distributed training.".format(datetime.datetime.now() - start))
            d.barrier()                    # every rank is done stepping before the chief stops its services
        d.barrier()                        # the chief's last timer checkpoint (if any) has been served
        ckpt.stop_service()
        try:
            export_wide_deep(tr, d, flags.saved_model_path, flags.model_version)
        except ConnectionError:
            raise
        except Exception as e:  # noqa: BLE001  (reference logs export failures and carries on)
            print(e)
        if auto_stop and d.cluster is not None and "ps" in d.cluster.jobs:
            # auto_stop_ps/task.py:270-272: one token per training task on every ps task's queue
            from ..session import PSClient
            from ..trainer_lib import done_queue_name
            for i in range(d.cluster.num_tasks("ps")):
                c = PSClient(d.cluster.task_address("ps", i), timeout_ms=timeout_ms, connect_retry_secs=10.0)
                c.enqueue(done_queue_name(i), 1)
                c.close()
    except ConnectionError as e:
        # a lost peer is fatal for the GPU job (its shard of every variable is gone): report and exit non-zero
        print(e)
        failed = e
    except Exception as e:  # noqa: BLE001  (reference task.py:260-261 swallows everything else)
        print(e)
    finally:
        try:
            ckpt.stop_service()
        except Exception as e:  # noqa: BLE001
            print(e)
        if failed is None:
            tr.close()
        d.close()
    if failed is not None:
        os._exit(1)                        # peers' mappings may be gone; skip CUDA teardown that could block
