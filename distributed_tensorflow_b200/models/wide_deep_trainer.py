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
import time
from typing import Dict, Optional

import numpy as np
import torch

from ..checkpoint import bundle as _bundle
from ..checkpoint.saver import Saver, checkpoint_exists, latest_checkpoint, update_checkpoint_state
from ..summary import FileWriter, Summary
from .wide_deep import WideDeepConfig, WideDeepTrainer


class _Dist:
    """Thin wrapper so that single-process runs need no process group."""

    def __init__(self):
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.pg = None
        if self.world > 1:
            import torch.distributed as dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            if not dist.is_initialized():
                dist.init_process_group("gloo", rank=self.rank, world_size=self.world)
            self.pg = dist

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


def make_trainer(cfg: WideDeepConfig, d: _Dist) -> WideDeepTrainer:
    torch.cuda.set_device(d.local_rank)
    tr = WideDeepTrainer(cfg, rank=d.rank, world=d.world, device=torch.device("cuda", d.local_rank))
    if d.world > 1:
        tr.attach_peers(d.pg)
    return tr


class WideDeepCheckpointer:
    """Collective save / restore of the sharded model in TF checkpoint-V2 layout."""

    def __init__(self, trainer: WideDeepTrainer, d: _Dist, logdir: str, max_to_keep: int = 5):
        self.tr, self.d, self.logdir = trainer, d, logdir
        self._saver = Saver(max_to_keep=max_to_keep)      # bookkeeping of retained checkpoints (chief)
        self._side = torch.cuda.Stream(device=trainer.device)
        self._pinned: Dict[str, torch.Tensor] = {}
        if d.rank == 0:
            os.makedirs(logdir, exist_ok=True)
            self._saver.recover_last_checkpoints(logdir)

    def _snapshot_to_host(self) -> Dict[str, np.ndarray]:
        shards = self.tr.named_shards()
        self._side.wait_stream(torch.cuda.current_stream(self.tr.device))
        with torch.cuda.stream(self._side):
            for k, t in shards.items():
                if k not in self._pinned:
                    self._pinned[k] = torch.empty(t.shape, dtype=t.dtype).pin_memory()
                self._pinned[k].copy_(t, non_blocking=True)
        self._side.synchronize()
        return {k: v.numpy() for k, v in self._pinned.items()}

    def save(self, step: int) -> str:
        """Every rank must call this at the same step."""
        prefix = os.path.join(self.logdir, "model.ckpt-%d" % step)
        part = "%s_temp/part-%05d" % (prefix, self.d.rank)
        if self.d.rank == 0:
            os.makedirs(os.path.dirname(part), exist_ok=True)
        self.d.barrier()
        host = self._snapshot_to_host()
        if self.d.rank == 0:
            host["global_step"] = np.array(step, dtype=np.int64)
        _bundle.write_bundle(part, host)
        self.d.barrier()
        if self.d.rank == 0:
            _bundle.merge_bundles(["%s_temp/part-%05d" % (prefix, r) for r in range(self.d.world)], prefix)
            try:
                os.rmdir(prefix + "_temp")
            except OSError:
                pass
            s = self._saver
            s._last_checkpoints.append(prefix)
            while len(s._last_checkpoints) > s.max_to_keep:
                s._delete(s._last_checkpoints.pop(0))
            update_checkpoint_state(self.logdir, prefix, s._last_checkpoints)
        self.d.barrier()
        return prefix

    def restore_latest(self) -> Optional[int]:
        """Chief semantics of Supervisor.prepare_session: restore if a checkpoint exists, else None."""
        path = self.d.broadcast_object(latest_checkpoint(self.logdir) if self.d.rank == 0 else None)
        if path is None:
            return None
        with _bundle.BundleReader(path) as r:
            if r.num_shards != self.d.world:
                raise RuntimeError("checkpoint %s was written by %d ranks, this job has %d" % (path, r.num_shards, self.d.world))
            names = [k for k in r.keys() if ("/part_%d" % self.d.rank) in k + "/"]
            tensors = {k: r.get_tensor(k) for k in names}
            step = int(r.get_tensor("global_step"))
        self.tr.load_shards(tensors, step=step)
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


def run_wide_deep(flags) -> None:
    if flags.device != "cuda" or not torch.cuda.is_available():
        raise SystemExit("--model=wide_deep needs --device=cuda and a visible GPU")
    start = datetime.datetime.now()
    d = _Dist()
    from .wide_deep import OPT_IDS
    opt = flags.optimizer if flags.optimizer in OPT_IDS else None
    print("Use the optimizer: {}".format(flags.optimizer))
    if opt is None:
        print("Unknow optimizer: {}, exit now".format(flags.optimizer))
        raise SystemExit(1)
    cfg = WideDeepConfig(emb_rows=flags.embedding_rows, emb_dim=flags.embedding_dim, batch=flags.batch_size,
                         optimizer=opt, lr=flags.learning_rate, sync=bool(flags.sync_replicas))
    tr = make_trainer(cfg, d)
    is_chief = d.rank == 0
    logdir = getattr(flags, "checkpoint_path", None) or getattr(flags, "checkpoint_dir")
    out_dir = getattr(flags, "output_path", None) or getattr(flags, "tensorboard_dir")
    ckpt = WideDeepCheckpointer(tr, d, logdir)
    restored = ckpt.restore_latest()
    if restored is None:
        tr.init_params()
        torch.cuda.synchronize()
        d.barrier()
    step0 = restored or 0
    print("Save tensorboard files into: {}".format(out_dir))
    writer = FileWriter(out_dir, filename_suffix=".rank%d" % d.rank)
    print("Run training with epoch number: {}".format(flags.max_epochs))
    gen = torch.Generator().manual_seed(100 + d.rank + 7919 * step0)
    steps_per_epoch = max(1, flags.train_steps // max(1, flags.max_epochs))
    last_save = time.time()
    ckpt.save(step0)                                       # Supervisor: first save right away
    step = step0
    torch.cuda.set_stream(torch.cuda.Stream(device=tr.device, priority=-1))   # compute outranks the ps-side apply
    for epoch in range(flags.max_epochs):
        for _ in range(steps_per_epoch):
            batch = synthetic_batch(cfg, gen, torch.device("cpu"))
            if d.world > 1:
                # copy-engine transfers are kept away from the device-barrier kernels of the overlapped apply
                # (they stalled intermittently when mixed inside CUDA graphs); the host-side
                # batch generation above is far longer than the apply, so this wait is normally already over
                tr.join()
            tr.step(*(t.to(tr.device) for t in batch))
            step += 1
        if epoch % flags.checkpoint_period == 0:
            loss = float(tr.last_loss_sum()) / cfg.batch
            print("Epoch: {}, loss: {}".format(epoch, loss))
            if is_chief:
                writer.add_summary(Summary([("loss", loss)]), step * d.world)
        want = d.broadcast_object(time.time() - last_save >= flags.save_model_secs)
        if want:
            torch.cuda.synchronize()
            ckpt.save(step)
            last_save = time.time()
    writer.close()
    torch.cuda.synchronize()
    print("[{}] End of distributed training.".format(datetime.datetime.now() - start))
    try:
        export_wide_deep(tr, d, flags.saved_model_path, flags.model_version)
    except Exception as e:  # noqa: BLE001  (reference logs export failures and carries on)
        print(e)
    tr.close()
