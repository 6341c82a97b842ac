"""One task of the wide-and-deep job with the GPU trainer replaced by a CPU stand-in, so that the whole
lifecycle (TF_CONFIG roles, one Server per task, control-plane collectives, Supervisor restore-or-init,
timer checkpoints through the checkpoint service, master-only export, ps auto-stop, dead-peer exit) runs
in the CPU test-suite.  Started by tests/test_wide_deep_lifecycle.py as
``python tests/helpers/wd_lifecycle_task.py [--flag=value ...]`` with TF_CONFIG in the environment."""
import os
import sys
import time

import numpy as np
import torch

from distributed_tensorflow_b200 import extra_flags, flags as flags_lib


class FakeTrainer:
    """Shard state of one rank as CPU tensors, with the trainer methods the lifecycle calls."""
    _SLOT_NAMES = {"adagrad": ("Adagrad",)}

    def __init__(self, cfg, d):
        self.cfg, self.rank, self.world = cfg, d.rank, d.world
        self.device = torch.device("cpu")
        n = (cfg.emb_rows - d.rank + d.world - 1) // d.world
        self.emb = torch.zeros(n, cfg.emb_dim)
        self.wide = torch.zeros(n)
        self.dense = torch.zeros(64)
        self.acc = torch.zeros(n, cfg.emb_dim)
        self.applied = 0
        self.die_at = int(os.environ.get("FAKE_DIE_AT_STEP", "-1"))
        self.peer_dead_at = int(os.environ.get("FAKE_PEER_DEAD_AT_STEP", "-1"))

    def set_barrier_timeout_ms(self, ms):
        pass

    def init_params(self, seed=None):
        self.emb.fill_(1.0 + self.rank)
        self.acc.fill_(0.1)

    def _named_shards(self):
        r = self.rank
        return {"embedding/part_%d" % r: self.emb, "wide_weights/part_%d" % r: self.wide, "dense_flat/part_%d" % r: self.dense,
                "embedding/part_%d/Adagrad" % r: self.acc}

    def named_shards(self):
        return self._named_shards()

    def load_shards(self, tensors, step=None):
        for k, dst in self._named_shards().items():
            dst.copy_(torch.from_numpy(np.ascontiguousarray(tensors[k])).view_as(dst))
        if step is not None:
            self.applied = int(step) // self.world

    def step(self, ids, dense, labels):
        self.applied += 1
        self.emb += 0.5
        self.wide += 1.0
        if self.die_at >= 0 and self.applied >= self.die_at:
            os._exit(9)                                   # this rank dies mid-run
        time.sleep(float(os.environ.get("FAKE_STEP_SECS", "0.002")))

    def join(self):
        pass

    def check_error(self):
        if self.peer_dead_at >= 0 and self.applied >= self.peer_dead_at:
            raise ConnectionError("rank %d: device barrier timed out waiting for rank 1 (peer dead or stalled)" % self.rank)

    def last_loss_sum(self):
        return 0.5 * self.cfg.batch

    def export_config(self):
        c = self.cfg
        return dict(emb_rows=c.emb_rows, emb_dim=c.emb_dim, slots=c.slots, dense_in=c.dense_in, dense_pad=c.dense_pad,
                    h1=c.h1, h2=c.h2, h3=c.h3, world=self.world)

    def close(self):
        pass


def main():
    from distributed_tensorflow_b200.trainer_lib import run
    FLAGS = flags_lib.FLAGS
    flags_lib.DEFINE_integer("max_epochs", 4, "")
    flags_lib.DEFINE_string("checkpoint_path", "./checkpoint/", "")
    flags_lib.DEFINE_string("output_path", "./tensorboard/", "")
    flags_lib.DEFINE_integer("checkpoint_period", 1, "")
    flags_lib.DEFINE_string("model_path", "./model/", "")
    flags_lib.DEFINE_float("learning_rate", 0.01, "")
    flags_lib.DEFINE_string("optimizer", "adagrad", "")
    flags_lib.DEFINE_string("saved_model_path", "./saved_model/", "")
    flags_lib.DEFINE_integer("model_version", 1, "")
    extra_flags.define()
    import distributed_tensorflow_b200.models.wide_deep_trainer as wdt
    real = wdt.run_wide_deep
    wdt.run_wide_deep = lambda flags, auto_stop=False: real(flags, auto_stop=auto_stop, trainer_factory=FakeTrainer)
    run(FLAGS, auto_stop=os.environ.get("FAKE_AUTO_STOP", "1") == "1")


if __name__ == "__main__":
    sys.exit(main())
