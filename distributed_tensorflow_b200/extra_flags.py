"""Flags added on top of the reference's set; every default leaves reference behaviour unchanged
(SURVEY.md 5.6)."""
from . import flags


def define() -> None:
    flags.DEFINE_string("model", "linear", "linear (the reference's toy model) or wide_deep")
    flags.DEFINE_string("device", "cpu", "cpu or cuda (wide_deep needs cuda)")
    flags.DEFINE_boolean("sync_replicas", False, "wide_deep: aggregate all replicas' gradients, apply once")
    flags.DEFINE_integer("batch_size", 8192, "wide_deep: examples per rank per step")
    flags.DEFINE_integer("embedding_rows", 10000000, "wide_deep: rows of the sharded embedding table")
    flags.DEFINE_integer("embedding_dim", 128, "wide_deep: embedding width")
    flags.DEFINE_integer("train_steps", 100, "wide_deep: steps to run")
    flags.DEFINE_float("save_model_secs", 60, "Supervisor checkpoint period (reference hard-codes 60)")
    flags.DEFINE_float("recovery_wait_secs", 30, "non-chief readiness poll period (TF default 30)")
    flags.DEFINE_float("connect_retry_secs", 60, "how long a task keeps retrying to reach the ps tasks")
    flags.DEFINE_integer("ps_timeout_ms", 30000, "socket timeout towards ps tasks; a dead ps becomes an error")
    flags.DEFINE_boolean("export_standalone", True, "export a SavedModel at the end of standalone training")
