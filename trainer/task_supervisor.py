#!/usr/bin/env python
"""``python -m trainer.task_supervisor`` -- the older variant of the trainer (interface of
/root/reference/trainer/task_supervisor.py, which is Python-2 only upstream): different flag names
(:21-33), ``Saver(sharded=True)`` (:145), legacy session_bundle export with aliases ``X`` /
``predict`` in both standalone (:107) and master (:185-200) mode.
"""
from distributed_tensorflow_b200 import flags, trainer_lib
from distributed_tensorflow_b200.extra_flags import define as _define_extra

flags.DEFINE_integer("max_epochs", 100, "Number of steps to run trainer.")
flags.DEFINE_string("checkpoint_dir", "./checkpoint/", "indicates the checkpoint dirctory")
flags.DEFINE_string("tensorboard_dir", "./tensorboard/", "indicates training output")
flags.DEFINE_integer("checkpoint_period", 1, "Number of epochs to save checkpoint.")
flags.DEFINE_integer("export_version", 1, "Version number of the model.")
flags.DEFINE_string("export_dir", "/tmp/linear_model/", "The model directory")
flags.DEFINE_float("learning_rate", 0.01, "Initial learning rate.")
flags.DEFINE_string("optimizer", "sgd", "Optimizer to train")
_define_extra()
FLAGS = flags.FLAGS


def main():
    trainer_lib.run(FLAGS, legacy=True, auto_stop=False)


if __name__ == "__main__":
    main()
