#!/usr/bin/env python
"""``python -m trainer.task`` -- main entry point (interface of /root/reference/trainer/task.py).

Standalone when ``TF_CONFIG`` is empty; otherwise the role (``ps`` / ``worker`` / ``master``) comes
from ``TF_CONFIG``.  The nine reference flags keep their names and defaults
(/root/reference/trainer/task.py:17-31); the extra flags below default to reference behaviour.
"""
from distributed_tensorflow_b200 import flags, trainer_lib

flags.DEFINE_integer("max_epochs", 10, "Number of steps to run trainer.")
flags.DEFINE_string("checkpoint_path", "./checkpoint/", "The checkpoint directory")
flags.DEFINE_string("output_path", "./tensorboard/", "indicates training output")
flags.DEFINE_integer("checkpoint_period", 1, "Number of epochs to save checkpoint.")
flags.DEFINE_string("model_path", "./model/", "The model directory")
flags.DEFINE_float("learning_rate", 0.01, "Initial learning rate.")
flags.DEFINE_string("optimizer", "sgd", "Optimizer to train")
flags.DEFINE_string("saved_model_path", "./saved_model/", "The path of the saved model")
flags.DEFINE_integer("model_version", 1, "The version of the model")
from distributed_tensorflow_b200.extra_flags import define as _define_extra
_define_extra()
FLAGS = flags.FLAGS


def main():
    trainer_lib.run(FLAGS, legacy=False, auto_stop=False)


if __name__ == "__main__":
    main()
