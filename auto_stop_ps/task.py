#!/usr/bin/env python
"""``python -m auto_stop_ps.task`` -- trainer.task plus "ps tasks exit once every worker and the
master are done" (interface of /root/reference/auto_stop_ps/task.py; ``max_epochs`` defaults to 20
there, :18).  The reference ships this directory without an ``__init__.py`` (a namespace module); here it has one so that setup.py can list it as a package."""
from distributed_tensorflow_b200 import flags, trainer_lib
from distributed_tensorflow_b200.extra_flags import define as _define_extra

flags.DEFINE_integer("max_epochs", 20, "Number of steps to run trainer.")
flags.DEFINE_string("checkpoint_path", "./checkpoint/", "The checkpoint directory")
flags.DEFINE_string("output_path", "./tensorboard/", "indicates training output")
flags.DEFINE_integer("checkpoint_period", 1, "Number of epochs to save checkpoint.")
flags.DEFINE_string("model_path", "./model/", "The model directory")
flags.DEFINE_float("learning_rate", 0.01, "Initial learning rate.")
flags.DEFINE_string("optimizer", "sgd", "Optimizer to train")
flags.DEFINE_string("saved_model_path", "./saved_model/", "The path of the saved model")
flags.DEFINE_integer("model_version", 1, "The version of the model")
_define_extra()
FLAGS = flags.FLAGS


def main():
    trainer_lib.run(FLAGS, legacy=False, auto_stop=True)


if __name__ == "__main__":
    main()
