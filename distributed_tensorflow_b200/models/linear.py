"""The reference's toy model: scalar linear regression with a ``keys`` pass-through.

Graph being restated (/root/reference/trainer/task.py:62-71, 130-139)::

    keys    = identity(keys_placeholder)            int32 [None, 1]
    predict = X * w + b                             float32 [None, 1]
    loss    = reduce_sum(square(Y - X * w - b))
    vars    : weight (f32 scalar, 0.0), bias (f32 scalar, 0.0), global_step (int32 scalar, 0)

The math is three scalars, so it runs in NumPy on the host in both standalone and
parameter-server mode (exactly where the reference ran it: README.md:14 sets
``CUDA_VISIBLE_DEVICES=''``).
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import numpy as np


class VariableSpec:
    def __init__(self, name: str, initial, dtype, trainable: bool = True):
        self.name, self.dtype, self.trainable = name, np.dtype(dtype), trainable
        self.initial = np.array(initial, dtype=dtype)


class LinearModel:
    name = "linear"

    # creation order matters: it drives round-robin placement (weight->ps0, bias->ps1, global_step->ps0)
    variables: List[VariableSpec] = [
        VariableSpec("weight", 0.0, np.float32),
        VariableSpec("bias", 0.0, np.float32),
        VariableSpec("global_step", 0, np.int32, trainable=False),
    ]
    global_step_name = "global_step"

    # serving signature (reference task.py:164-173): alias -> (tensor name, dtype enum, shape)
    signature_inputs = {"keys": ("Placeholder:0", np.int32, (-1, 1)), "features": ("Placeholder_1:0", np.float32, (-1, 1))}
    signature_outputs = {"keys": ("Identity:0", np.int32, (-1, 1)), "prediction": ("add:0", np.float32, (-1, 1))}
    # legacy session_bundle aliases (reference task.py:104-105, task_supervisor.py:189-198)
    legacy_inputs = {"keys": "keys", "X": "features"}
    legacy_outputs = {"keys": "keys", "predict": "prediction"}

    def trainable_names(self) -> List[str]:
        return [v.name for v in self.variables if v.trainable]

    def initial_values(self) -> Dict[str, np.ndarray]:
        return {v.name: v.initial.copy() for v in self.variables}

    @staticmethod
    def loss(values: Dict[str, np.ndarray], x, y) -> np.float32:
        x = np.asarray(x, np.float32)
        y = np.asarray(y, np.float32)
        r = y - x * np.float32(values["weight"]) - np.float32(values["bias"])
        return np.float32(np.sum(r * r))

    @staticmethod
    def loss_and_grads(values: Dict[str, np.ndarray], x, y) -> Tuple[np.float32, Dict[str, np.ndarray]]:
        x = np.asarray(x, np.float32)
        y = np.asarray(y, np.float32)
        r = y - x * np.float32(values["weight"]) - np.float32(values["bias"])
        loss = np.float32(np.sum(r * r))
        dw = np.float32(np.sum(np.float32(-2) * x * r))
        db = np.float32(np.sum(np.float32(-2) * r))
        return loss, {"weight": np.array(dw, np.float32), "bias": np.array(db, np.float32)}

    @staticmethod
    def predict(values: Dict[str, np.ndarray], inputs: Dict[str, np.ndarray]) -> Dict[str, np.ndarray]:
        keys = np.asarray(inputs["keys"]).astype(np.int32)        # floats are cast (README.md:40 sends 11.0)
        feats = np.asarray(inputs["features"], np.float32)
        pred = feats * np.float32(values["weight"]) + np.float32(values["bias"])
        return {"keys": keys, "prediction": pred.astype(np.float32)}


def make_toy_data(n: int = 100, rng=None):
    """``train_X = linspace(-1, 1, 100)``, ``train_Y = 2 X + randn * 0.33 + 10`` -- unseeded in the
    reference (task.py:36-37); pass ``rng`` for reproducible tests."""
    rng = np.random if rng is None else rng
    train_x = np.linspace(-1, 1, n)
    train_y = 2 * train_x + rng.randn(*train_x.shape) * 0.33 + 10
    return train_x, train_y
