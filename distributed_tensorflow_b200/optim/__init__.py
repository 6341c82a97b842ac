"""The six optimizers the reference can select (/root/reference/trainer/task.py:41-56), with the
TensorFlow-1.x update rules, default hyper-parameters, slot names and slot initial values, so that
checkpoints carry the same tensors as a TF run would.

=========  ==============================  =============================================
flag       TF class                        slots (name -> initial value)
=========  ==============================  =============================================
sgd        GradientDescentOptimizer        --
adadelta   AdadeltaOptimizer               ``<v>/Adadelta`` 0, ``<v>/Adadelta_1`` 0
adagrad    AdagradOptimizer                ``<v>/Adagrad`` 0.1
adam       AdamOptimizer                   ``<v>/Adam`` 0, ``<v>/Adam_1`` 0, ``beta1_power``, ``beta2_power``
ftrl       FtrlOptimizer                   ``<v>/Ftrl`` 0.1, ``<v>/Ftrl_1`` 0
rmsprop    RMSPropOptimizer                ``<v>/RMSProp`` 1.0, ``<v>/RMSProp_1`` 0
=========  ==============================  =============================================

These NumPy implementations run the standalone (single-process) path and serve as oracles for the
C++ apply kernels (``csrc/native/optimizers.cc``) and the fused CUDA push kernels.
"""
from __future__ import annotations

from typing import Dict, Mapping

import numpy as np

OPTIMIZER_IDS = {"sgd": 0, "adadelta": 1, "adagrad": 2, "adam": 3, "ftrl": 4, "rmsprop": 5}


class Optimizer:
    """Stateless rule object; state (slots) lives in the variable store passed to ``apply``."""
    name = "optimizer"
    opt_id = -1

    def __init__(self, learning_rate: float):
        self.learning_rate = float(learning_rate)

    def slot_specs(self, var_name: str) -> Dict[str, float]:
        """slot tensor name -> initial fill value for a variable."""
        return {}

    def global_specs(self) -> Dict[str, float]:
        """non-slot scalar state (Adam beta powers): name -> initial value."""
        return {}

    def ensure_slots(self, store: Dict[str, np.ndarray], var_name: str) -> None:
        v = store[var_name]
        for s, init in self.slot_specs(var_name).items():
            if s not in store:
                store[s] = np.full_like(v, init, dtype=np.float32)
        for s, init in self.global_specs().items():
            if s not in store:
                store[s] = np.array(init, dtype=np.float32)

    def apply(self, store: Dict[str, np.ndarray], grads: Mapping[str, np.ndarray]) -> None:
        """In-place update of every variable named in ``grads`` (and its slots)."""
        for name in grads:
            self.ensure_slots(store, name)
        self._begin(store)
        for name, g in grads.items():
            self._apply_one(store, name, np.asarray(g, dtype=np.float32))
        self._finish(store)

    def _begin(self, store):
        pass

    def _finish(self, store):
        pass

    def _apply_one(self, store, name, g):
        raise NotImplementedError


class GradientDescentOptimizer(Optimizer):
    name, opt_id = "sgd", 0

    def _apply_one(self, store, name, g):
        store[name] -= np.float32(self.learning_rate) * g


class AdagradOptimizer(Optimizer):
    name, opt_id = "adagrad", 2

    def __init__(self, learning_rate, initial_accumulator_value: float = 0.1):
        super().__init__(learning_rate)
        self.initial_accumulator_value = initial_accumulator_value

    def slot_specs(self, v):
        return {v + "/Adagrad": self.initial_accumulator_value}

    def _apply_one(self, store, name, g):
        a = store[name + "/Adagrad"]
        a += g * g
        store[name] -= np.float32(self.learning_rate) * g / np.sqrt(a)


class AdamOptimizer(Optimizer):
    name, opt_id = "adam", 3

    def __init__(self, learning_rate, beta1=0.9, beta2=0.999, epsilon=1e-8):
        super().__init__(learning_rate)
        self.beta1, self.beta2, self.epsilon = beta1, beta2, epsilon

    def slot_specs(self, v):
        return {v + "/Adam": 0.0, v + "/Adam_1": 0.0}

    def global_specs(self):
        return {"beta1_power": self.beta1, "beta2_power": self.beta2}

    def _begin(self, store):
        b1p, b2p = np.float32(store["beta1_power"]), np.float32(store["beta2_power"])
        self._lr_t = np.float32(self.learning_rate) * np.sqrt(np.float32(1) - b2p) / (np.float32(1) - b1p)

    def _apply_one(self, store, name, g):
        b1, b2 = np.float32(self.beta1), np.float32(self.beta2)
        m, v = store[name + "/Adam"], store[name + "/Adam_1"]
        m[...] = b1 * m + (np.float32(1) - b1) * g
        v[...] = b2 * v + (np.float32(1) - b2) * g * g
        store[name] -= self._lr_t * m / (np.sqrt(v) + np.float32(self.epsilon))

    def _finish(self, store):
        store["beta1_power"] = np.array(np.float32(store["beta1_power"]) * np.float32(self.beta1), dtype=np.float32)
        store["beta2_power"] = np.array(np.float32(store["beta2_power"]) * np.float32(self.beta2), dtype=np.float32)


class AdadeltaOptimizer(Optimizer):
    name, opt_id = "adadelta", 1

    def __init__(self, learning_rate, rho=0.95, epsilon=1e-8):
        super().__init__(learning_rate)
        self.rho, self.epsilon = rho, epsilon

    def slot_specs(self, v):
        return {v + "/Adadelta": 0.0, v + "/Adadelta_1": 0.0}

    def _apply_one(self, store, name, g):
        rho, eps = np.float32(self.rho), np.float32(self.epsilon)
        acc, accu = store[name + "/Adadelta"], store[name + "/Adadelta_1"]
        acc[...] = rho * acc + (np.float32(1) - rho) * g * g
        u = np.sqrt(accu + eps) / np.sqrt(acc + eps) * g
        accu[...] = rho * accu + (np.float32(1) - rho) * u * u
        store[name] -= np.float32(self.learning_rate) * u


class FtrlOptimizer(Optimizer):
    name, opt_id = "ftrl", 4

    def __init__(self, learning_rate, learning_rate_power=-0.5, initial_accumulator_value=0.1,
                 l1_regularization_strength=0.0, l2_regularization_strength=0.0):
        super().__init__(learning_rate)
        self.lr_power, self.init_acc = learning_rate_power, initial_accumulator_value
        self.l1, self.l2 = l1_regularization_strength, l2_regularization_strength

    def slot_specs(self, v):
        return {v + "/Ftrl": self.init_acc, v + "/Ftrl_1": 0.0}

    def _apply_one(self, store, name, g):
        lr, l1, l2, p = map(np.float32, (self.learning_rate, self.l1, self.l2, self.lr_power))
        n, z, var = store[name + "/Ftrl"], store[name + "/Ftrl_1"], store[name]
        n_new = n + g * g
        sigma = (np.power(n_new, -p) - np.power(n, -p)) / lr
        z += g - sigma * var
        quad = np.power(n_new, -p) / lr + np.float32(2) * l2
        var[...] = np.where(np.abs(z) > l1, (np.sign(z) * l1 - z) / quad, np.float32(0))
        n[...] = n_new


class RMSPropOptimizer(Optimizer):
    name, opt_id = "rmsprop", 5

    def __init__(self, learning_rate, decay=0.9, momentum=0.0, epsilon=1e-10):
        super().__init__(learning_rate)
        self.decay, self.momentum, self.epsilon = decay, momentum, epsilon

    def slot_specs(self, v):
        return {v + "/RMSProp": 1.0, v + "/RMSProp_1": 0.0}

    def _apply_one(self, store, name, g):
        d, mo, eps = np.float32(self.decay), np.float32(self.momentum), np.float32(self.epsilon)
        ms, mom = store[name + "/RMSProp"], store[name + "/RMSProp_1"]
        ms += (g * g - ms) * (np.float32(1) - d)
        mom[...] = mom * mo + np.float32(self.learning_rate) * g / np.sqrt(ms + eps)
        store[name] -= mom


_CLASSES = {
    "sgd": GradientDescentOptimizer, "adadelta": AdadeltaOptimizer, "adagrad": AdagradOptimizer,
    "adam": AdamOptimizer, "ftrl": FtrlOptimizer, "rmsprop": RMSPropOptimizer,
}


def get_optimizer(name: str, learning_rate: float) -> Optimizer:
    """Factory keyed by the reference's ``--optimizer`` values; ``KeyError`` for anything else (the
    entry points turn that into ``Unknow optimizer: ..., exit now`` + exit status 1)."""
    return _CLASSES[name](learning_rate)
