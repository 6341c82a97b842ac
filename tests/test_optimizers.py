"""Six optimizers: NumPy rules vs an independent restatement of the TF-1 formulas
and vs the C++ apply kernels that run on the parameter server."""
import ctypes as C

import numpy as np
import pytest

from distributed_tensorflow_b200 import _native
from distributed_tensorflow_b200.models.linear import LinearModel
from distributed_tensorflow_b200.optim import OPTIMIZER_IDS, get_optimizer

NAMES = ["sgd", "adadelta", "adagrad", "adam", "ftrl", "rmsprop"]


def _oracle_step(name, lr, var, slots, g, t):
    """float64 restatement, one step; slots is a dict mutated in place."""
    if name == "sgd":
        return var - lr * g
    if name == "adagrad":
        slots["a"] = slots.get("a", 0.1) + g * g
        return var - lr * g / np.sqrt(slots["a"])
    if name == "adam":
        b1, b2, eps = 0.9, 0.999, 1e-8
        slots["m"] = b1 * slots.get("m", 0.0) + (1 - b1) * g
        slots["v"] = b2 * slots.get("v", 0.0) + (1 - b2) * g * g
        lr_t = lr * np.sqrt(1 - b2 ** t) / (1 - b1 ** t)
        return var - lr_t * slots["m"] / (np.sqrt(slots["v"]) + eps)
    if name == "adadelta":
        rho, eps = 0.95, 1e-8
        slots["acc"] = rho * slots.get("acc", 0.0) + (1 - rho) * g * g
        u = np.sqrt(slots.get("accu", 0.0) + eps) / np.sqrt(slots["acc"] + eps) * g
        slots["accu"] = rho * slots.get("accu", 0.0) + (1 - rho) * u * u
        return var - lr * u
    if name == "ftrl":
        n = slots.get("n", 0.1)
        n2 = n + g * g
        sigma = (np.sqrt(n2) - np.sqrt(n)) / lr
        slots["z"] = slots.get("z", 0.0) + g - sigma * var
        slots["n"] = n2
        return -slots["z"] / (np.sqrt(n2) / lr) if abs(slots["z"]) > 0 else 0.0
    if name == "rmsprop":
        slots["ms"] = 0.9 * slots.get("ms", 1.0) + 0.1 * g * g
        slots["mom"] = 0.0 * slots.get("mom", 0.0) + lr * g / np.sqrt(slots["ms"] + 1e-10)
        return var - slots["mom"]
    raise KeyError(name)


@pytest.mark.parametrize("name", NAMES)
def test_numpy_rule_matches_oracle(name):
    rng = np.random.RandomState(0)
    opt = get_optimizer(name, 0.05)
    store = {"w": np.array(0.3, np.float32)}
    var, slots = 0.3, {}
    for t in range(1, 6):
        g = float(rng.randn())
        opt.apply(store, {"w": np.array(g, np.float32)})
        var = _oracle_step(name, 0.05, var, slots, g, t)
        assert np.allclose(store["w"], var, rtol=2e-4, atol=1e-6), (name, t)


def test_slot_names_and_initial_values():
    expect = {
        "sgd": {}, "adagrad": {"w/Adagrad": 0.1}, "adam": {"w/Adam": 0.0, "w/Adam_1": 0.0},
        "adadelta": {"w/Adadelta": 0.0, "w/Adadelta_1": 0.0}, "ftrl": {"w/Ftrl": 0.1, "w/Ftrl_1": 0.0},
        "rmsprop": {"w/RMSProp": 1.0, "w/RMSProp_1": 0.0},
    }
    for name, slots in expect.items():
        opt = get_optimizer(name, 0.01)
        store = {"w": np.zeros((2, 3), np.float32)}
        opt.ensure_slots(store, "w")
        for s, init in slots.items():
            assert store[s].shape == (2, 3) and np.all(store[s] == np.float32(init)), (name, s)
    store = {"w": np.zeros(1, np.float32)}
    get_optimizer("adam", 0.01).ensure_slots(store, "w")
    assert np.isclose(store["beta1_power"], 0.9) and np.isclose(store["beta2_power"], 0.999)


def test_unknown_optimizer_raises_keyerror():
    with pytest.raises(KeyError):
        get_optimizer("momentum", 0.1)


@pytest.mark.parametrize("name", NAMES)
def test_cpp_apply_kernels_match_numpy(name):
    L = _native.lib()
    rng = np.random.RandomState(1)
    n = 257
    fp = C.POINTER(C.c_float)
    p = lambda a: a.ctypes.data_as(fp)
    opt = get_optimizer(name, 0.03)
    store = {"w": rng.randn(n).astype(np.float32)}
    opt.ensure_slots(store, "w")
    cw = store["w"].copy()
    cs = {k: v.copy() for k, v in store.items() if k.startswith("w/")}
    b1p, b2p = 0.9, 0.999
    for _ in range(4):
        g = rng.randn(n).astype(np.float32)
        opt.apply(store, {"w": g})
        if name == "sgd":
            L.dtfn_apply_sgd(p(cw), p(g), n, 0.03)
        elif name == "adagrad":
            L.dtfn_apply_adagrad(p(cw), p(cs["w/Adagrad"]), p(g), n, 0.03)
        elif name == "adam":
            L.dtfn_apply_adam(p(cw), p(cs["w/Adam"]), p(cs["w/Adam_1"]), p(g), n, 0.03, 0.9, 0.999, 1e-8, b1p, b2p)
            b1p, b2p = b1p * 0.9, b2p * 0.999
        elif name == "adadelta":
            L.dtfn_apply_adadelta(p(cw), p(cs["w/Adadelta"]), p(cs["w/Adadelta_1"]), p(g), n, 0.03, 0.95, 1e-8)
        elif name == "ftrl":
            L.dtfn_apply_ftrl(p(cw), p(cs["w/Ftrl"]), p(cs["w/Ftrl_1"]), p(g), n, 0.03, 0.0, 0.0, -0.5)
        elif name == "rmsprop":
            L.dtfn_apply_rmsprop(p(cw), p(cs["w/RMSProp"]), p(cs["w/RMSProp_1"]), p(g), n, 0.03, 0.9, 0.0, 1e-10)
    assert np.allclose(cw, store["w"], rtol=1e-4, atol=1e-5)
    for k, v in cs.items():
        assert np.allclose(v, store[k], rtol=1e-4, atol=1e-6), k
    assert set(OPTIMIZER_IDS) == set(NAMES)


def test_toy_model_loss_and_grads_closed_form():
    m = LinearModel()
    vals = {"weight": np.float32(1.5), "bias": np.float32(-0.5)}
    x, y = np.array([[0.25]]), np.array([[3.0]])
    loss, g = m.loss_and_grads(vals, x, y)
    r = 3.0 - 0.25 * 1.5 + 0.5
    assert np.isclose(loss, r * r) and np.isclose(g["weight"], -2 * 0.25 * r) and np.isclose(g["bias"], -2 * r)
    out = m.predict({"weight": np.float32(2), "bias": np.float32(10)}, {"keys": [[11.0], [2.0]], "features": [[1], [2]]})
    assert out["keys"].dtype == np.int32 and out["keys"].tolist() == [[11], [2]]
    assert out["prediction"].tolist() == [[12.0], [14.0]]


@pytest.mark.parametrize("name", NAMES)
def test_device_apply_rule_matches_numpy_optimizer(name):
    """The per-element rule the CUDA ps-side apply kernels use (csrc/kernels/wd_optim.cuh, compiled for the
    host into libdtf_kernels.so) against the NumPy optimizer of the same name: weights and both slots, five
    steps, vectors with repeated sign changes.  Covers `--model=wide_deep --optimizer=<any of the six>`
    (reference trainer/task.py:41-56) without needing a GPU."""
    from distributed_tensorflow_b200.ops import _lib
    L = _lib.lib()
    L.dtf_opt_update_host.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_float,
                                      C.c_longlong]
    L.dtf_opt_slot1_init.restype = C.c_float
    rng = np.random.RandomState(1)
    n, lr = 257, 0.05
    opt = get_optimizer(name, lr)
    oid = OPTIMIZER_IDS[name]
    store = {"w": (rng.randn(n) * 0.3).astype(np.float32)}
    opt.ensure_slots(store, "w")
    specs = opt.slot_specs("w")
    slot_names = list(specs)
    assert L.dtf_opt_num_slots(oid) == len(slot_names)
    if slot_names:
        assert abs(L.dtf_opt_slot1_init(oid) - specs[slot_names[0]]) < 1e-7
    w = store["w"].copy()
    s = [np.full(n, specs[k], np.float32) for k in slot_names]
    ptr = lambda a: a.ctypes.data_as(C.c_void_p)
    for t in range(1, 6):
        g = (rng.randn(n) * (1.0 if t % 2 else 0.01)).astype(np.float32)
        opt.apply(store, {"w": g})
        rc = L.dtf_opt_update_host(oid, ptr(w), ptr(g), ptr(s[0]) if len(s) > 0 else None, ptr(s[1]) if len(s) > 1 else None,
                                   n, lr, t)
        assert rc == 0
        assert np.allclose(w, store["w"], rtol=1e-4, atol=1e-5), (name, t, np.abs(w - store["w"]).max())
        for k, arr in zip(slot_names, s):
            assert np.allclose(arr, store[k], rtol=1e-4, atol=1e-7), (name, k, t)
