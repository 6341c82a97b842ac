"""Native JSON fast path of the serving endpoint (csrc/native/json_fast.cc) against Python's json module:
same arrays for every rectangular numeric body, a clean fall-back for everything else, and formatted
responses that json.loads reads back to the same float32 / int values."""
import json

import numpy as np
import pytest

from distributed_tensorflow_b200.serving.server import _FastJson


@pytest.fixture(scope="module")
def fast():
    f = _FastJson.get()
    assert f is not None, "libdtf_native.so must provide the JSON fast path"
    return f


def test_readme_request_and_random_bodies_parse_like_json(fast):
    body = b'{"keys": [[11.0], [2.0]], "features": [[1], [2]]}'
    assert fast.parse(body, "keys").tolist() == [[11.0], [2.0]]
    assert fast.parse(body, "features").tolist() == [[1.0], [2.0]]
    assert fast.parse(body, "absent") is None
    rng = np.random.RandomState(0)
    for shape in [(1,), (7,), (3, 4), (5, 1), (2, 3, 4), (64, 39)]:
        a = (rng.randn(*shape) * 10 ** rng.randint(-8, 8)).astype(np.float64)
        ints = rng.randint(-10**6, 10**6, size=shape)
        txt = json.dumps({"other": {"x": [1, "s]", {"k": "}"}]}, "a": a.tolist(), "s": 'quote" [', "b": ints.tolist()},
                         indent=rng.choice([None, 1])).encode()
        np.testing.assert_array_equal(fast.parse(txt, "a"), np.asarray(json.loads(txt)["a"], np.float64))   # bit-exact doubles
        np.testing.assert_array_equal(fast.parse(txt, "b"), ints.astype(np.float64))
    exps = b'{"v": [1e3, -2.5E-3, 1.7976931348623157e308, 5e-324, 0.1, 123456789012345678901234567890, -0.0]}'
    np.testing.assert_array_equal(fast.parse(exps, "v"), np.asarray(json.loads(exps)["v"], np.float64))
    assert fast.parse(b'{"e": [], "f": [[], []]}', "e").shape == (0,)
    assert fast.parse(b'{"e": [], "f": [[], []]}', "f").shape == (2, 0)


@pytest.mark.parametrize("body", [
    b'{"a": [[1, 2], [3]]}',                 # ragged
    b'{"a": [1, [2]]}',                      # mixed depth
    b'{"a": [1, "x"]}',                      # not a number
    b'{"a": [true]}', b'{"a": [null]}', b'{"a": 3}', b'{"a": {"b": [1]}}',
    b'{"a": [1, 2', b'[1, 2]', b'{"a": [01x]}', b'{"a": [1.]}', b'{"a": [.5]}', b'{"a": [[[[[1]]]]]}',
])
def test_anything_else_falls_back(fast, body):
    assert fast.parse(body, "a") is None


def test_formatting_round_trips_float32_and_ints(fast):
    rng = np.random.RandomState(1)
    for shape in [(1,), (4, 1), (3, 5), (2, 2, 2)]:
        f = (rng.randn(*shape) * 10.0 ** rng.randint(-6, 6, size=shape)).astype(np.float32)
        back = np.asarray(json.loads(fast.format(f)), np.float32)
        np.testing.assert_array_equal(back, f)
        i = rng.randint(-2**40, 2**40, size=shape)
        assert json.loads(fast.format(i)) == i.tolist()
    assert fast.format(np.array([[12.0], [14.0]], np.float32)) == b"[[12.0], [14.0]]"     # floats stay floats
    assert fast.format(np.array([[11], [2]], np.int32)) == b"[[11], [2]]"
    assert fast.format(np.zeros((0,), np.float32)) == b"[]"
    assert fast.format(np.array(["a"])) is None


def test_model_server_uses_the_fast_path_and_matches_the_generic_path(tmp_path):
    from distributed_tensorflow_b200.export import export_saved_model
    from distributed_tensorflow_b200.models.linear import LinearModel
    from distributed_tensorflow_b200.optim import get_optimizer
    from distributed_tensorflow_b200.serving import ModelServer
    from distributed_tensorflow_b200.session import LocalSession
    s = LocalSession(LinearModel(), get_optimizer("sgd", 0.01))
    s.run_init()
    s.assign("weight", np.array(2.0, np.float32))
    s.assign("bias", np.array(10.0, np.float32))
    export_saved_model(s, str(tmp_path / "m"), 1)
    ms = ModelServer(str(tmp_path / "m"))
    body = b'{"keys": [[11.0], [2.0]], "features": [[1], [2]]}'
    assert json.loads(ms.predict_json(body)) == {"keys": [[11], [2]], "prediction": [[12.0], [14.0]]}
    assert json.loads(ms.predict_json(body)) == ms.predict(json.loads(body))
    enveloped = b'{"data": {"keys": [[1]], "features": [[0.5]]}}'                      # generic path
    assert json.loads(ms.predict_json(enveloped)) == {"keys": [[1]], "prediction": [[11.0]]}
    with pytest.raises(ValueError, match="missing input"):
        ms.predict_json(b'{"keys": [[1]]}')
    big = {"keys": [[float(i)] for i in range(4096)], "features": [[i * 0.001] for i in range(4096)]}
    out = json.loads(ms.predict_json(json.dumps(big).encode()))
    np.testing.assert_allclose(np.asarray(out["prediction"])[:, 0], 2.0 * np.arange(4096) * 0.001 + 10.0, rtol=1e-6)
    assert out["keys"][4095] == [4095]
