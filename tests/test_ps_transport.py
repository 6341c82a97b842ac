"""C++ TCP parameter server: variable store, fused push/pull, Hogwild apply, named queues, ps-side
save, failure surfacing (reference trainer/task.py:118-122,138; auto_stop_ps/task.py:127-150)."""
import threading
import time

import numpy as np
import pytest

from distributed_tensorflow_b200 import _native
from distributed_tensorflow_b200.checkpoint.bundle import BundleReader
from distributed_tensorflow_b200.cluster import ClusterSpec, Server
from distributed_tensorflow_b200.models.linear import LinearModel
from distributed_tensorflow_b200.optim import get_optimizer
from distributed_tensorflow_b200.session import LocalSession, PSClient, PSSession


@pytest.fixture()
def ps():
    srv = Server(ClusterSpec({"ps": ["127.0.0.1:0"]}), "ps", 0)
    yield srv, "127.0.0.1:%d" % srv.port
    srv.stop()


def test_init_pull_list_info(ps):
    srv, addr = ps
    c = PSClient(addr)
    c.ping()
    assert c.init_var("weight", np.array(0.5, np.float32)) is True
    assert c.init_var("weight", np.array(9.0, np.float32)) is False          # exists, not overwritten
    assert float(c.pull("weight")) == 0.5
    c.init_var("weight", np.array(9.0, np.float32), overwrite=True)
    assert float(c.pull("weight")) == 9.0
    c.init_var("emb", np.arange(12, dtype=np.float32).reshape(3, 4))
    c.init_var("global_step", np.array(0, np.int32))
    assert c.list() == ["emb", "global_step", "weight"]
    assert c.var_info("emb") == (np.dtype(np.float32), (3, 4), 48)
    assert c.var_info("nope") is None and srv.num_variables() == 3
    with pytest.raises(KeyError):
        c.pull("nope")
    c.close()


@pytest.mark.parametrize("name", ["sgd", "adagrad", "adam", "adadelta", "ftrl", "rmsprop"])
def test_push_pull_matches_local_optimizer(ps, name):
    _, addr = ps
    c = PSClient(addr)
    opt = get_optimizer(name, 0.05)
    rng = np.random.RandomState(3)
    local = {"w": rng.randn(5).astype(np.float32), "b": np.array(0.1, np.float32)}
    for k, v in local.items():
        c.init_var(k, v)
    c.init_var("global_step", np.array(0, np.int32))
    for t in range(1, 5):
        g = {"w": rng.randn(5).astype(np.float32), "b": np.array(rng.randn(), np.float32)}
        opt.apply(local, g)
        vals, step = c.push_pull(["w", "b"], [g["w"], g["b"]], opt.opt_id, 0.05, "global_step")
        assert step == t
        assert np.allclose(vals[0], local["w"], rtol=1e-4, atol=1e-6)
        assert np.allclose(vals[1], local["b"], rtol=1e-4, atol=1e-6)
    slots = [n for n in c.list() if "/" in n]
    assert sorted(slots) == sorted(k for k in local if "/" in k)
    for s in slots:
        assert np.allclose(c.pull(s), local[s], rtol=1e-4, atol=1e-6), s
    if name == "adam":
        assert np.isclose(float(c.pull("beta1_power")), 0.9 ** 5) and np.isclose(float(c.pull("beta2_power")), 0.999 ** 5)
    c.close()


def test_hogwild_updates_from_many_clients_all_land(ps):
    """No locks, no barrier: with SGD every pushed gradient must still be applied exactly once and
    global_step must count every step."""
    _, addr = ps
    boot = PSClient(addr)
    boot.init_var("w", np.zeros(1, np.float32))
    boot.init_var("global_step", np.array(0, np.int64))
    n_threads, n_steps = 4, 200

    def work():
        c = PSClient(addr)
        for _ in range(n_steps):
            c.push_pull(["w"], [np.ones(1, np.float32)], 0, 0.5, "global_step")
        c.close()
    ts = [threading.Thread(target=work) for _ in range(n_threads)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert int(boot.pull("global_step")) == n_threads * n_steps
    assert np.isclose(float(boot.pull("w")[0]), -0.5 * n_threads * n_steps, rtol=0.02)   # races may lose a few
    boot.close()


def test_named_queue_blocks_until_tokens_arrive(ps):
    _, addr = ps
    a, b = PSClient(addr), PSClient(addr)
    a.queue_create("worker_done_queue_0", 2)
    assert a.dequeue("worker_done_queue_0", timeout_ms=150) is None            # nothing yet -> timeout
    got = []
    t = threading.Thread(target=lambda: got.extend(a.dequeue("worker_done_queue_0", timeout_ms=5000) for _ in range(2)))
    t.start()
    time.sleep(0.2)
    assert t.is_alive()                                                        # still blocked
    b.enqueue("worker_done_queue_0", 1)
    b.enqueue("worker_done_queue_0", 1)
    t.join(5)
    assert got == [1, 1]
    a.close(), b.close()


def test_ps_side_save_writes_bundle(ps, tmp_path):
    _, addr = ps
    c = PSClient(addr)
    c.init_var("weight", np.array(2.0, np.float32))
    c.init_var("global_step", np.array(11, np.int32))
    c.init_var("skipme", np.zeros(2, np.float32))
    prefix = str(tmp_path / "part-00000")
    c.save(prefix, exclude=["skipme"])
    with BundleReader(prefix) as r:
        assert r.keys() == ["global_step", "weight"]
        assert int(r.get_tensor("global_step")) == 11 and r.get_tensor("global_step").dtype == np.int32
    c.close()


def test_dead_or_mute_ps_surfaces_as_error(ps):
    srv, addr = ps
    c = PSClient(addr, timeout_ms=500)
    c.init_var("w", np.zeros(1, np.float32))
    srv.inject_drop(1)                                                         # next request vanishes
    with pytest.raises(ConnectionError):
        c.pull("w")
    c.close()
    with pytest.raises(ConnectionError):
        PSClient("127.0.0.1:1", timeout_ms=300)                                # nothing listens there


def test_shutdown_request_releases_join(ps):
    srv, addr = ps
    done = threading.Event()
    t = threading.Thread(target=lambda: (srv.join(), done.set()))
    t.start()
    time.sleep(0.1)
    assert not done.is_set()
    c = PSClient(addr)
    c.shutdown_server()
    assert done.wait(5)
    c.close()


def test_ps_session_two_ps_tasks_round_robin_and_convergence():
    s0 = Server(ClusterSpec({"ps": ["127.0.0.1:0"]}), "ps", 0)
    s1 = Server(ClusterSpec({"ps": ["127.0.0.1:0"]}), "ps", 0)
    cluster = ClusterSpec({"ps": ["127.0.0.1:%d" % s0.port, "127.0.0.1:%d" % s1.port], "worker": ["127.0.0.1:1"]})
    try:
        model = LinearModel()
        sess = PSSession(model, get_optimizer("sgd", 0.01), cluster, connect_retry_secs=1)
        assert not sess.is_ready()
        sess.run_init()
        assert sess.is_ready()
        listed = lambda c: sorted(n for n in c.list() if n != PSSession.READY_MARKER)    # the chief's ready flag sits on every ps
        assert listed(sess.clients[0]) == ["global_step", "weight"] and listed(sess.clients[1]) == ["bias"]
        ref = LocalSession(model, get_optimizer("sgd", 0.01))
        ref.run_init()
        rng = np.random.RandomState(0)
        xs = np.linspace(-1, 1, 100)
        ys = 2 * xs + rng.randn(100) * 0.33 + 10
        for _ in range(5):
            for x, y in zip(xs, ys):
                sess.train_step(np.array([[x]]), np.array([[y]]))
                ref.train_step(np.array([[x]]), np.array([[y]]))
        assert sess.global_step() == 500
        assert np.isclose(float(sess.value("weight")), float(ref.value("weight")), rtol=1e-5)
        assert np.isclose(float(sess.value("bias")), float(ref.value("bias")), rtol=1e-5)
        snap = sess.snapshot()
        assert sorted(snap) == ["bias", "global_step", "weight"] and snap["global_step"].dtype == np.int32
        sess.close()
    finally:
        s0.stop(), s1.stop()


def test_client_recovers_on_a_fresh_connection_after_a_lost_exchange(ps):
    """A request that got no answer in time must not leave the socket one reply out of step: the client
    drops that connection and the next call runs on a new one, with its own reply."""
    srv, addr = ps
    c = PSClient(addr, timeout_ms=400)
    c.init_var("w", np.array([1.0, 2.0], np.float32))
    c.init_var("global_step", np.array(5, np.int32))
    srv.inject_drop(1)                                                         # the server vanishes mid-request
    with pytest.raises(ConnectionError):
        c.pull("w")
    # same client object: reconnects, and replies line up with requests again
    assert c.pull("w").tolist() == [1.0, 2.0]
    assert c.assign_add("global_step", 2) == 7
    assert sorted(c.list()) == ["global_step", "w"]
    c.close()


def test_full_bounded_queue_holds_an_enqueue_only_for_the_client_timeout(ps):
    _, addr = ps
    a = PSClient(addr, timeout_ms=700)
    a.queue_create("q", 1)
    a.enqueue("q", 1)
    t0 = time.time()
    with pytest.raises(_native.NativeError):
        a.enqueue("q", 2)                                                      # full: gives up after ~0.7 s
    assert 0.5 < time.time() - t0 < 3.0
    assert a.dequeue("q", timeout_ms=500) == 1                                 # connection still in step
    a.enqueue("q", 3)
    assert a.dequeue("q", timeout_ms=500) == 3
    a.close()


def test_init_var_rejects_sizes_that_do_not_match_the_shape_and_non_integer_step(ps):
    import ctypes as C
    _, addr = ps
    c = PSClient(addr)
    L = _native.lib()
    data = np.zeros(3, np.float32)
    dims = (C.c_longlong * 1)(5)                                               # claims 5 floats, sends 3
    rc = L.dtfn_client_init_var(c._h, b"bad", 1, 1, dims, data.ctypes.data_as(C.c_void_p), data.nbytes, 1)
    assert rc == -12
    rc = L.dtfn_client_init_var(c._h, b"bad", 19, 1, (C.c_longlong * 1)(3), data.ctypes.data_as(C.c_void_p), data.nbytes, 1)
    assert rc == -11                                                           # DT_HALF is not a ps dtype
    c.init_var("fstep", np.array(0.0, np.float32))
    with pytest.raises(_native.NativeError):
        c.assign_add("fstep", 1)
    with pytest.raises(_native.NativeError):
        c.push_pull([], [], 0, 0.1, "fstep")
    c.close()


def test_many_short_connections_are_reaped_and_stop_is_clean():
    srv = Server(ClusterSpec({"ps": ["127.0.0.1:0"]}), "ps", 0)
    addr = "127.0.0.1:%d" % srv.port
    for i in range(50):
        c = PSClient(addr)
        c.ping()
        c.close()
    keep = PSClient(addr)
    keep.init_var("w", np.zeros(1, np.float32))
    other = Server(ClusterSpec({"ps": ["127.0.0.1:0"]}), "ps", 0)              # its sockets may reuse freed fd numbers
    oc = PSClient("127.0.0.1:%d" % other.port)
    srv.stop()                                                                 # must not touch `other`'s descriptors
    oc.ping()
    oc.close()
    other.stop()
    keep.close()
