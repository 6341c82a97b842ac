"""Control-plane collectives of the GPU job over the C++ ps transport (reference trainer/task.py:110-127:
cluster and roles come from TF_CONFIG, every task talks through the servers it lists)."""
import threading

import pytest

from distributed_tensorflow_b200.cluster import ClusterSpec, Server
from distributed_tensorflow_b200.parallel.control import TransportGroup, rank_layout


def test_rank_layout_master_first_then_workers_and_store_on_ps0():
    c = ClusterSpec({"ps": ["h:1", "h:2"], "worker": ["h:3", "h:4"], "master": ["h:5"]})
    assert rank_layout(c, "master", 0) == (0, 3, "h:1")
    assert rank_layout(c, "worker", 0) == (1, 3, "h:1")
    assert rank_layout(c, "worker", 1) == (2, 3, "h:1")
    assert rank_layout(c, "ps", 1) == (None, 3, "h:1")
    c2 = ClusterSpec({"worker": ["h:3"], "master": ["h:5"]})
    assert rank_layout(c2, "worker", 0) == (1, 2, "h:5")          # no ps listed: the master's server hosts the store
    with pytest.raises(ValueError):
        rank_layout(ClusterSpec({"ps": ["h:1"]}), "ps", 0)


def test_collectives_across_four_ranks():
    srv = Server(ClusterSpec({"ps": ["127.0.0.1:0"]}), "ps", 0)
    addr = "127.0.0.1:%d" % srv.port
    world, results, errors = 4, {}, []

    def run(rank):
        try:
            g = TransportGroup(addr, rank, world, timeout_s=20)
            log = []
            for round_ in range(3):
                out = [None] * world
                g.all_gather_object(out, {"rank": rank, "round": round_, "blob": b"x" * (rank * 7 + round_)})
                log.append([(o["rank"], o["round"], len(o["blob"])) for o in out])
                box = [("tag", round_, rank)] if rank == round_ % world else [None]
                g.broadcast_object_list(box, src=round_ % world)
                log.append(box[0])
                g.barrier()
            results[rank] = log
            g.close()
        except Exception as e:  # noqa: BLE001
            errors.append((rank, repr(e)))

    ts = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    [t.start() for t in ts]
    [t.join(60) for t in ts]
    srv.stop()
    assert not errors, errors
    for rank in range(world):
        assert results[rank] == results[0]
    assert results[0][0] == [(r, 0, r * 7) for r in range(world)]
    assert results[0][1] == ("tag", 0, 0) and results[0][3] == ("tag", 1, 1)


def test_missing_peer_surfaces_as_connection_error_not_a_hang():
    srv = Server(ClusterSpec({"ps": ["127.0.0.1:0"]}), "ps", 0)
    g = TransportGroup("127.0.0.1:%d" % srv.port, 0, 2, timeout_s=0.6)
    with pytest.raises(ConnectionError, match="missing or dead"):
        g.barrier()
    g.close()
    srv.stop()
