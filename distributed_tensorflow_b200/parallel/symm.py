"""Host control plane of the symmetric heap: passing each rank's slab file descriptor to every peer.

A VMM allocation is shared as a POSIX file descriptor (``cuMemExportToShareableHandle``); descriptors
only travel between processes over Unix-domain sockets (``SCM_RIGHTS``).  Every rank listens on
``$TMPDIR/dtf_symm_<tag>_<rank>.sock`` and serves its descriptor to whoever connects; after a
barrier each rank fetches the descriptors of all peers.  This happens once at start-up; the data
plane afterwards is device code only (``csrc/kernels/symm.cu``).
"""
from __future__ import annotations

import os
import socket
import tempfile
import threading
from typing import List, Optional


def _path(tag: str, rank: int) -> str:
    return os.path.join(tempfile.gettempdir(), "dtf_symm_%s_%d.sock" % (tag, rank))


def exchange_fds(my_fd: int, rank: int, world: int, dist) -> List[Optional[int]]:
    """Collective: returns a list of length ``world`` with the received descriptors (own entry None)."""
    tag_box = [("%s_%d" % (os.environ.get("MASTER_PORT", "0"), os.getpid())) if rank == 0 else None]
    dist.broadcast_object_list(tag_box, src=0)
    tag = tag_box[0]
    path = _path(tag, rank)
    if os.path.exists(path):
        os.unlink(path)
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(path)
    srv.listen(world)

    def serve():
        for _ in range(world - 1):
            conn, _ = srv.accept()
            with conn:
                conn.recv(4)
                socket.send_fds(conn, [b"fd"], [my_fd])
    t = threading.Thread(target=serve, daemon=True)
    t.start()
    dist.barrier()
    out: List[Optional[int]] = [None] * world
    for peer in range(world):
        if peer == rank:
            continue
        c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        c.connect(_path(tag, peer))
        c.sendall(rank.to_bytes(4, "little"))
        _, fds, _, _ = socket.recv_fds(c, 16, 1)
        c.close()
        if not fds:
            raise RuntimeError("rank %d sent no descriptor" % peer)
        out[peer] = fds[0]
    t.join(30)
    dist.barrier()
    srv.close()
    try:
        os.unlink(path)
    except OSError:
        pass
    return out


def close_fds(fds) -> None:
    for fd in fds:
        try:
            os.close(fd)
        except OSError:
            pass
