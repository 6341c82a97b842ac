"""Host control plane of the symmetric heap: file descriptors passed between ranks over Unix
sockets (parallel/symm.py).  Runs two real processes with a gloo group on the CPU; the descriptors
are temporary files standing in for exported VMM handles."""
import os
import subprocess
import sys
import textwrap

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.timeout(120)

WORKER = textwrap.dedent("""
    import os, sys
    sys.path.insert(0, %r)
    import torch.distributed as dist
    from distributed_tensorflow_b200.parallel import symm
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import tempfile
    f = tempfile.TemporaryFile()
    f.write(b"hello from rank %%d" %% rank); f.flush()
    r = f.fileno()
    fds = symm.exchange_fds(r, rank, world, dist)          # every rank offers a descriptor of its own file
    assert fds[rank] is None and all(f is not None for i, f in enumerate(fds) if i != rank)
    for peer, fd in enumerate(fds):
        if fd is not None:
            got = os.pread(fd, 64, 0)
            assert got == b"hello from rank %%d" %% peer, (peer, got)
    symm.close_fds([x for x in fds if x is not None])
    dist.barrier()
    dist.destroy_process_group()
    print("rank", rank, "ok")
""") % REPO


@pytest.mark.parametrize("world", [2, 3])
def test_fds_reach_every_peer(tmp_path, world):
    script = tmp_path / "w.py"
    script.write_text(WORKER)
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                   CUDA_VISIBLE_DEVICES="")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    for r, p in enumerate(procs):
        out, _ = p.communicate(timeout=100)
        assert p.returncode == 0 and ("rank %d ok" % r) in out, out
