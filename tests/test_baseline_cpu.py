"""The PyTorch baseline (baseline/wide_deep_torch.py) on the CPU: its fixed-capacity routing, sort/segment
sparse Adagrad and flat-buffer dense Adagrad against a dense single-process autograd + torch.optim.Adagrad
restatement, with one rank and with two gloo ranks.  (The timed runs use NCCL + CUDA graphs on the GPU.)"""
import os
import subprocess
import sys

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
KW = dict(slots=3, dense_in=5, dense_pad=8, hidden=(16, 8), emb_dim=4, compute_dtype=torch.float32, lr=0.05)


def _batch(seed, rows, batch):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(0, rows, (batch, 3), generator=g, dtype=torch.int32)
    ids[:, 0] = ids[0, 0]                                  # duplicates inside the batch
    dense = torch.randn(batch, 5, generator=g)
    labels = (torch.rand(batch, generator=g) < 0.4).float()
    return ids, dense, labels


def _dense_reference(emb, wide, master, batches, lr=0.05):
    """Plain autograd over full tables + torch.optim.Adagrad (zero gradients leave untouched rows alone)."""
    emb, wide, flat = (t.clone().requires_grad_(True) for t in (emb, wide, master))
    opt = torch.optim.Adagrad([emb, wide, flat], lr=lr, initial_accumulator_value=0.1)
    shapes = [(16, 3 * 4 + 8), (16,), (8, 16), (8,), (1, 8), (1,), (1,)]
    for ids, dense, labels in batches:
        o, v = 0, []
        for s in shapes:
            n = 1
            for d in s:
                n *= d
            v.append(flat[o:o + n].view(s))
            o += n
        B = ids.shape[0]
        x = torch.zeros(B, 8)
        x[:, :5] = dense
        h = torch.cat([emb[ids.long()].reshape(B, -1), x], 1)
        h = torch.relu(torch.nn.functional.linear(h, v[0], v[1]))
        h = torch.relu(torch.nn.functional.linear(h, v[2], v[3]))
        logit = torch.nn.functional.linear(h, v[4], v[5]).squeeze(1) + wide[ids.long()].sum(1) + v[6]
        loss = torch.nn.functional.binary_cross_entropy_with_logits(logit, labels, reduction="sum") / B
        opt.zero_grad()
        loss.backward()
        opt.step()
    return emb.detach(), wide.detach(), flat.detach()


def test_single_rank_matches_dense_autograd_adagrad():
    from baseline.wide_deep_torch import TorchWideDeep
    rows, B = 200, 16
    m = TorchWideDeep(rows, B, 0, 1, torch.device("cpu"), **KW)
    e0, w0, f0 = m.emb[:rows].clone(), m.wide[:rows].clone(), m.master.clone()
    batches = [_batch(s, rows, B) for s in range(3)]
    for b in batches:
        m.step(*b)
    e1, w1, f1 = _dense_reference(e0, w0, f0, batches)
    assert torch.allclose(m.emb[:rows], e1, atol=1e-6) and not torch.equal(e1, e0)
    assert torch.allclose(m.wide[:rows], w1, atol=1e-6) and torch.allclose(m.master, f1, atol=1e-6)
    assert int(m.overflow) == 0 and torch.equal(m.emb[rows], torch.zeros(4))      # the dummy row never moves


def test_two_gloo_ranks_sync_equal_one_rank_on_the_global_batch(tmp_path):
    script = tmp_path / "b.py"
    script.write_text('''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, %r)
sys.path.insert(0, %r)
from baseline.wide_deep_torch import TorchWideDeep
from test_baseline_cpu import KW, _batch, _dense_reference
r, w = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=r, world_size=w)
rows, B = 201, 16
m = TorchWideDeep(rows, B, r, w, torch.device("cpu"), mode="sync", capacity=42, **KW)   # 42 < 48: real padding / discard slots
def gather():
    out = [None] * w
    dist.all_gather_object(out, (m.emb[: m.rows_local].clone(), m.wide[: m.rows_local].clone()))
    emb, wide = torch.zeros(rows, 4), torch.zeros(rows)
    for k, (e, x) in enumerate(out):
        emb[k::w], wide[k::w] = e, x
    return emb, wide
e0, w0 = gather()
f0 = m.master.clone()
for s in range(3):
    m.step(*_batch(10 * s + r, rows, B))
e1, w1 = gather()
if r == 0:
    batches = [tuple(torch.cat([_batch(10 * s + k, rows, B)[i] for k in range(w)]) for i in range(3)) for s in range(3)]
    er, wr, fr = _dense_reference(e0, w0, f0, batches)
    assert torch.allclose(e1, er, atol=1e-6), (e1 - er).abs().max()
    assert torch.allclose(w1, wr, atol=1e-6) and torch.allclose(m.master, fr, atol=1e-6)
    assert m.cap == 42 and int(m.overflow) == 0
    print("BASELINE_OK")
dist.destroy_process_group()
''' % (REPO, os.path.join(REPO, "tests")))
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", str(port), str(script)], capture_output=True, text=True, timeout=300, cwd=REPO)
    assert p.returncode == 0 and "BASELINE_OK" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]
