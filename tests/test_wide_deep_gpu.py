"""Wide-and-deep training step on the GPU vs a plain PyTorch fp32 restatement of the same math
(fused pull / MLP / loss / backward / fused push / ps-side applies); single rank and two ranks
emulated on one device (peers attached by pointer, host-ordered phases instead of device barriers)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cfg(**kw):
    from distributed_tensorflow_b200.models.wide_deep import WideDeepConfig
    base = dict(emb_rows=5000, slots=4, dense_in=5, dense_pad=64, h1=128, h2=64, h3=256, batch=300,
                optimizer="adagrad", lr=0.05, sync=True, emb_init_scale=0.05, seed=7)
    base.update(kw)
    return WideDeepConfig(**base)


def _batch(cfg, seed, dev):
    g = torch.Generator(device="cpu").manual_seed(seed)
    ids = torch.randint(0, cfg.emb_rows, (cfg.batch, cfg.slots), generator=g, dtype=torch.int32)
    ids[:, 0] = ids[0, 0]                       # force duplicates of one row inside the batch
    dense = torch.randn(cfg.batch, cfg.dense_in, generator=g)
    labels = (torch.rand(cfg.batch, generator=g) < 0.3).float()
    return ids.to(dev), dense.to(dev), labels.to(dev)


def _gather_params(trainers, cfg):
    """Global fp32 parameter set assembled from the ranks' shards."""
    world = len(trainers)
    dev = trainers[0].device
    emb = torch.zeros(cfg.emb_rows, cfg.emb_dim, device=dev)
    wide = torch.zeros(cfg.emb_rows, device=dev)
    flat = torch.zeros(trainers[0].info.P_pad, device=dev)
    for r, t in enumerate(trainers):
        n = t.info.rows_local
        emb[r::world] = t.buffer("emb_master", torch.float32)[: n * cfg.emb_dim].view(n, cfg.emb_dim)
        wide[r::world] = t.buffer("wide", torch.float32)[:n]
        flat[t.info.dense_lo: t.info.dense_hi] = t.buffer("dense_master", torch.float32)
    inf = trainers[0].info
    c = cfg
    dense = {"w1": flat[inf.off_w1: inf.off_w1 + c.h1 * inf.K0].view(c.h1, inf.K0).clone(),
             "w2": flat[inf.off_w2: inf.off_w2 + c.h2 * c.h1].view(c.h2, c.h1).clone(),
             "w3": flat[inf.off_w3: inf.off_w3 + c.h3 * c.h2].view(c.h3, c.h2).clone(),
             "b1": flat[inf.off_b1: inf.off_b1 + c.h1].clone(), "b2": flat[inf.off_b2: inf.off_b2 + c.h2].clone(),
             "b3": flat[inf.off_b3: inf.off_b3 + c.h3].clone(), "w4": flat[inf.off_w4: inf.off_w4 + c.h3].clone(),
             "b4": flat[inf.off_b4: inf.off_b4 + 1].clone(), "wide_b": flat[inf.off_wide_b: inf.off_wide_b + 1].clone()}
    return emb, wide, dense


def _close(a, b, rtol, atol, what):
    a, b = a.float(), b.float()
    err = (a - b).abs().max().item()
    scale = b.abs().max().item()
    assert err <= atol + rtol * scale, "%s: max err %.3e vs scale %.3e" % (what, err, scale)


def _close_frac(a, b, tol, max_frac, what):
    """Adam normalises every update to about +-lr, so gradients that are zero up to bf16 noise move
    a weight by a full step in either direction; compare by the fraction of outliers instead."""
    bad = ((a.float() - b.float()).abs() > tol).float().mean().item()
    assert bad <= max_frac, "%s: %.4f of the elements differ by more than %g" % (what, bad, tol)


def _make(cfg, world):
    from distributed_tensorflow_b200.models.wide_deep import WideDeepTrainer
    ts = [WideDeepTrainer(cfg, rank=r, world=world, no_device_barrier=True) for r in range(world)]
    if world > 1:
        for t in ts:
            t.attach_local(ts)
    for t in ts:
        t.init_params()
    torch.cuda.synchronize()
    return ts


def _run_sync_steps(ts, batches_per_step):
    for batches in batches_per_step:
        for t, b in zip(ts, batches):
            t.step_compute(*b)
        for t in ts:
            t.barrier(True)
        torch.cuda.synchronize()
        for t in ts:
            t.step_apply()
        torch.cuda.synchronize()


@pytest.mark.parametrize("opt", ["adagrad", "sgd", "adam"])
def test_single_rank_step_matches_torch(opt):
    from distributed_tensorflow_b200.models.wide_deep import TorchReference
    cfg = _cfg(optimizer=opt, lr=0.05 if opt != "adam" else 0.01)
    ts = _make(cfg, 1)
    t = ts[0]
    emb, wide, dense = _gather_params(ts, cfg)
    assert emb.abs().max() > 0 and dense["w1"].abs().max() > 0 and dense["b1"].abs().max() == 0
    ref = TorchReference(cfg, emb, wide, dense)
    b = _batch(cfg, 1, t.device)
    # forward pieces and gradients
    t.step_compute(*b)
    torch.cuda.synchronize()
    a, h1, h2, h3, logit = ref.forward(b[0], b[1])
    _close(t.buffer("a_buf", torch.bfloat16, (cfg.batch, cfg.k0)), a, 0, 1e-6, "gathered operand")
    _close(t.buffer("h1", torch.bfloat16, (cfg.batch, cfg.h1)), h1, 2e-2, 1e-3, "h1")
    _close(t.buffer("h3", torch.bfloat16, (cfg.batch, cfg.h3)), h3, 2e-2, 1e-3, "h3")
    loss_ref, g, da, dl = ref.grads(b[0], b[1], b[2], float(cfg.batch))
    assert abs(float(t.last_loss_sum()) - float(loss_ref)) < 1e-2 * float(loss_ref)
    _close(t.buffer("dlogit", torch.float32)[: cfg.batch], dl, 2e-2, 1e-6, "dlogit")
    gv = t.dense_views()
    for k in ("w1", "w2", "w3", "b1", "b2", "b3", "w4", "b4", "wide_b"):
        _close(gv["grad_" + k], g[k], 3e-2, 1e-6, "grad " + k)
    # pushed rows: every (example, slot) has a mailbox entry holding its gradient row
    n = cfg.batch * cfg.slots
    cnt = t.buffer("step_scratch", torch.int32)[:1]
    assert int(cnt[0]) == n
    packed = t.buffer("mbx_rows", torch.int64)[:n]                 # local row | example index << 32
    rows = packed & 0xFFFFFFFF
    assert torch.equal(torch.sort(packed >> 32).values, torch.arange(cfg.batch, device=t.device).repeat_interleave(cfg.slots))
    pay = t.buffer("mbx_payload", torch.bfloat16)[: n * cfg.emb_dim].view(n, cfg.emb_dim).float()
    got = torch.zeros(cfg.emb_rows, cfg.emb_dim, device=t.device).index_add_(0, rows, pay)
    want = torch.zeros(cfg.emb_rows, cfg.emb_dim, device=t.device).index_add_(0, b[0].long().reshape(-1), da.reshape(n, -1))
    _close(got, want, 3e-2, 1e-6, "pushed embedding gradients")
    # apply and compare the updated parameters after two full steps
    t.barrier(True)
    t.step_apply()
    ref.sync_step([b])
    b2 = _batch(cfg, 2, t.device)
    _run_sync_steps(ts, [[b2]])
    ref.sync_step([b2])
    emb2, wide2, dense2 = _gather_params(ts, cfg)
    if opt == "adam":
        _close_frac(emb2, ref.emb, 2e-3, 0.02, "embedding after 2 steps")
        _close_frac(wide2, ref.wide, 2e-3, 0.02, "wide after 2 steps")
        for k in dense2:
            _close_frac(dense2[k], ref.p[k], 2e-3, 0.05, "dense %s after 2 steps" % k)
        untouched = torch.ones(cfg.emb_rows, dtype=torch.bool, device=t.device)
        untouched[b[0].long().reshape(-1)] = False
        untouched[b2[0].long().reshape(-1)] = False
        assert torch.equal(emb2[untouched], emb[untouched])                    # lazy Adam: only touched rows move
    else:
        _close(emb2, ref.emb, 2e-2, 2e-4, "embedding after 2 steps")
        _close(wide2, ref.wide, 2e-2, 2e-4, "wide after 2 steps")
        for k in dense2:
            _close(dense2[k], ref.p[k], 2e-2, 3e-4, "dense %s after 2 steps" % k)
    assert t.global_step() == 2
    # working copies follow the masters
    _close(t.buffer("emb_bf16", torch.bfloat16)[: 64 * cfg.emb_dim], emb2[:64].reshape(-1), 1e-2, 1e-4, "bf16 copy")
    for x in ts:
        x.close()


def test_two_emulated_ranks_sync_equals_one_rank_on_the_global_batch():
    from distributed_tensorflow_b200.models.wide_deep import TorchReference
    cfg = _cfg(batch=256)
    ts = _make(cfg, 2)
    emb, wide, dense = _gather_params(ts, cfg)
    ref = TorchReference(cfg, emb, wide, dense)
    dev = ts[0].device
    steps = [[_batch(cfg, 10 + 2 * s + r, dev) for r in range(2)] for s in range(2)]
    _run_sync_steps(ts, steps)
    for s in steps:
        ref.sync_step(s)
    emb2, wide2, dense2 = _gather_params(ts, cfg)
    _close(emb2, ref.emb, 2e-2, 2e-4, "embedding")
    _close(wide2, ref.wide, 2e-2, 2e-4, "wide")
    for k in dense2:
        _close(dense2[k], ref.p[k], 2e-2, 3e-4, "dense " + k)
    # both ranks hold identical dense working copies (owners broadcast their ranges)
    w0 = ts[0].buffer("dense_bf16", torch.int16)
    w1 = ts[1].buffer("dense_bf16", torch.int16)
    assert torch.equal(w0, w1)
    for x in ts:
        x.close()


def test_async_hogwild_converges_and_predict_agrees():
    cfg = _cfg(sync=False, emb_rows=2000, lr=0.1)
    ts = _make(cfg, 1)
    t = ts[0]
    dev = t.device
    g = torch.Generator().manual_seed(0)
    w_true = torch.randn(cfg.dense_in, generator=g)
    losses = []
    for s in range(60):
        ids = torch.randint(0, cfg.emb_rows, (cfg.batch, cfg.slots), generator=g, dtype=torch.int32)
        dense = torch.randn(cfg.batch, cfg.dense_in, generator=g)
        labels = ((dense @ w_true) > 0).float()
        t.step(ids.to(dev), dense.to(dev), labels.to(dev))
        losses.append(float(t.last_loss_sum()) / cfg.batch)
    assert losses[-1] < 0.6 * losses[0], losses[::10]
    p = t.predict(ids.to(dev), dense.to(dev))[: cfg.batch].clone()
    acc = ((p > 0.5).float().cpu() == labels).float().mean().item()
    assert acc > 0.8
    t.close()


def test_async_serial_and_overlapped_paths_both_train_and_sync_refuses_overlap():
    from distributed_tensorflow_b200.ops import _lib
    g = torch.Generator().manual_seed(1)
    finals = {}
    for overlap in (False, True):
        cfg = _cfg(sync=False, emb_rows=2000, lr=0.1)
        t = _make(cfg, 1)[0]
        w_true = torch.tensor([1.0, -1.0, 0.5, 0.0, 0.25])
        gg = torch.Generator().manual_seed(3)
        losses = []
        for s in range(60):
            ids = torch.randint(0, cfg.emb_rows, (cfg.batch, cfg.slots), generator=gg, dtype=torch.int32)
            dense = torch.randn(cfg.batch, cfg.dense_in, generator=gg)
            labels = ((dense @ w_true) > 0).float()
            t.step(ids.to(t.device), dense.to(t.device), labels.to(t.device), overlap=overlap)
            losses.append(float(t.last_loss_sum()) / cfg.batch)
        t.join()
        torch.cuda.synchronize()
        assert losses[-1] < 0.7 * losses[0], (overlap, losses[::8])
        assert t.global_step() == 60
        finals[overlap] = losses[-1]
        t.close()
    assert abs(finals[True] - finals[False]) < 0.15         # one step of staleness does not change the outcome
    t = _make(_cfg(sync=True), 1)[0]
    b = _batch(t.cfg, 1, t.device)
    with pytest.raises(_lib.KernelError, match="sync-replica"):
        t.step(*b, overlap=True)
    t.step(*b)                                             # the plain path still works
    torch.cuda.synchronize()
    t.close()


@pytest.mark.parametrize("sync", [False, True])
def test_graphed_train_loop_trains_from_pinned_host_batches(sync):
    from distributed_tensorflow_b200.models.wide_deep_loop import GraphedTrainLoop
    cfg = _cfg(sync=sync, emb_rows=2000, lr=0.1)
    t = _make(cfg, 1)[0]
    g = torch.Generator().manual_seed(11)
    w_true = torch.randn(cfg.dense_in, generator=g)
    G = 4
    host = []
    for _ in range(G):
        ids = torch.randint(0, cfg.emb_rows, (cfg.batch, cfg.slots), generator=g, dtype=torch.int32).pin_memory()
        dense = torch.randn(cfg.batch, cfg.dense_in, generator=g).pin_memory()
        labels = ((dense @ w_true) > 0).float().pin_memory()
        host.append((ids, dense, labels))
    with pytest.raises(ValueError, match="pinned"):
        GraphedTrainLoop(t, [(host[0][0].clone(), host[0][1], host[0][2])])
    loop = GraphedTrainLoop(t, host, warmup=2)
    step0 = t.global_step()
    first = None
    for r in range(12):
        for ids, dense, labels in host:            # refill the pinned buffers in place between replays
            ids.copy_(torch.randint(0, cfg.emb_rows, ids.shape, generator=g, dtype=torch.int32))
            dense.copy_(torch.randn(dense.shape, generator=g))
            labels.copy_(((dense @ w_true) > 0).float())
        loop.run()
        ls = loop.losses()
        assert len(ls) == G and all(l == l and l > 0 for l in ls)
        first = first or ls[0]
    assert ls[-1] < 0.7 * first, (first, ls)
    assert t.global_step() == step0 + 12 * G
    assert loop.h2d_bytes_per_step == cfg.batch * (cfg.slots * 4 + cfg.dense_in * 4 + 4)
    # async: every step of a replay but the first runs layer 1 on rows prefetched during the step before (the
    # counter is kept by the host, so it sees the capture, not the replays)
    assert t.prefetched_steps() == (0 if sync else G - 1)
    t.close()


def test_async_parameter_prefetch_feeds_layer_1_the_rows_the_fused_pull_would_fetch():
    """``step(next_batch=...)``: the next step's rows are gathered on a low-priority stream and layer 1 runs as
    a plain GEMM on them.  With a zero learning rate the shards never change, so the prefetched operand, the
    wide values and the loss must equal those of the fused pull; a step on a batch other than the announced one
    drops the prefetched rows."""
    cfg = _cfg(sync=False, emb_rows=2000, lr=0.0, optimizer="sgd")
    seen = {}
    for pf in (False, True):
        t = _make(cfg, 1)[0]
        batches = [_batch(cfg, 100 + s, t.device) for s in range(5)]
        rec = []
        for s, b in enumerate(batches):
            nxt = batches[s + 1][:2] if (pf and s + 1 < len(batches)) else None
            t.step(*b, next_batch=nxt)
            t.join()
            torch.cuda.synchronize()
            par = int(t._L.dtf_wd_last_parity(t._h))
            a = t.buffer("a_buf1" if par else "a_buf", torch.bfloat16)[: cfg.batch * t.info.K0].view(cfg.batch, t.info.K0).float().clone()
            w = t.buffer("wide_vals1" if par else "wide_vals", torch.float32)[: cfg.batch * cfg.slots].clone()
            h1 = t.buffer("h1", torch.bfloat16)[: cfg.batch * cfg.h1].float().clone()
            rec.append((float(t.last_loss_sum()), a, w, h1))
        assert t.prefetched_steps() == (4 if pf else 0)
        seen[pf] = rec
        if pf:                                   # announce one batch, step on another: falls back to the fused pull
            t.step(*batches[0], next_batch=batches[1][:2])
            t.step(*batches[2])
            t.join()
            torch.cuda.synchronize()
            assert t.prefetched_steps() == 4
            assert abs(float(t.last_loss_sum()) - seen[False][2][0]) <= 1e-4 * abs(seen[False][2][0])
            # a small predict between the announcement and the step writes this parity's wide values: the pending
            # gather is waited for and dropped, and the step pulls again
            t.step(*batches[0], next_batch=batches[1][:2])
            pr = t.predict(batches[3][0], batches[3][1], n=2).clone()
            t.step(*batches[1])
            t.join()
            torch.cuda.synchronize()
            assert t.prefetched_steps() == 4 and bool(((pr > 0) & (pr < 1)).all())
            assert abs(float(t.last_loss_sum()) - seen[False][1][0]) <= 1e-4 * abs(seen[False][1][0])
        t.close()
    for s, ((l0, a0, w0, h0), (l1, a1, w1, h1)) in enumerate(zip(seen[False], seen[True])):
        assert torch.equal(a0, a1), "step %d: prefetched operand differs from the fused pull's" % s
        assert torch.equal(w0, w1), "step %d: wide values differ" % s
        _close(h1, h0, 1e-3, 1e-3, "layer-1 output, step %d" % s)
        assert abs(l0 - l1) <= 1e-4 * abs(l0), (s, l0, l1)


@pytest.mark.parametrize("opt", ["adadelta", "ftrl", "rmsprop"])
def test_remaining_reference_optimizers_apply_the_host_checked_rule(opt):
    """`--optimizer` accepts six names in the reference (trainer/task.py:41-56).  The three that the first
    round left out run the same per-element rule as tests/test_optimizers.py checks on the host
    (csrc/kernels/wd_optim.cuh): one ps-side apply on the GPU must reproduce that rule on the dense shard
    (master weights and both slots), move the touched embedding rows only, and keep training finite."""
    import ctypes as C
    from distributed_tensorflow_b200.models.wide_deep import OPT_IDS
    from distributed_tensorflow_b200.ops import _lib
    cfg = _cfg(optimizer=opt, lr=0.05)
    ts = _make(cfg, 1)
    t = ts[0]
    b = _batch(cfg, 3, t.device)
    names = t._SLOT_NAMES[opt]
    assert len(names) == 2 and all(("dense_flat/part_0/" + n) in t.named_shards() for n in names)
    w0 = t.buffer("dense_master", torch.float32).clone()
    s1_0 = t.buffer("dense_slot1", torch.float32).clone()
    s2_0 = t.buffer("dense_slot2", torch.float32).clone()
    emb0 = t.buffer("emb_master", torch.float32).clone()
    t.step_compute(*b)
    t.barrier(True)
    torch.cuda.synchronize()
    n = w0.numel()
    g = t.buffer("dense_grad", torch.float32)[t.info.dense_lo: t.info.dense_hi].clone()
    t.step_apply()
    torch.cuda.synchronize()
    L = _lib.lib()
    L.dtf_opt_update_host.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_float,
                                      C.c_longlong]
    hw, hg, h1, h2 = (x.cpu().numpy().copy() for x in (w0, g, s1_0, s2_0))
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    assert L.dtf_opt_update_host(OPT_IDS[opt], p(hw), p(hg), p(h1), p(h2), n, cfg.lr, 1) == 0
    _close(t.buffer("dense_master", torch.float32), torch.from_numpy(hw).to(t.device), 1e-4, 1e-6, "dense weights")
    _close(t.buffer("dense_slot1", torch.float32), torch.from_numpy(h1).to(t.device), 1e-4, 1e-7, "dense slot 1")
    _close(t.buffer("dense_slot2", torch.float32), torch.from_numpy(h2).to(t.device), 1e-4, 1e-7, "dense slot 2")
    emb1 = t.buffer("emb_master", torch.float32)
    nrow = t.info.rows_local
    moved = (emb1[: nrow * cfg.emb_dim].view(nrow, -1) != emb0[: nrow * cfg.emb_dim].view(nrow, -1)).any(1)
    touched = torch.zeros(nrow, dtype=torch.bool, device=t.device)
    touched[b[0].long().reshape(-1)] = True
    assert moved.any() and not (moved & ~touched).any()                        # only rows of this batch's ids move
    for s in range(4, 10):                                                      # keeps training without blowing up
        _run_sync_steps(ts, [[_batch(cfg, s, t.device)]])
    assert np.isfinite(float(t.last_loss_sum())) and torch.isfinite(t.buffer("dense_master", torch.float32)).all()
    assert t.global_step() == 7
    t.close()
