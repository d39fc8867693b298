"""GPU: K4/K5 (device replay: Philox sampling, row gather, ring append) bit-exact vs the oracle, and the
ImpalaLearner / PPOLearner drop-ins vs a reference-shaped torch step on the CPU."""
import copy
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from gpu_util import assert_close
from impala_b200 import losses, models, ops, optim, replay, synth
from impala_b200.agents import ImpalaLearner, PPOLearner
from oracle import c_oracle, losses_ref, replay_ref

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _no_tf32():
    """cuDNN convolutions default to TF32; the CPU reference step is fp32."""
    old = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


def _check_step(learner, model, ref_model):
    """Clipped gradients (what Adam consumed) and parameters after one step.  Adam's first step moves every
    weight by ~lr * sign(g), so parameters are compared to a quarter of lr = 1e-4."""
    ref_g = torch.cat([q.grad.reshape(-1) for q in ref_model.parameters()])
    assert_close(learner._bucket.grads, ref_g, "clipped grads", rtol=1e-3, atol=1e-6, rel_to_max=2e-4)
    for (n1, p), (_, q) in zip(model.named_parameters(), ref_model.named_parameters()):
        assert_close(p, q, n1, rtol=0, atol=2.5e-5)


@pytest.mark.parametrize("seed,stream,off,n,size", [(123, 0, 0, 4096, 1_000_000), (2**40 + 7, 5, 2**33, 1000, 3),
                                                    (0, 0, 0, 17, 1), (9, 1, 10**12, 100_000, 2**40), (1, 0, 0, 0, 5)])
def test_sampler_bit_exact(cuda, seed, stream, off, n, size):
    got = ops.sample_uniform(seed, stream, off, n, size, cuda).cpu().numpy()
    assert np.array_equal(got, replay_ref.sample_uniform(seed, stream, off, n, size))
    assert np.array_equal(got, c_oracle.sample_uniform(seed, stream, off, n, size))


def test_sampler_golden_and_stream_continuity(cuda):
    z = np.load(os.path.join(GOLDEN, "replay_indices.npz"))
    s = replay.UniformSampler()
    s.reset(123)
    a = torch.cat([s.sample(1000, 1_000_000, cuda), s.sample(3096, 1_000_000, cuda)]).cpu().numpy()
    assert np.array_equal(a, z["idx_seed123_1M"])
    s3 = replay.UniformSampler(stream_id=3)
    s3.reset(123)
    s3.draws = 10_000_000_000
    assert np.array_equal(s3.sample(257, 1000, cuda).cpu().numpy(), z["idx_stream3_big_offset"])
    with pytest.raises(RuntimeError):
        s.sample(4, 0, cuda)


@pytest.mark.parametrize("shape,dtype", [((3, 64, 64), torch.uint8), ((1,), torch.int64), ((1,), torch.float32),
                                         ((15,), torch.float32), ((7,), torch.uint8), ((3,), torch.int16),
                                         ((20, 3, 64, 64), torch.uint8)])
def test_gather_bit_exact(cuda, shape, dtype):
    nrows, n = 257, 600
    g = torch.Generator().manual_seed(1)
    src = (torch.randint(0, 255, (nrows,) + shape, generator=g).to(dtype) if dtype != torch.float32
           else torch.randn((nrows,) + shape, generator=g))
    idx = torch.from_numpy(replay_ref.sample_uniform(7, 0, 0, n, nrows))
    got = ops.gather_rows(src.cuda(), idx.cuda()).cpu()
    assert torch.equal(got, src.index_select(0, idx))
    assert np.array_equal(got.numpy(), c_oracle.gather_rows(src.numpy(), idx.numpy()))
    assert ops.gather_rows(src.cuda(), idx[:0].cuda()).shape == (0,) + shape       # empty batch
    # unaligned base pointer (odd element offset) -> narrower vector path, same bytes
    flat = src.cuda().reshape(nrows, -1)
    if flat.shape[1] > 1:
        sub = flat[:, 1:].contiguous()
        assert torch.equal(ops.gather_rows(sub, idx.cuda()).cpu(), sub.cpu().index_select(0, idx))


def test_ring_buffer_matches_oracle(cuda):
    cap = 37
    rb = replay.ReplayBuffer(replay.CircularBuffer(cap, device=cuda), replay.UniformSampler())
    rb.sampler.reset(11)
    ob = replay_ref.ReplayBuffer(replay_ref.CircularBuffer(cap), replay_ref.UniformSampler())
    ob.sampler.reset(11)
    g = torch.Generator().manual_seed(0)
    with pytest.raises(RuntimeError):
        rb.warm_up(5)
    for step, n in enumerate([5, 1, 20, 30, 37, 3, 50]):          # wraps, exact fill, n > capacity
        obs = torch.randint(0, 255, (n, 3, 8, 8), generator=g, dtype=torch.uint8)
        act = torch.randint(0, 15, (n, 1), generator=g)
        rew = torch.randn(n, 1, generator=g)
        if step % 2 == 0:
            rb.extend((obs, act, rew))
        else:
            rb.extend([[obs[i], act[i], rew[i]] for i in range(n)])
        if n <= cap:
            ob.extend([obs.numpy(), act.numpy(), rew.numpy()])
        else:                                                       # oracle: only the survivors matter
            ob.storage.count += n - cap
            ob.extend([obs[n - cap:].numpy(), act[n - cap:].numpy(), rew[n - cap:].numpy()])
        assert rb.size == ob.storage.size and rb.info() == (cap, ob.storage.size)
        for f, of in zip(rb.storage.fields, ob.storage.fields):
            occupied = min(rb.storage.count, cap)
            assert np.array_equal(f.cpu().numpy()[:occupied] if rb.storage.count < cap else f.cpu().numpy(),
                                  of[:occupied] if rb.storage.count < cap else of)
        keys, batch, probs = rb.sample(64)
        okeys, obatch, oprobs = ob.sample(64)
        assert np.array_equal(keys.cpu().numpy(), okeys)
        assert all(np.array_equal(b.cpu().numpy(), o) for b, o in zip(batch, obatch))
        assert np.allclose(probs.cpu().numpy(), oprobs)
    rb.append([torch.zeros(3, 8, 8, dtype=torch.uint8), torch.zeros(1, dtype=torch.int64), torch.zeros(1)])
    rb.warm_up(10)


def test_config4_shape_gather_roundtrip(cuda):
    """BASELINE config 4 row shape (64x64x3 u8), capacity scaled to 20k rows: rows are tagged with their
    index, so a gather is checked by content at any size."""
    cap, n = 20_000, 4096
    store = torch.empty(cap, 3, 64, 64, dtype=torch.uint8, device=cuda)
    tag = torch.arange(cap, device=cuda)
    store.view(cap, -1)[:, 0] = (tag % 251).to(torch.uint8)
    store.view(cap, -1)[:, 1] = ((tag // 251) % 251).to(torch.uint8)
    store.view(cap, -1)[:, -1] = (tag % 13).to(torch.uint8)
    idx = ops.sample_uniform(123, 0, 0, n, cap, cuda)
    out = ops.gather_rows(store, idx).view(n, -1)
    assert torch.equal(out[:, 0].long() + 251 * out[:, 1].long(), idx % (251 * 251))
    assert torch.equal(out[:, -1].long(), idx % 13)
    assert torch.equal(out, store.view(cap, -1).index_select(0, idx))


def test_config4_full_size_gather_and_append(cuda):
    """BASELINE config 4 at FULL size: 1,000,000 rows x 12,288 B = 12.29 GB ring, batch 4,096, seed 123.  Byte offsets
    pass 2^32 from row 349,526 on, which only this size exercises.  Size-independent checks: every gathered row
    carries its own index tag; the whole batch equals torch's index_select; appending the batch back (ring wrap
    included) and re-gathering is the identity; the index stream equals the committed golden stream."""
    import numpy as np
    from conftest import GOLDEN
    cap, row, n = 1_000_000, 3 * 64 * 64, 4096
    free, _ = torch.cuda.mem_get_info()
    if free < 14 * 2**30:
        pytest.skip("needs 14 GB of free HBM")
    store = torch.empty(cap, row, dtype=torch.uint8, device=cuda)
    store.view(torch.int32).random_()                                   # arbitrary bytes everywhere
    store[:, :8] = torch.arange(cap, device=cuda).view(cap, 1).view(torch.uint8).view(cap, 8)   # little-endian tag
    idx = ops.sample_uniform(123, 0, 0, n, cap, cuda)
    z = np.load(os.path.join(GOLDEN, "replay_indices.npz"))
    assert np.array_equal(idx.cpu().numpy(), z["idx_seed123_1M"])
    assert (idx > 349_526).any() and int(idx.max()) < cap
    out = ops.gather_rows(store, idx)
    assert torch.equal(out[:, :8].contiguous().view(torch.int64).view(-1), idx)
    assert torch.equal(out, store.index_select(0, idx))
    # ring append of the batch across the wrap point: count = cap - 100 -> rows land in slots cap-100.., 0..n-101
    count = cap - 100
    ops.ring_append(store, count, out)
    slots = (torch.arange(n, device=cuda) + count) % cap
    back = ops.gather_rows(store, slots)
    assert torch.equal(back, out)
    del store, out, back
    torch.cuda.empty_cache()


def _fill_impala_replay(rb, n_items, T, A, seed):
    g = torch.Generator().manual_seed(seed)
    for _ in range(n_items):
        rb.append([torch.randint(0, 255, (T, 3, 64, 64), generator=g, dtype=torch.uint8),
                   torch.randint(0, A, (T, 1), generator=g), torch.randn(T, 1, generator=g),
                   (torch.rand(T, 1, generator=g) > .05).float() * .99, torch.randn(T, A, generator=g)])


@pytest.mark.parametrize("body,fused_opt", [("nature", False), ("impala", False), ("impala", True)])
def test_impala_learner_step_matches_reference_shaped_step(cuda, body, fused_opt):
    torch.manual_seed(0)
    T, A, Bsz = 20, 15, 8
    model = models.ActorCritic((3, 64, 64), A, body=body).to(cuda)
    ref_model = copy.deepcopy(model).cpu()
    rb = replay.ReplayBuffer(replay.CircularBuffer(32, device=cuda), replay.UniformSampler())
    rb.sampler.reset(123)
    _fill_impala_replay(rb, 12, T, A, seed=5)
    opt = (optim.FusedClipAdam if fused_opt else torch.optim.Adam)(model.parameters(), lr=1e-4, eps=1e-5)
    learner = ImpalaLearner(model, rb, opt, batch_size=Bsz, learning_starts=10)
    learner.prepare()
    draws0 = rb.sampler.draws
    metrics = learner.train_step()
    assert {"train/loss", "train/entropy", "train/td", "train/pg", "train/kl", "train/ratio", "train/grad_norm",
            "debug/total_time"} <= set(metrics)
    # the same step as the reference writes it (agents/impala/learning.py:140-177), on the CPU
    slots = replay_ref.sample_uniform(123, 0, draws0, Bsz, rb.size)
    s, a, r, d, pi_ref = [f.cpu()[slots] for f in rb.storage.fields]
    a, r, d = a.squeeze(-1), r.squeeze(-1), d.squeeze(-1)
    ropt = torch.optim.Adam(ref_model.parameters(), lr=1e-4, eps=1e-5)
    ropt.zero_grad(set_to_none=True)
    pi, values = ref_model(s.flatten(0, 1))
    loss, m = losses_ref.impala_train_loss(pi.reshape(Bsz, T, -1), values.reshape(Bsz, T), a, r, d, pi_ref)
    loss.backward()
    gn = torch.nn.utils.clip_grad_norm_(ref_model.parameters(), .5)
    ropt.step()
    for k in ("train/loss", "train/entropy", "train/td", "train/pg", "train/kl", "train/ratio"):
        assert_close(metrics[k], m[k], k, rtol=1e-4, atol=1e-5)       # cuDNN fp32 vs CPU convs upstream
    assert_close(metrics["train/grad_norm"], gn, "grad_norm", rtol=1e-3, atol=1e-5)
    _check_step(learner, model, ref_model)


def test_ppo_learner_step(cuda):
    torch.manual_seed(0)
    A, N = 15, 256
    model = models.ActorCritic((3, 64, 64), A).to(cuda)
    ref_model = copy.deepcopy(model).cpu()
    rb = replay.ReplayBuffer(replay.CircularBuffer(1000, collate_fn=torch.cat, device=cuda), replay.UniformSampler())
    rb.sampler.reset(7)
    g = torch.Generator().manual_seed(1)
    rb.extend((torch.randint(0, 255, (600, 3, 64, 64), generator=g, dtype=torch.uint8),
               torch.randint(0, A, (600,), generator=g), torch.randn(600, generator=g), torch.randn(600, A, generator=g)))
    learner = PPOLearner(model, rb, torch.optim.Adam(model.parameters(), lr=1e-4, eps=1e-5), batch_size=N)
    learner.prepare()
    metrics = learner.train_step()
    slots = replay_ref.sample_uniform(7, 0, 0, N, 600)
    s, a, vt, pi_ref = [f.cpu()[slots] for f in rb.storage.fields]
    ropt = torch.optim.Adam(ref_model.parameters(), lr=1e-4, eps=1e-5)
    loss, m = losses_ref.ppo_loss(ref_model, (s, a, vt, pi_ref), entropy_cost=.01)
    loss.backward()
    torch.nn.utils.clip_grad_norm_(ref_model.parameters(), .5)
    ropt.step()
    for k in ("train/loss", "train/entropy", "train/td", "train/pg", "train/target", "train/kl", "train/ratio"):
        assert_close(metrics[k], m[k], k, rtol=1e-4, atol=1e-5)
    _check_step(learner, model, ref_model)


def test_rollout_ingestion_fused_discount_and_multifield_gather(cuda):
    """N2: stacked actor outputs -> ring storage in one launch with discount = logical_not(done) * gamma
    (reference agents/impala/learning.py:77-80), wrap-around included; then the multi-field gather."""
    T, A, cap, gamma = 5, 6, 7, 0.99
    rb = replay.ReplayBuffer(replay.CircularBuffer(cap, device=cuda), replay.UniformSampler())
    rb.sampler.reset(11)
    g = torch.Generator().manual_seed(3)
    want = None
    total = 0
    for n in (3, 2, 4, 7, 1):                                   # crosses the end of the ring several times
        s = torch.randint(0, 256, (n, T, 3, 8, 8), generator=g, dtype=torch.uint8)
        a = torch.randint(0, A, (n, T, 1), generator=g)
        r = torch.randn(n, T, 1, generator=g)
        d = torch.rand(n, T, 1, generator=g) < 0.3
        pi = torch.randn(n, T, A, generator=g)
        keys = rb.extend_rollouts(s, a, r, d, pi, gamma)
        assert list(keys) == list(range(total, total + n))
        ref_fields = [s, a, r, torch.logical_not(d) * gamma, pi]          # the reference's _make_replay
        if want is None:
            want = [torch.zeros((cap,) + tuple(f.shape[1:]), dtype=f.dtype) for f in ref_fields]
        for j in range(n):
            for w, f in zip(want, ref_fields):
                w[(total + j) % cap] = f[j]
        total += n
        occupied = min(total, cap)
        for f, w in zip(rb.storage.fields, want):
            assert f.dtype == w.dtype
            got = f.cpu()
            if total < cap:
                assert torch.equal(got[:occupied], w[:occupied])
            else:
                assert torch.equal(got, w)
    keys, batch, probs = rb.sample(33)
    slots = torch.from_numpy(replay_ref.sample_uniform(11, 0, 0, 33, cap).astype(np.int64))
    for b, w in zip(batch, want):
        assert torch.equal(b.cpu(), w[slots])
    # per-field launches and the fused one agree
    singles = [ops.gather_rows(f, slots.to(cuda)) for f in rb.storage.fields]
    for b, s1 in zip(batch, singles):
        assert torch.equal(b, s1)


def test_sumtree_bit_exact_vs_spec(cuda):
    """N3: device sum tree + inverse-CDF sampling vs the specification in oracle/replay_ref.py (bit-exact:
    tree nodes, sampled indices and probabilities), incremental and bulk updates."""
    rng = np.random.default_rng(0)
    # leaves 1..2^20: one subtree CTA, 2, 16 and 128 subtree CTAs (with and without level-by-level part)
    for cap in (1, 5, 600, 1000, 5000, 40_000, 1_000_000):
        ot = replay_ref.SumTree(cap)
        tree = torch.zeros(2 * ot.leaves, dtype=torch.float64, device=cuda)
        for n in (min(cap, 7), min(cap, 333), min(cap, 4096), min(cap, 8192), cap):   # n > 8192 takes the bulk path
            idx = rng.choice(cap, size=n, replace=False).astype(np.int64)
            val = rng.random(n) ** 3 + 1e-3
            if n == 4096:                                             # duplicates (same priority, as the learner's)
                idx[100:200] = idx[:100]
                val[100:200] = val[:100]
            ot.update(idx, val)
            ops.sumtree_update(tree, torch.from_numpy(idx).to(cuda), torch.from_numpy(val).to(cuda))
            assert np.array_equal(tree.cpu().numpy(), ot.tree)
        size = cap
        want_idx, want_p = ot.sample(123, 2, 10**9, 4096, size)
        got_idx, got_p = ops.sumtree_sample(tree, size, 123, 2, 10**9, 4096)
        assert np.array_equal(got_idx.cpu().numpy(), want_idx)
        assert np.array_equal(got_p.cpu().numpy(), want_p)
    # proportionality: slot frequencies follow the priorities
    tree = torch.zeros(16, dtype=torch.float64, device=cuda)
    pri = torch.tensor([1., 2., 3., 4., 0., 10.], dtype=torch.float64, device=cuda)
    ops.sumtree_update(tree, torch.arange(6, device=cuda), pri)
    idx, p = ops.sumtree_sample(tree, 6, 7, 0, 0, 200_000)
    freq = torch.bincount(idx, minlength=6).double() / idx.numel()
    assert torch.allclose(freq, pri / pri.sum(), atol=5e-3) and freq[4] == 0
    assert torch.allclose(p.double(), (pri / pri.sum())[idx], rtol=1e-6)


def test_prioritized_replay_buffer_surface(cuda):
    """rlmeta surface used by the Ape-X agents (reference agents/dqn/learning.py:54,122,184): extend with
    priorities, sample -> (keys, batch, probabilities), update(keys, priorities) ignoring evicted keys."""
    cap = 8
    rb = replay.ReplayBuffer(replay.CircularBuffer(cap, collate_fn=torch.cat, device=cuda),
                             replay.PrioritizedSampler(priority_exponent=0.6))
    rb.sampler.reset(5)
    obs = torch.arange(12, dtype=torch.float32).reshape(12, 1)
    keys = rb.extend((obs[:6], torch.arange(6)), priorities=torch.arange(1, 7, dtype=torch.float32))
    assert list(keys) == list(range(6))
    ot = replay_ref.SumTree(cap)
    vals = rb.sampler.tree[rb.sampler.tree.numel() // 2:][:6].cpu().numpy()          # device pow(alpha), as stored
    assert np.allclose(vals, np.arange(1, 7, dtype=np.float64) ** 0.6, rtol=1e-14)
    ot.update(np.arange(6), vals)
    k, batch, probs = rb.sample(64)
    want_idx, want_p = ot.sample(5, 0, 0, 64, 6)
    assert np.array_equal(k.cpu().numpy(), want_idx) and np.array_equal(probs.cpu().numpy(), want_p)
    assert torch.equal(batch[0].cpu(), obs[:6][torch.from_numpy(want_idx)])
    rb.extend((obs[6:], torch.arange(6, 12)))                                         # wraps: keys 0..3 evicted
    rb.update(torch.tensor([0, 3, 4, 11]), torch.tensor([100., 100., 2., 3.]))        # 0 and 3 are stale
    leaves = rb.sampler.tree[rb.sampler.tree.numel() // 2:].cpu().numpy()
    assert np.allclose(leaves[4], 2. ** 0.6) and np.allclose(leaves[11 % cap], 3. ** 0.6)
    assert np.allclose(leaves[0], 1.0) and np.allclose(leaves[1], 1.0)                # new rows entered at priority 1; stale keys 0, 3 ignored
    assert np.isclose(rb.sampler.tree[1].item(), leaves.sum())


def test_act_epilogue_matches_spec(cuda):
    """N5: fused softmax -> categorical draw -> argmax select (reference models/distributed_models.py:26-31).
    Deterministic rows are exact argmax; sampled rows must fall in the CDF bracket of the specified uniform
    (up to fp32 rounding at bracket edges) and follow softmax(logits) in frequency."""
    g = torch.Generator().manual_seed(0)
    n, A = 4096, 15
    logits = 2 * torch.randn(n, A, generator=g)
    det = torch.rand(n, generator=g) < 0.3
    got = ops.act_epilogue(logits.to(cuda), det, 99, 0, 12345).cpu().reshape(-1)
    a, u, cdf = losses_ref.act_epilogue(logits, det.numpy(), 99, 0, 12345)
    assert torch.equal(got[det], logits.argmax(-1)[det])
    lo = np.concatenate([np.zeros((n, 1)), cdf[:, :-1]], 1)
    gi = got.numpy()
    ok = (lo[np.arange(n), gi] - 1e-5 <= u) & (u <= cdf[np.arange(n), gi] + 1e-5)
    assert ok[~det.numpy()].all()
    assert (gi == a).mean() > 0.999
    one = torch.tensor([[0., 1., 2., 0.5]]).repeat(100_000, 1).to(cuda)
    draws = ops.act_epilogue(one, torch.tensor([False]), 7, 0, 0).reshape(-1)
    freq = torch.bincount(draws, minlength=4).double() / draws.numel()
    assert torch.allclose(freq, one[0].double().softmax(-1), atol=5e-3)
    assert torch.equal(ops.act_epilogue(one[:8], torch.tensor([True]), 7, 0, 0).reshape(-1), torch.full((8,), 2, device=cuda))
    model = models.ActorCritic((3, 64, 64), 15).to(cuda)                              # the drop-in act()
    model.reset_act_rng(3)
    act, lg, v = model.act(torch.zeros(5, 3, 64, 64, dtype=torch.uint8), torch.tensor([False]))
    assert act.shape == (5, 1) and act.dtype == torch.int64 and lg.shape == (5, 15) and v.shape == (5, 1)


@pytest.mark.parametrize("shapes", [[(3, 64, 64)], [(20, 3, 8, 8), (20, 1), (20, 1), (20, 15)], [(1,), (5,)], [(7, 3)]])
def test_sample_gather_in_one_launch_matches_two_launches(cuda, shapes):
    """b200rl_sample_gather_fields (K4 + K5 fused) == b200rl_sample_uniform then b200rl_gather_fields: same Philox draws,
    same rows, bit for bit -- incl. rows that straddle CTAs, tiny rows (many rows per CTA) and odd row sizes."""
    cap, size, n = 997, 613, 777
    g = torch.Generator(device=cuda).manual_seed(5)
    dts = [torch.uint8, torch.int64, torch.float32, torch.float32]
    stores = [(torch.randint(0, 250, (cap,) + sh, generator=g, device=cuda).to(dts[i % 4])) for i, sh in enumerate(shapes)]
    for call in range(2):
        off = 1000 * call + 3
        idx = ops.sample_uniform(99, 2, off, n, size, cuda)
        want = ops.gather_fields(stores, idx)
        got_idx, got = ops.sample_gather_fields(stores, 99, 2, off, n, size)
        torch.cuda.synchronize()
        assert torch.equal(got_idx, idx)
        assert int(got_idx.max()) < size
        for a, b in zip(got, want):
            assert torch.equal(a, b)
    # and through the drop-in buffer: keys / batch of the fused path == oracle sampler stream
    rb = replay.ReplayBuffer(replay.CircularBuffer(64, device=cuda), replay.UniformSampler(stream_id=1))
    rb.sampler.reset(7)
    rb.extend((torch.arange(40, device=cuda).reshape(40, 1).float(), torch.arange(40, device=cuda).reshape(40, 1)))
    keys, batch, probs = rb.sample(16)
    want_idx = ops.sample_uniform(7, 1, 0, 16, 40, cuda)
    assert torch.equal(keys, want_idx) and torch.equal(batch[1].reshape(-1), want_idx)
    keys2, _, _ = rb.sample(16)
    assert torch.equal(keys2, ops.sample_uniform(7, 1, 16, 16, 40, cuda))


@pytest.mark.parametrize("body,kw", [("impala", dict(channels_last=True, fused_glue=True)), ("impala", {})])
def test_learner_step_as_one_cuda_graph_matches_eager_steps(cuda, body, kw):
    """ImpalaLearner.capture_graph: the whole step (sample -> CNN -> K1 -> CNN bwd -> clip + Adam) replayed as ONE graph
    launch follows the eager learner step for step: the sampler's draw count and Adam's step live in device memory, so
    every replay draws new rollouts and uses the next bias correction."""
    T, A, Bsz, steps = 20, 15, 8, 5

    def make():
        torch.manual_seed(0)
        model = models.ActorCritic((3, 64, 64), A, body=body, **kw).to(cuda)
        if kw.get("channels_last"):
            model = model.to(memory_format=torch.channels_last)
        rb = replay.ReplayBuffer(replay.CircularBuffer(32, device=cuda), replay.UniformSampler())
        rb.sampler.reset(123)
        _fill_impala_replay(rb, 32, T, A, seed=5)
        # eps = 1e-2: with Adam's usual eps a gradient element at rounding-noise level still moves its parameter by +-lr
        # (m / sqrt(v) = +-1), so two runs whose convolutions round differently (cuDNN picks its algorithm per stream and
        # workspace) could not be compared elementwise; a large eps makes the update proportional to the gradient
        opt = optim.FusedClipAdam(model.parameters(), lr=1e-3, eps=1e-2)
        learner = ImpalaLearner(model, rb, opt, batch_size=Bsz, learning_starts=10, model_push_period=2)
        learner.prepare()
        model.pushes = []
        model.push = lambda: model.pushes.append(learner._step_count)       # the reference's RemotableModel.push
        return model, rb, opt, learner

    model_e, rb_e, opt_e, eager = make()
    want = []
    for _ in range(steps):
        m = eager.train_step()
        want.append(({k: float(m[k]) for k in ("train/loss", "train/td", "train/grad_norm")}, opt_e.flat.clone()))
    model_g, rb_g, opt_g, graphed = make()
    graphed.capture_graph(warmup=2)                     # two real (eager) steps, then the capture (which executes nothing)
    torch.cuda.synchronize()
    assert_close(opt_g.flat, want[1][1], "parameters after the 2 warm-up steps", rtol=1e-4, atol=2e-5)
    losses_seen = []
    for k in range(2, steps):
        m = graphed.train_step()                        # ONE graph launch
        torch.cuda.synchronize()
        losses_seen.append(float(m["train/loss"]))
        # the first replay starts from (numerically) the same state as eager step 3: tight.  Later steps: two TF32 training
        # runs drift apart step by step, so only their agreement to a few percent is meaningful; what the graph must get
        # right -- a NEW batch and the NEXT Adam step on every replay -- is asserted directly below and, bit for bit, in
        # tests/test_gpu_optim.py::test_device_resident_step_counter... and test_sample_gather_in_one_launch...
        rel, atol = (2e-3, 2e-5) if k == 2 else (5e-2, 5e-4)
        for name, v in want[k][0].items():
            assert float(m[name]) == pytest.approx(v, rel=rel, abs=1e-6), (k, name)
        assert_close(opt_g.flat, want[k][1], "parameters after step %d" % (k + 1), rtol=1e-3, atol=atol)
    assert len(set(losses_seen)) == len(losses_seen), "every replay must draw a new batch"
    assert opt_g.sync_step_count() == steps
    assert model_g.pushes == model_e.pushes == [2, 4]                       # warm-up steps count; replays push too
    assert set(m) >= {"debug/replay_sample_per_second", "debug/gradient_per_second", "debug/total_time", "debug/forward_dt",
                      "debug/update_time"}
    assert int(rb_g.sampler.draw_counter.item()) == steps * Bsz == rb_e.sampler.draws


def test_ppo_learner_step_as_one_cuda_graph_matches_eager_steps(cuda):
    """PPOLearner.capture_graph (agents/_graph.py): sample -> model -> K3 -> backward -> clip + Adam as ONE graph launch per
    step follows the eager learner (same seeds): new minibatch and next Adam step on every replay."""
    A, N, steps = 15, 256, 5

    def make():
        torch.manual_seed(0)
        model = models.ActorCritic((3, 64, 64), A).to(cuda)
        rb = replay.ReplayBuffer(replay.CircularBuffer(1000, collate_fn=torch.cat, device=cuda), replay.UniformSampler())
        rb.sampler.reset(7)
        g = torch.Generator().manual_seed(1)
        rb.extend((torch.randint(0, 255, (600, 3, 64, 64), generator=g, dtype=torch.uint8),
                   torch.randint(0, A, (600,), generator=g), torch.randn(600, generator=g), torch.randn(600, A, generator=g)))
        opt = optim.FusedClipAdam(model.parameters(), lr=1e-3, eps=1e-2)      # eps: see the IMPALA graph test above
        learner = PPOLearner(model, rb, opt, batch_size=N, model_push_period=2)
        learner.prepare()
        model.pushes = []
        model.push = lambda: model.pushes.append(learner._step_counter)
        return model, rb, opt, learner

    model_e, rb_e, opt_e, eager = make()
    want = []
    for _ in range(steps):
        m = eager.train_step()
        want.append(({k: float(m[k]) for k in ("train/loss", "train/td", "train/target", "train_step/grad_norm")}, opt_e.flat.clone()))
    model_g, rb_g, opt_g, graphed = make()
    graphed.capture_graph(warmup=2)
    torch.cuda.synchronize()
    assert_close(opt_g.flat, want[1][1], "parameters after the 2 warm-up steps", rtol=1e-4, atol=2e-5)
    targets_seen = []
    for k in range(2, steps):
        m = graphed.train_step()
        torch.cuda.synchronize()
        targets_seen.append(float(m["train/target"]))
        rel, atol = (2e-3, 2e-5) if k == 2 else (5e-2, 5e-4)
        for name, v in want[k][0].items():
            assert float(m[name]) == pytest.approx(v, rel=rel, abs=1e-5), (k, name)
        assert_close(opt_g.flat, want[k][1], "parameters after step %d" % (k + 1), rtol=1e-3, atol=atol)
    assert len(set(targets_seen)) == len(targets_seen), "every replay must draw a new minibatch"
    assert [float(t) for t in targets_seen] == pytest.approx([w[0]["train/target"] for w in want[2:]], rel=1e-5)   # mean(v_target): data only
    assert opt_g.sync_step_count() == steps and graphed._step_counter == steps
    assert model_g.pushes == model_e.pushes == [2, 4]
    assert int(rb_g.sampler.draw_counter.item()) == steps * N == rb_e.sampler.draws


@pytest.mark.parametrize("N,A,beta", [(512, 18, 0.4), (1, 1, 0.4), (33, 6, 0.0), (4096, 4, 1.0), (100_000, 15, 0.6), (257, 51, 0.4)])
def test_per_td_loss_matches_the_reference_lines(cuda, N, A, beta):
    """N3b: the PER-weighted TD loss of ApexLearner._train_step (reference agents/dqn/learning.py:161-172) in one launch:
    loss, the gradient the Q network receives, the new priorities and the logged means against the oracle; then the whole
    prioritized loop on the device -- sample (N3) -> loss -> update(keys, priorities)."""
    g = torch.Generator().manual_seed(N + A)
    q = torch.randn(N, A, generator=g)
    a = torch.randint(0, A, (N,), generator=g)
    tgt = torch.randn(N, generator=g) * 2
    p = (torch.rand(N, generator=g) ** 3) * 0.01 + 1e-6
    qc = q.clone().requires_grad_(True)
    want_loss, want_pri, want_m = losses_ref.apex_td_loss(qc, a, tgt, p, beta)
    want_loss.backward()
    qg = q.to(cuda).requires_grad_(True)
    loss, pri, m = losses.per_td_loss(qg, a.to(cuda), tgt.to(cuda), p.to(cuda), beta)
    loss.backward()
    assert_close(loss, want_loss, "loss", rtol=1e-5, atol=1e-6)
    for k in want_m:
        assert_close(m[k], want_m[k], k, rtol=1e-5, atol=1e-6)
    assert_close(pri, want_pri, "priorities", rtol=1e-6, atol=1e-7)
    assert_close(qg.grad * N, qc.grad * N, "dq x N", rtol=1e-5, atol=1e-6)
    assert int((qg.grad != 0).sum()) <= N                                    # one element per row at most
    sc, dq, pr2, w = ops.per_td_loss_fwdbwd(q.to(cuda), a.to(cuda), tgt.to(cuda), p.to(cuda), beta, want_weights=True)
    wref = p.pow(-beta)
    assert_close(w, wref / wref.max(), "weights", rtol=1e-5, atol=1e-7)
    sc2, dq2, _, _ = ops.per_td_loss_fwdbwd(q.to(cuda), a.to(cuda), tgt.to(cuda), p.to(cuda), beta)
    assert torch.equal(sc, sc2) and torch.equal(dq, dq2)                      # deterministic
    if N != 512:
        return
    rb = replay.ReplayBuffer(replay.CircularBuffer(2048, collate_fn=torch.cat, device=cuda), replay.PrioritizedSampler(0.6))
    rb.sampler.reset(3)
    rb.extend((torch.randn(2048, 8), torch.randint(0, A, (2048,)), torch.randn(2048)), priorities=torch.rand(2048) + 0.1)
    keys, (obs, act, target), probs = rb.sample(N)
    net = torch.nn.Linear(8, A).to(cuda)
    loss, priorities, metrics = losses.per_td_loss(net(obs), act, target, probs, 0.4)
    loss.backward()
    assert net.weight.grad is not None and torch.isfinite(net.weight.grad).all()
    before = rb.sampler.tree[1].item()
    rb.update(keys, priorities)                                               # :184
    assert rb.sampler.tree[1].item() != before
