"""GPU: K1 (fused V-trace + IMPALA loss fwd/bwd) through the C ABI vs the CPU oracle."""
import os

import pytest
import torch

from conftest import GOLDEN
from gpu_util import assert_grad_parity, assert_close, npz_tensors, oracle_impala, to_cuda
from impala_b200 import losses, ops, synth

pytestmark = pytest.mark.gpu
KEYS = ("train/loss", "train/entropy", "train/td", "train/pg", "train/kl", "train/ratio")
SLOTS = (ops.S_LOSS, ops.S_ENTROPY, ops.S_TD, ops.S_PG, ops.S_KL, ops.S_RATIO)


def run_k1(d, time_major=False, **kw):
    g = to_cuda(d)
    if time_major:
        g = {k: v.transpose(0, 1).contiguous() for k, v in g.items()}
    sc, dl, dv = ops.impala_loss_fwdbwd(g["logits"], g["values"], g["a"], g["r"], g["discount"], g["pi_ref"],
                                        time_major=time_major, **kw)
    torch.cuda.synchronize()
    if time_major:
        dl, dv = dl.transpose(0, 1), dv.transpose(0, 1)
    return sc.cpu(), dl.cpu(), dv.cpu()


def check(d, okw=None, kkw=None, time_major=False):
    okw, kkw = okw or {}, kkw or {}
    loss, m, dl_ref, dv_ref = oracle_impala(d, **okw)
    sc, dl, dv = run_k1(d, time_major, **kkw)
    for k, s in zip(KEYS, SLOTS):
        assert_close(sc[s], m[k], k, rtol=1e-5, atol=1e-6)
    assert_close(dl, dl_ref, "dlogits", rel_to_max=2e-6)
    assert_close(dv, dv_ref, "dvalues", rel_to_max=2e-6)
    _, _, dl64, dv64 = oracle_impala(d, dtype=torch.float64, **okw)
    n = d["values"].numel()
    assert_grad_parity(dl, dl_ref, n, "dlogits x N", dl64)
    assert_grad_parity(dv, dv_ref, n, "dvalues x N", dv64)


def test_golden_c1_matches_committed_vectors():
    d, z = npz_tensors(os.path.join(GOLDEN, "impala_c1.npz"), ("logits", "values", "a", "r", "discount", "pi_ref"))
    sc, dl, dv = run_k1(d)
    for i, s in enumerate(SLOTS):
        assert sc[s].item() == pytest.approx(z["metrics"][i], rel=1e-5, abs=1e-6)
    assert_close(dl, torch.from_numpy(z["dlogits"]), "dlogits", rel_to_max=2e-6)
    assert_close(dv, torch.from_numpy(z["dvalues"]), "dvalues", rel_to_max=2e-6)


def test_golden_small_lambda_clip_variants():
    d, z = npz_tensors(os.path.join(GOLDEN, "impala_small_lambda.npz"), ("logits", "values", "a", "r", "discount", "pi_ref"))
    sc, dl, dv = run_k1(d, lambda_=.9, clip_rho=.8, clip_pg_rho=1.2, entropy_coeff=.05)
    assert sc[ops.S_LOSS].item() == pytest.approx(z["metrics"][0], rel=1e-5, abs=1e-6)
    assert_close(dl, torch.from_numpy(z["dlogits"]), "dlogits", rel_to_max=2e-6)
    assert_close(dv, torch.from_numpy(z["dvalues"]), "dvalues", rel_to_max=2e-6)


@pytest.mark.parametrize("B,T,A", [(32, 20, 15), (1, 2, 1), (5, 7, 6), (3, 20, 15), (33, 20, 15), (64, 5, 8),
                                   (17, 33, 16), (9, 12, 17), (4, 6, 32), (2, 300, 4), (130, 20, 18)])
@pytest.mark.parametrize("time_major", [False, True])
def test_parity_shapes(B, T, A, time_major):
    """Aligned tiles take the TMA bulk path, ragged ones (odd B, A, tail tiles) the cooperative path."""
    check(synth.impala_rollout(B, T, A, seed=B * 1000 + T * 10 + A), time_major=time_major)


@pytest.mark.parametrize("B,T,A", [(32, 20, 15), (33, 20, 15), (5, 7, 6), (130, 20, 18), (4096, 20, 15)])
@pytest.mark.parametrize("impl", ["1", "2"])
def test_both_implementations_batch_major(B, T, A, impl, monkeypatch):
    """Batch-major launches dispatch on B200RL_INPUTS_STABLE (csrc/impala_loss.cu); both kernels are held to the oracle
    on the same shapes: B200RL_K1_TILE=1 forces the tile kernel, =2 the persistent one (falls back to the tile kernel for
    shapes it does not take, e.g. B*T not a multiple of 4)."""
    monkeypatch.setenv("B200RL_K1_TILE", impl)
    check(synth.impala_rollout(B, T, A, seed=B + T + A))


@pytest.mark.parametrize("okw,kkw", [
    (dict(lambda_=.9), dict(lambda_=.9)),
    (dict(lambda_=0.), dict(lambda_=0.)),
    (dict(clip_rho_threshold=.7, clip_pg_rho_threshold=1.5), dict(clip_rho=.7, clip_pg_rho=1.5)),
    (dict(entropy_coeff=.2), dict(entropy_coeff=.2)),
    (dict(detach_adv=True), dict(flags=ops.DETACH_ADV)),
    (dict(detach_rho=True, lambda_=.8), dict(flags=ops.DETACH_RHO, lambda_=.8)),
    (dict(detach_adv=True, detach_rho=True), dict(flags=ops.DETACH_ADV | ops.DETACH_RHO)),
])
def test_parity_semantic_switches(okw, kkw):
    check(synth.impala_rollout(24, 20, 15, seed=7), okw, kkw)
    check(synth.impala_rollout(7, 11, 6, seed=8), okw, kkw, time_major=True)


def test_all_done_and_no_done_and_extreme_rho():
    d = synth.impala_rollout(16, 20, 15, seed=3)
    d["discount"] = torch.zeros_like(d["discount"])                  # every step terminal
    check(d)
    d["discount"] = torch.full_like(d["discount"], .99)              # no terminal
    check(d)
    d["pi_ref"] = d["logits"] + 3. * torch.randn_like(d["logits"])   # rho from ~0 to >> 1
    check(d)
    d["pi_ref"] = d["logits"].clone()                                # rho == 1 exactly: the clip tie
    check(d)


def test_full_size_c2_properties():
    """BASELINE config 2 (T=20, B=4096, A=15): determinism, column-permutation equivariance, and
    sharded-sum == global (what the multi-GPU path relies on); plus oracle parity at full size."""
    B, T, A = 4096, 20, 15
    d = synth.impala_rollout(B, T, A, seed=2000)
    sc, dl, dv = run_k1(d)
    sc2, dl2, dv2 = run_k1(d)
    assert torch.equal(sc, sc2) and torch.equal(dl, dl2) and torch.equal(dv, dv2)
    perm = torch.randperm(B, generator=torch.Generator().manual_seed(0))
    scp, dlp, dvp = run_k1({k: v[perm] for k, v in d.items()})
    assert torch.equal(dlp, dl[perm]) and torch.equal(dvp, dv[perm])
    assert_close(scp[:6], sc[:6], "permuted scalars", rtol=1e-5, atol=1e-6)
    for n_shards in (2, 4, 8):      # the B-sharding of the multi-GPU path: shard sums == the global launch
        w = B // n_shards
        parts = [run_k1({k: v[i * w:(i + 1) * w] for k, v in d.items()}, n_global_b=B) for i in range(n_shards)]
        assert torch.equal(torch.cat([p[1] for p in parts]), dl) and torch.equal(torch.cat([p[2] for p in parts]), dv)
        assert_close(sum(p[0] for p in parts)[:6], sc[:6], "sharded scalars", rtol=1e-5, atol=1e-6)
    check(d)


def test_autograd_dropin_and_reference_signature(cuda):
    d = to_cuda(synth.impala_rollout(8, 20, 15, seed=5))
    lin = torch.nn.Linear(4, 16).to(cuda)
    feats = torch.randn(8 * 20, 4, device=cuda)

    def model(s):
        out = lin(feats)
        return out[:, :15], out[:, 15:]

    batch = (torch.zeros(8, 20, 1, device=cuda), d["a"].unsqueeze(-1), d["r"].unsqueeze(-1),
             d["discount"].unsqueeze(-1), d["pi_ref"])
    loss, m = losses.impala_loss(batch, model, lambda_=1., entropy_cost=.01)
    assert list(m) == list(KEYS)
    (2 * loss).backward()
    g_fused = lin.weight.grad.clone()
    # same thing through the oracle on the CPU
    lin_c = torch.nn.Linear(4, 16)
    lin_c.load_state_dict({k: v.cpu() for k, v in lin.state_dict().items()})
    out = lin_c(feats.cpu())
    from oracle import losses_ref
    c = {k: v.cpu() for k, v in d.items()}
    l2, _ = losses_ref.impala_train_loss(out[:, :15].reshape(8, 20, 15), out[:, 15:].reshape(8, 20), c["a"], c["r"],
                                         c["discount"], c["pi_ref"])
    (2 * l2).backward()
    assert_close(loss, l2, "loss", rtol=1e-5, atol=1e-6)
    assert_close(g_fused, lin_c.weight.grad, "dW", rtol=1e-4, atol=1e-6, rel_to_max=1e-4)


def test_errors_raise_runtime_error(cuda):
    d = to_cuda(synth.impala_rollout(4, 20, 40, seed=1))
    with pytest.raises(RuntimeError, match="32 actions"):
        ops.impala_loss_fwdbwd(d["logits"], d["values"], d["a"], d["r"], d["discount"], d["pi_ref"])
    d = to_cuda(synth.impala_rollout(4, 20, 15, seed=1))
    with pytest.raises(RuntimeError, match="shape"):
        ops.impala_loss_fwdbwd(d["logits"], d["values"][:, :-1], d["a"], d["r"], d["discount"], d["pi_ref"])


@pytest.mark.parametrize("impl", ["1", "2", "0"])
def test_pipelined_launches_match_serial(cuda, impl, monkeypatch):
    """PDL (B200RL_INPUTS_STABLE) and multi-stream issue are scheduling only: within one implementation (B200RL_K1_TILE=1
    the tile kernel, =2 the persistent kernel) every launch of a pipeline over disjoint buffers reproduces the serial
    result bit for bit.  With the default dispatch (=0) the serial launches run the persistent kernel and the pipelined
    ones the tile kernel: gradients still bit for bit (same per-sample code), scalars to the tolerance (another
    summation order).  No launch may poison the workspace."""
    monkeypatch.setenv("B200RL_K1_TILE", impl)
    sets = [to_cuda(synth.impala_rollout(96, 20, 15, seed=900 + i)) for i in range(6)]

    def same(a, b):
        if impl != "0":
            return all(torch.equal(x, y) for x, y in zip(a, b))
        torch.testing.assert_close(a[0][:6], b[0][:6], rtol=1e-5, atol=1e-6)
        return torch.equal(a[1], b[1]) and torch.equal(a[2], b[2])

    def run(g, flags=0):
        return ops.impala_loss_fwdbwd(g["logits"], g["values"], g["a"], g["r"], g["discount"], g["pi_ref"], flags=flags)

    serial = []
    for g in sets:
        serial.append(run(g))
        torch.cuda.synchronize()
    # (1) back-to-back PDL launches on one stream, inputs independent of the previous launch
    piped = [run(g, ops.INPUTS_STABLE) for g in sets for _ in range(3)]
    torch.cuda.synchronize()
    for k, out in enumerate(piped):
        assert same(out, serial[k // 3])
    # (2) round-robin over 3 streams (each stream has its own workspace), PDL on
    streams = [torch.cuda.Stream(cuda) for _ in range(3)]
    cur = torch.cuda.current_stream()
    outs = []
    for s_ in streams:
        s_.wait_stream(cur)
    for k in range(18):
        with torch.cuda.stream(streams[k % 3]):
            outs.append(run(sets[k % 6], ops.INPUTS_STABLE))
    for s_ in streams:
        cur.wait_stream(s_)
    torch.cuda.synchronize()
    for k, out in enumerate(outs):
        assert same(out, serial[k % 6])
    for ws in ops._workspaces.values():
        hdr = ws[:8].view(torch.int32).tolist()     # [launch generation, poisoned] (csrc/common.cuh grid_collect)
        assert hdr[1] == 0, "a collector timed out and poisoned the workspace"
    assert sum(ws[:4].view(torch.int32).item() for ws in ops._workspaces.values()) >= 18


def _cabi_trace_offset():
    from impala_b200 import _cabi
    return _cabi.load().b200rl_trace_offset_bytes()


@pytest.mark.parametrize("flags", [0, ops.INPUTS_STABLE])
def test_graph_capture_and_replay(cuda, flags):
    """The C ABI is CUDA-graph capturable (no allocation, no host reads): a captured pipeline replays to the
    same bits as eager launches of the same kind (flags = 0: the persistent kernel as a dependent chain;
    INPUTS_STABLE: the tile kernel with PDL overlap)."""
    sets = [to_cuda(synth.impala_rollout(64, 20, 15, seed=950 + i)) for i in range(4)]
    outs = [(torch.empty(8, device=cuda), torch.empty_like(g["logits"]), torch.empty_like(g["values"])) for g in sets]
    want = []
    for g in sets:
        want.append([x.clone() for x in ops.impala_loss_fwdbwd(g["logits"], g["values"], g["a"], g["r"],
                                                               g["discount"], g["pi_ref"], flags=flags)])
        torch.cuda.synchronize()
    torch.cuda.synchronize()
    side = torch.cuda.Stream(cuda)
    with torch.cuda.stream(side):       # workspace for the capture stream must exist before capture
        ops.workspace(cuda)
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=side):
        for g, out in zip(sets, outs):
            ops.impala_loss_fwdbwd(g["logits"], g["values"], g["a"], g["r"], g["discount"], g["pi_ref"],
                                   flags=flags, out=out)
    for _ in range(3):
        gr.replay()
    torch.cuda.synchronize()
    for out, w in zip(outs, want):
        for a, b in zip(out, w):
            assert torch.equal(a, b)


def test_trace_flag_stamps_phases(cuda):
    g = to_cuda(synth.impala_rollout(64, 20, 15, seed=77))
    base = ops.impala_loss_fwdbwd(g["logits"], g["values"], g["a"], g["r"], g["discount"], g["pi_ref"])
    traced = ops.impala_loss_fwdbwd(g["logits"], g["values"], g["a"], g["r"], g["discount"], g["pi_ref"], flags=ops.TRACE)
    torch.cuda.synchronize()
    for a, b in zip(base, traced):
        assert torch.equal(a, b)
    off = _cabi_trace_offset()
    tr = ops.workspace(cuda)[off:off + 8 * 16 * 8].view(torch.int64).cpu().view(8, 16)   # 64 columns / 8 per CTA
    assert (tr[:, 0] > 0).all()
    assert ((tr[:, 9:15] - tr[:, 8:14]) >= 0).all()                                # clock64 monotone per CTA
    ops.workspace(cuda)[off:off + 8 * 16 * 8].zero_()


@pytest.mark.parametrize("B,T,A,stages", [(32768, 20, 15, None), (20000, 20, 6, None), (3000, 128, 15, None),
                                            (4096, 20, 15, "2"), (1500, 20, 15, "3"), (65536, 20, 15, None),
                                            # beyond 16384 CTAs of the tuned tile: the tile kernel widens its tiles (fuzz find)
                                            (20421, 128, 15, None), (42594, 64, 6, None), (17000, 128, 18, None)])
def test_persistent_kernel_ring_refill_matches_tile_kernel(cuda, B, T, A, stages, monkeypatch):
    """Large batches: a CTA of the persistent kernel owns more chunks than fit in shared memory, so stages are refilled
    after their bulk store has drained (mbarrier phase flips, per-stage completion counts).  The tile kernel computes every
    sample with the same code: gradients bit for bit, scalars to the tolerance (another summation order); plus the oracle
    on a column sample through the sharding identity (columns are independent).  `stages` forces a short ring
    (B200RL_K1_STAGES) so that small batches go through many rounds too.  Regression test for a round-aliasing bug: a group
    two rounds ahead on a stage matched the parity of the long-completed round 0 (fixed by the per-stage `issued` counter)."""
    if stages:
        monkeypatch.setenv("B200RL_K1_STAGES", stages)
    g = to_cuda(synth.impala_rollout(B, T, A, seed=B + T))
    outs = {}
    for impl in ("1", "2"):
        monkeypatch.setenv("B200RL_K1_TILE", impl)
        outs[impl] = [x.clone() for x in ops.impala_loss_fwdbwd(g["logits"], g["values"], g["a"], g["r"], g["discount"], g["pi_ref"])]
        torch.cuda.synchronize()
    assert torch.equal(outs["1"][1], outs["2"][1]) and torch.equal(outs["1"][2], outs["2"][2])
    torch.testing.assert_close(outs["1"][0][:6], outs["2"][0][:6], rtol=1e-5, atol=1e-6)
    # the oracle on the first and last 64 columns (n_global_b = B keeps the scaling of the full launch)
    for sl in (slice(0, 64), slice(B - 64, B)):
        d = {k: v[sl].cpu() for k, v in g.items()}
        _, _, dl_ref, dv_ref = oracle_impala(d)
        scale = 64.0 / B
        assert_close(outs["2"][1][sl], dl_ref * scale, "dlogits (column sample)", rel_to_max=2e-6)
        assert_close(outs["2"][2][sl], dv_ref * scale, "dvalues (column sample)", rel_to_max=2e-6)
