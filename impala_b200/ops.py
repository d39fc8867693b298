"""Thin functional wrappers: torch tensors in, C-ABI launches on the current stream."""
import ctypes

import torch

from . import _cabi

DETACH_ADV, DETACH_RHO, ORDER_RLAX, INPUTS_STABLE, TRACE = 1, 2, 4, 8, 0x80000000
S_LOSS, S_ENTROPY, S_TD, S_PG, S_KL, S_RATIO, S_TARGET = range(7)

_workspaces = {}


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("impala_b200 ops need CUDA tensors (no CPU fallback); got %s" % t.device)


def workspace(device):
    """Zero-initialised scratch for the fused losses, one per (device, stream)."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), _stream())
    ws = _workspaces.get(key)
    if ws is None:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("impala_b200: the loss workspace of this stream must exist before CUDA-graph capture "
                               "(call ops.workspace(device) on the capture stream first); allocating it inside the "
                               "capture would tie it to the graph's private memory pool")
        n = _cabi.load().b200rl_loss_workspace_bytes()
        ws = torch.zeros(n, dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def _f32c(x, name="tensor"):
    """Kernel inputs are contiguous float32.  The library makes NO hidden copies (a silent ``.float().contiguous()``
    is HBM traffic the size of the kernel's own): anything else raises, like every other contract violation here."""
    if x.dtype == torch.float32 and x.is_contiguous():
        return x
    raise RuntimeError("impala_b200: %s must be a contiguous float32 tensor (got dtype %s, shape %s, strides %s); convert it "
                       "explicitly -- the ops never copy behind the caller's back" % (name, x.dtype, tuple(x.shape), tuple(x.stride())))


def _i64c(x, name="tensor"):
    if x.dtype == torch.int64 and x.is_contiguous():
        return x
    raise RuntimeError("impala_b200: %s must be a contiguous int64 tensor (got dtype %s, shape %s, strides %s); convert it "
                       "explicitly -- the ops never copy behind the caller's back" % (name, x.dtype, tuple(x.shape), tuple(x.stride())))


def impala_loss_fwdbwd(logits, values, actions, rewards, discounts, pi_ref, entropy_coeff=0.01,
                       lambda_=1.0, clip_rho=1.0, clip_pg_rho=1.0, flags=0, n_global_b=None,
                       time_major=False, out=None):
    """K1.  Batch-major ``[B,T,A]`` / ``[B,T]`` tensors (or ``[T,B,A]`` / ``[T,B]`` with
    ``time_major=True``).  Returns ``(scalars[8], dlogits, dvalues)``; nothing is synchronised."""
    _require_cuda(logits, values, actions, rewards, discounts, pi_ref)
    logits, pi_ref = _f32c(logits.detach(), "logits"), _f32c(pi_ref.detach(), "pi_ref")
    values, rewards, discounts = _f32c(values.detach(), "values"), _f32c(rewards, "rewards"), _f32c(discounts, "discounts")
    actions = _i64c(actions, "actions")
    if time_major:
        T, B, A = logits.shape
        sb, st = 1, B
    else:
        B, T, A = logits.shape
        sb, st = T, 1
    for name, x in (("values", values), ("actions", actions), ("rewards", rewards), ("discounts", discounts)):
        if tuple(x.shape) != tuple(logits.shape[:2]):
            raise RuntimeError("impala_loss: %s has shape %s, expected %s" % (name, tuple(x.shape), tuple(logits.shape[:2])))
    if pi_ref.shape != logits.shape:
        raise RuntimeError("impala_loss: pi_ref shape %s != logits shape %s" % (tuple(pi_ref.shape), tuple(logits.shape)))
    if out is None:
        scalars = torch.empty(8, dtype=torch.float32, device=logits.device)
        dlogits = torch.empty_like(logits)
        dvalues = torch.empty_like(values)
    else:
        scalars, dlogits, dvalues = out
    ws = workspace(logits.device)
    _cabi.call("b200rl_vtrace_impala_loss_fwdbwd",
               logits.data_ptr(), pi_ref.data_ptr(), actions.data_ptr(), rewards.data_ptr(),
               discounts.data_ptr(), values.data_ptr(), B, T, A, sb, st,
               B if n_global_b is None else n_global_b,
               lambda_, clip_rho, clip_pg_rho, entropy_coeff, flags,
               dlogits.data_ptr(), dvalues.data_ptr(), scalars.data_ptr(), ws.data_ptr(), _stream())
    return scalars, dlogits, dvalues


def _major(x):
    """(B, T, stride_b, stride_t, time_major) of a 2-D [B,T] view in either major order."""
    B, T = x.shape
    sb, st = x.stride()
    if st == 1 and sb >= T:
        return B, T, sb, st, False
    if sb == 1 and st >= B:
        return B, T, sb, st, True
    if T == 1 or B == 1:
        return B, T, T, 1, False
    return None


def _same_view(xs):
    """Make 2-D float32 inputs share one (stride_b, stride_t); copy only when needed."""
    xs = [x if x.dtype == torch.float32 else x.float() for x in xs]
    m = [_major(x) for x in xs]
    if any(v is None for v in m) or len({v[2:] for v in m}) != 1:
        xs = [x.contiguous() for x in xs]
        m = [_major(x) for x in xs]
    return xs, m[0]


def _compact_like(B, T, time_major, device):
    """Compact [B,T] float32 output in the requested major order."""
    if time_major:
        return torch.empty((T, B), dtype=torch.float32, device=device).t()
    return torch.empty((B, T), dtype=torch.float32, device=device)


def vtrace_fwd(v_tm1, v_t, r_t, discount_t, rho_tm1, lambda_=1.0, clip_rho=1.0, clip_pg_rho=1.0):
    """Stand-alone V-trace on ``[B,Tm]`` views -> compact ``(adv, err, q)`` and the layout used."""
    _require_cuda(v_tm1, v_t, r_t, discount_t, rho_tm1)
    xs, (B, Tm, sb, st, tm) = _same_view([v_tm1.detach(), v_t.detach(), r_t.detach(),
                                          discount_t.detach(), rho_tm1.detach()])
    adv, err, q = (_compact_like(B, Tm, tm, xs[0].device) for _ in range(3))
    _cabi.call("b200rl_vtrace", *[x.data_ptr() for x in xs], B, Tm, sb, st, lambda_, clip_rho, clip_pg_rho,
               adv.data_ptr(), err.data_ptr(), q.data_ptr(), _stream())
    return adv, err, q, xs, (B, Tm, sb, st, tm)


def vtrace_bwd(saved, layout, q, g_adv, g_err, g_q, lambda_, clip_pg_rho):
    v_tm1, _, _, disc, rho = saved
    B, Tm, sb, st, tm = layout

    def compact(g):
        if g is None:
            return None
        g = g.float()
        want = (1, B) if tm else (Tm, 1)
        if g.stride() != want and not (B == 1 or Tm == 1):
            g = g.t().contiguous().t() if tm else g.contiguous()
        return g

    g_adv, g_err, g_q = compact(g_adv), compact(g_err), compact(g_q)
    d_vtm1, d_vt, d_rho = (_compact_like(B, Tm, tm, q.device) for _ in range(3))
    _cabi.call("b200rl_vtrace_bwd", v_tm1.data_ptr(), disc.data_ptr(), rho.data_ptr(), q.data_ptr(),
               _cabi.ptr(g_adv), _cabi.ptr(g_err), _cabi.ptr(g_q), B, Tm, sb, st, lambda_, clip_pg_rho,
               d_vtm1.data_ptr(), d_vt.data_ptr(), d_rho.data_ptr(), _stream())
    return d_vtm1, d_vt, d_rho


def lambda_returns(r_t, discount_t, v_t, lambda_=1.0):
    """K2.  ``[B,T]`` views (batch- or time-major) -> returns in the same major order."""
    _require_cuda(r_t, discount_t, v_t)
    xs, (B, T, sb, st, tm) = _same_view([r_t.detach(), discount_t.detach(), v_t.detach()])
    if (tm and st != B) or (not tm and sb != T):  # output must be addressable with the input strides
        xs = [x.t().contiguous().t() if tm else x.contiguous() for x in xs]
        B, T, sb, st, tm = _major(xs[0])
    out = _compact_like(B, T, tm, xs[0].device)
    _cabi.call("b200rl_lambda_returns", *[x.data_ptr() for x in xs], B, T, sb, st, lambda_,
               out.data_ptr(), _stream())
    return out


def gae(r_t, discount_t, values, lambda_=1.0):
    """``r_t, discount_t [B,T]``, ``values [B,T+1]`` -> advantages ``[B,T]``."""
    _require_cuda(r_t, discount_t, values)
    (r, d), (B, T, sb, st, tm) = _same_view([r_t.detach(), discount_t.detach()])
    v = values.detach().float()
    mv = _major(v)
    if mv is None or mv[4] != tm or (tm and st != B) or (not tm and sb != T):
        r, d, v = ((x.t().contiguous().t() if tm else x.contiguous()) for x in (r, d, v))
        B, T, sb, st, tm = _major(r)
        mv = _major(v)
    out = _compact_like(B, T, tm, r.device)
    _cabi.call("b200rl_gae", r.data_ptr(), d.data_ptr(), v.data_ptr(), B, T, sb, st, mv[2], mv[3],
               lambda_, out.data_ptr(), _stream())
    return out


def ppo_loss_fwdbwd(logits, v, actions, v_target, pi_ref, entropy_cost=0.01, clip_coeff=0.1,
                    n_global=None, out=None):
    """K3.  ``logits, pi_ref [N,A]``; ``v, actions, v_target [N]``.  Returns ``(scalars[8], dlogits, dv)``."""
    _require_cuda(logits, v, actions, v_target, pi_ref)
    logits, pi_ref = _f32c(logits.detach(), "logits"), _f32c(pi_ref.detach(), "pi_ref")
    v, v_target = _f32c(v.detach(), "v").reshape(-1), _f32c(v_target, "v_target").reshape(-1)
    actions = _i64c(actions, "actions").reshape(-1)
    N, A = logits.shape
    if pi_ref.shape != logits.shape or v.numel() != N or v_target.numel() != N or actions.numel() != N:
        raise RuntimeError("ppo_loss: inconsistent shapes")
    if out is None:
        scalars = torch.empty(8, dtype=torch.float32, device=logits.device)
        dlogits = torch.empty_like(logits)
        dv = torch.empty_like(v)
    else:
        scalars, dlogits, dv = out
    ws = workspace(logits.device)
    _cabi.call("b200rl_ppo_loss_fwdbwd", logits.data_ptr(), pi_ref.data_ptr(), actions.data_ptr(),
               v_target.data_ptr(), v.data_ptr(), N, A, N if n_global is None else n_global,
               clip_coeff, entropy_cost, 0, dlogits.data_ptr(), dv.data_ptr(), scalars.data_ptr(),
               ws.data_ptr(), _stream())
    return scalars, dlogits, dv


def ppo_is_loss_fwdbwd(logits, values, v_boot, actions, rewards, done, pi_ref, entropy_cost=0.01, clip_coeff=0.1,
                       lambda_=0.95, gamma=0.99, clip_rho=1.0, clip_pg_rho=1.0, flags=0, n_global_b=None, out=None):
    """K7 (reference ``losses.py:62-89`` intent).  ``logits, pi_ref [B,T,A]``; ``values, actions, rewards, done
    [B,T]``; ``v_boot [B]``.  Returns ``(scalars[8], dlogits, dvalues)``."""
    _require_cuda(logits, values, v_boot, actions, rewards, done, pi_ref)
    logits, pi_ref = _f32c(logits.detach(), "logits"), _f32c(pi_ref.detach(), "pi_ref")
    values, rewards, v_boot = _f32c(values.detach(), "values"), _f32c(rewards, "rewards"), _f32c(v_boot.detach(), "v_boot").reshape(-1)
    if logits.dim() != 3:
        raise RuntimeError("ppo_is_loss: logits must be [B,T,A]")
    B, T, A = logits.shape
    actions = _i64c(actions, "actions")
    if done.dtype == torch.bool and done.is_contiguous():
        done = done.view(torch.uint8)
    elif done.dtype != torch.uint8 or not done.is_contiguous():
        done = (done != 0).contiguous().view(torch.uint8)
    if pi_ref.shape != logits.shape or v_boot.numel() != B or \
            any(tuple(x.shape) != (B, T) for x in (values, actions, rewards, done)):
        raise RuntimeError("ppo_is_loss: inconsistent shapes")
    if out is None:
        scalars = torch.empty(8, dtype=torch.float32, device=logits.device)
        dlogits = torch.empty_like(logits)
        dvalues = torch.empty_like(values)
    else:
        scalars, dlogits, dvalues = out
    ws = workspace(logits.device)
    _cabi.call("b200rl_ppo_is_loss_fwdbwd", logits.data_ptr(), pi_ref.data_ptr(), actions.data_ptr(),
               rewards.data_ptr(), done.data_ptr(), values.data_ptr(), v_boot.data_ptr(), B, T, A,
               B if n_global_b is None else n_global_b, gamma, lambda_, clip_rho, clip_pg_rho, clip_coeff,
               entropy_cost, flags, dlogits.data_ptr(), dvalues.data_ptr(), scalars.data_ptr(), ws.data_ptr(),
               _stream())
    return scalars, dlogits, dvalues


def sample_uniform(seed, stream_id, draw_offset, n, size, device, out=None):
    """K4.  ``n`` int64 indices in ``[0,size)`` on ``device``."""
    if out is None:
        out = torch.empty(n, dtype=torch.int64, device=device)
    _require_cuda(out)
    _cabi.call("b200rl_sample_uniform", seed, stream_id, draw_offset, n, size, out.data_ptr(), _stream())
    return out


def gather_rows(src, idx, out=None):
    """K5.  ``out[i] = src[idx[i]]`` for a contiguous ``src [nrows, ...]``."""
    _require_cuda(src, idx)
    if not src.is_contiguous():
        raise RuntimeError("gather_rows: storage must be contiguous")
    idx = _i64c(idx, "idx")
    n, nrows = idx.numel(), src.shape[0]
    row_bytes = src[0].numel() * src.element_size() if nrows else 0
    if out is None:
        out = torch.empty((n,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    _cabi.call("b200rl_gather_rows", src.data_ptr(), row_bytes, idx.data_ptr(), n, nrows, out.data_ptr(), _stream())
    return out


def ring_append(storage, count, rows):
    """Rows ``j`` go to slot ``(count + j) % capacity`` of the contiguous ``storage [capacity, ...]``."""
    _require_cuda(storage, rows)
    rows = rows.contiguous()
    if rows.dtype != storage.dtype or tuple(rows.shape[1:]) != tuple(storage.shape[1:]):
        raise RuntimeError("ring_append: rows %s/%s do not match storage %s/%s"
                           % (tuple(rows.shape[1:]), rows.dtype, tuple(storage.shape[1:]), storage.dtype))
    row_bytes = storage[0].numel() * storage.element_size()
    _cabi.call("b200rl_ring_append", storage.data_ptr(), row_bytes, storage.shape[0], count,
               rows.data_ptr(), rows.shape[0], _stream())


def _counters_ptr(device_counters):
    if device_counters is None:
        return 0
    if device_counters.dtype != torch.int64 or device_counters.numel() < 2 or not device_counters.is_cuda or \
            not device_counters.is_contiguous():
        raise RuntimeError("device_counters must be a contiguous int64[2] CUDA tensor (Adam step, flag generation)")
    return device_counters.data_ptr()


def clip_adam_step(params, grads, exp_avg, exp_avg_sq, step, lr, beta1=0.9, beta2=0.999, eps=1e-8,
                   max_norm=0.0, total_norm=None, device_counters=None):
    """K6.  Flat fp32 buffers, updated in place; returns the 1-element ``total_norm`` tensor (pre-clip).  With
    ``device_counters`` (int64[2] on the device) the step number lives in device memory and the launch is graph-replayable."""
    _require_cuda(params, grads, exp_avg, exp_avg_sq)
    n = params.numel()
    for name, x in (("params", params), ("grads", grads), ("exp_avg", exp_avg), ("exp_avg_sq", exp_avg_sq)):
        if x.dtype != torch.float32 or not x.is_contiguous() or x.numel() != n:
            raise RuntimeError("clip_adam_step: %s must be a contiguous float32 buffer of %d elements" % (name, n))
    if total_norm is None:
        total_norm = torch.empty(1, dtype=torch.float32, device=params.device)
    ws = workspace(params.device)
    _cabi.call("b200rl_clip_adam_step", params.data_ptr(), grads.data_ptr(), exp_avg.data_ptr(),
               exp_avg_sq.data_ptr(), n, max_norm, lr, beta1, beta2, eps, step, _counters_ptr(device_counters),
               total_norm.data_ptr(), ws.data_ptr(), _stream())
    return total_norm


def peer_allreduce_clip_adam_step(params, peer_bucket_ptrs, peer_flag_ptrs, rank, grads_out, extra_out, exp_avg,
                                  exp_avg_sq, step, lr, beta1=0.9, beta2=0.999, eps=1e-8, max_norm=0.0,
                                  total_norm=None, peer_reduced_ptrs=None, peer_cta_flag_ptrs=None, mc_bucket_ptr=0,
                                  mc_reduced_ptr=0, flag_generation=None, device_counters=None):
    """K8.  ``peer_bucket_ptrs`` / ``peer_flag_ptrs``: device addresses (ints) of every rank's bucket and flag area as
    mapped into THIS process (``parallel.PeerGradBucket``).  Returns the 1-element pre-clip norm of the reduced gradient."""
    _require_cuda(params, grads_out, exp_avg, exp_avg_sq)
    n, world = params.numel(), len(peer_bucket_ptrs)
    for name, x in (("params", params), ("grads_out", grads_out), ("exp_avg", exp_avg), ("exp_avg_sq", exp_avg_sq)):
        if x.dtype != torch.float32 or not x.is_contiguous() or x.numel() != n:
            raise RuntimeError("peer_allreduce_clip_adam_step: %s must be a contiguous float32 buffer of %d elements" % (name, n))
    if len(peer_flag_ptrs) != world:
        raise RuntimeError("peer_allreduce_clip_adam_step: %d buckets vs %d flag areas" % (world, len(peer_flag_ptrs)))
    n_extra = 0 if extra_out is None else extra_out.numel()
    if total_norm is None:
        total_norm = torch.empty(1, dtype=torch.float32, device=params.device)
    buckets = (ctypes.c_void_p * world)(*peer_bucket_ptrs)
    flags = (ctypes.c_void_p * world)(*peer_flag_ptrs)
    two = peer_reduced_ptrs is not None and peer_cta_flag_ptrs is not None
    red = (ctypes.c_void_p * world)(*peer_reduced_ptrs) if two else None
    cfl = (ctypes.c_void_p * world)(*peer_cta_flag_ptrs) if two else None
    ws = workspace(params.device)
    _cabi.call("b200rl_peer_allreduce_clip_adam_step", params.data_ptr(), ctypes.addressof(buckets), ctypes.addressof(flags),
               ctypes.addressof(red) if two else 0, ctypes.addressof(cfl) if two else 0, mc_bucket_ptr or 0,
               mc_reduced_ptr or 0, world, rank, grads_out.data_ptr(), 0 if extra_out is None else extra_out.data_ptr(), n_extra,
               exp_avg.data_ptr(), exp_avg_sq.data_ptr(), n, max_norm, lr, beta1, beta2, eps, step,
               step if flag_generation is None else flag_generation, _counters_ptr(device_counters),
               total_norm.data_ptr(), ws.data_ptr(), _stream())
    return total_norm


def ipc_export(tensor):
    """``(handle: bytes[64], offset: int)`` naming ``tensor``'s memory for another process (K8 plumbing)."""
    _require_cuda(tensor)
    handle = ctypes.create_string_buffer(64)
    offset = ctypes.c_int64(0)
    _cabi.call("b200rl_ipc_export", tensor.data_ptr(), ctypes.addressof(handle), ctypes.addressof(offset))
    return handle.raw, offset.value


def ipc_import(handle, offset):
    """Device address (int) of a peer's buffer in this process (the mapping lives until ``ipc_close`` / exit)."""
    buf = ctypes.create_string_buffer(bytes(handle), 64)
    out = ctypes.c_void_p(0)
    _cabi.call("b200rl_ipc_import", ctypes.addressof(buf), offset, ctypes.addressof(out))
    return out.value


def ipc_close(ptr, offset):
    _cabi.call("b200rl_ipc_close", ptr, offset)


def ring_append_fields(storages, count, rows, done_field=None, gamma=0.99):
    """N2.  Append ``n`` stacked rows of every field to its ring storage in ONE launch.  ``storages[i]`` is
    ``[capacity, ...]``, ``rows[i]`` is ``[n, ...]`` (device, same trailing shape and dtype) -- except field
    ``done_field``: its ``rows`` hold bool/uint8 done flags and its storage receives float32
    ``logical_not(done) * gamma`` (reference ``agents/impala/learning.py:79``)."""
    _require_cuda(*storages, *rows)
    if len(storages) != len(rows) or not storages:
        raise RuntimeError("ring_append_fields: %d storages vs %d row tensors" % (len(storages), len(rows)))
    n, cap = rows[0].shape[0], storages[0].shape[0]
    table = (_cabi.Field * len(rows))()
    keep = []
    for i, (st, r) in enumerate(zip(storages, rows)):
        if r.shape[0] != n or st.shape[0] != cap or tuple(r.shape[1:]) != tuple(st.shape[1:]) or not st.is_contiguous():
            raise RuntimeError("ring_append_fields: field %d rows %s do not match storage %s" % (i, tuple(r.shape), tuple(st.shape)))
        r = r.contiguous()
        if i == done_field:
            if st.dtype != torch.float32 or r.dtype not in (torch.bool, torch.uint8):
                raise RuntimeError("ring_append_fields: done field needs bool/uint8 rows and float32 storage")
            r = r.view(torch.uint8)
            table[i] = _cabi.Field(r.data_ptr(), st.data_ptr(), r[0].numel() if n else 1, _cabi.FIELD_DONE_TO_DISCOUNT, gamma)
        else:
            if r.dtype != st.dtype:
                raise RuntimeError("ring_append_fields: field %d dtype %s != storage %s" % (i, r.dtype, st.dtype))
            table[i] = _cabi.Field(r.data_ptr(), st.data_ptr(), (r[0].numel() * r.element_size()) if n else 1, _cabi.FIELD_COPY, 0.0)
        keep.append(r)
    _cabi.call("b200rl_ring_append_fields", table, len(rows), cap, count, n, _stream())
    return keep  # the caller keeps the (possibly re-laid-out) sources alive until the stream has consumed them


def gather_fields(storages, idx, outs=None):
    """Multi-field K5: ``outs[f][i] = storages[f][idx[i]]`` for all fields in ONE launch."""
    _require_cuda(*storages, idx)
    idx = idx if (idx.dtype == torch.int64 and idx.is_contiguous()) else idx.long().contiguous()
    n, nrows = idx.numel(), storages[0].shape[0]
    if outs is None:
        outs = [torch.empty((n,) + tuple(st.shape[1:]), dtype=st.dtype, device=st.device) for st in storages]
    table = (_cabi.Field * len(storages))()
    for i, (st, o) in enumerate(zip(storages, outs)):
        if not st.is_contiguous() or st.shape[0] != nrows or not o.is_contiguous():
            raise RuntimeError("gather_fields: field %d must be contiguous with %d rows" % (i, nrows))
        table[i] = _cabi.Field(st.data_ptr(), o.data_ptr(), (st[0].numel() * st.element_size()) if nrows else 1,
                               _cabi.FIELD_COPY, 0.0)
    _cabi.call("b200rl_gather_fields", table, len(storages), idx.data_ptr(), n, nrows, _stream())
    return outs


def counter_add(counter, delta):
    """``counter[0] += delta`` in stream order (device-resident counters of graph-replayed launches)."""
    _require_cuda(counter)
    if counter.dtype != torch.int64 or not counter.is_contiguous():
        raise RuntimeError("counter_add: int64 CUDA tensor expected")
    _cabi.call("b200rl_counter_add", counter.data_ptr(), delta, _stream())


def sample_gather_fields(storages, seed, stream_id, draw_offset, n, size, outs=None, out_idx=None, draw_counter=None):
    """K4 + K5 in ONE launch: draws the ``n`` indices of ``sample_uniform(seed, stream_id, draw_offset, n, size)`` (same
    Philox stream, bit for bit) and gathers those rows of every field.  Returns ``(idx, outs)``."""
    _require_cuda(*storages)
    nrows, dev = storages[0].shape[0], storages[0].device
    if out_idx is None:
        out_idx = torch.empty(n, dtype=torch.int64, device=dev)
    if outs is None:
        outs = [torch.empty((n,) + tuple(st.shape[1:]), dtype=st.dtype, device=st.device) for st in storages]
    table = (_cabi.Field * len(storages))()
    for i, (st, o) in enumerate(zip(storages, outs)):
        if not st.is_contiguous() or st.shape[0] != nrows or not o.is_contiguous():
            raise RuntimeError("sample_gather_fields: field %d must be contiguous with %d rows" % (i, nrows))
        table[i] = _cabi.Field(st.data_ptr(), o.data_ptr(), (st[0].numel() * st.element_size()) if nrows else 1,
                               _cabi.FIELD_COPY, 0.0)
    _cabi.call("b200rl_sample_gather_fields", table, len(storages), seed, stream_id, draw_offset,
               0 if draw_counter is None else _i64c(draw_counter, "draw_counter").data_ptr(), n, size, nrows,
               _i64c(out_idx, "out_idx").data_ptr(), _stream())
    return out_idx, outs


def n_step_returns(r_t, discount_t, v_t, n, lambda_=1.0):
    """N4.  ``[B,T]`` views (batch- or time-major) -> n-step bootstrapped returns, same major order."""
    _require_cuda(r_t, discount_t, v_t)
    xs, (B, T, sb, st, tm) = _same_view([r_t.detach(), discount_t.detach(), v_t.detach()])
    if (tm and st != B) or (not tm and sb != T):
        xs = [x.t().contiguous().t() if tm else x.contiguous() for x in xs]
        B, T, sb, st, tm = _major(xs[0])
    out = _compact_like(B, T, tm, xs[0].device)
    _cabi.call("b200rl_n_step_returns", *[x.data_ptr() for x in xs], B, T, sb, st, int(n), lambda_,
               out.data_ptr(), _stream())
    return out


def sumtree_update(tree, idx, value):
    """N3.  ``tree``: float64 ``[2 * leaves]`` (leaves a power of two); leaf ``idx[i] = value[i]``, sums refreshed."""
    _require_cuda(tree, idx, value)
    if tree.dtype != torch.float64 or not tree.is_contiguous():
        raise RuntimeError("sumtree_update: tree must be a contiguous float64 tensor")
    idx = idx if (idx.dtype == torch.int64 and idx.is_contiguous()) else idx.long().contiguous()
    value = value if (value.dtype == torch.float64 and value.is_contiguous()) else value.double().contiguous()
    if idx.numel() != value.numel():
        raise RuntimeError("sumtree_update: %d indices vs %d values" % (idx.numel(), value.numel()))
    _cabi.call("b200rl_sumtree_update", tree.data_ptr(), tree.numel() // 2, idx.data_ptr(), value.data_ptr(),
               idx.numel(), _stream())


def sumtree_sample(tree, size, seed, stream_id, draw_offset, n, with_prob=True):
    """N3.  ``n`` indices in ``[0,size)`` drawn proportionally to the leaves, and ``leaf / total`` (float32)."""
    _require_cuda(tree)
    idx = torch.empty(n, dtype=torch.int64, device=tree.device)
    prob = torch.empty(n, dtype=torch.float32, device=tree.device) if with_prob else None
    _cabi.call("b200rl_sumtree_sample", tree.data_ptr(), tree.numel() // 2, size, seed, stream_id, draw_offset, n,
               idx.data_ptr(), _cabi.ptr(prob), _stream())
    return idx, prob


def quantile_huber_fwdbwd(dist_src, tau_src, dist_target, huber_param=1.0):
    """N4b.  ``dist_src, tau_src [B,N]``, ``dist_target [B,M]`` -> ``(loss [B], d loss_b / d dist_src [B,N])``."""
    _require_cuda(dist_src, tau_src, dist_target)
    src, tau, tgt = _f32c(dist_src.detach()), _f32c(tau_src.detach()), _f32c(dist_target.detach())
    B, N = src.shape
    if tau.shape != src.shape or tgt.dim() != 2 or tgt.shape[0] != B:
        raise RuntimeError("quantile_huber: shapes %s %s %s" % (tuple(src.shape), tuple(tau.shape), tuple(tgt.shape)))
    loss = torch.empty(B, dtype=torch.float32, device=src.device)
    d_src = torch.empty_like(src)
    _cabi.call("b200rl_quantile_huber_fwdbwd", src.data_ptr(), tau.data_ptr(), tgt.data_ptr(), B, N, tgt.shape[1],
               float(huber_param), loss.data_ptr(), d_src.data_ptr(), _stream())
    return loss, d_src


def per_td_loss_fwdbwd(q, actions, target, probabilities, importance_sampling_exponent, want_weights=False):
    """N3b (reference ``agents/dqn/learning.py:161-172``).  ``q [N,A]``, ``actions, target, probabilities [N]`` ->
    ``(scalars[8], dq [N,A], priorities [N], weights [N] or None)``; ``scalars`` = loss, mean q, mean target, mean
    priority, the weight normaliser."""
    _require_cuda(q, actions, target, probabilities)
    q, target, probabilities = _f32c(q.detach(), "q"), _f32c(target.detach(), "target"), _f32c(probabilities, "probabilities")
    actions = _i64c(actions, "actions")
    if q.dim() != 2 or any(tuple(x.shape) != (q.shape[0],) for x in (actions, target, probabilities)):
        raise RuntimeError("per_td_loss: expected q [N,A] and actions / target / probabilities [N], got %s %s %s %s"
                           % (tuple(q.shape), tuple(actions.shape), tuple(target.shape), tuple(probabilities.shape)))
    N, A = q.shape
    scalars = torch.empty(8, dtype=torch.float32, device=q.device)
    dq, priorities = torch.empty_like(q), torch.empty_like(target)
    weights = torch.empty_like(target) if want_weights else None
    ws = workspace(q.device)
    _cabi.call("b200rl_per_td_loss_fwdbwd", q.data_ptr(), actions.data_ptr(), target.data_ptr(), probabilities.data_ptr(),
               N, A, float(importance_sampling_exponent), dq.data_ptr(), priorities.data_ptr(), _cabi.ptr(weights),
               scalars.data_ptr(), ws.data_ptr(), _stream())
    return scalars, dq, priorities, weights


def act_epilogue(logits, deterministic, seed, stream_id, draw_offset):
    """N5.  ``logits [n,A]`` -> int64 actions ``[n,1]``; ``deterministic``: bool tensor of 1 or n elements."""
    _require_cuda(logits)
    logits = _f32c(logits.detach())
    n, A = logits.shape
    det = deterministic.to(logits.device).reshape(-1).to(torch.uint8).contiguous()
    if det.numel() not in (1, n):
        raise RuntimeError("act_epilogue: deterministic has %d elements for %d rows" % (det.numel(), n))
    action = torch.empty((n, 1), dtype=torch.int64, device=logits.device)
    _cabi.call("b200rl_act_epilogue", logits.data_ptr(), n, A, det.data_ptr(), 0 if det.numel() == 1 and n != 1 else 1,
               seed, stream_id, draw_offset, action.data_ptr(), _stream())
    return action


# ---- K9: memory-bound glue of the IMPALA-CNN (fp32 NHWC = torch channels_last) ------------------------------------
GLUE_MAX_CTAS = 148 * 8          # grid of the two grid-stride glue kernels: 8 CTAs of 256 threads per SM


def _nhwc(x, name):
    if x.dim() != 4 or x.dtype != torch.float32 or not x.is_contiguous(memory_format=torch.channels_last):
        raise RuntimeError("%s: expected a float32 channels_last [N,C,H,W] tensor, got %s %s strides %s"
                           % (name, tuple(x.shape), x.dtype, x.stride()))
    return x


def _empty_nhwc(N, C, H, W, device, dtype=torch.float32):
    return torch.empty((N, C, H, W), dtype=dtype, device=device, memory_format=torch.channels_last)


def _glue_ctas(n_float4):
    return max(1, min(GLUE_MAX_CTAS, (n_float4 + 255) // 256))


def obs_to_f32_nhwc(obs, divisor=255.0):
    """``obs.float() / divisor`` (reference ``models/models.py:73``) for uint8 ``[N,3,H,W]`` NCHW frames, written
    directly in channels_last memory order: one pass instead of cast + divide + layout copy."""
    _require_cuda(obs)
    if obs.dtype != torch.uint8 or obs.dim() != 4 or not obs.is_contiguous():
        raise RuntimeError("obs_to_f32_nhwc: expected contiguous uint8 [N,C,H,W], got %s %s" % (tuple(obs.shape), obs.dtype))
    N, C, H, W = obs.shape
    out = _empty_nhwc(N, C, H, W, obs.device)
    _cabi.call("b200rl_obs_u8_to_f32_nhwc", obs.data_ptr(), out.data_ptr(), N, C, H, W, float(divisor), _stream())
    return out


def maxpool3x3s2_fwd(u, bias=None, with_relu=False):
    """``p = max_pool2d(u + bias, 3, 2, 1)``; returns ``(p, relu(p) or None, code)`` (``code``: uint8 argmax tap)."""
    _require_cuda(u)
    _nhwc(u, "maxpool3x3s2_fwd")
    N, C, H, W = u.shape
    OH, OW = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    p = _empty_nhwc(N, C, OH, OW, u.device)
    r = _empty_nhwc(N, C, OH, OW, u.device) if with_relu else None
    code = _empty_nhwc(N, C, OH, OW, u.device, torch.uint8)
    b = None if bias is None else _f32c(bias.detach())
    _cabi.call("b200rl_maxpool3x3s2_fwd", u.data_ptr(), _cabi.ptr(b), p.data_ptr(), _cabi.ptr(r), code.data_ptr(),
               N, H, W, C, _stream())
    return p, r, code


def maxpool3x3s2_bwd(g, code, in_shape):
    """Gradient w.r.t. the pool input ``[N,C,H,W]`` from ``g`` (pooled size) and the forward's ``code``."""
    _require_cuda(g, code)
    _nhwc(g, "maxpool3x3s2_bwd")
    N, C, H, W = in_shape
    du = _empty_nhwc(N, C, H, W, g.device)
    _cabi.call("b200rl_maxpool3x3s2_bwd", g.data_ptr(), code.data_ptr(), du.data_ptr(), N, H, W, C, _stream())
    return du


def bias_act_fwd(u, bias, res=None, want_y=True, want_r=False):
    """``v = u + bias[c] (+ res)``; returns ``(v or None, relu(v) or None)`` as requested."""
    _require_cuda(u, bias)
    _nhwc(u, "bias_act_fwd")
    if res is not None:
        _nhwc(res, "bias_act_fwd res")
        if res.shape != u.shape:
            raise RuntimeError("bias_act_fwd: res %s vs u %s" % (tuple(res.shape), tuple(u.shape)))
    if not (want_y or want_r):
        raise RuntimeError("bias_act_fwd: nothing to compute")
    N, C, H, W = u.shape
    y = torch.empty_like(u) if want_y else None
    r = torch.empty_like(u) if want_r else None
    b = _f32c(bias.detach())
    _cabi.call("b200rl_bias_act_fwd", u.data_ptr(), b.data_ptr(), _cabi.ptr(res), _cabi.ptr(y), _cabi.ptr(r),
               N * H * W, C, _glue_ctas(N * H * W * C // 4), _stream())
    return y, r


def relu_bwd_add(dy, dr, r):
    """``g = dy + dr * (r > 0)`` (either term may be ``None``) and the per-channel sum of ``g`` (the bias gradient of
    the op whose output ``g`` belongs to).  With ``dr is None`` no tensor is written: ``g`` is ``dy`` itself."""
    ref = dy if dy is not None else dr
    _require_cuda(ref)
    for t, nm in ((dy, "dy"), (dr, "dr"), (r if dr is not None else None, "r")):
        if t is not None:
            _nhwc(t, "relu_bwd_add " + nm)
            if t.shape != ref.shape:
                raise RuntimeError("relu_bwd_add: shape mismatch %s vs %s" % (tuple(t.shape), tuple(ref.shape)))
    N, C, H, W = ref.shape
    n_ctas = _glue_ctas(N * H * W * C // 4)
    partials = torch.empty((n_ctas, C), dtype=torch.float32, device=ref.device)
    g = torch.empty_like(ref) if dr is not None else None
    _cabi.call("b200rl_relu_bwd_add", _cabi.ptr(dy), _cabi.ptr(dr), _cabi.ptr(r if dr is not None else None),
               _cabi.ptr(g), partials.data_ptr(), N * H * W, C, n_ctas, _stream())
    return (g if g is not None else dy), partials.sum(0)


# ---- K10: 3x3 convolution weight gradient (TF32 tensor cores, fp32 accumulation) -----------------------------------
WGRAD_MAX_CTAS = 148              # one persistent CTA (24 warps, double-buffered tiles) per SM


def conv3x3_wgrad_supported(x_shape, cout):
    """Shapes ``b200rl_conv3x3_wgrad`` / ``_fwd`` take (the IMPALA-CNN's): 3x3/s1/p1, Cin 3/16/32, Cout 16/32, W in 8/16/32/64
    (W = 64: see ``_conv3x3_shape_ok``)."""
    return len(x_shape) == 4 and _conv3x3_shape_ok(x_shape[1], cout, x_shape[3])


def _conv3x3_shape_ok(cin, cout, W, wgrad=True):
    """Channel counts and image widths K10/K11 are instantiated for AND whose tiles fit one CTA's shared memory.  At
    W = 64 the weight gradient (K10) fits for the layers the IMPALA-CNN has there (3 -> 16/32 and 16 -> 16; the others
    would need 290-375 KB), the forward / input-gradient form (K11) for everything but 32 -> 32 (236 KB); the library
    refuses the rest with status -2 and they stay on cuDNN."""
    if cin not in (3, 16, 32) or cout not in (16, 32) or W not in (8, 16, 32, 64):
        return False
    if W < 64:
        return True
    return (cin == 3 or (cin, cout) == (16, 16)) if wgrad else (cin, cout) != (32, 32)


def conv3x3_wgrad(x, dy):
    """``d loss / d weight`` of ``conv2d(x, weight, padding=1)`` with a 3x3 kernel: ``x [N,Cin,H,W]``, ``dy [N,Cout,H,W]``
    (float32 channels_last) -> ``[Cout,Cin,3,3]`` in channels_last strides (physical ``[Cout,3,3,Cin]``)."""
    _require_cuda(x, dy)
    _nhwc(x, "conv3x3_wgrad x")
    _nhwc(dy, "conv3x3_wgrad dy")
    if not conv3x3_wgrad_supported(x.shape, dy.shape[1]) or x.shape[0] != dy.shape[0] or x.shape[2:] != dy.shape[2:]:
        raise RuntimeError("conv3x3_wgrad: unsupported shapes x %s dy %s" % (tuple(x.shape), tuple(dy.shape)))
    N, Cin, H, W = x.shape
    Cout = dy.shape[1]
    n_ctas = max(1, min(WGRAD_MAX_CTAS, N * ((H + 7) // 8)))
    partials = torch.empty((n_ctas, Cout, 3, 3, Cin), dtype=torch.float32, device=x.device)
    _cabi.call("b200rl_conv3x3_wgrad", x.data_ptr(), dy.data_ptr(), partials.data_ptr(), N, H, W, Cin, Cout, n_ctas,
               _stream())
    return partials.sum(0).permute(0, 3, 1, 2)


# ---- K11: the 3x3 convolution itself, forward and input gradient ----------------------------------------------------
def _w_cl(weight):
    if weight.dim() != 4 or tuple(weight.shape[2:]) != (3, 3) or weight.dtype != torch.float32:
        raise RuntimeError("conv3x3: expected a float32 [O,I,3,3] weight, got %s %s" % (tuple(weight.shape), weight.dtype))
    return weight.detach().contiguous(memory_format=torch.channels_last)        # physical [O][3][3][I]


def conv3x3_fwd(x, weight):
    """``conv2d(x, weight, None, padding=1)`` for a 3x3 kernel: float32 channels_last in and out."""
    _require_cuda(x, weight)
    _nhwc(x, "conv3x3_fwd x")
    w = _w_cl(weight)
    N, Cin, H, W = x.shape
    Cout = w.shape[0]
    if w.shape[1] != Cin or not _conv3x3_shape_ok(Cin, Cout, W, wgrad=False):
        raise RuntimeError("conv3x3_fwd: unsupported shapes x %s weight %s" % (tuple(x.shape), tuple(w.shape)))
    y = _empty_nhwc(N, Cout, H, W, x.device)
    _cabi.call("b200rl_conv3x3_fwd", x.data_ptr(), w.data_ptr(), y.data_ptr(), N, H, W, Cin, Cout, 0, _stream())
    return y


def conv3x3_dgrad(dy, weight):
    """Gradient of ``conv2d(x, weight, None, padding=1)`` w.r.t. ``x`` from ``dy [N,O,H,W]`` (channels_last)."""
    _require_cuda(dy, weight)
    _nhwc(dy, "conv3x3_dgrad dy")
    w = _w_cl(weight)
    N, O, H, W = dy.shape
    I = w.shape[1]
    if w.shape[0] != O or I == 3 or not _conv3x3_shape_ok(O, I, W, wgrad=False):   # runs as a forward O -> I
        raise RuntimeError("conv3x3_dgrad: unsupported shapes dy %s weight %s" % (tuple(dy.shape), tuple(w.shape)))
    dx = _empty_nhwc(N, I, H, W, dy.device)
    _cabi.call("b200rl_conv3x3_fwd", dy.data_ptr(), w.data_ptr(), dx.data_ptr(), N, H, W, O, I, 1, _stream())
    return dx


def conv3x3_fwd_fused(x, weight, bias, res=None, want_y=True, want_r=False):
    """K11 forward with the residual block's tail in the epilogue: ``v = conv2d(x, weight, bias, padding=1) (+ res)``;
    returns ``(v or None, relu(v) or None)`` as requested (reference ``models/common.py:30-34``)."""
    _require_cuda(x, weight, bias)
    _nhwc(x, "conv3x3_fwd_fused x")
    w = _w_cl(weight)
    N, Cin, H, W = x.shape
    Cout = w.shape[0]
    if w.shape[1] != Cin or not _conv3x3_shape_ok(Cin, Cout, W, wgrad=False) or not (want_y or want_r):
        raise RuntimeError("conv3x3_fwd_fused: unsupported shapes x %s weight %s" % (tuple(x.shape), tuple(w.shape)))
    if res is not None:
        _nhwc(res, "conv3x3_fwd_fused res")
        if tuple(res.shape) != (N, Cout, H, W):
            raise RuntimeError("conv3x3_fwd_fused: res %s vs output %s" % (tuple(res.shape), (N, Cout, H, W)))
    b = _f32c(bias.detach())
    y = _empty_nhwc(N, Cout, H, W, x.device) if want_y else None
    r = _empty_nhwc(N, Cout, H, W, x.device) if want_r else None
    _cabi.call("b200rl_conv3x3_fwd_fused", x.data_ptr(), w.data_ptr(), b.data_ptr(), _cabi.ptr(res), _cabi.ptr(y),
               _cabi.ptr(r), N, H, W, Cin, Cout, _stream())
    return y, r
