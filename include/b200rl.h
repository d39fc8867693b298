/* b200rl -- C ABI of the B200-native learner hot path (drop-in for d3sm0/impala).
 *
 * The reference has no FFI: its "operator API" is plain Python calls (SURVEY.md
 * sec 8b).  Every entry point below therefore names the reference Python site it
 * replaces; impala_b200/_cabi.py is the ctypes binding and INTEGRATION.md shows
 * the stub a reference maintainer would add.
 *
 * Conventions
 *  - all pointers are DEVICE pointers unless named host_*; the caller (PyTorch)
 *    owns and allocates every buffer, outputs and workspace included; the
 *    library never allocates, frees or retains a pointer;
 *  - launches go to `stream` (a cudaStream_t passed as void*), are asynchronous
 *    and CUDA-graph capturable;
 *  - return 0 = ok, <0 = argument error (checked on the host before launch),
 *    >0 = cudaError_t of the launch; text via b200rl_last_error() (thread local);
 *  - layouts: a [B,T] field is addressed as b*stride_b + t*stride_t (elements);
 *    supported: batch-major (stride_b=T, stride_t=1, the reference's layout,
 *    agents/impala/learning.py:142-145) and time-major (stride_b=1, stride_t=B).
 *    Per-action tensors ([.., A]) use the same sample index times A.
 *  - lambda_ is a double so that (1 - lambda_) is rounded to fp32 once, exactly as the
 *    reference's Python-float arithmetic does (bit-exact scans against the oracle);
 *  - there is no CPU fallback and no other backend: sm_100a only.
 */
#ifndef B200RL_H
#define B200RL_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define B200RL_VERSION 100

/* flags */
#define B200RL_DETACH_ADV 1u  /* treat pg advantages as constants (textbook IMPALA); reference: not detached, learning.py:155 */
#define B200RL_DETACH_RHO 2u  /* treat rho as a constant inside V-trace; reference: not detached, learning.py:148 */
#define B200RL_ORDER_RLAX 4u  /* b200rl_vtrace output order (errors, pg_adv, q) instead of call-site (pg_adv, errors, q) */
/* Fused losses are launched with programmatic dependent launch (PDL).  By default the kernel waits
 * for the preceding kernel of the stream before it reads anything (it may ask L2 to PREFETCH its inputs
 * earlier -- a hint, never a read).  With B200RL_INPUTS_STABLE the caller promises that NO buffer of
 * this call except `workspace` and `scalars` is written or read by ANY earlier launch of the stream
 * that may still be running (every fused loss lets its successor start early, so launch N+2 can
 * overlap launch N; e.g. a pipeline over independent rollout batches): loads, arithmetic and gradient
 * stores then overlap the earlier launches' tails and only the scalar reduction waits.  K1 also takes
 * the flag as "throughput matters, not latency" and runs its many-small-CTAs implementation. */
#define B200RL_INPUTS_STABLE 8u
/* Profiling aid (K1): stamps of %globaltimer (ns) at 7 phase boundaries into workspace + b200rl_trace_offset_bytes(),
 * layout uint64 [row][16] (slots 0-6 %globaltimer, 7 %smid, 8-14 clock64 of the same points).  Tile kernel: one row per
 * CTA < 4096 (thread 0).  Persistent kernel: one row per warp group (row = CTA * groups + group < 2048, the group's
 * first thread, first unit): 0 past the dependency wait, 1 chunk landed in shared memory, 2 softmax statistics done,
 * 3 V-trace recursion done, 4 gradients written and the chunk's bulk store issued, 5 reduction done, 6 bulk stores drained;
 * plus CTA rows 2048 + 1024 * (launch generation & 1) + CTA: 0 last warp at the reduction, 1 published, 2 kernel entry,
 * 3 past the dependency wait, 4 thread 0's last instruction (scripts/k1_trace.py prints the timeline). */
#define B200RL_TRACE 0x80000000u

/* slots of the float[8] `scalars` output of the fused losses */
#define B200RL_S_LOSS 0
#define B200RL_S_ENTROPY 1
#define B200RL_S_TD 2
#define B200RL_S_PG 3
#define B200RL_S_KL 4
#define B200RL_S_RATIO 5
#define B200RL_S_TARGET 6 /* PPO only: mean(v_target), losses.py:152 */

int32_t b200rl_version(void);
const char *b200rl_last_error(void);

/* Bytes of zero-initialised device workspace the fused losses need.  Zero it ONCE and never touch it again: it holds
 * the launch generation the scalar reduction tags its partial sums with.  One workspace per stream.  A reduction that
 * times out (~1 s without progress: the GPU is wedged) reports NaN scalars and poisons the workspace -- every later launch
 * on it reports NaN until it is zeroed again. */
int64_t b200rl_loss_workspace_bytes(void);
int64_t b200rl_trace_offset_bytes(void);

/* K1 -- replaces ImpalaLearner._train_step's loss + backward w.r.t. model outputs
 * (reference agents/impala/learning.py:146-160,167-168) and, inside it,
 * batched_vtrace = vmap(rlego.vtrace_td_error_and_advantage) (learning.py:15-26).
 * Inputs [B,T,A] / [B,T] (see layouts); T >= 2.  n_global_b = number of columns of
 * the GLOBAL batch (== B unless sharded along B across ranks): means are taken
 * over n_global_b*(T-1) resp. n_global_b*T, so per-rank scalars and gradients sum
 * to the single-GPU result.  Outputs: dlogits [B,T,A], dvalues [B,T] =
 * d loss / d logits, d loss / d values; scalars[8] (B200RL_S_*).
 * Limits (status -3 and a message, nothing launched): A <= 32; T <= 640 (256 when A > 16); and, for the tile kernel
 * (time-major inputs, B200RL_INPUTS_STABLE), B <= 16384 * floor(640 / T) columns per call (256 / T when A > 16) -- a CTA
 * holds one thread per sample and the workspace one partial sum per CTA for 16384 CTAs. */
int32_t b200rl_vtrace_impala_loss_fwdbwd(
    const float *logits, const float *pi_ref, const int64_t *actions, const float *rewards,
    const float *discounts, const float *values,
    int64_t B, int64_t T, int64_t A, int64_t stride_b, int64_t stride_t, int64_t n_global_b,
    double lambda_, float clip_rho, float clip_pg_rho, float entropy_coeff, uint32_t flags,
    float *dlogits, float *dvalues, float *scalars, void *workspace, void *stream);

/* K1 without any allocation.  NO counterpart in the reference: added for callers
 * that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same
 * arguments, kernels and results as b200rl_vtrace_impala_loss_fwdbwd minus `workspace`: the scalar reduction's scratch is a
 * static device array inside the library, B200RL_NOALLOC_SCRATCH_BYTES (1 MiB + 256 B) on every GPU that loads the library,
 * held until the process exits.  Limits (rejected on the host with status -3 and a message): B <= B200RL_NOALLOC_MAX_COLUMNS
 * batch columns per call; no B200RL_TRACE.  All calls on a device share the scratch and find the previous call's contents:
 * the entry point itself clears what must start at zero (one 1-thread launch ahead of the loss kernel, in stream order).
 * Calls that OVERLAP on the GPU (different streams or threads) race on the scratch and get undefined scalars[] -- such
 * callers use the ordinary entry point with one workspace per stream.  Calls no cudaMalloc / cudaFree / allocating routine. */
#define B200RL_NOALLOC_SCRATCH_BYTES 1048832
#define B200RL_NOALLOC_MAX_COLUMNS 16384
int32_t b200rl_vtrace_impala_loss_fwdbwd_noalloc(
    const float *logits, const float *pi_ref, const int64_t *actions, const float *rewards,
    const float *discounts, const float *values,
    int64_t B, int64_t T, int64_t A, int64_t stride_b, int64_t stride_t, int64_t n_global_b,
    double lambda_, float clip_rho, float clip_pg_rho, float entropy_coeff, uint32_t flags,
    float *dlogits, float *dvalues, float *scalars, void *stream);

/* Stand-alone drop-in for batched_vtrace(v_tm1, v_t, r_t, discount_t, rho_tm1)
 * (reference agents/impala/learning.py:18,150-153).  The five inputs are [B,Tm]
 * views sharing (stride_b, stride_t): batch-major (stride_t=1, stride_b>=Tm -- e.g.
 * the reference's values[:, :-1] slices of a [B,T] tensor) or time-major
 * (stride_b=1, stride_t>=B).  Outputs adv, err, q are COMPACT [B,Tm] in the same
 * major order (order fixed here; the Python wrapper applies B200RL_ORDER_RLAX). */
int32_t b200rl_vtrace(
    const float *v_tm1, const float *v_t, const float *r_t, const float *discount_t,
    const float *rho_tm1, int64_t B, int64_t Tm, int64_t stride_b, int64_t stride_t,
    double lambda_, float clip_rho, float clip_pg_rho,
    float *adv, float *err, float *q, void *stream);

/* Backward of b200rl_vtrace under stop_target_gradients=True: given upstream
 * g_adv, g_err, g_q (compact, any may be NULL = zero) and the saved compact q,
 * produce compact d v_tm1, d v_t, d rho.  v_tm1, discount_t, rho_tm1 use the
 * forward's input strides. */
int32_t b200rl_vtrace_bwd(
    const float *v_tm1, const float *discount_t, const float *rho_tm1, const float *q,
    const float *g_adv, const float *g_err, const float *g_q,
    int64_t B, int64_t Tm, int64_t stride_b, int64_t stride_t,
    double lambda_, float clip_pg_rho,
    float *d_v_tm1, float *d_v_t, float *d_rho, void *stream);

/* K2 -- replaces rlego.lambda_returns(r_t, discount_t, v_t, lambda_) per rollout
 * (reference agents/ppo/learning.py:65-69), batched over B columns.
 * G_t = r_t + g_t((1-l) v_t[t] + l G_{t+1}), bootstrap v_t[T-1]. */
int32_t b200rl_lambda_returns(
    const float *r_t, const float *discount_t, const float *v_t,
    int64_t B, int64_t T, int64_t stride_b, int64_t stride_t, double lambda_,
    float *out, void *stream);

/* GAE (rlego.truncated_generalized_advantage_estimation, reference losses.py:11).
 * values holds T+1 steps per column with its own strides. */
int32_t b200rl_gae(
    const float *r_t, const float *discount_t, const float *values,
    int64_t B, int64_t T, int64_t stride_b, int64_t stride_t,
    int64_t vstride_b, int64_t vstride_t, double lambda_, float *out, void *stream);

/* N4 -- rlego.n_step_bootstrapped_returns(r_t, discount_t, v_t, n, lambda_t) (reference
 * agents/dqn/learning.py:62,79), batched over B columns; out[b,t] bootstraps from v_t[b, t+n-1]
 * (v_t[b, T-1] past the end). */
int32_t b200rl_n_step_returns(
    const float *r_t, const float *discount_t, const float *v_t,
    int64_t B, int64_t T, int64_t stride_b, int64_t stride_t, int64_t n, double lambda_,
    float *out, void *stream);

/* N4b -- quantile_regression_loss(dist_src [B,N], tau_src [B,N], dist_target [B,M], huber_param) of reference
 * models/quantile_layers.py:83-93 (caller agents/dqn/learning.py:209) and its gradient w.r.t. dist_src:
 * loss [B] = mean_j sum_i |tau_i - 1[delta<0]| * huber(delta_ij), delta_ij = target_j - src_i;
 * d_src [B,N] = d loss_b / d src_i (the wrapper scales it by the upstream gradient). */
int32_t b200rl_quantile_huber_fwdbwd(
    const float *dist_src, const float *tau_src, const float *dist_target, int64_t B, int64_t N, int64_t M,
    float huber_param, float *loss, float *d_src, void *stream);

/* N3b -- the prioritized-replay-weighted TD loss of ApexLearner._train_step after q = model(obs) (reference
 * agents/dqn/learning.py:161-170) with its backward w.r.t. q (:169) and the new priorities (:172), one launch:
 * q [N,A], actions / target / probabilities [N] (the probabilities the PER sampler returned, N3);
 * w_i = p_i^(-exponent) / max_j p_j^(-exponent); loss = 0.5 mean((target - q[i,a_i])^2 w_i);
 * dq [N,A] = d loss / d q (one non-zero per row), priorities [N] = |target - q[i,a_i]|, weights [N] (may be NULL);
 * scalars[8] = loss, mean q (:188), mean target (:189), mean priority (:190), max_j p_j^(-exponent), 0, 0, 0.
 * `workspace` as for the fused losses. */
int32_t b200rl_per_td_loss_fwdbwd(
    const float *q, const int64_t *actions, const float *target, const float *probabilities, int64_t N, int64_t A,
    float importance_sampling_exponent, float *dq, float *priorities, float *weights, float *scalars, void *workspace,
    void *stream);

/* K3 -- replaces losses.ppo_loss after model(s) + backward w.r.t. model outputs
 * (reference losses.py:134-155, caller agents/ppo/learning.py:132-134).
 * Flat minibatch: logits, pi_ref [N,A]; actions, v_target, v [N].
 * scalars[B200RL_S_KL] is the UNclamped mean KL (clamp_min(0) belongs after the
 * cross-rank sum). */
int32_t b200rl_ppo_loss_fwdbwd(
    const float *logits, const float *pi_ref, const int64_t *actions, const float *v_target,
    const float *v, int64_t N, int64_t A, int64_t n_global,
    float clip_coeff, float entropy_cost, uint32_t flags,
    float *dlogits, float *dv, float *scalars, void *workspace, void *stream);

/* K7 -- restates the intent of reference losses.py:62-89 (`ppo_loss_is`: the PPO clip on
 * V-trace advantages; not executable as committed -- functorch is not imported and
 * batched_vtrace is undefined in that module) after the two model calls, plus its backward.
 * Batch-major [B,T,...]: logits, pi_ref [B,T,A]; actions i64, rewards f32, done u8 (0/1),
 * values (= v_tm1) f32 [B,T]; v_boot [B] = the value of s1[:, -1] (no gradient).  In-kernel:
 * discount = (1-done)*gamma (:65), mask = roll(1-done, 1, dim=1) (:64), v_t = cat(values[:,1:],
 * v_boot) (:70-72), V-trace with rho detached over all T steps (:73), td = 0.5 mean (mask err)^2
 * (:74), pg = mean max(-adv rho, -clamp(rho, 1-c, 1+c) adv) with adv = (adv mask).detach()
 * (:75-78), loss = pg + td - entropy_cost * H (:81).  Means run over n_global_b * T samples.
 * scalars[B200RL_S_KL] is the UNclamped mean KL as in K3. */
int32_t b200rl_ppo_is_loss_fwdbwd(
    const float *logits, const float *pi_ref, const int64_t *actions, const float *rewards,
    const uint8_t *done, const float *values, const float *v_boot, int64_t B, int64_t T, int64_t A,
    int64_t n_global_b, float gamma, double lambda_, float clip_rho, float clip_pg_rho,
    float clip_coeff, float entropy_cost, uint32_t flags,
    float *dlogits, float *dvalues, float *scalars, void *workspace, void *stream);

/* K4 -- replaces rlmeta UniformSampler (reference agents/impala/builder.py:31-32,
 * consumer learning.py:121).  Philox4x32-10, see oracle/replay_ref.py for the
 * stream definition; out_idx[n] int64 in [0,size). */
int32_t b200rl_sample_uniform(
    uint64_t seed, uint32_t stream_id, uint64_t draw_offset, int64_t n, int64_t size,
    int64_t *out_idx, void *stream);

/* K5 -- replaces CircularBuffer.__getitem__ + collate_nested(torch.stack) + .to(device)
 * (reference agents/impala/learning.py:121-123,142): dst[i] = src[idx[i]], rows of
 * row_bytes bytes.  Out-of-range indices are clamped into [0,nrows) (never faults). */
int32_t b200rl_gather_rows(
    const void *src, int64_t row_bytes, const int64_t *idx, int64_t n, int64_t nrows,
    void *dst, void *stream);

/* Replaces ReplayBuffer.append/extend storage writes (reference
 * agents/impala/learning.py:75, agents/ppo/learning.py:63): rows j=0..n-1 go to
 * slot (count + j) % capacity.  n <= capacity. */
int32_t b200rl_ring_append(
    void *storage, int64_t row_bytes, int64_t capacity, int64_t count,
    const void *rows, int64_t n, void *stream);

/* N2 -- device-side rollout ingestion: replaces ImpalaActor._make_replay + replay.async_append
 * (reference agents/impala/learning.py:71-80; PPO: agents/ppo/learning.py:59-69): all fields of n stacked
 * rollouts go to their ring storages (slot (count + j) % capacity) in ONE launch, and a field of kind
 * B200RL_FIELD_DONE_TO_DISCOUNT converts uint8/bool done flags to float32 `logical_not(done) * gamma`
 * on the way (row_bytes = flags per row; the destination row holds row_bytes floats).
 * `fields` is a HOST array (copied into the launch parameters). */
#define B200RL_FIELD_COPY 0
#define B200RL_FIELD_DONE_TO_DISCOUNT 1
typedef struct {
    const void *src;    /* device, n rows of row_bytes bytes */
    void *dst;          /* device ring storage of `capacity` rows */
    int64_t row_bytes;  /* bytes per source row */
    int32_t kind;       /* B200RL_FIELD_* */
    float gamma;        /* DONE_TO_DISCOUNT only */
} b200rl_field_t;
int32_t b200rl_ring_append_fields(
    const b200rl_field_t *fields, int32_t n_fields, int64_t capacity, int64_t count, int64_t n, void *stream);

/* Multi-field K5: dst_f[i] = src_f[idx[i]] for every field f in ONE launch (src = ring storage of nrows
 * rows, dst = the [n, ...] batch tensor; kind must be B200RL_FIELD_COPY).  Replaces the per-field
 * CircularBuffer.__getitem__ + collate of a sampled batch (reference agents/impala/learning.py:121-123,142). */
int32_t b200rl_gather_fields(
    const b200rl_field_t *fields, int32_t n_fields, const int64_t *idx, int64_t n, int64_t nrows, void *stream);

/* K4 + K5 in ONE launch -- replaces ReplayBuffer.sample(n) with the uniform sampler (reference agents/impala/learning.py:121,
 * agents/ppo/learning.py:112; rlmeta UniformSampler + CircularBuffer built at agents/impala/builder.py:30-36): draws the n
 * indices of b200rl_sample_uniform(seed, stream_id, draw_offset, n, size) -- the same Philox stream, bit for bit -- writes
 * them to out_idx and gathers those rows of every field (fields[f].src = storage [nrows, row_bytes], .dst = batch
 * [n, row_bytes]).  1 <= size <= nrows. */
int32_t b200rl_sample_gather_fields(const b200rl_field_t *fields, int32_t n_fields, uint64_t seed, uint32_t stream_id,
                                    uint64_t draw_offset, const uint64_t *draw_counter, int64_t n, int64_t size,
                                    int64_t nrows, int64_t *out_idx, void *stream);

/* `draw_counter` above (nullable, device memory) is added to draw_offset inside the kernel: a draw offset passed as an
 * argument would be frozen into a CUDA graph.  b200rl_counter_add(counter, delta) advances such a counter in stream order
 * (one thread); a captured learner step is sample_gather_fields(draw_counter) -> counter_add(counter, n) -> ... */
int32_t b200rl_counter_add(uint64_t *counter, int64_t delta, void *stream);

/* N3 -- prioritized sampling (replaces rlmeta PrioritizedSampler, reference agents/dqn/builder.py:83-90;
 * extend/sample/update at agents/dqn/learning.py:54,122,184).  `tree` is a caller-owned, zero-initialised
 * fp64 array of 2*leaves nodes (leaves = power of two >= capacity; node 1 = root, leaf j at leaves + j).
 * update: tree leaf idx[i] = value[i] (the already exponentiated priority), ancestors re-summed.
 * sample: n draws of stream (seed, stream_id) starting at draw_offset -> out_idx in [0,size), out_prob =
 * leaf / total (may be NULL).  Stream definition: oracle/replay_ref.py (bit-exact). */
int32_t b200rl_sumtree_update(
    double *tree, int64_t leaves, const int64_t *idx, const double *value, int64_t n, void *stream);
int32_t b200rl_sumtree_sample(
    const double *tree, int64_t leaves, int64_t size, uint64_t seed, uint32_t stream_id, uint64_t draw_offset,
    int64_t n, int64_t *out_idx, float *out_prob, void *stream);

/* N5 -- actor epilogue of AtariPPOModel.act (reference models/distributed_models.py:26-31): per row of
 * logits [n,A]: action = deterministic ? argmax : categorical sample of softmax(logits).  `deterministic` is a
 * uint8/bool array read with stride 0 (one flag for all rows, as the reference's 1-element tensor) or 1.
 * Draw specification: oracle/losses_ref.py:act_epilogue (Philox4x32-10, counter word 3 = 1). */
int32_t b200rl_act_epilogue(
    const float *logits, int64_t n, int64_t A, const uint8_t *deterministic, int64_t deterministic_stride,
    uint64_t seed, uint32_t stream_id, uint64_t draw_offset, int64_t *action, void *stream);

/* K6 -- replaces clip_grad_norm_(params, max_norm) + optimizer.step() (torch.optim.Adam, no weight decay,
 * no amsgrad) at reference agents/impala/learning.py:172-176 and agents/ppo/learning.py:136-140 for
 * parameters, gradients and Adam moments held in FLAT fp32 buffers of n elements.  In place: grads are
 * scaled by min(1, max_norm / (||g||_2 + 1e-6)) (max_norm <= 0: no clipping), then exp_avg, exp_avg_sq and
 * params take one Adam step; `step` is the 1-based step number (bias correction).  total_norm[0] receives
 * ||g||_2 before clipping.  `workspace` is the zero-initialised loss workspace (a separate region of it).
 * One cooperative launch (grid <= co-resident CTAs; CTAs loop over their chunk, so any n works).
 * `device_counters` (nullable; int64[2] in device memory, zeroed once by the caller) makes the launch CUDA-graph
 * replayable: a step number passed as an argument would be frozen into the graph, so with the pointer given the kernel
 * uses device_counters[0] + 1 as the step (`step` is then ignored), forms the bias corrections itself and stores the new
 * step back; K8 does the same with device_counters[1] for its flag generation. */
int32_t b200rl_clip_adam_step(
    float *params, float *grads, float *exp_avg, float *exp_avg_sq, int64_t n,
    float max_norm, double lr, double beta1, double beta2, double eps, int64_t step, int64_t *device_counters,
    float *total_norm, void *workspace, void *stream);

/* K8 -- K6 with the gradient allreduce fused in (SURVEY sec 8f N1 "fused with the allreduce epilogue"; replaces the
 * dist.all_reduce + clip_grad_norm_ + optimizer.step() of a data-parallel learner step).  One process per GPU;
 * every rank's bucket (n gradients followed by n_extra floats, e.g. the loss scalars) and flag area
 * (2 * 8 uint32, zeroed once) are mapped into every rank (b200rl_ipc_export / _import below).  The kernel
 * exchanges "gradients final" flags, reads its chunk from ALL buckets over NVLink with system-scope loads, adds
 * them in rank order (bit-identical on every rank), then norm -> clip -> Adam as K6; grads_out receives the
 * reduced, clipped gradient (must not alias a bucket), extra_out the n_extra tail sums.  The kernel returns only
 * after every peer has finished reading this rank's bucket.  `step` is Adam's step (bias correction);
 * `flag_generation` is the value the cross-rank flags carry: it belongs to the flag areas (zeroed once when they are
 * created), starts at 1 and every rank advances it by exactly one per call -- independent of the optimizer's step, so an
 * optimizer that is recreated or resumed from a checkpoint cannot fall behind stale flags.  A rank that never arrives
 * turns total_norm into NaN after ~18 s instead of hanging, and the NaN propagates into the update (as torch's
 * clip_grad_norm_ would) instead of applying partial gradients.  peer_buckets / peer_flags are HOST arrays of `world` device pointers (<= 8).
 * Routes: one-shot (every rank reads every bucket in full; default for 2 ranks) and two-shot (reduce-scatter +
 * all-gather through every rank's peer-mapped reduced buffer; default for more ranks).  Two-shot needs
 * peer_reduced (every rank's reduced-gradient buffer of n floats, peer_reduced[rank] == grads_out) and
 * peer_cta_flags (every rank's per-CTA flag area, 1024 * 8 uint32, zeroed once); pass NULL for both to force
 * one-shot (grads_out is then any local buffer).  NVLS: when the buffers come from a multicast-capable symmetric
 * allocation, mc_buckets / mc_reduced are the MULTICAST addresses of the buckets / reduced buffers; the two-shot
 * route then fetches each part with one multimem.ld_reduce (the NVSwitch adds) and broadcasts it with one
 * multimem.st.  The switch's summation order is its own: results are identical on all ranks but equal the
 * rank-ordered sum only to rounding.  Pass NULL for unicast loads / stores. */
int32_t b200rl_peer_allreduce_clip_adam_step(
    float *params, const float *const *peer_buckets, uint32_t *const *peer_flags, float *const *peer_reduced,
    uint32_t *const *peer_cta_flags, const float *mc_buckets, float *mc_reduced, int32_t world, int32_t rank,
    float *grads_out, float *extra_out, int64_t n_extra, float *exp_avg, float *exp_avg_sq, int64_t n,
    float max_norm, double lr, double beta1, double beta2, double eps, int64_t step, int64_t flag_generation,
    int64_t *device_counters, float *total_norm, void *workspace, void *stream);

/* CUDA IPC plumbing for K8.  export: 64-byte handle of the cudaMalloc block `ptr` lives in + its byte offset in
 * that block.  import: map a peer's block (peer access enabled lazily) and return the peer's buffer address.
 * close: unmap (pass the pointer and offset import returned / was given). */
int32_t b200rl_ipc_export(const void *ptr, void *handle64, int64_t *offset);
int32_t b200rl_ipc_import(const void *handle64, int64_t offset, void **out_ptr);
int32_t b200rl_ipc_close(void *ptr, int64_t offset);

/* K9 -- memory-bound glue of the IMPALA-CNN between its cuDNN convolutions (reference models/common.py:8-83:
 * ResidualBlock :28-34, ImpalaCNNBlock :52-58, ImpalaCNNLarge :71-78; input scaling models/models.py:73).  All
 * activations are fp32 NHWC (torch channels_last), C a multiple of 4, 16-byte aligned.
 *  obs_u8_to_f32_nhwc: out[n,h,w,c] = (float)obs[n,c,h,w] * (1.0f / divisor) -- what torch's CUDA `x.float() / 255.`
 *                      computes (scalar divisors become reciprocal multiplies)  (C == 3, H*W % 4 == 0).
 *  maxpool3x3s2_fwd:   p = max_pool2d(u + bias[c], kernel 3, stride 2, padding 1) with torch's scan order and tie /
 *                      NaN rule; u is [N,H,W,C], p (and r = relu(p) when r != NULL) are [N,OH,OW,C],
 *                      OH = (H-1)/2+1; code holds one byte per output element: kh*3+kw of the selected tap.
 *                      bias may be NULL.
 *  maxpool3x3s2_bwd:   du[N,H,W,C] from g[N,OH,OW,C] and code; gather form, deterministic.
 *  bias_act_fwd:       v = u + bias[c] (+ res when res != NULL); y = v when y != NULL; r = relu(v) when r != NULL.
 *  relu_bwd_add:       v = (dy ? dy : 0) + (dr ? dr * (r > 0) : 0); g = v when g != NULL; partials[cta][C] =
 *                      per-CTA channel sums of v (the caller adds the n_ctas rows in order: the bias gradient).
 *  n_ctas = grid size of the two grid-stride kernels (<= 65535); C/4 must divide 256. */
int32_t b200rl_obs_u8_to_f32_nhwc(
    const uint8_t *obs, float *out, int64_t N, int64_t C, int64_t H, int64_t W, float divisor, void *stream);
int32_t b200rl_maxpool3x3s2_fwd(
    const float *u, const float *bias, float *p, float *r, uint8_t *code, int64_t N, int64_t H, int64_t W, int64_t C,
    void *stream);
int32_t b200rl_maxpool3x3s2_bwd(
    const float *g, const uint8_t *code, float *du, int64_t N, int64_t H, int64_t W, int64_t C, void *stream);
int32_t b200rl_bias_act_fwd(
    const float *u, const float *bias, const float *res, float *y, float *r, int64_t n_pix, int64_t C, int64_t n_ctas,
    void *stream);
int32_t b200rl_relu_bwd_add(
    const float *dy, const float *dr, const float *r, float *g, float *partials, int64_t n_pix, int64_t C,
    int64_t n_ctas, void *stream);

/* K10 -- weight gradient of a 3x3 / stride 1 / padding 1 convolution (every convolution of the reference's
 * ImpalaCNNLarge, models/common.py:12-26,41-45) for fp32 NHWC activations: x [N,H,W,Cin], dy [N,H,W,Cout];
 * TF32 tensor-core products, fp32 accumulation (torch's default arithmetic for these layers).  Each of the n_ctas
 * CTAs writes its partial dW as partials[cta][Cout][3][3][Cin]; the caller adds the rows in order.
 * Supported: Cin in {3,16,32}, Cout in {16,32}, W in {8,16,32,64} (the tile must fit 3x in 227 KB of shared memory). */
int32_t b200rl_conv3x3_wgrad(
    const float *x, const float *dy, float *partials, int64_t N, int64_t H, int64_t W, int64_t Cin, int64_t Cout,
    int64_t n_ctas, void *stream);

/* K11 -- the bias-free 3x3 / stride 1 / padding 1 convolution of the same layers (reference models/common.py:12-26,
 * 41-45), forward (dgrad = 0: y[N,H,W,Cout] from x[N,H,W,Cin]) and input gradient (dgrad = 1: x is dy [N,H,W,Cin'],
 * y is dx [N,H,W,Cout'] with Cin' = the layer's output and Cout' = its input channels).  `weight` is ALWAYS the
 * layer's weight as torch holds it in channels_last: physical [O][3][3][I]; dgrad reads it transposed and flipped.
 * fp32 NHWC, TF32 tensor-core products, fp32 accumulation.  Supported: in 3/16/32 (dgrad: 16/32), out 16/32,
 * W in {8,16,32,64}. */
int32_t b200rl_conv3x3_fwd(
    const float *x, const float *weight, float *y, int64_t N, int64_t H, int64_t W, int64_t Cin, int64_t Cout,
    int32_t dgrad, void *stream);

/* K11 forward with the residual block's elementwise tail in the epilogue (reference models/common.py:30-34):
 * v = conv(x, weight) + bias[o] (+ res when res != NULL);  y = v when y != NULL;  relu_out = max(v, 0) when
 * relu_out != NULL (at least one of the two).  `relu(conv0(.))` is (y, relu_out) = (NULL, r1); `conv1(.) + inputs`
 * passes res = inputs and may also ask for the ReLU that opens the next block. */
int32_t b200rl_conv3x3_fwd_fused(
    const float *x, const float *weight, const float *bias, const float *res, float *y, float *relu_out, int64_t N,
    int64_t H, int64_t W, int64_t Cin, int64_t Cout, void *stream);

#ifdef __cplusplus
}
#endif
#endif
