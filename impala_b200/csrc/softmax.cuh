// Per-sample categorical statistics shared by the fused IMPALA (K1) and PPO (K3) losses.
// One thread owns one (b,t) sample: its A logits and A behaviour logits sit in shared
// memory (row stride A words; A odd, e.g. 15, is bank-conflict free), are pulled into
// registers once and stay there for the backward half.
//
// The kernels are issue-slot bound at BASELINE sizes (16.7 MB per launch), so the
// arithmetic is arranged for the fewest instructions per logit.  With
//   m = max x,  e_j = exp(x_j - m),  se = sum e,  p_j = e_j / se,  lse = m + ln se,  xbar = sum p_j x_j
// everything the losses need follows from three running sums (no per-logit log-probability is formed):
//   log p_j            = x_j - lse
//   entropy H          = lse - xbar                           (torch Categorical.entropy)
//   KL(p || p_ref)     = sum p_j (x_j - y_j) - lse_x + lse_y  (torch kl_divergence)
//   -c dH/dx_j         = c p_j (log p_j + H) = c p_j (x_j - xbar)
#pragma once
#include "common.cuh"

namespace b200rl {

template <int AT>
struct SampleStats {
    float d[AT];   // x_j (the logits, register copy)
    float e[AT];   // exp(x_j - max x)
    float inv_se;  // 1 / sum e
    float dbar;    // xbar = sum p_j x_j
    float H;       // entropy
    float kl;      // KL(p || p_ref)
    float lp;      // log pi(a)
    float rho;     // exp(log pi(a) - log pi_ref(a))
    int a;
};

// sx, sy: this sample's rows in shared memory.  EXACT: A == AT at compile time (no padding
// predicates); otherwise columns j >= A are skipped.
template <int AT, bool EXACT>
__device__ __forceinline__ void sample_stats(SampleStats<AT> &st, const float *sx, const float *sy,
                                             int A_rt, long long action) {
    const int A = EXACT ? AT : A_rt;
    float y[AT];
    float m = -INFINITY, mr = -INFINITY;
#pragma unroll
    for (int j = 0; j < AT; ++j) {
        if (EXACT || j < A) {
            st.d[j] = sx[j];
            y[j] = sy[j];
            m = fmaxf(m, st.d[j]);
            mr = fmaxf(mr, y[j]);
        } else {
            st.d[j] = 0.f;
            y[j] = 0.f;
        }
    }
    float se = 0.f, ser = 0.f, s1 = 0.f, s2 = 0.f;
    const float mk = m * kLog2e, mrk = mr * kLog2e;  // both exponents formed the same way: identical rows give rho == 1 exactly
#pragma unroll
    for (int j = 0; j < AT; ++j) {
        if (EXACT || j < A) {
            const float x = st.d[j];
            const float ej = fast_exp2(fmaf(x, kLog2e, -mk));
            se += ej;
            ser += fast_exp2(fmaf(y[j], kLog2e, -mrk));
            s1 = fmaf(ej, x, s1);
            s2 = fmaf(ej, x - y[j], s2);
            st.e[j] = ej;
        } else {
            st.e[j] = 0.f;
        }
    }
    // __fmul_rn: never contracted into an FMA, so log pi(a) and log pi_ref(a) are formed by the
    // same instruction sequence and identical rows give rho == 1 exactly (the clip tie)
    const float lnse = __fmul_rn(fast_log2(se), kLn2), lnser = __fmul_rn(fast_log2(ser), kLn2);
    st.inv_se = __fdividef(1.f, se);
    st.dbar = s1 * st.inv_se;
    st.H = (m + lnse) - st.dbar;
    st.kl = fmaf(s2, st.inv_se, (mr - m) + (lnser - lnse));
    int a = (int)action;
    a = a < 0 ? 0 : (a >= A ? A - 1 : a);  // never index outside the row
    st.a = a;
    st.lp = (sx[a] - m) - lnse;
    const float lpr = (sy[a] - mr) - lnser;
    st.rho = fast_exp2((st.lp - lpr) * kLog2e);
}

// d loss / d logits for one sample, written over its shared-memory row:
//   g_lp * (onehot(a) - p) + c_ent * p * (log p + H)  =  e_j * (c1 * x_j + k1)  (+ g_lp at j = a)
// (c_ent = entropy_coeff / N; the second term is -c_ent * dH/dlogits.)
template <int AT, bool EXACT>
__device__ __forceinline__ void write_dlogits(const SampleStats<AT> &st, float *srow, int A_rt,
                                              float g_lp, float c_ent) {
    const int A = EXACT ? AT : A_rt;
    if (!EXACT && A == 1) {  // one action: p = 1, every term cancels; exactly 0 instead of the factored form's rounding residue
        srow[0] = 0.f;
        return;
    }
    const float c1 = c_ent * st.inv_se;
    const float k1 = -fmaf(c_ent, st.dbar, g_lp) * st.inv_se;
#pragma unroll
    for (int j = 0; j < AT; ++j)
        if (EXACT || j < A) srow[j] = st.e[j] * fmaf(c1, st.d[j], k1);
    srow[st.a] += g_lp;
}

}  // namespace b200rl
