// Per-sample categorical statistics shared by the fused IMPALA (K1) and PPO (K3) losses.
// One thread owns one (b,t) sample: its A logits and A behaviour logits sit in shared
// memory (row stride A words; A odd, e.g. 15, is bank-conflict free), are pulled into
// registers once and stay there for the backward half.
#pragma once
#include "common.cuh"

namespace b200rl {

template <int AT>
struct SampleStats {
    float x[AT];   // logits (register copy)
    float e[AT];   // exp(x - max)
    float lse;     // logsumexp(x)
    float inv_se;  // 1 / sum(e)
    float H;       // entropy  -sum p log p      (torch Categorical.entropy)
    float kl;      // sum p (log p - log p_ref)  (torch kl_divergence(Categorical, Categorical))
    float lp;      // log pi(a)
    float rho;     // exp(log pi(a) - log pi_ref(a))
    int a;
};

// sx, sy: this sample's rows in shared memory.
template <int AT>
__device__ __forceinline__ void sample_stats(SampleStats<AT> &st, const float *sx, const float *sy,
                                             int A, long long action) {
    float y[AT];
    float m = -INFINITY, mr = -INFINITY;
#pragma unroll
    for (int j = 0; j < AT; ++j) {
        st.x[j] = j < A ? sx[j] : -INFINITY;
        y[j] = j < A ? sy[j] : -INFINITY;
        m = fmaxf(m, st.x[j]);
        mr = fmaxf(mr, y[j]);
    }
    float se = 0.f, ser = 0.f;
    const float mk = m * kLog2e, mrk = mr * kLog2e;
#pragma unroll
    for (int j = 0; j < AT; ++j) {
        st.e[j] = fast_exp2(fmaf(st.x[j], kLog2e, -mk));  // padded lanes: exp2(-inf) = 0
        se += st.e[j];
        ser += fast_exp2(fmaf(y[j], kLog2e, -mrk));
    }
    st.lse = fmaf(fast_log2(se), kLn2, m);
    const float lser = fmaf(fast_log2(ser), kLn2, mr);
    st.inv_se = __fdividef(1.f, se);
    float H = 0.f, kl = 0.f;
#pragma unroll
    for (int j = 0; j < AT; ++j) {
        if (j < A) {
            const float p = st.e[j] * st.inv_se;
            const float L = st.x[j] - st.lse;
            const float Lr = y[j] - lser;
            H = fmaf(-p, L, H);
            kl = fmaf(p, L - Lr, kl);
        }
    }
    st.H = H;
    st.kl = kl;
    int a = (int)action;
    a = a < 0 ? 0 : (a >= A ? A - 1 : a);  // never index outside the row
    st.a = a;
    st.lp = sx[a] - st.lse;
    const float lpr = sy[a] - lser;
    st.rho = fast_exp2((st.lp - lpr) * kLog2e);
}

// d loss / d logits for one sample, written over its shared-memory row:
//   g_lp * (onehot(a) - p) + c_ent * p * (log p + H)
// (c_ent = entropy_coeff / N; the second term is -c_ent * dH/dlogits.)
template <int AT>
__device__ __forceinline__ void write_dlogits(const SampleStats<AT> &st, float *srow, int A,
                                              float g_lp, float c_ent) {
#pragma unroll
    for (int j = 0; j < AT; ++j) {
        if (j < A) {
            const float p = st.e[j] * st.inv_se;
            const float L = st.x[j] - st.lse;
            const float onehot = (j == st.a) ? 1.f : 0.f;
            srow[j] = fmaf(g_lp, onehot - p, c_ent * p * (L + st.H));
        }
    }
}

}  // namespace b200rl
