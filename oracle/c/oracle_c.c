/* Plain-C scalar restatement of the integer/byte side of the hot path and of the
 * time scans (TEST ORACLE -- never linked into the product; PARITY UNPINNED, see
 * oracle/__init__.py).  Built by oracle/c/Makefile into oracle/_build/liboracle_c.so.
 *
 * Follows: replay sample -> gather (reference agents/impala/builder.py:30-36,
 * agents/impala/learning.py:121-123,142); V-trace per sequence (rlego call site
 * agents/impala/learning.py:150-153, algorithm: rlax vtrace_td_error_and_advantage);
 * lambda-returns (agents/ppo/learning.py:68); GAE (losses.py:11, comment only).
 */
#include <stdint.h>
#include <string.h>
#include <math.h>

#define M0 0xD2511F53u
#define M1 0xCD9E8D57u
#define W0 0x9E3779B9u
#define W1 0xBB67AE85u

void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)M0 * c0, p1 = (uint64_t)M1 * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* draws draw_offset .. draw_offset+n-1 of stream (seed, stream_id); index = mulhi64(x64, size) */
int oracle_sample_uniform(uint64_t seed, uint32_t stream_id, uint64_t draw_offset,
                          int64_t n, int64_t size, int64_t *out) {
    if (n < 0 || size <= 0) return -1;
    uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    for (int64_t i = 0; i < n; ++i) {
        uint64_t g = draw_offset + (uint64_t)i;
        uint32_t ctr[4] = {(uint32_t)g, (uint32_t)(g >> 32), stream_id, 0u}, o[4];
        oracle_philox4x32_10(ctr, key, o);
        uint64_t x = (uint64_t)o[0] | ((uint64_t)o[1] << 32);
        out[i] = (int64_t)(((unsigned __int128)x * (unsigned __int128)(uint64_t)size) >> 64);
    }
    return 0;
}

int oracle_gather_rows(const uint8_t *src, int64_t row_bytes, const int64_t *idx, int64_t n,
                       int64_t nrows, uint8_t *dst) {
    for (int64_t i = 0; i < n; ++i) {
        if (idx[i] < 0 || idx[i] >= nrows) return -2;
        memcpy(dst + i * row_bytes, src + idx[i] * row_bytes, (size_t)row_bytes);
    }
    return 0;
}

/* ring append: rows keyed count..count+n-1 go to slot key % capacity */
int oracle_ring_append(uint8_t *storage, int64_t row_bytes, int64_t capacity, int64_t count,
                       const uint8_t *rows, int64_t n) {
    for (int64_t j = 0; j < n; ++j)
        memcpy(storage + ((count + j) % capacity) * row_bytes, rows + j * row_bytes, (size_t)row_bytes);
    return 0;
}

/* V-trace on B sequences of length Tm (= T-1), element strides (sb, st) shared by all arrays.
 * fp32, operation order of the rlax loop: err = td + (disc * c) * err. */
int oracle_vtrace(const float *v_tm1, const float *v_t, const float *r_t, const float *disc,
                  const float *rho, int64_t B, int64_t Tm, int64_t sb, int64_t st,
                  float lambda_, float clip_rho, float clip_pg_rho,
                  float *adv, float *err, float *q) {
    for (int64_t b = 0; b < B; ++b) {
        float e = 0.0f, tgt_next = 0.0f;
        for (int64_t t = Tm - 1; t >= 0; --t) {
            int64_t i = b * sb + t * st;
            float rh = rho[i];
            float c = fminf(rh, 1.0f) * lambda_;
            float td = fminf(rh, clip_rho) * (r_t[i] + disc[i] * v_t[i] - v_tm1[i]);
            e = td + disc[i] * c * e;
            float tgt = e + v_tm1[i];           /* stop-gradient target */
            float er = tgt - v_tm1[i];
            float tgt2 = er + v_tm1[i];         /* as recomputed by vtrace_td_error_and_advantage */
            float qb = (t == Tm - 1) ? v_t[i]
                                     : lambda_ * tgt_next + (1.0f - lambda_) * v_tm1[i + st];
            float qq = r_t[i] + disc[i] * qb;
            err[i] = er;
            q[i] = qq;
            adv[i] = fminf(rh, clip_pg_rho) * (qq - v_tm1[i]);
            tgt_next = tgt2;
        }
    }
    return 0;
}

/* G_t = r_t + g_t((1-l) v_t[t] + l G_{t+1}), G_T = v_t[T-1]; arrays [B x T] with strides */
int oracle_lambda_returns(const float *r_t, const float *disc, const float *v_t,
                          int64_t B, int64_t T, int64_t sb, int64_t st, float lambda_, float *out) {
    for (int64_t b = 0; b < B; ++b) {
        float g = v_t[b * sb + (T - 1) * st];
        for (int64_t t = T - 1; t >= 0; --t) {
            int64_t i = b * sb + t * st;
            g = r_t[i] + disc[i] * ((1.0f - lambda_) * v_t[i] + lambda_ * g);
            out[i] = g;
        }
    }
    return 0;
}

/* GAE: values has T+1 entries per column (stride st, column stride sbv); out has T */
int oracle_gae(const float *r_t, const float *disc, const float *values,
               int64_t B, int64_t T, int64_t sb, int64_t st, int64_t sbv, int64_t stv,
               float lambda_, float *out) {
    for (int64_t b = 0; b < B; ++b) {
        float gae = 0.0f;
        for (int64_t t = T - 1; t >= 0; --t) {
            int64_t i = b * sb + t * st;
            float delta = r_t[i] + disc[i] * values[b * sbv + (t + 1) * stv] - values[b * sbv + t * stv];
            gae = delta + lambda_ * disc[i] * gae;
            out[i] = gae;
        }
    }
    return 0;
}
