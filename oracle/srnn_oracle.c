/* srnn_oracle.c -- CPU ORACLE (test infrastructure, NOT the product).
 *
 * A plain-C restatement of the population hot path of illiumst/self-replicating-neural-networks
 * (code/network.py, code/soup.py, code/experiment.py, code/setups/known-fixpoint-variation.py).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * call it, and only as the checker or the timed CPU baseline.  The shipped path is
 * libsrnn_b200.so (CUDA); it never links or calls anything in this file.
 *
 * The reference delegates all arithmetic to TensorFlow-1.x Keras (un-vendored, un-pinned; evidence of
 * vintage: keras 2.2.4 / numpy 1.15.4 in code/fixpoint-2.ipynb).  TensorFlow is not installable
 * here, so parity is anchored on the reference's committed Keras trajectories (tests/golden npz files,
 * extracted by tests/golden/make_golden.py).  Pin status per path (tests/test_oracle_golden.py):
 *   WW forward, AGG forward, run_net stop rule, 5-way classifier ... pinned bit-for-bit by goldens
 *   SGD train / learn_from ... semantics pinned (lr, MSE, batch 1, frozen targets) to ~1e-3 by
 *                              goldens; exact rounding / Keras' hidden shuffle: PARITY UNPINNED
 *   RNN forward / training ... Keras-documented semantics only: PARITY UNPINNED
 *
 * Pinned arithmetic, all FP32, MXCSR flush-to-zero + denormals-are-zero:
 *   R2  Dense with >= 2 output units:  acc = rn(x0*W0j); acc = fma(xk, Wkj, acc), k ascending
 *   R3  Dense with 1 output unit:      acc = +0; acc = rn(acc + rn(xk*Wk)), k ascending (not fused; the
 *       +0 start is pinned by 4 golden values where two negative underflows sum to +0.0, not -0.0)
 *   R4  position codes: Python float64 i/max cast to f32
 *   R5  AGG block mean: f64 sequential sum of the f32 values, / count in f64, cast to f32
 *   R6  predicates compare f32 values (and f32 differences) against the f64 epsilon
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off -mfma -fopenmp).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <xmmintrin.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/srnn_types.h"

#define MAXW SRNN_MAX_WEIGHTS
#define MAXU SRNN_MAX_WIDTH
#define MAXL SRNN_MAX_LAYERS

/* ------------------------------------------------------------------ FTZ/DAZ scope */
static inline unsigned ftz_on(void) {
    unsigned old = _mm_getcsr();
    _mm_setcsr(old | 0x8040u); /* FTZ (bit 15) | DAZ (bit 6) -- rule R1 */
    return old;
}
static inline void ftz_restore(unsigned old) { _mm_setcsr(old); }

/* ------------------------------------------------------------------ shapes */
typedef struct {
    int kind, n_layers;          /* dense layers (WW/AGG) or SimpleRNN layers (RNN)        */
    int in[MAXL], out[MAXL];     /* per layer                                             */
    int off[MAXL];               /* flat offset of the (input) kernel                     */
    int offU[MAXL];              /* RNN: flat offset of the recurrent kernel (out x out)  */
    int n_weights, aggregates;
} shape_t;

/* Layer stacks: code/network.py:226-230 (WW), :329-333 (AGG), :531-535 (RNN). */
static int make_shape(const srnn_net_desc *d, shape_t *s) {
    memset(s, 0, sizeof *s);
    if (!d || d->width < 1 || d->width > MAXU || d->depth < 1 || d->depth + 1 > MAXL) return SRNN_ERR_SHAPE;
    s->kind = d->kind;
    s->n_layers = d->depth + 1;
    int io;
    if (d->kind == SRNN_KIND_WW) io = 4;
    else if (d->kind == SRNN_KIND_AGG) io = d->aggregates;
    else if (d->kind == SRNN_KIND_RNN) io = 1;
    else return SRNN_ERR_SHAPE;
    if (io < 1 || io > MAXU) return SRNN_ERR_SHAPE;
    s->aggregates = d->kind == SRNN_KIND_AGG ? d->aggregates : 0;
    int off = 0;
    for (int l = 0; l < s->n_layers; ++l) {
        s->in[l] = l == 0 ? io : d->width;
        s->out[l] = l == s->n_layers - 1 ? (d->kind == SRNN_KIND_WW ? 1 : io) : d->width;
        s->off[l] = off;
        off += s->in[l] * s->out[l];
        if (d->kind == SRNN_KIND_RNN) { s->offU[l] = off; off += s->out[l] * s->out[l]; }
    }
    s->n_weights = off;
    if (off > MAXW) return SRNN_ERR_LIMIT;
    if (d->kind == SRNN_KIND_AGG) {
        int cs = off / d->aggregates;
        if (cs < 1 || off - cs * d->aggregates >= cs) return SRNN_ERR_SHAPE; /* reference would IndexError */
    }
    return SRNN_OK;
}

int srnn_oracle_desc_init(srnn_net_desc *d) {
    shape_t s;
    int rc = make_shape(d, &s);
    if (rc) return rc;
    d->n_weights = s.n_weights;
    return SRNN_OK;
}

/* ------------------------------------------------------------------ dense (R2 / R3) */
/* Test-only switch (srnn_oracle_self_apply_short_form): emulate the CUDA kernels' short form of the single-output layer,
 * whose first add "+0 + p0" is spelled fma(x0, W0, +0) -- the same number, except that a product flushed to -0 stays -0
 * (csrc/srnn_device.cuh, ww_apply ANYZERO).  Never set on the reference-restating paths. */
static __thread int g_short_form = 0;
static inline void dense(const float *x, const float *W, int in, int out, float *y) {
    if (out >= 2) {
        for (int j = 0; j < out; ++j) {
            float acc = x[0] * W[j];
            for (int k = 1; k < in; ++k) acc = fmaf(x[k], W[k * out + j], acc);
            y[j] = acc;
        }
    } else {
        float acc = 0.0f;
        int k0 = 0;
        if (g_short_form) { acc = fmaf(x[0], W[0], 0.0f); k0 = 1; }
        for (int k = k0; k < in; ++k) { float p = x[k] * W[k]; acc = acc + p; }
        y[0] = acc;
    }
}
/* g_prev[i] = sum_j g[j] * W[i][j]  (TF: matmul(g, W, transpose_b)).  Rule R7a, pinned by the recovered Keras sample
 * orders (tests/golden/ww_sgd_perms.npz): every result element is a DOT PRODUCT -- products rounded one by one and
 * added in ascending j, no FMA -- whatever the number of units (the forward GEMM rule R2 does NOT apply here). */
static inline void dense_T(const float *g, const float *W, int in, int out, float *gp) {
    for (int i = 0; i < in; ++i) {
        float acc = g[0] * W[i * out];
        for (int j = 1; j < out; ++j) { float p = g[j] * W[i * out + j]; acc = acc + p; }
        gp[i] = acc;
    }
}

static inline void mlp_forward(const shape_t *s, const float *w, const float *x, float *y) {
    float a[MAXU], b[MAXU];
    const float *cur = x;
    for (int l = 0; l < s->n_layers; ++l) {
        float *dst = (l == s->n_layers - 1) ? y : ((l & 1) ? b : a);
        dense(cur, w + s->off[l], s->in[l], s->out[l], dst);
        cur = dst;
    }
}

/* ------------------------------------------------------------------ position codes (R4) */
/* WeightwiseNeuralNetwork.normalize_id, code/network.py:215-220 */
static inline float norm_id(int v, int mx) { return (float)(mx > 1 ? (double)v / (double)mx : (double)v); }

/* compute_all_duplex_weight_points, code/network.py:239-255: flat k -> (layer, row, col) codes */
static void ww_codes(const shape_t *s, float *L, float *R, float *C) {
    int k = 0;
    for (int l = 0; l < s->n_layers; ++l)
        for (int r = 0; r < s->in[l]; ++r)
            for (int c = 0; c < s->out[l]; ++c, ++k) {
                L[k] = norm_id(l, s->n_layers - 1);
                R[k] = norm_id(r, s->in[l] - 1);
                C[k] = norm_id(c, s->out[l] - 1);
            }
}

/* ------------------------------------------------------------------ the three operators */
/* WeightwiseNeuralNetwork.apply_to_weights, code/network.py:265-279 (+ apply :235-237) */
static void ww_apply(const shape_t *s, const float *op, const float *in, float *out) {
    float L[MAXW], R[MAXW], C[MAXW], res[MAXW];
    ww_codes(s, L, R, C);
    for (int k = 0; k < s->n_weights; ++k) {
        float x[4] = {in[k], L[k], R[k], C[k]};
        mlp_forward(s, op, x, &res[k]);
    }
    memcpy(out, res, sizeof(float) * s->n_weights);
}

/* AggregatingNeuralNetwork: collect_weights :388-403, aggregate_average :294-301 (R5) */
static void agg_means(const shape_t *s, const float *in, float *agg) {
    int a = s->aggregates, n = s->n_weights, cs = n / a;
    for (int g = 0; g < a; ++g) {
        int lo = g * cs, hi = (g == a - 1) ? n : lo + cs;
        double total = 0.0;
        for (int k = lo; k < hi; ++k) total += (double)in[k];
        agg[g] = (float)(total / (double)(hi - lo));
    }
}
/* AggregatingNeuralNetwork.apply_to_weights, code/network.py:359-386 (deaggregate_identically :310) */
static void agg_apply(const shape_t *s, const float *op, const float *in, float *out) {
    float agg[MAXU], res[MAXU];
    int a = s->aggregates, n = s->n_weights, cs = n / a;
    agg_means(s, in, agg);
    mlp_forward(s, op, agg, res);
    for (int g = 0; g < a; ++g) {
        int lo = g * cs, hi = (g == a - 1) ? n : lo + cs;
        for (int k = lo; k < hi; ++k) out[k] = res[g];
    }
}

/* RecurrentNeuralNetwork.apply_to_weights, code/network.py:544-564: the flat weight list is the
 * input sequence of stacked linear SimpleRNN(return_sequences=True) layers, h_0 = 0,
 * h_t = rn( dense(x_t, K) + dense(h_{t-1}, U) )   (Keras SimpleRNNCell.call; order unpinned). */
static void rnn_forward_seq(const shape_t *s, const float *op, const float *seq_in, int T, float *seq_out,
                            float *hist /* optional [n_layers][T][MAXU] */) {
    float h[MAXL][MAXU];
    memset(h, 0, sizeof h);
    for (int t = 0; t < T; ++t) {
        float x[MAXU];
        x[0] = seq_in[t];
        for (int l = 0; l < s->n_layers; ++l) {
            float a[MAXU], b[MAXU];
            dense(x, op + s->off[l], s->in[l], s->out[l], a);
            dense(h[l], op + s->offU[l], s->out[l], s->out[l], b);
            for (int j = 0; j < s->out[l]; ++j) {
                h[l][j] = a[j] + b[j];
                x[j] = h[l][j];
                if (hist) hist[(l * T + t) * MAXU + j] = h[l][j];
            }
        }
        seq_out[t] = x[0];
    }
}
static void rnn_apply(const shape_t *s, const float *op, const float *in, float *out) {
    float res[MAXW];
    rnn_forward_seq(s, op, in, s->n_weights, res, NULL);
    memcpy(out, res, sizeof(float) * s->n_weights);
}

static inline void net_apply(const shape_t *s, const float *op, const float *in, float *out) {
    if (s->kind == SRNN_KIND_WW) ww_apply(s, op, in, out);
    else if (s->kind == SRNN_KIND_AGG) agg_apply(s, op, in, out);
    else rnn_apply(s, op, in, out);
}

/* ------------------------------------------------------------------ predicates (R6) */
/* are_weights_diverged, code/network.py:43-52 */
static inline int is_diverged(const float *w, int n) {
    for (int k = 0; k < n; ++k) if (isnan(w[k]) || isinf(w[k])) return 1;
    return 0;
}
/* are_weights_within / is_zero, code/network.py:54-62, :136-138 (inclusive, NaN fails) */
static inline int is_zero(const float *w, int n, double eps) {
    for (int k = 0; k < n; ++k) if (!(-eps <= (double)w[k] && (double)w[k] <= eps)) return 0;
    return 1;
}
/* is_fixpoint(degree), code/network.py:140-157: operator stays at w, data iterated `degree` times */
static int is_fixpoint(const shape_t *s, const float *w, int degree, double eps) {
    float cur[MAXW], nxt[MAXW];
    memcpy(cur, w, sizeof(float) * s->n_weights);
    for (int d = 0; d < degree; ++d) { net_apply(s, w, cur, nxt); memcpy(cur, nxt, sizeof(float) * s->n_weights); }
    if (is_diverged(cur, s->n_weights)) return 0;
    for (int k = 0; k < s->n_weights; ++k) {
        float diff = cur[k] - w[k];
        if ((double)fabsf(diff) >= eps) return 0;
    }
    return 1;
}
/* decision list of FixpointExperiment.count / Soup.count, code/experiment.py:79-91, code/soup.py:89-103;
 * with_fix_sec = 0 gives the setups' soup-level variant (code/setups/mixed-soup.py:27-53). */
static int classify(const shape_t *s, const float *w, double eps, int with_fix_sec) {
    if (is_diverged(w, s->n_weights)) return SRNN_CLS_DIVERGENT;
    if (is_fixpoint(s, w, 1, eps)) return is_zero(w, s->n_weights, eps) ? SRNN_CLS_FIX_ZERO : SRNN_CLS_FIX_OTHER;
    if (with_fix_sec && is_fixpoint(s, w, 2, eps)) return SRNN_CLS_FIX_SEC;
    return SRNN_CLS_OTHER;
}

/* ------------------------------------------------------------------ arena records (include/srnn_types.h) */
static inline void load_net(const float *w, int64_t ld, int64_t i, int n, float *dst) {
    for (int k = 0; k < n; ++k) dst[k] = w[i * ld + k];
}
static inline void store_net(float *w, int64_t ld, int64_t i, int n, const float *src) {
    for (int k = 0; k < n; ++k) w[i * ld + k] = src[k];
}

/* ================================================================== exported: apply */
/* out[out_idx[p]] <- f_{op[op_idx[p]]}( in[in_idx[p]] ), pairs processed in array order
 * (NeuralNetwork.apply_to_network / attack / fuck, code/network.py:112-122).  NULL idx = identity. */
int srnn_oracle_apply(const srnn_net_desc *d, const float *op_w, int64_t op_ld, const int32_t *op_idx,
                      const float *in_w, int64_t in_ld, const int32_t *in_idx, float *out_w, int64_t out_ld,
                      const int32_t *out_idx, int64_t n) {
    shape_t s;
    int rc = make_shape(d, &s);
    if (rc) return rc;
    unsigned csr = ftz_on();
    for (int64_t p = 0; p < n; ++p) {
        float op[MAXW], in[MAXW], out[MAXW];
        load_net(op_w, op_ld, op_idx ? op_idx[p] : p, s.n_weights, op);
        load_net(in_w, in_ld, in_idx ? in_idx[p] : p, s.n_weights, in);
        net_apply(&s, op, in, out);
        store_net(out_w, out_ld, out_idx ? out_idx[p] : p, s.n_weights, out);
    }
    ftz_restore(csr);
    return SRNN_OK;
}

/* ================================================================== exported: raw forward */
/* NeuralNetwork.apply(*inputs): model.predict on one input row (code/network.py:235-237, :355-357, :540-542).
 * WW/AGG: x[n][io] -> y[n][out];  RNN: x[n][T] -> y[n][T] with T = n_weights. */
int srnn_oracle_forward(const srnn_net_desc *d, const float *w, int64_t ld, const int32_t *idx, const float *x,
                        float *y, int64_t n) {
    shape_t s;
    int rc = make_shape(d, &s);
    if (rc) return rc;
    unsigned csr = ftz_on();
    for (int64_t p = 0; p < n; ++p) {
        float op[MAXW];
        load_net(w, ld, idx ? idx[p] : p, s.n_weights, op);
        if (s.kind == SRNN_KIND_RNN) rnn_forward_seq(&s, op, x + p * s.n_weights, s.n_weights, y + p * s.n_weights, NULL);
        else mlp_forward(&s, op, x + p * s.in[0], y + p * s.out[s.n_layers - 1]);
    }
    ftz_restore(csr);
    return SRNN_OK;
}

/* ================================================================== exported: self-application */
/* mode SRNN_RUN_FIXED: n_steps x self_attack (code/network.py:124-127), steps_out = n_steps.
 * mode SRNN_RUN_UNTIL: FixpointExperiment.run_net (code/experiment.py:70-77):
 *     while i < limit and not is_diverged() and not is_fixpoint(): self_attack(); i += 1
 * traj (optional): state after executed step s (1-based) at traj[((s-1)*n + i)*ld + k]. */
int srnn_oracle_self_apply(const srnn_net_desc *d, float *w, int64_t ld, int64_t n, int32_t n_steps, int32_t mode,
                           double eps, int32_t *steps_out, float *traj) {
    shape_t s;
    int rc = make_shape(d, &s);
    if (rc) return rc;
#pragma omp parallel
    {
        unsigned csr = ftz_on();
#pragma omp for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
            float cur[MAXW], nxt[MAXW];
            load_net(w, ld, i, s.n_weights, cur);
            int32_t it = 0;
            while (it < n_steps) {
                if (mode == SRNN_RUN_UNTIL && (is_diverged(cur, s.n_weights) || is_fixpoint(&s, cur, 1, eps))) break;
                net_apply(&s, cur, cur, nxt);
                memcpy(cur, nxt, sizeof(float) * s.n_weights);
                ++it;
                if (traj) store_net(traj + (int64_t)(it - 1) * n * ld, ld, i, s.n_weights, cur);
            }
            store_net(w, ld, i, s.n_weights, cur);
            if (steps_out) steps_out[i] = it;
        }
        ftz_restore(csr);
    }
    return SRNN_OK;
}

/* Test infrastructure for the zero-sign lemma (tests/test_oracle_zero_signs.py): n_steps self-applications the way the
 * fixed-length CUDA kernels run them -- steps 1 .. n_steps-1 in the short form (zeros of either sign), the last one exact
 * (exact_last = 0: every step in the short form, to show that the short form alone does differ). */
int srnn_oracle_self_apply_short_form(const srnn_net_desc *d, float *w, int64_t ld, int64_t n, int32_t n_steps,
                                      int32_t exact_last) {
    shape_t s;
    int rc = make_shape(d, &s);
    if (rc) return rc;
#pragma omp parallel
    {
        unsigned csr = ftz_on();
#pragma omp for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
            float cur[MAXW], nxt[MAXW];
            load_net(w, ld, i, s.n_weights, cur);
            for (int32_t it = 0; it < n_steps; ++it) {
                g_short_form = !exact_last || it < n_steps - 1;
                net_apply(&s, cur, cur, nxt);
                memcpy(cur, nxt, sizeof(float) * s.n_weights);
            }
            g_short_form = 0;
            store_net(w, ld, i, s.n_weights, cur);
        }
        ftz_restore(csr);
    }
    return SRNN_OK;
}

/* ================================================================== exported: classify */
int srnn_oracle_classify(const srnn_net_desc *d, const float *w, int64_t ld, int64_t n, double eps,
                         int32_t with_fix_sec, uint8_t *cls_out, int64_t *counters /* [5], accumulated */) {
    shape_t s;
    int rc = make_shape(d, &s);
    if (rc) return rc;
    int64_t cnt[5] = {0, 0, 0, 0, 0};
#pragma omp parallel
    {
        unsigned csr = ftz_on();
        int64_t loc[5] = {0, 0, 0, 0, 0};
#pragma omp for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
            float cur[MAXW];
            load_net(w, ld, i, s.n_weights, cur);
            int c = classify(&s, cur, eps, with_fix_sec);
            if (cls_out) cls_out[i] = (uint8_t)c;
            loc[c]++;
        }
#pragma omp critical
        for (int c = 0; c < 5; ++c) cnt[c] += loc[c];
        ftz_restore(csr);
    }
    if (counters) for (int c = 0; c < 5; ++c) counters[c] += cnt[c];
    return SRNN_OK;
}

/* per-net predicates, for the facade-level tests: bit0 diverged, bit1 zero, bit2 fixpoint(1), bit3 fixpoint(2) */
int srnn_oracle_predicates(const srnn_net_desc *d, const float *w, int64_t ld, int64_t n, double eps, uint8_t *flags) {
    shape_t s;
    int rc = make_shape(d, &s);
    if (rc) return rc;
    unsigned csr = ftz_on();
    for (int64_t i = 0; i < n; ++i) {
        float cur[MAXW];
        load_net(w, ld, i, s.n_weights, cur);
        flags[i] = (uint8_t)(is_diverged(cur, s.n_weights) | (is_zero(cur, s.n_weights, eps) << 1) |
                             (is_fixpoint(&s, cur, 1, eps) << 2) | (is_fixpoint(&s, cur, 2, eps) << 3));
    }
    ftz_restore(csr);
    return SRNN_OK;
}

/* ================================================================== Philox4x32-10 */
static inline void philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1,
                              uint32_t out[4]) {
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
int srnn_oracle_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t *out) {
    philox4x32(c0, c1, c2, c3, k0, k1, out);
    return SRNN_OK;
}

/* RNG streams: counter = (slot_lo, slot_hi | stream << 24, a, b), key = (seed_lo, seed_hi).
 * `slot` is the GLOBAL net index (so sharding over ranks is invisible). */
enum { ST_INIT = 1, ST_PERM = 2, ST_SOUP = 3, ST_KFV = 4, ST_ORTHO = 5 };
static inline void rng4(uint64_t seed, int stream, uint64_t slot, uint32_t a, uint32_t b, uint32_t out[4]) {
    philox4x32((uint32_t)slot, (uint32_t)((slot >> 32) & 0xFFFFFFu) | ((uint32_t)stream << 24), a, b, (uint32_t)seed,
               (uint32_t)(seed >> 32), out);
}
static inline float u01f(uint32_t x) { return (float)(x >> 8) * 5.9604644775390625e-08f; } /* [0,1), 24 bit */

/* ================================================================== exported: init (respawn) */
/* Keras defaults implied by Dense / SimpleRNN: kernels glorot_uniform U(+-sqrt(6/(in+out)));
 * SimpleRNN recurrent kernels orthogonal.  Deterministic stand-in for the unseeded Keras RNG:
 *   kernel weight k of slot i, generation g:  (2u-1)*limit, u from word (k&3) of rng4(ST_INIT, i, g, k>>2)
 *   recurrent kernel: Gram-Schmidt of rows drawn uniformly in the unit ball (rejection), i.e. Haar on O(n),
 *   using only + - * / sqrt so CPU and GPU agree bit-for-bit. */
static void ortho_rows(uint64_t seed, uint64_t slot, uint32_t gen, int layer, int n, float *U) {
    uint32_t blk = 0;
    for (int r = 0; r < n; ++r) {
        int placed = 0;
        for (int attempt = 0; attempt < 64; ++attempt) {
            float v[MAXU];
            float nn = 0.f;
            for (int c = 0; c < n; c += 4) {
                uint32_t x[4];
                rng4(seed, ST_ORTHO, slot, gen, ((uint32_t)layer << 24) | blk++, x);
                for (int q = 0; q < 4 && c + q < n; ++q) v[c + q] = 2.0f * u01f(x[q]) - 1.0f;
            }
            for (int c = 0; c < n; ++c) nn = nn + v[c] * v[c];
            if (n == 1) { U[0] = v[0] < 0.f ? -1.0f : 1.0f; placed = 1; break; } /* 1x1 orthogonal = +-1 exactly */
            if (nn > 1.0f || nn < 1e-4f) continue;           /* outside ball / too short: reject  */
            for (int p = 0; p < r; ++p) {                     /* remove components along earlier rows */
                float dot = 0.f;
                for (int c = 0; c < n; ++c) dot = dot + v[c] * U[p * n + c];
                for (int c = 0; c < n; ++c) v[c] = v[c] - dot * U[p * n + c];
            }
            nn = 0.f;
            for (int c = 0; c < n; ++c) nn = nn + v[c] * v[c];
            if (nn < 1e-2f) continue;
            float inv = 1.0f / sqrtf(nn);
            for (int c = 0; c < n; ++c) U[r * n + c] = v[c] * inv;
            placed = 1;
            break;
        }
        if (!placed) { /* 64 rejections in a row (probability ~ 1e-40): deterministic fallback, the identity matrix */
            for (int i = 0; i < n; ++i)
                for (int c = 0; c < n; ++c) U[i * n + c] = i == c ? 1.0f : 0.0f;
            return;
        }
    }
}
static void init_net(const shape_t *s, uint64_t seed, uint64_t slot, uint32_t gen, float *w) {
    for (int l = 0; l < s->n_layers; ++l) {
        float limit = (float)sqrt(6.0 / (double)(s->in[l] + s->out[l]));
        int cnt = s->in[l] * s->out[l];
        for (int q = 0; q < cnt; ++q) {
            int k = s->off[l] + q;
            uint32_t x[4];
            rng4(seed, ST_INIT, slot, gen, (uint32_t)(k >> 2), x);
            w[k] = (2.0f * u01f(x[k & 3]) - 1.0f) * limit;
        }
        if (s->kind == SRNN_KIND_RNN) ortho_rows(seed, slot, gen, l, s->out[l], w + s->offU[l]);
    }
}
int srnn_oracle_init(const srnn_net_desc *d, float *w, int64_t ld, int64_t n, const int32_t *idx, uint64_t seed,
                     uint64_t slot0, uint32_t generation) {
    shape_t s;
    int rc = make_shape(d, &s);
    if (rc) return rc;
    unsigned csr = ftz_on();
    for (int64_t p = 0; p < n; ++p) {
        int64_t i = idx ? idx[p] : p;
        float cur[MAXW];
        init_net(&s, seed, slot0 + (uint64_t)i, generation, cur);
        store_net(w, ld, i, s.n_weights, cur);
    }
    ftz_restore(csr);
    return SRNN_OK;
}

/* ================================================================== SGD (R7, proposed / unpinned) */
/* One Keras epoch with batch_size = 1 (TrainingNeuralNetworkDecorator.train / learn_from,
 * code/network.py:613-626): for each sample in `order`: forward, MSE, backward, W <- rn(W - rn(lr*g)).
 * Returns the f64 mean of the f32 pre-update sample losses (Keras History 'loss'). */
static int n_samples_of(const shape_t *s) { return s->kind == SRNN_KIND_WW ? s->n_weights : 1; }

static double mlp_sgd_sample(const shape_t *s, float *w, const float *x, const float *y, float lr) {
    float act[MAXL + 1][MAXU];
    int nl = s->n_layers;
    memcpy(act[0], x, sizeof(float) * s->in[0]);
    for (int l = 0; l < nl; ++l) dense(act[l], w + s->off[l], s->in[l], s->out[l], act[l + 1]);
    int no = s->out[nl - 1];
    float g[MAXU], gp[MAXU], dW[MAXW];
    float inv = 1.0f / (float)no, tot = 0.f;
    for (int j = 0; j < no; ++j) {
        float dlt = act[nl][j] - y[j];
        float sq = dlt * dlt;
        tot = tot + sq;
        g[j] = inv * (2.0f * dlt);
    }
    float loss = tot / (float)no;
    for (int l = nl - 1; l >= 0; --l) {
        const float *W = w + s->off[l];
        for (int i = 0; i < s->in[l]; ++i)
            for (int j = 0; j < s->out[l]; ++j) dW[s->off[l] + i * s->out[l] + j] = act[l][i] * g[j];
        if (l > 0) { dense_T(g, W, s->in[l], s->out[l], gp); memcpy(g, gp, sizeof(float) * s->in[l]); }
    }
    for (int k = 0; k < s->n_weights; ++k) { float stp = lr * dW[k]; w[k] = w[k] - stp; }
    return (double)loss;
}

/* BPTT through the stacked linear SimpleRNN; one sample = the whole sequence (compute_samples :566-574);
 * loss = mean_t (out_t - y_t)^2; gradient contributions accumulated for t descending. */
static double rnn_sgd_sample(const shape_t *s, float *w, const float *seq, const float *y, int T, float lr) {
    float hist[MAXL * MAXW * MAXU];
    float out[MAXW];
    rnn_forward_seq(s, w, seq, T, out, hist);
    int nl = s->n_layers;
    float dW[MAXW];
    memset(dW, 0, sizeof dW);
    float dh_next[MAXL][MAXU];
    memset(dh_next, 0, sizeof dh_next);
    float inv = 1.0f / (float)T, tot = 0.f;
    for (int t = 0; t < T; ++t) { float dlt = out[t] - y[t]; float sq = dlt * dlt; tot = tot + sq; }
    float loss = tot / (float)T;
    for (int t = T - 1; t >= 0; --t) {
        float dx[MAXU];
        dx[0] = inv * (2.0f * (out[t] - y[t]));
        for (int l = nl - 1; l >= 0; --l) {
            int u = s->out[l], in = s->in[l];
            const float *K = w + s->off[l], *U = w + s->offU[l];
            float dh[MAXU], rec[MAXU];
            dense_T(dh_next[l], U, u, u, rec);              /* dh_{t+1} . U^T (zero at t = T-1) */
            for (int j = 0; j < u; ++j) dh[j] = dx[j] + rec[j];
            const float *xin = l == 0 ? NULL : &hist[((l - 1) * T + t) * MAXU];
            for (int i = 0; i < in; ++i) {
                float xi = l == 0 ? seq[t] : xin[i];
                for (int j = 0; j < u; ++j) { float c = xi * dh[j]; dW[s->off[l] + i * u + j] = dW[s->off[l] + i * u + j] + c; }
            }
            if (t > 0) {
                const float *hp = &hist[(l * T + t - 1) * MAXU];
                for (int i = 0; i < u; ++i)
                    for (int j = 0; j < u; ++j) { float c = hp[i] * dh[j]; dW[s->offU[l] + i * u + j] = dW[s->offU[l] + i * u + j] + c; }
            }
            float dxin[MAXU];
            if (l > 0) dense_T(dh, K, in, u, dxin);
            memcpy(dh_next[l], dh, sizeof(float) * u);
            if (l > 0) memcpy(dx, dxin, sizeof(float) * in);
        }
    }
    for (int k = 0; k < s->n_weights; ++k) { float stp = lr * dW[k]; w[k] = w[k] - stp; }
    return (double)loss;
}

/* PERM_PHILOX: the engine's deterministic stand-in for Keras' unseeded per-epoch shuffle (np.random.shuffle inside
 * model.fit, code/network.py:617).  Fisher-Yates: for i = n-1 .. 1: j = draw(i + 1); swap(perm[i], perm[j]).
 * Draws come from 32-bit Philox words by exact fixed-point decomposition: j = hi32(r * m), r = lo32(r * m); a word
 * serves consecutive draws while the product of their ranges stays <= 2^22 (bias <= 2^-10), then the next word is
 * taken.  An epoch needs W(n) words (2 for n = 14); when W <= 4 one Philox block serves 4 / W consecutive epochs
 * (block index epoch / EPB, words (epoch % EPB) * W ...), else ceil(W / 4) blocks per epoch. */
static int perm_words_needed(int n) {
    int wi = -1;
    uint32_t prod = 1;
    for (int i = n - 1; i >= 1; --i) {
        uint32_t m = (uint32_t)i + 1;
        if (wi < 0 || (uint64_t)prod * m > (1u << 22)) { ++wi; prod = 1; }
        prod *= m;
    }
    return wi + 1;
}
static void philox_perm(uint64_t seed, uint64_t slot, uint64_t epoch, int n, uint8_t *perm) {
    for (int i = 0; i < n; ++i) perm[i] = (uint8_t)i;
    if (n < 2) return;
    const int W = perm_words_needed(n);
    const int EPB = W <= 4 ? 4 / W : 1;
    const uint64_t E = epoch / (uint64_t)EPB;
    const int o = (int)(epoch % (uint64_t)EPB) * W;
    uint32_t x[4], r = 0, prod = 1;
    int wi = -1, cur_blk = -1;
    for (int i = n - 1; i >= 1; --i) {
        uint32_t m = (uint32_t)i + 1;
        if (wi < 0 || (uint64_t)prod * m > (1u << 22)) {
            ++wi;
            prod = 1;
            int g = o + wi, blk = g >> 2;
            if (blk != cur_blk) {
                rng4(seed, ST_PERM, slot, (uint32_t)E, ((uint32_t)(E >> 32) << 8) | (uint32_t)blk, x);
                cur_blk = blk;
            }
            r = x[g & 3];
        }
        prod *= m;
        uint64_t wide = (uint64_t)r * m;
        int j = (int)(wide >> 32);
        r = (uint32_t)wide;
        uint8_t tmp = perm[i]; perm[i] = perm[j]; perm[j] = tmp;
    }
}
int srnn_oracle_perm(uint64_t seed, uint64_t slot, uint64_t epoch, int32_t n, uint8_t *perm) {
    philox_perm(seed, slot, epoch, n, perm);
    return SRNN_OK;
}

/* one epoch of `student` on the samples of `teacher` (teacher == student weights copy for self-train;
 * targets are frozen when the epoch starts: compute_samples is called before fit, code/network.py:615,623) */
static double train_epoch(const shape_t *s, float *student, const float *teacher, float lr, const uint8_t *order) {
    double acc = 0.0;
    if (s->kind == SRNN_KIND_WW) {
        float L[MAXW], R[MAXW], C[MAXW];
        ww_codes(s, L, R, C);                                /* compute_samples, code/network.py:281-289 */
        for (int q = 0; q < s->n_weights; ++q) {
            int k = order ? order[q] : q;
            float x[4] = {teacher[k], L[k], R[k], C[k]};
            float y = teacher[k];
            acc += mlp_sgd_sample(s, student, x, &y, lr);
        }
        return acc / (double)s->n_weights;
    } else if (s->kind == SRNN_KIND_AGG) {
        float agg[MAXU];
        agg_means(s, teacher, agg);                          /* compute_samples, code/network.py:414-417 */
        return mlp_sgd_sample(s, student, agg, agg, lr);
    } else {
        return rnn_sgd_sample(s, student, teacher, teacher, s->n_weights, lr);
    }
}

/* n_epochs of train() (teacher_w == NULL) or learn_from() (teacher arena + index) on every net.
 * perm_mode: NONE | EXPLICIT (perm[e*n_samples + q], shared) | PHILOX (slot = slot0 + i, epoch0 + e). */
int srnn_oracle_train(const srnn_net_desc *d, float *w, int64_t ld, int64_t n, int32_t n_epochs, float lr,
                      int32_t perm_mode, const uint8_t *perm, uint64_t seed, uint64_t slot0, uint64_t epoch0,
                      const float *teacher_w, int64_t teacher_ld, const int32_t *teacher_idx, double *loss_out) {
    shape_t s;
    int rc = make_shape(d, &s);
    if (rc) return rc;
    int ns = n_samples_of(&s);
    if (perm_mode == SRNN_PERM_EXPLICIT && !perm) return SRNN_ERR_ARG;
#pragma omp parallel
    {
        unsigned csr = ftz_on();
#pragma omp for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
            float cur[MAXW], tch[MAXW];
            uint8_t order[MAXW];
            load_net(w, ld, i, s.n_weights, cur);
            if (teacher_w) load_net(teacher_w, teacher_ld, teacher_idx ? teacher_idx[i] : i, s.n_weights, tch);
            double loss = 0.0;
            for (int e = 0; e < n_epochs; ++e) {
                if (!teacher_w) memcpy(tch, cur, sizeof(float) * s.n_weights);
                const uint8_t *ord = NULL;
                if (ns > 1 && perm_mode == SRNN_PERM_EXPLICIT) ord = perm + (int64_t)e * ns;
                else if (ns > 1 && perm_mode == SRNN_PERM_PHILOX) {
                    philox_perm(seed, slot0 + (uint64_t)i, epoch0 + (uint64_t)e, ns, order);
                    ord = order;
                }
                loss = train_epoch(&s, cur, tch, lr, ord);
            }
            store_net(w, ld, i, s.n_weights, cur);
            if (loss_out) loss_out[i] = loss;
        }
        ftz_restore(csr);
    }
    return SRNN_OK;
}

/* ================================================================== known-fixpoint variation */
/* code/setups/known-fixpoint-variation.py:37-46 (vary) and :60-89 (trial loop).  Deterministic
 * stand-in for Python's `random`: weights 2m, 2m+1 of trial i use words (0,1) and (2,3) of rng4(ST_KFV, slot, level, m)
 * as (coin, magnitude): coin < 0.5 ? w + u*e : w - u*e in float64, stored as float32. */
static inline double u01d(uint32_t x) { return (double)x * 2.3283064365386963e-10; } /* x / 2^32 */

int srnn_oracle_kfv(const srnn_net_desc *d, const float *base /* n_weights */, int64_t n, double scale,
                    uint32_t level, int32_t max_steps, double eps, uint64_t seed, uint64_t slot0,
                    int32_t *time_to_vergence, int32_t *time_as_fixpoint, float *w_out, int64_t ld) {
    shape_t s;
    int rc = make_shape(d, &s);
    if (rc) return rc;
#pragma omp parallel
    {
        unsigned csr = ftz_on();
#pragma omp for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
            float cur[MAXW], nxt[MAXW];
            for (int k = 0; k < s.n_weights; ++k) {
                uint32_t x[4];
                rng4(seed, ST_KFV, slot0 + (uint64_t)i, level, (uint32_t)(k >> 1), x);   /* one Philox block per 2 weights */
                double mag = u01d(x[(k & 1) * 2 + 1]) * scale;
                double v = u01d(x[(k & 1) * 2]) < 0.5 ? (double)base[k] + mag : (double)base[k] - mag;
                cur[k] = (float)v;
            }
            int32_t tts = 0, taf = 0;
            int still = 1;
            for (int st = 0; st < max_steps; ++st) {
                net_apply(&s, cur, cur, nxt);
                memcpy(cur, nxt, sizeof(float) * s.n_weights);
                if (is_zero(cur, s.n_weights, eps) || is_diverged(cur, s.n_weights)) break;
                if (is_fixpoint(&s, cur, 1, eps)) { if (still) ++taf; else still = 1; }
                else still = 0;
                ++tts;
            }
            time_to_vergence[i] = tts;
            time_as_fixpoint[i] = taf;
            if (w_out) store_net(w_out, ld, i, s.n_weights, cur);
        }
        ftz_restore(csr);
    }
    return SRNN_OK;
}

/* ================================================================== soup */
/* Pairing draws for one generation, snapshot-synchronous mode: particle `slot` draws
 * (attack coin, attack target, learn coin, learn target) = the reference's per-particle draw order
 * (code/soup.py:56-63) from rng4(ST_SOUP, slot, generation_lo, generation_hi). */
static inline void soup_draw(const srnn_soup_params *p, uint64_t seed, uint64_t slot, uint64_t gen, int64_t n_total,
                             int64_t *att, int64_t *lrn) {
    uint32_t x[4];
    rng4(seed, ST_SOUP, slot, (uint32_t)gen, (uint32_t)(gen >> 32), x);
    *att = (u01d(x[0]) < p->attacking_rate) ? (int64_t)(((uint64_t)x[1] * (uint64_t)n_total) >> 32) : -1;
    *lrn = (u01d(x[2]) < p->learn_from_rate) ? (int64_t)(((uint64_t)x[3] * (uint64_t)n_total) >> 32) : -1;
}
int srnn_oracle_soup_draw(const srnn_soup_params *p, uint64_t seed, uint64_t gen, int64_t n_total, int64_t slot0,
                          int64_t n, int64_t *att, int64_t *lrn) {
    for (int64_t i = 0; i < n; ++i) soup_draw(p, seed, (uint64_t)(slot0 + i), gen, n_total, &att[i], &lrn[i]);
    return SRNN_OK;
}

/* Shared tail of both soup modes for particle i (code/soup.py:62-87): learn_from x severity, train x t,
 * cull tests on the OLD particle, respawn keyed by (slot, generation). */
static void soup_particle_tail(const shape_t *s, const srnn_soup_params *p, float *cur, const float *teacher,
                               int teacher_is_self, uint64_t seed, uint64_t slot, uint64_t gen, uint8_t *action, double *loss,
                               float *logged /* weights of the old particle after its events */,
                               const float *spares, int64_t spare_ld, int32_t n_spares, int32_t *spares_used) {
    uint8_t order[MAXW];
    int ns = n_samples_of(s);
    int sev = p->learn_from_severity > 0 ? p->learn_from_severity : 0;
    int tr = p->train > 0 ? p->train : 0;
    uint64_t ep = gen * (uint64_t)(sev + tr);
    if (teacher || teacher_is_self) {
        float self_copy[MAXW];
        for (int e = 0; e < sev; ++e) {
            const uint8_t *ord = NULL;
            if (teacher_is_self) { memcpy(self_copy, cur, sizeof(float) * s->n_weights); teacher = self_copy; }
            if (ns > 1 && p->perm_mode == SRNN_PERM_PHILOX) { philox_perm(seed, slot, ep + (uint64_t)e, ns, order); ord = order; }
            train_epoch(s, cur, teacher, p->lr, ord);
        }
        *action = SRNN_ACT_LEARN;
    }
    for (int e = 0; e < tr; ++e) {
        float tch[MAXW];
        memcpy(tch, cur, sizeof(float) * s->n_weights);
        const uint8_t *ord = NULL;
        if (ns > 1 && p->perm_mode == SRNN_PERM_PHILOX) { philox_perm(seed, slot, ep + (uint64_t)(sev + e), ns, order); ord = order; }
        *loss = train_epoch(s, cur, tch, p->lr, ord);
        *action = SRNN_ACT_TRAIN;
    }
    if (logged) memcpy(logged, cur, sizeof(float) * s->n_weights);
    int dead = 0;
    if (p->remove_divergent && is_diverged(cur, s->n_weights)) { dead = 1; *action = SRNN_ACT_DIV_DEAD; }
    if (p->remove_zero && is_zero(cur, s->n_weights, p->epsilon)) { dead = 1; *action = SRNN_ACT_ZERO_DEAD; }
    if (dead) {
        /* respawn: the next caller-supplied spare record (generator() output) if any is left, else Philox init */
        if (spares && spares_used && *spares_used < n_spares) {
            load_net(spares, spare_ld, *spares_used, s->n_weights, cur);
            ++*spares_used;
        } else {
            init_net(s, seed, slot, (uint32_t)gen, cur);
        }
    }
}

/* Exact-order generation = Soup.evolve (code/soup.py:51-87): particles visited in index order, every
 * write visible to later particles.  att[i] / lrn[i] = target index or -1 (host draws them, e.g. with
 * Python's `random` in the reference's draw order).  Outputs per particle: action code, loss of the last
 * train epoch (NaN if none), logged[i*ld+k] = weights save_state would log for the old particle. */
int srnn_oracle_soup_ordered(const srnn_net_desc *d, const srnn_soup_params *p, float *w, int64_t ld, int64_t n,
                             const int64_t *att, const int64_t *lrn, uint64_t seed, uint64_t gen, uint8_t *action,
                             double *loss, float *logged, const float *spares, int64_t spare_ld, int32_t n_spares,
                             int32_t *spares_used) {
    shape_t s;
    int rc = make_shape(d, &s);
    if (rc) return rc;
    unsigned csr = ftz_on();
    if (spares_used) *spares_used = 0;
    for (int64_t i = 0; i < n; ++i) {
        float cur[MAXW], oth[MAXW], res[MAXW], lg[MAXW];
        uint8_t act = SRNN_ACT_NONE;
        double ls = NAN;
        load_net(w, ld, i, s.n_weights, cur);
        if (att && att[i] >= 0) {
            load_net(w, ld, att[i], s.n_weights, oth);
            net_apply(&s, cur, oth, res);
            store_net(w, ld, att[i], s.n_weights, res);
            if (att[i] == i) memcpy(cur, res, sizeof(float) * s.n_weights);
            act = SRNN_ACT_ATTACK;
        }
        int has_teacher = lrn && lrn[i] >= 0 && lrn[i] != i;
        int self_teacher = lrn && lrn[i] == i; /* samples recomputed per call: code/network.py:623 */
        if (has_teacher) load_net(w, ld, lrn[i], s.n_weights, oth);
        soup_particle_tail(&s, p, cur, has_teacher ? oth : NULL, self_teacher, seed, (uint64_t)i, gen, &act, &ls, lg,
                           spares, spare_ld, n_spares, spares_used);
        store_net(w, ld, i, s.n_weights, cur);
        if (action) action[i] = act;
        if (loss) loss[i] = ls;
        if (logged) store_net(logged, ld, i, s.n_weights, lg);
    }
    ftz_restore(csr);
    return SRNN_OK;
}

/* Snapshot-synchronous generation (the large-soup mode; a documented departure from the reference's
 * Gauss-Seidel order): all partners are read from the start-of-generation snapshot `snap`
 * (global arena of n_total nets); this call owns particles [slot0, slot0+n) and writes them to w_out.
 * A victim hit by several attackers applies them in ascending attacker index. */
int srnn_oracle_soup_sync(const srnn_net_desc *d, const srnn_soup_params *p, const float *snap, int64_t snap_ld,
                          int64_t n_total, float *w_out, int64_t out_ld, int64_t slot0, int64_t n, uint64_t seed,
                          uint64_t gen, uint8_t *action, double *loss, int64_t *n_attacks_out) {
    shape_t s;
    int rc = make_shape(d, &s);
    if (rc) return rc;
    /* inverse attack map by counting sort over all global draws (every owner evaluates the same function) */
    int64_t *cnt = (int64_t *)calloc((size_t)n + 1, sizeof(int64_t));
    int64_t n_att = 0;
    for (int64_t a = 0; a < n_total; ++a) {
        int64_t at, lr_;
        soup_draw(p, seed, (uint64_t)a, gen, n_total, &at, &lr_);
        if (at >= slot0 && at < slot0 + n) { cnt[at - slot0 + 1]++; n_att++; }
    }
    for (int64_t i = 0; i < n; ++i) cnt[i + 1] += cnt[i];
    int64_t *lst = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n_att ? n_att : 1));
    int64_t *pos = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n ? n : 1));
    memcpy(pos, cnt, sizeof(int64_t) * (size_t)n);
    for (int64_t a = 0; a < n_total; ++a) {
        int64_t at, lr_;
        soup_draw(p, seed, (uint64_t)a, gen, n_total, &at, &lr_);
        if (at >= slot0 && at < slot0 + n) lst[pos[at - slot0]++] = a; /* ascending a by construction */
    }
#pragma omp parallel
    {
        unsigned csr = ftz_on();
#pragma omp for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
            float cur[MAXW], oth[MAXW], res[MAXW];
            uint8_t act = SRNN_ACT_NONE;
            double ls = NAN;
            int64_t at, lr_;
            load_net(snap, snap_ld, slot0 + i, s.n_weights, cur);
            soup_draw(p, seed, (uint64_t)(slot0 + i), gen, n_total, &at, &lr_);
            if (at >= 0) act = SRNN_ACT_ATTACK;
            for (int64_t q = cnt[i]; q < cnt[i + 1]; ++q) {
                load_net(snap, snap_ld, lst[q], s.n_weights, oth);
                net_apply(&s, oth, cur, res);
                memcpy(cur, res, sizeof(float) * s.n_weights);
            }
            if (lr_ >= 0) load_net(snap, snap_ld, lr_, s.n_weights, oth);
            soup_particle_tail(&s, p, cur, lr_ >= 0 ? oth : NULL, 0, seed, (uint64_t)(slot0 + i), gen, &act, &ls, NULL,
                               NULL, 0, 0, NULL);
            store_net(w_out, out_ld, i, s.n_weights, cur);
            if (action) action[i] = act;
            if (loss) loss[i] = ls;
        }
        ftz_restore(csr);
    }
    if (n_attacks_out) *n_attacks_out = n_att;
    free(cnt); free(lst); free(pos);
    return SRNN_OK;
}

/* ================================================================== fast CPU baseline */
/* WW(2,2) self-application, same pinned arithmetic, restructured for the host's SIMD units: blocks of 64
 * records are transposed into planes so the inner loop runs over nets and gcc vectorises the fma chains.  Used as the "strong CPU baseline"
 * of bench.py; checked against srnn_oracle_self_apply in the CPU tests. */
enum { FAST_B = 64 };
/* n_steps self-applications of a block of 64 nets held as weight planes.  Compiled three times (AVX-512, AVX2+FMA,
 * baseline) and dispatched on the CPU the code runs on, so the library built in the container uses the widest SIMD
 * of the GPU box's host without -march=native.  fmaf stays one IEEE fma in every clone (libm call in the baseline). */
__attribute__((target_clones("arch=x86-64-v4", "arch=x86-64-v3", "default")))
static void ww22_block_steps(float (*cur)[FAST_B], float (*nxt)[FAST_B], int n_steps) {
    static const float L[14] = {0, 0, 0, 0, 0, 0, 0, 0, .5f, .5f, .5f, .5f, 1, 1};
    const float third = (float)(1.0 / 3.0), two3 = (float)(2.0 / 3.0);
    const float R[14] = {0, 0, third, third, two3, two3, 1, 1, 0, 0, 1, 1, 0, 1};
    static const float C[14] = {0, 1, 0, 1, 0, 1, 0, 1, 0, 1, 0, 1, 0, 0};
    for (int st = 0; st < n_steps; ++st) {
        for (int k = 0; k < 14; ++k) {
            const float l = L[k], r = R[k], c = C[k];
#pragma omp simd
            for (int i = 0; i < FAST_B; ++i) {
                float x = cur[k][i];
                float h0 = x * cur[0][i], h1 = x * cur[1][i];
                h0 = fmaf(l, cur[2][i], h0); h1 = fmaf(l, cur[3][i], h1);
                h0 = fmaf(r, cur[4][i], h0); h1 = fmaf(r, cur[5][i], h1);
                h0 = fmaf(c, cur[6][i], h0); h1 = fmaf(c, cur[7][i], h1);
                float g0 = h0 * cur[8][i], g1 = h0 * cur[9][i];
                g0 = fmaf(h1, cur[10][i], g0); g1 = fmaf(h1, cur[11][i], g1);
                float p0 = g0 * cur[12][i], p1 = g1 * cur[13][i];
                nxt[k][i] = (0.0f + p0) + p1;
            }
        }
        memcpy(cur, nxt, sizeof(float) * 14 * FAST_B);
    }
}
int srnn_oracle_ww22_self_apply_fast(float *w, int64_t ld, int64_t n, int32_t n_steps) {
    enum { B = FAST_B };
#pragma omp parallel
    {
        unsigned csr = ftz_on();
#pragma omp for schedule(static)
        for (int64_t i0 = 0; i0 < n; i0 += B) {
            int m = (int)((n - i0) < B ? (n - i0) : B);
            float cur[14][B], nxt[14][B];
            for (int k = 0; k < 14; ++k)
                for (int i = 0; i < m; ++i) cur[k][i] = w[(i0 + i) * ld + k];
            for (int k = 0; k < 14; ++k)
                for (int i = m; i < B; ++i) cur[k][i] = 0.f;
            ww22_block_steps(cur, nxt, n_steps);
            for (int k = 0; k < 14; ++k)
                for (int i = 0; i < m; ++i) w[(i0 + i) * ld + k] = cur[k][i];
        }
        ftz_restore(csr);
    }
    return SRNN_OK;
}

/* bench.py --impl reference under torchrun: the launcher exports OMP_NUM_THREADS=1, which would time one core */
void srnn_oracle_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#endif
}

int srnn_oracle_num_threads(void) {
    int nt = 1;
#ifdef _OPENMP
#pragma omp parallel
    {
#pragma omp master
        nt = omp_get_num_threads();
    }
#endif
    return nt;
}
