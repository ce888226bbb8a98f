// srnn_generic.cuh -- run-time-shape fallback of the per-net operations (any width / depth / aggregates that fits
// SRNN_MAX_WEIGHTS), for API completeness: WeightwiseNeuralNetwork(4, 3), AggregatingNeuralNetwork(6, 3, 2) ... work
// through srnn_apply / srnn_forward / srnn_self_apply / srnn_classify / srnn_train / srnn_init.
//
// Same pinned arithmetic as the templated kernels (R1-R8 in srnn_device.cuh), written with run-time loops and
// per-thread local arrays: correct and bit-identical to the oracle, not fast.  The compile-time shapes listed in
// SRNN_SHAPES (srnn_b200.cu) stay the fast path; soups and the KFV sweep exist for those only.
#pragma once
#include "srnn_device.cuh"

namespace srnn {
namespace gen {

constexpr int MAXW = SRNN_MAX_WEIGHTS, MAXU = SRNN_MAX_WIDTH, MAXL = SRNN_MAX_LAYERS;
constexpr int MAX_HIST = 1024;  // floats of SimpleRNN history one thread may keep for BPTT (layers x T x width)

struct DynShape {
    int kind, n_layers, n_weights, aggregates, n_samples, hist_stride;
    int in[MAXL], out[MAXL], off[MAXL], offU[MAXL];
    float limit[MAXL];                  // glorot limits, (float)sqrt(6 / (in + out))
    float L[MAXW], R[MAXW], C[MAXW];    // weightwise position codes (R4)
};

// host: layer stacks of code/network.py:226-230 (WW), :329-333 (AGG), :531-535 (RNN)
static inline int make_dyn(const srnn_net_desc *d, DynShape *s) {
    memset(s, 0, sizeof *s);
    if (!d || d->width < 1 || d->width > MAXU || d->depth < 1 || d->depth + 1 > MAXL) return SRNN_ERR_SHAPE;
    const int io = d->kind == SRNN_KIND_WW ? 4 : (d->kind == SRNN_KIND_AGG ? d->aggregates : (d->kind == SRNN_KIND_RNN ? 1 : 0));
    if (io < 1 || io > MAXU) return SRNN_ERR_SHAPE;
    s->kind = d->kind;
    s->n_layers = d->depth + 1;
    s->aggregates = d->kind == SRNN_KIND_AGG ? d->aggregates : 0;
    int off = 0;
    for (int l = 0; l < s->n_layers; ++l) {
        s->in[l] = l == 0 ? io : d->width;
        s->out[l] = l == s->n_layers - 1 ? (d->kind == SRNN_KIND_WW ? 1 : io) : d->width;
        s->off[l] = off;
        off += s->in[l] * s->out[l];
        if (d->kind == SRNN_KIND_RNN) { s->offU[l] = off; off += s->out[l] * s->out[l]; }
        s->limit[l] = (float)sqrt(6.0 / (double)(s->in[l] + s->out[l]));
    }
    if (off > MAXW) return SRNN_ERR_LIMIT;
    s->n_weights = off;
    s->n_samples = d->kind == SRNN_KIND_WW ? off : 1;
    if (d->kind == SRNN_KIND_AGG) {
        const int cs = off / io;
        if (cs < 1 || off - cs * io >= cs) return SRNN_ERR_SHAPE;
    }
    if (d->kind == SRNN_KIND_WW) {  // compute_all_duplex_weight_points, code/network.py:239-255
        int k = 0;
        for (int l = 0; l < s->n_layers; ++l)
            for (int r = 0; r < s->in[l]; ++r)
                for (int c = 0; c < s->out[l]; ++c, ++k) {
                    const int ml = s->n_layers - 1, mr = s->in[l] - 1, mc = s->out[l] - 1;
                    s->L[k] = (float)(ml > 1 ? (double)l / (double)ml : (double)l);
                    s->R[k] = (float)(mr > 1 ? (double)r / (double)mr : (double)r);
                    s->C[k] = (float)(mc > 1 ? (double)c / (double)mc : (double)c);
                }
    }
    s->hist_stride = d->width > io ? d->width : io;
    return SRNN_OK;
}

// ---------------------------------------------------------------------------------------------- arithmetic
__device__ inline void dense(const float *x, const float *W, int in, int out, float *y) {
    if (out >= 2) {  // R2
        for (int j = 0; j < out; ++j) {
            float acc = __fmul_rn(x[0], W[j]);
            for (int k = 1; k < in; ++k) acc = __fmaf_rn(x[k], W[k * out + j], acc);
            y[j] = acc;
        }
    } else {  // R3
        float acc = 0.0f;
        for (int k = 0; k < in; ++k) acc = __fadd_rn(acc, __fmul_rn(x[k], W[k]));
        y[0] = acc;
    }
}
__device__ inline void dense_T(const float *g, const float *W, int in, int out, float *gp) {   // rule R7a: dot products
    for (int i = 0; i < in; ++i) {
        float acc = 0.0f;
        for (int j = 0; j < out; ++j) acc = __fadd_rn(acc, __fmul_rn(g[j], W[i * out + j]));
        gp[i] = acc;
    }
}
__device__ inline void mlp_forward(const DynShape &s, const float *w, const float *x, float *y) {
    float a[MAXU], b[MAXU];
    const float *cur = x;
    for (int l = 0; l < s.n_layers; ++l) {
        float *dst = (l == s.n_layers - 1) ? y : ((l & 1) ? b : a);
        dense(cur, w + s.off[l], s.in[l], s.out[l], dst);
        cur = dst;
    }
}
__device__ inline void agg_means(const DynShape &s, const float *in, float *agg) {  // R5
    const int a = s.aggregates, n = s.n_weights, cs = n / a;
    for (int g = 0; g < a; ++g) {
        const int lo = g * cs, hi = (g == a - 1) ? n : lo + cs;
        double tot = 0.0;
        for (int k = lo; k < hi; ++k) tot = __dadd_rn(tot, (double)in[k]);
        agg[g] = (float)__ddiv_rn(tot, (double)(hi - lo));
    }
}
// stacked linear SimpleRNN over a sequence of length T; hist (optional): h of (layer l, step t) at hist[(l*T + t)*stride + j]
__device__ inline void rnn_forward_seq(const DynShape &s, const float *op, const float *seq, int T, float *out, float *hist) {
    float h[MAXL][MAXU];
    for (int l = 0; l < s.n_layers; ++l)
        for (int j = 0; j < MAXU; ++j) h[l][j] = 0.0f;
    for (int t = 0; t < T; ++t) {
        float x[MAXU];
        x[0] = seq[t];
        for (int l = 0; l < s.n_layers; ++l) {
            float a[MAXU], b[MAXU];
            dense(x, op + s.off[l], s.in[l], s.out[l], a);
            dense(h[l], op + s.offU[l], s.out[l], s.out[l], b);
            for (int j = 0; j < s.out[l]; ++j) {
                h[l][j] = __fadd_rn(a[j], b[j]);
                x[j] = h[l][j];
                if (hist) hist[(l * T + t) * s.hist_stride + j] = h[l][j];
            }
        }
        out[t] = x[0];
    }
}
// out <- f_op(in); out must not alias in / op
__device__ inline void net_apply(const DynShape &s, const float *op, const float *in, float *out) {
    if (s.kind == SRNN_KIND_WW) {
        for (int k = 0; k < s.n_weights; ++k) {
            const float x[4] = {in[k], s.L[k], s.R[k], s.C[k]};
            mlp_forward(s, op, x, &out[k]);
        }
    } else if (s.kind == SRNN_KIND_AGG) {
        float agg[MAXU], res[MAXU];
        agg_means(s, in, agg);
        mlp_forward(s, op, agg, res);
        const int a = s.aggregates, n = s.n_weights, cs = n / a;
        for (int g = 0; g < a; ++g) {
            const int lo = g * cs, hi = (g == a - 1) ? n : lo + cs;
            for (int k = lo; k < hi; ++k) out[k] = res[g];
        }
    } else {
        rnn_forward_seq(s, op, in, s.n_weights, out, nullptr);
    }
}
__device__ inline bool is_diverged(const float *w, int n) {
    bool d = false;
    for (int k = 0; k < n; ++k) d |= !(fabsf(w[k]) <= 3.402823466e+38f);
    return d;
}
__device__ inline bool is_zero(const float *w, int n, float eps_le) {
    bool z = true;
    for (int k = 0; k < n; ++k) z &= (fabsf(w[k]) <= eps_le);
    return z;
}
__device__ inline bool is_close(const float *nw, const float *w, int n, float eps_ge) {
    bool ok = !is_diverged(nw, n);
    for (int k = 0; k < n; ++k) ok &= !(fabsf(__fsub_rn(nw[k], w[k])) >= eps_ge);
    return ok;
}
__device__ inline int classify_with(const DynShape &s, const float *w, const float *nw, float eps_ge, float eps_le,
                                    int with_fix_sec) {
    const int n = s.n_weights;
    if (is_diverged(w, n)) return SRNN_CLS_DIVERGENT;
    if (is_close(nw, w, n, eps_ge)) return is_zero(w, n, eps_le) ? SRNN_CLS_FIX_ZERO : SRNN_CLS_FIX_OTHER;
    if (with_fix_sec) {
        float nnw[MAXW];
        net_apply(s, w, nw, nnw);
        if (is_close(nnw, w, n, eps_ge)) return SRNN_CLS_FIX_SEC;
    }
    return SRNN_CLS_OTHER;
}

// ---------------------------------------------------------------------------------------------- init
__device__ inline void ortho_rows(uint64_t seed, uint64_t slot, uint32_t gen, int layer, int n, float *U) {
    uint32_t blk = 0;
    for (int r = 0; r < n; ++r) {
        bool placed = false;
        for (int attempt = 0; attempt < 64; ++attempt) {
            float v[MAXU];
            for (int c = 0; c < n; c += 4) {
                const U4 x = rng4(seed, ST_ORTHO, slot, gen, ((uint32_t)layer << 24) | blk++);
                for (int q = 0; q < 4 && c + q < n; ++q) v[c + q] = __fsub_rn(__fmul_rn(2.0f, u01f(word_of(x, q))), 1.0f);
            }
            float nn = 0.f;
            for (int c = 0; c < n; ++c) nn = __fadd_rn(nn, __fmul_rn(v[c], v[c]));
            if (n == 1) { U[0] = v[0] < 0.f ? -1.0f : 1.0f; placed = true; break; }
            if (nn > 1.0f || nn < 1e-4f) continue;
            for (int p = 0; p < r; ++p) {
                float dot = 0.f;
                for (int c = 0; c < n; ++c) dot = __fadd_rn(dot, __fmul_rn(v[c], U[p * n + c]));
                for (int c = 0; c < n; ++c) v[c] = __fsub_rn(v[c], __fmul_rn(dot, U[p * n + c]));
            }
            nn = 0.f;
            for (int c = 0; c < n; ++c) nn = __fadd_rn(nn, __fmul_rn(v[c], v[c]));
            if (nn < 1e-2f) continue;
            const float inv = __fdiv_rn(1.0f, __fsqrt_rn(nn));
            for (int c = 0; c < n; ++c) U[r * n + c] = __fmul_rn(v[c], inv);
            placed = true;
            break;
        }
        if (!placed) {   // 64 rejections in a row (probability ~ 1e-40): deterministic fallback, the identity matrix
            for (int i = 0; i < n; ++i)
                for (int c = 0; c < n; ++c) U[i * n + c] = i == c ? 1.0f : 0.0f;
            return;
        }
    }
}
__device__ inline void init_net(const DynShape &s, uint64_t seed, uint64_t slot, uint32_t gen, float *w) {
    for (int l = 0; l < s.n_layers; ++l) {
        for (int q = 0; q < s.in[l] * s.out[l]; ++q) {
            const int k = s.off[l] + q;
            const U4 x = rng4(seed, ST_INIT, slot, gen, (uint32_t)(k >> 2));
            w[k] = __fmul_rn(__fsub_rn(__fmul_rn(2.0f, u01f(word_of(x, k & 3))), 1.0f), s.limit[l]);
        }
        if (s.kind == SRNN_KIND_RNN) ortho_rows(seed, slot, gen, l, s.out[l], w + s.offU[l]);
    }
}

// ---------------------------------------------------------------------------------------------- SGD (R7 / R8)
__device__ inline float mlp_sgd_sample(const DynShape &s, float *w, const float *x, const float *y, float lr) {
    float act[MAXL + 1][MAXU];
    const int nl = s.n_layers;
    for (int i = 0; i < s.in[0]; ++i) act[0][i] = x[i];
    for (int l = 0; l < nl; ++l) dense(act[l], w + s.off[l], s.in[l], s.out[l], act[l + 1]);
    const int no = s.out[nl - 1];
    float g[MAXU], gp[MAXU];
    const float inv = __fdiv_rn(1.0f, (float)no);
    float tot = 0.f;
    for (int j = 0; j < no; ++j) {
        const float dlt = __fsub_rn(act[nl][j], y[j]);
        tot = __fadd_rn(tot, __fmul_rn(dlt, dlt));
        g[j] = __fmul_rn(inv, __fmul_rn(2.0f, dlt));
    }
    const float loss = __fdiv_rn(tot, (float)no);
    for (int l = nl - 1; l >= 0; --l) {
        float *W = w + s.off[l];
        const int in = s.in[l], out = s.out[l];
        if (l > 0) dense_T(g, W, in, out, gp);
        for (int i = 0; i < in; ++i)
            for (int j = 0; j < out; ++j)
                W[i * out + j] = __fsub_rn(W[i * out + j], __fmul_rn(lr, __fmul_rn(act[l][i], g[j])));
        if (l > 0)
            for (int i = 0; i < in; ++i) g[i] = gp[i];
    }
    return loss;
}
__device__ inline float rnn_sgd_sample(const DynShape &s, float *w, const float *seq, float lr) {
    const int T = s.n_weights, nl = s.n_layers, hs = s.hist_stride;
    float hist[MAX_HIST];
    float out[MAXW], dW[MAXW];
    rnn_forward_seq(s, w, seq, T, out, hist);
    for (int k = 0; k < s.n_weights; ++k) dW[k] = 0.f;
    float dh_next[MAXL][MAXU];
    for (int l = 0; l < nl; ++l)
        for (int j = 0; j < MAXU; ++j) dh_next[l][j] = 0.f;
    const float inv = __fdiv_rn(1.0f, (float)T);
    float tot = 0.f;
    for (int t = 0; t < T; ++t) {
        const float dlt = __fsub_rn(out[t], seq[t]);
        tot = __fadd_rn(tot, __fmul_rn(dlt, dlt));
    }
    const float loss = __fdiv_rn(tot, (float)T);
    for (int t = T - 1; t >= 0; --t) {
        float dx[MAXU];
        dx[0] = __fmul_rn(inv, __fmul_rn(2.0f, __fsub_rn(out[t], seq[t])));
        for (int l = nl - 1; l >= 0; --l) {
            const int u = s.out[l], in = s.in[l];
            const float *K = w + s.off[l], *U = w + s.offU[l];
            float dh[MAXU], rec[MAXU], dxin[MAXU];
            dense_T(dh_next[l], U, u, u, rec);
            for (int j = 0; j < u; ++j) dh[j] = __fadd_rn(dx[j], rec[j]);
            for (int i = 0; i < in; ++i) {
                const float xi = l == 0 ? seq[t] : hist[((l - 1) * T + t) * hs + i];
                for (int j = 0; j < u; ++j)
                    dW[s.off[l] + i * u + j] = __fadd_rn(dW[s.off[l] + i * u + j], __fmul_rn(xi, dh[j]));
            }
            if (t > 0)
                for (int i = 0; i < u; ++i)
                    for (int j = 0; j < u; ++j)
                        dW[s.offU[l] + i * u + j] = __fadd_rn(dW[s.offU[l] + i * u + j],
                                                              __fmul_rn(hist[(l * T + t - 1) * hs + i], dh[j]));
            if (l > 0) {
                dense_T(dh, K, in, u, dxin);
                for (int i = 0; i < in; ++i) dx[i] = dxin[i];
            }
            for (int j = 0; j < u; ++j) dh_next[l][j] = dh[j];
        }
    }
    for (int k = 0; k < s.n_weights; ++k) w[k] = __fsub_rn(w[k], __fmul_rn(lr, dW[k]));
    return loss;
}
// same definition as PermPlan / philox_perm of srnn_device.cuh and the oracle, with the word schedule found at run time
__device__ inline void philox_perm(uint64_t seed, uint64_t slot, uint64_t epoch, int n, uint8_t *ord) {
    for (int i = 0; i < n; ++i) ord[i] = (uint8_t)i;
    if (n < 2) return;
    int W = 0;
    {
        int wi = -1;
        uint32_t prod = 1;
        for (int i = n - 1; i >= 1; --i) {
            const uint32_t m = (uint32_t)i + 1;
            if (wi < 0 || (unsigned long long)prod * m > (1u << 22)) { ++wi; prod = 1; }
            prod *= m;
        }
        W = wi + 1;
    }
    const int EPB = W <= 4 ? 4 / W : 1;
    const uint64_t E = epoch / (uint64_t)EPB;
    const int o = (int)(epoch % (uint64_t)EPB) * W;
    U4 x{0, 0, 0, 0};
    uint32_t r = 0, prod = 1;
    int wi = -1, cur_blk = -1;
    for (int i = n - 1; i >= 1; --i) {
        const uint32_t m = (uint32_t)i + 1;
        if (wi < 0 || (unsigned long long)prod * m > (1u << 22)) {
            ++wi;
            prod = 1;
            const int g = o + wi, blk = g >> 2;
            if (blk != cur_blk) {
                x = rng4(seed, ST_PERM, slot, (uint32_t)E, ((uint32_t)(E >> 32) << 8) | (uint32_t)blk);
                cur_blk = blk;
            }
            r = word_of(x, g & 3);
        }
        prod *= m;
        const int j = (int)__umulhi(r, m);
        r *= m;
        const uint8_t tmp = ord[i];
        ord[i] = ord[j];
        ord[j] = tmp;
    }
}
// one Keras epoch of the student on the samples of `tch` (frozen); f64 mean of the f32 sample losses
__device__ inline double train_epoch(const DynShape &s, float *w, const float *tch, const uint8_t *ord, float lr) {
    if (s.kind == SRNN_KIND_WW) {
        double acc = 0.0;
        for (int q = 0; q < s.n_weights; ++q) {
            const int k = ord ? (int)ord[q] : q;
            const float v = tch[k];
            const float x[4] = {v, s.L[k], s.R[k], s.C[k]};
            acc = __dadd_rn(acc, (double)mlp_sgd_sample(s, w, x, &v, lr));
        }
        return __ddiv_rn(acc, (double)s.n_weights);
    } else if (s.kind == SRNN_KIND_AGG) {
        float agg[MAXU];
        agg_means(s, tch, agg);
        return (double)mlp_sgd_sample(s, w, agg, agg, lr);
    }
    return (double)rnn_sgd_sample(s, w, tch, lr);
}

// ---------------------------------------------------------------------------------------------- kernels
constexpr int GBS = 128;

__device__ inline void load_rec(const float *w, int64_t ld, int64_t i, int n, float *dst) {
    const float *src = w + i * ld;
    for (int k = 0; k < n; ++k) dst[k] = src[k];
}
__device__ inline void store_rec(float *w, int64_t ld, int64_t i, int n, const float *src) {
    float *dst = w + i * ld;
    for (int k = 0; k < n; ++k) dst[k] = src[k];
}

__global__ void __launch_bounds__(GBS) apply_kernel(DynShape s, const float *op_w, int64_t op_ld, const int32_t *op_idx,
                                                    const float *in_w, int64_t in_ld, const int32_t *in_idx, float *out_w,
                                                    int64_t out_ld, const int32_t *out_idx, int64_t n) {
    const int64_t p = (int64_t)blockIdx.x * GBS + threadIdx.x;
    if (p >= n) return;
    float op[MAXW], in[MAXW], out[MAXW];
    load_rec(op_w, op_ld, op_idx ? op_idx[p] : p, s.n_weights, op);
    load_rec(in_w, in_ld, in_idx ? in_idx[p] : p, s.n_weights, in);
    net_apply(s, op, in, out);
    store_rec(out_w, out_ld, out_idx ? out_idx[p] : p, s.n_weights, out);
}

// NeuralNetwork.apply on one input row.  Two kernels (dense nets / recurrent net) rather than one with a run-time
// branch on the kind: the single-kernel form was miscompiled by nvcc 12.9 -O3 for the recurrent branch (right with
// a printf in the kernel, wrong without); the parity tests caught it.
__global__ void __launch_bounds__(GBS) forward_mlp_kernel(DynShape s, const float *w, int64_t ld, const int32_t *idx,
                                                          const float *x, float *y, int64_t n) {
    const int64_t p = (int64_t)blockIdx.x * GBS + threadIdx.x;
    if (p >= n) return;
    float op[MAXW], in[MAXU], out[MAXU];
    load_rec(w, ld, idx ? idx[p] : p, s.n_weights, op);
    const int io = s.in[0], no = s.out[s.n_layers - 1];
    for (int q = 0; q < io; ++q) in[q] = x[p * io + q];
    mlp_forward(s, op, in, out);
    for (int q = 0; q < no; ++q) y[p * no + q] = out[q];
}
__global__ void __launch_bounds__(GBS) forward_rnn_kernel(DynShape s, const float *w, int64_t ld, const int32_t *idx,
                                                          const float *x, float *y, int64_t n) {
    const int64_t p = (int64_t)blockIdx.x * GBS + threadIdx.x;
    if (p >= n) return;
    float op[MAXW], seq[MAXW], res[MAXW];
    load_rec(w, ld, idx ? idx[p] : p, s.n_weights, op);
    load_rec(x, s.n_weights, p, s.n_weights, seq);
    rnn_forward_seq(s, op, seq, s.n_weights, res, nullptr);
    store_rec(y, s.n_weights, p, s.n_weights, res);
}

__global__ void __launch_bounds__(GBS) self_apply_kernel(DynShape s, float *w, int64_t ld, int64_t n, int n_steps, int mode,
                                                         float eps_ge, float eps_le, int with_fix_sec, int *steps_out,
                                                         uint8_t *cls_out, unsigned long long *counters, float *traj) {
    const int64_t i = (int64_t)blockIdx.x * GBS + threadIdx.x;
    if (i >= n) return;
    const int nw = s.n_weights;
    float cur[MAXW], nxt[MAXW];
    load_rec(w, ld, i, nw, cur);
    net_apply(s, cur, cur, nxt);
    int it = 0;
    while (it < n_steps) {
        if (mode == SRNN_RUN_UNTIL && (is_diverged(cur, nw) || is_close(nxt, cur, nw, eps_ge))) break;
        for (int k = 0; k < nw; ++k) cur[k] = nxt[k];
        ++it;
        if (traj) store_rec(traj + (int64_t)(it - 1) * n * ld, ld, i, nw, cur);
        net_apply(s, cur, cur, nxt);
    }
    if (steps_out) steps_out[i] = it;
    if (cls_out || counters) {
        const int cls = classify_with(s, cur, nxt, eps_ge, eps_le, with_fix_sec);
        if (cls_out) cls_out[i] = (uint8_t)cls;
        if (counters) atomicAdd(&counters[cls], 1ull);
    }
    store_rec(w, ld, i, nw, cur);
}

__global__ void __launch_bounds__(GBS) classify_kernel(DynShape s, const float *w, int64_t ld, int64_t n, float eps_ge,
                                                       float eps_le, int with_fix_sec, uint8_t *cls_out, uint8_t *flags_out,
                                                       unsigned long long *counters) {
    const int64_t i = (int64_t)blockIdx.x * GBS + threadIdx.x;
    if (i >= n) return;
    const int nw = s.n_weights;
    float cur[MAXW], nxt[MAXW];
    load_rec(w, ld, i, nw, cur);
    net_apply(s, cur, cur, nxt);
    const int cls = classify_with(s, cur, nxt, eps_ge, eps_le, with_fix_sec);
    if (cls_out) cls_out[i] = (uint8_t)cls;
    if (counters) atomicAdd(&counters[cls], 1ull);
    if (flags_out) {
        float nn[MAXW];
        net_apply(s, cur, nxt, nn);
        flags_out[i] = (uint8_t)((is_diverged(cur, nw) ? 1 : 0) | (is_zero(cur, nw, eps_le) ? 2 : 0) |
                                 (is_close(nxt, cur, nw, eps_ge) ? 4 : 0) | (is_close(nn, cur, nw, eps_ge) ? 8 : 0));
    }
}

__global__ void __launch_bounds__(GBS) init_kernel(DynShape s, float *w, int64_t ld, int64_t n, const int32_t *idx,
                                                   uint64_t seed, uint64_t slot0, uint32_t gen) {
    const int64_t p = (int64_t)blockIdx.x * GBS + threadIdx.x;
    if (p >= n) return;
    const int64_t i = idx ? idx[p] : p;
    float cur[MAXW];
    init_net(s, seed, slot0 + (uint64_t)i, gen, cur);
    store_rec(w, ld, i, s.n_weights, cur);
}

__global__ void __launch_bounds__(GBS) train_kernel(DynShape s, float *w, int64_t ld, int64_t n, int n_epochs, float lr,
                                                    int perm_mode, const uint8_t *perm, uint64_t seed, uint64_t slot0,
                                                    uint64_t epoch0, const float *teacher_w, int64_t teacher_ld,
                                                    const int32_t *teacher_idx, double *loss_out) {
    const int64_t i = (int64_t)blockIdx.x * GBS + threadIdx.x;
    if (i >= n) return;
    const int nw = s.n_weights, ns = s.n_samples;
    float cur[MAXW], tch[MAXW];
    uint8_t ord[MAXW];
    load_rec(w, ld, i, nw, cur);
    if (teacher_w) load_rec(teacher_w, teacher_ld, teacher_idx ? teacher_idx[i] : i, nw, tch);
    double loss = 0.0;
    for (int e = 0; e < n_epochs; ++e) {
        if (!teacher_w)
            for (int k = 0; k < nw; ++k) tch[k] = cur[k];
        const uint8_t *o = nullptr;
        if (ns > 1 && perm_mode == SRNN_PERM_EXPLICIT) {
            for (int q = 0; q < ns; ++q) ord[q] = perm[(int64_t)e * ns + q];
            o = ord;
        } else if (ns > 1 && perm_mode == SRNN_PERM_PHILOX) {
            philox_perm(seed, slot0 + (uint64_t)i, epoch0 + (uint64_t)e, ns, ord);
            o = ord;
        }
        loss = train_epoch(s, cur, tch, o, lr);
    }
    store_rec(w, ld, i, nw, cur);
    if (loss_out) loss_out[i] = loss;
}

}  // namespace gen
}  // namespace srnn
