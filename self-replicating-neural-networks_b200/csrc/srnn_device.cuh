// srnn_device.cuh -- device-side building blocks of libsrnn_b200.so (sm_100a).
//
// Everything a single net needs is held in registers by ONE thread: its weight record (14 / 20 / 17
// floats for the benchmarked shapes), the operator's temporaries and, for SGD, the activations.  Net
// shapes are compile-time (template Shape<>) so every weight index is a constant after unrolling and
// the position codes of the weightwise net become FFMA immediates.
//
// Arithmetic is pinned to the reference's committed Keras trajectories (DESIGN.md "Pinned arithmetic"):
// compile with -fmad=false -ftz=true -prec-div=true -prec-sqrt=true and spell every fusion explicitly.
//   R2  Dense, >= 2 output units:  acc = rn(x0*W0j); acc = fma(xk, Wkj, acc), k ascending
//   R3  Dense, 1 output unit:      acc = +0; acc = rn(acc + rn(xk*Wk)), k ascending
//   R4  position codes = (float)((double)i / (double)max)
//   R5  AGG block mean = f64 sequential sum / count, cast to f32
//   R6  predicates on f32 values against f32-rounded epsilons (host rounds eps up / down, see srnn_b200.cu)
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <type_traits>
#include <utility>

#include "../../include/srnn_types.h"

namespace srnn {

#define SRNN_DEV __device__ __forceinline__
#define SRNN_HD __host__ __device__ constexpr

template <int I, int N, class F>
SRNN_DEV void static_for(F &&f) {
    if constexpr (I < N) {
        f(std::integral_constant<int, I>{});
        static_for<I + 1, N>(f);
    }
}
// counts down: I = N-1 .. 0
template <int N, class F>
SRNN_DEV void static_rfor(F &&f) {
    if constexpr (N > 0) {
        f(std::integral_constant<int, N - 1>{});
        static_rfor<N - 1>(f);
    }
}

// ------------------------------------------------------------------------------------------------ shapes
// Keras layer stacks of code/network.py:226-230 (WW: 4 -> width x depth -> 1), :329-333 (AGG: a -> ... -> a),
// :531-535 (RNN: SimpleRNN 1 -> width x depth -> 1, each with input kernel K and recurrent kernel U).
template <int KIND_, int WIDTH_, int DEPTH_, int AGG_>
struct ShapeFn {  // helpers live in a base so they are complete when Shape's constants are evaluated
    static constexpr int NL = DEPTH_ + 1;
    static constexpr int IO = KIND_ == SRNN_KIND_WW ? 4 : (KIND_ == SRNN_KIND_AGG ? AGG_ : 1);
    static constexpr int OUTW = KIND_ == SRNN_KIND_WW ? 1 : IO;
    static SRNN_HD int in(int l) { return l == 0 ? IO : WIDTH_; }
    static SRNN_HD int out(int l) { return l == NL - 1 ? OUTW : WIDTH_; }
    static SRNN_HD int layer_size(int l) { return in(l) * out(l) + (KIND_ == SRNN_KIND_RNN ? out(l) * out(l) : 0); }
    static SRNN_HD int off(int l) {
        int o = 0;
        for (int q = 0; q < l; ++q) o += layer_size(q);
        return o;
    }
    static SRNN_HD int offU(int l) { return off(l) + in(l) * out(l); }
    // weightwise position codes: compute_all_duplex_weight_points, code/network.py:239-255 (rule R4)
    static SRNN_HD int layer_of(int k) {
        int l = 0;
        while (l < NL - 1 && k >= off(l + 1)) ++l;
        return l;
    }
    static SRNN_HD float norm_id(int v, int mx) { return (float)(mx > 1 ? (double)v / (double)mx : (double)v); }
    static SRNN_HD float codeL(int k) { return norm_id(layer_of(k), NL - 1); }
    static SRNN_HD float codeR(int k) { return norm_id((k - off(layer_of(k))) / out(layer_of(k)), in(layer_of(k)) - 1); }
    static SRNN_HD float codeC(int k) { return norm_id((k - off(layer_of(k))) % out(layer_of(k)), out(layer_of(k)) - 1); }
};
template <int KIND_, int WIDTH_, int DEPTH_, int AGG_>
struct Shape : ShapeFn<KIND_, WIDTH_, DEPTH_, AGG_> {
    using Fn = ShapeFn<KIND_, WIDTH_, DEPTH_, AGG_>;
    static constexpr int KIND = KIND_, WIDTH = WIDTH_, DEPTH = DEPTH_, AGGS = AGG_;
    static constexpr int NL = Fn::NL, IO = Fn::IO, OUTW = Fn::OUTW;
    static constexpr int MAXU = IO > WIDTH_ ? IO : WIDTH_;
    static constexpr int NW = Fn::off(NL);
    static constexpr int NS = KIND_ == SRNN_KIND_WW ? NW : 1;  // SGD samples per epoch
    // aggregating net: collect_weights, code/network.py:388-403 -- block g = [g*CS, (g+1)*CS), last takes the rest
    static constexpr int CS = KIND_ == SRNN_KIND_AGG ? NW / (AGG_ > 0 ? AGG_ : 1) : 0;
};

// ------------------------------------------------------------------------------------------------ packed FP32
// Blackwell's packed FP32 instructions (PTX fma/mul .f32x2 -> SASS FFMA2/FMUL2): two IEEE-rn, flush-to-zero
// lanes per thread-instruction, i.e. the SAME per-lane results as the scalar forms with half the instruction
// issue and 64-bit register-pair operands (a scalar operand is broadcast to both lanes by the hardware).
// The two lanes are the two units of a width-2 layer, whose kernel columns W[k][0], W[k][1] are adjacent.
#ifndef SRNN_PACKED_F32X2
#define SRNN_PACKED_F32X2 1
#endif
typedef unsigned long long f32x2_t;
SRNN_DEV f32x2_t pack2(float lo, float hi) {
    f32x2_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
SRNN_DEV void unpack2(f32x2_t v, float &lo, float &hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
SRNN_DEV f32x2_t mul2(f32x2_t a, f32x2_t b) {
    f32x2_t r;
    asm("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
// 1.0f that ptxas cannot see through.  ptxas (12.9) contracts mul.rn.f32x2 followed by add.rn.f32x2 -- or by an
// fma.f32x2 whose multiplier is the literal 1 -- into ONE FFMA2, i.e. a single rounding where the reference has
// two (the explicit .rn does not stop it for the packed forms; it does for scalars).  Multiplying by a 1 that is
// only known at run time keeps the two roundings and costs the same instruction.
static __device__ float g_opaque_one = 1.0f;
SRNN_DEV float opaque_one() { return *reinterpret_cast<volatile float *>(&g_opaque_one); }
SRNN_DEV f32x2_t fma2(f32x2_t a, f32x2_t b, f32x2_t c) {
    f32x2_t r;
    asm("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

// ------------------------------------------------------------------------------------------------ dense
template <int IN, int OUT>
SRNN_DEV void dense(const float *x, const float *W, float *y) {
#if SRNN_PACKED_F32X2
    if constexpr (OUT % 2 == 0) {  // R2, two units per instruction: acc = rn(x0*W0j); acc = fma(xk, Wkj, acc)
#pragma unroll
        for (int j = 0; j < OUT; j += 2) {
            f32x2_t acc = mul2(pack2(x[0], x[0]), pack2(W[j], W[j + 1]));
#pragma unroll
            for (int k = 1; k < IN; ++k) acc = fma2(pack2(x[k], x[k]), pack2(W[k * OUT + j], W[k * OUT + j + 1]), acc);
            unpack2(acc, y[j], y[j + 1]);
        }
        return;
    } else if constexpr (OUT == 1 && IN == 2) {  // R3: the two products in one instruction, adds unfused from +0
        float p0, p1;
        unpack2(mul2(pack2(x[0], x[1]), pack2(W[0], W[1])), p0, p1);
        y[0] = __fadd_rn(__fadd_rn(0.0f, p0), p1);
        return;
    }
#endif
    if constexpr (OUT >= 2) {  // R2
#pragma unroll
        for (int j = 0; j < OUT; ++j) {
            float acc = __fmul_rn(x[0], W[j]);
#pragma unroll
            for (int k = 1; k < IN; ++k) acc = __fmaf_rn(x[k], W[k * OUT + j], acc);
            y[j] = acc;
        }
    } else {  // R3
        float acc = 0.0f;
#pragma unroll
        for (int k = 0; k < IN; ++k) acc = __fadd_rn(acc, __fmul_rn(x[k], W[k]));
        y[0] = acc;
    }
}
// gp[i] = sum_j g[j] * W[i][j]   (backward of Dense).  Rule R7a, pinned by the Keras sample orders recovered from the
// reference's self-training trajectories: a dot product per result element -- products rounded one by one and added in
// ascending j, never fused -- whatever the number of units.
template <int IN, int OUT>
SRNN_DEV void dense_T(const float *g, const float *W, float *gp) {
#pragma unroll
    for (int i = 0; i < IN; ++i) {
        float acc = __fmul_rn(g[0], W[i * OUT]);
#pragma unroll
        for (int j = 1; j < OUT; ++j) acc = __fadd_rn(acc, __fmul_rn(g[j], W[i * OUT + j]));
        gp[i] = acc;
    }
}

template <class S, int L = 0>
SRNN_DEV void mlp_forward(const float *w, const float *x, float *y) {
    if constexpr (L == S::NL - 1) {
        dense<S::in(L), S::out(L)>(x, w + S::off(L), y);
    } else {
        float h[S::MAXU];
        dense<S::in(L), S::out(L)>(x, w + S::off(L), h);
        mlp_forward<S, L + 1>(w, h, y);
    }
}

// ------------------------------------------------------------------------------------------------ operators
// out <- f_op(in);  `out` must not alias `in` or `op`.
// WeightwiseNeuralNetwork.apply_to_weights, code/network.py:265-279
#if SRNN_PACKED_F32X2
// Layers L.. of the MLP for TWO weight points at once: lane = point, the kernel entries are scalar operands the
// hardware broadcasts (SASS `.F32`).  Per lane the operations and their order are those of dense<>: R2 for layers of
// two or more units (first product rounded, then FMAs), R3 for the single output unit (products rounded one by one,
// unfused adds from +0: each add is fma(1, product, acc) with the run-time 1 so that ptxas cannot contract it with its
// multiply.  The first one may NOT be shortened to fma(x, W, +0): a product that flushes to -0 must still add up to +0,
// as the reference's goldens have it).  Every instruction is packed: no scalar FADD between packed FMAs, which cost
// two FMA-pipe cycles each (profiles/r02_issue_mix.txt).
template <class S, int L, bool ANYZERO>
SRNN_DEV void mlp_forward_pair(const float *w, const f32x2_t *x, float &y0, float &y1, float one) {
    constexpr int IN = S::in(L), OUT = S::out(L);
    const float *W = w + S::off(L);
    f32x2_t h[OUT];
    if constexpr (OUT >= 2) {
#pragma unroll
        for (int j = 0; j < OUT; ++j) {
            f32x2_t acc = mul2(x[0], pack2(W[j], W[j]));
#pragma unroll
            for (int k = 1; k < IN; ++k) acc = fma2(x[k], pack2(W[k * OUT + j], W[k * OUT + j]), acc);
            h[j] = acc;
        }
    } else {
        // ANYZERO: (+0 + p0) spelled fma(x0, W0, +0) -- one instruction less; the same number, except that a product
        // flushed to -0 stays -0 where the reference gets +0 (see ww_apply for when that is allowed).
        f32x2_t acc = ANYZERO ? fma2(x[0], pack2(W[0], W[0]), pack2(0.0f, 0.0f))
                              : fma2(pack2(one, one), mul2(x[0], pack2(W[0], W[0])), pack2(0.0f, 0.0f));
#pragma unroll
        for (int k = 1; k < IN; ++k) acc = fma2(pack2(one, one), mul2(x[k], pack2(W[k], W[k])), acc);
        h[0] = acc;
    }
    if constexpr (L == S::NL - 1) unpack2(h[0], y0, y1);
    else mlp_forward_pair<S, L + 1, ANYZERO>(w, h, y0, y1, one);
}
#endif

// ANYZERO = true is the form for steps whose resulting zeros nobody sees with their sign: its results equal the exact
// form's in every bit except that a zero may carry either sign.  (A) The sign of a zero never reaches a non-zero value,
// a NaN or the zero / non-zero pattern of anything computed from it (a zero times anything finite is a zero, a zero plus
// x is x, 0 * Inf is NaN whatever the sign) -- so tests that look at magnitudes only (the KFV loop) may use it freely.
// (B) When the single output unit has at most TWO inputs, the exact form's own zeros do not depend on the signs of the
// zeros it was given: out = (+0 + p0) + p1, and +0 + p0 is +0 or non-zero, never -0.  Hence, for such shapes
// (exact_step_ignores_zero_signs), any number of ANYZERO applications followed by ONE exact application is
// bit-identical to the all-exact chain -- the fixed-length kernels run their inner steps this way.  With three or more
// inputs (B) fails: a partial sum can underflow to -0 and the next product may be a zero of either sign
// (tests/test_oracle_zero_signs.py shows both; tests/test_gpu_parity.py::test_self_apply_fixed_zero_signs).
template <class S>
SRNN_HD bool exact_step_ignores_zero_signs() {
    return S::KIND == SRNN_KIND_WW && S::in(S::NL - 1) <= 2;
}
template <class S, bool ANYZERO = false>
SRNN_DEV void ww_apply(const float *op, const float *in, float *out, float one) {
#if SRNN_PACKED_F32X2
    if constexpr (S::OUTW == 1) {
        // two weight points per instruction; a position code of exactly 1 multiplies by the run-time 1 (see opaque_one:
        // with the literal, points with layer code 1 lost a rounding -- caught by the parity tests)
        static_for<0, S::NW / 2>([&](auto M) {
            constexpr int k = 2 * decltype(M)::value;
            auto code_pair = [&](float c0, float c1) { return pack2(c0 == 1.0f ? one : c0, c1 == 1.0f ? one : c1); };
            const f32x2_t x[4] = {pack2(in[k], in[k + 1]), code_pair(S::codeL(k), S::codeL(k + 1)),
                                  code_pair(S::codeR(k), S::codeR(k + 1)), code_pair(S::codeC(k), S::codeC(k + 1))};
            mlp_forward_pair<S, 0, ANYZERO>(op, x, out[k], out[k + 1], one);
        });
        if constexpr (S::NW % 2 == 1) {
            constexpr int k = S::NW - 1;
            const float x[4] = {in[k], S::codeL(k), S::codeR(k), S::codeC(k)};
            mlp_forward<S>(op, x, &out[k]);
        }
        return;
    }
#endif
    static_for<0, S::NW>([&](auto K) {
        constexpr int k = decltype(K)::value;
#if SRNN_PACKED_F32X2
        if constexpr (S::WIDTH == 2) {
            // first layer with the position codes as compile-time constants; a code of exactly 1 multiplies by the
            // run-time 1 (see opaque_one: with the literal, points with layer code 1 lost a rounding -- caught by the
            // parity tests).
            f32x2_t acc = mul2(pack2(in[k], in[k]), pack2(op[0], op[1]));
            static_for<0, 3>([&](auto R) {
                constexpr int r = decltype(R)::value;
                const f32x2_t W = pack2(op[2 * (r + 1)], op[2 * (r + 1) + 1]);
                constexpr float code = r == 0 ? S::codeL(k) : (r == 1 ? S::codeR(k) : S::codeC(k));
                if constexpr (code == 1.0f) acc = fma2(pack2(one, one), W, acc);
                else acc = fma2(pack2(code, code), W, acc);
            });
            float h[2];
            unpack2(acc, h[0], h[1]);
            mlp_forward<S, 1>(op, h, &out[k]);
            return;
        }
#endif
        const float x[4] = {in[k], S::codeL(k), S::codeR(k), S::codeC(k)};
        mlp_forward<S>(op, x, &out[k]);
    });
}
// a / N for a compile-time count N, correctly rounded like __ddiv_rn(a, N) but in three instructions instead of the
// ~20 of the general division: with y = rn(1/N) and q0 = rn(a*y) (within one ulp of a/N), the remainder
// r = a - N*q0 is exact in one FMA and rn(q0 + r*y) is the correctly rounded quotient (Markstein).  The operands here
// are sums of float32 values, far from the double range limits; an infinite sum stays infinite, NaN stays NaN, and
// the sums are never -0.  Checked against IEEE division for N = 1..64 on 1.3e9 operands; tests/test_div_count_cpu.py
// repeats the check on 1.2e7 with a C restatement (tests/div_count_check.c).
template <int N>
SRNN_DEV double div_count(double a) {
    if constexpr ((N & (N - 1)) == 0) {
        return __dmul_rn(a, 1.0 / (double)N);   // power of two: exact
    } else {
        constexpr double y = 1.0 / (double)N;
        const double q0 = __dmul_rn(a, y);
        const double r = __fma_rn(-(double)N, q0, a);
        const double q = __fma_rn(r, y, q0);
        return isinf(q0) ? q0 : q;
    }
}
// AggregatingNeuralNetwork.apply_to_weights, code/network.py:359-386 (aggregate_average :294-301, R5)
template <class S>
SRNN_DEV void agg_means(const float *in, float *agg) {
    static_for<0, S::AGGS>([&](auto G) {
        constexpr int g = decltype(G)::value;
        constexpr int lo = g * S::CS, hi = g == S::AGGS - 1 ? S::NW : lo + S::CS;
        double tot = 0.0;
#pragma unroll
        for (int k = lo; k < hi; ++k) tot = __dadd_rn(tot, (double)in[k]);
        agg[g] = (float)div_count<hi - lo>(tot);
    });
}
template <class S>
SRNN_DEV void agg_apply(const float *op, const float *in, float *out) {
    float agg[S::AGGS > 0 ? S::AGGS : 1], res[S::AGGS > 0 ? S::AGGS : 1];
    agg_means<S>(in, agg);
    mlp_forward<S>(op, agg, res);
    static_for<0, S::AGGS>([&](auto G) {
        constexpr int g = decltype(G)::value;
        constexpr int lo = g * S::CS, hi = g == S::AGGS - 1 ? S::NW : lo + S::CS;
#pragma unroll
        for (int k = lo; k < hi; ++k) out[k] = res[g];
    });
}
// RecurrentNeuralNetwork.apply_to_weights, code/network.py:544-564: h_t = rn(dense(x_t, K) + dense(h_{t-1}, U)),
// h_{-1} = 0, layers chained per time step.  HIST = true also records every h (for BPTT): hist[(l*T + t)*MAXU + j].
template <class S, bool HIST>
SRNN_DEV void rnn_forward_seq(const float *op, const float *seq, float *out, float *hist) {
    constexpr int T = S::NW;
    float h[S::NL][S::MAXU];
#pragma unroll
    for (int l = 0; l < S::NL; ++l)
#pragma unroll
        for (int j = 0; j < S::MAXU; ++j) h[l][j] = 0.0f;
    static_for<0, T>([&](auto TT) {
        constexpr int t = decltype(TT)::value;
        float x[S::MAXU];
        x[0] = seq[t];
        static_for<0, S::NL>([&](auto LL) {
            constexpr int l = decltype(LL)::value;
            float a[S::MAXU], b[S::MAXU];
            dense<S::in(l), S::out(l)>(x, op + S::off(l), a);
            dense<S::out(l), S::out(l)>(h[l], op + S::offU(l), b);
#pragma unroll
            for (int j = 0; j < S::out(l); ++j) {
                h[l][j] = __fadd_rn(a[j], b[j]);
                x[j] = h[l][j];
                if constexpr (HIST) hist[(l * T + t) * S::MAXU + j] = h[l][j];
            }
        });
        out[t] = x[0];
    });
}
// `one` = opaque_one(), fetched once per thread by the caller (hoisted out of the step loops)
template <class S, bool ANYZERO = false>
SRNN_DEV void net_apply(const float *op, const float *in, float *out, float one) {
    if constexpr (S::KIND == SRNN_KIND_WW) ww_apply<S, ANYZERO>(op, in, out, one);
    else if constexpr (S::KIND == SRNN_KIND_AGG) agg_apply<S>(op, in, out);
    else rnn_forward_seq<S, false>(op, in, out, nullptr);
}

// ------------------------------------------------------------------------------------------------ predicates
// All three predicates are reductions to ONE number and one compare (7 three-input FMNMX for 14 weights instead of 14
// compares and 7 predicate merges each -- ALU instructions next to packed FMAs are not free, profiles/r02_issue_mix.txt):
//   m = NaN-propagating max_k |w_k|:   diverged <=> !(m <= FLT_MAX)   (m is NaN or Inf iff some weight is)
//                                      zero     <=> m <= eps_le       (NaN compares false, like every element would)
//   md = NaN-ignoring max_k |nw_k - w_k|:  every !(|d_k| >= eps) holds <=> !(md >= eps)  (a NaN d_k passes its own test
//                                      and is skipped by fmaxf; all NaN gives NaN, which passes too)
// -- the same booleans as the element-wise forms for every input, NaN and Inf included.
SRNN_DEV float max_nan(float a, float b) {
    float d;
    asm("max.NaN.ftz.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
    return d;
}
template <int N>
SRNN_DEV float max_abs_nan(const float *w) {
    float m = fabsf(w[0]);
#pragma unroll
    for (int k = 1; k < N; ++k) m = max_nan(m, fabsf(w[k]));
    return m;
}
template <int N>
SRNN_DEV float max_abs_diff(const float *nw, const float *w) {
    float d[N];
#if SRNN_PACKED_F32X2
#pragma unroll
    for (int k = 0; k + 1 < N; k += 2) {
        f32x2_t r;
        asm("sub.rn.ftz.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pack2(nw[k], nw[k + 1])), "l"(pack2(w[k], w[k + 1])));
        unpack2(r, d[k], d[k + 1]);
    }
    if constexpr (N % 2 == 1) d[N - 1] = __fsub_rn(nw[N - 1], w[N - 1]);
#else
#pragma unroll
    for (int k = 0; k < N; ++k) d[k] = __fsub_rn(nw[k], w[k]);
#endif
    float m = fabsf(d[0]);
#pragma unroll
    for (int k = 1; k < N; ++k) m = fmaxf(m, fabsf(d[k]));
    return m;
}
SRNN_DEV bool diverged_from_max(float m) { return !(m <= 3.402823466e+38f); }
// are_weights_diverged, code/network.py:43-52
template <int N>
SRNN_DEV bool is_diverged(const float *w) { return diverged_from_max(max_abs_nan<N>(w)); }
// is_zero, code/network.py:136-138: all -eps <= w <= eps  (eps_le = largest f32 <= eps)
template <int N>
SRNN_DEV bool is_zero(const float *w, float eps_le) { return max_abs_nan<N>(w) <= eps_le; }
// tail of is_fixpoint, code/network.py:149-157: nw = iterated application result (eps_ge = smallest f32 >= eps)
template <int N>
SRNN_DEV bool is_close(const float *nw, const float *w, float eps_ge) {
    return !is_diverged<N>(nw) && !(max_abs_diff<N>(nw, w) >= eps_ge);
}
// decision list of FixpointExperiment.count, code/experiment.py:79-91; nw = f_w(w) already known
template <class S>
SRNN_DEV int classify_with(const float *w, const float *nw, float eps_ge, float eps_le, int with_fix_sec) {
    if (is_diverged<S::NW>(w)) return SRNN_CLS_DIVERGENT;
    if (is_close<S::NW>(nw, w, eps_ge)) return is_zero<S::NW>(w, eps_le) ? SRNN_CLS_FIX_ZERO : SRNN_CLS_FIX_OTHER;
    if (with_fix_sec) {
        float nnw[S::NW];
        net_apply<S>(w, nw, nnw, opaque_one());  // is_fixpoint(2): the operator stays at w
        if (is_close<S::NW>(nnw, w, eps_ge)) return SRNN_CLS_FIX_SEC;
    }
    return SRNN_CLS_OTHER;
}

// ------------------------------------------------------------------------------------------------ Philox4x32-10
struct U4 { uint32_t x, y, z, w; };
SRNN_DEV U4 philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return U4{c0, c1, c2, c3};
}
// streams: counter = (slot_lo, slot_hi(24b) | stream << 24, a, b), key = seed.  slot = GLOBAL net index.
enum { ST_INIT = 1, ST_PERM = 2, ST_SOUP = 3, ST_KFV = 4, ST_ORTHO = 5 };
SRNN_DEV U4 rng4(uint64_t seed, int stream, uint64_t slot, uint32_t a, uint32_t b) {
    return philox4x32((uint32_t)slot, (uint32_t)((slot >> 32) & 0xFFFFFFu) | ((uint32_t)stream << 24), a, b,
                      (uint32_t)seed, (uint32_t)(seed >> 32));
}
SRNN_DEV uint32_t word_of(const U4 &v, int q) { return q == 0 ? v.x : (q == 1 ? v.y : (q == 2 ? v.z : v.w)); }
SRNN_DEV float u01f(uint32_t x) { return __fmul_rn((float)(x >> 8), 5.9604644775390625e-08f); }
SRNN_DEV double u01d(uint32_t x) { return __dmul_rn((double)x, 2.3283064365386963e-10); }

// ------------------------------------------------------------------------------------------------ init / respawn
// glorot_uniform kernels + orthogonal recurrent kernels (Gram-Schmidt of rejection-sampled unit-ball rows);
// only + - * / sqrt, so it matches the oracle bit-for-bit.  limits[l] = (float)sqrt(6/(in+out)) from the host.
template <class S>
SRNN_DEV void ortho_rows(uint64_t seed, uint64_t slot, uint32_t gen, int layer, float *U) {
    constexpr int NMAX = S::MAXU;
    const int n = layer == S::NL - 1 ? S::OUTW : S::WIDTH;
    uint32_t blk = 0;
    for (int r = 0; r < n; ++r) {
        bool placed = false;
        for (int attempt = 0; attempt < 64; ++attempt) {
            float v[NMAX];
            for (int c = 0; c < n; c += 4) {
                const U4 x = rng4(seed, ST_ORTHO, slot, gen, ((uint32_t)layer << 24) | blk++);
                for (int q = 0; q < 4 && c + q < n; ++q)
                    v[c + q] = __fsub_rn(__fmul_rn(2.0f, u01f(word_of(x, q))), 1.0f);
            }
            float nn = 0.f;
            for (int c = 0; c < n; ++c) nn = __fadd_rn(nn, __fmul_rn(v[c], v[c]));
            if (n == 1) { U[0] = v[0] < 0.f ? -1.0f : 1.0f; placed = true; break; }  // 1x1 orthogonal = +-1 exactly
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
struct Limits { float v[SRNN_MAX_LAYERS]; };
template <class S>
SRNN_DEV void init_net(uint64_t seed, uint64_t slot, uint32_t gen, const Limits &lim, float *w) {
    static_for<0, S::NL>([&](auto LL) {
        constexpr int l = decltype(LL)::value;
        const float limit = lim.v[l];
#pragma unroll
        for (int q = 0; q < S::in(l) * S::out(l); ++q) {
            const int k = S::off(l) + q;
            const U4 x = rng4(seed, ST_INIT, slot, gen, (uint32_t)(k >> 2));
            w[k] = __fmul_rn(__fsub_rn(__fmul_rn(2.0f, u01f(word_of(x, k & 3))), 1.0f), limit);
        }
        if constexpr (S::KIND == SRNN_KIND_RNN) {
            float U[S::MAXU * S::MAXU];
            ortho_rows<S>(seed, slot, gen, l, U);
#pragma unroll
            for (int q = 0; q < S::out(l) * S::out(l); ++q) w[S::offU(l) + q] = U[q];
        }
    });
}

// ------------------------------------------------------------------------------------------------ SGD
// One sample of Keras SGD(lr) on MSE, batch 1 (TrainingNeuralNetworkDecorator.train / learn_from,
// code/network.py:613-626): forward, g = (1/n_out) * 2 (out - y), per layer dW = act^T g (single products),
// g_prev = dense_T(g, W) with the PRE-update W, then W <- rn(W - rn(lr * dW)).  Returns the f32 sample loss.
template <class S>
SRNN_DEV float mlp_sgd_sample(float *w, const float *x, const float *y, float lr, float one) {
    float act[S::NL + 1][S::MAXU];
#pragma unroll
    for (int i = 0; i < S::IO; ++i) act[0][i] = x[i];
    static_for<0, S::NL>([&](auto LL) {
        constexpr int l = decltype(LL)::value;
        dense<S::in(l), S::out(l)>(act[l], w + S::off(l), act[l + 1]);
    });
    constexpr int NO = S::OUTW;
    float g[S::MAXU];
    constexpr float inv = 1.0f / (float)NO;   // rn(1/n_out), as TF's mean gradient
    float tot = 0.f;
#pragma unroll
    for (int j = 0; j < NO; ++j) {
        const float dlt = __fsub_rn(act[S::NL][j], y[j]);
        tot = __fadd_rn(tot, __fmul_rn(dlt, dlt));
        // one output unit: 1/n_out = 1 and rn(1 * x) = x for the flushed x = rn(2 * dlt), so the multiply is dropped
        g[j] = NO == 1 ? __fmul_rn(2.0f, dlt) : __fmul_rn(inv, __fmul_rn(2.0f, dlt));
    }
    // tot / n_out: for a power-of-two n_out the multiplication by 1/n_out is the same rounding and avoids the
    // division subroutine in the sample loop
    constexpr bool POW2 = (NO & (NO - 1)) == 0;
    const float loss = NO == 1 ? tot : (POW2 ? __fmul_rn(tot, 1.0f / (float)NO) : __fdiv_rn(tot, (float)NO));
    static_rfor<S::NL>([&](auto LL) {
        constexpr int l = decltype(LL)::value;
        constexpr int IN = S::in(l), OUT = S::out(l);
        float *W = w + S::off(l);
        float gp[S::MAXU];
        if constexpr (l > 0) dense_T<IN, OUT>(g, W, gp);
#if SRNN_PACKED_F32X2
        // W <- rn(W - rn(lr * rn(act*g))) two weights at a time.  The subtraction is spelled fma(1, -step, W) with the
        // run-time 1 (opaque_one) so that ptxas cannot contract it with the multiply before it.
        if constexpr (OUT % 2 == 0) {
            const f32x2_t nlr = pack2(-lr, -lr), ones = pack2(one, one);
#pragma unroll
            for (int j = 0; j < OUT; j += 2) {
                const f32x2_t gpair = pack2(g[j], g[j + 1]);
#pragma unroll
                for (int i = 0; i < IN; ++i) {
                    const f32x2_t step = mul2(nlr, mul2(pack2(act[l][i], act[l][i]), gpair));
                    unpack2(fma2(ones, step, pack2(W[i * OUT + j], W[i * OUT + j + 1])), W[i * OUT + j], W[i * OUT + j + 1]);
                }
            }
        } else if constexpr (OUT == 1 && IN == 2) {
            const f32x2_t step = mul2(pack2(-lr, -lr), mul2(pack2(act[l][0], act[l][1]), pack2(g[0], g[0])));
            unpack2(fma2(pack2(one, one), step, pack2(W[0], W[1])), W[0], W[1]);
        } else
#endif
        {
#pragma unroll
            for (int i = 0; i < IN; ++i)
#pragma unroll
                for (int j = 0; j < OUT; ++j)
                    W[i * OUT + j] = __fsub_rn(W[i * OUT + j], __fmul_rn(lr, __fmul_rn(act[l][i], g[j])));
        }
        if constexpr (l > 0) {
#pragma unroll
            for (int i = 0; i < IN; ++i) g[i] = gp[i];
        }
    });
    return loss;
}

// BPTT through the stacked linear SimpleRNN; one sample = the whole weight sequence (compute_samples,
// code/network.py:566-574); loss = mean_t (out_t - y_t)^2; contributions accumulated for t descending.
template <class S>
SRNN_DEV float rnn_sgd_sample(float *w, const float *seq, float lr) {
    constexpr int T = S::NW;
    float hist[S::NL * T * S::MAXU];
    float out[T];
    rnn_forward_seq<S, true>(w, seq, out, hist);
    float dW[S::NW];
#pragma unroll
    for (int k = 0; k < S::NW; ++k) dW[k] = 0.f;
    float dh_next[S::NL][S::MAXU];
#pragma unroll
    for (int l = 0; l < S::NL; ++l)
#pragma unroll
        for (int j = 0; j < S::MAXU; ++j) dh_next[l][j] = 0.f;
    const float inv = __fdiv_rn(1.0f, (float)T);
    float tot = 0.f;
#pragma unroll
    for (int t = 0; t < T; ++t) {
        const float dlt = __fsub_rn(out[t], seq[t]);
        tot = __fadd_rn(tot, __fmul_rn(dlt, dlt));
    }
    const float loss = __fdiv_rn(tot, (float)T);
    static_rfor<T>([&](auto TT) {
        constexpr int t = decltype(TT)::value;
        float dx[S::MAXU];
        dx[0] = __fmul_rn(inv, __fmul_rn(2.0f, __fsub_rn(out[t], seq[t])));
        static_rfor<S::NL>([&](auto LL) {
            constexpr int l = decltype(LL)::value;
            constexpr int u = S::out(l), in = S::in(l);
            const float *K = w + S::off(l), *U = w + S::offU(l);
            float dh[S::MAXU], rec[S::MAXU];
            dense_T<u, u>(dh_next[l], U, rec);
#pragma unroll
            for (int j = 0; j < u; ++j) dh[j] = __fadd_rn(dx[j], rec[j]);
#pragma unroll
            for (int i = 0; i < in; ++i) {
                float xi;
                if constexpr (l == 0) xi = seq[t];
                else xi = hist[((l - 1) * T + t) * S::MAXU + i];
#pragma unroll
                for (int j = 0; j < u; ++j)
                    dW[S::off(l) + i * u + j] = __fadd_rn(dW[S::off(l) + i * u + j], __fmul_rn(xi, dh[j]));
            }
            if constexpr (t > 0) {
#pragma unroll
                for (int i = 0; i < u; ++i)
#pragma unroll
                    for (int j = 0; j < u; ++j)
                        dW[S::offU(l) + i * u + j] = __fadd_rn(dW[S::offU(l) + i * u + j],
                                                               __fmul_rn(hist[(l * T + t - 1) * S::MAXU + i], dh[j]));
            }
            if constexpr (l > 0) {
                float dxin[S::MAXU];
                dense_T<in, u>(dh, K, dxin);
#pragma unroll
                for (int i = 0; i < in; ++i) dx[i] = dxin[i];
            }
#pragma unroll
            for (int j = 0; j < u; ++j) dh_next[l][j] = dh[j];
        });
    });
#pragma unroll
    for (int k = 0; k < S::NW; ++k) w[k] = __fsub_rn(w[k], __fmul_rn(lr, dW[k]));
    return loss;
}

// Per-thread SGD scratch in shared memory, one column per thread (conflict-free: address = row*stride + tid):
//   tgt[k]   frozen sample source (the teacher's weights, or a copy of the net's own weights)   NW rows
//   ord[q]   sample order of the current epoch                                                  NS rows (bytes)
struct SgdScratch {
    float *tgt;
    uint8_t *ord;
    int stride;       // threads per block (column stride)
    const float *codes;  // WW: [3][NW] position codes (L, R, C), shared by the block
    float one;        // opaque_one(), fetched once per thread
};

// PERM_PHILOX sample orders: the deterministic stand-in for Keras' unseeded per-epoch shuffle.  Fisher-Yates,
// for i = NS-1 .. 1: j = draw(i + 1); swap(perm[i], perm[j]).  Draws come from 32-bit Philox words by exact fixed-point
// decomposition, j = hi32(r * m), r = lo32(r * m); a word serves consecutive draws while the product of their ranges
// stays <= 2^22, then the next word is taken (oracle: philox_perm in oracle/srnn_oracle.c).  An epoch needs W words
// (2 for the 14 samples of WW(2,2)); one Philox block serves EPB = 4 / W consecutive epochs.
template <int NS>
struct PermPlan {
    static constexpr int NDRAW = NS > 0 ? NS - 1 : 0;
    static SRNN_HD int word_of_draw(int d) {   // draw d = 0 .. NDRAW-1 has range m = NS - d
        int wi = 0;
        unsigned prod = 1;
        for (int dd = 0; dd <= d; ++dd) {
            const unsigned m = (unsigned)(NS - dd);
            if (dd > 0 && (unsigned long long)prod * m > (1u << 22)) { ++wi; prod = 1; }
            prod *= m;
        }
        return wi;
    }
    static constexpr int W = NDRAW > 0 ? word_of_draw(NDRAW - 1) + 1 : 1;
    static_assert(W <= 4, "fixed shapes need at most one Philox block per epoch");
    static constexpr int EPB = 4 / W;
};
// the epoch's W words; `blk` caches the Philox block across the EPB epochs it serves (blk_id = its index, ~0 = none)
template <int NS>
SRNN_DEV void perm_words(uint64_t seed, uint64_t slot, uint64_t epoch, uint32_t *r, U4 &blk, uint64_t &blk_id) {
    using P = PermPlan<NS>;
    const uint64_t E = epoch / (uint64_t)P::EPB;
    if (E != blk_id) {
        blk = rng4(seed, ST_PERM, slot, (uint32_t)E, (uint32_t)(E >> 32) << 8);
        blk_id = E;
    }
    const int o = (int)(epoch % (uint64_t)P::EPB) * P::W;
#pragma unroll
    for (int wi = 0; wi < P::W; ++wi) r[wi] = P::EPB == 1 ? word_of(blk, wi) : word_of(blk, o + wi);
}
// orders of up to 16 samples live in ONE 64-bit register, 4 bits per position (sample q runs weight point
// (p >> 4q) & 15): no shared memory, the 13 swaps are shifts and xors
template <int NS>
SRNN_DEV uint64_t perm_nibbles(const uint32_t *r_in) {
    using P = PermPlan<NS>;
    static_assert(NS <= 16, "nibble-packed orders hold 16 samples");
    uint32_t r[P::W];
#pragma unroll
    for (int wi = 0; wi < P::W; ++wi) r[wi] = r_in[wi];
    uint64_t p = 0xFEDCBA9876543210ull;
    static_for<0, P::NDRAW>([&](auto D) {
        constexpr int d = decltype(D)::value, i = NS - 1 - d, wi = P::word_of_draw(d);
        constexpr uint32_t m = (uint32_t)(i + 1);
        const uint32_t sj = __umulhi(r[wi], m) << 2;
        r[wi] *= m;
        const uint32_t x = ((uint32_t)(p >> (4 * i)) ^ (uint32_t)(p >> sj)) & 15u;
        p ^= ((uint64_t)x << (4 * i)) ^ ((uint64_t)x << sj);
    });
    return p;
}
// larger sample counts keep the order as bytes in shared memory, one column per thread
template <int NS>
SRNN_DEV void philox_perm(uint64_t seed, uint64_t slot, uint64_t epoch, uint8_t *ord, int stride) {
    using P = PermPlan<NS>;
    uint32_t r[P::W];
    U4 blk;
    uint64_t blk_id = ~0ull;
    perm_words<NS>(seed, slot, epoch, r, blk, blk_id);
    if constexpr (NS <= 16) {
        const uint64_t p = perm_nibbles<NS>(r);
#pragma unroll
        for (int q = 0; q < NS; ++q) ord[q * stride] = (uint8_t)((uint32_t)(p >> (4 * q)) & 15u);
    } else {
#pragma unroll
        for (int i = 0; i < NS; ++i) ord[i * stride] = (uint8_t)i;
        static_for<0, P::NDRAW>([&](auto D) {
            constexpr int d = decltype(D)::value, i = NS - 1 - d, wi = P::word_of_draw(d);
            constexpr uint32_t m = (uint32_t)(i + 1);
            const int j = (int)__umulhi(r[wi], m);
            r[wi] *= m;
            const uint8_t tmp = ord[i * stride];
            ord[i * stride] = ord[j * stride];
            ord[j * stride] = tmp;
        });
    }
}

// One epoch of a weightwise net with the sample loop unrolled: sample q's weight point comes out of the nibble-packed
// order with one shift (ORDERED) or is q itself; target and position codes are fetched from shared memory by that
// index.  The f64 loss accumulation only exists in the WANT_LOSS instantiation (the last epoch of a call).
template <class S, bool ORDERED, bool WANT_LOSS>
SRNN_DEV double ww_epoch_unrolled(float *w, const float *tgt, int stride, const float *codes, uint64_t order, float lr,
                                  float one) {
    static_assert(S::KIND == SRNN_KIND_WW && S::NW <= 16, "unrolled epoch: weightwise nets of up to 16 weights");
    double acc = 0.0;
    static_for<0, S::NW>([&](auto Q) {
        constexpr int q = decltype(Q)::value;
        const uint32_t k = ORDERED ? ((uint32_t)(order >> (4 * q)) & 15u) : (uint32_t)q;
        const float v = tgt[k * stride];
        const float x[4] = {v, codes[k], codes[S::NW + k], codes[2 * S::NW + k]};
        const float l = mlp_sgd_sample<S>(w, x, &v, lr, one);
        if constexpr (WANT_LOSS) acc = __dadd_rn(acc, (double)l);
    });
    if constexpr (WANT_LOSS) return div_count<S::NW>(acc);
    return 0.0;
}

// One Keras epoch of the student `w` on the samples frozen in sc.tgt; returns the f64 mean sample loss.
template <class S>
SRNN_DEV double train_epoch(float *w, const SgdScratch &sc, bool ordered, float lr, bool want_loss = true) {
    if constexpr (S::KIND == SRNN_KIND_WW) {  // compute_samples, code/network.py:281-289
        double acc = 0.0;
        for (int q = 0; q < S::NW; ++q) {
            const int k = ordered ? (int)sc.ord[q * sc.stride] : q;
            const float v = sc.tgt[k * sc.stride];
            const float x[4] = {v, sc.codes[k], sc.codes[S::NW + k], sc.codes[2 * S::NW + k]};
            const float l = mlp_sgd_sample<S>(w, x, &v, lr, sc.one);
            if (want_loss) acc = __dadd_rn(acc, (double)l);   // only the last epoch's loss is ever reported
        }
        return div_count<S::NW>(acc);
    } else if constexpr (S::KIND == SRNN_KIND_AGG) {  // compute_samples, code/network.py:414-417
        float t[S::NW], agg[S::AGGS];
#pragma unroll
        for (int k = 0; k < S::NW; ++k) t[k] = sc.tgt[k * sc.stride];
        agg_means<S>(t, agg);
        return (double)mlp_sgd_sample<S>(w, agg, agg, lr, sc.one);
    } else {
        float t[S::NW];
#pragma unroll
        for (int k = 0; k < S::NW; ++k) t[k] = sc.tgt[k * sc.stride];
        return (double)rnn_sgd_sample<S>(w, t, lr);
    }
}

template <class S>
SRNN_DEV void fill_codes(float *codes) {  // by one thread
    if constexpr (S::KIND == SRNN_KIND_WW) {
        static_for<0, S::NW>([&](auto K) {
            constexpr int k = decltype(K)::value;
            codes[k] = S::codeL(k);
            codes[S::NW + k] = S::codeR(k);
            codes[2 * S::NW + k] = S::codeC(k);
        });
    }
}

}  // namespace srnn
