/* sgd_pin.c -- TEST INFRASTRUCTURE (never linked into the product): recover the hidden sample order of a Keras
 * epoch and probe the rounding of its SGD step, for WeightwiseNeuralNetwork(2,2).
 *
 * TrainingNeuralNetworkDecorator.train (code/network.py:613-618) is `model.fit(x, y, batch_size=1)` with Keras'
 * default shuffle=True: the 14 samples of an epoch are visited in an order drawn from the unseeded global NumPy
 * RNG.  The reference committed the weights after every one of 1010 such epochs
 * (code/results/self_training_weightwise_network/trajectorys.dill -> tests/golden/ww_self_training.npz), not the
 * orders.  The order matters at O(lr^2) ~ 1e-4, i.e. ~1000 ulp, so pinning the SGD arithmetic to the north_star's
 * 1e-5 needs the order back.  tests/golden/make_sgd_perms.py does that with the two functions below:
 *
 *   sgd_pin_search_exact : exhaustive meet-in-the-middle over all 14! orders on the dynamics themselves (below).
 *   sgd_pin_epoch  : one epoch in the oracle's FP32 arithmetic (FTZ/DAZ, R2/R3 forward) for a given order and a
 *                    given choice of the two roundings the forward goldens cannot pin:
 *                      gp : g1 = g2 . W2^T   0: fma(g2_1,W[i][1], rn(g2_0*W[i][0]))   (GEMM-style, rule R2)
 *                                            1: rn(rn(g2_0*W[i][0]) + rn(g2_1*W[i][1])) (dot-product style, rule R3)
 *                                            2: fma(g2_0,W[i][0], rn(g2_1*W[i][1]))
 *                      up : update           0: W - rn(lr*dW)      1: fma(-lr, dW, W)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <xmmintrin.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define NW 14

static const float CODE_L[NW] = {0, 0, 0, 0, 0, 0, 0, 0, .5f, .5f, .5f, .5f, 1, 1};
static const double CODE_R_D[NW] = {0, 0, 1.0 / 3, 1.0 / 3, 2.0 / 3, 2.0 / 3, 1, 1, 0, 0, 1, 1, 0, 1};
static const float CODE_C[NW] = {0, 1, 0, 1, 0, 1, 0, 1, 0, 1, 0, 1, 0, 0};

/* one SGD sample, WW(2,2): w = [W1 4x2 | W2 2x2 | W3 2x1], sample (x[4], y) */
static float sample_step(float *w, const float *x, float y, float lr, int gp, int up) {
    const float *W1 = w, *W2 = w + 8, *W3 = w + 12;
    float h1[2], h2[2];
    for (int j = 0; j < 2; ++j) {
        float acc = x[0] * W1[j];
        for (int k = 1; k < 4; ++k) acc = fmaf(x[k], W1[k * 2 + j], acc);
        h1[j] = acc;
    }
    for (int j = 0; j < 2; ++j) h2[j] = fmaf(h1[1], W2[2 + j], h1[0] * W2[j]);
    float out = 0.0f;
    { float p0 = h2[0] * W3[0]; out = out + p0; float p1 = h2[1] * W3[1]; out = out + p1; }
    float dlt = out - y;
    float loss = dlt * dlt;
    float g3 = 2.0f * dlt;
    float dW[NW];
    dW[12] = h2[0] * g3; dW[13] = h2[1] * g3;
    float g2[2] = {g3 * W3[0], g3 * W3[1]};
    for (int i = 0; i < 2; ++i) for (int j = 0; j < 2; ++j) dW[8 + i * 2 + j] = h1[i] * g2[j];
    float g1[2];
    for (int i = 0; i < 2; ++i) {
        float a = g2[0] * W2[i * 2], b = g2[1] * W2[i * 2 + 1];
        if (gp == 0) g1[i] = fmaf(g2[1], W2[i * 2 + 1], a);
        else if (gp == 1) g1[i] = a + b;
        else g1[i] = fmaf(g2[0], W2[i * 2], b);
    }
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 2; ++j) dW[i * 2 + j] = x[i] * g1[j];
    for (int k = 0; k < NW; ++k) {
        if (up == 0) { float s = lr * dW[k]; w[k] = w[k] - s; }
        else w[k] = fmaf(-lr, dW[k], w[k]);
    }
    return loss;
}

/* one epoch of `w` on the samples of `teacher` in the given order; returns the f64 mean of the f32 sample losses */
double sgd_pin_epoch(float *w, const float *teacher, const uint8_t *order, float lr, int gp, int up) {
    unsigned csr = _mm_getcsr();
    _mm_setcsr(csr | 0x8040u);
    double acc = 0.0;
    for (int q = 0; q < NW; ++q) {
        int k = order[q];
        float x[4] = {teacher[k], CODE_L[k], (float)CODE_R_D[k], CODE_C[k]};
        acc += (double)sample_step(w, x, teacher[k], lr, gp, up);
    }
    _mm_setcsr(csr);
    return acc / NW;
}

typedef struct { double key; int idx; } keyed_t;
static int cmp_keyed(const void *a, const void *b) {
    double x = ((const keyed_t *)a)->key, y = ((const keyed_t *)b)->key;
    return (x > y) - (x < y);
}

/* ---------------------------------------------------------------- exact meet-in-the-middle on the dynamics */
/* A perturbative (pairwise-interaction) model of the order dependence is only good to ~20 % of the interaction terms
 * (lr*J is 0.1..0.3 for these nets), not enough to single out an order.  So the search uses the dynamics themselves: for every split of the 14 samples
 * into a first half A and a second half B, run all orders of A FORWARD from the golden state `a` (FP32, the
 * oracle's arithmetic) and all orders of B BACKWARD from the golden next state `b` (each SGD step inverted in
 * float64 by fixed-point iteration, W_prev = W_next + lr*g_k(W_prev); a contraction because lr*|J| < 1), and match
 * the two clouds of mid-epoch states in all 14 coordinates.  14!/7! = 17.3 M states per side, exhaustive. */
static void grad64(const double *w, const double *x, double y, double *g) {
    const double *W1 = w, *W2 = w + 8, *W3 = w + 12;
    double h1[2], h2[2];
    for (int j = 0; j < 2; ++j) h1[j] = x[0] * W1[j] + x[1] * W1[2 + j] + x[2] * W1[4 + j] + x[3] * W1[6 + j];
    for (int j = 0; j < 2; ++j) h2[j] = h1[0] * W2[j] + h1[1] * W2[2 + j];
    double out = h2[0] * W3[0] + h2[1] * W3[1];
    double g3 = 2.0 * (out - y);
    g[12] = h2[0] * g3; g[13] = h2[1] * g3;
    double g2[2] = {g3 * W3[0], g3 * W3[1]};
    for (int i = 0; i < 2; ++i) for (int j = 0; j < 2; ++j) g[8 + i * 2 + j] = h1[i] * g2[j];
    double g1[2];
    for (int i = 0; i < 2; ++i) g1[i] = g2[0] * W2[i * 2] + g2[1] * W2[i * 2 + 1];
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 2; ++j) g[i * 2 + j] = x[i] * g1[j];
}
/* invert one SGD step: find w_prev with w_prev - lr * g_k(w_prev) = w_next.  Fixed-point iteration first (a
 * contraction when lr * |Hessian| < 1); if it has not settled after 60 rounds, Newton on the 14-dim system with a
 * finite-difference Jacobian (early epochs have samples whose lr * |Hessian| exceeds 1). */
static int solve14(double A[NW][NW], double *b) {   /* Gaussian elimination with partial pivoting, in place */
    for (int c = 0; c < NW; ++c) {
        int piv = c;
        for (int r = c + 1; r < NW; ++r) if (fabs(A[r][c]) > fabs(A[piv][c])) piv = r;
        if (fabs(A[piv][c]) < 1e-300) return 0;
        if (piv != c) {
            for (int k = 0; k < NW; ++k) { double t = A[c][k]; A[c][k] = A[piv][k]; A[piv][k] = t; }
            double t = b[c]; b[c] = b[piv]; b[piv] = t;
        }
        for (int r = c + 1; r < NW; ++r) {
            double f = A[r][c] / A[c][c];
            if (f == 0.0) continue;
            for (int k = c; k < NW; ++k) A[r][k] -= f * A[c][k];
            b[r] -= f * b[c];
        }
    }
    for (int c = NW - 1; c >= 0; --c) {
        double t = b[c];
        for (int k = c + 1; k < NW; ++k) t -= A[c][k] * b[k];
        b[c] = t / A[c][c];
    }
    return 1;
}
static void unstep64(const double *w_next, const double *x, double y, double lr, double *w_prev) {
    double g[NW];
    memcpy(w_prev, w_next, sizeof(double) * NW);
    for (int it = 0; it < 60; ++it) {
        grad64(w_prev, x, y, g);
        double worst = 0;
        for (int k = 0; k < NW; ++k) {
            double nv = w_next[k] + lr * g[k];
            double e = fabs(nv - w_prev[k]);
            if (e > worst) worst = e;
            w_prev[k] = nv;
        }
        if (worst < 1e-13) return;
        if (!(worst < 1e3)) break;
    }
    memcpy(w_prev, w_next, sizeof(double) * NW);
    for (int it = 0; it < 50; ++it) {
        double F[NW], A[NW][NW], gp[NW], wp[NW];
        grad64(w_prev, x, y, g);
        double worst = 0;
        for (int k = 0; k < NW; ++k) { F[k] = -(w_prev[k] - lr * g[k] - w_next[k]); if (fabs(F[k]) > worst) worst = fabs(F[k]); }
        if (worst < 1e-14) return;
        for (int c = 0; c < NW; ++c) {
            memcpy(wp, w_prev, sizeof wp);
            const double h = 1e-7 * (fabs(wp[c]) + 1e-3);
            wp[c] += h;
            grad64(wp, x, y, gp);
            for (int r = 0; r < NW; ++r) A[r][c] = (r == c ? 1.0 : 0.0) - lr * (gp[r] - g[r]) / h;
        }
        if (!solve14(A, F)) return;
        for (int k = 0; k < NW; ++k) w_prev[k] += F[k];
    }
}

typedef struct {
    const float *teacher; float lr; int gp, up;
    const int *el; int len; int count;
    float *fst; double *bst; uint8_t *ord;     /* [count][NW], [count][NW], [count][len] */
} enum_t;

static void fwd_rec(enum_t *E, int depth, int used, const float *w, uint8_t *seq) {
    if (depth == E->len) {
        memcpy(E->fst + (size_t)E->count * NW, w, sizeof(float) * NW);
        memcpy(E->ord + (size_t)E->count * E->len, seq, (size_t)E->len);
        ++E->count;
        return;
    }
    for (int u = 0; u < E->len; ++u) {
        if (used >> u & 1) continue;
        int k = E->el[u];
        float nw[NW];
        memcpy(nw, w, sizeof nw);
        float x[4] = {E->teacher[k], CODE_L[k], (float)CODE_R_D[k], CODE_C[k]};
        sample_step(nw, x, E->teacher[k], E->lr, E->gp, E->up);
        seq[depth] = (uint8_t)k;
        fwd_rec(E, depth + 1, used | 1 << u, nw, seq);
    }
}
static void bwd_rec(enum_t *E, int depth, int used, const double *w, uint8_t *seq) {
    if (depth == E->len) {
        memcpy(E->bst + (size_t)E->count * NW, w, sizeof(double) * NW);
        for (int u = 0; u < E->len; ++u) E->ord[(size_t)E->count * E->len + u] = seq[E->len - 1 - u];   /* execution order */
        ++E->count;
        return;
    }
    for (int u = 0; u < E->len; ++u) {
        if (used >> u & 1) continue;
        int k = E->el[u];
        double x[4] = {(double)E->teacher[k], (double)CODE_L[k], (double)(float)CODE_R_D[k], (double)CODE_C[k]};
        double pw[NW];
        unstep64(w, x, (double)E->teacher[k], (double)E->lr, pw);
        seq[depth] = (uint8_t)k;
        bwd_rec(E, depth + 1, used | 1 << u, pw, seq);
    }
}
static int factorial(int n) { int f = 1; for (int i = 2; i <= n; ++i) f *= i; return f; }

/* a: golden state before the epoch (also the frozen targets), b: golden state after it.  n_fwd of the 14 samples run
 * forward from a, the other 14 - n_fwd backward from b (7 = balanced and fastest; 9 or 10 when the epoch has steps
 * whose inversion amplifies the rounding of b too much -- backward steps expand what forward steps contract).
 * Returns up to topk orders whose mid-epoch states agree best (max-norm `resid`, ascending); tol = match window. */
int sgd_pin_search_exact(const float *a, const float *b, float lr, int gp, int up, int n_fwd, double tol, int topk,
                         uint8_t *out_orders, double *out_resid) {
    if (n_fwd < 1 || n_fwd > 10 || NW - n_fwd > 8) return -1;
    const int n_bwd = NW - n_fwd, NF = factorial(n_fwd), NB = factorial(n_bwd);
    int n_found = 0;
    for (int i = 0; i < topk; ++i) out_resid[i] = INFINITY;
    double b64[NW];
    for (int k = 0; k < NW; ++k) b64[k] = (double)b[k];
#pragma omp parallel
    {
        unsigned csr = _mm_getcsr();
        _mm_setcsr(csr | 0x8040u);
        float *fst = malloc(sizeof(float) * NW * (size_t)NF);
        double *bst = malloc(sizeof(double) * NW * (size_t)NB);
        uint8_t *of = malloc((size_t)NF * n_fwd), *ob = malloc((size_t)NB * n_bwd);
        keyed_t *keys = malloc(sizeof(keyed_t) * (size_t)NF);
#pragma omp for schedule(dynamic, 4)
        for (int mask = 0; mask < (1 << NW); ++mask) {
            if (__builtin_popcount(mask) != n_fwd) continue;
            int A[NW], B[NW], na = 0, nb = 0;
            for (int k = 0; k < NW; ++k) { if (mask >> k & 1) A[na++] = k; else B[nb++] = k; }
            uint8_t seq[NW];
            enum_t EF = {a, lr, gp, up, A, n_fwd, 0, fst, NULL, of};
            fwd_rec(&EF, 0, 0, a, seq);
            enum_t EB = {a, lr, gp, up, B, n_bwd, 0, NULL, bst, ob};
            bwd_rec(&EB, 0, 0, b64, seq);
            /* sort coordinate: largest spread among the forward states of this split */
            int sc = 0; double best = -1;
            for (int k = 0; k < NW; ++k) {
                float lo = fst[k], hi = fst[k];
                for (int p = 1; p < NF; p += 7) { float v = fst[(size_t)p * NW + k]; if (v < lo) lo = v; if (v > hi) hi = v; }
                if ((double)(hi - lo) > best) { best = hi - lo; sc = k; }
            }
            for (int p = 0; p < NF; ++p) { keys[p].key = (double)fst[(size_t)p * NW + sc]; keys[p].idx = p; }
            qsort(keys, (size_t)NF, sizeof(keyed_t), cmp_keyed);
            for (int q = 0; q < NB; ++q) {
                const double *bq = bst + (size_t)q * NW;
                double want = bq[sc];
                int lo = 0, hi = NF;
                while (lo < hi) { int mid = (lo + hi) / 2; if (keys[mid].key < want - tol) lo = mid + 1; else hi = mid; }
                for (int r = lo; r < NF && keys[r].key <= want + tol; ++r) {
                    int p = keys[r].idx;
                    const float *fp = fst + (size_t)p * NW;
                    double worst = 0;
                    for (int k = 0; k < NW; ++k) {
                        double e = fabs((double)fp[k] - bq[k]);
                        if (e > worst) { worst = e; if (worst > tol) break; }
                    }
                    if (worst > tol) continue;
#pragma omp critical
                    {
                        if (worst < out_resid[topk - 1]) {
                            int pos = topk - 1;
                            while (pos > 0 && out_resid[pos - 1] > worst) {
                                out_resid[pos] = out_resid[pos - 1];
                                memcpy(out_orders + (size_t)pos * NW, out_orders + (size_t)(pos - 1) * NW, NW);
                                --pos;
                            }
                            out_resid[pos] = worst;
                            memcpy(out_orders + (size_t)pos * NW, of + (size_t)p * n_fwd, (size_t)n_fwd);
                            memcpy(out_orders + (size_t)pos * NW + n_fwd, ob + (size_t)q * n_bwd, (size_t)n_bwd);
                            if (n_found < topk) ++n_found;
                        }
                    }
                }
            }
        }
        free(fst); free(bst); free(of); free(ob); free(keys);
        _mm_setcsr(csr);
    }
    return n_found;
}
