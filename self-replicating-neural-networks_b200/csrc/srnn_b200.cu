// srnn_b200.cu -- kernels and C-ABI entry points of libsrnn_b200.so (sm_100a only).
//
// Build (see __graft_entry__.build):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -ftz=true
//        -prec-div=true -prec-sqrt=true -Xcompiler -fPIC -shared --split-compile 0 -o libsrnn_b200.so srnn_b200.cu
//
// Execution model: one thread owns one net.  A CTA of 256 threads stages its tile of 256 contiguous
// records (256 * n_weights floats) between HBM and shared memory with fully coalesced accesses, then
// every thread pulls its own record into registers, runs the whole episode (100 self-applications,
// 1000 SGD epochs, one soup generation ...) without touching memory again, and the tile goes back the
// same way.  FP32 CUDA cores only: the per-net matrices are 4x2 / 2x2 / 2x1 -- nothing a tcgen05 tile
// could chew on -- so the bound is the FP32 issue rate for iterated kernels and HBM for one-pass ones.
#include <cuda_runtime.h>

#include <cfloat>
#include <cmath>
#include <cstdio>
#include <algorithm>
#include <cstring>
#include <mutex>
#include <dlfcn.h>

#include "../../include/srnn_b200.h"
#include "srnn_device.cuh"
#include "srnn_generic.cuh"

namespace srnn {

constexpr int BS = 256;  // threads per CTA = records per tile
#ifndef SRNN_SOUP_BS
#define SRNN_SOUP_BS 128
#endif
// CTA size / minimum resident CTAs of the fused soup generation kernel.  The kernel is latency-bound on its short
// serial phases (one warp per CTA walking attack lists or running a learner's SGD while the others wait at the
// barrier), so it wants many resident CTAs: A/B on a B200 at 10 M particles, ms per generation (train = 0 / 1):
// 256 thr x>=1: 0.93 / 1.38;  128 x>=1: 0.98 / 1.45;  64 x>=1: 1.11 / 1.66;  256 x>=6: 0.79 / 1.24;  128 x>=10: 0.73 / 1.18.
constexpr int SOUP_BS = SRNN_SOUP_BS;
// engine_path = 0 (auto) takes the fused single kernel; the multi-pass (globally compacted) generation is kept
// selectable (engine_path = 2): it measured 0.91 / 1.33 ms against 0.73 / 1.18 ms for the fused kernel at 10 M particles.
constexpr int64_t SOUP_MULTIPASS_MIN = INT64_MAX;
// srnn_soup_sync_run chains generations (one launch each) up to this population size
#ifndef SRNN_SOUP_CHAIN_MAX
#define SRNN_SOUP_CHAIN_MAX (1 << 21)
#endif
constexpr int64_t SOUP_CHAIN_MAX = SRNN_SOUP_CHAIN_MAX;
// largest unsharded soup the opt-in persistent cooperative variant of srnn_soup_sync_run accepts
constexpr int64_t SOUP_PERSISTENT_MAX = 1 << 20;
#ifndef SRNN_EXACT_SPLIT_MIN
#define SRNN_EXACT_SPLIT_MIN (1 << 22)   // measured: 10 M particles 1.444 -> 1.392 ms per generation, 1.1 M 0.226 -> 0.238 ms
#endif
constexpr int64_t SOUP_EXACT_SPLIT_MIN = SRNN_EXACT_SPLIT_MIN;

// ------------------------------------------------------------------------------------------------ errors
static thread_local char g_err[512] = "";
static int fail(int code, const char *fmt, const char *detail = "") {
    snprintf(g_err, sizeof g_err, fmt, detail);
    return code;
}
static int check_launch(const char *what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(g_err, sizeof g_err, "%s: %s", what, cudaGetErrorString(e));
        return SRNN_ERR_CUDA;
    }
    return SRNN_OK;
}

// R6: predicates compare f32 values with the f64 epsilon.  For an f32 value d:  d >= eps  <=>  d >= eps_ge
// (smallest f32 >= eps) and d <= eps  <=>  d <= eps_le (largest f32 <= eps), so the kernels compare in f32.
struct Eps { float ge, le; };
static int make_eps(double eps, Eps *e) {
    if (!(eps >= (double)FLT_MIN) || !(eps <= (double)FLT_MAX))
        return fail(SRNN_ERR_ARG, "epsilon must be a normal float32 magnitude (FLT_MIN <= eps <= FLT_MAX)%s");
    float f = (float)eps;
    e->ge = ((double)f < eps) ? nextafterf(f, INFINITY) : f;
    e->le = ((double)f > eps) ? nextafterf(f, 0.0f) : f;
    return SRNN_OK;
}

// ------------------------------------------------------------------------------------------------ tile staging
// A CTA's tile = NT consecutive records = one contiguous run of NT*NW floats in a packed arena.  It moves between HBM
// and shared memory as a flat copy (16-byte vectors when the run is 16-byte aligned) -- no per-element index maths --
// and each thread reads / writes its own record with VEC-wide shared-memory accesses: VEC = 4, 2 or 1 floats,
// the widest that divides NW, which is bank-conflict free for the benchmarked records (14 -> LDS.64, 20 -> LDS.128,
// 17 -> LDS.32 at an odd stride).
template <int NW> struct TileVec { static constexpr int V = NW % 4 == 0 ? 4 : (NW % 2 == 0 ? 2 : 1); };

// CG = true: loads bypass L1 (ld.global.cg) -- for the persistent soup kernel, which re-reads arenas that other CTAs
// rewrote earlier in the SAME launch.
template <int NT, bool CG = false>
__device__ __forceinline__ void flat_copy(float *__restrict__ dst, const float *__restrict__ src, int total) {
    const int tid = threadIdx.x;
    if ((((uintptr_t)dst | (uintptr_t)src) & 15) == 0) {
        const int nv = total >> 2;
        const float4 *s4 = reinterpret_cast<const float4 *>(src);
        float4 *d4 = reinterpret_cast<float4 *>(dst);
        for (int e = tid; e < nv; e += NT) d4[e] = CG ? __ldcg(s4 + e) : s4[e];
        for (int e = (nv << 2) + tid; e < total; e += NT) dst[e] = CG ? __ldcg(src + e) : src[e];
    } else {
        for (int e = tid; e < total; e += NT) dst[e] = CG ? __ldcg(src + e) : src[e];
    }
}
// Asynchronous global -> shared copies (LDGSTS): no staging registers, and the issuing thread carries on -- the soup
// tile kernel computes its pairing draws while its records are in flight.  Completion: async_copy_wait() + barrier.
__device__ __forceinline__ void async_copy16(float *dst_smem, const float *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void async_copy8(float *dst_smem, const float *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void async_copy4(float *dst_smem, const float *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void async_copy_wait() {
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}
// flat_copy's asynchronous twin for a packed arena (caller checked ld == NW); dst is 16-byte aligned shared memory
template <int NT>
__device__ __forceinline__ void flat_copy_async(float *dst, const float *__restrict__ src, int total) {
    const int tid = threadIdx.x;
    if (((uintptr_t)src & 15) == 0) {
        const int nv = total >> 2;
        for (int e = tid; e < nv; e += NT) async_copy16(dst + 4 * e, src + 4 * e);
        for (int e = (nv << 2) + tid; e < total; e += NT) async_copy4(dst + e, src + e);
    } else {
        for (int e = tid; e < total; e += NT) async_copy4(dst + e, src + e);
    }
}
// one record global -> shared, asynchronously, in 64-bit words when its address allows (else 32-bit)
template <int NW>
__device__ __forceinline__ void record_copy_async(float *dst_smem, const float *src) {
    if (NW % 2 == 0 && (reinterpret_cast<uintptr_t>(src) & 7) == 0) {
#pragma unroll
        for (int k = 0; k < NW; k += 2) async_copy8(dst_smem + k, src + k);
    } else {
#pragma unroll
        for (int k = 0; k < NW; ++k) async_copy4(dst_smem + k, src + k);
    }
}
template <int NW>
__device__ __forceinline__ void tile_get(const float *tile, int r, float *reg) {
    constexpr int V = TileVec<NW>::V;
    const float *row = tile + r * NW;
    if constexpr (V == 4) {
#pragma unroll
        for (int k = 0; k < NW; k += 4) {
            const float4 v = *reinterpret_cast<const float4 *>(row + k);
            reg[k] = v.x; reg[k + 1] = v.y; reg[k + 2] = v.z; reg[k + 3] = v.w;
        }
    } else if constexpr (V == 2) {
#pragma unroll
        for (int k = 0; k < NW; k += 2) {
            const float2 v = *reinterpret_cast<const float2 *>(row + k);
            reg[k] = v.x; reg[k + 1] = v.y;
        }
    } else {
#pragma unroll
        for (int k = 0; k < NW; ++k) reg[k] = row[k];
    }
}
template <int NW>
__device__ __forceinline__ void tile_put(float *tile, int r, const float *reg) {
    constexpr int V = TileVec<NW>::V;
    float *row = tile + r * NW;
    if constexpr (V == 4) {
#pragma unroll
        for (int k = 0; k < NW; k += 4) *reinterpret_cast<float4 *>(row + k) = make_float4(reg[k], reg[k + 1], reg[k + 2], reg[k + 3]);
    } else if constexpr (V == 2) {
#pragma unroll
        for (int k = 0; k < NW; k += 2) *reinterpret_cast<float2 *>(row + k) = make_float2(reg[k], reg[k + 1]);
    } else {
#pragma unroll
        for (int k = 0; k < NW; ++k) row[k] = reg[k];
    }
}
// whole-tile staging, HBM -> shared (any record stride)
// NT threads stage TILE consecutive records (TILE a multiple of NT)
template <int NW, int NT, bool CG = false, int TILE = NT>
__device__ __forceinline__ void tile_stage_in(const float *__restrict__ w, int64_t ld, int64_t base, int64_t n, float *tile) {
    if constexpr (TILE == NT) {   // one record per thread: kept apart, the general form below costs registers
        const int cnt = (int)min((int64_t)NT, n - base);   // (RNN train kernel 127 -> 144, 7 % slower)
        if (ld == NW) {
            flat_copy<NT, CG>(tile, w + base * NW, cnt * NW);
        } else if ((int)threadIdx.x < cnt) {
            const float *src = w + (base + threadIdx.x) * ld;
#pragma unroll
            for (int k = 0; k < NW; ++k) tile[threadIdx.x * NW + k] = CG ? __ldcg(src + k) : __ldg(src + k);
        }
    } else {
        const int cnt = (int)max((int64_t)0, min((int64_t)TILE, n - base));
        if (ld == NW) {
            flat_copy<NT, CG>(tile, w + base * NW, cnt * NW);
        } else {
            for (int v = threadIdx.x; v < cnt; v += NT) {
                const float *src = w + (base + v) * ld;
#pragma unroll
                for (int k = 0; k < NW; ++k) tile[v * NW + k] = CG ? __ldcg(src + k) : __ldg(src + k);
            }
        }
    }
}
template <int NW, int NT, int TILE = NT>
__device__ __forceinline__ void tile_stage_out(float *__restrict__ w, int64_t ld, int64_t base, int64_t n, const float *tile) {
    if constexpr (TILE == NT) {
        const int cnt = (int)min((int64_t)NT, n - base);
        if (ld == NW) {
            flat_copy<NT>(w + base * NW, tile, cnt * NW);
        } else if ((int)threadIdx.x < cnt) {
            float *dst = w + (base + threadIdx.x) * ld;
#pragma unroll
            for (int k = 0; k < NW; ++k) dst[k] = tile[threadIdx.x * NW + k];
        }
    } else {
        const int cnt = (int)max((int64_t)0, min((int64_t)TILE, n - base));
        if (ld == NW) {
            flat_copy<NT>(w + base * NW, tile, cnt * NW);
        } else {
            for (int v = threadIdx.x; v < cnt; v += NT) {
                float *dst = w + (base + v) * ld;
#pragma unroll
                for (int k = 0; k < NW; ++k) dst[k] = tile[v * NW + k];
            }
        }
    }
}
// tile + registers in one go (the register-resident kernels)
template <int NW>
__device__ __forceinline__ void tile_load(const float *__restrict__ w, int64_t ld, int64_t base, int64_t n,
                                          float *tile, float *reg) {
    tile_stage_in<NW, BS>(w, ld, base, n, tile);
    __syncthreads();
    if (base + threadIdx.x < n) tile_get<NW>(tile, threadIdx.x, reg);
}
template <int NW>
__device__ __forceinline__ void tile_store(float *__restrict__ w, int64_t ld, int64_t base, int64_t n, float *tile,
                                           const float *reg) {
    __syncthreads();  // tile may still be read by slower threads of a previous phase
    if (base + threadIdx.x < n) tile_put<NW>(tile, threadIdx.x, reg);
    __syncthreads();
    tile_stage_out<NW, BS>(w, ld, base, n, tile);
}
// One record to / from registers by one thread (gathers and scatters by index), in the widest words its address
// allows: H leading floats bring the address to a 16-byte boundary (one 32-bit and / or one 64-bit word), then 128-bit
// words, then the tail.  A WW(2,2) record (56 B) moves in 4 instructions instead of 14, an AGG(4,2,2) record in 5
// instead of 20 (pair-apply kernel, 2^24 pairs in arena order: 0.86 -> 0.46 ms).
enum { LD_PLAIN = 0, LD_RO = 1, LD_CG = 2 };   // ordinary load / read-only path / L1-bypassing
template <int MODE, class T>
__device__ __forceinline__ T ld_as(const float *p) {
    const T *q = reinterpret_cast<const T *>(p);
    return MODE == LD_RO ? __ldg(q) : MODE == LD_CG ? __ldcg(q) : *q;
}
template <int NW, int H, int MODE>
__device__ __forceinline__ void record_get_h(const float *src, float *reg) {
    int k = 0;
    if (H & 1) { reg[0] = ld_as<MODE, float>(src); k = 1; }
    if (H & 2) {
        const float2 v = ld_as<MODE, float2>(src + k);
        reg[k] = v.x; reg[k + 1] = v.y;
        k += 2;
    }
#pragma unroll
    for (int q = 0; q < (NW - H) / 4; ++q, k += 4) {
        const float4 v = ld_as<MODE, float4>(src + k);
        reg[k] = v.x; reg[k + 1] = v.y; reg[k + 2] = v.z; reg[k + 3] = v.w;
    }
    if ((NW - H) & 2) {
        const float2 v = ld_as<MODE, float2>(src + k);
        reg[k] = v.x; reg[k + 1] = v.y;
        k += 2;
    }
    if ((NW - H) & 1) reg[k] = ld_as<MODE, float>(src + k);
}
template <int NW, int H>
__device__ __forceinline__ void record_put_h(float *dst, const float *reg) {
    int k = 0;
    if (H & 1) { dst[0] = reg[0]; k = 1; }
    if (H & 2) { *reinterpret_cast<float2 *>(dst + k) = make_float2(reg[k], reg[k + 1]); k += 2; }
#pragma unroll
    for (int q = 0; q < (NW - H) / 4; ++q, k += 4)
        *reinterpret_cast<float4 *>(dst + k) = make_float4(reg[k], reg[k + 1], reg[k + 2], reg[k + 3]);
    if ((NW - H) & 2) { *reinterpret_cast<float2 *>(dst + k) = make_float2(reg[k], reg[k + 1]); k += 2; }
    if ((NW - H) & 1) dst[k] = reg[k];
}
template <int NW, int MODE>
__device__ __forceinline__ void record_get(const float *src, float *reg) {
    static_assert(NW >= 3, "record shorter than the longest alignment head");
    switch ((4 - (int)((reinterpret_cast<uintptr_t>(src) >> 2) & 3)) & 3) {   // floats up to the next 16-byte boundary
    case 0: record_get_h<NW, 0, MODE>(src, reg); break;
    case 1: record_get_h<NW, 1, MODE>(src, reg); break;
    case 2: record_get_h<NW, 2, MODE>(src, reg); break;
    default: record_get_h<NW, 3, MODE>(src, reg); break;
    }
}
// The soup's partner gathers sit on the latency-critical path of a tile: there the four-way alignment switch costs
// more (lanes of a warp diverge by record parity, 0.68 -> 0.88 ms per generation at 10 M particles) than the wider
// words save.  This variant takes 64-bit words when the record is 8-byte aligned -- the same answer for every lane of a
// packed arena with an even number of weights, so the branch is uniform -- and 32-bit words otherwise.
template <int NW, int MODE>
__device__ __forceinline__ void record_get_uniform(const float *src, float *reg) {
    if (NW % 2 == 0 && (reinterpret_cast<uintptr_t>(src) & 7) == 0) {
#pragma unroll
        for (int k = 0; k < NW; k += 2) {
            const float2 v = ld_as<MODE, float2>(src + k);
            reg[k] = v.x; reg[k + 1] = v.y;
        }
    } else {
#pragma unroll
        for (int k = 0; k < NW; ++k) reg[k] = ld_as<MODE, float>(src + k);
    }
}
template <int NW>
__device__ __forceinline__ void record_load(const float *__restrict__ w, int64_t ld, int64_t i, float *reg) {
    record_get<NW, LD_RO>(w + i * ld, reg);
}
// 32-bit words only: for stores inside hot loops that are usually skipped (trajectory logging), where the alignment
// switch of record_store would cost the loop registers (general self-apply kernel: 60 -> 72)
template <int NW>
__device__ __forceinline__ void record_store_scalar(float *__restrict__ w, int64_t ld, int64_t i, const float *reg) {
    float *dst = w + i * ld;
#pragma unroll
    for (int k = 0; k < NW; ++k) dst[k] = reg[k];
}
template <int NW>
__device__ __forceinline__ void record_store(float *__restrict__ w, int64_t ld, int64_t i, const float *reg) {
    static_assert(NW >= 3, "record shorter than the longest alignment head");
    float *dst = w + i * ld;
    switch ((4 - (int)((reinterpret_cast<uintptr_t>(dst) >> 2) & 3)) & 3) {
    case 0: record_put_h<NW, 0>(dst, reg); break;
    case 1: record_put_h<NW, 1>(dst, reg); break;
    case 2: record_put_h<NW, 2>(dst, reg); break;
    default: record_put_h<NW, 3>(dst, reg); break;
    }
}

__device__ __forceinline__ void block_histogram(int cls, bool active, unsigned int *sh /*[5]*/,
                                                unsigned long long *counters) {
    if (threadIdx.x < 5) sh[threadIdx.x] = 0;
    __syncthreads();
    if (active) atomicAdd(&sh[cls], 1u);
    __syncthreads();
    if (threadIdx.x < 5 && sh[threadIdx.x]) atomicAdd(&counters[threadIdx.x], (unsigned long long)sh[threadIdx.x]);
}

// ================================================================================================ K1 self-application
// FAST: n_steps unconditional self-applications, nothing else in the loop (the benchmarked kernel).
template <class S>
__global__ void __launch_bounds__(BS) self_apply_fixed_kernel(float *__restrict__ w, int64_t ld, int64_t n,
                                                              int n_steps) {
    __shared__ __align__(16) float tile[BS * S::NW];
    const int64_t base = (int64_t)blockIdx.x * BS;
    float a[S::NW], b[S::NW];
    tile_load<S::NW>(w, ld, base, n, tile, a);
    if (base + threadIdx.x < n) {
        const float one = opaque_one();
        // all applications but the last may leave zeros with either sign (ww_apply, ANYZERO): nobody sees those states,
        // and -- for the shapes where that holds -- the exact last application restores the reference's result bit for bit
        constexpr bool AZ = exact_step_ignores_zero_signs<S>();
        int s = 0;
        for (; s + 2 <= n_steps - 1; s += 2) {  // ping-pong: no register copies between steps
            net_apply<S, AZ>(a, a, b, one);
            net_apply<S, AZ>(b, b, a, one);
        }
        if (s < n_steps - 1) {
            net_apply<S, AZ>(a, a, b, one);
#pragma unroll
            for (int k = 0; k < S::NW; ++k) a[k] = b[k];
        }
        if (n_steps > 0) {
            net_apply<S>(a, a, b, one);
#pragma unroll
            for (int k = 0; k < S::NW; ++k) a[k] = b[k];
        }
    }
    tile_store<S::NW>(w, ld, base, n, tile, a);
}

// GENERAL: optional run_net stop rule, per-step trajectory, fused classification.
template <class S>
__global__ void __launch_bounds__(BS) self_apply_general_kernel(float *__restrict__ w, int64_t ld, int64_t n,
                                                                int n_steps, int mode, Eps eps, int with_fix_sec,
                                                                int *__restrict__ steps_out,
                                                                uint8_t *__restrict__ cls_out,
                                                                unsigned long long *counters,
                                                                float *__restrict__ traj) {
    __shared__ __align__(16) float tile[BS * S::NW];
    __shared__ unsigned int hist[5];
    const int64_t base = (int64_t)blockIdx.x * BS, i = base + threadIdx.x;
    const bool active = i < n;
    float cur[S::NW], nxt[S::NW];
    tile_load<S::NW>(w, ld, base, n, tile, cur);
    int it = 0, cls = SRNN_CLS_OTHER;
    if (active && mode == SRNN_RUN_FIXED && !traj) {
        // applying-fixpoints.py flow (n_steps x self_attack, then count): the bare ping-pong loop of the fixed kernel,
        // and one more application afterwards for the classifier's fixpoint test
        const float one = opaque_one();
        int s = 0;
        constexpr bool AZ = exact_step_ignores_zero_signs<S>();
        for (; s + 2 <= n_steps - 1; s += 2) {   // inner steps: zeros of either sign (ww_apply, ANYZERO); last step exact
            net_apply<S, AZ>(cur, cur, nxt, one);
            net_apply<S, AZ>(nxt, nxt, cur, one);
        }
        if (s < n_steps - 1) {
            net_apply<S, AZ>(cur, cur, nxt, one);
#pragma unroll
            for (int k = 0; k < S::NW; ++k) cur[k] = nxt[k];
        }
        if (n_steps > 0) {
            net_apply<S>(cur, cur, nxt, one);
#pragma unroll
            for (int k = 0; k < S::NW; ++k) cur[k] = nxt[k];
        }
        it = n_steps;
        if (steps_out) steps_out[i] = it;
        if (cls_out || counters) {
            net_apply<S>(cur, cur, nxt, one);
            cls = classify_with<S>(cur, nxt, eps.ge, eps.le, with_fix_sec);
            if (cls_out) cls_out[i] = (uint8_t)cls;
        }
    } else if (active) {
        const float one = opaque_one();
        net_apply<S>(cur, cur, nxt, one);  // look-ahead: is_fixpoint's application IS the next self_attack
        // max |w_k| of a state is reduced once, when the state is produced: it is the divergence test of is_fixpoint's
        // result in this step and the divergence test of the net itself in the next
        float m_cur = max_abs_nan<S::NW>(cur), m_nxt = max_abs_nan<S::NW>(nxt);
        while (it < n_steps) {
            if (mode == SRNN_RUN_UNTIL &&
                (diverged_from_max(m_cur) ||
                 (!diverged_from_max(m_nxt) && !(max_abs_diff<S::NW>(nxt, cur) >= eps.ge))))
                break;
#pragma unroll
            for (int k = 0; k < S::NW; ++k) cur[k] = nxt[k];   // (a ping-pong of the two buffers measured 1.5 % slower here)
            m_cur = m_nxt;
            ++it;
            if (traj) record_store_scalar<S::NW>(traj + (int64_t)(it - 1) * n * ld, ld, i, cur);
            net_apply<S>(cur, cur, nxt, one);
            m_nxt = max_abs_nan<S::NW>(nxt);
        }
        if (steps_out) steps_out[i] = it;
        if (cls_out || counters) {
            cls = classify_with<S>(cur, nxt, eps.ge, eps.le, with_fix_sec);
            if (cls_out) cls_out[i] = (uint8_t)cls;
        }
    }
    if (counters) block_histogram(cls, active, hist, counters);
    tile_store<S::NW>(w, ld, base, n, tile, cur);
}

// ================================================================================================ K2/K3/K4 apply
template <class S>
__global__ void __launch_bounds__(BS) apply_kernel(const float *op_w, int64_t op_ld,
                                                   const int32_t *__restrict__ op_idx, const float *in_w,
                                                   int64_t in_ld, const int32_t *__restrict__ in_idx, float *out_w,
                                                   int64_t out_ld, const int32_t *__restrict__ out_idx, int64_t n) {
    const int64_t p = (int64_t)blockIdx.x * BS + threadIdx.x;
    if (p >= n) return;
    float op[S::NW], in[S::NW], out[S::NW];
    {   // op / in may alias out (attack on self): plain loads, no read-only path
        const float *so = op_w + (op_idx ? (int64_t)op_idx[p] : p) * op_ld;
        const float *si = in_w + (in_idx ? (int64_t)in_idx[p] : p) * in_ld;
        record_get<S::NW, LD_PLAIN>(so, op);
        record_get<S::NW, LD_PLAIN>(si, in);
    }
    net_apply<S>(op, in, out, opaque_one());
    record_store<S::NW>(out_w, out_ld, out_idx ? out_idx[p] : p, out);
}

// ================================================================================================ raw forward
// NeuralNetwork.apply(*inputs) = model.predict on one row (code/network.py:235-237, :355-357, :540-542)
template <class S>
__global__ void __launch_bounds__(BS) forward_kernel(const float *__restrict__ w, int64_t ld,
                                                     const int32_t *__restrict__ idx, const float *__restrict__ x,
                                                     float *__restrict__ y, int64_t n) {
    const int64_t p = (int64_t)blockIdx.x * BS + threadIdx.x;
    if (p >= n) return;
    float op[S::NW];
    record_load<S::NW>(w, ld, idx ? idx[p] : p, op);
    if constexpr (S::KIND == SRNN_KIND_RNN) {
        float seq[S::NW], out[S::NW];
        record_get<S::NW, LD_RO>(x + p * S::NW, seq);
        rnn_forward_seq<S, false>(op, seq, out, nullptr);
        record_store<S::NW>(y, S::NW, p, out);
    } else {
        float in[S::IO], out[S::OUTW];
        if constexpr (S::IO >= 3) {
            record_get<S::IO, LD_RO>(x + p * S::IO, in);
        } else {
#pragma unroll
            for (int q = 0; q < S::IO; ++q) in[q] = __ldg(x + p * S::IO + q);
        }
        mlp_forward<S>(op, in, out);
        if constexpr (S::OUTW >= 3) {
            record_store<S::OUTW>(y, S::OUTW, p, out);
        } else {
#pragma unroll
            for (int q = 0; q < S::OUTW; ++q) y[p * S::OUTW + q] = out[q];
        }
    }
}

// ================================================================================================ K6 classify
template <class S>
__global__ void __launch_bounds__(BS) classify_kernel(const float *__restrict__ w, int64_t ld, int64_t n, Eps eps,
                                                      int with_fix_sec, uint8_t *__restrict__ cls_out,
                                                      uint8_t *__restrict__ flags_out, unsigned long long *counters) {
    __shared__ __align__(16) float tile[BS * S::NW];
    __shared__ unsigned int hist[5];
    const int64_t base = (int64_t)blockIdx.x * BS, i = base + threadIdx.x;
    const bool active = i < n;
    float cur[S::NW], nxt[S::NW];
    tile_load<S::NW>(w, ld, base, n, tile, cur);
    int cls = SRNN_CLS_OTHER;
    if (active) {
        net_apply<S>(cur, cur, nxt, opaque_one());
        cls = classify_with<S>(cur, nxt, eps.ge, eps.le, with_fix_sec);
        if (cls_out) cls_out[i] = (uint8_t)cls;
        if (flags_out) {
            float nn[S::NW];
            net_apply<S>(cur, nxt, nn, opaque_one());
            flags_out[i] = (uint8_t)((is_diverged<S::NW>(cur) ? 1 : 0) | (is_zero<S::NW>(cur, eps.le) ? 2 : 0) |
                                     (is_close<S::NW>(nxt, cur, eps.ge) ? 4 : 0) |
                                     (is_close<S::NW>(nn, cur, eps.ge) ? 8 : 0));
        }
    }
    if (counters) block_histogram(cls, active, hist, counters);
}

// ================================================================================================ K7 init / respawn
template <class S>
__global__ void __launch_bounds__(BS) init_kernel(float *__restrict__ w, int64_t ld, int64_t n,
                                                  const int32_t *__restrict__ idx, uint64_t seed, uint64_t slot0,
                                                  uint32_t gen, Limits lim) {
    __shared__ __align__(16) float tile[BS * S::NW];
    const int64_t base = (int64_t)blockIdx.x * BS, p = base + threadIdx.x;
    float cur[S::NW];
    if (!idx && ld == S::NW) {  // whole population: coalesced through the tile
        if (p < n) init_net<S>(seed, slot0 + (uint64_t)p, gen, lim, cur);
        tile_store<S::NW>(w, ld, base, n, tile, cur);
        return;
    }
    if (p >= n) return;
    const int64_t i = idx ? idx[p] : p;
    init_net<S>(seed, slot0 + (uint64_t)i, gen, lim, cur);
    record_store<S::NW>(w, ld, i, cur);
}

// ================================================================================================ K5 train / learn_from
// SGD scratch columns live in dynamic shared memory: tgt float[NW][BS] | ord uint8[NS][BS] | codes float[3*NW]
template <class S> __host__ __device__ constexpr size_t sgd_smem_bytes(int threads) {
    return sizeof(float) * S::NW * threads + (size_t)((S::NS * threads + 15) / 16 * 16) + sizeof(float) * 3 * S::NW;
}
template <class S>
__device__ __forceinline__ SgdScratch sgd_scratch(unsigned char *smem, int threads) {
    SgdScratch sc;
    sc.tgt = reinterpret_cast<float *>(smem) + threadIdx.x;
    sc.ord = smem + sizeof(float) * S::NW * threads + threadIdx.x;
    sc.stride = threads;
    float *codes = reinterpret_cast<float *>(smem + sizeof(float) * S::NW * threads + (S::NS * threads + 15) / 16 * 16);
    if (threadIdx.x == 0) fill_codes<S>(codes);
    sc.codes = codes;
    sc.one = opaque_one();
    return sc;
}

template <class S>
__global__ void __launch_bounds__(BS) train_kernel(float *__restrict__ w, int64_t ld, int64_t n, int n_epochs,
                                                   float lr, int perm_mode, const uint8_t *__restrict__ perm,
                                                   uint64_t seed, uint64_t slot0, uint64_t epoch0,
                                                   const float *__restrict__ teacher_w, int64_t teacher_ld,
                                                   const int32_t *__restrict__ teacher_idx,
                                                   double *__restrict__ loss_out) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ __align__(16) float tile[BS * S::NW];
    const int64_t base = (int64_t)blockIdx.x * BS, i = base + threadIdx.x;
    const bool active = i < n;
    SgdScratch sc = sgd_scratch<S>(smem, BS);
    float cur[S::NW];
    tile_load<S::NW>(w, ld, base, n, tile, cur);
    __syncthreads();  // codes visible
    if (active) {
        if (teacher_w) {
            // scalar on purpose: the wide-word variants cost registers here (RNN: 127 -> 144) and gain nothing, the
            // epoch that follows dominates.  Plain loads: the teacher may live in the arena being trained.
            const float *src = teacher_w + (teacher_idx ? (int64_t)teacher_idx[i] : i) * teacher_ld;
#pragma unroll
            for (int k = 0; k < S::NW; ++k) sc.tgt[k * sc.stride] = src[k];
        }
        double loss = 0.0;
        const bool ordered = S::NS > 1 && perm_mode != SRNN_PERM_NONE;
        if constexpr (S::KIND == SRNN_KIND_WW && S::NW <= 16) {
            // fast path: sample loop unrolled, order in one register, Philox block shared by EPB epochs
            U4 blk;
            uint64_t blk_id = ~0ull;
            for (int e = 0; e < n_epochs; ++e) {
                if (!teacher_w) {  // targets frozen at the start of the epoch, code/network.py:615
#pragma unroll
                    for (int k = 0; k < S::NW; ++k) sc.tgt[k * sc.stride] = cur[k];
                }
                uint64_t order = 0;
                if (perm_mode == SRNN_PERM_EXPLICIT) {
#pragma unroll
                    for (int q = 0; q < S::NS; ++q) order |= (uint64_t)(perm[(int64_t)e * S::NS + q] & 15u) << (4 * q);
                } else if (perm_mode == SRNN_PERM_PHILOX) {
                    uint32_t r[PermPlan<S::NS>::W];
                    perm_words<S::NS>(seed, slot0 + (uint64_t)i, epoch0 + (uint64_t)e, r, blk, blk_id);
                    order = perm_nibbles<S::NS>(r);
                }
                const bool last = loss_out != nullptr && e == n_epochs - 1;
                if (!ordered) loss = ww_epoch_unrolled<S, false, true>(cur, sc.tgt, sc.stride, sc.codes, 0, lr, sc.one);
                else if (last) loss = ww_epoch_unrolled<S, true, true>(cur, sc.tgt, sc.stride, sc.codes, order, lr, sc.one);
                else ww_epoch_unrolled<S, true, false>(cur, sc.tgt, sc.stride, sc.codes, order, lr, sc.one);
            }
        } else {
            for (int e = 0; e < n_epochs; ++e) {
                if (!teacher_w) {  // targets frozen at the start of the epoch, code/network.py:615
#pragma unroll
                    for (int k = 0; k < S::NW; ++k) sc.tgt[k * sc.stride] = cur[k];
                }
                if (ordered) {
                    if (perm_mode == SRNN_PERM_EXPLICIT) {
                        for (int q = 0; q < S::NS; ++q) sc.ord[q * sc.stride] = perm[(int64_t)e * S::NS + q];
                    } else {
                        philox_perm<S::NS>(seed, slot0 + (uint64_t)i, epoch0 + (uint64_t)e, sc.ord, sc.stride);
                    }
                }
                loss = train_epoch<S>(cur, sc, ordered, lr, loss_out != nullptr && e == n_epochs - 1);
            }
        }
        if (loss_out) loss_out[i] = loss;
    }
    tile_store<S::NW>(w, ld, base, n, tile, cur);
}

// ================================================================================================ K9 known-fixpoint variation
struct BaseNet { float v[SRNN_MAX_WEIGHTS]; };

// One thread per trial, the warp runs until its slowest lane ends.  (A variant that refills finished lanes with new
// trials of a per-warp chunk -- ballot + prefix rank, batched refills -- was measured SLOWER, 5.5e10 vs 7.2e10
// executed self-applications/s: starting a trial costs as much as several steps and the refill logic defeats the
// straight-line step loop.)
// ANYZERO: no weights are returned (w_out == nullptr), only step counts -- every test in the loop looks at magnitudes, so
// the applications may leave zeros with either sign (ww_apply) and save one instruction per pair of weight points.
template <class S, bool ANYZERO>
__global__ void __launch_bounds__(BS) kfv_kernel(BaseNet basenet, int64_t n, double scale, uint32_t level,
                                                 int max_steps, Eps eps, uint64_t seed, uint64_t slot0,
                                                 int *__restrict__ ys, int *__restrict__ zs,
                                                 unsigned long long *hist, float *__restrict__ w_out, int64_t ld) {
    extern __shared__ unsigned int hist_s[];  // [2][max_steps + 1], only when hist != nullptr
    const int64_t i = (int64_t)blockIdx.x * BS + threadIdx.x;
    const int nbins = 2 * (max_steps + 1);
    if (hist) {
        for (int b = threadIdx.x; b < nbins; b += BS) hist_s[b] = 0;
        __syncthreads();
    }
    int tts = 0, taf = 0;
    if (i < n) {
    float cur[S::NW], nxt[S::NW];
    // vary(), known-fixpoint-variation.py:37-46: float64 perturbation stored as float32
#pragma unroll
    for (int k = 0; k < S::NW; k += 2) {   // one Philox block per two weights
        const U4 x = rng4(seed, ST_KFV, slot0 + (uint64_t)i, level, (uint32_t)(k >> 1));
        const double m0 = __dmul_rn(u01d(x.y), scale), b0 = (double)basenet.v[k];
        cur[k] = (float)(u01d(x.x) < 0.5 ? __dadd_rn(b0, m0) : __dsub_rn(b0, m0));
        if (k + 1 < S::NW) {
            const double m1 = __dmul_rn(u01d(x.w), scale), b1 = (double)basenet.v[k + 1];
            cur[k + 1] = (float)(u01d(x.z) < 0.5 ? __dadd_rn(b1, m1) : __dsub_rn(b1, m1));
        }
    }
    // trial loop :69-82; after each self_attack the fixpoint test's application is the next step's result
    // (the max |w_k| of a state is computed once, when the state is produced, and serves the fixpoint test of this step and
    // the zero / divergence tests of the next)
    // The two buffers swap roles every step ("cur = nxt" costs no moves): `state` is the net after the self_attack,
    // `ahead` receives the next application.
    bool still = true;
    const float one = opaque_one();
    net_apply<S, ANYZERO>(cur, cur, nxt, one);
    float m_nxt = max_abs_nan<S::NW>(nxt);
    auto trial_step = [&](const float *state, float *ahead) -> bool {
        if (m_nxt <= eps.le || diverged_from_max(m_nxt)) return false;
        net_apply<S, ANYZERO>(state, state, ahead, one);
        m_nxt = max_abs_nan<S::NW>(ahead);
        if (!diverged_from_max(m_nxt) && !(max_abs_diff<S::NW>(ahead, state) >= eps.ge)) {
            if (still) ++taf;
            else still = true;
        } else {
            still = false;
        }
        ++tts;
        return true;
    };
    bool state_in_nxt = false;
    for (int st = 0; st < max_steps; st += 2) {
        state_in_nxt = true;
        if (!trial_step(nxt, cur)) break;
        if (st + 1 >= max_steps) break;
        state_in_nxt = false;
        if (!trial_step(cur, nxt)) break;
    }
    if (ys) ys[i] = tts;
    if (zs) zs[i] = taf;
    if (w_out) {
#pragma unroll
        for (int k = 0; k < S::NW; ++k) cur[k] = state_in_nxt ? nxt[k] : cur[k];
        record_store<S::NW>(w_out, ld, i, cur);
    }
    if (hist) {  // a few hot bins: count in shared memory, one global atomic per non-empty bin and CTA
        atomicAdd(&hist_s[tts], 1u);
        atomicAdd(&hist_s[(max_steps + 1) + taf], 1u);
    }
    }
    if (hist) {
        __syncthreads();
        for (int b = threadIdx.x; b < nbins; b += BS)
            if (hist_s[b]) atomicAdd(&hist[b], (unsigned long long)hist_s[b]);
    }
}

// ================================================================================================ K8 soup
struct SoupP {
    double attacking_rate, learn_from_rate;
    int train, severity, remove_divergent, remove_zero, perm_mode;
    float lr;
    Eps eps;
    int path;  // 0 auto, 1 fused single kernel, 2 multi-pass
    int dying_cap;          // trajectory side channel (include/srnn_types.h): pre-respawn records of the particles that die
    float *dying_w;
    int64_t *dying_idx;
    int *dying_count;
};
// a particle is about to be respawned: keep what save_state would have logged for it (code/soup.py:77-87)
template <int NW>
__device__ __forceinline__ void log_dying(const SoupP &p, int64_t global_index, const float *rec) {
    if (!p.dying_count) return;
    const int pos = atomicAdd(p.dying_count, 1);
    if (pos < p.dying_cap) {
        p.dying_idx[pos] = global_index;
#pragma unroll
        for (int k = 0; k < NW; ++k) p.dying_w[(int64_t)pos * NW + k] = rec[k];
    }
}

__device__ __forceinline__ void soup_draw(const SoupP &p, uint64_t seed, uint64_t slot, uint64_t gen,
                                          int64_t n_total, int64_t *att, int64_t *lrn) {
    const U4 x = rng4(seed, ST_SOUP, slot, (uint32_t)gen, (uint32_t)(gen >> 32));
    *att = (u01d(x.x) < p.attacking_rate) ? (int64_t)(((uint64_t)x.y * (uint64_t)n_total) >> 32) : -1;
    *lrn = (u01d(x.z) < p.learn_from_rate) ? (int64_t)(((uint64_t)x.w * (uint64_t)n_total) >> 32) : -1;
}

// Shared tail of both soup modes for one particle (code/soup.py:62-87): learn_from x severity, train x t,
// cull tests on the OLD particle, respawn keyed by (slot, generation).  Teacher weights (if any) are already
// in sc.tgt unless teacher_is_self (then samples are re-frozen per call, network.py:623).
template <class S>
__device__ __forceinline__ void soup_tail(const SoupP &p, float *cur, const SgdScratch &sc, bool has_teacher,
                                          bool teacher_is_self, uint64_t seed, uint64_t slot, uint64_t gen,
                                          const Limits &lim, uint8_t *action, double *loss, float *logged,
                                          int64_t logged_ld, int64_t logged_i, bool *dead, const float *spares = nullptr,
                                          int64_t spare_ld = 0, int n_spares = 0, int *spares_used = nullptr) {
    const int sev = p.severity > 0 ? p.severity : 0, tr = p.train > 0 ? p.train : 0;
    const uint64_t ep = gen * (uint64_t)(sev + tr);
    const bool ordered = S::NS > 1 && p.perm_mode == SRNN_PERM_PHILOX;
    if (has_teacher || teacher_is_self) {
        for (int e = 0; e < sev; ++e) {
            if (teacher_is_self) {
#pragma unroll
                for (int k = 0; k < S::NW; ++k) sc.tgt[k * sc.stride] = cur[k];
            }
            if (ordered) philox_perm<S::NS>(seed, slot, ep + (uint64_t)e, sc.ord, sc.stride);
            train_epoch<S>(cur, sc, ordered, p.lr);
        }
        *action = SRNN_ACT_LEARN;
    }
    for (int e = 0; e < tr; ++e) {
#pragma unroll
        for (int k = 0; k < S::NW; ++k) sc.tgt[k * sc.stride] = cur[k];
        if (ordered) philox_perm<S::NS>(seed, slot, ep + (uint64_t)(sev + e), sc.ord, sc.stride);
        *loss = train_epoch<S>(cur, sc, ordered, p.lr);
        *action = SRNN_ACT_TRAIN;
    }
    if (logged) record_store<S::NW>(logged, logged_ld, logged_i, cur);
    bool d = false;
    if (p.remove_divergent && is_diverged<S::NW>(cur)) { d = true; *action = SRNN_ACT_DIV_DEAD; }
    if (p.remove_zero && is_zero<S::NW>(cur, p.eps.le)) { d = true; *action = SRNN_ACT_ZERO_DEAD; }
    if (d) {  // respawn: next caller-supplied spare record (generator() output) if one is left, else Philox init
        log_dying<S::NW>(p, (int64_t)slot, cur);
        if (spares && spares_used && *spares_used < n_spares) {
            const float *sp = spares + (int64_t)(*spares_used) * spare_ld;
#pragma unroll
            for (int k = 0; k < S::NW; ++k) cur[k] = sp[k];
            ++*spares_used;
        } else {
            init_net<S>(seed, slot, (uint32_t)gen, lim, cur);
        }
    }
    *dead = d;
}

// exact order: ONE thread walks the particles in index order (Gauss-Seidel, code/soup.py:54-87)
template <class S>
__global__ void soup_ordered_kernel(float *w, int64_t ld, int64_t n, const int64_t *__restrict__ att,
                                    const int64_t *__restrict__ lrn, SoupP p, uint64_t seed, uint64_t gen,
                                    Limits lim, uint8_t *action, double *loss, float *logged,
                                    const float *__restrict__ spares, int64_t spare_ld, int n_spares,
                                    int *spares_used_out) {
    extern __shared__ __align__(16) unsigned char smem[];
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    SgdScratch sc = sgd_scratch<S>(smem, 1);
    int used = 0;
    for (int64_t i = 0; i < n; ++i) {
        float cur[S::NW], oth[S::NW], res[S::NW];
        uint8_t act = SRNN_ACT_NONE;
        double ls = nan("");
        bool dead = false;
        const float *rec = w + i * ld;
#pragma unroll
        for (int k = 0; k < S::NW; ++k) cur[k] = rec[k];
        const int64_t a = att ? att[i] : -1, l = lrn ? lrn[i] : -1;
        if (a >= 0) {  // particle.attack(other): other <- f_particle(other), soup.py:56-61
            const float *vr = w + a * ld;
#pragma unroll
            for (int k = 0; k < S::NW; ++k) oth[k] = vr[k];
            net_apply<S>(cur, oth, res, opaque_one());
            record_store<S::NW>(w, ld, a, res);
            if (a == i) {
#pragma unroll
                for (int k = 0; k < S::NW; ++k) cur[k] = res[k];
            }
            act = SRNN_ACT_ATTACK;
        }
        const bool has_teacher = l >= 0 && l != i, self_teacher = l == i;
        if (has_teacher) {
            const float *tr = w + l * ld;
#pragma unroll
            for (int k = 0; k < S::NW; ++k) sc.tgt[k * sc.stride] = tr[k];
        }
        soup_tail<S>(p, cur, sc, has_teacher, self_teacher, seed, (uint64_t)i, gen, lim, &act, &ls, logged, ld, i, &dead,
                     spares, spare_ld, n_spares, &used);
        record_store<S::NW>(w, ld, i, cur);
        if (action) action[i] = act;
        if (loss) loss[i] = ls;
    }
    if (spares_used_out) *spares_used_out = used;
}

// ---------------------------------------------------------------------------------------------- exact order at scale
// Soup.evolve's sequential semantics (particle i sees every write of particles < i) WITHOUT serialising the soup: a
// dependency-levelled wavefront.  Event j (particle j's turn) touches particle j, its victim a_j and its teacher l_j.
// For every particle the events touching it -- its own turn, its attackers, its learners -- form a chain ordered by
// event index; event j may run as soon as its immediate predecessor in each of its (up to three) chains has run.
// Events that become ready in the same round cannot share a particle (the earlier one would be an unfinished
// predecessor of the later one), so a round is one fully parallel kernel working in place on the arena; the number
// of rounds is the depth of the dependency DAG (a few tens for random pairings).
// One cooperative launch runs whole generations: all CTAs are co-resident and meet at grid barriers between the phases
//   P0 reset list heads / done flags   P1 draw pairings, thread events into the per-particle attacker / learner lists
//   P2 every event finds its (<= 3) immediate predecessors; events with none run at once (round 0, ~2/3 of a soup at
//      rates 0.1 / 0.1), the others go to a work list with their predecessors stored
//   rounds: only the work list is visited -- ready events run, the rest are pushed to the next list -- until it is
//      empty.  A round costs O(pending), the host never looks, the number of rounds is left in device memory.
// Everything another CTA may have written earlier in the launch is read with L1-bypassing loads.
#ifndef SRNN_EXACT_MINB
#define SRNN_EXACT_MINB 3
#endif
struct ExactBufs {
    int32_t *headA, *nextA, *headL, *nextL, *att32, *lrn32, *pred /* [3n] */, *wl[2] /* [n] each */;
    uint16_t *stamp;     // [n] 0 = not run yet, else 1 + the round the event ran in
    int32_t *cnt;        // [0..2] rotating work-list counters, [3] = rounds of the last generation, [4] = interacting particles
    unsigned *barrier;
};

__device__ __forceinline__ void grid_barrier(unsigned *counter, unsigned &target);

// the latest event before j that touches particle p (its own turn, an attack on it, a learn_from it), or -1;
// ha / hl = the already loaded list heads headA[p], headL[p]
__device__ __forceinline__ int32_t chain_pred(int32_t p, int32_t j, int32_t ha, int32_t hl, const int32_t *nextA,
                                              const int32_t *nextL) {
    int32_t best = p < j ? p : -1;
    for (int32_t e = ha; e >= 0; e = __ldcg(nextA + e))
        if (e < j && e > best) best = e;
    for (int32_t e = hl; e >= 0; e = __ldcg(nextL + e))
        if (e < j && e > best) best = e;
    return best;
}

__device__ __forceinline__ void warp_push(int32_t *counter, int32_t *list, int32_t value, bool pred);

// particle j's turn, exactly as code/soup.py:55-87 (same body as the single-thread ordered kernel), in place
template <class S>
__device__ __forceinline__ void exact_event(float *w, int64_t ld, int32_t j, int32_t a, int32_t l, const SoupP &p,
                                            const SgdScratch &sc, uint64_t seed, uint64_t gen, const Limits &lim,
                                            uint8_t *action, double *loss, uint8_t *dead_out, uint16_t *stamp,
                                            uint16_t my_stamp, const float *preloaded = nullptr) {
    float cur[S::NW], oth[S::NW], res[S::NW];
    uint8_t act = SRNN_ACT_NONE;
    double ls = nan("");
    bool dead = false;
    if (preloaded) {
#pragma unroll
        for (int k = 0; k < S::NW; ++k) cur[k] = preloaded[k];
    } else {
        record_get<S::NW, LD_CG>(w + (int64_t)j * ld, cur);
    }
    if (a >= 0) {
        record_get<S::NW, LD_CG>(w + (int64_t)a * ld, oth);
        net_apply<S>(cur, oth, res, sc.one);
        record_store<S::NW>(w, ld, a, res);
        if (a == j) {
#pragma unroll
            for (int k = 0; k < S::NW; ++k) cur[k] = res[k];
        }
        act = SRNN_ACT_ATTACK;
    }
    const bool has_teacher = l >= 0 && l != j, self_teacher = l == j;
    if (has_teacher) {
        float tch[S::NW];
        record_get<S::NW, LD_CG>(w + (int64_t)l * ld, tch);
#pragma unroll
        for (int k = 0; k < S::NW; ++k) sc.tgt[k * sc.stride] = tch[k];
    }
    const bool touched = has_teacher || self_teacher || p.train > 0;
    soup_tail<S>(p, cur, sc, has_teacher, self_teacher, seed, (uint64_t)j, gen, lim, &act, &ls, nullptr, 0, 0, &dead);
    if (touched || dead) record_store<S::NW>(w, ld, j, cur);   // an idle particle's record is left alone (a == j was stored above)
    if (action) action[j] = act;
    if (loss) loss[j] = ls;
    if (dead_out) dead_out[j] = dead ? 1 : 0;
    // No fence: a stamp only counts for readers of LATER rounds, and the grid barrier between rounds orders everything
    // written before it (an event that sees a same-round stamp ignores it).
    stamp[j] = my_stamp;
}

// One global atomic per CTA and loop iteration instead of one per warp: at 10 M particles the work-list counter is a
// single L2 address, and 300 k returning atomics on it were the bottleneck of the streaming pass.  Must be reached by
// every thread of the CTA (uniform loop bounds).
__device__ __forceinline__ void block_push(int32_t *counter, int32_t *list, int32_t value, bool pred, int *s_cnt, int *s_base) {
    if (threadIdx.x == 0) *s_cnt = 0;
    __syncthreads();
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    const int lane = threadIdx.x & 31;
    int wbase = 0;
    if (m && lane == __ffs(m) - 1) wbase = atomicAdd(s_cnt, __popc(m));
    wbase = __shfl_sync(0xffffffffu, wbase, m ? __ffs(m) - 1 : 0);
    __syncthreads();
    if (threadIdx.x == 0 && *s_cnt) *s_base = atomicAdd(counter, *s_cnt);
    __syncthreads();
    if (pred) list[*s_base + wbase + __popc(m & ((1u << lane) - 1u))] = value;
}

// P1 of an exact-order generation for particle i: its pairings (given or drawn) and its place in its victim's / teacher's lists
__device__ __forceinline__ void exact_prepare_one(const SoupP &p, uint64_t seed, uint64_t gen, int64_t n,
                                                  const int64_t *__restrict__ att_in, const int64_t *__restrict__ lrn_in,
                                                  const ExactBufs &b, int64_t i) {
    int64_t a, l;
    if (att_in) { a = att_in[i]; l = lrn_in[i]; }
    else soup_draw(p, seed, (uint64_t)i, gen, n, &a, &l);
    b.att32[i] = (int32_t)a;
    b.lrn32[i] = (int32_t)l;
    if (a >= 0) b.nextA[i] = atomicExch(&b.headA[a], (int32_t)i);
    if (l >= 0) b.nextL[i] = atomicExch(&b.headL[l], (int32_t)i);
}
// The same as an ordinary launch at full occupancy (large soups: the cooperative kernel holds 3 CTAs per SM, and this
// atomics-bound pass takes 169 us of a 10 M-particle generation in there); heads / stamps / counters are reset by memsets.
__global__ void __launch_bounds__(BS) soup_exact_prepare_kernel(SoupP p, uint64_t seed, uint64_t gen, int64_t n,
                                                                const int64_t *__restrict__ att_in,
                                                                const int64_t *__restrict__ lrn_in, ExactBufs b) {
    const int64_t i = (int64_t)blockIdx.x * BS + threadIdx.x;
    if (i < n) exact_prepare_one(p, seed, gen, n, att_in, lrn_in, b, i);
}

template <class S>
__global__ void __launch_bounds__(BS, SRNN_EXACT_MINB) soup_exact_coop_kernel(float *w, int64_t ld, int64_t n, SoupP p, uint64_t seed,
                                                             uint64_t gen0, int n_gens, Limits lim,
                                                             const int64_t *__restrict__ att_in,
                                                             const int64_t *__restrict__ lrn_in, ExactBufs b,
                                                             uint8_t *action, double *loss, uint8_t *dead_out,
                                                             int32_t *rounds_out, int prepared) {
    // prepared != 0: P0 and P1 of the (single) generation were done by memsets + soup_exact_prepare_kernel
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ __align__(16) float tile[BS * S::NW];
    __shared__ int s_cnt, s_base;
    SgdScratch sc = sgd_scratch<S>(smem, BS);
    __syncthreads();  // position codes visible
    unsigned target = 0;
    const int64_t gtid = (int64_t)blockIdx.x * BS + threadIdx.x, gsize = (int64_t)gridDim.x * BS;
    const int64_t cta0 = (int64_t)blockIdx.x * BS;     // CTA-uniform loop bounds wherever a block_push sits in the loop
#ifdef SRNN_EXACT_TIMING
    unsigned long long *tlog = reinterpret_cast<unsigned long long *>(b.barrier + 2);
    int tl = 0;
#define TSTAMP() do { if (gtid == 0 && tl < 24) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); tlog[tl++] = t_; } } while (0)
#else
#define TSTAMP() do { } while (0)
#endif
    for (int g = 0; g < n_gens; ++g) {
        const uint64_t gen = gen0 + (uint64_t)g;
#ifdef SRNN_EXACT_TIMING
        tl = 0;
#endif
        TSTAMP();
        if (!prepared) {
            // ---- P0
            for (int64_t i = gtid; i < n; i += gsize) { b.headA[i] = -1; b.headL[i] = -1; b.stamp[i] = 0; }
            if (gtid < 5 && gtid != 3) b.cnt[gtid] = 0;
            grid_barrier(b.barrier, target);
            TSTAMP();
            // ---- P1
            for (int64_t i = gtid; i < n; i += gsize) exact_prepare_one(p, seed, gen, n, att_in, lrn_in, b, i);
            grid_barrier(b.barrier, target);
        }
        TSTAMP();
        // ---- P2a: streaming pass.  A particle nobody touches and that touches nobody (about 2/3 of a soup at rates 0.1 / 0.1)
        //      has no predecessor: its turn runs here, with coalesced traffic only.  The others are compacted.
        for (int64_t base = cta0; base < n; base += gsize) {
            const int64_t j64 = base + threadIdx.x;
            bool busy = false;
            // the CTA's 256 records move as one contiguous, fully coalesced copy (per-thread record loads reach a third
            // of the HBM bandwidth here: 56-byte records straddle the 16-byte vectors differently from lane to lane)
            tile_stage_in<S::NW, BS, true>(w, ld, base, n, tile);
            __syncthreads();
            if (j64 < n) {
                const int32_t j = (int32_t)j64;
                float rec[S::NW];
                tile_get<S::NW>(tile, threadIdx.x, rec);
                const int32_t a = __ldcg(b.att32 + j), l = __ldcg(b.lrn32 + j), ha = __ldcg(b.headA + j), hl = __ldcg(b.headL + j);
                if ((a & l & ha & hl) < 0)
                    exact_event<S>(w, ld, j, -1, -1, p, sc, seed, gen, lim, action, loss, dead_out, b.stamp, (uint16_t)1, rec);
                else busy = true;
            }
            block_push(&b.cnt[4], b.wl[1], (int32_t)j64, busy, &s_cnt, &s_base);
        }
        grid_barrier(b.barrier, target);
        TSTAMP();
        // ---- P2b: dense pass over the interacting particles: find the (<= 3) immediate predecessors of each event; events
        //      without any run now (round 0), the others are listed with their predecessors
        {
            const int32_t mx = __ldcg(b.cnt + 4);
            for (int64_t base = cta0; base < mx; base += gsize) {
                const int64_t idx = base + threadIdx.x;
                bool wait = false;
                int32_t j = -1;
                if (idx < mx) {
                    j = __ldcg(b.wl[1] + idx);
                    // the first-level loads are independent: issue them together
                    const int32_t a = __ldcg(b.att32 + j), l = __ldcg(b.lrn32 + j), ha = __ldcg(b.headA + j), hl = __ldcg(b.headL + j);
                    const bool va = a >= 0 && a != j, vl = l >= 0 && l != j && l != a;
                    int32_t haa = -1, hla = -1, hal = -1, hll = -1;
                    if (va) { haa = __ldcg(b.headA + a); hla = __ldcg(b.headL + a); }
                    if (vl) { hal = __ldcg(b.headA + l); hll = __ldcg(b.headL + l); }
                    const int32_t p0 = chain_pred(j, j, ha, hl, b.nextA, b.nextL);
                    const int32_t p1 = va ? chain_pred(a, j, haa, hla, b.nextA, b.nextL) : -1;
                    const int32_t p2 = vl ? chain_pred(l, j, hal, hll, b.nextA, b.nextL) : -1;
                    if ((p0 & p1 & p2) < 0) {
                        exact_event<S>(w, ld, j, a, l, p, sc, seed, gen, lim, action, loss, dead_out, b.stamp, (uint16_t)1);
                    } else {
                        b.pred[3 * (int64_t)j] = p0; b.pred[3 * (int64_t)j + 1] = p1; b.pred[3 * (int64_t)j + 2] = p2;
                        wait = true;
                    }
                }
                block_push(&b.cnt[0], b.wl[0], j, wait, &s_cnt, &s_base);
            }
        }
        grid_barrier(b.barrier, target);
        TSTAMP();
        // ---- rounds over the pending events only
        int rounds = 1;
        for (int r = 0;; ++r, ++rounds) {
            const int32_t m = __ldcg(b.cnt + r % 3);
            if (m == 0) break;
            if (gtid == 0) b.cnt[(r + 2) % 3] = 0;         // read in round r-1, pushed to in round r+1
            const int32_t *src = b.wl[r & 1];
            int32_t *dst = b.wl[(r + 1) & 1];
            for (int64_t base = cta0; base < m; base += gsize) {
                const int64_t idx = base + threadIdx.x;
                bool wait = false;
                int32_t j = -1;
                if (idx < m) {
                    j = __ldcg(src + idx);
                    const int32_t p0 = __ldcg(b.pred + 3 * (int64_t)j), p1 = __ldcg(b.pred + 3 * (int64_t)j + 1),
                                  p2 = __ldcg(b.pred + 3 * (int64_t)j + 2);
                    const int32_t a = __ldcg(b.att32 + j), l = __ldcg(b.lrn32 + j);
                    // a predecessor counts once it ran in an EARLIER round: stamp in 1 .. r + 1 (this round writes r + 2)
                    const unsigned s0 = p0 < 0 ? 1u : __ldcg(b.stamp + p0), s1 = p1 < 0 ? 1u : __ldcg(b.stamp + p1),
                                   s2 = p2 < 0 ? 1u : __ldcg(b.stamp + p2);
                    const unsigned lim_stamp = (unsigned)r + 1u;
                    if (s0 - 1u < lim_stamp && s1 - 1u < lim_stamp && s2 - 1u < lim_stamp) {
                        exact_event<S>(w, ld, j, a, l, p, sc, seed, gen, lim, action, loss, dead_out, b.stamp,
                                       (uint16_t)(r + 2));
                    } else {
                        wait = true;
                    }
                }
                block_push(&b.cnt[(r + 1) % 3], dst, j, wait, &s_cnt, &s_base);
            }
            grid_barrier(b.barrier, target);
            TSTAMP();
            if (rounds >= 65000) break;                   // stamps are 16 bits: deeper chains are refused by the host check below
        }
        if (gtid == 0) {
            b.cnt[3] = rounds;
            if (rounds_out) rounds_out[g] = rounds;
        }
    }
}

// snapshot-synchronous, phase A: every owner evaluates the same pairing function for ALL global particles and
// threads the attackers of its local victims into per-victim lists (head[n] | next[n_total]).
__global__ void __launch_bounds__(BS) soup_lists_kernel(SoupP p, uint64_t seed, uint64_t gen, int64_t n_total,
                                                        int64_t slot0, int64_t n, int32_t *head, int32_t *next) {
    const int64_t a = (int64_t)blockIdx.x * BS + threadIdx.x;
    if (a >= n_total) return;
    int64_t at, lr;
    soup_draw(p, seed, (uint64_t)a, gen, n_total, &at, &lr);
    if (at >= slot0 && at < slot0 + n) next[a] = atomicExch(&head[at - slot0], (int32_t)a);
}

// Where the start-of-generation records live: shard r holds global particles [r*len, (r+1)*len).  One shard =
// a plain (gathered or unsharded) arena; several = the ranks' own arenas mapped into this process (NVLink peer
// memory), so partner records are fetched straight from their owner GPU and no all-gather is needed.
struct ShardTable {
    const float *base[SRNN_MAX_SHARDS];
    int64_t len;
    int32_t n;
};
__device__ __forceinline__ const float *shard_record(const ShardTable &t, int64_t ld, int64_t g) {
    const uint32_t r = t.n == 1 ? 0u : (uint32_t)g / (uint32_t)t.len;  // n_total < 2^31
    return t.base[r] + (g - (int64_t)r * t.len) * ld;
}

template <class S, bool CG = false>
__device__ __forceinline__ void attack_chain(const ShardTable &snap, int64_t snap_ld, const int32_t *head,
                                             const int32_t *next, int64_t v, float *cur);

// phase B.  The CTA keeps its 256 records in shared memory and runs the generation as a sequence of phases; the
// rare events of each phase (a particle that is attacked: ~10 %, one that learns: ~10 %, one that dies) are
// COMPACTED into a queue and handed to the first threads of the CTA, so whole warps skip a phase instead of every
// warp executing every path for one or two of its lanes.  Results do not depend on who processes a particle.
#ifndef SRNN_SOUP_MINB
#define SRNN_SOUP_MINB 7   // x 128 threads x 2 particles = 1792 particles per SM: WW(2,2) fits 7 CTAs of 31 KB with the
#endif                     // partner-prefetch areas at TILE/6 records (SRNN_SOUP_QDIV) and 72 registers (16 B of spills).
                           // Same-box A/B against 6 CTAs with TILE/4: 10 M particles 0.558 -> 0.529 ms (train 0),
                           // 0.957 -> 0.946 (train 1), 1 M 0.063 -> 0.059 (profiles/r01_soup_async_ab.txt)
// Each thread OWNS PPT particles of the tile (particle tid + r*SBS) for the uniform work -- staging, pairing draw,
// training, cull test -- so the per-tile barriers are paid once per 256 particles and the compacted queues fill a
// whole warp (about 25 events per phase instead of 12).  77 registers, no spills, 6 CTAs = 1536 particles per SM.
// Same-box A/B against one particle per thread at 10 CTAs/SM (48 registers, 40 B of spills), 10 M particles,
// ms per generation: attack .1 + learn .1: 0.664 -> 0.677 (train 0), 1.006 -> 1.012 (train 1); attack .3:
// 0.700 -> 0.661; 1 M particles: 0.077 -> 0.074 / 0.111 -> 0.108.  A wash at the default rates, better where the
// event rates are higher; three particles per thread are slower (profiles/r01_soup_ppt_ab.txt).
#ifndef SRNN_SOUP_PPT
#define SRNN_SOUP_PPT 2
#endif
#ifndef SRNN_SOUP_QDIV
#define SRNN_SOUP_QDIV 6   // prefetch areas hold TILE/6 = 40 records per phase (mean 24, sd 5); the rest take the slow path
#endif
// Small soups do not fill the SMs, so residency buys nothing and the 72-register cap only costs (100k particles,
// train 1: 19 vs 20 us per generation): below SOUP_DENSE_MIN particles per launch the kernels are the 6-CTA build
// (80 registers, prefetch areas TILE/4), from there on the 7-CTA build.
constexpr int64_t SOUP_DENSE_MIN = 1 << 19;
constexpr int64_t SOUP_ONE_WAVE_MAX = 148 * 6 * 128;   // particles that fit ONE wave of one-particle-per-thread CTAs (6 per SM)
// launches of at least this many particles hand learn_from to the dense second kernel; below, the extra launches cost more
// than the tile kernel saves (mixed 10 M soup, generations/s without / with: 8 GPUs, 417 k particles per launch: 5777 / 4897;
// 4 GPUs, 833 k: 3263 / 3206; 2 GPUs, 1.67 M: 2020 / 2194; 1 GPU, 3.3 M: 1363 / 1737)
constexpr int64_t SOUP_DEFER_MIN = 1 << 20;
constexpr int SOUP_MINB_SMALL = 6;
__host__ __device__ constexpr int soup_qdiv(int mb) { return mb >= 7 ? SRNN_SOUP_QDIV : 4; }
#ifndef SRNN_SOUP_ASYNC
#define SRNN_SOUP_ASYNC 1   // cp.async staging of the tile and the partner records (0: through registers)
#endif
// shared memory of one CTA: tile + two partner-prefetch areas (TILE/4 records each) + queues (static) and the
// per-thread SGD scratch (dynamic); shapes whose records are too large for 48 KB keep one particle per thread
__host__ __device__ constexpr int soup_static_smem(int nw, int ppt) { return nw * 4 * SOUP_BS * ppt * 3 / 2 + 10 * SOUP_BS * ppt + 64; }   // upper bound (QDIV >= 4)
template <class S>
__host__ __device__ constexpr int soup_ppt() {
    return soup_static_smem(S::NW, SRNN_SOUP_PPT) + (int)sgd_smem_bytes<S>(SOUP_BS) <= 48 * 1024 ? SRNN_SOUP_PPT : 1;
}
template <class S>
__host__ __device__ constexpr int soup_tile() { return SOUP_BS * soup_ppt<S>(); }   // particles per CTA

// DEFER: the learn_from epochs are NOT run here.  The tile only lists its learners (global list, one atomic per CTA) and
// leaves their cull test to soup_deferred_learners_kernel, a dense kernel with its own register budget.  For the recurrent
// net, whose BPTT needs ~160 registers and sat behind this kernel's 72-register cap as 384 bytes of spills on a single
// warp per tile, that takes a third off the generation (profiles/r02_soup_by_kind.txt).
template <class S, int SBS, bool CG, int QDIV = SRNN_SOUP_QDIV, bool DEFER = false, int PPT_ = soup_ppt<S>()>
__device__ __forceinline__ void soup_tile_generation(int64_t tile_idx, const ShardTable &snap, const float *own,
                                                     int64_t snap_ld, int64_t n_total, float *w_out, int64_t out_ld,
                                                     int64_t slot0, int64_t n, const SoupP &p, uint64_t seed, uint64_t gen,
                                                     const Limits &lim, const int32_t *head, const int32_t *next,
                                                     uint8_t *action, double *loss, uint8_t *dead_out,
                                                     int32_t *defer_list = nullptr, int32_t *defer_count = nullptr) {
    static_assert(!(DEFER && CG), "deferred learners: not in the persistent kernel");
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int NW = S::NW, PPT = PPT_, TILE = SBS * PPT;
    constexpr int QCAP = TILE / QDIV / 4 * 4;   // partner records prefetched per phase and tile (mean TILE/10); beyond: slow path
    __shared__ __align__(16) float tile[TILE * NW];
    __shared__ __align__(16) float att_s[QCAP * NW];   // the first attacker's record of each queued victim
    __shared__ __align__(16) float tch_s[QCAP * NW];   // the teacher's record of each queued learner
    __shared__ int32_t lrn_s[TILE];
    __shared__ uint16_t queue[TILE], queue_l[TILE];
    __shared__ int32_t nxt_s[QCAP];     // second attacker of each queued victim (-1: the usual single attacker)
    __shared__ int q_len, ql_len, dl_base;
    const int tid = threadIdx.x;
    const int64_t base = tile_idx * TILE;
    const int sev = p.severity > 0 ? p.severity : 0, tr = p.train > 0 ? p.train : 0;
    const uint64_t ep = gen * (uint64_t)(sev + tr);
    const bool ordered = S::NS > 1 && p.perm_mode == SRNN_PERM_PHILOX;
    SgdScratch sc = sgd_scratch<S>(smem, SBS);
    // ---- everything that only needs the particle index is issued up front so the DRAM latencies overlap: the list
    //      heads, the tile itself (asynchronous copies: the thread goes on to its pairing draws while the records are in
    //      flight) and -- known as soon as the draw is -- the teacher's record of every learner
    const bool prefetch = tr == 0;
    constexpr bool ASYNC = !CG && SRNN_SOUP_ASYNC;   // the persistent kernel must bypass L1: it keeps the register path
    if (tid == 0) { q_len = 0; ql_len = 0; }
    if (ASYNC) __syncthreads();   // nothing is in flight yet: this barrier is free, and the queues can fill below
    int32_t my_head[PPT];
    int64_t at[PPT], lr[PPT];
#pragma unroll
    for (int r = 0; r < PPT; ++r) {
        const int64_t i = base + tid + r * SBS;
        my_head[r] = i < n ? (CG ? __ldcg(head + i) : __ldg(head + i)) : -1;
    }
    if (ASYNC && snap_ld == NW) {
        flat_copy_async<SBS>(tile, own + base * NW, (int)max((int64_t)0, min((int64_t)TILE, n - base)) * NW);
    } else {
        tile_stage_in<NW, SBS, CG, TILE>(own, snap_ld, base, n, tile);
    }
#pragma unroll
    for (int r = 0; r < PPT; ++r) {
        const int v = tid + r * SBS;
        const int64_t i = base + v;
        at[r] = -1; lr[r] = -1;
        if (i < n) soup_draw(p, seed, (uint64_t)(slot0 + i), gen, n_total, &at[r], &lr[r]);
        lrn_s[v] = (int32_t)lr[r];
        if (ASYNC && lr[r] >= 0 && sev > 0) {
            const int ql = atomicAdd(&ql_len, 1);
            queue_l[ql] = (uint16_t)v;
            if (!DEFER && prefetch && ql < QCAP) record_copy_async<NW>(tch_s + ql * NW, shard_record(snap, snap_ld, lr[r]));
        }
    }
    if (ASYNC) async_copy_wait();
    __syncthreads();
    if (DEFER && tid == 0 && ql_len > 0) dl_base = atomicAdd(defer_count, ql_len);   // visible after the next barrier
    // ---- owners enqueue their events and, in the memory-latency-bound regime (no self-training in the generation),
    //      PREFETCH the partner records into shared memory: one DRAM round trip for attackers and teachers together
    //      instead of one per serial phase (-9 % at 10 M particles; with training the extra instructions cost more
    //      than the latency they hide, so there the partners are fetched inside the phases)
#pragma unroll
    for (int r = 0; r < PPT; ++r) {
        const int v = tid + r * SBS;
        if (my_head[r] >= 0) {
            const int qa = atomicAdd(&q_len, 1);
            queue[qa] = (uint16_t)v;
            if (prefetch && qa < QCAP) {
                const float *rec = shard_record(snap, snap_ld, my_head[r]);
                if (ASYNC) {   // nothing here waits for memory
                    async_copy4(reinterpret_cast<float *>(nxt_s + qa), reinterpret_cast<const float *>(next + my_head[r]));
                    record_copy_async<NW>(att_s + qa * NW, rec);
                } else {
                    nxt_s[qa] = CG ? __ldcg(next + my_head[r]) : __ldg(next + my_head[r]);
                    float tmp[NW];
                    record_get_uniform<NW, CG ? LD_CG : LD_RO>(rec, tmp);
                    tile_put<NW>(att_s, qa, tmp);
                }
            }
        }
        if (!ASYNC && lr[r] >= 0 && sev > 0) {
            const int ql = atomicAdd(&ql_len, 1);
            queue_l[ql] = (uint16_t)v;
            if (prefetch && ql < QCAP) {
                float tmp[NW];
                record_get_uniform<NW, CG ? LD_CG : LD_RO>(shard_record(snap, snap_ld, lr[r]), tmp);
                tile_put<NW>(tch_s, ql, tmp);
            }
        }
    }
    if (ASYNC) async_copy_wait();
    __syncthreads();
    // ---- phase 1: particles hit by attackers (code/soup.py:56-61 seen from the victim)
    for (int q = tid; q < q_len; q += SBS) {
        const int v = queue[q];
        float cur[NW];
        tile_get<NW>(tile, v, cur);
        if (prefetch && q < QCAP && nxt_s[q] < 0) {   // the usual case: exactly one attacker, already in shared memory
            float op[NW], res[NW];
            tile_get<NW>(att_s, q, op);
            net_apply<S>(op, cur, res, sc.one);
            tile_put<NW>(tile, v, res);
        } else {
            attack_chain<S, CG>(snap, snap_ld, head, next, base + v, cur);
            tile_put<NW>(tile, v, cur);
        }
    }
    __syncthreads();
    // ---- phase 2: learn_from x severity (code/soup.py:62-68), teacher read from the snapshot.  (Running the learners
    //      that are not attacked BEFORE phase 1, under the attackers' gathers, was measured slower: 0.585 -> 0.598 ms.)
    if (DEFER) {
        for (int q = tid; q < ql_len; q += SBS) defer_list[dl_base + q] = (int32_t)(base + queue_l[q]);
    }
    for (int q = tid; q < (DEFER ? 0 : ql_len); q += SBS) {
        const int v = queue_l[q];
        float cur[NW];
        tile_get<NW>(tile, v, cur);
        if (prefetch && q < QCAP) {
#pragma unroll
            for (int k = 0; k < NW; ++k) sc.tgt[k * sc.stride] = tch_s[q * NW + k];
        } else {
            float tmp[NW];
            record_get_uniform<NW, CG ? LD_CG : LD_RO>(shard_record(snap, snap_ld, lrn_s[v]), tmp);
#pragma unroll
            for (int k = 0; k < NW; ++k) sc.tgt[k * sc.stride] = tmp[k];
        }
        for (int e = 0; e < sev; ++e) {
            if (ordered) philox_perm<S::NS>(seed, (uint64_t)(slot0 + base + v), ep + (uint64_t)e, sc.ord, sc.stride);
            train_epoch<S>(cur, sc, ordered, p.lr);
        }
        tile_put<NW>(tile, v, cur);
    }
    __syncthreads();
    // ---- phase 3: self-training (every particle, code/soup.py:69-76) and the cull tests (:77-86)
    double ls[PPT];
    int dead[PPT];
#pragma unroll
    for (int r = 0; r < PPT; ++r) {
        const int v = tid + r * SBS;
        const int64_t i = base + v;
        ls[r] = nan("");
        dead[r] = 0;
        if (i < n) {
            float cur[NW];
            tile_get<NW>(tile, v, cur);
            if constexpr (!DEFER && PPT == 1 && S::KIND == SRNN_KIND_WW && S::NW <= 16) {
                // the one-particle-per-thread build only serves small TRAINING soups: the self-training kernel's epoch
                // (samples unrolled, order in one register, one Philox block per EPB epochs) is worth its size here --
                // 30 epochs per generation 19 % faster; inside the large-soup builds it cost 4-5 %
                // (profiles/r02_soup_train_unrolled_ab.txt)
                U4 blk;
                uint64_t blk_id = ~0ull;
                for (int e = 0; e < tr; ++e) {
#pragma unroll
                    for (int k = 0; k < NW; ++k) sc.tgt[k * sc.stride] = cur[k];
                    const bool last = e == tr - 1;   // only the last epoch's loss is reported
                    if (ordered) {
                        uint32_t rw[PermPlan<S::NS>::W];
                        perm_words<S::NS>(seed, (uint64_t)(slot0 + i), ep + (uint64_t)(sev + e), rw, blk, blk_id);
                        const uint64_t order = perm_nibbles<S::NS>(rw);
                        if (last) ls[r] = ww_epoch_unrolled<S, true, true>(cur, sc.tgt, sc.stride, sc.codes, order, p.lr, sc.one);
                        else ww_epoch_unrolled<S, true, false>(cur, sc.tgt, sc.stride, sc.codes, order, p.lr, sc.one);
                    } else {
                        ls[r] = ww_epoch_unrolled<S, false, true>(cur, sc.tgt, sc.stride, sc.codes, 0, p.lr, sc.one);
                    }
                }
            } else
            for (int e = 0; e < (DEFER ? 0 : tr); ++e) {   // (the deferring build is only launched for generations without training)
#pragma unroll
                for (int k = 0; k < NW; ++k) sc.tgt[k * sc.stride] = cur[k];
                if (ordered) philox_perm<S::NS>(seed, (uint64_t)(slot0 + i), ep + (uint64_t)(sev + e), sc.ord, sc.stride);
                ls[r] = train_epoch<S>(cur, sc, ordered, p.lr);
            }
            if (tr > 0) tile_put<NW>(tile, v, cur);
            const bool deferred = DEFER && lr[r] >= 0 && sev > 0;   // its cull test follows its learn_from, elsewhere
            if (!deferred && p.remove_divergent && is_diverged<NW>(cur)) dead[r] = SRNN_ACT_DIV_DEAD;
            if (!deferred && p.remove_zero && is_zero<NW>(cur, p.eps.le)) dead[r] = SRNN_ACT_ZERO_DEAD;
        }
    }
    if (tid == 0) q_len = 0;
    __syncthreads();
    // ---- phase 4: respawn (generate_particle): a pure function of (seed, slot, generation)
#pragma unroll
    for (int r = 0; r < PPT; ++r)
        if (dead[r]) queue[atomicAdd(&q_len, 1)] = (uint16_t)(tid + r * SBS);
    __syncthreads();
    for (int q = tid; q < q_len; q += SBS) {
        const int v = queue[q];
        float fresh[NW];
        if (p.dying_count) {
            tile_get<NW>(tile, v, fresh);
            log_dying<NW>(p, slot0 + base + v, fresh);
        }
        init_net<S>(seed, (uint64_t)(slot0 + base + v), (uint32_t)gen, lim, fresh);
        tile_put<NW>(tile, v, fresh);
    }
#pragma unroll
    for (int r = 0; r < PPT; ++r) {
        const int64_t i = base + tid + r * SBS;
        if (i < n) {
            uint8_t act = SRNN_ACT_NONE;
            if (at[r] >= 0) act = SRNN_ACT_ATTACK;
            if (lr[r] >= 0) act = SRNN_ACT_LEARN;
            if (tr > 0) act = SRNN_ACT_TRAIN;
            if (dead[r]) act = (uint8_t)dead[r];
            if (action) action[i] = act;
            if (loss) loss[i] = ls[r];
            if (dead_out) dead_out[i] = dead[r] ? 1 : 0;
        }
    }
    __syncthreads();
    // ---- write the new generation
    tile_stage_out<NW, SBS, TILE>(w_out, out_ld, base, n, tile);
}

template <class S, int SBS, int MB, bool DEFER = false>
__global__ void __launch_bounds__(SBS, MB) soup_sync_kernel(ShardTable snap, const float *__restrict__ own, int64_t snap_ld,
                                                       int64_t n_total, float *__restrict__ w_out, int64_t out_ld,
                                                       int64_t slot0, int64_t n, SoupP p, uint64_t seed, uint64_t gen,
                                                       Limits lim, const int32_t *__restrict__ head,
                                                       const int32_t *__restrict__ next, uint8_t *action,
                                                       double *loss, uint8_t *dead_out, int32_t *defer_list,
                                                       int32_t *defer_count) {
    soup_tile_generation<S, SBS, false, soup_qdiv(MB), DEFER>(blockIdx.x, snap, own, snap_ld, n_total, w_out, out_ld, slot0, n, p,
                                                              seed, gen, lim, head, next, action, loss, dead_out, defer_list,
                                                              defer_count);
}
// which kinds hand their learn_from epochs to the dense kernel below (when the generation has no self-training, whose
// epochs would have to follow the learn_from of the same particle)
// Measured (generation of 3.3 M / 10 M particles, attack 0.1 + learn_from 0.1, ms): RNN 0.463 -> 0.303, WW 0.192 -> 0.175 /
// 0.535 -> 0.476; AGG 0.204 -> 0.211 (its SGD is one 4-value sample: cheaper inside the tile than a second pass)
template <class S> constexpr bool soup_defers_learning() { return S::KIND != SRNN_KIND_AGG; }

// The deferred learners of a generation, one thread each, densely: the record the tile kernel left in w_out (after the
// attacks on it), the teacher's start-of-generation record, learn_from x severity, then the cull tests and the respawn
// that the tile kernel skipped for them (code/soup.py:62-68, 77-86).
template <class S>
__global__ void __launch_bounds__(BS) soup_deferred_learners_kernel(ShardTable snap, int64_t snap_ld, int64_t n_total,
                                                                    float *w_out, int64_t out_ld, int64_t slot0, SoupP p,
                                                                    uint64_t seed, uint64_t gen, Limits lim,
                                                                    const int32_t *__restrict__ learners,
                                                                    const int32_t *__restrict__ count, uint8_t *action,
                                                                    uint8_t *dead_out) {
    extern __shared__ __align__(16) unsigned char smem[];
    SgdScratch sc = sgd_scratch<S>(smem, BS);
    __syncthreads();  // position codes visible
    const int m = *count;
    const int sev = p.severity > 0 ? p.severity : 0;
    const uint64_t ep = gen * (uint64_t)sev;               // train == 0 in this mode
    const bool ordered = S::NS > 1 && p.perm_mode == SRNN_PERM_PHILOX;
    for (int64_t t = (int64_t)blockIdx.x * BS + threadIdx.x; t < m; t += (int64_t)gridDim.x * BS) {
        const int64_t l = learners[t];
        float cur[S::NW];
        {
            const float *rec = w_out + l * out_ld;
#pragma unroll
            for (int k = 0; k < S::NW; ++k) cur[k] = rec[k];
        }
        int64_t at, lr;
        soup_draw(p, seed, (uint64_t)(slot0 + l), gen, n_total, &at, &lr);
        {
            float tch[S::NW];
            record_get_uniform<S::NW, LD_RO>(shard_record(snap, snap_ld, lr), tch);
#pragma unroll
            for (int k = 0; k < S::NW; ++k) sc.tgt[k * sc.stride] = tch[k];
        }
        for (int e = 0; e < sev; ++e) {
            if (ordered) philox_perm<S::NS>(seed, (uint64_t)(slot0 + l), ep + (uint64_t)e, sc.ord, sc.stride);
            train_epoch<S>(cur, sc, ordered, p.lr);
        }
        int dead = 0;
        if (p.remove_divergent && is_diverged<S::NW>(cur)) dead = SRNN_ACT_DIV_DEAD;
        if (p.remove_zero && is_zero<S::NW>(cur, p.eps.le)) dead = SRNN_ACT_ZERO_DEAD;
        if (dead) {
            log_dying<S::NW>(p, slot0 + l, cur);
            init_net<S>(seed, (uint64_t)(slot0 + l), (uint32_t)gen, lim, cur);
        }
        record_store<S::NW>(w_out, out_ld, l, cur);
        if (action) action[l] = dead ? (uint8_t)dead : (uint8_t)SRNN_ACT_LEARN;
        if (dead_out) dead_out[l] = dead ? 1 : 0;
    }
}

// Chained variant for srnn_soup_sync_run (unsharded, many generations): ONE launch per generation.  After its tile
// generation every thread resets its own list head (consumed) and threads next generation's attack into the OTHER
// pair of list arrays, so the memset and the separate lists kernel of the per-generation path disappear.
// PPT = particles per thread.  Soups that do not fill the GPU (one wave of CTAs) are bound by the latency of a tile, and a
// thread that trains two particles one after the other doubles the longest phase: they run one particle per thread.
template <class S, int SBS, int MB, int PPT = soup_ppt<S>()>
__global__ void __launch_bounds__(SBS, MB) soup_chain_kernel(const float *__restrict__ src, int64_t ld,
                                                                        float *__restrict__ dst, int64_t n, SoupP p,
                                                                        uint64_t seed, uint64_t gen, Limits lim,
                                                                        int32_t *head_cur, const int32_t *next_cur,
                                                                        int32_t *head_nxt, int32_t *next_nxt,
                                                                        uint8_t *action, uint8_t *dead_out, int make_next) {
    ShardTable tab;
    tab.base[0] = src;
    tab.len = n;
    tab.n = 1;
    soup_tile_generation<S, SBS, false, soup_qdiv(MB), false, PPT>(blockIdx.x, tab, src, ld, n, dst, ld, 0, n, p, seed, gen, lim,
                                                                   head_cur, next_cur, action, nullptr, dead_out);
#pragma unroll
    for (int r = 0; r < PPT; ++r) {
        const int64_t i = (int64_t)blockIdx.x * (SBS * PPT) + threadIdx.x + r * SBS;
        if (i < n) {
            head_cur[i] = -1;   // only this CTA reads its heads, and it is done with them
            if (make_next && p.attacking_rate > 0.0) {
                int64_t at, lr;
                soup_draw(p, seed, (uint64_t)i, gen + 1, n, &at, &lr);
                if (at >= 0) next_nxt[i] = atomicExch(&head_nxt[at], (int32_t)i);
            }
        }
    }
}

// ---- sharded soups: the attack lists are built by the ATTACKER's owner.  head / next of every rank live in peer-mapped
// memory (head[r]: one word per particle of rank r; next[r]: one word per GLOBAL particle); an attacker pushes itself
// onto its victim's list with one system-scope exchange on the victim owner's head word and stores the previous head
// into the owner's next[] at its own global index.  Generation g's kernel pushes generation g+1's events (other
// parity of the list arrays), so a generation is ONE launch per rank whose cost is proportional to the rank's own
// particles: no pass over all n_total particles, no memset.
struct PeerListTable {
    int32_t *head[SRNN_MAX_SHARDS];
    int32_t *next[SRNN_MAX_SHARDS];
};
// flag[r] = rank r's int32[n + 1] in peer-mapped memory: word s = last generation rank s completed, word n = scratch
struct PeerFlagTable {
    int32_t *flag[SRNN_MAX_SHARDS];
    int32_t n;       // 0: no fused signalling
};
__device__ __forceinline__ void push_attack(const SoupP &p, uint64_t seed, int64_t g, uint64_t gen, int64_t n_total,
                                            int64_t shard_len, const PeerListTable &t) {
    int64_t at, lr;
    soup_draw(p, seed, (uint64_t)g, gen, n_total, &at, &lr);
    if (at >= 0) {
        const uint32_t r = (uint32_t)at / (uint32_t)shard_len;        // n_total < 2^31
        const int64_t v = at - (int64_t)r * shard_len;
        const int32_t old = atomicExch_system(t.head[r] + v, (int32_t)g);
        t.next[r][g] = old;
    }
}
template <class S, int SBS, int MB, bool DEFER = false>
__global__ void __launch_bounds__(SBS, MB) soup_chain_peers_kernel(ShardTable snap, const float *__restrict__ own,
                                                                   int64_t ld, int64_t n_total, float *__restrict__ w_out,
                                                                   int64_t out_ld, int64_t slot0, int64_t n, SoupP p,
                                                                   uint64_t seed, uint64_t gen, Limits lim, int32_t *head_cur,
                                                                   const int32_t *next_cur, PeerListTable nxt, int make_next,
                                                                   uint8_t *action, double *loss, uint8_t *dead_out,
                                                                   int32_t *defer_list, int32_t *defer_count) {
    // Generation g+1's pushes go FIRST: the exchange on the victim owner's head word is a round trip over NVLink, and
    // issued here it travels under the whole tile generation instead of holding the finished CTA on the SM (at the tail
    // it cost 15 % at 2 GPUs).  Only the returned previous head has to be kept until the store below.
    constexpr int PPT = soup_ppt<S>();
    int32_t old[PPT];
    int32_t *slot[PPT];
#pragma unroll
    for (int r = 0; r < PPT; ++r) {
        const int64_t i = (int64_t)blockIdx.x * soup_tile<S>() + threadIdx.x + r * SBS;
        slot[r] = nullptr;
        old[r] = -1;
        if (i < n && make_next && p.attacking_rate > 0.0) {
            int64_t at, lr;
            soup_draw(p, seed, (uint64_t)(slot0 + i), gen + 1, n_total, &at, &lr);
            if (at >= 0) {
                const uint32_t rk = (uint32_t)at / (uint32_t)snap.len;        // n_total < 2^31
                old[r] = atomicExch_system(nxt.head[rk] + (at - (int64_t)rk * snap.len), (int32_t)(slot0 + i));
                slot[r] = nxt.next[rk] + (slot0 + i);
            }
        }
    }
    soup_tile_generation<S, SBS, false, soup_qdiv(MB), DEFER>(blockIdx.x, snap, own, ld, n_total, w_out, out_ld, slot0, n, p, seed,
                                                              gen, lim, head_cur, next_cur, action, loss, dead_out, defer_list,
                                                              defer_count);
#pragma unroll
    for (int r = 0; r < PPT; ++r) {
        const int64_t i = (int64_t)blockIdx.x * soup_tile<S>() + threadIdx.x + r * SBS;
        if (i < n) head_cur[i] = -1;   // only this CTA reads its heads, and it is done with them; pushes for g+2 come after a barrier
        if (slot[r]) *slot[r] = old[r];
    }
}
// ---- cross-GPU ordering without a host-side barrier: after the generation kernel a one-warp kernel stores `gen` into
// word my_shard of every rank's flag array (the kernel boundary in front of it makes the generation's records, pushes
// and head resets visible system-wide; a per-CTA __threadfence_system() inside the generation kernel was measured 30 %
// slower at 2 GPUs), and in front of the next generation a one-warp kernel waits for every rank's flag.
__global__ void soup_signal_flags_kernel(PeerFlagTable flags, int my_shard, int32_t gen) {
    const int r = threadIdx.x;
    if (r < flags.n) {
        __threadfence_system();
        *reinterpret_cast<volatile int32_t *>(flags.flag[r] + my_shard) = gen;
    }
}
// A kernel of its own rather than a spin at the top of the generation kernel: that would park a GPU-full of CTAs, and
// with several soups per GPU (the kinds of a mixed soup on their streams) parked CTAs of one kind could starve the kind
// the peers are waiting for.
__global__ void soup_wait_flags_kernel(const int32_t *my_flags, int n_shards, int32_t want) {
    const int r = threadIdx.x;
    if (r < n_shards) {
        while (*reinterpret_cast<const volatile int32_t *>(my_flags + r) < want) { }
        __threadfence_system();
    }
}
// the lists of the FIRST generation of a run: every rank pushes the attacks of its own particles
__global__ void __launch_bounds__(BS) soup_push_lists_kernel(SoupP p, uint64_t seed, uint64_t gen, int64_t n_total,
                                                             int64_t shard_len, int64_t slot0, PeerListTable t) {
    const int64_t i = (int64_t)blockIdx.x * BS + threadIdx.x;
    if (i < shard_len) push_attack(p, seed, slot0 + i, gen, n_total, shard_len, t);
}

// Persistent variant for small, launch-bound soups (one GPU, whole population): the generation loop runs ON the device.
// All CTAs are co-resident (cooperative launch) and meet at a grid barrier twice per generation:
//   [pairing draws -> attack lists]  barrier  [tile generations, reset own list heads]  barrier
// so 1000 generations of a 100k-particle soup are ONE launch instead of 3000 stream operations.  Arenas written
// earlier in the same launch are re-read with L1-bypassing loads (CG).
__device__ __forceinline__ void grid_barrier(unsigned *counter, unsigned &target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        target += gridDim.x;
        __threadfence();
        atomicAdd(counter, 1u);
        while (*reinterpret_cast<volatile unsigned *>(counter) < target) { }
        __threadfence();
    }
    __syncthreads();
}

template <class S, int SBS>
__global__ void __launch_bounds__(SBS, SRNN_SOUP_MINB) soup_persistent_kernel(float *w_a, float *w_b, int64_t ld, int64_t n,
                                                                             SoupP p, uint64_t seed, uint64_t gen0,
                                                                             int n_gens, Limits lim, int32_t *head,
                                                                             int32_t *next, uint8_t *action,
                                                                             uint8_t *dead_out, unsigned *barrier) {
    unsigned target = 0;
    const int64_t gtid = (int64_t)blockIdx.x * SBS + threadIdx.x, gsize = (int64_t)gridDim.x * SBS;
    const int64_t n_tiles = (n + soup_tile<S>() - 1) / soup_tile<S>();
    for (int64_t i = gtid; i < n; i += gsize) head[i] = -1;
    grid_barrier(barrier, target);
    for (int g = 0; g < n_gens; ++g) {
        const uint64_t gen = gen0 + (uint64_t)g;
        float *src = (g & 1) ? w_b : w_a, *dst = (g & 1) ? w_a : w_b;
        if (p.attacking_rate > 0.0) {
            for (int64_t a = gtid; a < n; a += gsize) {
                int64_t at, lr;
                soup_draw(p, seed, (uint64_t)a, gen, n, &at, &lr);
                if (at >= 0) next[a] = atomicExch(&head[at], (int32_t)a);
            }
        }
        grid_barrier(barrier, target);
        ShardTable tab;
        tab.base[0] = src;
        tab.len = n;
        tab.n = 1;
        const bool last = g == n_gens - 1;
        for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
            soup_tile_generation<S, SBS, true>(t, tab, src, ld, n, dst, ld, 0, n, p, seed, gen, lim, head, next,
                                               last ? action : nullptr, nullptr, last ? dead_out : nullptr);
#pragma unroll
            for (int r = 0; r < soup_ppt<S>(); ++r) {
                const int64_t i = t * soup_tile<S>() + threadIdx.x + r * SBS;
                if (i < n) head[i] = -1;   // consumed: ready for the next generation's lists
            }
            __syncthreads();           // the shared tile is reused by the next tile of this CTA
        }
        grid_barrier(barrier, target);
    }
}

// ---------------------------------------------------------------------------------------------- multi-pass path
// For large soups the generation is split into dense passes over GLOBALLY compacted work lists, so that every warp
// of every pass is full and the long-latency chains (list walk -> partner gather -> operator / SGD) of a million
// particles overlap each other instead of serialising inside a CTA:
//   events   : every global particle draws its pairing; local victims / learners are appended to work lists
//   victims  : one thread per attacked particle: apply its attackers (ascending index) -> w_out
//   learners : one thread per learning particle: learn_from x severity -> w_out
//   final    : streaming pass over all local particles: pick up touched records, train, cull, respawn -> w_out
// scratch layout (int32 words): head[n] | next[n_total] | victims[n] | learners[n] | flags (n bytes) | counters[4]
enum { FLAG_ATTACKS = 1, FLAG_LEARNS = 2 };

__device__ __forceinline__ void warp_push(int32_t *counter, int32_t *list, int32_t value, bool pred) {
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    if (!m) return;
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    int32_t start = 0;
    if (lane == leader) start = atomicAdd(counter, __popc(m));
    start = __shfl_sync(0xffffffffu, start, leader);
    if (pred) list[start + __popc(m & ((1u << lane) - 1u))] = value;
}

constexpr int EV_BS = 1024;
// One global atomic per CTA and list: the CTA counts its pushes in shared memory, reserves a range, then writes.
__global__ void __launch_bounds__(EV_BS) soup_events_kernel(SoupP p, uint64_t seed, uint64_t gen, int64_t n_total,
                                                            int64_t slot0, int64_t n, int32_t *head, int32_t *next,
                                                            int32_t *victims, int32_t *learners, uint8_t *flags,
                                                            int32_t *counters) {
    __shared__ int cnt_s[2], base_s[2];
    if (threadIdx.x < 2) cnt_s[threadIdx.x] = 0;
    __syncthreads();
    const int64_t a = (int64_t)blockIdx.x * EV_BS + threadIdx.x;
    int64_t at = -1, lr = -1;
    if (a < n_total) soup_draw(p, seed, (uint64_t)a, gen, n_total, &at, &lr);
    bool first_hit = false;
    if (at >= slot0 && at < slot0 + n) {
        const int32_t prev = atomicExch(&head[at - slot0], (int32_t)a);
        next[a] = prev;
        first_hit = prev < 0;
    }
    const bool local = a >= slot0 && a < slot0 + n;
    const bool learns = local && lr >= 0 && p.severity > 0;
    if (local) flags[a - slot0] = (uint8_t)((at >= 0 ? FLAG_ATTACKS : 0) | (lr >= 0 ? FLAG_LEARNS : 0));
    const int my_v = first_hit ? atomicAdd(&cnt_s[0], 1) : -1;
    const int my_l = learns ? atomicAdd(&cnt_s[1], 1) : -1;
    __syncthreads();
    if (threadIdx.x < 2 && cnt_s[threadIdx.x]) base_s[threadIdx.x] = atomicAdd(&counters[threadIdx.x], cnt_s[threadIdx.x]);
    __syncthreads();
    if (first_hit) victims[base_s[0] + my_v] = (int32_t)(at - slot0);
    if (learns) learners[base_s[1] + my_l] = (int32_t)(a - slot0);
}

// apply the attackers of local particle v (list head[v] -> next[...]) to `cur` in ascending attacker index
template <class S, bool CG>
__device__ __forceinline__ void attack_chain(const ShardTable &snap, int64_t snap_ld, const int32_t *head,
                                             const int32_t *next, int64_t v, float *cur) {
    int32_t last = -1;
    for (;;) {  // lists are short (Poisson(attacking_rate)): select the next-larger attacker by rescanning
        int32_t best = 0x7fffffff, larger = 0;
        for (int32_t a = CG ? __ldcg(head + v) : head[v]; a >= 0; a = CG ? __ldcg(next + a) : next[a])
            if (a > last) { ++larger; if (a < best) best = a; }
        if (larger == 0) break;
        float op[S::NW], res[S::NW];
        record_get_uniform<S::NW, CG ? LD_CG : LD_RO>(shard_record(snap, snap_ld, best), op);
        net_apply<S>(op, cur, res, opaque_one());
#pragma unroll
        for (int k = 0; k < S::NW; ++k) cur[k] = res[k];
        if (larger == 1) break;  // that was the last one: no need to walk the list again
        last = best;
    }
}

template <class S>
__global__ void __launch_bounds__(BS) soup_victims_kernel(ShardTable snap, const float *__restrict__ own, int64_t snap_ld,
                                                          float *__restrict__ w_out, int64_t out_ld,
                                                          const int32_t *__restrict__ head,
                                                          const int32_t *__restrict__ next,
                                                          const int32_t *__restrict__ victims,
                                                          const int32_t *__restrict__ counters) {
    const int64_t t = (int64_t)blockIdx.x * BS + threadIdx.x;
    if (t >= counters[0]) return;
    const int64_t v = victims[t];
    float cur[S::NW];
    record_load<S::NW>(own, snap_ld, v, cur);
    attack_chain<S>(snap, snap_ld, head, next, v, cur);
    record_store<S::NW>(w_out, out_ld, v, cur);
}

template <class S>
__global__ void __launch_bounds__(BS) soup_learners_kernel(ShardTable snap, const float *__restrict__ own, int64_t snap_ld,
                                                           int64_t n_total, float *w_out, int64_t out_ld, int64_t slot0,
                                                           SoupP p, uint64_t seed, uint64_t gen,
                                                           const int32_t *__restrict__ head,
                                                           const int32_t *__restrict__ learners,
                                                           const int32_t *__restrict__ counters) {
    extern __shared__ __align__(16) unsigned char smem[];
    SgdScratch sc = sgd_scratch<S>(smem, BS);
    __syncthreads();  // position codes visible
    const int64_t t = (int64_t)blockIdx.x * BS + threadIdx.x;
    if (t >= counters[1]) return;
    const int64_t l = learners[t];
    float cur[S::NW];
    if (head[l] >= 0) {  // attacked earlier in this generation: continue from the victims pass' result
        const float *rec = w_out + l * out_ld;
#pragma unroll
        for (int k = 0; k < S::NW; ++k) cur[k] = rec[k];
    } else {
        record_load<S::NW>(own, snap_ld, l, cur);
    }
    int64_t at, lr;
    soup_draw(p, seed, (uint64_t)(slot0 + l), gen, n_total, &at, &lr);
    const float *rec = shard_record(snap, snap_ld, lr);
#pragma unroll
    for (int k = 0; k < S::NW; ++k) sc.tgt[k * sc.stride] = __ldg(rec + k);
    const int sev = p.severity > 0 ? p.severity : 0, tr = p.train > 0 ? p.train : 0;
    const uint64_t ep = gen * (uint64_t)(sev + tr);
    const bool ordered = S::NS > 1 && p.perm_mode == SRNN_PERM_PHILOX;
    for (int e = 0; e < sev; ++e) {
        if (ordered) philox_perm<S::NS>(seed, (uint64_t)(slot0 + l), ep + (uint64_t)e, sc.ord, sc.stride);
        train_epoch<S>(cur, sc, ordered, p.lr);
    }
    record_store<S::NW>(w_out, out_ld, l, cur);
}

template <class S>
__global__ void __launch_bounds__(BS) soup_final_kernel(const float *__restrict__ own, int64_t snap_ld, float *w_out,
                                                        int64_t out_ld, int64_t slot0, int64_t n, SoupP p, uint64_t seed,
                                                        uint64_t gen, Limits lim, const int32_t *__restrict__ head,
                                                        const uint8_t *__restrict__ flags, uint8_t *action, double *loss,
                                                        uint8_t *dead_out) {
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int NW = S::NW;
    __shared__ __align__(16) float tile[BS * NW];
    __shared__ uint16_t queue[BS];
    __shared__ int q_len;
    const int tid = threadIdx.x;
    const int64_t base = (int64_t)blockIdx.x * BS, i = base + tid;
    const bool active = i < n;
    const int tr = p.train > 0 ? p.train : 0, sev = p.severity > 0 ? p.severity : 0;
    SgdScratch sc;
    if (tr > 0) sc = sgd_scratch<S>(smem, BS);
    tile_stage_in<NW, BS>(own, snap_ld, base, n, tile);
    if (tid == 0) q_len = 0;
    __syncthreads();
    float cur[NW];
    uint8_t fl = 0;
    double ls = nan("");
    int dead = 0;
    if (active) {
        fl = flags[i];
        const bool touched = head[i] >= 0 || ((fl & FLAG_LEARNS) && sev > 0);
        if (touched) {  // the victims / learners passes left this particle's record in w_out
            const float *rec = w_out + i * out_ld;
#pragma unroll
            for (int k = 0; k < NW; ++k) cur[k] = rec[k];
        } else {
            tile_get<NW>(tile, tid, cur);
        }
        const uint64_t ep = gen * (uint64_t)(sev + tr);
        const bool ordered = S::NS > 1 && p.perm_mode == SRNN_PERM_PHILOX;
        for (int e = 0; e < tr; ++e) {
#pragma unroll
            for (int k = 0; k < NW; ++k) sc.tgt[k * sc.stride] = cur[k];
            if (ordered) philox_perm<S::NS>(seed, (uint64_t)(slot0 + i), ep + (uint64_t)(sev + e), sc.ord, sc.stride);
            ls = train_epoch<S>(cur, sc, ordered, p.lr);
        }
        if (touched || tr > 0) tile_put<NW>(tile, tid, cur);
        if (p.remove_divergent && is_diverged<NW>(cur)) dead = SRNN_ACT_DIV_DEAD;
        if (p.remove_zero && is_zero<NW>(cur, p.eps.le)) dead = SRNN_ACT_ZERO_DEAD;
    }
    if (dead) queue[atomicAdd(&q_len, 1)] = (uint16_t)tid;
    __syncthreads();
    if (tid < q_len) {  // respawn, compacted within the CTA
        const int v = queue[tid];
        float fresh[NW];
        if (p.dying_count) {
            tile_get<NW>(tile, v, fresh);
            log_dying<NW>(p, slot0 + base + v, fresh);
        }
        init_net<S>(seed, (uint64_t)(slot0 + base + v), (uint32_t)gen, lim, fresh);
        tile_put<NW>(tile, v, fresh);
    }
    if (active) {
        uint8_t act = SRNN_ACT_NONE;
        if (fl & FLAG_ATTACKS) act = SRNN_ACT_ATTACK;
        if (fl & FLAG_LEARNS) act = SRNN_ACT_LEARN;
        if (tr > 0) act = SRNN_ACT_TRAIN;
        if (dead) act = (uint8_t)dead;
        if (action) action[i] = act;
        if (loss) loss[i] = ls;
        if (dead_out) dead_out[i] = dead ? 1 : 0;
    }
    __syncthreads();
    tile_stage_out<NW, BS>(w_out, out_ld, base, n, tile);
}

// ================================================================================================ uid scan
constexpr int SCAN_BS = 1024;
__global__ void __launch_bounds__(SCAN_BS) uid_count_kernel(const uint8_t *__restrict__ dead, int64_t n, int64_t *block_tmp) {
    const int64_t i = (int64_t)blockIdx.x * SCAN_BS + threadIdx.x;
    const int c = __syncthreads_count(i < n && dead[i] != 0);
    if (threadIdx.x == 0) block_tmp[blockIdx.x] = c;
}
__global__ void __launch_bounds__(SCAN_BS) uid_blockscan_kernel(int64_t *block_tmp, int64_t nb, int64_t *n_dead_out) {
    // exclusive scan of nb block counts by one CTA: each thread owns a contiguous chunk
    __shared__ int64_t part[SCAN_BS];
    const int64_t chunk = (nb + SCAN_BS - 1) / SCAN_BS, lo = (int64_t)threadIdx.x * chunk;
    const int64_t hi = lo + chunk < nb ? lo + chunk : nb;
    int64_t s = 0;
    for (int64_t b = lo; b < hi; ++b) s += block_tmp[b];
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int64_t run = 0;
        for (int t = 0; t < SCAN_BS; ++t) { const int64_t v = part[t]; part[t] = run; run += v; }
        block_tmp[nb] = run;
        if (n_dead_out) *n_dead_out = run;
    }
    __syncthreads();
    int64_t run = part[threadIdx.x];
    for (int64_t b = lo; b < hi; ++b) { const int64_t v = block_tmp[b]; block_tmp[b] = run; run += v; }
}
__global__ void __launch_bounds__(SCAN_BS) uid_assign_kernel(const uint8_t *__restrict__ dead, int64_t n, int64_t uid_base,
                                                             const int64_t *__restrict__ block_tmp, int64_t *uid) {
    __shared__ int warp_tot[SCAN_BS / 32];
    const int64_t i = (int64_t)blockIdx.x * SCAN_BS + threadIdx.x;
    const bool d = i < n && dead[i] != 0;
    const unsigned m = __ballot_sync(0xffffffffu, d);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) warp_tot[wid] = __popc(m);
    __syncthreads();
    int before = __popc(m & ((1u << lane) - 1u));
    for (int q = 0; q < wid; ++q) before += warp_tot[q];
    if (d) uid[i] = uid_base + block_tmp[blockIdx.x] + before;
}

// ================================================================================================ FP32 probe
template <int VARIANT>
__global__ void __launch_bounds__(BS) fp32_probe_kernel(float *out, int iters) {
    float acc[16];
    const float a = 1.0f + 1e-7f * (float)threadIdx.x, b = 1e-9f * (float)blockIdx.x;
#pragma unroll
    for (int q = 0; q < 16; ++q) acc[q] = (float)q;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int q = 0; q < 16; ++q) {
                if constexpr (VARIANT == 0) acc[q] = __fmaf_rn(acc[q], a, b);  // reg, reg, reg
                else if constexpr (VARIANT == 1) acc[q] = __fmaf_rn(acc[q], 0.99999f, b);  // reg, imm, reg
            }
        if constexpr (VARIANT == 2) {  // packed fma.rn.ftz.f32x2: 2 lanes-FMA per thread-instruction
            unsigned long long pa, pb;
            asm("mov.b64 %0, {%1, %1};" : "=l"(pa) : "f"(a));
            asm("mov.b64 %0, {%1, %1};" : "=l"(pb) : "f"(b));
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int q = 0; q < 16; q += 2) {
                    unsigned long long v;
                    asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(acc[q]), "f"(acc[q + 1]));
                    asm("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(v) : "l"(v), "l"(pa), "l"(pb));
                    asm("mov.b64 {%0, %1}, %2;" : "=f"(acc[q]), "=f"(acc[q + 1]) : "l"(v));
                }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < 16; ++q) s += acc[q];
    out[(int64_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ================================================================================================ host side
static Limits make_limits(const srnn_net_desc *d) {
    Limits lim;
    const int nl = d->depth + 1;
    const int io = d->kind == SRNN_KIND_WW ? 4 : (d->kind == SRNN_KIND_AGG ? d->aggregates : 1);
    for (int l = 0; l < SRNN_MAX_LAYERS; ++l) {
        const int in = l == 0 ? io : d->width;
        const int out = l == nl - 1 ? (d->kind == SRNN_KIND_WW ? 1 : io) : d->width;
        lim.v[l] = (float)sqrt(6.0 / (double)(in + out));
    }
    return lim;
}
static int make_soup(const srnn_soup_params *p, SoupP *q) {
    if (!p) return fail(SRNN_ERR_ARG, "soup params is NULL%s");
    q->attacking_rate = p->attacking_rate;
    q->learn_from_rate = p->learn_from_rate;
    q->train = p->train;
    q->severity = p->learn_from_severity;
    q->remove_divergent = p->remove_divergent;
    q->remove_zero = p->remove_zero;
    q->perm_mode = p->perm_mode;
    q->lr = p->lr;
    q->path = p->engine_path;
    q->dying_cap = p->dying_count ? p->dying_cap : 0;
    q->dying_w = p->dying_w;
    q->dying_idx = p->dying_idx;
    q->dying_count = (p->dying_count && p->dying_w && p->dying_idx && p->dying_cap > 0) ? p->dying_count : nullptr;
    if (p->perm_mode != SRNN_PERM_NONE && p->perm_mode != SRNN_PERM_PHILOX)
        return fail(SRNN_ERR_ARG, "soup perm_mode must be SRNN_PERM_NONE or SRNN_PERM_PHILOX%s");
    return make_eps(p->epsilon, &q->eps);
}

// the instantiated net shapes: (kind, width, depth, aggregates)
#ifdef SRNN_DEV_ONLY_WW22   // development builds: one shape compiles in a fraction of the time (never shipped)
#define SRNN_SHAPES(X) X(SRNN_KIND_WW, 2, 2, 0)
#elif defined(SRNN_DEV_ONLY_RNN22)
#define SRNN_SHAPES(X) X(SRNN_KIND_RNN, 2, 2, 0)
#else
#define SRNN_SHAPES(X)                                                                               \
    X(SRNN_KIND_WW, 2, 2, 0) X(SRNN_KIND_WW, 3, 2, 0) X(SRNN_KIND_WW, 2, 3, 0) X(SRNN_KIND_WW, 2, 1, 0) \
    X(SRNN_KIND_AGG, 2, 2, 4) X(SRNN_KIND_AGG, 2, 2, 3) X(SRNN_KIND_AGG, 3, 2, 4)                    \
    X(SRNN_KIND_RNN, 2, 2, 0) X(SRNN_KIND_RNN, 2, 1, 0)
#endif

static inline long shape_key(int kind, int width, int depth, int aggs) {
    return ((long)kind << 24) | ((long)width << 16) | ((long)depth << 8) | (long)(kind == SRNN_KIND_AGG ? aggs : 0);
}
#define SRNN_CASE(KIND, W, D, A)                    \
    case ((long)(KIND) << 24) | ((long)(W) << 16) | ((long)(D) << 8) | (long)(A): { \
        using S = Shape<KIND, W, D, A>;             \
        BODY(S)                                     \
    } break;
#define SRNN_DISPATCH(d)                                                                              \
    switch (shape_key((d)->kind, (d)->width, (d)->depth, (d)->aggregates)) {                          \
        SRNN_SHAPES(SRNN_CASE)                                                                        \
        default: return fail(SRNN_ERR_SHAPE, "this operation exists only for the net shapes instantiated in libsrnn_b200 (see SRNN_SHAPES)%s"); \
    }
// same, but shapes outside SRNN_SHAPES run GENERIC (a statement using `dyn`, the run-time shape) instead of failing
#define SRNN_DISPATCH_OR_GENERIC(d, ...)                                                              \
    switch (shape_key((d)->kind, (d)->width, (d)->depth, (d)->aggregates)) {                          \
        SRNN_SHAPES(SRNN_CASE)                                                                        \
        default: {                                                                                    \
            gen::DynShape dyn;                                                                        \
            const int grc = gen::make_dyn(d, &dyn);                                                   \
            if (grc) return fail(grc, "invalid or too large net shape%s");                            \
            __VA_ARGS__                                                                               \
        }                                                                                             \
    }
static inline unsigned ggrid(int64_t n) { return (unsigned)((n + gen::GBS - 1) / gen::GBS); }

static inline unsigned grid_for(int64_t n) { return (unsigned)((n + BS - 1) / BS); }

template <class K>
static int allow_smem(K kernel, size_t bytes) {
    if (bytes > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        if (e != cudaSuccess) return fail(SRNN_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    }
    return SRNN_OK;
}

}  // namespace srnn

// ---------------------------------------------------------------------------------------------- run-time-shape soups + KFV
// Population-scale Soup.evolve and the known-fixpoint sweep for ANY shape that fits SRNN_MAX_WEIGHTS:
// one thread per particle / trial, run-time loops over the layer table, same pinned arithmetic and the same Philox
// functions as the templated kernels -- bit-identical to the oracle, not tuned (the instantiated shapes keep the fast path).
namespace srnn {
namespace gen {
__device__ inline void log_dying_dyn(const SoupP &p, int nw, int64_t global_index, const float *rec) {
    if (!p.dying_count) return;
    const int pos = atomicAdd(p.dying_count, 1);
    if (pos < p.dying_cap) {
        p.dying_idx[pos] = global_index;
        for (int k = 0; k < nw; ++k) p.dying_w[(int64_t)pos * nw + k] = rec[k];
    }
}
// one snapshot-synchronous generation for particles [slot0, slot0 + n): the particle's whole turn in one thread
// (code/soup.py:55-87 seen from the particle: attacks on it in ascending attacker index, learn_from, train, cull, respawn)
__global__ void __launch_bounds__(GBS) soup_sync_kernel(DynShape s, ShardTable snap, int64_t snap_ld, int64_t n_total, float *w_out,
                                                       int64_t out_ld, int64_t slot0, int64_t n, SoupP p, uint64_t seed,
                                                       uint64_t gen, const int32_t *__restrict__ head,
                                                       const int32_t *__restrict__ next, uint8_t *action, double *loss,
                                                       uint8_t *dead_out) {
    const int64_t i = (int64_t)blockIdx.x * GBS + threadIdx.x;
    if (i >= n) return;
    const int nw = s.n_weights, ns = s.n_samples;
    float cur[MAXW], oth[MAXW], res[MAXW];
    {
        const float *rec = shard_record(snap, snap_ld, slot0 + i);
        for (int k = 0; k < nw; ++k) cur[k] = __ldg(rec + k);
    }
    int32_t last = -1;
    for (;;) {   // the attackers of this particle, ascending
        int32_t best = 0x7fffffff, larger = 0;
        for (int32_t a = head[i]; a >= 0; a = next[a])
            if (a > last) { ++larger; if (a < best) best = a; }
        if (larger == 0) break;
        const float *rec = shard_record(snap, snap_ld, best);
        for (int k = 0; k < nw; ++k) oth[k] = __ldg(rec + k);
        net_apply(s, oth, cur, res);
        for (int k = 0; k < nw; ++k) cur[k] = res[k];
        if (larger == 1) break;
        last = best;
    }
    int64_t at, lr;
    soup_draw(p, seed, (uint64_t)(slot0 + i), gen, n_total, &at, &lr);
    uint8_t act = at >= 0 ? SRNN_ACT_ATTACK : SRNN_ACT_NONE;
    const int sev = p.severity > 0 ? p.severity : 0, tr = p.train > 0 ? p.train : 0;
    const uint64_t ep = gen * (uint64_t)(sev + tr);
    const bool ordered = ns > 1 && p.perm_mode == SRNN_PERM_PHILOX;
    uint8_t ord[MAXW];
    double ls = nan("");
    if (lr >= 0 && sev > 0) {
        const float *rec = shard_record(snap, snap_ld, lr);
        for (int k = 0; k < nw; ++k) oth[k] = __ldg(rec + k);
        for (int e = 0; e < sev; ++e) {
            if (ordered) philox_perm(seed, (uint64_t)(slot0 + i), ep + (uint64_t)e, ns, ord);
            train_epoch(s, cur, oth, ordered ? ord : nullptr, p.lr);
        }
    }
    if (lr >= 0) act = SRNN_ACT_LEARN;
    for (int e = 0; e < tr; ++e) {
        for (int k = 0; k < nw; ++k) oth[k] = cur[k];
        if (ordered) philox_perm(seed, (uint64_t)(slot0 + i), ep + (uint64_t)(sev + e), ns, ord);
        ls = train_epoch(s, cur, oth, ordered ? ord : nullptr, p.lr);
        act = SRNN_ACT_TRAIN;
    }
    int dead = 0;
    if (p.remove_divergent && is_diverged(cur, nw)) dead = SRNN_ACT_DIV_DEAD;
    if (p.remove_zero && is_zero(cur, nw, p.eps.le)) dead = SRNN_ACT_ZERO_DEAD;
    if (dead) {
        log_dying_dyn(p, nw, slot0 + i, cur);
        init_net(s, seed, (uint64_t)(slot0 + i), (uint32_t)gen, cur);
        act = (uint8_t)dead;
    }
    store_rec(w_out, out_ld, i, nw, cur);
    if (action) action[i] = act;
    if (loss) loss[i] = ls;
    if (dead_out) dead_out[i] = dead ? 1 : 0;
}
// known-fixpoint-variation trial loop (code/setups/known-fixpoint-variation.py:37-46, 60-89), run-time shape
__global__ void __launch_bounds__(GBS) kfv_kernel(DynShape s, BaseNet basenet, int64_t n, double scale, uint32_t level, int max_steps,
                                                 Eps eps, uint64_t seed, uint64_t slot0, int *__restrict__ ys,
                                                 int *__restrict__ zs, unsigned long long *hist, float *__restrict__ w_out,
                                                 int64_t ld) {
    const int64_t i = (int64_t)blockIdx.x * GBS + threadIdx.x;
    if (i >= n) return;
    const int nw = s.n_weights;
    float cur[MAXW], nxt[MAXW];
    for (int k = 0; k < nw; k += 2) {   // one Philox block per two weights
        const U4 x = rng4(seed, ST_KFV, slot0 + (uint64_t)i, level, (uint32_t)(k >> 1));
        const double m0 = __dmul_rn(u01d(x.y), scale), b0 = (double)basenet.v[k];
        cur[k] = (float)(u01d(x.x) < 0.5 ? __dadd_rn(b0, m0) : __dsub_rn(b0, m0));
        if (k + 1 < nw) {
            const double m1 = __dmul_rn(u01d(x.w), scale), b1 = (double)basenet.v[k + 1];
            cur[k + 1] = (float)(u01d(x.z) < 0.5 ? __dadd_rn(b1, m1) : __dsub_rn(b1, m1));
        }
    }
    int tts = 0, taf = 0;
    bool still = true;
    net_apply(s, cur, cur, nxt);
    for (int st = 0; st < max_steps; ++st) {
        for (int k = 0; k < nw; ++k) cur[k] = nxt[k];
        if (is_zero(cur, nw, eps.le) || is_diverged(cur, nw)) break;
        net_apply(s, cur, cur, nxt);
        if (is_close(nxt, cur, nw, eps.ge)) {
            if (still) ++taf;
            else still = true;
        } else {
            still = false;
        }
        ++tts;
    }
    if (ys) ys[i] = tts;
    if (zs) zs[i] = taf;
    if (w_out) store_rec(w_out, ld, i, nw, cur);
    if (hist) {
        atomicAdd(&hist[tts], 1ull);
        atomicAdd(&hist[(max_steps + 1) + taf], 1ull);
    }
}
}  // namespace gen
}  // namespace srnn
using namespace srnn;

extern "C" {

int srnn_version(void) { return 100; }
const char *srnn_last_error_string(void) { return g_err; }

int srnn_set_device(int device) {
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(SRNN_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    return SRNN_OK;
}

int srnn_desc_init(srnn_net_desc *d) {
    if (!d) return fail(SRNN_ERR_ARG, "desc is NULL%s");
    if (d->width < 1 || d->width > SRNN_MAX_WIDTH || d->depth < 1 || d->depth + 1 > SRNN_MAX_LAYERS)
        return fail(SRNN_ERR_SHAPE, "width/depth out of range%s");
    int io;
    if (d->kind == SRNN_KIND_WW) io = 4;
    else if (d->kind == SRNN_KIND_AGG) io = d->aggregates;
    else if (d->kind == SRNN_KIND_RNN) io = 1;
    else return fail(SRNN_ERR_SHAPE, "unknown net kind%s");
    if (io < 1 || io > SRNN_MAX_WIDTH) return fail(SRNN_ERR_SHAPE, "aggregates out of range%s");
    int off = 0;
    for (int l = 0; l <= d->depth; ++l) {
        const int in = l == 0 ? io : d->width;
        const int out = l == d->depth ? (d->kind == SRNN_KIND_WW ? 1 : io) : d->width;
        off += in * out + (d->kind == SRNN_KIND_RNN ? out * out : 0);
    }
    if (off > SRNN_MAX_WEIGHTS) return fail(SRNN_ERR_LIMIT, "more than SRNN_MAX_WEIGHTS weights%s");
    if (d->kind == SRNN_KIND_AGG) {
        const int cs = off / io;
        if (cs < 1 || off - cs * io >= cs)
            return fail(SRNN_ERR_SHAPE, "aggregating shape whose leftovers fill another collection (reference raises)%s");
    }
    d->n_weights = off;
    return SRNN_OK;
}

int srnn_shape_supported(const srnn_net_desc *d) {
    if (!d) return 0;
#define BODY(S) return 1;
    switch (shape_key(d->kind, d->width, d->depth, d->aggregates)) {
        SRNN_SHAPES(SRNN_CASE)
        default: {
            gen::DynShape dyn;
            return gen::make_dyn(d, &dyn) == SRNN_OK ? 2 : 0;
        }
    }
#undef BODY
}

#define SRNN_CHECK_ARENA(w, ld, S)                                                              \
    if (!(w)) return fail(SRNN_ERR_ARG, "arena pointer is NULL%s");                              \
    if ((ld) < S::NW) return fail(SRNN_ERR_ARG, "record stride ld is smaller than n_weights%s");

int srnn_apply(const srnn_net_desc *d, const float *op_w, int64_t op_ld, const int32_t *op_idx, const float *in_w,
               int64_t in_ld, const int32_t *in_idx, float *out_w, int64_t out_ld, const int32_t *out_idx, int64_t n,
               void *stream) {
    if (!d) return fail(SRNN_ERR_ARG, "desc is NULL%s");
    if (n <= 0) return n == 0 ? SRNN_OK : fail(SRNN_ERR_ARG, "n < 0%s");
#define BODY(S)                                                                                              \
    SRNN_CHECK_ARENA(op_w, op_ld, S) SRNN_CHECK_ARENA(in_w, in_ld, S) SRNN_CHECK_ARENA(out_w, out_ld, S)     \
    apply_kernel<S><<<grid_for(n), BS, 0, (cudaStream_t)stream>>>(op_w, op_ld, op_idx, in_w, in_ld, in_idx, out_w, \
                                                                  out_ld, out_idx, n);
    SRNN_DISPATCH_OR_GENERIC(d, {
        if (!op_w || !in_w || !out_w) return fail(SRNN_ERR_ARG, "arena pointer is NULL%s");
        if (op_ld < dyn.n_weights || in_ld < dyn.n_weights || out_ld < dyn.n_weights)
            return fail(SRNN_ERR_ARG, "record stride ld is smaller than n_weights%s");
        gen::apply_kernel<<<ggrid(n), gen::GBS, 0, (cudaStream_t)stream>>>(dyn, op_w, op_ld, op_idx, in_w, in_ld, in_idx, out_w,
                                                                        out_ld, out_idx, n);
    })
#undef BODY
    return check_launch("srnn_apply");
}

int srnn_forward(const srnn_net_desc *d, const float *w, int64_t ld, const int32_t *idx, const float *x, float *y,
                 int64_t n, void *stream) {
    if (!d || !x || !y) return fail(SRNN_ERR_ARG, "desc, x or y is NULL%s");
    if (n <= 0) return n == 0 ? SRNN_OK : fail(SRNN_ERR_ARG, "n < 0%s");
#define BODY(S)                \
    SRNN_CHECK_ARENA(w, ld, S) \
    forward_kernel<S><<<grid_for(n), BS, 0, (cudaStream_t)stream>>>(w, ld, idx, x, y, n);
    SRNN_DISPATCH_OR_GENERIC(d, {
        if (!w || ld < dyn.n_weights) return fail(SRNN_ERR_ARG, "arena pointer is NULL or ld smaller than n_weights%s");
        if (dyn.kind == SRNN_KIND_RNN)
            gen::forward_rnn_kernel<<<ggrid(n), gen::GBS, 0, (cudaStream_t)stream>>>(dyn, w, ld, idx, x, y, n);
        else
            gen::forward_mlp_kernel<<<ggrid(n), gen::GBS, 0, (cudaStream_t)stream>>>(dyn, w, ld, idx, x, y, n);
    })
#undef BODY
    return check_launch("srnn_forward");
}

int srnn_self_apply(const srnn_net_desc *d, float *w, int64_t ld, int64_t n, int32_t n_steps, int32_t mode, double eps,
                    int32_t with_fix_sec, int32_t *steps_out, uint8_t *cls_out, int64_t *counters, float *traj,
                    void *stream) {
    if (!d) return fail(SRNN_ERR_ARG, "desc is NULL%s");
    if (n <= 0) return n == 0 ? SRNN_OK : fail(SRNN_ERR_ARG, "n < 0%s");
    if (n_steps < 0) return fail(SRNN_ERR_ARG, "n_steps < 0%s");
    if (mode != SRNN_RUN_FIXED && mode != SRNN_RUN_UNTIL) return fail(SRNN_ERR_ARG, "unknown run mode%s");
    const bool fast = mode == SRNN_RUN_FIXED && !steps_out && !cls_out && !counters && !traj;
    Eps e{0.f, 0.f};
    if (!fast) { int rc = make_eps(eps, &e); if (rc) return rc; }
#define BODY(S)                                                                                              \
    SRNN_CHECK_ARENA(w, ld, S)                                                                               \
    if (fast) self_apply_fixed_kernel<S><<<grid_for(n), BS, 0, (cudaStream_t)stream>>>(w, ld, n, n_steps);   \
    else self_apply_general_kernel<S><<<grid_for(n), BS, 0, (cudaStream_t)stream>>>(                         \
            w, ld, n, n_steps, mode, e, with_fix_sec, steps_out, cls_out, (unsigned long long *)counters, traj);
    SRNN_DISPATCH_OR_GENERIC(d, {
        if (!w || ld < dyn.n_weights) return fail(SRNN_ERR_ARG, "arena pointer is NULL or ld smaller than n_weights%s");
        if (fast) { int rc = make_eps(1e-14, &e); if (rc) return rc; }
        gen::self_apply_kernel<<<ggrid(n), gen::GBS, 0, (cudaStream_t)stream>>>(
            dyn, w, ld, n, n_steps, mode, e.ge, e.le, with_fix_sec, steps_out, cls_out, (unsigned long long *)counters, traj);
    })
#undef BODY
    return check_launch("srnn_self_apply");
}

int srnn_classify(const srnn_net_desc *d, const float *w, int64_t ld, int64_t n, double eps, int32_t with_fix_sec,
                  uint8_t *cls_out, uint8_t *flags_out, int64_t *counters, void *stream) {
    if (!d) return fail(SRNN_ERR_ARG, "desc is NULL%s");
    if (n <= 0) return n == 0 ? SRNN_OK : fail(SRNN_ERR_ARG, "n < 0%s");
    Eps e;
    int rc = make_eps(eps, &e);
    if (rc) return rc;
#define BODY(S)                                                                                              \
    SRNN_CHECK_ARENA(w, ld, S)                                                                               \
    classify_kernel<S><<<grid_for(n), BS, 0, (cudaStream_t)stream>>>(w, ld, n, e, with_fix_sec, cls_out, flags_out, \
                                                                     (unsigned long long *)counters);
    SRNN_DISPATCH_OR_GENERIC(d, {
        if (!w || ld < dyn.n_weights) return fail(SRNN_ERR_ARG, "arena pointer is NULL or ld smaller than n_weights%s");
        gen::classify_kernel<<<ggrid(n), gen::GBS, 0, (cudaStream_t)stream>>>(dyn, w, ld, n, e.ge, e.le, with_fix_sec, cls_out,
                                                                           flags_out, (unsigned long long *)counters);
    })
#undef BODY
    return check_launch("srnn_classify");
}

int srnn_train(const srnn_net_desc *d, float *w, int64_t ld, int64_t n, int32_t n_epochs, float lr, int32_t perm_mode,
               const uint8_t *perm, uint64_t seed, uint64_t slot0, uint64_t epoch0, const float *teacher_w,
               int64_t teacher_ld, const int32_t *teacher_idx, double *loss_out, void *stream) {
    if (!d) return fail(SRNN_ERR_ARG, "desc is NULL%s");
    if (n <= 0) return n == 0 ? SRNN_OK : fail(SRNN_ERR_ARG, "n < 0%s");
    if (n_epochs < 0) return fail(SRNN_ERR_ARG, "n_epochs < 0%s");
    if (perm_mode < SRNN_PERM_NONE || perm_mode > SRNN_PERM_PHILOX) return fail(SRNN_ERR_ARG, "unknown perm_mode%s");
    if (perm_mode == SRNN_PERM_EXPLICIT && !perm) return fail(SRNN_ERR_ARG, "SRNN_PERM_EXPLICIT needs perm%s");
#define BODY(S)                                                                                              \
    SRNN_CHECK_ARENA(w, ld, S)                                                                               \
    if (teacher_w && teacher_ld < S::NW) return fail(SRNN_ERR_ARG, "teacher_ld is smaller than n_weights%s"); \
    const size_t sm = sgd_smem_bytes<S>(BS);                                                                 \
    int rc = allow_smem(train_kernel<S>, sm + sizeof(float) * BS * S::NW);                    \
    if (rc) return rc;                                                                                       \
    train_kernel<S><<<grid_for(n), BS, sm, (cudaStream_t)stream>>>(w, ld, n, n_epochs, lr, perm_mode, perm, seed, slot0, \
                                                                   epoch0, teacher_w, teacher_ld, teacher_idx, loss_out);
    SRNN_DISPATCH_OR_GENERIC(d, {
        if (!w || ld < dyn.n_weights) return fail(SRNN_ERR_ARG, "arena pointer is NULL or ld smaller than n_weights%s");
        if (teacher_w && teacher_ld < dyn.n_weights) return fail(SRNN_ERR_ARG, "teacher_ld is smaller than n_weights%s");
        if (dyn.kind == SRNN_KIND_RNN && dyn.n_layers * dyn.n_weights * dyn.hist_stride > gen::MAX_HIST)
            return fail(SRNN_ERR_LIMIT, "recurrent net too large for the generic BPTT path%s");
        gen::train_kernel<<<ggrid(n), gen::GBS, 0, (cudaStream_t)stream>>>(dyn, w, ld, n, n_epochs, lr, perm_mode, perm, seed, slot0,
                                                                        epoch0, teacher_w, teacher_ld, teacher_idx, loss_out);
    })
#undef BODY
    return check_launch("srnn_train");
}

int srnn_init(const srnn_net_desc *d, float *w, int64_t ld, int64_t n, const int32_t *idx, uint64_t seed,
              uint64_t slot0, uint32_t generation, void *stream) {
    if (!d) return fail(SRNN_ERR_ARG, "desc is NULL%s");
    if (n <= 0) return n == 0 ? SRNN_OK : fail(SRNN_ERR_ARG, "n < 0%s");
    const Limits lim = make_limits(d);
#define BODY(S)                \
    SRNN_CHECK_ARENA(w, ld, S) \
    init_kernel<S><<<grid_for(n), BS, 0, (cudaStream_t)stream>>>(w, ld, n, idx, seed, slot0, generation, lim);
    SRNN_DISPATCH_OR_GENERIC(d, {
        if (!w || ld < dyn.n_weights) return fail(SRNN_ERR_ARG, "arena pointer is NULL or ld smaller than n_weights%s");
        gen::init_kernel<<<ggrid(n), gen::GBS, 0, (cudaStream_t)stream>>>(dyn, w, ld, n, idx, seed, slot0, generation);
    })
#undef BODY
    return check_launch("srnn_init");
}

int srnn_kfv(const srnn_net_desc *d, const float *base_host, int64_t n, double scale, uint32_t level, int32_t max_steps,
             double eps, uint64_t seed, uint64_t slot0, int32_t *time_to_vergence, int32_t *time_as_fixpoint,
             int64_t *hist, float *w_out, int64_t ld, void *stream) {
    if (!d || !base_host) return fail(SRNN_ERR_ARG, "desc or base is NULL%s");
    if (n <= 0) return n == 0 ? SRNN_OK : fail(SRNN_ERR_ARG, "n < 0%s");
    if (max_steps < 0 || max_steps > 4096) return fail(SRNN_ERR_ARG, "need 0 <= max_steps <= 4096%s");
    Eps e;
    int rc = make_eps(eps, &e);
    if (rc) return rc;
    BaseNet bn;
    memset(&bn, 0, sizeof bn);
#define BODY(S)                                                                                              \
    if (w_out && ld < S::NW) return fail(SRNN_ERR_ARG, "record stride ld is smaller than n_weights%s");       \
    memcpy(bn.v, base_host, sizeof(float) * S::NW);                                                          \
    const size_t hsm = hist ? sizeof(unsigned int) * 2 * (max_steps + 1) : 0;                                \
    if (w_out)                                                                                               \
        kfv_kernel<S, false><<<grid_for(n), BS, hsm, (cudaStream_t)stream>>>(                                \
            bn, n, scale, level, max_steps, e, seed, slot0, time_to_vergence, time_as_fixpoint,               \
            (unsigned long long *)hist, w_out, ld);                                                          \
    else                                                                                                     \
        kfv_kernel<S, true><<<grid_for(n), BS, hsm, (cudaStream_t)stream>>>(                                 \
            bn, n, scale, level, max_steps, e, seed, slot0, time_to_vergence, time_as_fixpoint,               \
            (unsigned long long *)hist, w_out, ld);
    SRNN_DISPATCH_OR_GENERIC(d, {
        if (w_out && ld < dyn.n_weights) return fail(SRNN_ERR_ARG, "record stride ld is smaller than n_weights%s");
        memcpy(bn.v, base_host, sizeof(float) * dyn.n_weights);
        gen::kfv_kernel<<<ggrid(n), gen::GBS, 0, (cudaStream_t)stream>>>(dyn, bn, n, scale, level, max_steps, e, seed, slot0,
                                                                         time_to_vergence, time_as_fixpoint,
                                                                         (unsigned long long *)hist, w_out, ld);
    })
#undef BODY
    return check_launch("srnn_kfv");
}

int srnn_soup_ordered(const srnn_net_desc *d, const srnn_soup_params *p, float *w, int64_t ld, int64_t n,
                      const int64_t *att, const int64_t *lrn, uint64_t seed, uint64_t gen, uint8_t *action,
                      double *loss, float *logged, const float *spares, int64_t spare_ld, int32_t n_spares,
                      int32_t *spares_used, void *stream) {
    if (!d) return fail(SRNN_ERR_ARG, "desc is NULL%s");
    if (n <= 0) return n == 0 ? SRNN_OK : fail(SRNN_ERR_ARG, "n < 0%s");
    SoupP q;
    int rc = make_soup(p, &q);
    if (rc) return rc;
    const Limits lim = make_limits(d);
#define BODY(S)                                                                                              \
    SRNN_CHECK_ARENA(w, ld, S)                                                                               \
    if (spares && spare_ld < S::NW) return fail(SRNN_ERR_ARG, "spare_ld is smaller than n_weights%s");        \
    soup_ordered_kernel<S><<<1, 32, sgd_smem_bytes<S>(1), (cudaStream_t)stream>>>(                            \
        w, ld, n, att, lrn, q, seed, gen, lim, action, loss, logged, spares, spare_ld, spares ? n_spares : 0, spares_used);
    SRNN_DISPATCH(d)
#undef BODY
    return check_launch("srnn_soup_ordered");
}

// [a, a + na*lda) and [b, b + nb*ldb) as address ranges (floats); used to refuse outputs that overlap the snapshot
static bool ranges_overlap(const float *a, int64_t na, int64_t lda, const float *b, int64_t nb, int64_t ldb) {
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(a), a1 = a0 + sizeof(float) * (size_t)(na * lda);
    const uintptr_t b0 = reinterpret_cast<uintptr_t>(b), b1 = b0 + sizeof(float) * (size_t)(nb * ldb);
    return a0 < b1 && b0 < a1;
}

extern "C++" {
template <class S>
static int launch_deferred_learners(const ShardTable &tab, int64_t snap_ld, int64_t n_total, float *w_out, int64_t out_ld,
                                    int64_t slot0, int64_t n, const SoupP &q, uint64_t seed, uint64_t gen, const Limits &lim,
                                    const int32_t *list, const int32_t *count, uint8_t *action, uint8_t *dead, cudaStream_t st) {
    const size_t sm = sgd_smem_bytes<S>(BS);
    int rc = allow_smem(soup_deferred_learners_kernel<S>, sm);
    if (rc) return rc;
    // sized for the expected number of learners (rate x n) with head-room; the kernel strides over whatever there is
    const double expect = (q.learn_from_rate < 1.0 ? q.learn_from_rate : 1.0) * (double)n * 1.25 + 4096.0;
    const int64_t want = std::min<int64_t>(n, (int64_t)expect);
    soup_deferred_learners_kernel<S><<<grid_for(want), BS, sm, st>>>(tab, snap_ld, n_total, w_out, out_ld, slot0, q, seed, gen,
                                                                      lim, list, count, action, dead);
    return SRNN_OK;
}
// (function templates so that `if constexpr` really discards the DEFER instantiations of the kinds that do not defer)
template <class S>
static int try_deferred_sync(bool *done, const ShardTable &tab, const float *own, int64_t snap_ld, int64_t n_total, float *w_out,
                             int64_t out_ld, int64_t slot0, int64_t n, const SoupP &q, uint64_t seed, uint64_t gen,
                             const Limits &lim, const int32_t *head, const int32_t *next, uint8_t *action, double *loss,
                             uint8_t *dead, int32_t *dl, int32_t *dc, unsigned grid, size_t sm, cudaStream_t st) {
    *done = false;
    if constexpr (soup_defers_learning<S>()) {
        if (dl && q.learn_from_rate > 0.0 && q.severity > 0 && q.train <= 0 && n >= SOUP_DEFER_MIN) {
            cudaError_t ce = cudaMemsetAsync(dc, 0, sizeof(int32_t), st);
            if (ce != cudaSuccess) return fail(SRNN_ERR_CUDA, "cudaMemsetAsync: %s", cudaGetErrorString(ce));
            soup_sync_kernel<S, SOUP_BS, SRNN_SOUP_MINB, true><<<grid, SOUP_BS, sm, st>>>(
                tab, own, snap_ld, n_total, w_out, out_ld, slot0, n, q, seed, gen, lim, head, next, action, loss, dead, dl, dc);
            int rc = launch_deferred_learners<S>(tab, snap_ld, n_total, w_out, out_ld, slot0, n, q, seed, gen, lim, dl, dc, action,
                                                 dead, st);
            if (rc) return rc;
            *done = true;
        }
    }
    return SRNN_OK;
}
template <class S>
static int try_deferred_chain(bool *done, const ShardTable &tab, const float *own, int64_t ld, int64_t n_total, float *w_out,
                              int64_t out_ld, int64_t slot0, int64_t n, const SoupP &q, uint64_t seed, uint64_t gen,
                              const Limits &lim, int32_t *head_cur, const int32_t *next_cur, const PeerListTable &nxt,
                              int make_next, uint8_t *action, double *loss, uint8_t *dead, int32_t *dl, int32_t *dc,
                              unsigned grid, size_t sm, cudaStream_t st) {
    *done = false;
    if constexpr (soup_defers_learning<S>()) {
        if (dl && q.learn_from_rate > 0.0 && q.severity > 0 && q.train <= 0 && n >= SOUP_DEFER_MIN) {
            cudaError_t ce = cudaMemsetAsync(dc, 0, sizeof(int32_t), st);
            if (ce != cudaSuccess) return fail(SRNN_ERR_CUDA, "cudaMemsetAsync: %s", cudaGetErrorString(ce));
            soup_chain_peers_kernel<S, SOUP_BS, SRNN_SOUP_MINB, true><<<grid, SOUP_BS, sm, st>>>(
                tab, own, ld, n_total, w_out, out_ld, slot0, n, q, seed, gen, lim, head_cur, next_cur, nxt, make_next, action, loss,
                dead, dl, dc);
            int rc = launch_deferred_learners<S>(tab, ld, n_total, w_out, out_ld, slot0, n, q, seed, gen, lim, dl, dc, action, dead, st);
            if (rc) return rc;
            *done = true;
        }
    }
    return SRNN_OK;
}
}  // extern "C++"

static int soup_sync_impl(const srnn_net_desc *d, const srnn_soup_params *p, const ShardTable &tab, const float *own,
                          int64_t snap_ld, int64_t n_total, float *w_out, int64_t out_ld, int64_t slot0, int64_t n,
                          uint64_t seed, uint64_t gen, int32_t *scratch, uint8_t *action, double *loss, uint8_t *dead,
                          void *stream) {
    if (!d || !scratch) return fail(SRNN_ERR_ARG, "desc or scratch is NULL%s");
    if (n <= 0) return n == 0 ? SRNN_OK : fail(SRNN_ERR_ARG, "n < 0%s");
    if (n_total >= 0x7fffffffLL || slot0 < 0 || slot0 + n > n_total)
        return fail(SRNN_ERR_ARG, "need 0 <= slot0, slot0 + n <= n_total < 2^31%s");
    SoupP q;
    int rc = make_soup(p, &q);
    if (rc) return rc;
    const Limits lim = make_limits(d);
    int32_t *head = scratch, *next = scratch + n;
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t ce = cudaMemsetAsync(head, 0xFF, sizeof(int32_t) * (size_t)n, st);
    if (ce != cudaSuccess) return fail(SRNN_ERR_CUDA, "cudaMemsetAsync: %s", cudaGetErrorString(ce));
    if (q.path < 0 || q.path > 3) return fail(SRNN_ERR_ARG, "engine_path must be 0..3%s");
    const bool fast_shape = srnn_shape_supported(d) == 1;
    const bool multipass = fast_shape && (q.path == 2 || (q.path == 0 && n >= SOUP_MULTIPASS_MIN));
    if (multipass) {
        int32_t *victims = next + n_total, *learners = victims + n;
        uint8_t *flags = reinterpret_cast<uint8_t *>(learners + n);
        int32_t *counters = learners + n + (n + 3) / 4;
        ce = cudaMemsetAsync(counters, 0, sizeof(int32_t) * 4, st);
        if (ce != cudaSuccess) return fail(SRNN_ERR_CUDA, "cudaMemsetAsync: %s", cudaGetErrorString(ce));
        soup_events_kernel<<<(unsigned)((n_total + EV_BS - 1) / EV_BS), EV_BS, 0, st>>>(q, seed, gen, n_total, slot0, n, head,
                                                                                        next, victims, learners, flags, counters);
#define BODY(S)                                                                                              \
    if (snap_ld < S::NW || out_ld < S::NW) return fail(SRNN_ERR_ARG, "record stride ld is smaller than n_weights%s"); \
    const size_t sm = sgd_smem_bytes<S>(BS);                                                                 \
    rc = allow_smem(soup_learners_kernel<S>, sm);                                                            \
    if (!rc) rc = allow_smem(soup_final_kernel<S>, sm + sizeof(float) * BS * S::NW + 1024);                  \
    if (rc) return rc;                                                                                       \
    if (q.attacking_rate > 0.0)                                                                              \
        soup_victims_kernel<S><<<grid_for(n), BS, 0, st>>>(tab, own, snap_ld, w_out, out_ld, head, next, victims, counters); \
    if (q.learn_from_rate > 0.0 && q.severity > 0)                                                           \
        soup_learners_kernel<S><<<grid_for(n), BS, sm, st>>>(tab, own, snap_ld, n_total, w_out, out_ld, slot0, q, seed, gen, \
                                                             head, learners, counters);                      \
    soup_final_kernel<S><<<grid_for(n), BS, q.train > 0 ? sm : 0, st>>>(own, snap_ld, w_out, out_ld, slot0, n, q, seed, gen, \
                                                                       lim, head, flags, action, loss, dead);
        SRNN_DISPATCH(d)
#undef BODY
        return check_launch("srnn_soup_sync (multi-pass)");
    }
    if (q.attacking_rate > 0.0)
        soup_lists_kernel<<<grid_for(n_total), BS, 0, (cudaStream_t)stream>>>(q, seed, gen, n_total, slot0, n, head, next);
#define BODY(S)                                                                                              \
    if (snap_ld < S::NW || out_ld < S::NW) return fail(SRNN_ERR_ARG, "record stride ld is smaller than n_weights%s"); \
    constexpr int SBS = SOUP_BS;                                                                             \
    const size_t sm = sgd_smem_bytes<S>(SBS);                                                                \
    const unsigned grid = (unsigned)((n + soup_tile<S>() - 1) / soup_tile<S>());                             \
    bool deferred_ = false;                                                                                  \
    rc = try_deferred_sync<S>(&deferred_, tab, own, snap_ld, n_total, w_out, out_ld, slot0, n, q, seed, gen, lim, head, next, \
                              action, loss, dead, next + n_total + n, scratch + n_total + 4 * n + 8, grid, sm, st); \
    if (rc) return rc;                                                                                       \
    if (deferred_) { }                                                                                       \
    else if (n >= SOUP_DENSE_MIN)                                                                            \
        soup_sync_kernel<S, SBS, SRNN_SOUP_MINB><<<grid, SBS, sm, (cudaStream_t)stream>>>(                   \
            tab, own, snap_ld, n_total, w_out, out_ld, slot0, n, q, seed, gen, lim, head, next, action, loss, dead, nullptr, nullptr); \
    else                                                                                                     \
        soup_sync_kernel<S, SBS, SOUP_MINB_SMALL><<<grid, SBS, sm, (cudaStream_t)stream>>>(                  \
            tab, own, snap_ld, n_total, w_out, out_ld, slot0, n, q, seed, gen, lim, head, next, action, loss, dead, nullptr, nullptr);
    SRNN_DISPATCH_OR_GENERIC(d, {
        if (snap_ld < dyn.n_weights || out_ld < dyn.n_weights) return fail(SRNN_ERR_ARG, "record stride ld is smaller than n_weights%s");
        gen::soup_sync_kernel<<<ggrid(n), gen::GBS, 0, st>>>(dyn, tab, snap_ld, n_total, w_out, out_ld, slot0, n, q, seed, gen, head,
                                                            next, action, loss, dead);
    })
#undef BODY
    return check_launch("srnn_soup_sync");
}

int srnn_soup_sync(const srnn_net_desc *d, const srnn_soup_params *p, const float *snap, int64_t snap_ld,
                   int64_t n_total, float *w_out, int64_t out_ld, int64_t slot0, int64_t n, uint64_t seed, uint64_t gen,
                   int32_t *scratch, uint8_t *action, double *loss, uint8_t *dead, void *stream) {
    if (!snap || !w_out) return fail(SRNN_ERR_ARG, "arena pointer is NULL%s");
    if (n > 0 && n_total > 0 && ranges_overlap(w_out, n, out_ld, snap, n_total, snap_ld))
        return fail(SRNN_ERR_ARG, "w_out must not overlap snap (partner records are read while w_out is written)%s");
    ShardTable tab;
    memset(&tab, 0, sizeof tab);
    tab.base[0] = snap;
    tab.len = n_total;
    tab.n = 1;
    return soup_sync_impl(d, p, tab, snap + slot0 * snap_ld, snap_ld, n_total, w_out, out_ld, slot0, n, seed, gen, scratch,
                          action, loss, dead, stream);
}

int srnn_soup_sync_peers(const srnn_net_desc *d, const srnn_soup_params *p, const float *const *shards, int32_t n_shards,
                         int64_t shard_len, int32_t my_shard, int64_t ld, float *w_out, int64_t out_ld, uint64_t seed,
                         uint64_t gen, int32_t *scratch, uint8_t *action, double *loss, uint8_t *dead, void *stream) {
    if (!shards || !w_out) return fail(SRNN_ERR_ARG, "shards or w_out is NULL%s");
    if (n_shards < 1 || n_shards > SRNN_MAX_SHARDS || my_shard < 0 || my_shard >= n_shards)
        return fail(SRNN_ERR_ARG, "need 1 <= n_shards <= SRNN_MAX_SHARDS and 0 <= my_shard < n_shards%s");
    ShardTable tab;
    memset(&tab, 0, sizeof tab);
    for (int r = 0; r < n_shards; ++r) {
        if (!shards[r]) return fail(SRNN_ERR_ARG, "a shard pointer is NULL%s");
        if (shard_len > 0 && ranges_overlap(w_out, shard_len, out_ld, shards[r], shard_len, ld))
            return fail(SRNN_ERR_ARG, "w_out must not overlap a shard%s");
        tab.base[r] = shards[r];
    }
    tab.len = shard_len;
    tab.n = n_shards;
    return soup_sync_impl(d, p, tab, shards[my_shard], ld, shard_len * n_shards, w_out, out_ld, (int64_t)my_shard * shard_len,
                          shard_len, seed, gen, scratch, action, loss, dead, stream);
}

// ---- trajectory ring: one row of the caller's ring buffer per logged generation.  One thread per watched
// particle gathers what particle.save_state(**description) records at code/soup.py:55-87 -- the record after the
// generation, the action that won the precedence, the loss of the last train epoch, the uid it ran under, the global
// index of its counterpart (attack victim / teacher, re-derived from the pairing function) -- and, when it died, the
// record it had before the respawn (looked up in the dying side channel; NaN when that overflowed).
__global__ void soup_log_watched_kernel(const float *__restrict__ w, int64_t ld, int nw, int64_t n_total, int64_t slot0,
                                        const int32_t *__restrict__ watch, int n_watch, SoupP p, uint64_t seed, uint64_t gen,
                                        const uint8_t *__restrict__ action, const double *__restrict__ loss,
                                        const int64_t *__restrict__ uid_local, float *ring_w, float *ring_dying,
                                        uint8_t *ring_action, int64_t *ring_uid, int64_t *ring_counterpart,
                                        double *ring_loss) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_watch) return;
    const int64_t v = watch[i];
    const uint8_t act = action[v];
    for (int k = 0; k < nw; ++k) ring_w[(int64_t)i * nw + k] = w[v * ld + k];
    ring_action[i] = act;
    ring_uid[i] = uid_local[v];
    ring_loss[i] = loss ? loss[v] : nan("");
    int64_t at, lr;
    soup_draw(p, seed, (uint64_t)(slot0 + v), gen, n_total, &at, &lr);
    // code/soup.py:56-68: the learn_from description overwrites the attack's; train_self sets counterpart None
    ring_counterpart[i] = lr >= 0 ? lr : at;
    const bool died = act == SRNN_ACT_DIV_DEAD || act == SRNN_ACT_ZERO_DEAD;
    int found = -1;
    if (died && p.dying_count) {
        const int cnt = min(*p.dying_count, p.dying_cap);
        for (int q = 0; q < cnt; ++q)
            if (p.dying_idx[q] == slot0 + v) { found = q; break; }
    }
    for (int k = 0; k < nw; ++k)
        ring_dying[(int64_t)i * nw + k] = found >= 0 ? p.dying_w[(int64_t)found * nw + k] : nanf("");
}

int srnn_soup_log_watched(const srnn_net_desc *d, const srnn_soup_params *p, const float *w, int64_t ld, int64_t n_total,
                          int64_t slot0, const int32_t *watch, int32_t n_watch, uint64_t seed, uint64_t gen,
                          const uint8_t *action, const double *loss, const int64_t *uid_local, float *ring_w,
                          float *ring_dying, uint8_t *ring_action, int64_t *ring_uid, int64_t *ring_counterpart,
                          double *ring_loss, void *stream) {
    if (!d || !w || !watch || !action || !uid_local || !ring_w || !ring_dying || !ring_action || !ring_uid ||
        !ring_counterpart || !ring_loss)
        return fail(SRNN_ERR_ARG, "srnn_soup_log_watched: a required pointer is NULL%s");
    if (n_watch <= 0) return n_watch == 0 ? SRNN_OK : fail(SRNN_ERR_ARG, "n_watch < 0%s");
    srnn_net_desc dd = *d;
    if (srnn_desc_init(&dd) != SRNN_OK || ld < dd.n_weights) return fail(SRNN_ERR_ARG, "bad descriptor or record stride%s");
    SoupP q;
    int rc = make_soup(p, &q);
    if (rc) return rc;
    soup_log_watched_kernel<<<(n_watch + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        w, ld, dd.n_weights, n_total, slot0, watch, n_watch, q, seed, gen, action, loss, uid_local, ring_w, ring_dying,
        ring_action, ring_uid, ring_counterpart, ring_loss);
    return check_launch("srnn_soup_log_watched");
}

static int fill_list_table(PeerListTable *t, int32_t *const *heads, int32_t *const *nexts, int32_t n_shards) {
    memset(t, 0, sizeof *t);
    for (int r = 0; r < n_shards; ++r) {
        if (!heads[r] || !nexts[r]) return fail(SRNN_ERR_ARG, "a list pointer is NULL%s");
        t->head[r] = heads[r];
        t->next[r] = nexts[r];
    }
    return SRNN_OK;
}

int srnn_soup_push_lists(const srnn_soup_params *p, int32_t n_shards, int64_t shard_len, int32_t my_shard, uint64_t seed,
                         uint64_t gen, int32_t *const *heads, int32_t *const *nexts, void *stream) {
    if (!heads || !nexts) return fail(SRNN_ERR_ARG, "heads or nexts is NULL%s");
    if (n_shards < 1 || n_shards > SRNN_MAX_SHARDS || my_shard < 0 || my_shard >= n_shards)
        return fail(SRNN_ERR_ARG, "need 1 <= n_shards <= SRNN_MAX_SHARDS and 0 <= my_shard < n_shards%s");
    if (shard_len <= 0) return shard_len == 0 ? SRNN_OK : fail(SRNN_ERR_ARG, "shard_len < 0%s");
    if (shard_len * n_shards >= 0x7fffffffLL) return fail(SRNN_ERR_ARG, "need n_total < 2^31%s");
    SoupP q;
    int rc = make_soup(p, &q);
    if (rc) return rc;
    PeerListTable t;
    rc = fill_list_table(&t, heads, nexts, n_shards);
    if (rc) return rc;
    if (q.attacking_rate > 0.0)
        soup_push_lists_kernel<<<grid_for(shard_len), BS, 0, (cudaStream_t)stream>>>(q, seed, gen, shard_len * n_shards, shard_len,
                                                                                     (int64_t)my_shard * shard_len, t);
    return check_launch("srnn_soup_push_lists");
}

int srnn_soup_chain_peers(const srnn_net_desc *d, const srnn_soup_params *p, const float *const *shards, int32_t n_shards,
                          int64_t shard_len, int32_t my_shard, int64_t ld, float *w_out, int64_t out_ld, uint64_t seed,
                          uint64_t gen, int32_t *const *heads_cur, int32_t *const *nexts_cur, int32_t *const *heads_nxt,
                          int32_t *const *nexts_nxt, int32_t *const *flags, int64_t wait_gen, int32_t *scratch,
                          uint8_t *action, double *loss, uint8_t *dead, void *stream) {
    if (!d || !shards || !w_out || !heads_cur || !nexts_cur) return fail(SRNN_ERR_ARG, "a required pointer is NULL%s");
    if (gen >= 0x7fffffffULL && flags) return fail(SRNN_ERR_ARG, "fused signalling needs gen < 2^31%s");
    if (n_shards < 1 || n_shards > SRNN_MAX_SHARDS || my_shard < 0 || my_shard >= n_shards)
        return fail(SRNN_ERR_ARG, "need 1 <= n_shards <= SRNN_MAX_SHARDS and 0 <= my_shard < n_shards%s");
    if (shard_len <= 0) return shard_len == 0 ? SRNN_OK : fail(SRNN_ERR_ARG, "shard_len < 0%s");
    const int64_t n_total = shard_len * n_shards, n = shard_len, slot0 = (int64_t)my_shard * shard_len;
    if (n_total >= 0x7fffffffLL) return fail(SRNN_ERR_ARG, "need n_total < 2^31%s");
    if ((heads_nxt == nullptr) != (nexts_nxt == nullptr)) return fail(SRNN_ERR_ARG, "pass both heads_nxt and nexts_nxt or neither%s");
    ShardTable tab;
    memset(&tab, 0, sizeof tab);
    for (int r = 0; r < n_shards; ++r) {
        if (!shards[r]) return fail(SRNN_ERR_ARG, "a shard pointer is NULL%s");
        if (ranges_overlap(w_out, shard_len, out_ld, shards[r], shard_len, ld))
            return fail(SRNN_ERR_ARG, "w_out must not overlap a shard%s");
        tab.base[r] = shards[r];
    }
    tab.len = shard_len;
    tab.n = n_shards;
    if (!heads_cur[my_shard] || !nexts_cur[my_shard]) return fail(SRNN_ERR_ARG, "this rank's list pointers are NULL%s");
    SoupP q;
    int rc = make_soup(p, &q);
    if (rc) return rc;
    const Limits lim = make_limits(d);
    PeerListTable nxt;
    memset(&nxt, 0, sizeof nxt);
    const int make_next = heads_nxt != nullptr;
    if (make_next) {
        rc = fill_list_table(&nxt, heads_nxt, nexts_nxt, n_shards);
        if (rc) return rc;
    }
    int32_t *head_cur = heads_cur[my_shard];
    const int32_t *next_cur = nexts_cur[my_shard];
    const float *own = shards[my_shard];
    PeerFlagTable ft;
    memset(&ft, 0, sizeof ft);
    if (flags) {
        for (int r = 0; r < n_shards; ++r) {
            if (!flags[r]) return fail(SRNN_ERR_ARG, "a flag pointer is NULL%s");
            ft.flag[r] = flags[r];
        }
        ft.n = n_shards;
        if (wait_gen > 0)
            soup_wait_flags_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(flags[my_shard], n_shards, (int32_t)wait_gen);
    }
#define BODY(S)                                                                                              \
    if (ld < S::NW || out_ld < S::NW) return fail(SRNN_ERR_ARG, "record stride ld is smaller than n_weights%s"); \
    constexpr int SBS = SOUP_BS;                                                                             \
    const size_t sm = sgd_smem_bytes<S>(SBS);                                                                \
    const unsigned grid = (unsigned)((n + soup_tile<S>() - 1) / soup_tile<S>());                             \
    bool deferred_ = false;                                                                                  \
    rc = try_deferred_chain<S>(&deferred_, tab, own, ld, n_total, w_out, out_ld, slot0, n, q, seed, gen, lim, head_cur, next_cur, \
                               nxt, make_next, action, loss, dead, scratch, scratch ? scratch + n + 8 : nullptr, grid, sm, \
                               (cudaStream_t)stream);                                                        \
    if (rc) return rc;                                                                                       \
    if (deferred_) { }                                                                                       \
    else if (n >= SOUP_DENSE_MIN)                                                                            \
        soup_chain_peers_kernel<S, SBS, SRNN_SOUP_MINB><<<grid, SBS, sm, (cudaStream_t)stream>>>(            \
            tab, own, ld, n_total, w_out, out_ld, slot0, n, q, seed, gen, lim, head_cur, next_cur, nxt, make_next, action, loss, dead, nullptr, nullptr); \
    else                                                                                                     \
        soup_chain_peers_kernel<S, SBS, SOUP_MINB_SMALL><<<grid, SBS, sm, (cudaStream_t)stream>>>(           \
            tab, own, ld, n_total, w_out, out_ld, slot0, n, q, seed, gen, lim, head_cur, next_cur, nxt, make_next, action, loss, dead, nullptr, nullptr);
    SRNN_DISPATCH(d)
#undef BODY
    if (ft.n > 0) soup_signal_flags_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(ft, my_shard, (int32_t)gen);
    return check_launch("srnn_soup_chain_peers");
}

// scratch layout (int32 words): headA[n] nextA[n] headL[n] nextL[n] att32[n] lrn32[n] pred[3n] wl0[n] wl1[n] stamp(n x 16 bit) cnt[4] barrier
static int soup_exact_launch(const srnn_net_desc *d, const srnn_soup_params *p, float *w, int64_t ld, int64_t n,
                             const int64_t *att, const int64_t *lrn, uint64_t seed, uint64_t gen0, int32_t n_gens,
                             int32_t *scratch, uint8_t *action, double *loss, uint8_t *dead, int32_t *rounds_dev,
                             int32_t **rounds_last, void *stream) {
    if (!d || !scratch) return fail(SRNN_ERR_ARG, "desc or scratch is NULL%s");
    if (n_gens < 0) return fail(SRNN_ERR_ARG, "n_gens < 0%s");
    if (n <= 0 || n_gens == 0) return n >= 0 ? SRNN_OK : fail(SRNN_ERR_ARG, "n < 0%s");
    if (n >= 0x7fffffffLL / 3) return fail(SRNN_ERR_ARG, "need 3 n < 2^31%s");
    if ((att == nullptr) != (lrn == nullptr))
        return fail(SRNN_ERR_ARG, "srnn_soup_exact: pass both att and lrn (explicit pairings) or neither (Philox draws)%s");
    if (att && n_gens != 1) return fail(SRNN_ERR_ARG, "explicit pairings describe ONE generation%s");
    SoupP q;
    int rc = make_soup(p, &q);
    if (rc) return rc;
    Limits lim = make_limits(d);
    cudaStream_t st = (cudaStream_t)stream;
    ExactBufs b;
    b.headA = scratch; b.nextA = b.headA + n; b.headL = b.nextA + n; b.nextL = b.headL + n;
    b.att32 = b.nextL + n; b.lrn32 = b.att32 + n; b.pred = b.lrn32 + n;
    b.wl[0] = b.pred + 3 * n; b.wl[1] = b.wl[0] + n;
    b.stamp = reinterpret_cast<uint16_t *>(b.wl[1] + n);
    b.cnt = b.wl[1] + n + (n + 1) / 2;
    b.barrier = reinterpret_cast<unsigned *>(b.cnt + 6);
    if (rounds_last) *rounds_last = b.cnt + 3;
    // Every generation of the call inside ONE cooperative launch -- except for very large soups (2^22 particles and more),
    // where the list pass is worth an ordinary launch of its own: per generation memsets + prepare + cooperative kernel.
    const bool split = n >= SOUP_EXACT_SPLIT_MIN;
    cudaError_t ce = cudaSuccess;
#define BODY(S)                                                                                              \
    SRNN_CHECK_ARENA(w, ld, S)                                                                               \
    const size_t sm = sgd_smem_bytes<S>(BS);                                                                 \
    rc = allow_smem(soup_exact_coop_kernel<S>, sm);                                                          \
    if (rc) return rc;                                                                                       \
    int per_sm = 0, dev = 0, sms = 0;                                                                        \
    cudaGetDevice(&dev);                                                                                     \
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);                                       \
    ce = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, soup_exact_coop_kernel<S>, BS, sm);          \
    if (ce != cudaSuccess || per_sm < 1) return fail(SRNN_ERR_CUDA, "occupancy query: %s", cudaGetErrorString(ce)); \
    const unsigned grid = (unsigned)std::min<int64_t>((n + BS - 1) / BS, (int64_t)per_sm * sms);              \
    for (int32_t g = 0; g < (split ? n_gens : 1); ++g) {                                                     \
        uint64_t gen_g = gen0 + (uint64_t)g;                                                                 \
        int32_t gens_g = split ? 1 : n_gens, prepared = split ? 1 : 0;                                       \
        int32_t *rounds_g = rounds_dev ? rounds_dev + g : nullptr;                                           \
        ce = cudaMemsetAsync(b.barrier, 0, sizeof(unsigned), st);                                            \
        if (split) {                                                                                         \
            if (ce == cudaSuccess) ce = cudaMemsetAsync(b.headA, 0xFF, sizeof(int32_t) * (size_t)n, st);     \
            if (ce == cudaSuccess) ce = cudaMemsetAsync(b.headL, 0xFF, sizeof(int32_t) * (size_t)n, st);     \
            if (ce == cudaSuccess) ce = cudaMemsetAsync(b.stamp, 0, sizeof(uint16_t) * (size_t)n, st);       \
            if (ce == cudaSuccess) ce = cudaMemsetAsync(b.cnt, 0, sizeof(int32_t) * 6, st);                  \
            if (ce == cudaSuccess)                                                                           \
                soup_exact_prepare_kernel<<<grid_for(n), BS, 0, st>>>(q, seed, gen_g, n, att, lrn, b);       \
        }                                                                                                    \
        if (ce != cudaSuccess) return fail(SRNN_ERR_CUDA, "cudaMemsetAsync: %s", cudaGetErrorString(ce));    \
        void *args[] = {&w, &ld, &n, &q, &seed, &gen_g, &gens_g, &lim, &att, &lrn, &b, &action, &loss, &dead, &rounds_g, \
                        &prepared};                                                                          \
        ce = cudaLaunchCooperativeKernel((void *)soup_exact_coop_kernel<S>, dim3(grid), dim3(BS), args, sm, st); \
        if (ce != cudaSuccess) return fail(SRNN_ERR_CUDA, "cooperative launch: %s", cudaGetErrorString(ce)); \
    }
    SRNN_DISPATCH(d)
#undef BODY
    return check_launch("srnn_soup_exact");
}

int srnn_soup_exact(const srnn_net_desc *d, const srnn_soup_params *p, float *w, int64_t ld, int64_t n, const int64_t *att,
                    const int64_t *lrn, uint64_t seed, uint64_t gen, int32_t *scratch, uint8_t *action, double *loss,
                    uint8_t *dead, int32_t *rounds_out, void *stream) {
    int32_t *rounds_last = nullptr;
    int rc = soup_exact_launch(d, p, w, ld, n, att, lrn, seed, gen, 1, scratch, action, loss, dead, nullptr, &rounds_last,
                               stream);
    if (rc || !rounds_out) return rc;
    *rounds_out = 0;
    if (n <= 0 || !rounds_last) return SRNN_OK;
    cudaError_t ce = cudaMemcpyAsync(rounds_out, rounds_last, sizeof(int32_t), cudaMemcpyDeviceToHost, (cudaStream_t)stream);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize((cudaStream_t)stream);
    if (ce != cudaSuccess) return fail(SRNN_ERR_CUDA, "srnn_soup_exact: %s", cudaGetErrorString(ce));
    return SRNN_OK;
}

int srnn_soup_exact_run(const srnn_net_desc *d, const srnn_soup_params *p, float *w, int64_t ld, int64_t n, uint64_t seed,
                        uint64_t gen0, int32_t n_gens, int32_t *scratch, uint8_t *action, uint8_t *dead,
                        int32_t *rounds_dev, void *stream) {
    return soup_exact_launch(d, p, w, ld, n, nullptr, nullptr, seed, gen0, n_gens, scratch, action, nullptr, dead,
                             rounds_dev, nullptr, stream);
}

int srnn_soup_sync_run(const srnn_net_desc *d, const srnn_soup_params *p, float *w_a, float *w_b, int64_t ld, int64_t n,
                       uint64_t seed, uint64_t gen0, int32_t n_gens, int32_t *scratch, uint8_t *action, uint8_t *dead,
                       void *stream) {
    if (n_gens < 0) return fail(SRNN_ERR_ARG, "n_gens < 0%s");
    if (n_gens == 0 || n <= 0) return SRNN_OK;
    const bool fast_shape = d && srnn_shape_supported(d) == 1;   // other shapes: one generic srnn_soup_sync call per generation
    if (fast_shape && p && scratch && w_a && w_b && n <= SOUP_PERSISTENT_MAX && p->engine_path == 3) {
        // opt-in (engine_path = 3): one cooperative launch runs all generations on the device.  Measured on a B200 it
        // buys little -- 22 vs 23.5 us per generation at 100k particles, slower at 1 M -- because a generation is bound
        // by the latency of its serial phases inside each tile, not by launch overhead; so it is not the default.
        if (n >= 0x7fffffffLL) return fail(SRNN_ERR_ARG, "n too large%s");
        if (w_a == w_b) return fail(SRNN_ERR_ARG, "w_a must not alias w_b%s");
        SoupP q;
        int rc = make_soup(p, &q);
        if (rc) return rc;
        Limits lim = make_limits(d);
        int32_t *head = scratch, *next = scratch + n;
        unsigned *barrier = reinterpret_cast<unsigned *>(scratch + 5 * n + 8);
        cudaStream_t st = (cudaStream_t)stream;
        cudaError_t ce = cudaMemsetAsync(barrier, 0, sizeof(unsigned), st);
        if (ce != cudaSuccess) return fail(SRNN_ERR_CUDA, "cudaMemsetAsync: %s", cudaGetErrorString(ce));
        int nn_gens = n_gens;
#define BODY(S)                                                                                              \
    if (ld < S::NW) return fail(SRNN_ERR_ARG, "record stride ld is smaller than n_weights%s");                \
    constexpr int SBS = SOUP_BS;                                                                             \
    const size_t sm = sgd_smem_bytes<S>(SBS);                                                                \
    int per_sm = 0, dev = 0, sms = 0;                                                                        \
    cudaGetDevice(&dev);                                                                                     \
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);                                       \
    ce = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, soup_persistent_kernel<S, SBS>, SBS, sm);    \
    if (ce != cudaSuccess || per_sm < 1) return fail(SRNN_ERR_CUDA, "occupancy query: %s", cudaGetErrorString(ce)); \
    const int64_t tiles = (n + soup_tile<S>() - 1) / soup_tile<S>();                                                   \
    const unsigned grid = (unsigned)std::min<int64_t>(tiles, (int64_t)per_sm * sms);                         \
    void *args[] = {&w_a, &w_b, &ld, &n, &q, &seed, &gen0, &nn_gens, &lim, &head, &next, &action, &dead, &barrier}; \
    ce = cudaLaunchCooperativeKernel((void *)soup_persistent_kernel<S, SBS>, dim3(grid), dim3(SBS), args, sm, st); \
    if (ce != cudaSuccess) return fail(SRNN_ERR_CUDA, "cooperative launch: %s", cudaGetErrorString(ce));
        SRNN_DISPATCH(d)
#undef BODY
        return check_launch("srnn_soup_sync_run (persistent)");
    }
    if (fast_shape && p && scratch && w_a && w_b && p->engine_path == 0 && n <= SOUP_CHAIN_MAX) {
        // default for small and medium soups, where launch overhead matters (20 vs 23.5 us per generation at 100k
        // particles): chained generations, one launch each -- the lists of generation g+1 are built at the end of
        // generation g.  Beyond SOUP_CHAIN_MAX the extra pairing draw per particle costs more than the launches save.
        if (w_a == w_b) return fail(SRNN_ERR_ARG, "w_a must not alias w_b%s");
        SoupP q;
        int rc = make_soup(p, &q);
        if (rc) return rc;
        const Limits lim = make_limits(d);
        cudaStream_t st = (cudaStream_t)stream;
        int32_t *headA = scratch, *nextA = headA + n, *headB = nextA + n, *nextB = headB + n;
        cudaError_t ce = cudaMemsetAsync(headA, 0xFF, sizeof(int32_t) * (size_t)n, st);
        if (ce == cudaSuccess) ce = cudaMemsetAsync(headB, 0xFF, sizeof(int32_t) * (size_t)n, st);
        if (ce != cudaSuccess) return fail(SRNN_ERR_CUDA, "cudaMemsetAsync: %s", cudaGetErrorString(ce));
        if (q.attacking_rate > 0.0)
            soup_lists_kernel<<<grid_for(n), BS, 0, st>>>(q, seed, gen0, n, 0, n, headA, nextA);
        for (int32_t g = 0; g < n_gens; ++g) {
            float *src = (g & 1) ? w_b : w_a, *dst = (g & 1) ? w_a : w_b;
            int32_t *hc = (g & 1) ? headB : headA, *nc = (g & 1) ? nextB : nextA;
            int32_t *hn = (g & 1) ? headA : headB, *nn = (g & 1) ? nextA : nextB;
            const bool last = g == n_gens - 1;
            const uint64_t gen = gen0 + (uint64_t)g;
#define BODY(S)                                                                                              \
    if (ld < S::NW) return fail(SRNN_ERR_ARG, "record stride ld is smaller than n_weights%s");                \
    constexpr int SBS = SOUP_BS;                                                                             \
    const unsigned grid = (unsigned)((n + soup_tile<S>() - 1) / soup_tile<S>());                             \
    if (n >= SOUP_DENSE_MIN)                                                                                 \
        soup_chain_kernel<S, SBS, SRNN_SOUP_MINB><<<grid, SBS, sgd_smem_bytes<S>(SBS), st>>>(                \
            src, ld, dst, n, q, seed, gen, lim, hc, nc, hn, nn, last ? action : nullptr, last ? dead : nullptr, last ? 0 : 1); \
    else if (n <= SOUP_ONE_WAVE_MAX && q.train > 0)                                                          \
        soup_chain_kernel<S, SBS, SOUP_MINB_SMALL, 1><<<(unsigned)((n + SBS - 1) / SBS), SBS, sgd_smem_bytes<S>(SBS), st>>>( \
            src, ld, dst, n, q, seed, gen, lim, hc, nc, hn, nn, last ? action : nullptr, last ? dead : nullptr, last ? 0 : 1); \
    else                                                                                                     \
        soup_chain_kernel<S, SBS, SOUP_MINB_SMALL><<<grid, SBS, sgd_smem_bytes<S>(SBS), st>>>(               \
            src, ld, dst, n, q, seed, gen, lim, hc, nc, hn, nn, last ? action : nullptr, last ? dead : nullptr, last ? 0 : 1);
            SRNN_DISPATCH(d)
#undef BODY
        }
        return check_launch("srnn_soup_sync_run");
    }
    for (int32_t g = 0; g < n_gens; ++g) {   // engine_path 1 / 2: one srnn_soup_sync call per generation
        float *src = (g & 1) ? w_b : w_a, *dst = (g & 1) ? w_a : w_b;
        int rc = srnn_soup_sync(d, p, src, ld, n, dst, ld, 0, n, seed, gen0 + (uint64_t)g, scratch, action, nullptr, dead,
                                stream);
        if (rc) return rc;
    }
    return SRNN_OK;
}

// ---------------------------------------------------------------- communicator family (NCCL resolved at run time)
extern "C++" {
namespace {
struct NcclId { char internal[SRNN_COMM_ID_BYTES]; };
struct NcclApi {
    int (*get_unique_id)(NcclId *) = nullptr;
    int (*comm_init_rank)(void **, int, NcclId, int) = nullptr;
    int (*all_gather)(const void *, void *, size_t, int, void *, cudaStream_t) = nullptr;
    int (*comm_destroy)(void *) = nullptr;
    const char *(*error_string)(int) = nullptr;
    bool ok = false;
};
const NcclApi &nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);     // the copy already in the process (e.g. torch's)
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) return;
        api.get_unique_id = reinterpret_cast<int (*)(NcclId *)>(dlsym(h, "ncclGetUniqueId"));
        api.comm_init_rank = reinterpret_cast<int (*)(void **, int, NcclId, int)>(dlsym(h, "ncclCommInitRank"));
        api.all_gather = reinterpret_cast<int (*)(const void *, void *, size_t, int, void *, cudaStream_t)>(
            dlsym(h, "ncclAllGather"));
        api.comm_destroy = reinterpret_cast<int (*)(void *)>(dlsym(h, "ncclCommDestroy"));
        api.error_string = reinterpret_cast<const char *(*)(int)>(dlsym(h, "ncclGetErrorString"));
        api.ok = api.get_unique_id && api.comm_init_rank && api.all_gather && api.comm_destroy && api.error_string;
    });
    return api;
}
int nccl_fail(const NcclApi &api, const char *what, int rc) {
    char msg[192];
    snprintf(msg, sizeof msg, "%s: %s", what, api.error_string ? api.error_string(rc) : "NCCL error");
    return fail(SRNN_ERR_CUDA, "%s", msg);
}
}  // namespace
}  // extern "C++"

int srnn_comm_unique_id(uint8_t *id_out) {
    if (!id_out) return fail(SRNN_ERR_ARG, "id_out is NULL%s");
    const NcclApi &api = nccl_api();
    if (!api.ok) return fail(SRNN_ERR_CUDA, "libnccl.so.2 could not be loaded%s");
    NcclId id;
    const int rc = api.get_unique_id(&id);
    if (rc) return nccl_fail(api, "ncclGetUniqueId", rc);
    memcpy(id_out, id.internal, SRNN_COMM_ID_BYTES);
    return SRNN_OK;
}

int srnn_comm_init(const uint8_t *id, int32_t rank, int32_t world, void **comm_out) {
    if (!id || !comm_out) return fail(SRNN_ERR_ARG, "id or comm_out is NULL%s");
    if (world < 1 || rank < 0 || rank >= world) return fail(SRNN_ERR_ARG, "need 0 <= rank < world%s");
    const NcclApi &api = nccl_api();
    if (!api.ok) return fail(SRNN_ERR_CUDA, "libnccl.so.2 could not be loaded%s");
    NcclId nid;
    memcpy(nid.internal, id, SRNN_COMM_ID_BYTES);
    void *comm = nullptr;
    const int rc = api.comm_init_rank(&comm, world, nid, rank);
    if (rc) return nccl_fail(api, "ncclCommInitRank", rc);
    *comm_out = comm;
    return SRNN_OK;
}

int srnn_comm_allgather(void *comm, const float *send, float *recv, int64_t count, void *stream) {
    if (!comm || !send || !recv) return fail(SRNN_ERR_ARG, "comm, send or recv is NULL%s");
    if (count < 0) return fail(SRNN_ERR_ARG, "count < 0%s");
    if (count == 0) return SRNN_OK;
    const NcclApi &api = nccl_api();
    if (!api.ok) return fail(SRNN_ERR_CUDA, "libnccl.so.2 could not be loaded%s");
    const int rc = api.all_gather(send, recv, (size_t)count, 7 /* ncclFloat32 */, comm, (cudaStream_t)stream);
    if (rc) return nccl_fail(api, "ncclAllGather", rc);
    return SRNN_OK;
}

int srnn_comm_destroy(void *comm) {
    if (!comm) return SRNN_OK;
    const NcclApi &api = nccl_api();
    if (!api.ok) return fail(SRNN_ERR_CUDA, "libnccl.so.2 could not be loaded%s");
    const int rc = api.comm_destroy(comm);
    if (rc) return nccl_fail(api, "ncclCommDestroy", rc);
    return SRNN_OK;
}

int srnn_assign_uids(const uint8_t *dead, int64_t n, int64_t uid_base, int64_t *uid, int64_t *block_tmp,
                     int64_t *n_dead_out, void *stream) {
    if (!dead || !uid || !block_tmp) return fail(SRNN_ERR_ARG, "dead, uid or block_tmp is NULL%s");
    if (n <= 0) return n == 0 ? SRNN_OK : fail(SRNN_ERR_ARG, "n < 0%s");
    const int64_t nb = (n + SCAN_BS - 1) / SCAN_BS;
    uid_count_kernel<<<(unsigned)nb, SCAN_BS, 0, (cudaStream_t)stream>>>(dead, n, block_tmp);
    uid_blockscan_kernel<<<1, SCAN_BS, 0, (cudaStream_t)stream>>>(block_tmp, nb, n_dead_out);
    uid_assign_kernel<<<(unsigned)nb, SCAN_BS, 0, (cudaStream_t)stream>>>(dead, n, uid_base, block_tmp, uid);
    return check_launch("srnn_assign_uids");
}

// ---- allocation-free variant of srnn_assign_uids ------------------------------------------------------------
// Some hosts of this library only ever hold device pointers that somebody else allocated: a ctypes / cgo style
// binding that passes through arenas owned by another framework and links no CUDA allocator of its own, or a
// soup loop whose device memory was carved up once at start-up (a population sized to fill HBM) and which must
// not touch cudaMalloc -- with its implicit device-wide synchronisation -- between generations.  For them the
// block-count scratch of the uid prefix sum lives inside the library image: one fixed array per GPU that
// loads libsrnn_b200.so, SRNN_UIDS_STATIC_MAX / 1024 + 1 counters = 128 KiB, held until the process exits.
static __device__ __align__(256) long long g_uid_scratch[SRNN_UIDS_STATIC_MAX / srnn::SCAN_BS + 1];

// In host code `g_uid_scratch` is only a handle for the symbol, not an address a kernel may use; the device
// address is looked up once per device (with that device current) and cached.
static int uid_scratch_of_current_device(int64_t **out) {
    constexpr int MAX_DEV = 64;
    static int64_t *cached[MAX_DEV] = {};
    static std::mutex mu;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return srnn::fail(SRNN_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= MAX_DEV) return srnn::fail(SRNN_ERR_LIMIT, "device ordinal beyond the static-scratch table%s");
    std::lock_guard<std::mutex> lock(mu);
    if (!cached[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_uid_scratch);
        if (e != cudaSuccess || !p)
            return srnn::fail(SRNN_ERR_CUDA, "cudaGetSymbolAddress(g_uid_scratch): %s", cudaGetErrorString(e));
        cached[dev] = static_cast<int64_t *>(p);
    }
    *out = cached[dev];
    return SRNN_OK;
}

int srnn_assign_uids_static(const uint8_t *dead, int64_t n, int64_t uid_base, int64_t *uid, int64_t *n_dead_out,
                            void *stream) {
    if (!dead || !uid) return fail(SRNN_ERR_ARG, "dead or uid is NULL%s");
    if (n <= 0) return n == 0 ? SRNN_OK : fail(SRNN_ERR_ARG, "n < 0%s");
    if (n > (int64_t)SRNN_UIDS_STATIC_MAX)
        return fail(SRNN_ERR_LIMIT, "srnn_assign_uids_static: n exceeds SRNN_UIDS_STATIC_MAX (2^24 particles); "
                                    "use srnn_assign_uids with caller-provided scratch%s");
    int64_t *scratch = nullptr;
    int rc = uid_scratch_of_current_device(&scratch);   // first call on a device resolves the symbol address
    if (rc) return rc;
    // The array that was just looked up belongs to the CURRENT device.  A stream created on another GPU would run the
    // kernels there, against that GPU's inputs but this GPU's scratch address: refuse it instead of corrupting memory.
    {
        int cur = -1, sdev = -1;
        if (cudaGetDevice(&cur) == cudaSuccess && cudaStreamGetDevice((cudaStream_t)stream, &sdev) == cudaSuccess &&
            sdev != cur)
            return fail(SRNN_ERR_ARG, "srnn_assign_uids_static: the stream belongs to another device than the current "
                                      "one; make the stream's device current before the call%s");
        (void)cudaGetLastError();   // a failed query (very old driver) is not an error of this call
    }
    // Nothing in the scratch has to start from zero: every counter that is read was written earlier in this
    // same call (counts -> scan -> offsets), so whatever an earlier call left behind is never observed.
    const int64_t nb = (n + SCAN_BS - 1) / SCAN_BS;
    uid_count_kernel<<<(unsigned)nb, SCAN_BS, 0, (cudaStream_t)stream>>>(dead, n, scratch);
    uid_blockscan_kernel<<<1, SCAN_BS, 0, (cudaStream_t)stream>>>(scratch, nb, n_dead_out);
    uid_assign_kernel<<<(unsigned)nb, SCAN_BS, 0, (cudaStream_t)stream>>>(dead, n, uid_base, scratch, uid);
    return check_launch("srnn_assign_uids_static");
}

int srnn_fp32_probe(float *out, int32_t grid, int32_t block, int32_t iters, void *stream) {
    if (!out || grid < 1 || block < 1 || block > BS) return fail(SRNN_ERR_ARG, "bad probe arguments%s");
    if (iters >= 0) fp32_probe_kernel<0><<<grid, block, 0, (cudaStream_t)stream>>>(out, iters);
    else if (iters > -(1 << 24)) fp32_probe_kernel<1><<<grid, block, 0, (cudaStream_t)stream>>>(out, -iters);
    else fp32_probe_kernel<2><<<grid, block, 0, (cudaStream_t)stream>>>(out, -iters - (1 << 24));
    return check_launch("srnn_fp32_probe");
}

}  // extern "C"
