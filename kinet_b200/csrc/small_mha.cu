// small_mha.cu -- attention core of nn.MultiheadAttention for the short token sequences of the `kine` path
// (include/kinet_layers_b200.h: kinet_b200_small_mha_forward / _backward).
//
// The reference's KinematicTransformer / DualKinematicTransformer (src/trackformer/models/transformer.py:85-185)
// runs plain nn.MultiheadAttention layers (:316, :368-369: self.self_attn / self.multihead_attn) over <= 150 object
// + T tracklet tokens of dimension 288 (cfgs/train_kinet.yaml:17-18: 8 heads of 36 channels).  Sequences of a few
// hundred tokens with a head dimension that is not a multiple of 8 are what library attention kernels handle
// worst (16-bit only, padded heads, or an fp32 "math" path of ~8 launches per call); here one kernel does
// softmax(q k^T * scale + key_padding_mask) v for every (batch, head, 16-query tile) with the keys / values of the
// head staged in shared memory as fp32 (up to 256 keys per pass: one global-memory round trip for the usual sequence
// lengths; longer ones are walked pass by pass with an online softmax), and two kernels do the backward (dK / dV per
// key tile, dQ per query tile; the probabilities are recomputed from the saved log-sum-exp, every output element has
// exactly one writer: no atomics, bit-reproducible).
//
// All arithmetic is fp32 whatever the storage type.  Nothing here is GEMM-sized (N * H * Lq * Lk * D is a few tens
// of MFLOP): the kernels are latency-bound, so what matters is ONE launch instead of a chain of tiny ones inside a
// CUDA-graphed frame, batched global loads, and enough independent shared-memory loads per FMA group.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/kinet_layers_b200.h"
#include "../../include/msda_b200.h"

namespace kinet_mha {

constexpr int kDMax = KINET_MHA_MAX_D;  // channels per head: lane l owns channels l and l + 32

template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __half from_f<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// four consecutive elements (the address is aligned to 4 elements)
template <typename T> __device__ __forceinline__ float4 ld4(const T* p);
template <> __device__ __forceinline__ float4 ld4<float>(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
template <> __device__ __forceinline__ float4 ld4<__half>(const __half* p)
{
    const uint2 u = __ldg(reinterpret_cast<const uint2*>(p));
    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&u.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
    return make_float4(a.x, a.y, b.x, b.y);
}
template <> __device__ __forceinline__ float4 ld4<__nv_bfloat16>(const __nv_bfloat16* p)
{
    const uint2 u = __ldg(reinterpret_cast<const uint2*>(p));
    return make_float4(__uint_as_float(u.x << 16), __uint_as_float(u.x & 0xffff0000u),
                       __uint_as_float(u.y << 16), __uint_as_float(u.y & 0xffff0000u));
}

__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// rows [r0, r0 + rows) x D of a strided (batch, row, head, channel) tensor -> shared fp32 [row][stride], scaled by
// `mul`; rows at or beyond `limit` are zero-filled.  VEC: the caller checked D % 4 == 0 and 4-element alignment of
// the base and the row stride.  The loads of U loop iterations are issued before the first store: with U = 9 the 256
// threads of the forward fetch 256 keys of 36 channels in ONE global-memory round trip (the kernel is latency-bound).
template <typename T, bool VEC, int THREADS, int U = 5>
__device__ __forceinline__ void stage_rows(float* dst, int stride, const T* __restrict__ src, int64_t row_stride,
                                           int r0, int rows, int limit, int D, float mul)
{
    if constexpr (VEC) {
        const int vpr = D >> 2, total = rows * vpr;
        for (int i0 = threadIdx.x; i0 < total; i0 += U * THREADS) {
            float4 x[U];
            int r[U], c[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = i0 + u * THREADS;
                r[u] = i / vpr; c[u] = (i - r[u] * vpr) * 4;
                x[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (i < total && r0 + r[u] < limit) x[u] = ld4<T>(src + (size_t)(r0 + r[u]) * row_stride + c[u]);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (i0 + u * THREADS < total) {
                    float* d = dst + r[u] * stride + c[u];
                    d[0] = x[u].x * mul; d[1] = x[u].y * mul; d[2] = x[u].z * mul; d[3] = x[u].w * mul;
                }
            }
        }
    } else {
        const int total = rows * D;
        for (int i0 = threadIdx.x; i0 < total; i0 += U * THREADS) {
            float x[U];
            int r[U], c[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = i0 + u * THREADS;
                r[u] = i / D; c[u] = i - r[u] * D;
                x[u] = (i < total && r0 + r[u] < limit) ? to_f<T>(src[(size_t)(r0 + r[u]) * row_stride + c[u]]) : 0.f;
            }
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (i0 + u * THREADS < total) dst[r[u] * stride + c[u]] = x[u] * mul;
        }
    }
}

// (the stride of a dimension of size 1 is never multiplied by a non-zero index: it may be anything)
static inline bool vec_ok(const void* p, int n, int64_t batch_stride, int rows, int64_t row_stride, int D, int elem)
{
    return (D & 3) == 0 && (n <= 1 || (batch_stride & 3) == 0) && (rows <= 1 || (row_stride & 3) == 0) &&
           ((uintptr_t)p % (4 * elem)) == 0;
}

// ---------------------------------------------------------------- forward ----------------------------------------------------------------
constexpr int kFwdWarps = 8;
constexpr int kFwdThreads = kFwdWarps * 32;
constexpr int kFwdQPW = 2;                          // queries per warp, processed together
constexpr int kFwdQT = kFwdWarps * kFwdQPW;         // queries per CTA

// shared memory (floats): k [KT][D | 1] (odd stride: lanes reading one column of 32 rows hit 32 banks), v [KT][D],
// q [QT][D] (scaled), p [QT][KT], dead [KT] bytes
static inline size_t fwd_smem_bytes(int D, int KT)
{
    return ((size_t)KT * ((D | 1) + D) + (size_t)kFwdQT * (D + KT)) * 4 + KT;
}

// KPL keys per lane and pass: KT = 32 * KPL keys are staged per pass
template <typename T, int KPL, bool VEC>
__global__ void __launch_bounds__(kFwdThreads)
small_mha_forward_kernel(const kinet_mha_args a)
{
    constexpr int KT = 32 * KPL, QPW = kFwdQPW, QT = kFwdQT;
    extern __shared__ float4 smem_raw[];
    const int D = a.D, ks = D | 1;
    float* sk = reinterpret_cast<float*>(smem_raw);
    float* sv = sk + KT * ks;
    float* sq = sv + KT * D;
    float* sp = sq + QT * D;
    unsigned char* sdead = reinterpret_cast<unsigned char*>(sp + QT * KT);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = blockIdx.x / a.H, h = blockIdx.x - n * a.H;
    const int q0 = blockIdx.y * QT;
    const T* qb = static_cast<const T*>(a.q) + (size_t)n * a.q_batch_stride + (size_t)h * D;
    const T* kb = static_cast<const T*>(a.k) + (size_t)n * a.k_batch_stride + (size_t)h * D;
    const T* vb = static_cast<const T*>(a.v) + (size_t)n * a.v_batch_stride + (size_t)h * D;
    stage_rows<T, VEC, kFwdThreads, 1>(sq, D, qb, a.q_row_stride, q0, QT, a.Lq, D, a.scale);

    float m_run[QPW], l_run[QPW], acc[QPW][2];
#pragma unroll
    for (int qi = 0; qi < QPW; ++qi) { m_run[qi] = -INFINITY; l_run[qi] = 0.f; acc[qi][0] = acc[qi][1] = 0.f; }
    const bool c0 = lane < D, c1 = lane + 32 < D;

    for (int k0 = 0; k0 < a.Lk; k0 += KT) {
        if (k0) __syncthreads();             // the previous pass is consumed
        stage_rows<T, VEC, kFwdThreads, 9>(sk, ks, kb, a.k_row_stride, k0, KT, a.Lk, D, 1.f);
        stage_rows<T, VEC, kFwdThreads, 9>(sv, D, vb, a.v_row_stride, k0, KT, a.Lk, D, 1.f);
        for (int j = threadIdx.x; j < KT; j += kFwdThreads)
            sdead[j] = (k0 + j >= a.Lk) || (a.key_padding_mask && a.key_padding_mask[(size_t)n * a.Lk + k0 + j]);
        __syncthreads();
        // scores of the warp's queries against this lane's KPL keys
        float s[QPW][KPL];
#pragma unroll
        for (int qi = 0; qi < QPW; ++qi)
#pragma unroll
            for (int t = 0; t < KPL; ++t) s[qi][t] = 0.f;
        const float* kr = sk + lane * ks;
        const float* qr = sq + warp * QPW * D;
#pragma unroll 2
        for (int c = 0; c < D; ++c) {
            float kk[KPL];
#pragma unroll
            for (int t = 0; t < KPL; ++t) kk[t] = kr[t * 32 * ks + c];
#pragma unroll
            for (int qi = 0; qi < QPW; ++qi) {
                const float qc = qr[qi * D + c];
#pragma unroll
                for (int t = 0; t < KPL; ++t) s[qi][t] = fmaf(qc, kk[t], s[qi][t]);
            }
        }
        float alpha[QPW];
#pragma unroll
        for (int qi = 0; qi < QPW; ++qi) {
            float mx = -INFINITY;
#pragma unroll
            for (int t = 0; t < KPL; ++t) {
                if (sdead[lane + 32 * t]) s[qi][t] = -INFINITY;
                mx = fmaxf(mx, s[qi][t]);
            }
            const float m_new = fmaxf(m_run[qi], warp_max(mx));
            alpha[qi] = 1.f;
            float sum = 0.f;
            float* pr = sp + (warp * QPW + qi) * KT;
            const bool live = m_new != -INFINITY;              // (warp-uniform) at least one live key so far
            if (live) alpha[qi] = __expf(m_run[qi] - m_new);   // exp(-inf) = 0 for the first live pass
#pragma unroll
            for (int t = 0; t < KPL; ++t) {
                const float p = live ? __expf(s[qi][t] - m_new) : 0.f;
                sum += p;
                pr[lane + 32 * t] = p;
            }
            l_run[qi] = l_run[qi] * alpha[qi] + warp_sum(sum);
            m_run[qi] = m_new;
        }
        __syncwarp();
        // out += p v: lane = channel; four keys per step (one 16-byte broadcast load of p per query)
        float t0[QPW], t1[QPW];
#pragma unroll
        for (int qi = 0; qi < QPW; ++qi) t0[qi] = t1[qi] = 0.f;
        const float* pr = sp + warp * QPW * KT;
        const float* v0 = sv + (c0 ? lane : 0);
        const float* v1 = sv + (c1 ? lane + 32 : 0);
        const int kmax = min(KT, a.Lk - k0);                // keys beyond the sequence carry p = 0 and v = 0: skip them
#pragma unroll 2
        for (int j = 0; j < kmax; j += 4) {
            float x0[4], x1[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) { x0[e] = v0[(j + e) * D]; x1[e] = v1[(j + e) * D]; }
#pragma unroll
            for (int qi = 0; qi < QPW; ++qi) {
                const float4 p = *reinterpret_cast<const float4*>(pr + qi * KT + j);
                t0[qi] = fmaf(p.x, x0[0], t0[qi]); t1[qi] = fmaf(p.x, x1[0], t1[qi]);
                t0[qi] = fmaf(p.y, x0[1], t0[qi]); t1[qi] = fmaf(p.y, x1[1], t1[qi]);
                t0[qi] = fmaf(p.z, x0[2], t0[qi]); t1[qi] = fmaf(p.z, x1[2], t1[qi]);
                t0[qi] = fmaf(p.w, x0[3], t0[qi]); t1[qi] = fmaf(p.w, x1[3], t1[qi]);
            }
        }
#pragma unroll
        for (int qi = 0; qi < QPW; ++qi) {
            acc[qi][0] = fmaf(acc[qi][0], alpha[qi], t0[qi]);
            acc[qi][1] = fmaf(acc[qi][1], alpha[qi], t1[qi]);
        }
    }
    T* ob = static_cast<T*>(a.out) + (size_t)n * a.out_batch_stride + (size_t)h * D;
#pragma unroll
    for (int qi = 0; qi < QPW; ++qi) {
        const int q = q0 + warp * QPW + qi;
        if (q >= a.Lq) continue;
        // every key masked: 0, as torch >= 2.5 computes it (its attention uses a "safe" softmax; torch 1.5, which the
        // reference was written for, returned NaN)
        const float inv = l_run[qi] > 0.f ? 1.f / l_run[qi] : 0.f;
        T* o = ob + (size_t)q * a.out_row_stride;
        if (c0) o[lane] = from_f<T>(acc[qi][0] * inv);
        if (c1) o[lane + 32] = from_f<T>(acc[qi][1] * inv);
        if (a.lse && lane == 0) a.lse[((size_t)n * a.H + h) * a.Lq + q] = l_run[qi] > 0.f ? m_run[qi] + logf(l_run[qi]) : INFINITY;
    }
}

// ---------------------------------------------------------------- backward ----------------------------------------------------------------
constexpr int kWarps = 4;
constexpr int kThreads = kWarps * 32;
constexpr int kQPW = 4;                 // queries per warp, processed together (K / V rows are read once for the four)
constexpr int kQT = kWarps * kQPW;      // queries per tile
constexpr int kKT = 64;                 // keys per tile: lane l owns keys l and l + 32

struct Shared {
    float k[kKT * (kDMax + 1)];     // row stride D | 1
    float v[kKT * (kDMax + 1)];     // row stride D | 1 (the backward reads columns of v rows, like k)
    float q[kQT * kDMax];           // scaled queries of the current query tile
    float g[kQT * kDMax];           // rows of d out
    float p[kQT * kKT];             // probabilities of the tile, [query][key]
    float ds[kQT * kKT];            // d logits of the tile
    float lse[kQT], delta[kQT];
    unsigned char dead[kKT];        // key masked or beyond Lk
};

// p and d logits of the warp's four queries against the staged key tile, written to sm.p / sm.ds ([query][key]).
// Needs sm.q (scaled queries), sm.g (d out rows), sm.lse, sm.delta, sm.k, sm.v, sm.dead.
__device__ __forceinline__ void tile_p_ds(Shared& sm, int D, int ks, int warp, int lane)
{
    float s[kQPW][2], dp[kQPW][2];
#pragma unroll
    for (int qi = 0; qi < kQPW; ++qi) s[qi][0] = s[qi][1] = dp[qi][0] = dp[qi][1] = 0.f;
    const float* k0 = sm.k + lane * ks;
    const float* k1 = sm.k + (lane + 32) * ks;
    const float* v0 = sm.v + lane * ks;
    const float* v1 = sm.v + (lane + 32) * ks;
    const float* q = sm.q + warp * kQPW * D;
    const float* g = sm.g + warp * kQPW * D;
#pragma unroll 2
    for (int c = 0; c < D; ++c) {
        const float a = k0[c], b = k1[c], x = v0[c], y = v1[c];
#pragma unroll
        for (int qi = 0; qi < kQPW; ++qi) {
            const float qc = q[qi * D + c], gc = g[qi * D + c];
            s[qi][0] = fmaf(qc, a, s[qi][0]);
            s[qi][1] = fmaf(qc, b, s[qi][1]);
            dp[qi][0] = fmaf(gc, x, dp[qi][0]);
            dp[qi][1] = fmaf(gc, y, dp[qi][1]);
        }
    }
    const bool d0 = sm.dead[lane], d1 = sm.dead[lane + 32];
#pragma unroll
    for (int qi = 0; qi < kQPW; ++qi) {
        const int r = warp * kQPW + qi;
        const float lse = sm.lse[r], dl = sm.delta[r];
        // rows beyond Lq (and rows whose keys were all masked) carry lse = +inf: p = 0
        const float p0 = d0 ? 0.f : __expf(s[qi][0] - lse), p1 = d1 ? 0.f : __expf(s[qi][1] - lse);
        sm.p[r * kKT + lane] = p0;
        sm.p[r * kKT + lane + 32] = p1;
        sm.ds[r * kKT + lane] = p0 * (dp[qi][0] - dl);
        sm.ds[r * kKT + lane + 32] = p1 * (dp[qi][1] - dl);
    }
}

// query rows [q0, q0 + kQT) -> sm.q (scaled), sm.g (d out), sm.lse, sm.delta = <d out, out>
template <typename T, bool VEC>
__device__ __forceinline__ void stage_query_tile(Shared& sm, const kinet_mha_bwd_args& b, int n, int h, int q0, int warp, int lane)
{
    const kinet_mha_args& a = b.fwd;
    const int D = a.D;
    const T* qb = static_cast<const T*>(a.q) + (size_t)n * a.q_batch_stride + (size_t)h * D;
    const T* gb = static_cast<const T*>(b.dout) + (size_t)n * b.dout_batch_stride + (size_t)h * D;
    const T* ob = static_cast<const T*>(a.out) + (size_t)n * a.out_batch_stride + (size_t)h * D;
    stage_rows<T, VEC, kThreads>(sm.q, D, qb, a.q_row_stride, q0, kQT, a.Lq, D, a.scale);
    stage_rows<T, VEC, kThreads>(sm.g, D, gb, b.dout_row_stride, q0, kQT, a.Lq, D, 1.f);
    float d[kQPW];
#pragma unroll
    for (int qi = 0; qi < kQPW; ++qi) {
        const int q = q0 + warp * kQPW + qi;
        d[qi] = 0.f;
        if (q < a.Lq) {
            const T* o = ob + (size_t)q * a.out_row_stride;
            const T* g = gb + (size_t)q * b.dout_row_stride;
            if (lane < D) d[qi] = to_f<T>(o[lane]) * to_f<T>(g[lane]);
            if (lane + 32 < D) d[qi] = fmaf(to_f<T>(o[lane + 32]), to_f<T>(g[lane + 32]), d[qi]);
        }
    }
#pragma unroll
    for (int qi = 0; qi < kQPW; ++qi) {
        const int r = warp * kQPW + qi, q = q0 + r;
        const float t = warp_sum(d[qi]);
        if (lane == 0) {
            sm.delta[r] = t;
            sm.lse[r] = q < a.Lq ? a.lse[((size_t)n * a.H + h) * a.Lq + q] : INFINITY;
        }
    }
}

template <typename T, bool VEC>
__device__ __forceinline__ void stage_key_tile(Shared& sm, const kinet_mha_args& a, int n, int h, int k0)
{
    const int D = a.D, ks = D | 1;
    const T* kb = static_cast<const T*>(a.k) + (size_t)n * a.k_batch_stride + (size_t)h * D;
    const T* vb = static_cast<const T*>(a.v) + (size_t)n * a.v_batch_stride + (size_t)h * D;
    stage_rows<T, VEC, kThreads>(sm.k, ks, kb, a.k_row_stride, k0, kKT, a.Lk, D, 1.f);
    stage_rows<T, VEC, kThreads>(sm.v, ks, vb, a.v_row_stride, k0, kKT, a.Lk, D, 1.f);
    if (threadIdx.x < kKT) {
        const int j = k0 + threadIdx.x;
        sm.dead[threadIdx.x] = (j >= a.Lk) || (a.key_padding_mask && a.key_padding_mask[(size_t)n * a.Lk + j]);
    }
}

// dK, dV of one key tile: loop over the query tiles; warp w then owns keys 16 w .. 16 w + 15 of the tile and lane l
// channels l, l + 32, and accumulates over the tile's queries from sm.p / sm.ds.
template <typename T, bool VEC>
__global__ void __launch_bounds__(kThreads)
small_mha_backward_kv_kernel(const kinet_mha_bwd_args b)
{
    extern __shared__ float4 smem_raw[];
    Shared& sm = *reinterpret_cast<Shared*>(smem_raw);
    const kinet_mha_args& a = b.fwd;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = blockIdx.x / a.H, h = blockIdx.x - n * a.H;
    const int k0 = blockIdx.y * kKT;
    const int D = a.D, ks = D | 1;
    constexpr int KPW = kKT / kWarps;
    stage_key_tile<T, VEC>(sm, a, n, h, k0);
    float dk[KPW][2], dv[KPW][2];
#pragma unroll
    for (int j = 0; j < KPW; ++j) dk[j][0] = dk[j][1] = dv[j][0] = dv[j][1] = 0.f;
    const bool c0 = lane < D, c1 = lane + 32 < D;
    for (int q0 = 0; q0 < a.Lq; q0 += kQT) {
        __syncthreads();
        stage_query_tile<T, VEC>(sm, b, n, h, q0, warp, lane);
        __syncthreads();
        tile_p_ds(sm, D, ks, warp, lane);
        __syncthreads();
        const int rmax = min(kQT, a.Lq - q0);
        for (int r = 0; r < rmax; ++r) {
            const float g0 = c0 ? sm.g[r * D + lane] : 0.f, g1 = c1 ? sm.g[r * D + lane + 32] : 0.f;
            const float x0 = c0 ? sm.q[r * D + lane] : 0.f, x1 = c1 ? sm.q[r * D + lane + 32] : 0.f;   // scale * q
#pragma unroll
            for (int j4 = 0; j4 < KPW; j4 += 4) {
                const float4 p = *reinterpret_cast<const float4*>(sm.p + r * kKT + warp * KPW + j4);
                const float4 d = *reinterpret_cast<const float4*>(sm.ds + r * kKT + warp * KPW + j4);
                const float pp[4] = {p.x, p.y, p.z, p.w}, dd[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    dv[j4 + e][0] = fmaf(pp[e], g0, dv[j4 + e][0]); dv[j4 + e][1] = fmaf(pp[e], g1, dv[j4 + e][1]);
                    dk[j4 + e][0] = fmaf(dd[e], x0, dk[j4 + e][0]); dk[j4 + e][1] = fmaf(dd[e], x1, dk[j4 + e][1]);
                }
            }
        }
    }
    T* dkb = static_cast<T*>(b.dk);
    T* dvb = static_cast<T*>(b.dv);
#pragma unroll
    for (int j = 0; j < KPW; ++j) {
        const int key = k0 + warp * KPW + j;
        if (key >= a.Lk) continue;
        const size_t o = ((size_t)n * a.Lk + key) * ((size_t)a.H * D) + (size_t)h * D;    // dk / dv: [N][Lk][H * D] contiguous
        if (c0) { dkb[o + lane] = from_f<T>(dk[j][0]); dvb[o + lane] = from_f<T>(dv[j][0]); }
        if (c1) { dkb[o + lane + 32] = from_f<T>(dk[j][1]); dvb[o + lane + 32] = from_f<T>(dv[j][1]); }
    }
}

// dQ of one query tile: loop over the key tiles; dq[q][c] = scale * sum_j ds[q][j] k[j][c]
template <typename T, bool VEC>
__global__ void __launch_bounds__(kThreads)
small_mha_backward_q_kernel(const kinet_mha_bwd_args b)
{
    extern __shared__ float4 smem_raw[];
    Shared& sm = *reinterpret_cast<Shared*>(smem_raw);
    const kinet_mha_args& a = b.fwd;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = blockIdx.x / a.H, h = blockIdx.x - n * a.H;
    const int q0 = blockIdx.y * kQT;
    const int D = a.D, ks = D | 1;
    stage_query_tile<T, VEC>(sm, b, n, h, q0, warp, lane);
    float dq[kQPW][2];
#pragma unroll
    for (int qi = 0; qi < kQPW; ++qi) dq[qi][0] = dq[qi][1] = 0.f;
    const bool c0 = lane < D, c1 = lane + 32 < D;
    for (int k0 = 0; k0 < a.Lk; k0 += kKT) {
        __syncthreads();
        stage_key_tile<T, VEC>(sm, a, n, h, k0);
        __syncthreads();
        tile_p_ds(sm, D, ks, warp, lane);
        __syncwarp();
        const float* dsr = sm.ds + warp * kQPW * kKT;
        const float* x0p = sm.k + (c0 ? lane : 0);
        const float* x1p = sm.k + (c1 ? lane + 32 : 0);
        const int kmax = min(kKT, a.Lk - k0);
#pragma unroll 2
        for (int j = 0; j < kmax; j += 4) {
            float x0[4], x1[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) { x0[e] = x0p[(j + e) * ks]; x1[e] = x1p[(j + e) * ks]; }
#pragma unroll
            for (int qi = 0; qi < kQPW; ++qi) {
                const float4 d = *reinterpret_cast<const float4*>(dsr + qi * kKT + j);
                dq[qi][0] = fmaf(d.x, x0[0], dq[qi][0]); dq[qi][1] = fmaf(d.x, x1[0], dq[qi][1]);
                dq[qi][0] = fmaf(d.y, x0[1], dq[qi][0]); dq[qi][1] = fmaf(d.y, x1[1], dq[qi][1]);
                dq[qi][0] = fmaf(d.z, x0[2], dq[qi][0]); dq[qi][1] = fmaf(d.z, x1[2], dq[qi][1]);
                dq[qi][0] = fmaf(d.w, x0[3], dq[qi][0]); dq[qi][1] = fmaf(d.w, x1[3], dq[qi][1]);
            }
        }
    }
    T* dqb = static_cast<T*>(b.dq);
#pragma unroll
    for (int qi = 0; qi < kQPW; ++qi) {
        const int q = q0 + warp * kQPW + qi;
        if (q >= a.Lq) continue;
        const size_t o = ((size_t)n * a.Lq + q) * ((size_t)a.H * D) + (size_t)h * D;      // dq: [N][Lq][H * D] contiguous
        if (c0) dqb[o + lane] = from_f<T>(dq[qi][0] * a.scale);
        if (c1) dqb[o + lane + 32] = from_f<T>(dq[qi][1] * a.scale);
    }
}

// ---------------------------------------------------------------- host ----------------------------------------------------------------
static int check_args(const kinet_mha_args& a)
{
    if (!a.q || !a.k || !a.v || !a.out) return MSDA_E_NULL;
    if (a.dtype != MSDA_F32 && a.dtype != MSDA_BF16 && a.dtype != MSDA_F16) return MSDA_E_DTYPE;
    if (a.N < 0 || a.H <= 0 || a.D <= 0 || a.D > KINET_MHA_MAX_D || a.Lq < 0 || a.Lk < 0) return MSDA_E_SHAPE;
    if ((long long)a.N * a.H > 0x7fffffffLL || a.Lq > 65535 * kQT || a.Lk > 65535 * kKT) return MSDA_E_SHAPE;
    return 0;
}

// kernels whose shared memory exceeds the 48 KB a kernel gets without asking
template <typename K> static int configure(K kern, int& configured_dev, size_t bytes)
{
    int dev = 0;
    cudaGetDevice(&dev);
    if (configured_dev != dev) {
        const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        if (e != cudaSuccess) return (int)e;
        configured_dev = dev;
    }
    return 0;
}

template <typename T, int KPL, bool VEC>
static int launch_forward_impl(const kinet_mha_args& a, cudaStream_t s)
{
    static thread_local int dev = -1;
    auto kern = small_mha_forward_kernel<T, KPL, VEC>;
    // configured for the widest head this instantiation is used with: the attribute is set once per device
    if (const int e = configure(kern, dev, fwd_smem_bytes(KPL == 8 ? 40 : kDMax, 32 * KPL))) return e;
    const dim3 grid((unsigned)(a.N * a.H), (unsigned)((a.Lq + kFwdQT - 1) / kFwdQT));
    kern<<<grid, kFwdThreads, fwd_smem_bytes(a.D, 32 * KPL), s>>>(a);
    return (int)cudaGetLastError();
}

template <typename T>
static int launch_forward(const kinet_mha_args& a, cudaStream_t s)
{
    const int e = (int)sizeof(T);
    const bool vec = vec_ok(a.q, a.N, a.q_batch_stride, a.Lq, a.q_row_stride, a.D, e) && vec_ok(a.k, a.N, a.k_batch_stride, a.Lk, a.k_row_stride, a.D, e) &&
                     vec_ok(a.v, a.N, a.v_batch_stride, a.Lk, a.v_row_stride, a.D, e);
    // 256 keys per pass for heads of <= 40 channels (<= 92 KB of shared memory), 128 beyond
    if (a.D <= 40 && a.Lk > 128) return vec ? launch_forward_impl<T, 8, true>(a, s) : launch_forward_impl<T, 8, false>(a, s);
    return vec ? launch_forward_impl<T, 4, true>(a, s) : launch_forward_impl<T, 4, false>(a, s);
}

template <typename T, bool VEC>
static int launch_backward_impl(const kinet_mha_bwd_args& b, cudaStream_t s)
{
    const kinet_mha_args& a = b.fwd;
    static thread_local int dev_kv = -1, dev_q = -1;
    auto kv = small_mha_backward_kv_kernel<T, VEC>;
    auto kq = small_mha_backward_q_kernel<T, VEC>;
    if (const int e = configure(kv, dev_kv, sizeof(Shared))) return e;
    if (const int e = configure(kq, dev_q, sizeof(Shared))) return e;
    if (a.Lk > 0) {
        const dim3 grid((unsigned)(a.N * a.H), (unsigned)((a.Lk + kKT - 1) / kKT));
        kv<<<grid, kThreads, sizeof(Shared), s>>>(b);
        const int e = (int)cudaGetLastError();
        if (e) return e;
    }
    if (a.Lq > 0) {
        const dim3 grid((unsigned)(a.N * a.H), (unsigned)((a.Lq + kQT - 1) / kQT));
        kq<<<grid, kThreads, sizeof(Shared), s>>>(b);
    }
    return (int)cudaGetLastError();
}

template <typename T>
static int launch_backward(const kinet_mha_bwd_args& b, cudaStream_t s)
{
    const kinet_mha_args& a = b.fwd;
    const int e = (int)sizeof(T);
    const bool vec = vec_ok(a.q, a.N, a.q_batch_stride, a.Lq, a.q_row_stride, a.D, e) && vec_ok(a.k, a.N, a.k_batch_stride, a.Lk, a.k_row_stride, a.D, e) &&
                     vec_ok(a.v, a.N, a.v_batch_stride, a.Lk, a.v_row_stride, a.D, e) && vec_ok(b.dout, a.N, b.dout_batch_stride, a.Lq, b.dout_row_stride, a.D, e);
    return vec ? launch_backward_impl<T, true>(b, s) : launch_backward_impl<T, false>(b, s);
}

}  // namespace kinet_mha

extern "C" {

int kinet_b200_small_mha_forward(const kinet_mha_args* args, void* stream)
{
    using namespace kinet_mha;
    if (!args) return MSDA_E_NULL;
    const int rc = check_args(*args);
    if (rc) return rc;
    const kinet_mha_args& a = *args;
    if (a.N == 0 || a.Lq == 0) return 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (a.dtype == MSDA_F32) return launch_forward<float>(a, s);
    if (a.dtype == MSDA_BF16) return launch_forward<__nv_bfloat16>(a, s);
    return launch_forward<__half>(a, s);
}

int kinet_b200_small_mha_backward(const kinet_mha_bwd_args* args, void* stream)
{
    using namespace kinet_mha;
    if (!args) return MSDA_E_NULL;
    const int rc = check_args(args->fwd);
    if (rc) return rc;
    const kinet_mha_bwd_args& b = *args;
    const kinet_mha_args& a = b.fwd;
    if (!a.lse || !b.dout || !b.dq || !b.dk || !b.dv) return MSDA_E_NULL;
    if (a.N == 0) return 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (a.dtype == MSDA_F32) return launch_backward<float>(b, s);
    if (a.dtype == MSDA_BF16) return launch_backward<__nv_bfloat16>(b, s);
    return launch_backward<__half>(b, s);
}

}  // extern "C"
