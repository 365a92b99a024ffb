// layer_ops.cu -- glue kernels of the deformable-transformer layers around MSDeformAttn
// (include/kinet_layers_b200.h): fused residual-add + LayerNorm (+ positional add, + low-precision
// companions for the next GEMMs), its backward, and deterministic column sums for bias gradients.
//
// These are row-streaming, HBM/L2-bound kernels: one warp owns a row (C <= 1024 channels sit in
// registers, 16 bytes per lane and access), rows are read once and written once, the statistics
// are warp-shuffle reductions.  Column reductions (dgamma, dbeta, bias gradients) are kept in
// registers across all the rows a warp / thread visits, folded once per CTA through shared memory
// into a [CTA][column] scratch and summed by a second tiny kernel in a fixed order.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/kinet_layers_b200.h"
#include "../../include/msda_b200.h"

namespace kinet {

constexpr int kWarps = 8;           // warps (= rows in flight) per CTA
constexpr int kMaxCtas = 148 * 4;   // partial-sum scratch is [kMaxCtas][columns] (backward / column sums)

// 4 consecutive elements at element index i (a multiple of 4) of an fp32 or bf16 array
__device__ __forceinline__ float4 load4(const void* p, int dtype, size_t i)
{
    if (dtype == MSDA_F32) return __ldcs(reinterpret_cast<const float4*>(static_cast<const float*>(p) + i));
    const uint2 u = __ldcs(reinterpret_cast<const uint2*>(static_cast<const unsigned short*>(p) + i));
    float4 f;
    if (dtype == MSDA_F16) {
        const float2 lo = __half22float2(*reinterpret_cast<const __half2*>(&u.x));
        const float2 hi = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
        return make_float4(lo.x, lo.y, hi.x, hi.y);
    }
    f.x = __uint_as_float(u.x << 16); f.y = __uint_as_float(u.x & 0xffff0000u);
    f.z = __uint_as_float(u.y << 16); f.w = __uint_as_float(u.y & 0xffff0000u);
    return f;
}

__device__ __forceinline__ void store4(void* p, int dtype, size_t i, float4 f)
{
    if (dtype == MSDA_F32) {
        *reinterpret_cast<float4*>(static_cast<float*>(p) + i) = f;
    } else if (dtype == MSDA_F16) {
        const __half2 a = __floats2half2_rn(f.x, f.y), b = __floats2half2_rn(f.z, f.w);
        uint2 u;
        u.x = *reinterpret_cast<const unsigned*>(&a);
        u.y = *reinterpret_cast<const unsigned*>(&b);
        *reinterpret_cast<uint2*>(static_cast<__half*>(p) + i) = u;
    } else {
        const __nv_bfloat162 a = __floats2bfloat162_rn(f.x, f.y), b = __floats2bfloat162_rn(f.z, f.w);
        uint2 u;
        u.x = *reinterpret_cast<const unsigned*>(&a);
        u.y = *reinterpret_cast<const unsigned*>(&b);
        *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(p) + i) = u;
    }
}

__device__ __forceinline__ float load1(const void* p, int dtype, size_t i)
{
    if (dtype == MSDA_F32) return static_cast<const float*>(p)[i];
    if (dtype == MSDA_F16) return __half2float(static_cast<const __half*>(p)[i]);
    return __bfloat162float(static_cast<const __nv_bfloat16*>(p)[i]);
}

__device__ __forceinline__ void store1(void* p, int dtype, size_t i, float v)
{
    if (dtype == MSDA_F32) static_cast<float*>(p)[i] = v;
    else if (dtype == MSDA_F16) static_cast<__half*>(p)[i] = __float2half_rn(v);
    else static_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(v);
}

__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- dropout of the branch inside the add + LayerNorm kernels ----------------------------------------------
// Counter-based: the keep / drop decision of an element is a hash of its index and a per-call key, so the backward
// regenerates the mask instead of reading one.  One 32-bit hash (two multiply-xorshift rounds: full avalanche) serves
// two elements, 16 bits each, compared with p * 65536.
struct Drop {
    uint32_t thr;      // drop when the element's 16 bits are below thr; 0 = no dropout
    uint32_t key;
    float scale;       // 1 / (1 - thr / 65536)
};
__device__ __forceinline__ uint32_t mix32(uint32_t x)
{
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}
__device__ __forceinline__ Drop make_drop(float p, uint32_t site, uint64_t seed, const int64_t* step)
{
    Drop d;
    d.thr = p > 0.f ? (uint32_t)fminf(p * 65536.f + 0.5f, 65535.f) : 0u;
    d.scale = 65536.f / (float)(65536u - d.thr);
    const uint64_t st = step ? (uint64_t)*step : 0ull;
    d.key = mix32((uint32_t)seed ^ mix32((uint32_t)(seed >> 32) ^ mix32(site * 0x9e3779b1U ^ mix32((uint32_t)st ^ (uint32_t)(st >> 32) * 0x85ebca6bU))));
    return d;
}
// keep bits (bit i = element i of the vector survives) of vector `vec` (= row * vectors per row + v, modulo 2^31)
__device__ __forceinline__ uint32_t drop_keep4(const Drop& d, uint32_t vec)
{
    const uint32_t h0 = mix32(vec * 2u + d.key), h1 = mix32(vec * 2u + 1u + d.key);
    return ((h0 & 0xffffu) >= d.thr ? 1u : 0u) | ((h0 >> 16) >= d.thr ? 2u : 0u) |
           ((h1 & 0xffffu) >= d.thr ? 4u : 0u) | ((h1 >> 16) >= d.thr ? 8u : 0u);
}
__device__ __forceinline__ float4 drop_scale4(float4 r, uint32_t keep, float scale)
{
    return make_float4((keep & 1u) ? r.x * scale : 0.f, (keep & 2u) ? r.y * scale : 0.f,
                       (keep & 4u) ? r.z * scale : 0.f, (keep & 8u) ? r.w * scale : 0.f);
}

// KMAX: 16-byte vectors per lane; a row has C / 4 vectors, lane l owns vectors l, l + 32, ...
template <int KMAX>
__global__ void __launch_bounds__(kWarps * 32, KMAX <= 3 ? 6 : (KMAX <= 4 ? 4 : 2))
add_layernorm_fwd_kernel(const kinet_ln_fwd_args a)
{
    const Drop drop = make_drop(a.r ? a.drop_p : 0.f, a.drop_site, a.drop_seed, a.drop_step);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nvec = a.C >> 2;
    const float inv_c = 1.f / (float)a.C;
    // gamma / beta are re-read per row (L1 hits) instead of living in registers: the kernel waits on DRAM
    // latency, and 40 instead of 56-64 registers put 6 instead of 4 CTAs (rows in flight) on an SM
    const float4* __restrict__ gamma4 = reinterpret_cast<const float4*>(a.gamma);
    const float4* __restrict__ beta4 = reinterpret_cast<const float4*>(a.beta);
    for (int64_t row = (int64_t)blockIdx.x * kWarps + warp; row < a.rows; row += (int64_t)gridDim.x * kWarps) {
        const size_t base = (size_t)row * a.C;
        float4 z[KMAX];
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < KMAX; ++k) {
            const int v = lane + 32 * k;
            z[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (v < nvec) {
                z[k] = load4(a.x, a.x_dtype, base + 4 * v);
                if (a.r) {
                    float4 r = load4(a.r, a.r_dtype, base + 4 * v);
                    if (drop.thr) r = drop_scale4(r, drop_keep4(drop, (uint32_t)row * (uint32_t)nvec + (uint32_t)v), drop.scale);
                    z[k].x += r.x; z[k].y += r.y; z[k].z += r.z; z[k].w += r.w;
                }
                s += (z[k].x + z[k].y) + (z[k].z + z[k].w);
            }
        }
        const float mean = warp_sum(s) * inv_c;
        float q = 0.f;
#pragma unroll
        for (int k = 0; k < KMAX; ++k) {
            if (lane + 32 * k < nvec) {
                const float dx = z[k].x - mean, dy = z[k].y - mean, dz = z[k].z - mean, dw = z[k].w - mean;
                q += (dx * dx + dy * dy) + (dz * dz + dw * dw);
            }
        }
        const float rstd = rsqrtf(warp_sum(q) * inv_c + a.eps);
        if (a.mean_rstd && lane == 0) reinterpret_cast<float2*>(a.mean_rstd)[row] = make_float2(mean, rstd);
        const size_t pbase = a.pos ? (size_t)(row % a.pos_rows) * a.C : 0;
#pragma unroll
        for (int k = 0; k < KMAX; ++k) {
            const int v = lane + 32 * k;
            if (v < nvec) {
                const float4 gam = __ldg(gamma4 + v), bet = __ldg(beta4 + v);
                float4 y;
                y.x = (z[k].x - mean) * rstd * gam.x + bet.x;
                y.y = (z[k].y - mean) * rstd * gam.y + bet.y;
                y.z = (z[k].z - mean) * rstd * gam.z + bet.z;
                y.w = (z[k].w - mean) * rstd * gam.w + bet.w;
                store4(a.y, a.y_dtype, base + 4 * v, y);
                if (a.y_lp) store4(a.y_lp, a.lp_dtype, base + 4 * v, y);
                if (a.ypos_lp) {
                    const float4 p = load4(a.pos, a.pos_dtype, pbase + 4 * v);
                    store4(a.ypos_lp, a.lp_dtype, base + 4 * v, make_float4(y.x + p.x, y.y + p.y, y.z + p.z, y.w + p.w));
                }
            }
        }
    }
}

// Fold per-warp column accumulators of a CTA into partial[blockIdx][col0 .. col0 + C).
template <int KMAX>
__device__ __forceinline__ void fold_columns(float4 (&acc)[KMAX], float* smem, float* partial_row, int C, int lane, int warp)
{
    const int nvec = C >> 2;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < KMAX; ++k) {
        const int v = lane + 32 * k;
        if (v < nvec) reinterpret_cast<float4*>(smem + (size_t)warp * C)[v] = acc[k];
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) t += smem[(size_t)w * C + c];
        partial_row[c] = t;
    }
}

// EXT: dropout of the branch and / or the column sums of dr (want_dbias); its own instantiation so that the plain
// kernel keeps its register budget (4 / 3 CTAs per SM).
template <int KMAX, bool EXT>
__global__ void __launch_bounds__(kWarps * 32, EXT ? (KMAX <= 2 ? 3 : (KMAX <= 4 ? 2 : 1)) : (KMAX <= 2 ? 4 : (KMAX <= 3 ? 3 : (KMAX <= 4 ? 2 : 1))))
add_layernorm_bwd_kernel(const kinet_ln_bwd_args a)
{
    extern __shared__ float smem[];   // [kWarps][C]
    const Drop drop = make_drop((EXT && a.r) ? a.drop_p : 0.f, a.drop_site, a.drop_seed, a.drop_step);
    float4 dbr[EXT ? KMAX : 1];
#pragma unroll
    for (int k = 0; k < (EXT ? KMAX : 1); ++k) dbr[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nvec = a.C >> 2;
    const float inv_c = 1.f / (float)a.C;
    // gamma is re-read per row (L1 hits), like in the forward: registers buy rows in flight
    const float4* __restrict__ gamma4 = reinterpret_cast<const float4*>(a.gamma);
    float4 dg[KMAX], db[KMAX];
#pragma unroll
    for (int k = 0; k < KMAX; ++k) dg[k] = db[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t row = (int64_t)blockIdx.x * kWarps + warp; row < a.rows; row += (int64_t)gridDim.x * kWarps) {
        const size_t base = (size_t)row * a.C;
        const float2 ms = reinterpret_cast<const float2*>(a.mean_rstd)[row];
        float4 xh[KMAX], wd[KMAX];
        uint32_t keep[EXT ? KMAX : 1];
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int k = 0; k < KMAX; ++k) {
            const int v = lane + 32 * k;
            xh[k] = wd[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (v < nvec) {
                float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
                if (a.dy) g = load4(a.dy, a.dy_dtype, base + 4 * v);
                if (a.dy_lp) {
                    const float4 t = load4(a.dy_lp, a.lp_dtype, base + 4 * v);
                    g.x += t.x; g.y += t.y; g.z += t.z; g.w += t.w;
                }
                if (a.dypos_lp) {
                    const float4 t = load4(a.dypos_lp, a.lp_dtype, base + 4 * v);
                    g.x += t.x; g.y += t.y; g.z += t.z; g.w += t.w;
                }
                float4 z = load4(a.x, a.x_dtype, base + 4 * v);
                if (a.r) {
                    float4 r = load4(a.r, a.r_dtype, base + 4 * v);
                    if constexpr (EXT) {
                        keep[k] = 15u;
                        if (drop.thr) {
                            keep[k] = drop_keep4(drop, (uint32_t)row * (uint32_t)nvec + (uint32_t)v);
                            r = drop_scale4(r, keep[k], drop.scale);
                        }
                    }
                    z.x += r.x; z.y += r.y; z.z += r.z; z.w += r.w;
                }
                xh[k] = make_float4((z.x - ms.x) * ms.y, (z.y - ms.x) * ms.y, (z.z - ms.x) * ms.y, (z.w - ms.x) * ms.y);
                dg[k].x += g.x * xh[k].x; dg[k].y += g.y * xh[k].y; dg[k].z += g.z * xh[k].z; dg[k].w += g.w * xh[k].w;
                db[k].x += g.x; db[k].y += g.y; db[k].z += g.z; db[k].w += g.w;
                const float4 gam = __ldg(gamma4 + v);
                wd[k] = make_float4(g.x * gam.x, g.y * gam.y, g.z * gam.z, g.w * gam.w);
                c1 += (wd[k].x + wd[k].y) + (wd[k].z + wd[k].w);
                c2 += (wd[k].x * xh[k].x + wd[k].y * xh[k].y) + (wd[k].z * xh[k].z + wd[k].w * xh[k].w);
            }
        }
        c1 = warp_sum(c1) * inv_c;
        c2 = warp_sum(c2) * inv_c;
#pragma unroll
        for (int k = 0; k < KMAX; ++k) {
            const int v = lane + 32 * k;
            if (v < nvec) {
                float4 d;
                d.x = ms.y * (wd[k].x - c1 - xh[k].x * c2);
                d.y = ms.y * (wd[k].y - c1 - xh[k].y * c2);
                d.z = ms.y * (wd[k].z - c1 - xh[k].z * c2);
                d.w = ms.y * (wd[k].w - c1 - xh[k].w * c2);
                store4(a.dx, a.x_dtype, base + 4 * v, d);
                if constexpr (EXT) {
                    if (a.r) {
                        const float4 e = drop.thr ? drop_scale4(d, keep[k], drop.scale) : d;     // d (x + dropout(r)) / d r
                        if (a.dr) store4(a.dr, a.r_dtype, base + 4 * v, e);
                        dbr[k].x += e.x; dbr[k].y += e.y; dbr[k].z += e.z; dbr[k].w += e.w;
                    }
                } else {
                    if (a.dr) store4(a.dr, a.r_dtype, base + 4 * v, d);
                }
            }
        }
    }
    const int ncol = (EXT && a.want_dbias) ? 3 : 2;
    float* prow = a.partial + (size_t)blockIdx.x * ncol * a.C;
    fold_columns<KMAX>(dg, smem, prow, a.C, lane, warp);
    fold_columns<KMAX>(db, smem, prow + a.C, a.C, lane, warp);
    if constexpr (EXT) {
        if (a.want_dbias) fold_columns<KMAX>(dbr, smem, prow + 2 * a.C, a.C, lane, warp);
    }
}

// out[c] = sum over the G rows of partial[G][n], fixed order.  A CTA of 32 x 32 threads owns 32 columns:
// thread (tx, ty) adds rows ty, ty + 32, ... of column tx (independent loads, a few latency rounds instead
// of G dependent ones), the 32 row lanes are folded through shared memory.
__global__ void __launch_bounds__(1024)
colsum_finalize_kernel(const float* __restrict__ partial, int G, int n, void* out, int out_dtype)
{
    __shared__ float tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    float t0 = 0.f, t1 = 0.f, t2 = 0.f, t3 = 0.f;
    if (c < n) {
        int g = ty;
        for (; g + 96 < G; g += 128) {
            t0 += partial[(size_t)g * n + c];
            t1 += partial[(size_t)(g + 32) * n + c];
            t2 += partial[(size_t)(g + 64) * n + c];
            t3 += partial[(size_t)(g + 96) * n + c];
        }
        for (; g < G; g += 32) t0 += partial[(size_t)g * n + c];
    }
    tile[ty][tx] = (t0 + t1) + (t2 + t3);
    __syncthreads();
    if (ty == 0 && c < n) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 32; ++w) t += tile[w][tx];
        store1(out, out_dtype, c, t);
    }
}

// Column sums of a [rows][C] matrix.  Thread layout: tx = column vector (4 columns), ty = row lane;
// a CTA walks rows ty, ty + RL, ... of its slab with 4 independent loads in flight per thread.
// RELU: the matrix summed is dh = dy * (h > 0) (ReLU backward), which is also written out -- the bias
// gradient of the Linear layer in front of a ReLU costs no pass of its own.
template <bool RELU>
__global__ void __launch_bounds__(256)
colsum_kernel(const void* __restrict__ a, const void* __restrict__ h, void* __restrict__ dh, int dtype,
              int64_t rows, int C, int RL, float* __restrict__ partial, float scale)
{
    extern __shared__ float smem[];   // [RL][C]
    const int nvec = C >> 2;
    const int tx = threadIdx.x % nvec, ty = threadIdx.x / nvec;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    auto fetch = [&](int64_t row) {
        const size_t i = (size_t)row * C + 4 * tx;
        float4 v = load4(a, dtype, i);
        if constexpr (RELU) {
            const float4 o = load4(h, dtype, i);
            v.x = o.x > 0.f ? v.x * scale : 0.f; v.y = o.y > 0.f ? v.y * scale : 0.f;
            v.z = o.z > 0.f ? v.z * scale : 0.f; v.w = o.w > 0.f ? v.w * scale : 0.f;
            store4(dh, dtype, i, v);
        }
        return v;
    };
    if (ty < RL) {
        const int64_t stride = (int64_t)gridDim.x * RL;
        int64_t row = (int64_t)blockIdx.x * RL + ty;
        for (; row + 3 * stride < rows; row += 4 * stride) {
            const float4 v0 = fetch(row), v1 = fetch(row + stride), v2 = fetch(row + 2 * stride), v3 = fetch(row + 3 * stride);
            acc.x += (v0.x + v1.x) + (v2.x + v3.x);
            acc.y += (v0.y + v1.y) + (v2.y + v3.y);
            acc.z += (v0.z + v1.z) + (v2.z + v3.z);
            acc.w += (v0.w + v1.w) + (v2.w + v3.w);
        }
        for (; row < rows; row += stride) {
            const float4 v0 = fetch(row);
            acc.x += v0.x; acc.y += v0.y; acc.z += v0.z; acc.w += v0.w;
        }
        reinterpret_cast<float4*>(smem + (size_t)ty * C)[tx] = acc;
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float t = 0.f;
        for (int w = 0; w < RL; ++w) t += smem[(size_t)w * C + c];
        partial[(size_t)blockIdx.x * C + c] = t;
    }
}

// Few rows (the decoder's 300-800 queries x batch): one launch, no scratch.  A CTA of 32 x 32 threads owns 32
// columns; thread (tx, ty) adds rows ty, ty + 32, ... of column tx, the row lanes fold through shared memory.
template <bool RELU>
__global__ void __launch_bounds__(1024)
colsum_small_kernel(const void* __restrict__ a, const void* __restrict__ h, void* __restrict__ dh, int dtype,
                    int rows, int C, void* out, int out_dtype, float scale)
{
    __shared__ float tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    float t0 = 0.f, t1 = 0.f;
    if (c < C) {
        for (int r = ty; r < rows; r += 64) {
            const int r1 = r + 32;
            float v0 = load1(a, dtype, (size_t)r * C + c);
            float v1 = r1 < rows ? load1(a, dtype, (size_t)r1 * C + c) : 0.f;
            if constexpr (RELU) {
                v0 = load1(h, dtype, (size_t)r * C + c) > 0.f ? v0 * scale : 0.f;
                if (r1 < rows) v1 = load1(h, dtype, (size_t)r1 * C + c) > 0.f ? v1 * scale : 0.f;
                store1(dh, dtype, (size_t)r * C + c, v0);
                if (r1 < rows) store1(dh, dtype, (size_t)r1 * C + c, v1);
            }
            t0 += v0;
            t1 += v1;
        }
    }
    tile[ty][tx] = t0 + t1;
    __syncthreads();
    if (ty == 0 && c < C) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 32; ++w) t += tile[w][tx];
        store1(out, out_dtype, c, t);
    }
}

// out = dropout(x): element e survives when drop_keep4's bit says so (the same hash as the add + LayerNorm kernels, vector
// index = e / 4), scaled by 1 / (1 - p').  No mask tensor: for the activations of a ReLU the backward needs none either --
// dropout(relu(a)) > 0 exactly where the element survived AND the ReLU was open (kinet_b200_relu_dropout_backward_colsum).
__global__ void __launch_bounds__(256)
dropout_fwd_kernel(const void* __restrict__ x, void* __restrict__ out, int dtype, int64_t nvec, float p, uint32_t site,
                   uint64_t seed, const int64_t* __restrict__ step)
{
    const Drop drop = make_drop(p, site, seed, step);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < nvec; i += 4 * stride) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = load4(x, dtype, (size_t)(i + u * stride) * 4);
#pragma unroll
        for (int u = 0; u < 4; ++u)
            store4(out, dtype, (size_t)(i + u * stride) * 4, drop_scale4(v[u], drop_keep4(drop, (uint32_t)(i + u * stride)), drop.scale));
    }
    for (; i < nvec; i += stride)
        store4(out, dtype, (size_t)i * 4, drop_scale4(load4(x, dtype, (size_t)i * 4), drop_keep4(drop, (uint32_t)i), drop.scale));
}

constexpr int kSmallRows = 2048;

inline bool ok_dtype(int d) { return d == MSDA_F32 || d == MSDA_BF16 || d == MSDA_F16; }
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
inline bool aligned_for(const void* p, int dtype) { return (reinterpret_cast<uintptr_t>(p) & (dtype == MSDA_F32 ? 15 : 7)) == 0; }

// One wave: as many CTAs as are resident at once (a grid of a few more CTAs than that runs a second, nearly
// empty wave and doubles the kernel's time), never more than the scratch has rows for.
template <typename K>
int resident_ctas(K kernel, int threads, size_t smem)
{
    int dev = 0, sms = 148, per_sm = 1;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem) != cudaSuccess || per_sm < 1) {
        (void)cudaGetLastError();
        per_sm = 1;
    }
    return sms * per_sm;
}

inline int row_grid(int64_t rows, int resident)
{
    const int64_t need = (rows + kWarps - 1) / kWarps;
    return (int)(need < resident ? (need > 0 ? need : 1) : resident);
}

template <int KMAX>
int launch_fwd(const kinet_ln_fwd_args& a, cudaStream_t st)
{
    static thread_local int resident = 0;     // per kernel instantiation and host thread; one device per process
    if (!resident) resident = resident_ctas(add_layernorm_fwd_kernel<KMAX>, kWarps * 32, 0);
    add_layernorm_fwd_kernel<KMAX><<<row_grid(a.rows, resident), kWarps * 32, 0, st>>>(a);
    return (int)cudaGetLastError();
}

template <int KMAX, bool EXT>
int launch_bwd_impl(const kinet_ln_bwd_args& a, int* grid, cudaStream_t st)
{
    const size_t smem = (size_t)kWarps * a.C * sizeof(float);
    const int resident = resident_ctas(add_layernorm_bwd_kernel<KMAX, EXT>, kWarps * 32, smem);
    *grid = row_grid(a.rows, resident < kMaxCtas ? resident : kMaxCtas);   // the scratch has kMaxCtas rows
    add_layernorm_bwd_kernel<KMAX, EXT><<<*grid, kWarps * 32, smem, st>>>(a);
    return (int)cudaGetLastError();
}

template <int KMAX>
int launch_bwd(const kinet_ln_bwd_args& a, int* grid, cudaStream_t st)
{
    const bool ext = a.want_dbias || (a.r && a.drop_p > 0.f);
    return ext ? launch_bwd_impl<KMAX, true>(a, grid, st) : launch_bwd_impl<KMAX, false>(a, grid, st);
}

}  // namespace kinet

using namespace kinet;

extern "C" {

int kinet_b200_add_layernorm_forward(const kinet_ln_fwd_args* args, void* stream)
{
    if (!args) return MSDA_E_NULL;
    const kinet_ln_fwd_args& a = *args;
    if (!a.x || !a.y || !a.gamma || !a.beta) return MSDA_E_NULL;
    if (a.rows < 0 || a.C <= 0 || (a.C & 3) || a.C > KINET_LN_MAX_C) return MSDA_E_SHAPE;
    if (!(a.drop_p >= 0.f && a.drop_p < 1.f)) return MSDA_E_SHAPE;
    if (!ok_dtype(a.x_dtype) || !ok_dtype(a.y_dtype) || (a.r && !ok_dtype(a.r_dtype))) return MSDA_E_DTYPE;
    if ((a.y_lp || a.ypos_lp) && !ok_dtype(a.lp_dtype)) return MSDA_E_DTYPE;
    if (a.ypos_lp && (!a.pos || !ok_dtype(a.pos_dtype) || a.pos_rows <= 0)) return MSDA_E_NULL;
    if (!aligned_for(a.x, a.x_dtype) || !aligned_for(a.y, a.y_dtype) || !aligned16(a.gamma) || !aligned16(a.beta) ||
        (a.r && !aligned_for(a.r, a.r_dtype)) || (a.y_lp && !aligned_for(a.y_lp, a.lp_dtype)) ||
        (a.ypos_lp && (!aligned_for(a.ypos_lp, a.lp_dtype) || !aligned_for(a.pos, a.pos_dtype))) ||
        (a.mean_rstd && (reinterpret_cast<uintptr_t>(a.mean_rstd) & 7)))
        return MSDA_E_ALIGN;
    if (a.rows == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const int k = (a.C + 127) / 128;
    if (k <= 2) return launch_fwd<2>(a, st);
    if (k <= 3) return launch_fwd<3>(a, st);
    if (k <= 4) return launch_fwd<4>(a, st);
    return launch_fwd<8>(a, st);
}

size_t kinet_b200_colsum_partial_floats(int32_t n_columns)
{
    return n_columns > 0 ? (size_t)kMaxCtas * (size_t)n_columns : 0;
}

int kinet_b200_add_layernorm_backward(const kinet_ln_bwd_args* args, void* stream)
{
    if (!args) return MSDA_E_NULL;
    const kinet_ln_bwd_args& a = *args;
    if (!a.x || !a.dx || !a.gamma || !a.mean_rstd || !a.dgamma_dbeta || !a.partial) return MSDA_E_NULL;
    if (!a.dy && !a.dy_lp && !a.dypos_lp) return MSDA_E_NULL;
    if (a.r == nullptr && (a.dr != nullptr || a.want_dbias)) return MSDA_E_NULL;
    if (a.rows < 0 || a.C <= 0 || (a.C & 3) || a.C > KINET_LN_MAX_C) return MSDA_E_SHAPE;
    if (!(a.drop_p >= 0.f && a.drop_p < 1.f)) return MSDA_E_SHAPE;
    if (!ok_dtype(a.x_dtype) || (a.r && !ok_dtype(a.r_dtype)) || (a.dy && !ok_dtype(a.dy_dtype)) ||
        ((a.dy_lp || a.dypos_lp) && !ok_dtype(a.lp_dtype)))
        return MSDA_E_DTYPE;
    if (!aligned_for(a.x, a.x_dtype) || !aligned_for(a.dx, a.x_dtype) || !aligned16(a.gamma) ||
        (a.r && !aligned_for(a.r, a.r_dtype)) || (a.dr && !aligned_for(a.dr, a.r_dtype)) ||
        (a.dy && !aligned_for(a.dy, a.dy_dtype)) || (a.dy_lp && !aligned_for(a.dy_lp, a.lp_dtype)) ||
        (a.dypos_lp && !aligned_for(a.dypos_lp, a.lp_dtype)) || !aligned16(a.partial))
        return MSDA_E_ALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    int grid = 0;
    if (a.rows > 0) {
        const int k = (a.C + 127) / 128;
        int rc;
        if (k <= 2) rc = launch_bwd<2>(a, &grid, st);
        else if (k <= 3) rc = launch_bwd<3>(a, &grid, st);
        else if (k <= 4) rc = launch_bwd<4>(a, &grid, st);
        else rc = launch_bwd<8>(a, &grid, st);
        if (rc) return rc;
    }
    const int n = (a.want_dbias ? 3 : 2) * a.C;
    colsum_finalize_kernel<<<(n + 31) / 32, 1024, 0, st>>>(a.partial, grid, n, a.dgamma_dbeta, MSDA_F32);
    return (int)cudaGetLastError();
}

static int colsum_common(const void* a, const void* h, void* dh, int32_t a_dtype, int64_t rows, int32_t C, void* out,
                         int32_t out_dtype, float* partial, void* stream, float scale = 1.f)
{
    const bool relu = h != nullptr;
    if (!out || !partial || (rows > 0 && !a)) return MSDA_E_NULL;
    if (relu && !dh) return MSDA_E_NULL;
    if (rows < 0 || C <= 0 || (C & 3) || C > KINET_LN_MAX_C) return MSDA_E_SHAPE;
    if (!ok_dtype(a_dtype) || !ok_dtype(out_dtype)) return MSDA_E_DTYPE;
    if ((rows > 0 && !aligned_for(a, a_dtype)) || !aligned16(partial) ||
        (relu && (!aligned_for(h, a_dtype) || !aligned_for(dh, a_dtype))))
        return MSDA_E_ALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    if (rows <= kSmallRows) {
        const int g = (C + 31) / 32;
        if (relu) colsum_small_kernel<true><<<g, 1024, 0, st>>>(a, h, dh, a_dtype, (int)rows, C, out, out_dtype, scale);
        else colsum_small_kernel<false><<<g, 1024, 0, st>>>(a, nullptr, nullptr, a_dtype, (int)rows, C, out, out_dtype, 1.f);
        return (int)cudaGetLastError();
    }
    const int nvec = C >> 2;
    const int RL = 256 / nvec;     // row lanes per CTA (nvec <= 256)
    int grid = 0;
    if (rows > 0) {
        // at least 8 rows per row lane, at most kMaxCtas CTAs
        const int64_t want = (rows + (int64_t)RL * 8 - 1) / ((int64_t)RL * 8);
        grid = (int)(want < kMaxCtas ? (want > 0 ? want : 1) : kMaxCtas);
        const size_t smem = (size_t)RL * C * sizeof(float);
        if (relu) colsum_kernel<true><<<grid, 256, smem, st>>>(a, h, dh, a_dtype, rows, C, RL, partial, scale);
        else colsum_kernel<false><<<grid, 256, smem, st>>>(a, nullptr, nullptr, a_dtype, rows, C, RL, partial, 1.f);
        if (int rc = (int)cudaGetLastError()) return rc;
    }
    colsum_finalize_kernel<<<(C + 31) / 32, 1024, 0, st>>>(partial, grid, C, out, out_dtype);
    return (int)cudaGetLastError();
}

int kinet_b200_colsum(const void* a, int32_t a_dtype, int64_t rows, int32_t C, void* out, int32_t out_dtype,
                      float* partial, void* stream)
{
    return colsum_common(a, nullptr, nullptr, a_dtype, rows, C, out, out_dtype, partial, stream);
}

int kinet_b200_relu_backward_colsum(const void* dy, const void* h, int32_t dtype, int64_t rows, int32_t C, void* dh,
                                    void* out, int32_t out_dtype, float* partial, void* stream)
{
    if (!h) return MSDA_E_NULL;
    return colsum_common(dy, h, dh, dtype, rows, C, out, out_dtype, partial, stream);
}

int kinet_b200_relu_dropout_backward_colsum(const void* dy, const void* h_dropped, float drop_p, int32_t dtype, int64_t rows,
                                            int32_t C, void* dh, void* out, int32_t out_dtype, float* partial, void* stream)
{
    if (!h_dropped) return MSDA_E_NULL;
    if (!(drop_p >= 0.f && drop_p < 1.f)) return MSDA_E_SHAPE;
    const uint32_t thr = drop_p > 0.f ? (uint32_t)fminf(drop_p * 65536.f + 0.5f, 65535.f) : 0u;
    return colsum_common(dy, h_dropped, dh, dtype, rows, C, out, out_dtype, partial, stream, 65536.f / (float)(65536u - thr));
}

int kinet_b200_dropout_forward(const void* x, void* out, int32_t dtype, int64_t n, float drop_p, uint32_t drop_site,
                               uint64_t drop_seed, const int64_t* drop_step, void* stream)
{
    if (n < 0 || (n & 3)) return MSDA_E_SHAPE;
    if (!(drop_p >= 0.f && drop_p < 1.f)) return MSDA_E_SHAPE;
    if (!ok_dtype(dtype)) return MSDA_E_DTYPE;
    if (n == 0) return 0;
    if (!x || !out) return MSDA_E_NULL;
    if (!aligned_for(x, dtype) || !aligned_for(out, dtype)) return MSDA_E_ALIGN;
    const int64_t nvec = n >> 2;
    int64_t grid = (nvec + 256 * 4 - 1) / (256 * 4);
    if (grid > 148 * 16) grid = 148 * 16;
    dropout_fwd_kernel<<<(int)grid, 256, 0, (cudaStream_t)stream>>>(x, out, dtype, nvec, drop_p, drop_site, drop_seed, drop_step);
    return (int)cudaGetLastError();
}

}  // extern "C"
