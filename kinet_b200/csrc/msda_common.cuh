// msda_common.cuh -- shared device helpers for the sm_100a MSDeformAttn kernels.
//
// Semantics follow the reference kernels (LuisCossioUOH/Kinet,
// src/trackformer/models/ops/src/cuda/ms_deform_im2col_cuda.cuh): pixel-centre
// shift x = loc_x*W - 0.5 (:227-228), open sample window -1 < x < W (:229),
// zero contribution from corners outside [0,W-1]x[0,H-1] (:24-67).
// Nothing here is derived from that file's code; it is a new design:
//   * a CTA owns one (batch, head) pair and a compact patch of queries, so the gather
//     footprint of the patch (one head's 64/128-byte rows) stays L1-resident;
//   * a group of GS lanes owns one (query, head) row, each lane VEC channels
//     (16-byte loads: fp32x4 / bf16x8), several rows per warp;
//   * spatial shapes / level starts are staged in shared memory by the CTA
//     prologue (level starts are computed there when the caller passes none).
#pragma once
#include <cassert>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/msda_b200.h"

namespace msda {

constexpr int kMaxLevels = MSDA_MAX_LEVELS;

// Per-level geometry staged in shared memory.
struct LevelTable {
    int H[kMaxLevels];
    int W[kMaxLevels];
    int start[kMaxLevels];       // first pixel of the level in the flattened pyramid (value rows)
    int qstart[kMaxLevels + 1];  // encoder form: first query of the level (prefix sums of H*W)
    int min_w;                   // smallest level width (W == 1 levels need a per-sample column step)
};

struct KernelDims {
    int N, S, M, D, L, Lq, P;
    int tile_q;   // query rows per work item
    int band_h;   // encoder form (Lq == S): rows of a band of the band-ordered query walk; 0: plain query order
    int acc16;    // backward, 16-bit values: grad_value_acc IS the 16-bit grad_value; packed 16-bit reductions (sparse query sets)
};

// Fills the level table (all threads call; ends with __syncthreads()).
__device__ __forceinline__ void stage_levels(LevelTable& t, const int64_t* __restrict__ shapes,
                                             const int64_t* __restrict__ level_start,
                                             const KernelDims& d)
{
    if (threadIdx.x == 0) {
        int run = 0, min_w = 1 << 30;
        for (int l = 0; l < d.L; ++l) {
            const int H = (int)shapes[2 * l], W = (int)shapes[2 * l + 1];
            t.H[l] = H;
            t.W[l] = W;
            t.start[l] = level_start ? (int)level_start[l] : run;
            t.qstart[l] = run;
            run += H * W;
            min_w = min(min_w, W);
        }
        t.qstart[d.L] = run;
        t.min_w = min_w;
    }
    __syncthreads();
}

// Work items.  An item is (chunk of tile_q consecutive positions of the query walk, batch, head); item
// ids run head-fastest, so the CTAs in flight at any time cover all heads of the same few chunks, whose
// loc / attn / grad rows are contiguous in DRAM.  ONE ITEM PER CTA, and several times more CTAs than
// resident slots: the hardware's CTA scheduler hands out items as SMs free up.  The SMs do not run at
// the same speed (profiles/r1k_per_sm_finish_times.txt: with equal static shares the same dozen SMs
// finish 15-20 % late in every run), so equal shares make the kernel as slow as its slowest SM.
//
// Encoder form: the walk visits the pixels of a level band by band (band_h rows), column by column
// inside a band, so tile_q consecutive positions are a band_h x (tile_q / band_h) patch of one level --
// the gather footprint of an item is compact and stays in L1 -- while the number of items depends only
// on Lq (the host never needs the level shapes, which live in device memory).
__device__ __forceinline__ int walk_query(const LevelTable& t, const KernelDims& d, int pos)
{
    if (d.band_h == 0) return pos;
    int l = 0;
    while (l + 1 < d.L && pos >= t.qstart[l + 1]) ++l;
    const int W = t.W[l], H = t.H[l];
    const int r = pos - t.qstart[l];
    const int band = r / (d.band_h * W);
    const int rr = r - band * d.band_h * W;
    const int hb = min(d.band_h, H - band * d.band_h);   // the last band of a level may be lower
    const int ix = rr / hb, iy = rr - ix * hb;
    const int q = t.qstart[l] + (band * d.band_h + iy) * W + ix;
    return q < d.Lq ? q : -1;   // only if the shapes do not add up to S (caller error): drop, never write out of range
}

// ---- tiles of the finest level (encoder form; shared by the dense-window kernel and the per-row kernel) -------
// The dense-window backward (msda_dense.cuh) takes the queries of the FINEST level (the one with the most pixels:
// its tiles look into windows of their own size in their own level and smaller ones in every coarser level) and
// cuts that level's query grid into nty x ntx balanced tiles of at most kTileQueries queries; the tile height is
// chosen to minimise tiles x window area (halo of kTileHalo pixels around the tile: offsets of a few pixels).
// The per-row kernel processes every other query, and the tiles the dense kernel leaves (done flags).
constexpr int kTileQueries = 128;
constexpr int kTileHalo = 12;
struct FineTiles {
    int level;          // -1: no split (tile table does not fit the flag array, or not the encoder form)
    int nty, ntx, total;
    int q_begin, q_end; // the level's queries
};

// All threads call (blockDim.x >= 16); ends with __syncthreads().  scratch: one int of shared memory.
__device__ __forceinline__ void build_fine_tiles(FineTiles& ft, int* scratch, const LevelTable& t, int L, int tiles_bound)
{
    int f = 0;
    for (int l = 1; l < L; ++l)
        if (t.H[l] * t.W[l] > t.H[f] * t.W[f]) f = l;
    const int H = t.H[f], W = t.W[f];
    if (threadIdx.x == 0) *scratch = 0x7fffffff;
    __syncthreads();
    if (threadIdx.x < 16) {
        const int cand = threadIdx.x;
        const int th = min(3 + cand, H), tw = min(kTileQueries / th, W);
        const long long count = (long long)((H + th - 1) / th) * ((W + tw - 1) / tw);
        const long long cost = count * (th + kTileHalo) * (tw + kTileHalo);
        const int score = (int)min(cost >> 4, (long long)((1 << 27) - 1));
        atomicMin(scratch, score * 16 + cand);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const int c = *scratch & 15;
        const int th = min(3 + c, H), tw = min(kTileQueries / th, W);
        ft.nty = (H + th - 1) / th;
        ft.ntx = (W + tw - 1) / tw;
        ft.total = ft.nty * ft.ntx;
        ft.level = ft.total <= tiles_bound ? f : -1;
        ft.q_begin = t.qstart[f];
        ft.q_end = t.qstart[f + 1];
    }
    __syncthreads();
}

// Query of position i (row-major inside the tile) of tile `tile`; -1 beyond the tile.
struct TileRect { int y0, x0, h, w; };
__device__ __forceinline__ TileRect fine_tile_rect(const FineTiles& ft, const LevelTable& t, int tile)
{
    const int ty = tile / ft.ntx, tx = tile - ty * ft.ntx;
    const int H = t.H[ft.level], W = t.W[ft.level];
    TileRect r;
    r.y0 = (ty * H) / ft.nty;
    r.x0 = (tx * W) / ft.ntx;
    r.h = ((ty + 1) * H) / ft.nty - r.y0;
    r.w = ((tx + 1) * W) / ft.ntx - r.x0;
    return r;
}
__device__ __forceinline__ int fine_tile_query(const FineTiles& ft, const LevelTable& t, const TileRect& r, int i, int Lq)
{
    const int iy = i / r.w, ix = i - iy * r.w;
    const int q = (iy < r.h) ? ft.q_begin + (r.y0 + iy) * t.W[ft.level] + r.x0 + ix : -1;
    return q < Lq ? q : -1;
}

// ---- element conversion -------------------------------------------------------------
template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

// ---- vector loads of VEC consecutive channels into fp32 registers ---------------------
// Read-only path with L1 allocation (gathered rows are re-used by neighbouring queries).
template <typename T, int VEC> struct VecIO;

template <> struct VecIO<float, 4> {
    static __device__ __forceinline__ void load(const float* p, float (&v)[4]) {
        const float4 r = __ldg(reinterpret_cast<const float4*>(p));
        v[0] = r.x; v[1] = r.y; v[2] = r.z; v[3] = r.w;
    }
    static __device__ __forceinline__ void load_stream(const float* p, float (&v)[4]) {
        float4 r;
        asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
        v[0] = r.x; v[1] = r.y; v[2] = r.z; v[3] = r.w;
    }
    static __device__ __forceinline__ void store(float* p, const float (&v)[4]) {
        asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
                     :: "l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]) : "memory");
    }
};

__device__ __forceinline__ void unpack_bf16x2(uint32_t w, float& lo, float& hi) {
    lo = __uint_as_float(w << 16);
    hi = __uint_as_float(w & 0xffff0000u);
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&h);
}
__device__ __forceinline__ void unpack_f16x2(uint32_t w, float& lo, float& hi) {
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w));
    lo = f.x; hi = f.y;
}
__device__ __forceinline__ uint32_t pack_f16x2(float lo, float hi) {
    const __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&h);
}

template <typename T> struct Half2Codec;
template <> struct Half2Codec<__nv_bfloat16> {
    static __device__ __forceinline__ void unpack(uint32_t w, float& a, float& b) { unpack_bf16x2(w, a, b); }
    static __device__ __forceinline__ uint32_t pack(float a, float b) { return pack_bf16x2(a, b); }
};
template <> struct Half2Codec<__half> {
    static __device__ __forceinline__ void unpack(uint32_t w, float& a, float& b) { unpack_f16x2(w, a, b); }
    static __device__ __forceinline__ uint32_t pack(float a, float b) { return pack_f16x2(a, b); }
};

template <typename T> struct VecIO<T, 8> {  // 16-bit types, 16-byte vectors
    static __device__ __forceinline__ void load(const T* p, float (&v)[8]) {
        const uint4 r = __ldg(reinterpret_cast<const uint4*>(p));
        Half2Codec<T>::unpack(r.x, v[0], v[1]); Half2Codec<T>::unpack(r.y, v[2], v[3]);
        Half2Codec<T>::unpack(r.z, v[4], v[5]); Half2Codec<T>::unpack(r.w, v[6], v[7]);
    }
    static __device__ __forceinline__ void load_stream(const T* p, float (&v)[8]) {
        uint4 r;
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
        Half2Codec<T>::unpack(r.x, v[0], v[1]); Half2Codec<T>::unpack(r.y, v[2], v[3]);
        Half2Codec<T>::unpack(r.z, v[4], v[5]); Half2Codec<T>::unpack(r.w, v[6], v[7]);
    }
    static __device__ __forceinline__ void store(T* p, const float (&v)[8]) {
        const uint32_t a = Half2Codec<T>::pack(v[0], v[1]), b = Half2Codec<T>::pack(v[2], v[3]);
        const uint32_t c = Half2Codec<T>::pack(v[4], v[5]), d = Half2Codec<T>::pack(v[6], v[7]);
        asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                     :: "l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
    }
};

template <typename T> struct VecIO<T, 4> {  // 16-bit types, 8-byte vectors (D = 36)
    static __device__ __forceinline__ void load(const T* p, float (&v)[4]) {
        const uint2 r = __ldg(reinterpret_cast<const uint2*>(p));
        Half2Codec<T>::unpack(r.x, v[0], v[1]); Half2Codec<T>::unpack(r.y, v[2], v[3]);
    }
    static __device__ __forceinline__ void load_stream(const T* p, float (&v)[4]) {
        uint2 r;
        asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];"
                     : "=r"(r.x), "=r"(r.y) : "l"(p));
        Half2Codec<T>::unpack(r.x, v[0], v[1]); Half2Codec<T>::unpack(r.y, v[2], v[3]);
    }
    static __device__ __forceinline__ void store(T* p, const float (&v)[4]) {
        const uint32_t a = Half2Codec<T>::pack(v[0], v[1]), b = Half2Codec<T>::pack(v[2], v[3]);
        asm volatile("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};"
                     :: "l"(p), "r"(a), "r"(b) : "memory");
    }
};

// Vector fp32 reduction into global memory (sm_90+): one instruction adds 16 bytes.
__device__ __forceinline__ void red_add_f32x4(float* p, float a, float b, float c, float d) {
    asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Packed 16-bit reductions (sm_90+): 8 bytes = 4 channels per instruction, straight into a bf16 / f16 tensor.
template <typename T> __device__ __forceinline__ void red_add_16x4(T* p, float a, float b, float c, float d);
template <> __device__ __forceinline__ void red_add_16x4<__nv_bfloat16>(__nv_bfloat16* p, float a, float b, float c, float d) {
    asm volatile("red.relaxed.gpu.global.add.noftz.v2.bf16x2 [%0], {%1,%2};"
                 :: "l"(p), "r"(pack_bf16x2(a, b)), "r"(pack_bf16x2(c, d)) : "memory");
}
template <> __device__ __forceinline__ void red_add_16x4<__half>(__half* p, float a, float b, float c, float d) {
    asm volatile("red.relaxed.gpu.global.add.noftz.v2.f16x2 [%0], {%1,%2};"
                 :: "l"(p), "r"(pack_f16x2(a, b)), "r"(pack_f16x2(c, d)) : "memory");
}
template <> __device__ __forceinline__ void red_add_16x4<float>(float*, float, float, float, float) {}

// Streaming scalar loads of the per-sample metadata (read exactly once).
template <typename A> __device__ __forceinline__ float2 load_loc(const A* p);
template <> __device__ __forceinline__ float2 load_loc<float>(const float* p) {
    float2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f32 {%0,%1}, [%2];" : "=f"(r.x), "=f"(r.y) : "l"(p));
    return r;
}
template <> __device__ __forceinline__ float2 load_loc<__nv_bfloat16>(const __nv_bfloat16* p) {
    uint32_t w;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(w) : "l"(p));
    float2 r; unpack_bf16x2(w, r.x, r.y); return r;
}
template <> __device__ __forceinline__ float2 load_loc<__half>(const __half* p) {
    uint32_t w;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(w) : "l"(p));
    float2 r; unpack_f16x2(w, r.x, r.y); return r;
}
template <typename A> __device__ __forceinline__ float load_attn(const A* p) { return to_f32<A>(__ldg(p)); }
// The same through L1 (the dense-window kernel reads a row's samples twice: window search, then level by level).
template <typename A> __device__ __forceinline__ float2 load_loc_cached(const A* p);
template <> __device__ __forceinline__ float2 load_loc_cached<float>(const float* p) { return __ldg(reinterpret_cast<const float2*>(p)); }
template <> __device__ __forceinline__ float2 load_loc_cached<__nv_bfloat16>(const __nv_bfloat16* p) {
    float2 r; unpack_bf16x2(__ldg(reinterpret_cast<const uint32_t*>(p)), r.x, r.y); return r;
}
template <> __device__ __forceinline__ float2 load_loc_cached<__half>(const __half* p) {
    float2 r; unpack_f16x2(__ldg(reinterpret_cast<const uint32_t*>(p)), r.x, r.y); return r;
}

template <typename A> __device__ __forceinline__ void store_loc(A* p, float x, float y);
template <> __device__ __forceinline__ void store_loc<float>(float* p, float x, float y) {
    *reinterpret_cast<float2*>(p) = make_float2(x, y);
}
template <> __device__ __forceinline__ void store_loc<__nv_bfloat16>(__nv_bfloat16* p, float x, float y) {
    *reinterpret_cast<uint32_t*>(p) = pack_bf16x2(x, y);
}
template <> __device__ __forceinline__ void store_loc<__half>(__half* p, float x, float y) {
    *reinterpret_cast<uint32_t*>(p) = pack_f16x2(x, y);
}

// One bilinear tap.  Corner order: (y0,x0) (y0,x1) (y1,x0) (y1,x1).
// Offsets are ALWAYS valid addresses (coordinates clamped into the level) so the four
// corner loads are unconditional and can be issued back to back; corners outside the
// level, and samples outside the open window (-1,H)x(-1,W), get weight 0 instead
// (same result as the reference's guarded loads, .cuh:36-60).
struct Tap {
    int off[4];        // element offsets from the (batch, head) base
    float wy0, wy1;    // (1-lh), lh   masked by row validity
    float wx0, wx1;    // (1-lw), lw   masked by column validity
    float my0, my1, mx0, mx1;   // 0/1 validity masks (needed by the derivatives)
    bool ok[4];
};

__device__ __forceinline__ Tap make_tap(float lx, float ly, int H, int W, int start, int row_stride)
{
    Tap t;
    const float x = fmaf(lx, (float)W, -0.5f);
    const float y = fmaf(ly, (float)H, -0.5f);
    const bool inside = (y > -1.f) && (x > -1.f) && (y < (float)H) && (x < (float)W);
    const float fy = floorf(y), fx = floorf(x);
    const int y0 = (int)fy, x0 = (int)fx;
    const float lh = y - fy, lw = x - fx;
    const bool by0 = inside && y0 >= 0, by1 = inside && y0 + 1 <= H - 1;
    const bool bx0 = inside && x0 >= 0, bx1 = inside && x0 + 1 <= W - 1;
    t.my0 = by0 ? 1.f : 0.f; t.my1 = by1 ? 1.f : 0.f;
    t.mx0 = bx0 ? 1.f : 0.f; t.mx1 = bx1 ? 1.f : 0.f;
    t.wy0 = by0 ? 1.f - lh : 0.f; t.wy1 = by1 ? lh : 0.f;
    t.wx0 = bx0 ? 1.f - lw : 0.f; t.wx1 = bx1 ? lw : 0.f;
    t.ok[0] = by0 && bx0; t.ok[1] = by0 && bx1; t.ok[2] = by1 && bx0; t.ok[3] = by1 && bx1;
    const int y0c = min(max(y0, 0), H - 1), y1c = min(max(y0 + 1, 0), H - 1);
    const int x0c = min(max(x0, 0), W - 1), x1c = min(max(x0 + 1, 0), W - 1);
    const int r0 = (start + y0c * W) * row_stride, r1 = (start + y1c * W) * row_stride;
    const int c0 = x0c * row_stride, c1 = x1c * row_stride;
    t.off[0] = r0 + c0; t.off[1] = r0 + c1; t.off[2] = r1 + c0; t.off[3] = r1 + c1;
    return t;
}

// ---- owner-lane taps ---------------------------------------------------------------------
// Each lane of a row group prepares ONE sample of the current chunk (instead of every lane
// redoing the same coordinate arithmetic for every sample); the loop over the chunk then
// only broadcasts the finished numbers with warp shuffles.
//
// A tap is expressed on a FIXED 2x2 footprint whose top-left pixel (yb, xb) is clamped into
// the level (xb = clamp(floor(x), 0, W-2)), so the four corner addresses always are
// base, base+dx, base+dy, base+dy+dx and are always valid memory.  Border handling (the
// reference's guarded corners, .cuh:36-60, and the open window -1 < x < W, :229) moves into
// the per-column / per-row weights: a corner that the reference skips gets weight 0, and when
// floor(x) = -1 (or W-1) the surviving corner's weight is assigned to the footprint column
// that holds it.
struct Axis {
    int b;          // clamped footprint origin along this axis
    float w0, w1;   // weights of footprint column b and b+1
    float d0, d1;   // d(w0)/dcoord, d(w1)/dcoord  (one-sided at the kinks, like .cuh:141-160)
};

__device__ __forceinline__ Axis make_axis(float c, int size, bool inside)
{
    Axis ax;
    const float f = floorf(c);
    const int i0 = (int)f;
    const float frac = c - f;
    const bool v0 = inside && i0 >= 0, v1 = inside && i0 + 1 <= size - 1;
    const float a0 = v0 ? 1.f - frac : 0.f, a1 = v1 ? frac : 0.f;   // weights of pixels i0, i0+1
    const float g0 = v0 ? -1.f : 0.f, g1 = v1 ? 1.f : 0.f;
    ax.b = min(max(i0, 0), max(size - 2, 0));
    const int sh = i0 - ax.b;                  // 0 normally, -1 at the low border, +1 at the high border
    ax.w0 = sh == 0 ? a0 : (sh == -1 ? a1 : 0.f);
    ax.w1 = sh == 0 ? a1 : (sh == 1 ? a0 : 0.f);
    ax.d0 = sh == 0 ? g0 : (sh == -1 ? g1 : 0.f);
    ax.d1 = sh == 0 ? g1 : (sh == 1 ? g0 : 0.f);
    return ax;
}

struct OwnerTap {
    int off0;         // offset of the footprint's top-left corner from the (batch, head) base,
                      // in units of `row_stride` (caller's choice: elements or VEC-element vectors)
    int dy;           // step to the second footprint row (0 when H == 1)
    int dx;           // step to the second footprint column (0 when W == 1)
    Axis ay, ax;
    float a;          // attention weight (0 for lanes without a sample)
};

__device__ __forceinline__ OwnerTap make_owner_tap(bool mine, float lx, float ly, float a, int H, int W,
                                                   int start, int row_stride)
{
    OwnerTap t;
    const float x = fmaf(lx, (float)W, -0.5f);
    const float y = fmaf(ly, (float)H, -0.5f);
    const bool inside = mine && (y > -1.f) && (x > -1.f) && (y < (float)H) && (x < (float)W);
    t.ay = make_axis(y, H, inside);
    t.ax = make_axis(x, W, inside);
    t.a = mine ? a : 0.f;
    t.off0 = mine ? (start + t.ay.b * W + t.ax.b) * row_stride : 0;
    t.dy = (mine && H > 1) ? W * row_stride : 0;
    t.dx = (mine && W > 1) ? row_stride : 0;
    return t;
}

// ---- where a row's samples come from ----------------------------------------------------------
// Plain entry points read materialised sampling locations and attention weights.  The fused entry
// points (msda_b200_forward_fused / _backward_fused) read what the module's Linear layers produce --
// raw sampling offsets and attention logits -- plus the reference points, and do the module's
// prologue (reference ops/modules/ms_deform_attn.py:69-82) in the owner lanes:
//   attention = softmax over the row's L*P logits, zeroed for masked queries (:71-74),
//   location  = ref + offset / norm[level]                         (2-d reference points, :76-79)
//             = ref_xy + offset / P * ref_wh * 0.5                 (4-d reference boxes,  :80-82)
struct FusedArgs {
    const float* ref;             // [N][Lq][L][ref_dim]
    const float* norm;            // [L][2] divisors of the (x, y) offsets (ref_dim == 2 only)
    const unsigned char* qmask;   // [N][Lq], nonzero = query masked; may be nullptr
    float* grad_loc;              // backward only, optional: d out / d location, [N][Lq][M][L][P][2] fp32
    int ref_dim;                  // 2 or 4
};

struct MaxOp { static __device__ __forceinline__ float f(float a, float b) { return fmaxf(a, b); } };
struct SumOp { static __device__ __forceinline__ float f(float a, float b) { return a + b; } };

// Reduction over the GS lanes of a row group; every lane of the group gets the result.
template <int GS, typename Op>
__device__ __forceinline__ float group_reduce(float v, int gl, int grp_base)
{
    if constexpr ((GS & (GS - 1)) == 0) {
#pragma unroll
        for (int o = GS / 2; o > 0; o >>= 1) v = Op::f(v, __shfl_xor_sync(0xffffffffu, v, o));
        return v;
    } else {
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
            const float t = __shfl_down_sync(0xffffffffu, v, o);
            if (gl < o && gl + o < GS) v = Op::f(v, t);
        }
        return __shfl_sync(0xffffffffu, v, grp_base);
    }
}

// Samples of a row are prepared CH at a time by the "owner" lanes gl < CH of its group, CH = the
// largest power of two <= GS (GS = 9 for D = 36: lane 8 only carries channels).  A power-of-two chunk
// divides the usual L*P (16, 32) without a ragged tail and keeps the backward's reduce-scatter a butterfly.
__host__ __device__ constexpr int chunk_of(int gs)
{
    int c = 1;
    while (c * 2 <= gs) c *= 2;
    return c;
}

// Tap records live in shared memory as [sample j][group], `record_stride(GS)` records per sample.
// The owner lanes of a warp write their records at once (lane (grp, gl) -> record (gl, grp)); with the
// natural stride 32/GS the eight 16-byte records of a quarter warp fall into one or two 16-byte bank
// groups (ncu: 21 data-pipe wavefronts per chunk for the two stores instead of 6).  A stride congruent
// to max(1, 8/GS) modulo 8 spreads them over all eight bank groups and keeps the broadcast reads of a
// step (one record per group, consecutive records) conflict-free.
__host__ __device__ constexpr int record_stride(int gs)
{
    const int gpw = 32 / gs;
    const int b = 8 / gs > 1 ? 8 / gs : 1;
    if (b == 1) return gpw | 1;
    int s = gpw;
    while (s % 8 != b) ++s;
    return s;
}
__host__ __device__ constexpr int record_count(int gs) { return chunk_of(gs) * record_stride(gs); }

template <typename AT, bool FUSED, int PRE> struct SampleSrc;

// PRE: chunks of a row whose location / weight the owner lane loads back to back when the row starts;
// the chunk loop then finds them in registers instead of waiting for a dependent global load (latency
// chain per chunk: load -> tap -> shared memory -> gathers).  3 registers per chunk: the backward
// (128 registers) uses PRE = 4 (C2-enc bf16 495 -> 474 us, C3-enc 392 -> 364; fp32 unchanged within
// noise), the forward, compiled for 64 registers, PRE = 0 (with 4 it spills: 126 -> 132 us).
template <typename AT, int PRE> struct SampleSrc<AT, false, PRE> {
    static constexpr int kPrefetchChunks = PRE;
    const AT* lrow;
    const AT* arow;
    float2 pxy[PRE > 0 ? PRE : 1];
    float pa[PRE > 0 ? PRE : 1];
    template <int GS, int CH>
    __device__ __forceinline__ void begin(const AT* loc, const AT* attn, const FusedArgs&, const KernelDims& d, size_t row,
                                          size_t, bool row_ok, int gl, int)
    {
        const int LP = d.L * d.P;
        lrow = loc + row * LP * 2;
        arow = attn + row * LP;
#pragma unroll
        for (int c = 0; c < kPrefetchChunks; ++c) {
            const int s = c * CH + gl;
            pxy[c] = make_float2(0.f, 0.f);
            pa[c] = 0.f;
            if (row_ok && gl < CH && s < LP) {
                pxy[c] = load_loc<AT>(lrow + s * 2);
                pa[c] = load_attn<AT>(arow + s);
            }
        }
    }
    __device__ __forceinline__ void get(int s, int c, int, bool mine, const KernelDims&, float& lx, float& ly, float& a) const
    {
        lx = 0.f; ly = 0.f; a = 0.f;
        if (PRE > 0 && c < kPrefetchChunks) {
            float2 xy = pxy[0];
            float w = pa[0];
#pragma unroll
            for (int k = 1; k < kPrefetchChunks; ++k) {
                xy.x = (c == k) ? pxy[k].x : xy.x;
                xy.y = (c == k) ? pxy[k].y : xy.y;
                w = (c == k) ? pa[k] : w;
            }
            if (mine) { lx = xy.x; ly = xy.y; a = w; }
        } else if (mine) {
            const float2 xy = load_loc<AT>(lrow + s * 2);
            lx = xy.x; ly = xy.y;
            a = load_attn<AT>(arow + s);
        }
    }
};

constexpr int kFusedMaxChunks = 4;   // fused: a lane keeps the softmax numerators of its <= 4 samples of a row

template <typename AT, int PRE> struct SampleSrc<AT, true, PRE> {
    const AT* orow;      // raw offsets of the row
    const float* rrow;   // reference points of the query, [L][ref_dim]
    const float* norm;
    float e[kFusedMaxChunks];   // exp(logit - max) of this lane's samples gl, gl + CH, ... of the row (gl < CH)
    float inv_sum;       // 1 / sum of the row's numerators
    float sx, sy;        // d location / d offset of the last get() (backward)
    float2 po[PRE > 0 ? PRE : 1];   // raw offsets of the first PRE chunks, loaded when the row starts (see above)
    bool masked;
    int ref_dim;

    // One pass over the row's logits (launcher guarantees L*P <= kFusedMaxChunks * CH).
    template <int GS, int CH>
    __device__ __forceinline__ void begin(const AT* off, const AT* logit, const FusedArgs& fa, const KernelDims& d,
                                          size_t row, size_t rowq, bool row_ok, int gl, int grp_base)
    {
        const int LP = d.L * d.P;
        orow = off + row * LP * 2;
        const AT* grow = logit + row * LP;
        ref_dim = fa.ref_dim;
        rrow = fa.ref + rowq * d.L * ref_dim;
        norm = fa.norm;
        masked = fa.qmask != nullptr && row_ok && fa.qmask[rowq] != 0;
        float m_ = -INFINITY;
#pragma unroll
        for (int c = 0; c < kFusedMaxChunks; ++c) {
            const int s = c * CH + gl;
            e[c] = (row_ok && gl < CH && s < LP) ? load_attn<AT>(grow + s) : -INFINITY;
            m_ = fmaxf(m_, e[c]);
        }
#pragma unroll
        for (int c = 0; c < PRE; ++c) {
            const int s = c * CH + gl;
            po[c] = (row_ok && gl < CH && s < LP) ? load_loc<AT>(orow + s * 2) : make_float2(0.f, 0.f);
        }
        const float mx = group_reduce<GS, MaxOp>(m_, gl, grp_base);
        float sum_ = 0.f;
#pragma unroll
        for (int c = 0; c < kFusedMaxChunks; ++c) {
            e[c] = row_ok ? expf(e[c] - mx) : 0.f;      // exp(-inf) = 0 for the slots without a sample
            sum_ += e[c];
        }
        const float sum = group_reduce<GS, SumOp>(sum_, gl, grp_base);
        inv_sum = row_ok ? 1.f / sum : 0.f;
    }
    // softmax probability of this lane's sample of chunk c
    __device__ __forceinline__ float prob(int c) const
    {
        float v = e[0];
#pragma unroll
        for (int k = 1; k < kFusedMaxChunks; ++k) v = (c == k) ? e[k] : v;
        return v * inv_sum;
    }
    __device__ __forceinline__ void get(int s, int c, int ls, bool mine, const KernelDims& d, float& lx, float& ly,
                                        float& a)
    {
        lx = 0.f; ly = 0.f; a = 0.f; sx = 0.f; sy = 0.f;
        if (mine) {
            float2 o;
            if (PRE > 0 && c < PRE) {
                o = po[0];
#pragma unroll
                for (int k = 1; k < PRE; ++k) {
                    o.x = (c == k) ? po[k].x : o.x;
                    o.y = (c == k) ? po[k].y : o.y;
                }
            } else {
                o = load_loc<AT>(orow + s * 2);
            }
            if (ref_dim == 2) {
                const float2 r = *reinterpret_cast<const float2*>(rrow + ls * 2);
                const float2 n = *reinterpret_cast<const float2*>(norm + ls * 2);
                // one reciprocal (MUFU.RCP, <= 1 ulp) instead of two IEEE divisions per sample: the divisor is
                // a level size, the offset term moves by < 1e-6 px against the reference's offset / size
                sx = __frcp_rn(n.x); sy = __frcp_rn(n.y);
                lx = fmaf(o.x, sx, r.x);
                ly = fmaf(o.y, sy, r.y);
            } else {
                const float4 r = *reinterpret_cast<const float4*>(rrow + ls * 4);
                const float hp = 0.5f * __frcp_rn((float)d.P);
                sx = r.z * hp; sy = r.w * hp;
                lx = fmaf(o.x, sx, r.x);
                ly = fmaf(o.y, sy, r.y);
            }
            a = masked ? 0.f : prob(c);
        }
    }
};

// acc[0..VEC) += w * v[0..VEC) with packed fp32x2 FMAs (sm_100 FFMA2).
template <int VEC>
__device__ __forceinline__ void axpy_vec(float (&acc)[VEC], float w, const float (&v)[VEC])
{
    const float2 ww = make_float2(w, w);
#pragma unroll
    for (int k = 0; k < VEC; k += 2) {
        const float2 r = __ffma2_rn(ww, make_float2(v[k], v[k + 1]), make_float2(acc[k], acc[k + 1]));
        acc[k] = r.x; acc[k + 1] = r.y;
    }
}

}  // namespace msda
