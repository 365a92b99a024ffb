// msda_dense.cuh -- encoder-form MSDeformAttn backward on the sm_100a tensor cores (tcgen05 + TMEM).
//
// Replaces, for the encoder form (the queries are the pyramid's pixels, Lq == S) and 16-bit values, the
// reference's col2im kernels (ms_deform_im2col_cuda.cuh:239-378: one global atomicAdd per channel and
// corner) and this library's own one-reduction-per-corner-row kernel (msda_bwd.cuh).  Nothing here follows
// the reference's code; the semantics (pixel-centre shift, open window, zero padding, one-sided
// derivatives at the kinks: .cuh:24-163, 227-229, 373-374) are those of make_axis (msda_common.cuh).
//
// Why.  In the encoder a tile of neighbouring queries samples a compact window of every level (reference
// point = the query's own position, offsets of a few pixels).  Per (batch, head) the four levels of the
// 800 x 1333 pyramid receive 21 / 85 / 339 / 1302 corner rows per pixel row: sending each of them to L2 as
// its own 128-byte reduction is what bounds msda_bwd.cuh (2.56 GB of reductions into a 91 MB accumulator).
// Here the per-corner work is O(1) scalar work and the fan-out over the D channels is done by two dense
// tile products per 128-pixel window chunk:
//
//   item      = (tile of <= 128 queries of one level, batch, head); one thread owns one query row
//   window_l  = bounding box of the item's 2x2 footprints in level l (found per item, data driven)
//   chunk     = 128 consecutive pixels of a window (row-major inside the window)
//   U  [q][pix] = G[q][:] . V[pix][:]     (128 x 128 x D,  tcgen05.mma, bf16/f16 x same -> fp32 in TMEM)
//       -> <grad_out, corner> for every (query, window pixel); the owner thread picks its <= 16 corners
//          of the level and forms grad_attn / grad_loc with its own fp32 bilinear weights
//   GV [pix][:] = sum_q W[pix][q] G[q][:] (128 x D x 128)  with W[pix][q] = bilinear x attention weight,
//       densified in shared memory by the owner threads (column q of W belongs to thread q: no atomics)
//       -> the item's whole contribution to the chunk's grad_value rows, flushed with ONE
//          red.global.add.v4.f32 row per touched pixel instead of one per (query, sample, corner).
//
// Numerics: G and V are exact in the operand type; W is rounded to the operand type (bf16: 2^-9 relative per
// weight, below the final rounding of the 16-bit grad_value), all accumulation is fp32.  grad_attn and
// grad_loc use fp32 weights on exact fp32 dot products.
//
// Division of labour with msda_bwd.cuh (launched right after this kernel, same stream, split mode): this kernel
// takes the tiles of the FINEST level's queries -- their windows are compact in every level -- and flags the
// (tile, batch, head) items it has done; items whose windows would still need more than kMaxItemChunks chunks (arbitrary
// sampling locations) are left unflagged; the per-row kernel processes the unflagged tiles and all queries of the coarser levels (a tile of a
// coarse level looks into windows 2x, 4x, ... its own size in the finer levels with ~1 corner per pixel: nothing to
// aggregate).  The result never depends on locality, only the speed does.
#pragma once
#include "msda_bwd.cuh"

namespace msda {
namespace dense {

// Development aid (-DMSDA_DENSE_TIMING): thread 0 of every CTA accumulates the cycles between phase marks and adds
// them to dense_phase_cycles[] at kernel end (read back with msda_b200_debug_dense_phase_cycles).
#ifdef MSDA_DENSE_TIMING
__device__ unsigned long long dense_phase_cycles[16];
#define DT_DECL long long dt_last = clock64(); unsigned long long dt_acc[16] = {0};
#define DT_MARK(i) { if (threadIdx.x == 0) { const long long t_ = clock64(); dt_acc[i] += (unsigned long long)(t_ - dt_last); dt_last = t_; } }
#define DT_FLUSH { if (threadIdx.x == 0) for (int i_ = 0; i_ < 16; ++i_) atomicAdd(&dense_phase_cycles[i_], dt_acc[i_]); }
#else
#define DT_DECL
#define DT_MARK(i)
#define DT_FLUSH
#endif

constexpr int kThreads = 128;        // = query rows per item = TMEM lanes
constexpr int kPix = 128;            // window pixels per chunk (MMA M of the scatter product, N of the dot product)
constexpr int kMaxItemChunks = 32;   // beyond this the per-row path is cheaper
constexpr int kMaxP = 4;             // samples per level a thread keeps in registers
constexpr int kTmemCols = 256;       // U: 128 columns, GV: D columns (<= 64)

// ---- tcgen05 / mbarrier wrappers (PTX ISA 8.6+, sm_100a) ---------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K-major, SWIZZLE_NONE matrix descriptor: core matrix = 8 rows x 16 bytes (rows 16 bytes apart);
// lbo = byte distance between the two 16-byte K chunks of one MMA, sbo = between 8-row groups.
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo)
{
    return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) |
           ((uint64_t)((sbo >> 4) & 0x3fff) << 32) | ((uint64_t)1 << 46);
}
// instruction descriptor: fp32 accumulate, A and B K-major, fmt 0 = f16, 1 = bf16
__host__ __device__ constexpr uint32_t umma_idesc(int M, int N, int fmt)
{
    return (1u << 4) | ((uint32_t)fmt << 7) | ((uint32_t)fmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, bool accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"((uint32_t)accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t mbar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(mbar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t mbar, uint32_t parity)
{
    asm volatile("{\n\t.reg .pred P1;\n\tWAIT_%=:\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
                 "@P1 bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}"
                 :: "r"(mbar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tmem_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void proxy_fence() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// 32 consecutive fp32 columns of this thread's TMEM lane (warp w reads lanes 32w .. 32w+31)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32])
{
    uint32_t r[32];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32"
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15,"
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                   "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
}

// 4 consecutive fp32 columns (the tail of a D = 36 accumulator row)
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, float (&v)[4])
{
    uint32_t r[4];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = __uint_as_float(r[j]);
}

// ---- operand type helpers --------------------------------------------------------------------------------------
template <typename T> struct Op16;
template <> struct Op16<__nv_bfloat16> {
    static constexpr int fmt = 1;
    static __device__ __forceinline__ float up(uint16_t h) { return __uint_as_float((uint32_t)h << 16); }
    static __device__ __forceinline__ uint16_t down(float f) { return __bfloat16_as_ushort(__float2bfloat16_rn(f)); }
};
template <> struct Op16<__half> {
    static constexpr int fmt = 0;
    static __device__ __forceinline__ float up(uint16_t h) { return __half2float(__ushort_as_half(h)); }
    static __device__ __forceinline__ uint16_t down(float f) { return __half_as_ushort(__float2half_rn(f)); }
};

// ---- shared-memory plan --------------------------------------------------------------------------------------
// All MMA operands are K-major SWIZZLE_NONE; strides are padded by 16 bytes where the owner threads' 2-byte /
// 4-byte accesses would otherwise fall into the same banks.
template <int D> struct Plan {
    static_assert(D % 4 == 0 && D >= 16 && D <= 64, "channels per head: rows of whole 8-byte pieces, <= 64 accumulator columns");
    // The tensor core wants K (dot product) and N (scatter product) in steps of 16: D = 36 is padded to DP = 48 -- in shared
    // memory only (the pad chunks of G and V are zeroed once and never written: they contribute exact zeros).
    static constexpr int DP = (D + 15) / 16 * 16;
    static constexpr int KCH = DP / 8;                      // 16-byte chunks of a (padded) channel row
    // A channel row moves between global and shared memory in pieces of 16 bytes when rows are 16-byte aligned (D % 8 == 0),
    // else of 8 bytes (D = 36: 72-byte rows); PPR pieces per row, RPI rows per warp instruction, NIT instructions per 32 rows
    static constexpr int PIECE = (D % 8 == 0) ? 16 : 8;
    static constexpr int PPR = D * 2 / PIECE, RPI = 32 / PPR, NIT = (32 + RPI - 1) / RPI;
    // G [128 q][DP] and V [128 pix][DP]: rows of DP channels
    // (chunk stride 160 / 144 bytes: the lanes that store one row's chunks hit different banks)
    static constexpr uint32_t G_LBO = KCH == 4 ? 160 : 144, G_SBO = KCH * G_LBO, G_BYTES = 16 * G_SBO;
    // byte offset of piece c of row r inside a G / V image
    static __device__ __forceinline__ uint32_t piece_off(int r, int c)
    {
        return (uint32_t)((r & 7) * 16 + (r >> 3) * G_SBO + ((c * PIECE) >> 4) * G_LBO + ((c * PIECE) & 15));
    }
    // W [128 pix][128 q]
    static constexpr uint32_t W_LBO = 144, W_SBO = 16 * 144 + 16, W_BYTES = 16 * W_SBO;
    // (the scatter product's B operand is G itself, read MN-major: no transposed copy)
    static constexpr int STAGE_STRIDE = 68;                 // floats per thread-private staging row (64 + 4: conflict-free 16-byte stores)
    static constexpr uint32_t STAGE_BYTES = kThreads * STAGE_STRIDE * 4;
    // (two V buffers: the next chunk's value rows are fetched with cp.async while the current chunk is processed)
    static constexpr uint32_t OFF_G = 0, OFF_V = OFF_G + G_BYTES, OFF_W = OFF_V + 2 * G_BYTES,
                              OFF_STAGE = OFF_W + W_BYTES, TOTAL = OFF_STAGE + STAGE_BYTES;
};

// The P samples of level l of one (query, head) row: sampling locations + attention weights in, their gradients out.
// A row's samples of one level are contiguous (P * 2 and P elements): with P == 4 they move as 16-byte vectors --
// neighbouring lanes own rows that are M * L * P elements apart, so every access of a lane is its own L1 wavefront
// and the instruction count is what matters.
template <typename AT> struct LevelIO;
template <> struct LevelIO<float> {
    static __device__ __forceinline__ void load(const float* lrow, const float* arow, int l, int P, bool have,
                                                float2 (&xy)[kMaxP], float (&a)[kMaxP])
    {
#pragma unroll
        for (int p = 0; p < kMaxP; ++p) { xy[p] = make_float2(-8.f, -8.f); a[p] = 0.f; }
        if (!have) return;
        if (P == 4) {
            const float4 u = __ldg(reinterpret_cast<const float4*>(lrow + l * 8));
            const float4 v = __ldg(reinterpret_cast<const float4*>(lrow + l * 8 + 4));
            const float4 w = __ldg(reinterpret_cast<const float4*>(arow + l * 4));
            xy[0] = make_float2(u.x, u.y); xy[1] = make_float2(u.z, u.w);
            xy[2] = make_float2(v.x, v.y); xy[3] = make_float2(v.z, v.w);
            a[0] = w.x; a[1] = w.y; a[2] = w.z; a[3] = w.w;
        } else {
#pragma unroll
            for (int p = 0; p < kMaxP; ++p)
                if (p < P) { xy[p] = load_loc_cached<float>(lrow + (l * P + p) * 2); a[p] = load_attn<float>(arow + l * P + p); }
        }
    }
    // plain (coherent) loads: the softmax backward re-reads what this thread stored
    static __device__ __forceinline__ void load_attn_only(const float* arow, int l, int P, bool have, float (&a)[kMaxP])
    {
#pragma unroll
        for (int p = 0; p < kMaxP; ++p) a[p] = 0.f;
        if (!have) return;
        if (P == 4) {
            const float4 w = *reinterpret_cast<const float4*>(arow + l * 4);
            a[0] = w.x; a[1] = w.y; a[2] = w.z; a[3] = w.w;
        } else {
#pragma unroll
            for (int p = 0; p < kMaxP; ++p) if (p < P) a[p] = arow[l * P + p];
        }
    }
    static __device__ __forceinline__ void store_attn_only(float* garow, int l, int P, const float (&ga)[kMaxP])
    {
        if (P == 4) *reinterpret_cast<float4*>(garow + l * 4) = make_float4(ga[0], ga[1], ga[2], ga[3]);
        else {
#pragma unroll
            for (int p = 0; p < kMaxP; ++p) if (p < P) garow[l * P + p] = ga[p];
        }
    }
    static __device__ __forceinline__ void store_loc_only(float* glrow, int l, int P, const float (&gx)[kMaxP],
                                                          const float (&gy)[kMaxP])
    {
        if (P == 4) {
            *reinterpret_cast<float4*>(glrow + l * 8) = make_float4(gx[0], gy[0], gx[1], gy[1]);
            *reinterpret_cast<float4*>(glrow + l * 8 + 4) = make_float4(gx[2], gy[2], gx[3], gy[3]);
        } else {
#pragma unroll
            for (int p = 0; p < kMaxP; ++p) if (p < P) store_loc<float>(glrow + (l * P + p) * 2, gx[p], gy[p]);
        }
    }
    static __device__ __forceinline__ void store(float* glrow, float* garow, int l, int P, const float (&gx)[kMaxP],
                                                 const float (&gy)[kMaxP], const float (&ga)[kMaxP])
    {
        if (P == 4) {
            *reinterpret_cast<float4*>(glrow + l * 8) = make_float4(gx[0], gy[0], gx[1], gy[1]);
            *reinterpret_cast<float4*>(glrow + l * 8 + 4) = make_float4(gx[2], gy[2], gx[3], gy[3]);
            *reinterpret_cast<float4*>(garow + l * 4) = make_float4(ga[0], ga[1], ga[2], ga[3]);
        } else {
#pragma unroll
            for (int p = 0; p < kMaxP; ++p)
                if (p < P) { store_loc<float>(glrow + (l * P + p) * 2, gx[p], gy[p]); garow[l * P + p] = ga[p]; }
        }
    }
};
template <typename H> struct LevelIO16 {
    static __device__ __forceinline__ void load(const H* lrow, const H* arow, int l, int P, bool have,
                                                float2 (&xy)[kMaxP], float (&a)[kMaxP])
    {
#pragma unroll
        for (int p = 0; p < kMaxP; ++p) { xy[p] = make_float2(-8.f, -8.f); a[p] = 0.f; }
        if (!have) return;
        if (P == 4) {
            const uint4 u = __ldg(reinterpret_cast<const uint4*>(lrow + l * 8));
            const uint2 w = __ldg(reinterpret_cast<const uint2*>(arow + l * 4));
            Half2Codec<H>::unpack(u.x, xy[0].x, xy[0].y); Half2Codec<H>::unpack(u.y, xy[1].x, xy[1].y);
            Half2Codec<H>::unpack(u.z, xy[2].x, xy[2].y); Half2Codec<H>::unpack(u.w, xy[3].x, xy[3].y);
            Half2Codec<H>::unpack(w.x, a[0], a[1]); Half2Codec<H>::unpack(w.y, a[2], a[3]);
        } else {
#pragma unroll
            for (int p = 0; p < kMaxP; ++p)
                if (p < P) { xy[p] = load_loc_cached<H>(lrow + (l * P + p) * 2); a[p] = load_attn<H>(arow + l * P + p); }
        }
    }
    static __device__ __forceinline__ void load_attn_only(const H* arow, int l, int P, bool have, float (&a)[kMaxP])
    {
#pragma unroll
        for (int p = 0; p < kMaxP; ++p) a[p] = 0.f;
        if (!have) return;
        if (P == 4) {
            const uint2 w = *reinterpret_cast<const uint2*>(arow + l * 4);
            Half2Codec<H>::unpack(w.x, a[0], a[1]); Half2Codec<H>::unpack(w.y, a[2], a[3]);
        } else {
#pragma unroll
            for (int p = 0; p < kMaxP; ++p) if (p < P) a[p] = to_f32<H>(arow[l * P + p]);
        }
    }
    static __device__ __forceinline__ void store_attn_only(H* garow, int l, int P, const float (&ga)[kMaxP])
    {
        if (P == 4) *reinterpret_cast<uint2*>(garow + l * 4) = make_uint2(Half2Codec<H>::pack(ga[0], ga[1]), Half2Codec<H>::pack(ga[2], ga[3]));
        else {
#pragma unroll
            for (int p = 0; p < kMaxP; ++p) if (p < P) garow[l * P + p] = from_f32<H>(ga[p]);
        }
    }
    static __device__ __forceinline__ void store(H* glrow, H* garow, int l, int P, const float (&gx)[kMaxP],
                                                 const float (&gy)[kMaxP], const float (&ga)[kMaxP])
    {
        if (P == 4) {
            *reinterpret_cast<uint4*>(glrow + l * 8) = make_uint4(Half2Codec<H>::pack(gx[0], gy[0]), Half2Codec<H>::pack(gx[1], gy[1]),
                                                                  Half2Codec<H>::pack(gx[2], gy[2]), Half2Codec<H>::pack(gx[3], gy[3]));
            *reinterpret_cast<uint2*>(garow + l * 4) = make_uint2(Half2Codec<H>::pack(ga[0], ga[1]), Half2Codec<H>::pack(ga[2], ga[3]));
        } else {
#pragma unroll
            for (int p = 0; p < kMaxP; ++p)
                if (p < P) { store_loc<H>(glrow + (l * P + p) * 2, gx[p], gy[p]); garow[l * P + p] = from_f32<H>(ga[p]); }
        }
    }
};
template <> struct LevelIO<__nv_bfloat16> : LevelIO16<__nv_bfloat16> {};
template <> struct LevelIO<__half> : LevelIO16<__half> {};

struct DenseArgs {
    int* claim;                    // zero-initialised work counter (in the caller's workspace)
    uint8_t* done;                 // zero-initialised, [tiles_bound * N * M]: 1 = item done here; the per-row kernel does the rest
    int tiles_bound;
};

template <typename VT, typename AT, int D, bool FUSED>
__global__ void __launch_bounds__(kThreads, 2)
msda_backward_dense_kernel(const VT* __restrict__ value, const int64_t* __restrict__ shapes,
                           const int64_t* __restrict__ level_start, const AT* __restrict__ loc,
                           const AT* __restrict__ attn, const VT* __restrict__ grad_out,
                           float* __restrict__ grad_value_acc, AT* __restrict__ grad_loc,
                           AT* __restrict__ grad_attn, const KernelDims d, const FusedArgs fa, const DenseArgs da)
{
    static_assert(D % 32 == 0 || D % 32 == 4, "accumulator rows are read 32 columns at a time (+ one group of 4)");
    using P = Plan<D>;
    using O = Op16<VT>;
    constexpr int NW = kThreads / 32;
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ LevelTable lt;
    __shared__ FineTiles ft;
    __shared__ int s_scratch;
    __shared__ int s_box[NW][kMaxLevels][4];
    __shared__ int s_win[kMaxLevels][5];       // ymin, xmin, wh, ww, chunks
    __shared__ int s_item[2];
    __shared__ int s_bail[2];                  // decided when the item in the same slot was claimed: stop (see below)
    __shared__ uint32_t s_tmem;
    __shared__ __align__(8) uint64_t s_mbar[2];   // dot products ready / scatter product ready

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // The per-row kernel behind this one is its programmatic dependent (launch_backward_vec_impl): release it now, so its
    // CTAs move in as soon as CTAs of this grid retire.  It waits for this grid before it reads the done-item flags.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    stage_levels(lt, shapes, level_start, d);
    build_fine_tiles(ft, &s_scratch, lt, d.L, da.tiles_bound);

    uint8_t* sG = smem + P::OFF_G;
    uint8_t* sV = smem + P::OFF_V;
    uint8_t* sW = smem + P::OFF_W;
    float* stage = reinterpret_cast<float*>(smem + P::OFF_STAGE) + tid * P::STAGE_STRIDE;   // this thread's row

    // one-time setup: zero W (the owner threads restore the zeros after every chunk), TMEM, barrier
    for (uint32_t i = tid * 16; i < P::W_BYTES; i += kThreads * 16) *reinterpret_cast<uint4*>(sW + i) = make_uint4(0, 0, 0, 0);
    if constexpr (P::DP != D)      // padded channels of G and of both V buffers: zero for good
        for (uint32_t i = tid * 16; i < 3 * P::G_BYTES; i += kThreads * 16) *reinterpret_cast<uint4*>(sG + i) = make_uint4(0, 0, 0, 0);
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(&s_tmem)), "r"((uint32_t)kTmemCols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(&s_mbar[0])), "r"(1u));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(&s_mbar[1])), "r"(1u));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        s_item[0] = atomicAdd(da.claim, 1);
        s_bail[0] = 0;
    }
    tmem_fence_before();
    __syncthreads();
    tmem_fence_after();
    const uint32_t tmem_u = s_tmem, tmem_gv = s_tmem + kPix;
    const uint32_t lane_sel = (uint32_t)(warp * 32) << 16;
    const uint32_t mbar = smem_u32(&s_mbar[0]), mbar2 = smem_u32(&s_mbar[1]);
    uint32_t phase = 0;
    const int LP = d.L * d.P;
    const int items = ft.level >= 0 ? ft.total * d.N * d.M : 0;
    constexpr uint32_t idesc_u = umma_idesc(128, kPix, O::fmt);
    // B = G as staged for the dot product, [K = 128 query rows][N = D channels]: MN-major (instruction descriptor bit 16);
    // for an MN-major SWIZZLE_NONE operand the descriptor's LBO is the stride between 8-row groups along K and its SBO
    // the stride between 16-byte chunks along N (tools/micro/umma_probe.cu tests 8 / 9, profiles/r2_umma_probe.txt)
    constexpr uint32_t idesc_gv = umma_idesc(128, P::DP, O::fmt) | (1u << 16);

    int slot = 0;
    DT_DECL
    for (int item = s_item[0]; item < items; item = s_item[slot]) {
        DT_MARK(0)
        // claim the next item now; it is read after the barrier that ends this one
        if (tid == 0) {
            s_item[slot ^ 1] = atomicAdd(da.claim, 1);
            // Sampling locations without locality (windows beyond kMaxItemChunks chunks everywhere): once at least 64
            // items were left to the per-row kernel and they outnumber the items done here 8 : 1, this CTA stops --
            // items nobody marks done are the per-row kernel's -- and the grid is gone after about one window search
            // per CTA, which frees its SMs (a dense CTA pair holds every register of its SM) for the per-row kernel.
            // (da.claim[2] / [3]: items done / left, zeroed with the work counter.)
            const int n_back = *reinterpret_cast<volatile int*>(da.claim + 3), n_done = *reinterpret_cast<volatile int*>(da.claim + 2);
            s_bail[slot ^ 1] = n_back >= 64 && n_back >= 8 * n_done;
        }
        const bool bail = s_bail[slot] != 0;
        slot ^= 1;
        if (bail) break;

        // ---- item -> (tile, batch, head) -> this thread's query -------------------------------------------------
        const int tile = item / (d.N * d.M);
        const int bm = item - tile * (d.N * d.M);
        const int b = bm / d.M, m = bm - b * d.M;
        const TileRect rect = fine_tile_rect(ft, lt, tile);
        const int q = fine_tile_query(ft, lt, rect, tid, d.Lq);
        const bool row_ok = q >= 0;
        const size_t rowq = (size_t)b * d.Lq + (row_ok ? q : 0);
        const size_t row = rowq * d.M + m;
        const AT* lrow = loc + row * LP * 2;       // sampling locations (FUSED: raw sampling offsets)
        const AT* arow = attn + row * LP;          // attention weights  (FUSED: attention logits)
        // FUSED: the module prologue (reference ops/modules/ms_deform_attn.py:69-82) in the owner thread:
        //   attention = softmax over the row's L*P logits, zero for masked queries; location = ref + offset * scale
        float sm_max = 0.f, sm_inv = 0.f;
        bool masked = false;
        const float* rrow = nullptr;
        if constexpr (FUSED) {
            rrow = fa.ref + rowq * d.L * fa.ref_dim;
            masked = fa.qmask != nullptr && row_ok && fa.qmask[rowq] != 0;
            float mx = -INFINITY;
            for (int l = 0; l < d.L; ++l) {
                float lg[kMaxP];
                LevelIO<AT>::load_attn_only(arow, l, d.P, row_ok, lg);
#pragma unroll
                for (int p = 0; p < kMaxP; ++p) if (p < d.P) mx = fmaxf(mx, lg[p]);
            }
            float sum = 0.f;
            for (int l = 0; l < d.L; ++l) {
                float lg[kMaxP];
                LevelIO<AT>::load_attn_only(arow, l, d.P, row_ok, lg);
#pragma unroll
                for (int p = 0; p < kMaxP; ++p) if (p < d.P) sum += expf(lg[p] - mx);
            }
            sm_max = row_ok ? mx : 0.f;
            sm_inv = row_ok ? 1.f / sum : 0.f;
        }
        // raw (offset, logit) of a level's samples -> (location, attention weight); sx / sy = d location / d offset
        auto prologue = [&](int l, float2 (&xy)[kMaxP], float (&a)[kMaxP], float& sx, float& sy) {
            sx = sy = 1.f;
            if constexpr (FUSED) {
                float rx, ry;
                if (fa.ref_dim == 2) {
                    const float2 r = *reinterpret_cast<const float2*>(rrow + l * 2);
                    const float2 n = *reinterpret_cast<const float2*>(fa.norm + l * 2);
                    sx = __frcp_rn(n.x); sy = __frcp_rn(n.y);
                    rx = r.x; ry = r.y;
                } else {
                    const float4 r = *reinterpret_cast<const float4*>(rrow + l * 4);
                    const float hp = 0.5f * __frcp_rn((float)d.P);
                    sx = r.z * hp; sy = r.w * hp;
                    rx = r.x; ry = r.y;
                }
#pragma unroll
                for (int p = 0; p < kMaxP; ++p) {
                    const bool have = row_ok && p < d.P;
                    xy[p] = have ? make_float2(fmaf(xy[p].x, sx, rx), fmaf(xy[p].y, sy, ry)) : make_float2(-8.f, -8.f);
                    a[p] = (have && !masked) ? expf(a[p] - sm_max) * sm_inv : 0.f;
                }
            }
        };

        // samples of the level being processed next (requested one level ahead)
        float2 nxt_xy[kMaxP];
        float nxt_a[kMaxP];
        DT_MARK(1)
        // ---- windows: bounding box of the live corner pixels, per level ---------------------------------------------
        // (a corner is live when its pixel exists: the reference's guarded corner loads, .cuh:36-60, and the open
        // sample window -1 < x < W, :229)
        // the loads of four levels are issued before the first use (one memory round trip per group, not per level)
        for (int l0 = 0; l0 < d.L; l0 += 4) {
            float2 bxy[4][kMaxP];
            float baw[4][kMaxP];
#pragma unroll
            for (int j = 0; j < 4; ++j) LevelIO<AT>::load(lrow, arow, min(l0 + j, d.L - 1), d.P, row_ok && l0 + j < d.L, bxy[j], baw[j]);
            if (l0 == 0) {      // the chunk phase starts with level 0: keep its (raw) samples
#pragma unroll
                for (int p = 0; p < kMaxP; ++p) { nxt_xy[p] = bxy[0][p]; nxt_a[p] = baw[0][p]; }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int l = l0 + j;
                if (l < d.L) {
                    const int H = lt.H[l], W = lt.W[l];
                    int ymin = 1 << 30, ymax = -1, xmin = 1 << 30, xmax = -1;
                    {
                        float sx, sy;
                        prologue(l, bxy[j], baw[j], sx, sy);
                    }
#pragma unroll
                    for (int p = 0; p < kMaxP; ++p) {
                        const float x = fmaf(bxy[j][p].x, (float)W, -0.5f), y = fmaf(bxy[j][p].y, (float)H, -0.5f);
                        if ((y > -1.f) && (x > -1.f) && (y < (float)H) && (x < (float)W)) {
                            const int iy = (int)floorf(y), ix = (int)floorf(x);
                            ymin = min(ymin, max(iy, 0)); ymax = max(ymax, min(iy + 1, H - 1));
                            xmin = min(xmin, max(ix, 0)); xmax = max(xmax, min(ix + 1, W - 1));
                        }
                    }
                    ymin = __reduce_min_sync(0xffffffffu, ymin); ymax = __reduce_max_sync(0xffffffffu, ymax);
                    xmin = __reduce_min_sync(0xffffffffu, xmin); xmax = __reduce_max_sync(0xffffffffu, xmax);
                    if (lane == 0) { s_box[warp][l][0] = ymin; s_box[warp][l][1] = ymax; s_box[warp][l][2] = xmin; s_box[warp][l][3] = xmax; }
                }
            }
        }
        __syncthreads();
        if (tid < d.L) {
            const int l = tid;
            int ymin = 1 << 30, ymax = -1, xmin = 1 << 30, xmax = -1;
            for (int w = 0; w < NW; ++w) {
                ymin = min(ymin, s_box[w][l][0]); ymax = max(ymax, s_box[w][l][1]);
                xmin = min(xmin, s_box[w][l][2]); xmax = max(xmax, s_box[w][l][3]);
            }
            int wh = 0, ww = 0;
            if (ymax >= 0) {
                wh = ymax - ymin + 1;
                ww = xmax - xmin + 1;
            } else { ymin = 0; xmin = 0; }
            s_win[l][0] = ymin; s_win[l][1] = xmin; s_win[l][2] = wh; s_win[l][3] = ww;
            s_win[l][4] = (wh * ww + kPix - 1) / kPix;
        }
        __syncthreads();
        int total_chunks = 0;
        for (int l = 0; l < d.L; ++l) total_chunks += s_win[l][4];
        if (total_chunks > kMaxItemChunks) {
            // windows too large for dense chunks: the per-row kernel does this item
            if (tid == 0) atomicAdd(da.claim + 3, 1);
            __syncthreads();
            continue;
        }

        // The next item's rows (sampling locations, attention weights, grad_out) are requested into L2 now: its window
        // search and its operand staging would otherwise start with a trip to DRAM (the claim made at the top of this
        // item became visible with the barriers above).
        {
            const int nitem = s_item[slot];
            if (nitem < items) {
                const int ntile = nitem / (d.N * d.M);
                const int nbm = nitem - ntile * (d.N * d.M);
                const int nb = nbm / d.M, nm = nbm - nb * d.M;
                const TileRect nrect = fine_tile_rect(ft, lt, ntile);
                const int nq = fine_tile_query(ft, lt, nrect, tid, d.Lq);
                if (nq >= 0) {
                    const size_t nrow = ((size_t)nb * d.Lq + nq) * d.M + nm;
                    asm volatile("prefetch.global.L2 [%0];" :: "l"(loc + nrow * LP * 2));
                    if (LP * 2 * sizeof(AT) > 128) asm volatile("prefetch.global.L2 [%0];" :: "l"(reinterpret_cast<const char*>(loc + nrow * LP * 2) + 128));
                    asm volatile("prefetch.global.L2 [%0];" :: "l"(attn + nrow * LP));
                    asm volatile("prefetch.global.L2 [%0];" :: "l"(grad_out + nrow * D));
                }
            }
        }
        DT_MARK(2)
        // ---- G: the item's rows of grad_out -> operand A of the dot product (K-major) and, the same bytes read MN-major,
        // operand B of the scatter product.  D / 8 lanes per row (16 bytes each): a warp instruction reads 32 / (D / 8) whole rows.
        {
#pragma unroll
            for (int i = 0; i < P::NIT; ++i) {
                const int rl = i * P::RPI + lane / P::PPR, c = lane % P::PPR;      // row inside the warp's 32, piece of the row
                const bool act = lane / P::PPR < P::RPI && rl < 32;
                const int qq = __shfl_sync(0xffffffffu, q, rl & 31);
                const int trow = warp * 32 + rl;
                const char* src = reinterpret_cast<const char*>(grad_out + (((size_t)b * d.Lq + max(qq, 0)) * d.M + m) * D) + c * P::PIECE;
                if constexpr (P::PIECE == 16) {
                    uint4 g4 = make_uint4(0, 0, 0, 0);
                    if (act && qq >= 0) g4 = __ldg(reinterpret_cast<const uint4*>(src));
                    if (act) *reinterpret_cast<uint4*>(sG + P::piece_off(trow, c)) = g4;
                } else {
                    uint2 g2 = make_uint2(0, 0);
                    if (act && qq >= 0) g2 = __ldg(reinterpret_cast<const uint2*>(src));
                    if (act) *reinterpret_cast<uint2*>(sG + P::piece_off(trow, c)) = g2;
                }
            }
        }

        const VT* vbase = value + ((size_t)b * d.S * d.M + m) * D;
        float* gvbase = grad_value_acc + ((size_t)b * d.S * d.M + m) * D;
        const size_t pix_stride = (size_t)d.M * D;

        float sm_dot = 0.f;
        for (int l = 0; l < d.L; ++l) {
            const int nch = s_win[l][4];
            if (nch == 0) {
                // nothing of this item samples inside level l: zero gradients for its samples
                if (row_ok) {
                    const float z[kMaxP] = {0.f, 0.f, 0.f, 0.f};
                    LevelIO<AT>::store(grad_loc + row * LP * 2, grad_attn + row * LP, l, d.P, z, z, z);
                    if constexpr (FUSED) {
                        if (fa.grad_loc) LevelIO<float>::store_loc_only(fa.grad_loc + row * LP * 2, l, d.P, z, z);
                    }
                }
                if (l + 1 < d.L) {
                    LevelIO<AT>::load(lrow, arow, l + 1, d.P, row_ok, nxt_xy, nxt_a);
                }
                continue;
            }
            const int H = lt.H[l], W = lt.W[l];
            const int ymin = s_win[l][0], xmin = s_win[l][1], wh = s_win[l][2], ww = s_win[l][3];
            const int npix = wh * ww;
            // ---- this row's 4 * P corner targets of the level, everything the chunk loop needs precomputed ----
            // tr   = window-linear pixel index of the corner (kNone: the corner's pixel does not exist / no sample)
            // wa   = byte address of the target's W element inside a chunk (column `tid`)
            // tw   = bilinear x attention weight (scatter product)
            // cga  = bilinear weight: coefficient of <grad_out, corner> in grad_attn (those of d / dx, d / dy are the
            //        sample's row / column weights with a sign)
            constexpr int NT = 4 * kMaxP;
            constexpr int kNone = -(1 << 20);
            int tr[NT];
            uint32_t wa[NT];
            float tw[NT], cga[NT];
            float swy[kMaxP][2], swx[kMaxP][2];     // per-sample row / column weights (their derivatives are -1 / +1)
            float acc_a[kMaxP], acc_x[kMaxP], acc_y[kMaxP], att[kMaxP];
            const uint32_t w_col = smem_u32(sW) + (tid >> 3) * P::W_LBO + (tid & 7) * 2;
            const uint32_t st_row = smem_u32(stage);
            // the level's samples were requested while the previous level was being processed
            float lev_sx, lev_sy;
            prologue(l, nxt_xy, nxt_a, lev_sx, lev_sy);
#pragma unroll
            for (int p = 0; p < kMaxP; ++p) {
                acc_a[p] = acc_x[p] = acc_y[p] = 0.f;
                const float2 xy = nxt_xy[p];
                const float a = nxt_a[p];
                att[p] = a;
                const float x = fmaf(xy.x, (float)W, -0.5f), y = fmaf(xy.y, (float)H, -0.5f);
                const bool inside = (y > -1.f) && (x > -1.f) && (y < (float)H) && (x < (float)W);
                const float fy = floorf(y), fx = floorf(x);
                const int iy = (int)fy, ix = (int)fx;
                const float ly = y - fy, lx = x - fx;
                const bool vy[2] = {inside && iy >= 0, inside && iy + 1 <= H - 1};
                const bool vx[2] = {inside && ix >= 0, inside && ix + 1 <= W - 1};
                const float wy[2] = {1.f - ly, ly}, wx[2] = {1.f - lx, lx};
                const int r00 = (iy - ymin) * ww + (ix - xmin);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int t = p * 4 + c, cy = c >> 1, cx = c & 1;
                    const bool live = vy[cy] && vx[cx];
                    const int r = r00 + cy * ww + cx;
                    tr[t] = live ? r : kNone;
                    const int idx = r & (kPix - 1);
                    wa[t] = w_col + (idx & 7) * 16 + (idx >> 3) * P::W_SBO;
                    // (dead corners never match a chunk: their weights need no zeroing)
                    cga[t] = wy[cy] * wx[cx];
                    tw[t] = cga[t] * a;
                }
                swy[p][0] = wy[0]; swy[p][1] = wy[1]; swx[p][0] = wx[0]; swx[p][1] = wx[1];
            }
            // request the next level's samples now
            if (l + 1 < d.L) {
                LevelIO<AT>::load(lrow, arow, l + 1, d.P, row_ok, nxt_xy, nxt_a);
            }
            // window coordinates of this thread's pixel of the current chunk, advanced incrementally
            // (r / ww for r < 2^16 through the reciprocal: exact with the half-pixel offset)
            const float inv_ww = 1.f / (float)ww;
            int pwy = (int)(((float)tid + 0.5f) * inv_ww), pwx = tid - pwy * ww;
            const int step_y = (int)(((float)kPix + 0.5f) * inv_ww), step_x = kPix - step_y * ww;
            // value rows of chunk kk -> V buffer kk & 1 (operand B of the dot product): PPR lanes per pixel row, one piece
            // each, global -> shared without passing through registers; rows beyond the window are zero-filled
            auto fetch_values = [&](int kk) {
                uint8_t* vb = sV + (kk & 1) * P::G_BYTES;
#pragma unroll
                for (int i = 0; i < P::NIT; ++i) {
                    const int rl = i * P::RPI + lane / P::PPR, c = lane % P::PPR;
                    const int trow = warp * 32 + rl;
                    const int r = kk * kPix + trow;                               // window-linear pixel
                    const bool act = lane / P::PPR < P::RPI && rl < 32;
                    const bool ok = r < npix;
                    const int wy = (int)(((float)r + 0.5f) * inv_ww), wx = r - wy * ww;
                    const size_t vp = ok ? (size_t)lt.start[l] + (size_t)(ymin + wy) * W + (xmin + wx) : 0;
                    const uint32_t dst = smem_u32(vb + P::piece_off(trow, c));
                    const void* src = reinterpret_cast<const char*>(vbase + vp * pix_stride) + c * P::PIECE;
                    if (act) {
                        if constexpr (P::PIECE == 16)
                            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(dst), "l"(src), "r"(ok ? 16 : 0) : "memory");
                        else
                            asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" :: "r"(dst), "l"(src), "r"(ok ? 8 : 0) : "memory");
                    }
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
            };
            fetch_values(0);

            for (int k = 0; k < nch; ++k) {
                const int r0 = k * kPix;
                DT_MARK(3)
                // (a) this thread's window pixel (the epilogue's TMEM lane); the NEXT chunk's value rows start moving
                const bool pix_ok = r0 + tid < npix;
                const size_t pix = (size_t)lt.start[l] + (size_t)(ymin + pwy) * W + (xmin + pwx);
                pwy += step_y; pwx += step_x;
                if (pwx >= ww) { pwx -= ww; ++pwy; }
                if (k + 1 < nch) fetch_values(k + 1);
                DT_MARK(4)
                // (b) densify: column `tid` of W gets this row's corner weights that fall into the chunk.  The four
                // corners of one sample are distinct pixels: four independent read-modify-writes; corners of
                // different samples may coincide, so the samples go one after the other.
#pragma unroll
                for (int p = 0; p < kMaxP; ++p) {
                    uint16_t h[4];
                    bool hit[4];
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const int t = p * 4 + c;
                        hit[c] = (tr[t] >> 7) == k;
                        h[c] = 0;
                        if (hit[c]) asm volatile("ld.shared.u16 %0, [%1];" : "=h"(h[c]) : "r"(wa[t]) : "memory");
                    }
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const int t = p * 4 + c;
                        if (hit[c]) {
                            const uint16_t nh = O::down(O::up(h[c]) + tw[t]);
                            asm volatile("st.shared.u16 [%0], %1;" :: "r"(wa[t]), "h"(nh) : "memory");
                        }
                    }
                }
                DT_MARK(5)
                // (c) this chunk's value rows (operand B of the dot product) have landed
                if (k + 1 < nch) asm volatile("cp.async.wait_group 1;" ::: "memory");
                else asm volatile("cp.async.wait_group 0;" ::: "memory");
                DT_MARK(6)
                // (d) operands visible to the tensor core; TMEM reads of the previous chunk are complete
                proxy_fence();
                tmem_fence_before();
                __syncthreads();
                DT_MARK(7)
                // (e) one thread issues both products; the dot products are committed first so that the picks
                // below overlap the scatter product
                if (tid == 0) {
                    tmem_fence_after();
                    const uint64_t dg = umma_desc(smem_u32(sG), P::G_LBO, P::G_SBO);
                    const uint64_t dv = umma_desc(smem_u32(sV) + (k & 1) * P::G_BYTES, P::G_LBO, P::G_SBO);
#pragma unroll
                    for (int j = 0; j < P::DP / 16; ++j)
                        umma(tmem_u, dg + (uint64_t)(j * 2 * P::G_LBO / 16), dv + (uint64_t)(j * 2 * P::G_LBO / 16), idesc_u, j > 0);
                    umma_commit(mbar);
                    const uint64_t dw = umma_desc(smem_u32(sW), P::W_LBO, P::W_SBO);
                    const uint64_t dt = umma_desc(smem_u32(sG), P::G_SBO, P::G_LBO);
#pragma unroll
                    for (int j = 0; j < kThreads / 16; ++j)
                        umma(tmem_gv, dw + (uint64_t)(j * 2 * P::W_LBO / 16), dt + (uint64_t)(j * 2 * P::G_SBO / 16), idesc_gv, j > 0);
                    umma_commit(mbar2);
                }
                DT_MARK(8)
                // (f) dot products ready
                mbar_wait(mbar, phase);
                tmem_fence_after();
                DT_MARK(9)
                // (g) this row's 128 columns -> private staging row (64 at a time) -> pick the corners of this chunk
#pragma unroll 1
                for (int half = 0; half < 2; ++half) {
                    if (half * 64 >= npix - r0) break;       // columns beyond the window
                    {
                        float u[32];
                        tmem_ld32(tmem_u + lane_sel + half * 64, u);
#pragma unroll
                        for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(stage + j) = make_float4(u[j], u[j + 1], u[j + 2], u[j + 3]);
                        tmem_ld32(tmem_u + lane_sel + half * 64 + 32, u);
#pragma unroll
                        for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(stage + 32 + j) = make_float4(u[j], u[j + 1], u[j + 2], u[j + 3]);
                    }
                    const int want = 2 * k + half;
#pragma unroll
                    for (int t = 0; t < NT; ++t) {
                        float u = 0.f;
                        if ((tr[t] >> 6) == want) asm volatile("ld.shared.f32 %0, [%1];" : "=f"(u) : "r"(st_row + (tr[t] & 63) * 4) : "memory");
                        // one-sided derivatives of the bilinear weights (.cuh:141-160): -1 / +1 for the lower / upper pixel
                        const int cy = (t >> 1) & 1, cx = t & 1;
                        acc_a[t >> 2] = fmaf(cga[t], u, acc_a[t >> 2]);
                        acc_x[t >> 2] = fmaf(cx ? swy[t >> 2][cy] : -swy[t >> 2][cy], u, acc_x[t >> 2]);
                        acc_y[t >> 2] = fmaf(cy ? swx[t >> 2][cx] : -swx[t >> 2][cx], u, acc_y[t >> 2]);
                    }
                }
                DT_MARK(10)
                // (h) the item's contribution to the chunk's grad_value rows: TMEM lane = window pixel r0 + tid
                mbar_wait(mbar2, phase);
                phase ^= 1;
                tmem_fence_after();
                DT_MARK(11)
                if (r0 + warp * 32 < npix) {         // (warp-uniform) this warp's 32 pixel rows exist
                    float gv[D];
#pragma unroll
                    for (int c0 = 0; c0 + 32 <= D; c0 += 32) tmem_ld32(tmem_gv + lane_sel + c0, *reinterpret_cast<float(*)[32]>(&gv[c0]));
                    if constexpr (D % 32 == 4) tmem_ld4(tmem_gv + lane_sel + (D - 4), *reinterpret_cast<float(*)[4]>(&gv[D - 4]));
                    uint32_t any = 0;
#pragma unroll
                    for (int j = 0; j < D; ++j) any |= __float_as_uint(gv[j]);
                    const bool nz = (any << 1) != 0 && pix_ok;          // -0.0 counts as zero
                    __syncwarp();
#pragma unroll
                    for (int j = 0; j < D; j += 4) *reinterpret_cast<float4*>(stage + j) = make_float4(gv[j], gv[j + 1], gv[j + 2], gv[j + 3]);
                    const unsigned live = __ballot_sync(0xffffffffu, nz);
                    __syncwarp();
                    // a warp instruction carries 32 / (D / 4) whole rows (16 bytes per lane)
                    constexpr int LPR = D / 4, RPI = 32 / LPR;
                    const float* wstage = reinterpret_cast<const float*>(smem + P::OFF_STAGE) + (warp * 32 + lane / LPR) * P::STAGE_STRIDE + (lane % LPR) * 4;
                    float* my_dst = gvbase + (lane % LPR) * 4;
                    const unsigned my_pix = (unsigned)pix;
#pragma unroll
                    for (int i = 0; i < 32; i += RPI) {
                        const int rr = i + lane / LPR;
                        const unsigned spix = __shfl_sync(0xffffffffu, my_pix, rr & 31);
                        if (lane / LPR < RPI && rr < 32 && ((live >> rr) & 1u)) {
                            const float4 x = *reinterpret_cast<const float4*>(wstage + i * P::STAGE_STRIDE);
                            red_add_f32x4(my_dst + (size_t)spix * pix_stride, x.x, x.y, x.z, x.w);
                        }
                    }
                    __syncwarp();
                }
                DT_MARK(12)
                // (i) restore the zeros of W
#pragma unroll
                for (int t = 0; t < NT; ++t)
                    if ((tr[t] >> 7) == k) asm volatile("st.shared.u16 [%0], %1;" :: "r"(wa[t]), "h"((uint16_t)0) : "memory");
            }

            DT_MARK(13)
            // ---- the level's samples are complete: gradients of the sampling locations and attention weights ----
            if (row_ok) {
                float gx[kMaxP], gy[kMaxP];
#pragma unroll
                for (int p = 0; p < kMaxP; ++p) {
                    // .cuh:373-374: pixel-space derivative times attention weight times W (x) / H (y)
                    gx[p] = acc_x[p] * att[p] * (float)W;
                    gy[p] = acc_y[p] * att[p] * (float)H;
                }
                if constexpr (FUSED) {
                    // d / d location kept if asked for (the caller reduces it to d / d reference point)
                    if (fa.grad_loc) LevelIO<float>::store_loc_only(fa.grad_loc + row * LP * 2, l, d.P, gx, gy);
#pragma unroll
                    for (int p = 0; p < kMaxP; ++p) {
                        gx[p] *= lev_sx; gy[p] *= lev_sy;                     // location = ref + offset * scale
                        if (masked) acc_a[p] = 0.f;                          // masked_fill(mask, 0): no gradient to the softmax
                        sm_dot = fmaf(att[p], acc_a[p], sm_dot);             // sum_s prob_s * d out / d attention_s
                    }
                }
                // FUSED: grad_attn holds d out / d attention until the row's softmax backward below
                LevelIO<AT>::store(grad_loc + row * LP * 2, grad_attn + row * LP, l, d.P, gx, gy, acc_a);
            }
        }
        if constexpr (FUSED) {
            // softmax backward of the row: d / d logit_s = prob_s * (g_s - sum_t prob_t g_t); g_s were parked in grad_attn
            if (row_ok) {
                for (int l = 0; l < d.L; ++l) {
                    float lg[kMaxP], g[kMaxP];
                    LevelIO<AT>::load_attn_only(arow, l, d.P, true, lg);
                    LevelIO<AT>::load_attn_only(grad_attn + row * LP, l, d.P, true, g);
#pragma unroll
                    for (int p = 0; p < kMaxP; ++p) g[p] = expf(lg[p] - sm_max) * sm_inv * (g[p] - sm_dot);
                    LevelIO<AT>::store_attn_only(grad_attn + row * LP, l, d.P, g);
                }
            }
        }
        // G / the window table are rewritten by the next item
        if (tid == 0) { da.done[item] = 1; atomicAdd(da.claim + 2, 1); }
        tmem_fence_before();
        __syncthreads();
    }

    DT_MARK(14)
    DT_FLUSH
    tmem_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(s_tmem), "r"((uint32_t)kTmemCols));
}

}  // namespace dense
}  // namespace msda
