// msda_fwd.cuh -- fused multi-scale deformable attention forward for sm_100a.
//
// Replaces the reference's forward pipeline (ms_deform_attn_cuda.cu:50-83:
// `columns` scratch of L*P x the output + at::sum, kernel .cuh:165-237) with ONE
// kernel that gathers, interpolates, weights and reduces in registers and writes
// every output element exactly once.
//
// Mapping.  A CTA handles one work item (chunk of the query walk, batch, head; see walk_query).
// Inside an item a group of GS lanes owns one (query, head) row; each lane owns
// VEC consecutive channels and issues one 16-byte load per bilinear corner, so a
// warp-wide load instruction touches 32/GS distinct rows.  Sample metadata
// (x, y, weight) is loaded once per row, coalesced across the group, and
// broadcast with warp shuffles.
#pragma once
#include "msda_common.cuh"

// Resident CTAs per SM the forward is compiled for (register cap 64).  With one short item per CTA, four
// CTAs hide the per-item prologue and the gather latency better than three 80-register ones:
// C2-enc bf16 138 -> 129 us, fp32 168 -> 164 us; the 300-query decoder form loses ~2 us.
#ifndef MSDA_FWD_MINB
#define MSDA_FWD_MINB 4
#endif
namespace msda {

// One gathered sample: four 16-byte corner loads issued back to back, then 4*VEC FMAs.
template <typename VT, int VEC>
__device__ __forceinline__ void gather_sample(float (&acc)[VEC], const VT* lane_base, int o0, int o2, int dx,
                                              float b00, float b01, float b10, float b11)
{
    float v00[VEC], v01[VEC], v10[VEC], v11[VEC];
    // offsets are in VEC-element vectors: one IMAD.WIDE per address
    VecIO<VT, VEC>::load(lane_base + (size_t)(unsigned)o0 * VEC, v00);
    VecIO<VT, VEC>::load(lane_base + (size_t)(unsigned)(o0 + dx) * VEC, v01);
    VecIO<VT, VEC>::load(lane_base + (size_t)(unsigned)o2 * VEC, v10);
    VecIO<VT, VEC>::load(lane_base + (size_t)(unsigned)(o2 + dx) * VEC, v11);
    axpy_vec<VEC>(acc, b00, v00);
    axpy_vec<VEC>(acc, b01, v01);
    axpy_vec<VEC>(acc, b10, v10);
    axpy_vec<VEC>(acc, b11, v11);
}

// WIDE: every level has W > 1, so the column step of the 2x2 footprint is the same for all samples.
// FUSED: `loc` / `attn` are the module's raw sampling offsets / attention logits and the prologue
// (softmax, location arithmetic) runs in the owner lanes (SampleSrc<AT, true>, msda_common.cuh).
// DCH: channels per head.  DCH == VEC * GS: the GS lanes of a row tile its channels exactly.  DCH < VEC * GS
// ("overhang"; 16-bit D = 36 with 5 lanes of 8 channels): a (pixel, head) row of 72 bytes starts on a 16-byte
// boundary for even heads and 8 bytes past one for odd heads (M even), so the five 16-byte lanes cover
// [row, row + 80) for even heads and [row - 8, row + 72) for odd heads -- always inside the pixel's M rows, always
// aligned -- and the 4 channels a group reads beyond its row (the neighbouring head's) are accumulated into
// registers that are never stored.  Six rows per warp instruction with 16-byte loads instead of three with 8-byte
// loads (GS = 9, VEC = 4).
template <typename VT, typename AT, int VEC, int GS, int NWARPS, bool WIDE, bool FUSED, int DCH>
__device__ __forceinline__ void forward_items(const LevelTable& lt, int2* stage_a, int* stage_dx,
                                              float4* stage_b,
                                              const VT* __restrict__ value,
                                              const AT* __restrict__ loc, const AT* __restrict__ attn,
                                              VT* __restrict__ out, const KernelDims& d, const FusedArgs& fa)
{
    constexpr int GPW = 32 / GS;           // rows per warp
    constexpr int RS = record_stride(GS);  // records per sample in the shared-memory tap arrays
    constexpr int CH = chunk_of(GS);       // samples prepared per chunk (owner lanes gl < CH)
    constexpr int D = DCH;
    constexpr bool OVERHANG = DCH != VEC * GS;
    static_assert(!OVERHANG || (VEC == 8 && VEC * GS - DCH == 4 && sizeof(VT) == 2), "overhang: one half vector beyond the row");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int grp = lane / GS, gl = lane - grp * GS;
    const bool lane_used = grp < GPW;
    const int LP = d.L * d.P;
    const int rsv = d.M * D / VEC;         // vectors between neighbouring pixels (overhang: M is even)
    const int chunks = (d.Lq + d.tile_q - 1) / d.tile_q;
    const int items = chunks * d.N * d.M;
    constexpr int PASS = NWARPS * GPW;     // query rows per CTA pass

    for (int item = blockIdx.x; item < items; item += gridDim.x) {   // one item per CTA (grid == items)
        const int chunk = item / (d.N * d.M);
        const int bm = item - chunk * (d.N * d.M);
        const int b = bm / d.M, m = bm - b * d.M;
        const int shift = (OVERHANG && (m & 1)) ? -(VEC / 2) : 0;   // odd heads: the aligned window starts 4 channels early
        const VT* lane_base = value + ((size_t)b * d.S * d.M + m) * D + shift + gl * VEC;

        const int pos0 = chunk * d.tile_q;
        const int count = min(d.tile_q, d.Lq - pos0);
        for (int i0 = warp * GPW; i0 < count; i0 += PASS) {
            const int q = (lane_used && i0 + grp < count) ? walk_query(lt, d, pos0 + i0 + grp) : -1;
            const bool row_ok = q >= 0;
            const size_t rowq = (size_t)b * d.Lq + (row_ok ? q : 0);
            const size_t row = rowq * d.M + m;
            SampleSrc<AT, FUSED, 0> src;
            src.template begin<GS, CH>(loc, attn, fa, d, row, rowq, row_ok, gl, grp * GS);

            float acc[VEC];
#pragma unroll
            for (int k = 0; k < VEC; ++k) acc[k] = 0.f;

            for (int s0 = 0; s0 < LP; s0 += CH) {
                // owner phase: lane gl < CH of the group prepares sample s0 + gl of its row
                const int s = s0 + gl;
                const bool mine = row_ok && gl < CH && s < LP;
                const int ls = min(s, LP - 1) / d.P;
                float2 xy;
                float a;
                src.get(s, s0 / CH, ls, mine, d, xy.x, xy.y, a);
                const OwnerTap t = make_owner_tap(mine, xy.x, xy.y, a, lt.H[ls], lt.W[ls], lt.start[ls], rsv);
#ifdef MSDA_BOUNDS_CHECK
                // debug builds (compute-sanitizer replacement): the whole 2x2 footprint lies inside the batch item
                if (mine) assert(t.off0 >= 0 && (long long)t.off0 + t.dy + (WIDE ? rsv : t.dx) <= (long long)(d.S - 1) * rsv);
                if (row_ok) assert(q < d.Lq && b < d.N);
#endif
                const float wa0 = t.ay.w0 * t.a, wa1 = t.ay.w1 * t.a;
                // publish the finished tap in shared memory, laid out [sample j][group] so the
                // broadcast reads of one step (one 16-byte record per group) are conflict-free.
                // (Shuffles would cost one LSU data-pipe wavefront per value; this is 3 per sample.)
                __syncwarp();
                if (lane_used && gl < CH) {
                    stage_a[gl * RS + grp] = make_int2(t.off0, t.off0 + t.dy);
                    if constexpr (!WIDE) stage_dx[gl * RS + grp] = t.dx;
                    stage_b[gl * RS + grp] = make_float4(wa0 * t.ax.w0, wa0 * t.ax.w1, wa1 * t.ax.w0, wa1 * t.ax.w1);
                }
                __syncwarp();
                // broadcast phase: all lanes of the group consume sample s0 + j.  A ragged last
                // chunk runs the full CH steps too: lanes without a sample publish weight 0 at
                // offset 0 (valid memory), so no guards are needed inside the unrolled body.
                const int gsel = lane_used ? grp : 0;
#pragma unroll
                for (int j = 0; j < CH; ++j) {
                    const int2 ra = stage_a[j * RS + gsel];
                    const float4 rb = stage_b[j * RS + gsel];
                    gather_sample<VT, VEC>(acc, lane_base, ra.x, ra.y, WIDE ? rsv : stage_dx[j * RS + gsel], rb.x, rb.y,
                                           rb.z, rb.w);
                }
            }
            if (row_ok) {
                VT* o = out + row * D + shift + gl * VEC;
                if constexpr (!OVERHANG) {
                    VecIO<VT, VEC>::store(o, acc);
                } else {
                    // the lane holding the neighbouring head's 4 channels stores its own half only
                    if (shift == 0 && gl == GS - 1) VecIO<VT, VEC / 2>::store(o, *reinterpret_cast<float(*)[VEC / 2]>(&acc[0]));
                    else if (shift != 0 && gl == 0) VecIO<VT, VEC / 2>::store(o + VEC / 2, *reinterpret_cast<float(*)[VEC / 2]>(&acc[VEC / 2]));
                    else VecIO<VT, VEC>::store(o, acc);
                }
            }
        }
    }
}

template <typename VT, typename AT, int VEC, int GS, int NWARPS, bool FUSED, int DCH = VEC * GS>
__global__ void __launch_bounds__(NWARPS * 32, MSDA_FWD_MINB)
msda_forward_kernel(const VT* __restrict__ value, const int64_t* __restrict__ shapes,
                    const int64_t* __restrict__ level_start, const AT* __restrict__ loc,
                    const AT* __restrict__ attn, VT* __restrict__ out, const KernelDims d, const FusedArgs fa)
{
    __shared__ LevelTable lt;
    stage_levels(lt, shapes, level_start, d);
    // With W > 1 everywhere and at least two pixels, base + column step is always valid memory.
    __shared__ int2 stage_a[NWARPS][record_count(GS)];
    __shared__ int stage_dx[NWARPS][record_count(GS)];
    __shared__ float4 stage_b[NWARPS][record_count(GS)];
    const int warp = threadIdx.x >> 5;
    if (lt.min_w > 1)
        forward_items<VT, AT, VEC, GS, NWARPS, true, FUSED, DCH>(lt, stage_a[warp], stage_dx[warp], stage_b[warp], value, loc,
                                                            attn, out, d, fa);
    else
        forward_items<VT, AT, VEC, GS, NWARPS, false, FUSED, DCH>(lt, stage_a[warp], stage_dx[warp], stage_b[warp], value, loc,
                                                             attn, out, d, fa);
}

// Generic fallback: any D, any supported type (incl. fp64).  One warp per
// (query, head) row, lanes stride over channels.  ACC is the accumulation type.
template <typename VT, typename AT, typename ACC>
__global__ void __launch_bounds__(256)
msda_forward_generic_kernel(const VT* __restrict__ value, const int64_t* __restrict__ shapes,
                            const int64_t* __restrict__ level_start, const AT* __restrict__ loc,
                            const AT* __restrict__ attn, VT* __restrict__ out, const KernelDims d)
{
    __shared__ LevelTable lt;
    stage_levels(lt, shapes, level_start, d);
    const int lane = threadIdx.x & 31;
    const size_t warps = (size_t)gridDim.x * (blockDim.x >> 5);
    const size_t rows = (size_t)d.N * d.Lq * d.M;
    const int LP = d.L * d.P;
    for (size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows; row += warps) {
        const int m = (int)(row % d.M);
        const size_t b = row / ((size_t)d.Lq * d.M);
        const VT* vb = value + (b * d.S * d.M + m) * d.D;
        for (int c = lane; c < d.D; c += 32) {
            ACC acc = 0;
            for (int s = 0; s < LP; ++s) {
                const int l = s / d.P;
                const int H = lt.H[l], W = lt.W[l];
                const ACC x = (ACC)loc[(row * LP + s) * 2] * (ACC)W - (ACC)0.5;
                const ACC y = (ACC)loc[(row * LP + s) * 2 + 1] * (ACC)H - (ACC)0.5;
                if (!(y > (ACC)-1 && x > (ACC)-1 && y < (ACC)H && x < (ACC)W)) continue;
                const ACC a = (ACC)attn[row * LP + s];
                const ACC fy = floor(y), fx = floor(x);
                const int y0 = (int)fy, x0 = (int)fx;
                const ACC lh = y - fy, lw = x - fx, hh = (ACC)1 - lh, hw = (ACC)1 - lw;
                const size_t pix = (size_t)lt.start[l] + (size_t)y0 * W + x0;
                const size_t rs = (size_t)d.M * d.D;
                ACC sum = 0;
                if (y0 >= 0 && x0 >= 0) sum += hh * hw * (ACC)vb[pix * rs + c];
                if (y0 >= 0 && x0 + 1 <= W - 1) sum += hh * lw * (ACC)vb[(pix + 1) * rs + c];
                if (y0 + 1 <= H - 1 && x0 >= 0) sum += lh * hw * (ACC)vb[(pix + W) * rs + c];
                if (y0 + 1 <= H - 1 && x0 + 1 <= W - 1) sum += lh * lw * (ACC)vb[(pix + W + 1) * rs + c];
                acc += sum * a;
            }
            out[row * d.D + c] = (VT)acc;
        }
    }
}

}  // namespace msda
