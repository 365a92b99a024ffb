// msda_fwd.cuh -- fused multi-scale deformable attention forward for sm_100a.
//
// Replaces the reference's forward pipeline (ms_deform_attn_cuda.cu:50-83:
// `columns` scratch of L*P x the output + at::sum, kernel .cuh:165-237) with ONE
// kernel that gathers, interpolates, weights and reduces in registers and writes
// every output element exactly once.
//
// Mapping.  A CTA loops (persistently) over work items (query tile, batch, head).
// Inside an item a group of GS lanes owns one (query, head) row; each lane owns
// VEC consecutive channels and issues one 16-byte load per bilinear corner, so a
// warp-wide load instruction touches 32/GS distinct rows.  Sample metadata
// (x, y, weight) is loaded once per row, coalesced across the group, and
// broadcast with warp shuffles.
#pragma once
#include "msda_common.cuh"

namespace msda {

template <typename VT, typename AT, int VEC, int GS, int NWARPS>
__global__ void __launch_bounds__(NWARPS * 32)
msda_forward_kernel(const VT* __restrict__ value, const int64_t* __restrict__ shapes,
                    const int64_t* __restrict__ level_start, const AT* __restrict__ loc,
                    const AT* __restrict__ attn, VT* __restrict__ out, const KernelDims d)
{
    constexpr int GPW = 32 / GS;           // rows per warp
    constexpr int D = VEC * GS;
    __shared__ LevelTable lt;
    stage_levels(lt, shapes, level_start, d);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int grp = lane / GS, gl = lane - grp * GS;
    const bool lane_used = grp < GPW;
    const int grp_base = grp * GS;         // first lane of this group
    const int LP = d.L * d.P;
    const int row_stride = d.M * D;        // elements between neighbouring pixels
    const int items = lt.total_tiles * d.N * d.M;

    for (int item = blockIdx.x; item < items; item += gridDim.x) {
        const int tile = item / (d.N * d.M);
        const int bm = item - tile * (d.N * d.M);
        const int b = bm / d.M, m = bm - b * d.M;
        const VT* vbase = value + ((size_t)b * d.S * d.M + m) * D + gl * VEC;

        for (int i0 = warp * GPW; i0 < d.tile_q; i0 += NWARPS * GPW) {
            const int q = (lane_used && i0 + grp < d.tile_q) ? tile_query(lt, d, tile, i0 + grp) : -1;
            const bool row_ok = q >= 0;
            const size_t row = ((size_t)b * d.Lq + (row_ok ? q : 0)) * d.M + m;
            const AT* lrow = loc + row * LP * 2;
            const AT* arow = attn + row * LP;

            float acc[VEC];
#pragma unroll
            for (int k = 0; k < VEC; ++k) acc[k] = 0.f;

            int l = 0, p = 0;
            int H = lt.H[0], W = lt.W[0], st = lt.start[0];
            for (int s0 = 0; s0 < LP; s0 += GS) {
                // each lane of the group fetches one sample's metadata
                float2 my_xy = make_float2(-8.f, -8.f);
                float my_a = 0.f;
                if (row_ok && s0 + gl < LP) {
                    my_xy = load_loc<AT>(lrow + (s0 + gl) * 2);
                    my_a = load_attn<AT>(arow + s0 + gl);
                }
                const int cnt = min(GS, LP - s0);
#pragma unroll 4
                for (int j = 0; j < cnt; ++j) {
                    const float lx = __shfl_sync(0xffffffffu, my_xy.x, grp_base + j);
                    const float ly = __shfl_sync(0xffffffffu, my_xy.y, grp_base + j);
                    const float a = __shfl_sync(0xffffffffu, my_a, grp_base + j);
                    const Tap t = make_tap(lx, ly, H, W, st, row_stride);
                    const float wa0 = t.wy0 * a, wa1 = t.wy1 * a;
                    const float w[4] = {wa0 * t.wx0, wa0 * t.wx1, wa1 * t.wx0, wa1 * t.wx1};
                    float v[4][VEC];
#pragma unroll
                    for (int c = 0; c < 4; ++c) VecIO<VT, VEC>::load(vbase + t.off[c], v[c]);
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
#pragma unroll
                        for (int k = 0; k < VEC; ++k) acc[k] = fmaf(w[c], v[c][k], acc[k]);
                    }
                    if (++p == d.P) {
                        p = 0;
                        ++l;
                        if (l < d.L) { H = lt.H[l]; W = lt.W[l]; st = lt.start[l]; }
                    }
                }
            }
            if (row_ok) VecIO<VT, VEC>::store(out + row * D + gl * VEC, acc);
        }
    }
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
