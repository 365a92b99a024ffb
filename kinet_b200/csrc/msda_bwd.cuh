// msda_bwd.cuh -- multi-scale deformable attention backward for sm_100a.
//
// Replaces the reference's two backward kernels (.cuh:239-306 grad_value with a
// 5x5 neighbourhood scan and one scalar global atomic per channel and corner;
// .cuh:308-378 grad_sampling_loc / grad_attn_weight with a serial, uncoalesced
// channel loop that re-reads every corner three times) with ONE pass over the
// samples that reads each corner row once (16-byte loads) and produces all three
// gradients:
//   * grad_attn_weight / grad_sampling_loc: per-lane partial sums over the lane's
//     VEC channels, reduced across the GS lanes of the row with warp shuffles;
//     every output element has exactly one writer (no atomics), written coalesced.
//   * grad_value: 16-byte vector reductions (red.global.add.v4.f32, sm_90+) --
//     one instruction per lane and corner instead of four scalar atomics.
//     16-bit value types accumulate into an fp32 workspace that a second small
//     kernel rounds once (accumulating in bf16 would lose the small terms).
#pragma once
#include "msda_common.cuh"

#ifndef MSDA_BWD_MINB
#define MSDA_BWD_MINB 2
#endif
namespace msda {

// Reduce-scatter over the first N lanes of a row group (N a power of two; the group starts at lane
// grp_base, which need not be a multiple of N): on entry every lane holds, for each of the N samples,
// 4 partial dot products; on exit lane gl < N holds the totals over those N lanes of sample gl.
// log2(N) butterfly steps that halve the number of live values, i.e. 4*(N-1) shuffles per chunk
// instead of 4*N*log2(N) for per-sample all-reduces.
template <int N>
__device__ __forceinline__ void reduce_scatter(float (&U)[N][4], int gl, int grp_base, float (&out)[4])
{
    if constexpr (N == 1) {
#pragma unroll
        for (int c = 0; c < 4; ++c) out[c] = U[0][c];
    } else {
        constexpr int HALF = N / 2;
        const bool upper = (gl & HALF) != 0;
        float K[HALF][4];
#pragma unroll
        for (int i = 0; i < HALF; ++i) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const float keep = upper ? U[i + HALF][c] : U[i][c];
                const float send = upper ? U[i][c] : U[i + HALF][c];
                // lanes outside the butterfly (gl >= its size; unused lanes of the warp) read garbage, never used
                K[i][c] = keep + __shfl_sync(0xffffffffu, send, (grp_base + (gl ^ HALF)) & 31);
            }
        }
        reduce_scatter<HALF>(K, gl, grp_base, out);
    }
}

// WIDE: every level has W > 1 (uniform column step of the 2x2 footprint).
// FUSED: loc / attn are raw sampling offsets / attention logits, grad_loc / grad_attn their gradients
// (through the location arithmetic and the softmax); see SampleSrc<AT, true> in msda_common.cuh.
//
// backward_rows: the `count` query rows qmap(0) .. qmap(count-1) (-1 = no row) of one (batch, head) pair, NWARPS
// warps.  Called once per work item by msda_backward_kernel (walk items; in split mode also the tiles the
// dense-window kernel of msda_dense.cuh handed back).
template <typename VT, typename AT, int VEC, int GS, int NWARPS, bool WIDE, bool FUSED, typename QMap>
__device__ __forceinline__ void backward_rows(const LevelTable& lt, int2* stage_a, int* stage_dx,
                                              float4* stage_b, float4* fold,
                                              const VT* __restrict__ value,
                                              const AT* __restrict__ loc, const AT* __restrict__ attn,
                                              const VT* __restrict__ grad_out,
                                              float* __restrict__ grad_value_acc, AT* __restrict__ grad_loc,
                                              AT* __restrict__ grad_attn, const KernelDims& d, const FusedArgs& fa,
                                              int b, int m, int count, const QMap& qmap)
{
    // fused: chunks of a row whose d out / d attention stay in registers until the row's softmax backward
    // (the launcher guarantees L*P <= MAXCH * CH)
    constexpr int MAXCH = kFusedMaxChunks;
    constexpr int GPW = 32 / GS;
    constexpr int D = VEC * GS;
    constexpr int RS = record_stride(GS);  // records per sample in the shared-memory tap arrays
    constexpr int CH = chunk_of(GS);   // samples per chunk, prepared by the owner lanes gl < CH
    // VEC = 4 for every value type: a lane's reduction is 16 contiguous bytes of the fp32 accumulator
    // row, the GS lanes of a row cover it without gaps, and one warp instruction carries whole rows.
    // The SM->L2 request port sends one packet (header + 32-byte sectors) per 128-byte line touched,
    // so fewer, fuller lines per instruction is what sets the scatter rate: with 8
    // channels per lane (bf16 x 8 loads) the same kernel measured 945 us instead of 538 us.
    static_assert(VEC == 4, "one 16-byte reduction per lane and corner");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int grp = lane / GS, gl = lane - grp * GS;
    const bool lane_used = grp < GPW;
    const int grp_base = grp * GS;
    const int LP = d.L * d.P;
    const int rsv = d.M * GS;
    constexpr int PASS = NWARPS * GPW;     // query rows per CTA pass
    {
        const size_t bbase = ((size_t)b * d.S * d.M + m) * D + gl * VEC;
        const VT* lane_base = value + bbase;
        float* gv_base = grad_value_acc + bbase;

        for (int i0 = warp * GPW; i0 < count; i0 += PASS) {
            const int q = (lane_used && i0 + grp < count) ? qmap(i0 + grp) : -1;
            const bool row_ok = q >= 0;
            const size_t rowq = (size_t)b * d.Lq + (row_ok ? q : 0);
            const size_t row = rowq * d.M + m;
            SampleSrc<AT, FUSED, 4> src;
            src.template begin<GS, CH>(loc, attn, fa, d, row, rowq, row_ok, gl, grp_base);
            float ga_keep[MAXCH] = {0.f, 0.f, 0.f, 0.f};   // fused: d out / d attention of this lane's samples
            float dot = 0.f;                                // fused: sum_s prob_s * d out / d attention_s

            float g[VEC];
#pragma unroll
            for (int k = 0; k < VEC; ++k) g[k] = 0.f;
            if (row_ok) VecIO<VT, VEC>::load_stream(grad_out + row * D + gl * VEC, g);
            for (int s0 = 0; s0 < LP; s0 += CH) {
                // owner phase: lane gl < CH prepares sample s0 + gl of its row
                const int s = s0 + gl;
                const bool mine = row_ok && gl < CH && s < LP;
                const int ls = min(s, LP - 1) / d.P;
                float2 xy;
                float a;
                src.get(s, s0 / CH, ls, mine, d, xy.x, xy.y, a);
                const int H = lt.H[ls], W = lt.W[ls];
                const OwnerTap t = make_owner_tap(mine, xy.x, xy.y, a, H, W, lt.start[ls], rsv);
#ifdef MSDA_BOUNDS_CHECK
                // debug builds (compute-sanitizer replacement): the whole 2x2 footprint lies inside the batch item
                if (mine) assert(t.off0 >= 0 && (long long)t.off0 + t.dy + (WIDE ? rsv : t.dx) <= (long long)(d.S - 1) * rsv);
                if (row_ok) assert(q < d.Lq && b < d.N);
#endif
                const float wa0 = t.ay.w0 * t.a, wa1 = t.ay.w1 * t.a;
                // publish the tap (offsets; corner weight * attn) in shared memory, [sample][group]
                __syncwarp();
                if (lane_used && gl < CH) {
                    stage_a[gl * RS + grp] = make_int2(t.off0, t.off0 + t.dy);
                    if constexpr (!WIDE) stage_dx[gl * RS + grp] = t.dx;
                    stage_b[gl * RS + grp] = make_float4(wa0 * t.ax.w0, wa0 * t.ax.w1, wa1 * t.ax.w0, wa1 * t.ax.w1);
                }
                __syncwarp();
                const int gsel = lane_used ? grp : 0;

                // broadcast phase: every lane of the group handles its VEC channels of sample s0+j
                float U[CH][4];
#pragma unroll
                for (int j = 0; j < CH; ++j) {
                    const int2 ra = stage_a[j * RS + gsel];
                    const float4 rb = stage_b[j * RS + gsel];
                    const int o0 = ra.x, o2 = ra.y, dx = WIDE ? rsv : stage_dx[j * RS + gsel];
                    const float bw[4] = {rb.x, rb.y, rb.z, rb.w};
                    const unsigned o[4] = {(unsigned)o0, (unsigned)(o0 + dx), (unsigned)o2, (unsigned)(o2 + dx)};
                    float v[4][VEC];
#pragma unroll
                    for (int c = 0; c < 4; ++c) VecIO<VT, VEC>::load(lane_base + (size_t)o[c] * VEC, v[c]);
                    // u[c] = <grad_out slice, corner c>: all three scalar gradients are linear in u
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        float acc2[2] = {0.f, 0.f};
#pragma unroll
                        for (int k = 0; k < VEC; k += 2) {
                            const float2 r = __ffma2_rn(make_float2(g[k], g[k + 1]),
                                                        make_float2(v[c][k], v[c][k + 1]),
                                                        make_float2(acc2[0], acc2[1]));
                            acc2[0] = r.x; acc2[1] = r.y;
                        }
                        U[j][c] = acc2[0] + acc2[1];
                    }
                    // scatter: grad_value[corner] += (corner weight * attn) * grad_out
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        if (bw[c] != 0.f) {
                            if (sizeof(VT) == 2 && d.acc16) {
                                // sparse query sets (~1 contribution per pixel row): straight into the 16-bit output
                                VT* dst = reinterpret_cast<VT*>(grad_value_acc) + bbase + (size_t)o[c] * VEC;
#pragma unroll
                                for (int k = 0; k < VEC; k += 4)
                                    red_add_16x4<VT>(dst + k, bw[c] * g[k], bw[c] * g[k + 1], bw[c] * g[k + 2], bw[c] * g[k + 3]);
                            } else {
                                float* dst = gv_base + (size_t)o[c] * VEC;
#pragma unroll
                                for (int k = 0; k < VEC; k += 4)
                                    red_add_f32x4(dst + k, bw[c] * g[k], bw[c] * g[k + 1], bw[c] * g[k + 2],
                                                  bw[c] * g[k + 3]);
                            }
                        }
                    }
                }
                // bring the group totals of sample s0+gl to its owner lane
                float u[4];
                if constexpr (GS != CH) {
                    // lanes gl >= CH (D = 36: lane 8) hand their partials to the owners through shared memory
                    static_assert(GS == CH + 1, "one extra channel lane per group");
                    __syncwarp();
                    if (lane_used && gl == CH) {
#pragma unroll
                        for (int j = 0; j < CH; ++j) fold[grp * CH + j] = make_float4(U[j][0], U[j][1], U[j][2], U[j][3]);
                    }
                    __syncwarp();
                }
                reduce_scatter<CH>(U, gl, grp_base, u);
                if constexpr (GS != CH) {
                    const float4 f = fold[(lane_used ? grp : 0) * CH + (gl & (CH - 1))];
                    u[0] += f.x; u[1] += f.y; u[2] += f.z; u[3] += f.w;
                }
                if (mine) {
                    // owner lane: its own (unbroadcast) tap turns u into the three gradients
                    const float ga = t.ay.w0 * (t.ax.w0 * u[0] + t.ax.w1 * u[1]) +
                                     t.ay.w1 * (t.ax.w0 * u[2] + t.ax.w1 * u[3]);
                    const float gx = t.ay.w0 * (t.ax.d0 * u[0] + t.ax.d1 * u[1]) +
                                     t.ay.w1 * (t.ax.d0 * u[2] + t.ax.d1 * u[3]);
                    const float gy = t.ay.d0 * (t.ax.w0 * u[0] + t.ax.w1 * u[1]) +
                                     t.ay.d1 * (t.ax.w0 * u[2] + t.ax.w1 * u[3]);
                    // .cuh:373-374: pixel-space derivative times attn times W (x) / H (y)
                    const float glx = gx * t.a * (float)W, gly = gy * t.a * (float)H;
                    if constexpr (!FUSED) {
                        store_loc<AT>(grad_loc + (row * LP + s) * 2, glx, gly);
                        grad_attn[row * LP + s] = from_f32<AT>(ga);
                    } else {
                        // location = ref + offset * (sx, sy)  ->  d/d offset; d/d location kept if asked for
                        store_loc<AT>(grad_loc + (row * LP + s) * 2, glx * src.sx, gly * src.sy);
                        if (fa.grad_loc) store_loc<float>(fa.grad_loc + (row * LP + s) * 2, glx, gly);
                        const float gam = src.masked ? 0.f : ga;   // masked_fill(mask, 0): no gradient to the softmax
                        const int c = s0 / CH;
#pragma unroll
                        for (int k = 0; k < MAXCH; ++k)
                            if (c == k) ga_keep[k] = gam;
                        dot = fmaf(src.prob(c), gam, dot);
                    }
                }
            }
            if constexpr (FUSED) {
                // softmax backward of the row: d/d logit_s = prob_s * (g_s - sum_t prob_t * g_t)
                const float tot = group_reduce<GS, SumOp>(dot, gl, grp_base);
                if (row_ok) {
#pragma unroll
                    for (int k = 0; k < MAXCH; ++k) {
                        const int s = k * CH + gl;
                        if (gl < CH && s < LP) grad_attn[row * LP + s] = from_f32<AT>(src.prob(k) * (ga_keep[k] - tot));
                    }
                }
            }
        }
    }
}

// One work item per CTA (grid == items; the loop only guards against a capped grid): item = (chunk of tile_q
// positions of the query walk, batch, head), see walk_query.
//
// Split mode (dense_done != nullptr; the dense-window kernel of msda_dense.cuh ran before on the same stream): the
// queries of the finest level belong to the dense kernel, which flags the (tile, batch, head) items it has done
// (it leaves those whose windows are too large for its chunks).  This kernel is then a persistent grid that claims, from a
// zero-initialised counter, (1) chunks of tile_q walk positions of the OTHER levels' queries and (2) the unflagged tiles.  (The host
// does not know where the finest level's queries start -- the shapes live in device memory -- so it cannot size a
// one-item-per-CTA grid for them.)
template <typename VT, typename AT, int VEC, int GS, int NWARPS, bool WIDE, bool FUSED>
__device__ __forceinline__ void backward_items(const LevelTable& lt, const FineTiles& ft, int2* stage_a, int* stage_dx,
                                               float4* stage_b, float4* fold,
                                               const VT* __restrict__ value,
                                               const AT* __restrict__ loc, const AT* __restrict__ attn,
                                               const VT* __restrict__ grad_out,
                                               float* __restrict__ grad_value_acc, AT* __restrict__ grad_loc,
                                               AT* __restrict__ grad_attn, const KernelDims& d, const FusedArgs& fa,
                                               const uint8_t* dense_done, int* __restrict__ split_claim, int* s_claim)
{
    if (dense_done == nullptr) {
        const int chunks = (d.Lq + d.tile_q - 1) / d.tile_q;
        const int items = chunks * d.N * d.M;
        for (int item = blockIdx.x; item < items; item += gridDim.x) {
            const int chunk = item / (d.N * d.M);
            const int bm = item - chunk * (d.N * d.M);
            const int b = bm / d.M, m = bm - b * d.M;
            const int pos0 = chunk * d.tile_q;
            const int count = min(d.tile_q, d.Lq - pos0);
            backward_rows<VT, AT, VEC, GS, NWARPS, WIDE, FUSED>(
                lt, stage_a, stage_dx, stage_b, fold, value, loc, attn, grad_out, grad_value_acc, grad_loc, grad_attn, d, fa,
                b, m, count, [&](int i) { return walk_query(lt, d, pos0 + i); });
        }
        return;
    }
    const bool split = ft.level >= 0;
    // walk positions and queries of a level share one index range: the positions outside [q_begin, q_end)
    const int skip_begin = split ? ft.q_begin : d.Lq, skip_len = split ? ft.q_end - ft.q_begin : 0;
    const int coarse = d.Lq - skip_len;
    const int walk_items = ((coarse + d.tile_q - 1) / d.tile_q) * d.N * d.M;
    const int tile_items = split ? ft.total * d.N * d.M : 0;
    bool waited = false;
    for (;;) {
        if (threadIdx.x == 0) *s_claim = atomicAdd(split_claim, 1);
        __syncthreads();
        const int item = *s_claim;
        __syncthreads();
        if (item >= walk_items + tile_items) break;
        if (item < walk_items) {
            const int chunk = item / (d.N * d.M);
            const int bm = item - chunk * (d.N * d.M);
            const int b = bm / d.M, m = bm - b * d.M;
            const int c0 = chunk * d.tile_q;
            const int count = min(d.tile_q, coarse - c0);
            backward_rows<VT, AT, VEC, GS, NWARPS, WIDE, FUSED>(
                lt, stage_a, stage_dx, stage_b, fold, value, loc, attn, grad_out, grad_value_acc, grad_loc, grad_attn, d, fa,
                b, m, count, [&](int i) {
                    const int c = c0 + i;
                    return walk_query(lt, d, c < skip_begin ? c : c + skip_len);
                });
        } else {
            const int it = item - walk_items;
            // the flags are the dense-window kernel's: wait for it (no-op unless this grid was launched as its
            // programmatic dependent, and after the first time)
            if (!waited) { asm volatile("griddepcontrol.wait;" ::: "memory"); waited = true; }
            // (written by the dense-window grid while THIS grid is already resident: a coherent load -- the read-only
            // data path, which `const __restrict__` invites, may serve a line cached before that grid finished)
            if (*reinterpret_cast<const volatile uint8_t*>(dense_done + it) != 0) continue;
            const int tile = it / (d.N * d.M);
            const int bm = it - tile * (d.N * d.M);
            const int b = bm / d.M, m = bm - b * d.M;
            const TileRect r = fine_tile_rect(ft, lt, tile);
            backward_rows<VT, AT, VEC, GS, NWARPS, WIDE, FUSED>(
                lt, stage_a, stage_dx, stage_b, fold, value, loc, attn, grad_out, grad_value_acc, grad_loc, grad_attn, d, fa,
                b, m, kTileQueries, [&](int i) { return fine_tile_query(ft, lt, r, i, d.Lq); });
        }
    }
    // this grid must not complete before the dense-window kernel has (the rounding pass is ordered behind this grid)
    if (!waited) asm volatile("griddepcontrol.wait;" ::: "memory");
}

template <typename VT, typename AT, int VEC, int GS, int NWARPS, bool FUSED>
__global__ void __launch_bounds__(NWARPS * 32, MSDA_BWD_MINB)
msda_backward_kernel(const VT* __restrict__ value, const int64_t* __restrict__ shapes,
                     const int64_t* __restrict__ level_start, const AT* __restrict__ loc,
                     const AT* __restrict__ attn, const VT* __restrict__ grad_out,
                     float* __restrict__ grad_value_acc, AT* __restrict__ grad_loc,
                     AT* __restrict__ grad_attn, const KernelDims d, const FusedArgs fa,
                     const uint8_t* dense_done, int tiles_bound, int* __restrict__ split_claim)
{
    __shared__ LevelTable lt;
    __shared__ FineTiles ft;
    __shared__ int ft_scratch, s_claim;
    stage_levels(lt, shapes, level_start, d);
    if (dense_done != nullptr) build_fine_tiles(ft, &ft_scratch, lt, d.L, tiles_bound);
    __shared__ int2 stage_a[NWARPS][record_count(GS)];
    __shared__ int stage_dx[NWARPS][record_count(GS)];
    __shared__ float4 stage_b[NWARPS][record_count(GS)];
    constexpr int FOLD = GS == chunk_of(GS) ? 1 : (32 / GS) * chunk_of(GS);   // only groups with a non-owner lane fold
    __shared__ float4 fold[NWARPS][FOLD];
    const int warp = threadIdx.x >> 5;
    if (lt.min_w > 1)
        backward_items<VT, AT, VEC, GS, NWARPS, true, FUSED>(lt, ft, stage_a[warp], stage_dx[warp], stage_b[warp], fold[warp], value, loc,
                                                             attn, grad_out, grad_value_acc, grad_loc, grad_attn, d, fa, dense_done, split_claim, &s_claim);
    else
        backward_items<VT, AT, VEC, GS, NWARPS, false, FUSED>(lt, ft, stage_a[warp], stage_dx[warp], stage_b[warp], fold[warp], value, loc,
                                                              attn, grad_out, grad_value_acc, grad_loc, grad_attn, d, fa, dense_done, split_claim, &s_claim);
}

// fp32 workspace -> 16-bit grad_value (one rounding of the full sum).
template <typename VT>
__global__ void __launch_bounds__(256)
msda_round_grad_value_kernel(const float* __restrict__ acc, VT* __restrict__ gv, size_t n8)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8;
         i += (size_t)gridDim.x * blockDim.x) {
        float v[8];
        VecIO<float, 4>::load_stream(acc + i * 8, *reinterpret_cast<float(*)[4]>(&v[0]));
        VecIO<float, 4>::load_stream(acc + i * 8 + 4, *reinterpret_cast<float(*)[4]>(&v[4]));
        VecIO<VT, 8>::store(gv + i * 8, v);
    }
}

template <typename T> __device__ __forceinline__ void atomic_add_any(T* p, T v) { atomicAdd(p, v); }

// Generic fallback (any D / fp64): one warp per (query, head) row.
// GV: accumulation buffer element type (double for fp64, float otherwise).
template <typename VT, typename AT, typename ACC>
__global__ void __launch_bounds__(256)
msda_backward_generic_kernel(const VT* __restrict__ value, const int64_t* __restrict__ shapes,
                             const int64_t* __restrict__ level_start, const AT* __restrict__ loc,
                             const AT* __restrict__ attn, const VT* __restrict__ grad_out,
                             ACC* __restrict__ grad_value_acc, AT* __restrict__ grad_loc,
                             AT* __restrict__ grad_attn, const KernelDims d)
{
    __shared__ LevelTable lt;
    stage_levels(lt, shapes, level_start, d);
    const int lane = threadIdx.x & 31;
    const size_t warps = (size_t)gridDim.x * (blockDim.x >> 5);
    const size_t rows = (size_t)d.N * d.Lq * d.M;
    const int LP = d.L * d.P;
    const size_t rs = (size_t)d.M * d.D;
    for (size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows; row += warps) {
        const int m = (int)(row % d.M);
        const size_t b = row / ((size_t)d.Lq * d.M);
        const size_t bbase = (b * d.S * d.M + m) * d.D;
        for (int s = 0; s < LP; ++s) {
            const int l = s / d.P;
            const int H = lt.H[l], W = lt.W[l];
            const ACC x = (ACC)loc[(row * LP + s) * 2] * (ACC)W - (ACC)0.5;
            const ACC y = (ACC)loc[(row * LP + s) * 2 + 1] * (ACC)H - (ACC)0.5;
            const ACC a = (ACC)attn[row * LP + s];
            ACC sa = 0, sx = 0, sy = 0;
            if (y > (ACC)-1 && x > (ACC)-1 && y < (ACC)H && x < (ACC)W) {
                const ACC fy = floor(y), fx = floor(x);
                const int y0 = (int)fy, x0 = (int)fx;
                const ACC lh = y - fy, lw = x - fx, hh = (ACC)1 - lh, hw = (ACC)1 - lw;
                const size_t pix = (size_t)lt.start[l] + (size_t)y0 * W + x0;
                const bool ok[4] = {y0 >= 0 && x0 >= 0, y0 >= 0 && x0 + 1 <= W - 1,
                                    y0 + 1 <= H - 1 && x0 >= 0, y0 + 1 <= H - 1 && x0 + 1 <= W - 1};
                const size_t off[4] = {pix * rs, (pix + 1) * rs, (pix + W) * rs, (pix + W + 1) * rs};
                const ACC w[4] = {hh * hw, hh * lw, lh * hw, lh * lw};
                for (int c = lane; c < d.D; c += 32) {
                    const ACC g = (ACC)grad_out[row * d.D + c];
                    ACC v[4];
                    for (int k = 0; k < 4; ++k) v[k] = ok[k] ? (ACC)value[bbase + off[k] + c] : (ACC)0;
                    sa += g * (w[0] * v[0] + w[1] * v[1] + w[2] * v[2] + w[3] * v[3]);
                    sx += g * (hh * (v[1] - v[0]) + lh * (v[3] - v[2]));
                    sy += g * (hw * (v[2] - v[0]) + lw * (v[3] - v[1]));
                    for (int k = 0; k < 4; ++k)
                        if (ok[k]) atomic_add_any<ACC>(grad_value_acc + bbase + off[k] + c, w[k] * a * g);
                }
            }
            for (int o = 16; o > 0; o >>= 1) {
                sa += __shfl_xor_sync(0xffffffffu, sa, o);
                sx += __shfl_xor_sync(0xffffffffu, sx, o);
                sy += __shfl_xor_sync(0xffffffffu, sy, o);
            }
            if (lane == 0) {
                grad_loc[(row * LP + s) * 2] = (AT)(sx * a * (ACC)W);
                grad_loc[(row * LP + s) * 2 + 1] = (AT)(sy * a * (ACC)H);
                grad_attn[row * LP + s] = (AT)sa;
            }
        }
    }
}

// ACC workspace -> VT grad_value for the generic path (scalar).
template <typename VT, typename ACC>
__global__ void __launch_bounds__(256)
msda_round_generic_kernel(const ACC* __restrict__ acc, VT* __restrict__ gv, size_t n)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (size_t)gridDim.x * blockDim.x)
        gv[i] = (VT)acc[i];
}

}  // namespace msda
