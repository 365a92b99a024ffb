// msda_launch.cuh -- host-side launchers: kernel selection from (dtype, D), work-item length and
// grid sizing.  Included by msda_abi.cu (entry points) and by the msda_inst_*.cu files,
// which hold the explicit instantiations so the dtype families compile in parallel.
#pragma once
#include <cstdlib>
#include <cstring>

#include "msda_bwd.cuh"
#include "msda_dense.cuh"
#include "msda_fwd.cuh"

namespace msda {

// per-thread call state, defined in msda_abi.cu
extern thread_local int g_last_launches;
extern thread_local cudaEvent_t g_trace_start, g_trace_stop;

// One-shot event bracket around the dominant kernel (msda_b200_trace_next).
struct TraceScope {
    cudaStream_t st;
    cudaEvent_t stop;
    explicit TraceScope(cudaStream_t s) : st(s), stop(g_trace_stop)
    {
        if (g_trace_start) cudaEventRecord(g_trace_start, st);
        g_trace_start = nullptr;
        g_trace_stop = nullptr;
    }
    ~TraceScope() { if (stop) cudaEventRecord(stop, st); }
};

struct DeviceInfo {
    int sms = 0;
};

inline DeviceInfo device_info()
{
    static thread_local int cached_dev = -1;
    static thread_local DeviceInfo info;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return info;
    if (dev != cached_dev) {
        cudaDeviceGetAttribute(&info.sms, cudaDevAttrMultiProcessorCount, dev);
        cached_dev = dev;
    }
    return info;
}

inline int env_int(const char* name, int dflt)
{
    const char* v = std::getenv(name);
    return (v && *v) ? std::atoi(v) : dflt;
}

inline int check_common(const void* value, const int64_t* shapes, const void* loc, const void* attn,
                 const msda_dims* d)
{
    if (!d) return MSDA_E_NULL;
    if (d->N <= 0 || d->S <= 0 || d->M <= 0 || d->D <= 0 || d->L <= 0 || d->P <= 0 || d->Lq < 0)
        return MSDA_E_SHAPE;
    if (d->L > MSDA_MAX_LEVELS) return MSDA_E_SHAPE;
    // 32-bit element offsets inside one batch item
    if ((long long)d->S * d->M * d->D >= (1ll << 31)) return MSDA_E_SHAPE;
    if ((long long)d->L * d->P >= (1 << 20)) return MSDA_E_SHAPE;
    // 32-bit item / row indices in the kernels (forward_items, backward_items: item = chunk * N * M + b * M + m)
    if ((long long)d->N * d->M * d->Lq >= (1ll << 30)) return MSDA_E_SHAPE;
    if (!value || !shapes) return MSDA_E_NULL;
    if (d->Lq > 0 && (!loc || !attn)) return MSDA_E_NULL;
    return 0;
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Work items (see walk_query in msda_common.cuh).  Encoder form (Lq == S: the queries are the pyramid's
// pixels and their samples cluster around the query's own position): band-ordered walk, so an item is a
// compact band_h x (tile_q / band_h) patch.  Otherwise plain query order.  The item length trades the
// scheduling tail (one item per CTA, handed out by the hardware as SMs free up: shorter items, shorter
// tail) against the per-CTA prologue; it shrinks further when there are too few items to cover every SM.
constexpr int kForwardTileQ = 128;
constexpr int kBackwardTileQ = 128;

inline KernelDims make_dims(const msda_dims* d, int rows_per_pass, int default_tile_q)
{
    KernelDims k;
    k.N = d->N; k.S = d->S; k.M = d->M; k.D = d->D; k.L = d->L; k.Lq = d->Lq; k.P = d->P;
    k.acc16 = 0;
    const int sms = device_info().sms > 0 ? device_info().sms : 148;
    k.band_h = (d->Lq == d->S && !env_int("MSDA_B200_TILE_1D", 0)) ? env_int("MSDA_B200_TILE_H", 8) : 0;
    if (k.band_h < 0) k.band_h = 0;
    int tq = env_int("MSDA_B200_TILE_Q", 0);
    if (tq <= 0) {
        tq = default_tile_q;
        // shrink until there are at least ~4 items per SM (or one pass per item)
        while (tq > rows_per_pass && (long long)d->N * d->M * ((d->Lq + tq - 1) / tq) < 4ll * sms) tq >>= 1;
        // whole passes: a ragged last pass leaves warps of the CTA without rows (D = 36: 24 rows per pass)
        tq = tq / rows_per_pass * rows_per_pass;
    }
    if (tq < rows_per_pass) tq = rows_per_pass;
    k.tile_q = tq;
    return k;
}

// Generic kernels: a persistent grid of warps striding over the rows.
template <typename Kern>
int query_occupancy(Kern kern, int threads)
{
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, threads, 0);
    return occ < 1 ? 1 : occ;
}

inline int persistent_grid(int occ, long long items_upper)
{
    const int sms = device_info().sms > 0 ? device_info().sms : 148;
    long long g = (long long)sms * occ;
    if (items_upper < g) g = items_upper;
    return (int)(g < 1 ? 1 : g);
}

// One CTA per item; gridDim.x is capped only by what the hardware accepts (the kernels stride).
inline int item_grid(const KernelDims& k)
{
    const long long items = (long long)((k.Lq + k.tile_q - 1) / k.tile_q) * k.N * k.M;
    const long long cap = 1ll << 30;
    return (int)(items < 1 ? 1 : (items > cap ? cap : items));
}

// fa == nullptr: loc / attn are materialised sampling locations / attention weights;
// otherwise they are raw offsets / logits and *fa carries the rest of the module prologue's inputs.
template <typename VT, typename AT, int VEC, int GS, bool FUSED, int DCH = VEC * GS>
int launch_forward_vec_impl(const void* value, const int64_t* shapes, const int64_t* ls, const void* loc,
                            const void* attn, void* out, const msda_dims* d, cudaStream_t st, const FusedArgs& fa)
{
    constexpr int NW = 8;
    const KernelDims k = make_dims(d, NW * (32 / GS), kForwardTileQ);
    auto kern = msda_forward_kernel<VT, AT, VEC, GS, NW, FUSED, DCH>;
    const int grid = item_grid(k);
    {
        TraceScope trace(st);
        kern<<<grid, NW * 32, 0, st>>>((const VT*)value, shapes, ls, (const AT*)loc, (const AT*)attn,
                                       (VT*)out, k, fa);
    }
    ++g_last_launches;
    return (int)cudaGetLastError();
}

template <typename VT, typename AT, int VEC, int GS, int DCH = VEC * GS>
int launch_forward_vec(const void* value, const int64_t* shapes, const int64_t* ls, const void* loc,
                       const void* attn, void* out, const msda_dims* d, cudaStream_t st, const FusedArgs* fa)
{
    if (fa) {
        if (d->L * d->P > kFusedMaxChunks * chunk_of(GS)) return MSDA_E_SHAPE;   // a lane keeps <= 4 softmax numerators per row
        return launch_forward_vec_impl<VT, AT, VEC, GS, true, DCH>(value, shapes, ls, loc, attn, out, d, st, *fa);
    }
    return launch_forward_vec_impl<VT, AT, VEC, GS, false, DCH>(value, shapes, ls, loc, attn, out, d, st, FusedArgs{});
}

template <typename VT, typename AT, typename ACC>
int launch_forward_generic(const void* value, const int64_t* shapes, const int64_t* ls,
                           const void* loc, const void* attn, void* out, const msda_dims* d,
                           cudaStream_t st)
{
    KernelDims k = make_dims(d, 8, 8);
    k.band_h = 0; k.tile_q = 8;
    const long long rows = (long long)d->N * d->Lq * d->M;
    auto kern = msda_forward_generic_kernel<VT, AT, ACC>;
    static const int occ = query_occupancy(kern, 256);
    const int grid = persistent_grid(occ, (rows + 7) / 8);
    kern<<<grid, 256, 0, st>>>((const VT*)value, shapes, ls, (const AT*)loc, (const AT*)attn,
                               (VT*)out, k);
    ++g_last_launches;
    return (int)cudaGetLastError();
}

// MSDA_E_SHAPE from a fused call = "no fused kernel for this shape": the caller runs the prologue
// itself and uses the plain entry point.
template <typename VT, typename AT>
int forward_typed(const void* value, const int64_t* shapes, const int64_t* ls, const void* loc,
                  const void* attn, void* out, const msda_dims* d, cudaStream_t st, const FusedArgs* fa)
{
    const bool al = aligned16(value) && aligned16(out) && !env_int("MSDA_B200_FORCE_GENERIC", 0);
    if constexpr (sizeof(VT) == 4) {
        if (al) switch (d->D) {
            case 16: return launch_forward_vec<VT, AT, 4, 4>(value, shapes, ls, loc, attn, out, d, st, fa);
            case 32: return launch_forward_vec<VT, AT, 4, 8>(value, shapes, ls, loc, attn, out, d, st, fa);
            case 36: return launch_forward_vec<VT, AT, 4, 9>(value, shapes, ls, loc, attn, out, d, st, fa);
            case 64: return launch_forward_vec<VT, AT, 4, 16>(value, shapes, ls, loc, attn, out, d, st, fa);
            default: break;
        }
    } else {
        if (al) switch (d->D) {
            case 16: return launch_forward_vec<VT, AT, 8, 2>(value, shapes, ls, loc, attn, out, d, st, fa);
            case 32: return launch_forward_vec<VT, AT, 8, 4>(value, shapes, ls, loc, attn, out, d, st, fa);
            case 36:
                // 5 lanes of 8 channels on 16-byte windows (msda_fwd.cuh, "overhang") when the heads come in pairs, fused,
                // the row's samples fit four chunks of 4, and there are enough rows to fill the machine with 48-row passes
                // (C3-enc 109 -> 97 us; the 560-query decoder form is latency-bound and loses with half as many CTAs:
                // 20.6 -> 26.6 us); else 9 lanes of 4 channels
                if ((d->M & 1) == 0 && (!fa || d->L * d->P <= kFusedMaxChunks * chunk_of(5)) &&
                    ((long long)d->N * d->M * d->Lq >= 48ll * 8 * (device_info().sms > 0 ? device_info().sms : 148) ||
                     env_int("MSDA_B200_D36_WIDE", 0)) &&
                    !env_int("MSDA_B200_D36_NARROW", 0))
                    return launch_forward_vec<VT, AT, 8, 5, 36>(value, shapes, ls, loc, attn, out, d, st, fa);
                return launch_forward_vec<VT, AT, 4, 9>(value, shapes, ls, loc, attn, out, d, st, fa);
            case 64: return launch_forward_vec<VT, AT, 8, 8>(value, shapes, ls, loc, attn, out, d, st, fa);
            default: break;
        }
    }
    if (fa) return MSDA_E_SHAPE;
    return launch_forward_generic<VT, AT, float>(value, shapes, ls, loc, attn, out, d, st);
}

// dense_done != nullptr: split mode, after the dense-window kernel (see backward_items); the grid then also covers
// the tile items.  traced: this launch is the one msda_b200_trace_next brackets.
template <typename VT, typename AT, int VEC, int GS, bool FUSED>
int launch_backward_vec_impl(const void* value, const int64_t* shapes, const int64_t* ls, const void* loc,
                             const void* attn, const void* go, float* gv_acc, void* gl, void* ga,
                             const msda_dims* d, cudaStream_t st, const FusedArgs& fa,
                             const uint8_t* dense_done = nullptr, int tiles_bound = 0, bool traced = true,
                             int* split_claim = nullptr, bool acc16 = false)
{
    constexpr int NW = 8;
    KernelDims k = make_dims(d, NW * (32 / GS), kBackwardTileQ);
    k.acc16 = acc16 ? 1 : 0;
    auto kern = msda_backward_kernel<VT, AT, VEC, GS, NW, FUSED>;
    const int sms = device_info().sms > 0 ? device_info().sms : 148;
    // split mode: a persistent grid claims (64-row chunks of the coarse levels' queries, then the flagged tiles)
    if (dense_done) k.tile_q = 64 / (NW * (32 / GS)) * (NW * (32 / GS)) > 0 ? 64 / (NW * (32 / GS)) * (NW * (32 / GS)) : NW * (32 / GS);
    const int grid = dense_done ? MSDA_BWD_MINB * sms : item_grid(k);
    if (traced) {
        TraceScope trace(st);
        kern<<<grid, NW * 32, 0, st>>>((const VT*)value, shapes, ls, (const AT*)loc, (const AT*)attn,
                                       (const VT*)go, gv_acc, (AT*)gl, (AT*)ga, k, fa, dense_done, tiles_bound, split_claim);
    } else if (dense_done && env_int("MSDA_B200_PDL", 1)) {
        // Programmatic dependent launch behind the dense-window kernel, which releases its dependents as soon as it
        // starts: CTAs of this grid become resident as dense CTAs retire and do the coarse levels' queries (independent
        // of the dense kernel: commutative reductions, rows of their own) while the last dense items finish; they wait
        // for the dense kernel (griddepcontrol.wait) only before they read its done-item flags.
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid);
        cfg.blockDim = dim3(NW * 32);
        cfg.dynamicSmemBytes = 0;
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        cudaError_t e = cudaLaunchKernelEx(&cfg, kern, (const VT*)value, shapes, ls, (const AT*)loc, (const AT*)attn,
                                           (const VT*)go, gv_acc, (AT*)gl, (AT*)ga, k, fa, dense_done, tiles_bound, split_claim);
        ++g_last_launches;
        return (int)e;
    } else {
        kern<<<grid, NW * 32, 0, st>>>((const VT*)value, shapes, ls, (const AT*)loc, (const AT*)attn,
                                       (const VT*)go, gv_acc, (AT*)gl, (AT*)ga, k, fa, dense_done, tiles_bound, split_claim);
    }
    ++g_last_launches;
    return (int)cudaGetLastError();
}

template <typename VT, typename AT, int VEC, int GS>
int launch_backward_vec(const void* value, const int64_t* shapes, const int64_t* ls, const void* loc,
                        const void* attn, const void* go, float* gv_acc, void* gl, void* ga,
                        const msda_dims* d, cudaStream_t st, const FusedArgs* fa, bool acc16 = false)
{
    if (fa) {
        if (d->L * d->P > kFusedMaxChunks * chunk_of(GS)) return MSDA_E_SHAPE;   // a lane keeps <= 4 samples of a row in registers
        return launch_backward_vec_impl<VT, AT, VEC, GS, true>(value, shapes, ls, loc, attn, go, gv_acc, gl, ga, d, st, *fa,
                                                               nullptr, 0, true, nullptr, acc16);
    }
    return launch_backward_vec_impl<VT, AT, VEC, GS, false>(value, shapes, ls, loc, attn, go, gv_acc, gl, ga, d, st,
                                                            FusedArgs{}, nullptr, 0, true, nullptr, acc16);
}

template <typename VT, typename AT, typename ACC>
int launch_backward_generic(const void* value, const int64_t* shapes, const int64_t* ls,
                            const void* loc, const void* attn, const void* go, ACC* gv_acc,
                            void* gl, void* ga, const msda_dims* d, cudaStream_t st)
{
    KernelDims k = make_dims(d, 8, 8);
    k.band_h = 0; k.tile_q = 8;
    const long long rows = (long long)d->N * d->Lq * d->M;
    auto kern = msda_backward_generic_kernel<VT, AT, ACC>;
    static const int occ = query_occupancy(kern, 256);
    const int grid = persistent_grid(occ, (rows + 7) / 8);
    kern<<<grid, 256, 0, st>>>((const VT*)value, shapes, ls, (const AT*)loc, (const AT*)attn,
                               (const VT*)go, gv_acc, (AT*)gl, (AT*)ga, k);
    ++g_last_launches;
    return (int)cudaGetLastError();
}

// 16-bit values and a sparse query set (the decoder form: a few hundred queries over a 20k-pixel pyramid): with at
// most ~1 corner per pixel row on average (4 * Lq * L * P <= S) accumulating in the 16-bit output itself loses nothing
// against the stated 16-bit tolerance, and the call needs no fp32 workspace, zero-fills half the bytes and has no
// rounding pass.  Denser query sets keep fp32 accumulation with one final rounding.
inline bool sparse_16bit_regime(const msda_dims* d)
{
    return 4ll * d->Lq * d->L * d->P <= (long long)d->S && (d->D == 16 || d->D == 32 || d->D == 36 || d->D == 64) &&
           env_int("MSDA_B200_ACC16", 1) && !env_int("MSDA_B200_FORCE_GENERIC", 0);
}

// Encoder-form backward on the tensor cores (msda_dense.cuh) + the per-row kernel in split mode for the rest.
// Behind the fp32 accumulator the caller's workspace holds the dense kernel's work counter and the dense_done-item
// flags (one byte per (tile, batch, head)); one memset zeroes all three.
inline size_t dense_counter_offset(size_t nval) { return (nval * sizeof(float) + 255) & ~(size_t)255; }
constexpr size_t kDenseCounterBytes = 256;
inline int dense_tiles_bound(const msda_dims* d) { return d->Lq / 32 + 64; }
inline size_t dense_extra_bytes(const msda_dims* d)
{
    return kDenseCounterBytes + (((size_t)dense_tiles_bound(d) * d->N * d->M + 255) & ~(size_t)255);
}

template <typename VT, typename AT, int D, bool FUSED>
int launch_backward_dense(const void* value, const int64_t* shapes, const int64_t* ls, const void* loc,
                          const void* attn, const void* go, float* gv_acc, void* gl, void* ga, int* claim,
                          uint8_t* dense_done, const msda_dims* d, cudaStream_t st, const FusedArgs& fa)
{
    KernelDims k = make_dims(d, 16, kBackwardTileQ);
    auto kern = dense::msda_backward_dense_kernel<VT, AT, D, FUSED>;
    constexpr int smem = (int)dense::Plan<D>::TOTAL;
    static thread_local int configured_dev = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (configured_dev != dev) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return (int)e;
        configured_dev = dev;
    }
    const int sms = device_info().sms > 0 ? device_info().sms : 148;
    dense::DenseArgs da;
    da.claim = claim;
    da.done = dense_done;
    da.tiles_bound = dense_tiles_bound(d);
    TraceScope trace(st);       // the group of launches is "the backward kernel" of msda_b200_trace_next
    kern<<<2 * sms, dense::kThreads, smem, st>>>((const VT*)value, shapes, ls, (const AT*)loc, (const AT*)attn,
                                                 (const VT*)go, gv_acc, (AT*)gl, (AT*)ga, k, fa, da);
    ++g_last_launches;
    if (cudaError_t e = cudaGetLastError()) return (int)e;
    // the queries of the coarser levels, and whatever the dense kernel handed back
    return launch_backward_vec_impl<VT, AT, 4, D / 4, FUSED>(value, shapes, ls, loc, attn, go, gv_acc, gl, ga, d, st, fa,
                                                             dense_done, da.tiles_bound, false, claim + 1);
}

// VT = float: gradients accumulate straight into grad_value (zero-filled first).
// 16-bit VT: accumulate in the fp32 workspace, then round once.
template <typename VT, typename AT>
int backward_typed(const void* value, const int64_t* shapes, const int64_t* ls, const void* loc,
                   const void* attn, const void* go, void* gv, void* gl, void* ga, void* ws,
                   size_t ws_bytes, const msda_dims* d, cudaStream_t st, const FusedArgs* fa)
{
    const size_t nval = (size_t)d->N * d->S * d->M * d->D;
    float* acc;
    bool acc16 = false;
    if constexpr (sizeof(VT) == 4) {
        acc = (float*)gv;
    } else {
        acc16 = sparse_16bit_regime(d) && aligned16(value) && aligned16(go) && aligned16(gv) &&
                (!fa || d->L * d->P <= kFusedMaxChunks * chunk_of(d->D / 4));
        if (acc16) {
            acc = (float*)gv;       // the kernels reinterpret it (KernelDims::acc16)
        } else {
            if (!ws || ws_bytes < nval * sizeof(float)) return MSDA_E_WORKSPACE;
            if (!aligned16(ws)) return MSDA_E_ALIGN;
            acc = (float*)ws;
        }
    }
    // dense-window tensor-core path: encoder form, 16-bit values, D = 32 / 36 / 64, and a workspace with room for
    // the work counter (msda_b200_backward_workspace_bytes)
    bool dense_ok = false;
    if constexpr (sizeof(VT) == 2) {
        dense_ok = !acc16 && d->Lq == d->S && d->Lq > 0 && d->P <= dense::kMaxP && (d->D == 32 || d->D == 36 || d->D == 64) &&
                   (!fa || d->L * d->P <= kFusedMaxChunks * chunk_of(d->D / 4)) &&     // the per-row kernel's fused limit
                   ws_bytes >= dense_counter_offset(nval) + dense_extra_bytes(d) && aligned16(value) && aligned16(go) &&
                   aligned16(loc) && aligned16(attn) && aligned16(gl) && aligned16(ga) &&
                   env_int("MSDA_B200_DENSE", 1) && !env_int("MSDA_B200_FORCE_GENERIC", 0) && !env_int("MSDA_B200_TILE_1D", 0);
    }
    cudaError_t e = cudaMemsetAsync(acc, 0, dense_ok ? dense_counter_offset(nval) + dense_extra_bytes(d)
                                                     : nval * (acc16 ? sizeof(VT) : sizeof(float)), st);
    if (e != cudaSuccess) return (int)e;
    int rc = 0;
    if (d->Lq > 0) {
        const bool al = aligned16(value) && aligned16(go) && aligned16(acc) &&
                        !env_int("MSDA_B200_FORCE_GENERIC", 0);
        bool done = false;
        if constexpr (sizeof(VT) == 2) {
            if (dense_ok) {
                int* claim = reinterpret_cast<int*>(reinterpret_cast<char*>(ws) + dense_counter_offset(nval));
                uint8_t* dense_done = reinterpret_cast<uint8_t*>(claim) + kDenseCounterBytes;
                if (fa) {
                    if (d->D == 32) rc = launch_backward_dense<VT, AT, 32, true>(value, shapes, ls, loc, attn, go, acc, gl, ga, claim, dense_done, d, st, *fa);
                    else if (d->D == 36) rc = launch_backward_dense<VT, AT, 36, true>(value, shapes, ls, loc, attn, go, acc, gl, ga, claim, dense_done, d, st, *fa);
                    else rc = launch_backward_dense<VT, AT, 64, true>(value, shapes, ls, loc, attn, go, acc, gl, ga, claim, dense_done, d, st, *fa);
                } else {
                    if (d->D == 32) rc = launch_backward_dense<VT, AT, 32, false>(value, shapes, ls, loc, attn, go, acc, gl, ga, claim, dense_done, d, st, FusedArgs{});
                    else if (d->D == 36) rc = launch_backward_dense<VT, AT, 36, false>(value, shapes, ls, loc, attn, go, acc, gl, ga, claim, dense_done, d, st, FusedArgs{});
                    else rc = launch_backward_dense<VT, AT, 64, false>(value, shapes, ls, loc, attn, go, acc, gl, ga, claim, dense_done, d, st, FusedArgs{});
                }
                done = true;
            }
        }
        // 4 channels per lane for every value type (16-bit rows are read with 8-byte loads): a lane's
        // reduction is then 16 contiguous bytes of the fp32 accumulator row and one instruction covers
        // whole 128-byte rows, i.e. the fewest SM->L2 request packets per byte
        if (al && !done) switch (d->D) {
            case 16: rc = launch_backward_vec<VT, AT, 4, 4>(value, shapes, ls, loc, attn, go, acc, gl, ga, d, st, fa, acc16); done = true; break;
            case 32: rc = launch_backward_vec<VT, AT, 4, 8>(value, shapes, ls, loc, attn, go, acc, gl, ga, d, st, fa, acc16); done = true; break;
            case 36: rc = launch_backward_vec<VT, AT, 4, 9>(value, shapes, ls, loc, attn, go, acc, gl, ga, d, st, fa, acc16); done = true; break;
            case 64: rc = launch_backward_vec<VT, AT, 4, 16>(value, shapes, ls, loc, attn, go, acc, gl, ga, d, st, fa, acc16); done = true; break;
            default: break;
        }
        if (!done) {
            if (fa) return MSDA_E_SHAPE;      // no fused generic kernel: the caller runs the prologue itself
            rc = launch_backward_generic<VT, AT, float>(value, shapes, ls, loc, attn, go, acc, gl, ga, d, st);
        }
        if (rc) return rc;
    }
    if constexpr (sizeof(VT) == 2) {
        if (acc16) return rc;       // accumulated in place
        const int sms = device_info().sms > 0 ? device_info().sms : 148;
        if (nval % 8 == 0 && aligned16(gv)) {
            msda_round_grad_value_kernel<VT><<<sms * 8, 256, 0, st>>>(acc, (VT*)gv, nval / 8);
        } else {
            msda_round_generic_kernel<VT, float><<<sms * 8, 256, 0, st>>>(acc, (VT*)gv, nval);
        }
        ++g_last_launches;
        rc = (int)cudaGetLastError();
    }
    return rc;
}

// fp64 (generic kernels only; parity with the reference's AT_DISPATCH_FLOATING_TYPES), defined in
// msda_inst_f32.cu
int forward_f64(const void* value, const int64_t* shapes, const int64_t* ls, const void* loc,
                const void* attn, void* out, const msda_dims* d, cudaStream_t st);
int backward_f64(const void* value, const int64_t* shapes, const int64_t* ls, const void* loc,
                 const void* attn, const void* go, void* gv, void* gl, void* ga,
                 const msda_dims* d, cudaStream_t st);

#define MSDA_DECLARE_TYPED(EXTERN, VT, AT)                                                                   \
    EXTERN template int forward_typed<VT, AT>(const void*, const int64_t*, const int64_t*, const void*,      \
                                              const void*, void*, const msda_dims*, cudaStream_t,            \
                                              const FusedArgs*);                                             \
    EXTERN template int backward_typed<VT, AT>(const void*, const int64_t*, const int64_t*, const void*,     \
                                               const void*, const void*, void*, void*, void*, void*, size_t, \
                                               const msda_dims*, cudaStream_t, const FusedArgs*);

}  // namespace msda
