// scatter_ceiling.cu -- measures what the B200 gives for the two access patterns that bound
// MSDeformAttn: (a) gathering 128-byte / 64-byte rows that hit in L2/L1, (b) accumulating
// 128-byte fp32 rows into scattered global rows (red.global.add.v4.f32, scalar red, TMA bulk reduce).
// Development tool (speed-of-light numbers for DESIGN.md); not part of the product library.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scatter_ceiling scatter_ceiling.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x;
}

// rows_per_instr groups of GS lanes; each group handles one 16*GS-byte row per step.
// mode 0: red.v4.f32   1: st.v4.f32   2: ld.v4 (gather)   3: red.v4.bf16x2 (16 B = 8 bf16 per lane: the same row
// of channels in HALF the sectors -- the 16-bit accumulator option of DESIGN.md section 10)    rows are `row_bytes` apart
template <int GS, int MODE>
__global__ void __launch_bounds__(256) k_rows(float* buf, uint32_t nrows, int row_floats, int steps, uint32_t seed,
                                              int local_span, float* sink)
{
    const int lane = threadIdx.x & 31;
    const int grp = lane / GS, gl = lane % GS;
    const uint32_t wid = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    float4 acc = make_float4(0, 0, 0, 0);
    uint32_t base_row = hash32(wid * 7919u + seed) % nrows;
    for (int s = 0; s < steps; ++s) {
        uint32_t r;
        if (local_span > 0) r = (base_row + hash32(wid * 131u + s * 32u + grp + seed) % (uint32_t)local_span) % nrows;
        else r = hash32((wid * (uint32_t)steps + s) * 32u + grp + seed) % nrows;
        float* p = buf + (size_t)r * row_floats + gl * 4;
        if (MODE == 0) {
            asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(1.f), "f"(2.f), "f"(3.f), "f"(4.f) : "memory");
        } else if (MODE == 3) {
            const unsigned one2 = 0x3f803f80u;      // bf16x2 {1.0, 1.0}
            asm volatile("red.relaxed.gpu.global.add.noftz.v4.bf16x2 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(one2), "r"(one2), "r"(one2), "r"(one2) : "memory");
        } else if (MODE == 1) {
            *reinterpret_cast<float4*>(p) = make_float4(1.f, 2.f, 3.f, 4.f);
        } else {
            const float4 v = __ldg(reinterpret_cast<const float4*>(p));
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
    }
    if (MODE == 2 && acc.x + acc.y + acc.z + acc.w == 123.456f) sink[0] = acc.x;
}

// scalar red: 32 lanes x 4 B = one 128-byte row per instruction
__global__ void __launch_bounds__(256) k_red_scalar(float* buf, uint32_t nrows, int steps, uint32_t seed)
{
    const int lane = threadIdx.x & 31;
    const uint32_t wid = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    for (int s = 0; s < steps; ++s) {
        const uint32_t r = hash32((wid * (uint32_t)steps + s) + seed) % nrows;
        atomicAdd(buf + (size_t)r * 32 + lane, 1.f);
    }
}

// TMA bulk reduce: each warp keeps a 128-byte row in shared memory and one lane issues
// cp.reduce.async.bulk (add.f32) of that row into a random global row.
__global__ void __launch_bounds__(256) k_bulk_red(float* buf, uint32_t nrows, int steps, uint32_t seed, int bytes)
{
    __shared__ __align__(128) float rows[8][128];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = lane; i < 128; i += 32) rows[warp][i] = 1.f;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    const uint32_t wid = blockIdx.x * (blockDim.x >> 5) + warp;
    const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(&rows[warp][0]);
    const int row_floats = bytes / 4;
    if (lane == 0) {
        for (int s = 0; s < steps; ++s) {
            const uint32_t r = hash32((wid * (uint32_t)steps + s) + seed) % nrows;
            float* p = buf + (size_t)r * row_floats;
            asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;"
                         :: "l"(p), "r"(saddr), "r"(bytes) : "memory");
            if ((s & 7) == 7) {
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory");
            }
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

template <typename F>
float time_ms(F f, int iters = 5)
{
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    f(); CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int i = 0; i < iters; ++i) {
        CK(cudaEventRecord(a)); f(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        float ms; CK(cudaEventElapsedTime(&ms, a, b)); if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

int main()
{
    const size_t bytes = 91ull << 20;           // fp32 grad_value of C2_enc: 2*22223*256*4 = 91 MB
    float* buf; CK(cudaMalloc(&buf, bytes)); CK(cudaMemset(buf, 0, bytes));
    float* sink; CK(cudaMalloc(&sink, 4));
    const int grid = 148 * 8, threads = 256, steps = 2048;
    const double warps = (double)grid * (threads / 32);
    struct { const char* name; int span; } regimes[] = {{"random over 91 MB", 0}, {"local span 2048 rows", 2048}};
    for (auto& rg : regimes) {
        const uint32_t nrows128 = (uint32_t)(bytes / 128), nrows64 = (uint32_t)(bytes / 64);
        float ms;
        ms = time_ms([&] { k_rows<8, 0><<<grid, threads>>>(buf, nrows128, 32, steps, 1u, rg.span, sink); });
        printf("%-22s red.v4.f32 128B rows (4 rows/instr): %8.3f ms  %7.1f GB/s  %6.2f Grow/s\n", rg.name, ms, warps * steps * 4 * 128 / ms / 1e6, warps * steps * 4 / ms / 1e6);
        ms = time_ms([&] { k_rows<8, 1><<<grid, threads>>>(buf, nrows128, 32, steps, 1u, rg.span, sink); });
        printf("%-22s st.v4.f32  128B rows (4 rows/instr): %8.3f ms  %7.1f GB/s\n", rg.name, ms, warps * steps * 4 * 128 / ms / 1e6);
        ms = time_ms([&] { k_rows<8, 2><<<grid, threads>>>(buf, nrows128, 32, steps, 1u, rg.span, sink); });
        printf("%-22s ld.v4      128B rows (4 rows/instr): %8.3f ms  %7.1f GB/s  %6.2f Grow/s\n", rg.name, ms, warps * steps * 4 * 128 / ms / 1e6, warps * steps * 4 / ms / 1e6);
        ms = time_ms([&] { k_rows<4, 2><<<grid, threads>>>(buf, nrows64, 16, steps, 1u, rg.span, sink); });
        printf("%-22s ld.v4       64B rows (8 rows/instr): %8.3f ms  %7.1f GB/s  %6.2f Grow/s\n", rg.name, ms, warps * steps * 8 * 64 / ms / 1e6, warps * steps * 8 / ms / 1e6);
        ms = time_ms([&] { k_rows<4, 0><<<grid, threads>>>(buf, nrows64, 16, steps, 1u, rg.span, sink); });
        printf("%-22s red.v4.f32  64B rows (8 rows/instr): %8.3f ms  %7.1f GB/s  %6.2f Grow/s\n", rg.name, ms, warps * steps * 8 * 64 / ms / 1e6, warps * steps * 8 / ms / 1e6);
        // 32 bf16 channels = one 64-byte row per 4 lanes: what a 16-bit grad_value accumulator would send
        ms = time_ms([&] { k_rows<4, 3><<<grid, threads>>>(buf, nrows64, 16, steps, 1u, rg.span, sink); });
        printf("%-22s red.v4.bf16x2 64B rows (8 rows/instr): %6.3f ms  %7.1f GB/s  %6.2f Grow/s\n", rg.name, ms, warps * steps * 8 * 64 / ms / 1e6, warps * steps * 8 / ms / 1e6);
    }
    {
        const uint32_t nrows128 = (uint32_t)(bytes / 128);
        float ms = time_ms([&] { k_red_scalar<<<grid, threads>>>(buf, nrows128, steps, 3u); });
        printf("%-22s red.f32 scalar 128B row/instr:        %8.3f ms  %7.1f GB/s  %6.2f Grow/s\n", "random over 91 MB", ms, warps * steps * 128 / ms / 1e6, warps * steps / ms / 1e6);
        for (int b : {128, 256, 512}) {
            ms = time_ms([&] { k_bulk_red<<<grid, threads>>>(buf, (uint32_t)(bytes / b), 512, 5u, b); });
            printf("%-22s TMA bulk reduce add.f32 %3d B rows:    %8.3f ms  %7.1f GB/s  %6.2f Grow/s\n", "random over 91 MB", b, ms, warps * 512 * b / ms / 1e6, warps * 512 / ms / 1e6);
        }
    }
    return 0;
}
