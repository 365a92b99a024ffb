// umma_probe.cu -- checks, on a real B200, the tcgen05 encodings the dense-window MSDeformAttn kernels
// rely on (development tool; not part of the product library):
//   * shared-memory matrix descriptors for K-major, SWIZZLE_NONE operands with free LBO / SBO
//     (core matrix = 8 rows x 16 bytes; LBO = stride between the two 16-byte K chunks of one MMA,
//     SBO = stride between 8-row groups), including padded strides (LBO = 144);
//   * the instruction descriptor (M = 128, N = 32 / 128, fp32 accumulate, bf16 x bf16 and f16 x bf16);
//   * accumulator layout in TMEM for M = 128 (row i -> lane i, column j -> column j) and M = 64;
//   * tcgen05.ld 32x32b.x32, tcgen05.commit -> mbarrier, alloc / dealloc;
// and times the pieces (cycles): an MMA batch, TMEM -> register loads, register -> shared staging.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_probe umma_probe.cu && ./umma_probe <test>
// Every wait is bounded, so a wrong encoding reports FAIL instead of hanging the GPU.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__); exit(2); } } while (0)

struct Params {
    uint32_t a_bytes, b_bytes;        // operand images (copied verbatim into shared memory)
    uint32_t a_lbo, a_sbo, b_lbo, b_sbo;
    uint32_t a_kstep, b_kstep;        // start-address advance per MMA (bytes)
    uint32_t ksteps;                  // MMAs per batch (first one overwrites, the rest accumulate)
    uint32_t idesc;
    uint32_t n;                       // accumulator columns to read back (multiple of 32)
    uint32_t m;                       // 128 or 64
    uint32_t repeat;                  // timing: the batch is issued `repeat` times before one commit
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo)
{
    uint64_t d = 0;
    d |= (uint64_t)((addr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;           // descriptor version (Blackwell)
    return d;                          // base offset 0, lbo mode 0, layout type 0 = SWIZZLE_NONE
}

__device__ __forceinline__ bool mbar_wait_bounded(uint32_t bar, uint32_t parity, int max_tries)
{
    for (int i = 0; i < max_tries; ++i) {
        uint32_t ok;
        asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (ok) return true;
    }
    return false;
}

extern __shared__ __align__(1024) uint8_t smem[];

__global__ void __launch_bounds__(128, 1)
probe_kernel(const uint8_t* __restrict__ a_img, const uint8_t* __restrict__ b_img, float* __restrict__ d_out,
             long long* __restrict__ cycles, int* __restrict__ status, Params p)
{
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(8) uint64_t mbar;
    uint8_t* sa = smem;
    uint8_t* sb = smem + ((p.a_bytes + 1023) & ~1023u);
    const int tid = threadIdx.x, warp = tid >> 5;

    for (uint32_t i = tid * 16; i < p.a_bytes; i += 128 * 16) *reinterpret_cast<uint4*>(sa + i) = *reinterpret_cast<const uint4*>(a_img + i);
    for (uint32_t i = tid * 16; i < p.b_bytes; i += 128 * 16) *reinterpret_cast<uint4*>(sb + i) = *reinterpret_cast<const uint4*>(b_img + i);

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(&tmem_base_s)), "r"(512u));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(&mbar)), "r"(1u));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy smem writes -> tensor core
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base_s;

    long long t0 = 0, t1 = 0, t2 = 0, t3 = 0;
    if (tid == 0) {
        t0 = clock64();
        for (uint32_t r = 0; r < p.repeat; ++r) {
            for (uint32_t k = 0; k < p.ksteps; ++k) {
                const uint64_t da = make_desc(smem_u32(sa) + k * p.a_kstep, p.a_lbo, p.a_sbo);
                const uint64_t db = make_desc(smem_u32(sb) + k * p.b_kstep, p.b_lbo, p.b_sbo);
                const uint32_t acc = k > 0 ? 1u : 0u;
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                             "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                             :: "r"(tmem), "l"(da), "l"(db), "r"(p.idesc), "r"(acc) : "memory");
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&mbar)) : "memory");
        t1 = clock64();
    }
    const bool ok = mbar_wait_bounded(smem_u32(&mbar), 0, 1 << 20);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (tid == 0) t2 = clock64();
    if (!ok) { if (tid == 0) status[0] = 1; }

    // read back: warp w reads TMEM lanes 32w .. 32w+31, 32 columns at a time; stage through shared memory
    // the way the kernel will (thread-private rows, 16-byte stores) to time that path too
    float* stage = reinterpret_cast<float*>(sb + ((p.b_bytes + 1023) & ~1023u));   // [128][36]
    float sum = 0.f;
    for (uint32_t c0 = 0; c0 < p.n; c0 += 32) {
        uint32_t v[32];
        const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + c0;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32"
                     "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15,"
                     "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                       "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                       "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                       "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                     : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
        for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<uint4*>(stage + tid * 36 + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
#pragma unroll
        for (int j = 0; j < 32; ++j) d_out[(size_t)tid * p.n + c0 + j] = __uint_as_float(v[j]);
        sum += stage[tid * 36 + (tid & 31)];
    }
    if (tid == 0) t3 = clock64();
    // timing only: 16 passes of (TMEM load of 32 columns, 8 16-byte stores to a private shared-memory row)
    __syncthreads();
    long long t4 = clock64();
    for (int it = 0; it < 16; ++it) {
        for (uint32_t c0 = 0; c0 < 128; c0 += 32) {
            uint32_t v[32];
            const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + c0;
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32"
                         "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15,"
                         "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                         : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                           "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                           "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                           "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                         : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (p.repeat != 7777) {
#pragma unroll
                for (int j = 0; j < 32; j += 4)
                    *reinterpret_cast<uint4*>(stage + tid * 36 + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) sum += __uint_as_float(v[j]);
            }
        }
    }
    __syncthreads();
    long long t5 = clock64();
    if (tid == 0) cycles[3] = t5 - t4;
    if (sum == 123.456f) status[1] = 1;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(512u));
    if (tid == 0) { cycles[0] = t1 - t0; cycles[1] = t2 - t0; cycles[2] = t3 - t2; }
}

// ---- host: K-major SWIZZLE_NONE image of a [rows][k] 16-bit matrix -----------------------------------------
static void put16(std::vector<uint8_t>& img, size_t off, uint16_t v) { if (off + 2 > img.size()) img.resize(off + 2, 0); memcpy(&img[off], &v, 2); }
static size_t kmajor_off(int r, int k, uint32_t lbo, uint32_t sbo) { return (size_t)(r % 8) * 16 + (size_t)(r / 8) * sbo + (size_t)(k / 8) * lbo + (size_t)(k % 8) * 2; }
static uint16_t bf16_bits(float f) { __nv_bfloat16 h = __float2bfloat16(f); uint16_t b; memcpy(&b, &h, 2); return b; }
static uint16_t f16_bits(float f) { __half h = __float2half(f); uint16_t b; memcpy(&b, &h, 2); return b; }

static uint32_t make_idesc(int M, int N, int a_fmt, int b_fmt)
{
    uint32_t d = 0;
    d |= 1u << 4;                      // accumulator format F32
    d |= (uint32_t)a_fmt << 7;         // 0 = F16, 1 = BF16
    d |= (uint32_t)b_fmt << 10;
    d |= (uint32_t)(N >> 3) << 17;
    d |= (uint32_t)(M >> 4) << 24;
    return d;                          // K-major A and B, no negate, dense
}

int main(int argc, char** argv)
{
    const int test = argc > 1 ? atoi(argv[1]) : 0;
    // test 0: M=128 N=128 K=32  bf16 x bf16, LBO=128 SBO=512                      (dot products U = G V^T)
    // test 1: M=128 N=32  K=128 f16 x bf16, padded LBO=144, SBO=2320 / 2304      (window scatter W^T G)
    // test 2: as 1 with bf16 x bf16
    // test 3: as 0 with LBO / SBO swapped (expected FAIL: tells the two interpretations apart)
    // test 4: M=64 N=32 K=128 f16 x bf16 (accumulator rows -> lanes (r%16) + 32*(r/16))
    // test 5: timing, test 1's batch issued 64 times;  test 6: timing, test 0's batch issued 64 times
    int M = 128, N = 128, K = 32, a_fmt = 1, b_fmt = 1, repeat = 1;
    uint32_t a_lbo = 128, a_sbo = 512, b_lbo = 128, b_sbo = 512;
    bool swap = false;
    if (test == 1 || test == 2 || test == 4 || test == 5) {
        N = 32; K = 128; a_fmt = (test == 2) ? 1 : 0;
        a_lbo = 144; a_sbo = 16 * 144 + 16; b_lbo = 144; b_sbo = 16 * 144;
        if (test == 4) M = 64;
        if (test == 5) repeat = 64;
    }
    if (test == 3) swap = true;
    if (test == 6) repeat = 64;
    // test 8 / 9: as test 2 (M=128 N=32 K=128, the scatter product), but B is MN-major: the image of G as the dot
    // product stages it ([K = 128 query rows][N = 32 channels], core matrix = 8 rows x 16 bytes, chunk stride 160, row
    // group stride 640) with idesc bit 16 (b_major) set -- would make the transposed copy Gt unnecessary.
    // 8: LBO = chunk stride (N direction), SBO = row-group stride (K direction); 9: the two swapped.
    bool b_mn = false;
    if (test == 8 || test == 9) {
        N = 32; K = 128; a_lbo = 144; a_sbo = 16 * 144 + 16; b_lbo = 160; b_sbo = 640; b_mn = true;
    }
    if (test == 7) {   // free form: 7 M N K a_fmt b_fmt repeat [a_lbo a_sbo b_lbo b_sbo]
        M = atoi(argv[2]); N = atoi(argv[3]); K = atoi(argv[4]); a_fmt = atoi(argv[5]); b_fmt = atoi(argv[6]); repeat = atoi(argv[7]);
        a_lbo = b_lbo = 128; a_sbo = b_sbo = (K / 8) * 128;
        if (argc > 11) { a_lbo = atoi(argv[8]); a_sbo = atoi(argv[9]); b_lbo = atoi(argv[10]); b_sbo = atoi(argv[11]); }
    }

    std::vector<float> A((size_t)M * K), B((size_t)N * K);
    srand(1234 + test);
    for (auto& v : A) v = (float)((rand() % 17) - 8) * 0.125f;
    for (auto& v : B) v = (float)((rand() % 17) - 8) * 0.25f;
    std::vector<uint8_t> ia, ib;
    for (int r = 0; r < M; ++r) for (int k = 0; k < K; ++k)
        put16(ia, kmajor_off(r, k, a_lbo, a_sbo), a_fmt ? bf16_bits(A[(size_t)r * K + k]) : f16_bits(A[(size_t)r * K + k]));
    for (int r = 0; r < N; ++r) for (int k = 0; k < K; ++k)
        put16(ib, b_mn ? kmajor_off(k, r, b_lbo, b_sbo) : kmajor_off(r, k, b_lbo, b_sbo),
              b_fmt ? bf16_bits(B[(size_t)r * K + k]) : f16_bits(B[(size_t)r * K + k]));
    ia.resize((ia.size() + 15) & ~15ull, 0);
    ib.resize((ib.size() + 15) & ~15ull, 0);

    Params p;
    p.a_bytes = (uint32_t)ia.size(); p.b_bytes = (uint32_t)ib.size();
    p.a_lbo = swap ? a_sbo : a_lbo; p.a_sbo = swap ? a_lbo : a_sbo;
    p.b_lbo = swap ? b_sbo : b_lbo; p.b_sbo = swap ? b_lbo : b_sbo;
    p.a_kstep = 2 * a_lbo; p.b_kstep = 2 * b_lbo;
    p.ksteps = K / 16;
    p.idesc = make_idesc(M, N, a_fmt, b_fmt);
    if (b_mn) {
        p.idesc |= 1u << 16;             // B is MN-major
        p.b_kstep = 2 * b_sbo;           // 16 K rows = two row groups
        if (test == 9) { p.b_lbo = b_sbo; p.b_sbo = b_lbo; }
    }
    p.n = (uint32_t)N; p.m = (uint32_t)M; p.repeat = (uint32_t)repeat;

    uint8_t *da, *db; float* dd; long long* dc; int* ds;
    CK(cudaMalloc(&da, ia.size())); CK(cudaMalloc(&db, ib.size()));
    CK(cudaMalloc(&dd, (size_t)128 * N * 4)); CK(cudaMalloc(&dc, 64)); CK(cudaMalloc(&ds, 64));
    CK(cudaMemcpy(da, ia.data(), ia.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(db, ib.data(), ib.size(), cudaMemcpyHostToDevice));
    CK(cudaMemset(dd, 0, (size_t)128 * N * 4)); CK(cudaMemset(dc, 0, 64)); CK(cudaMemset(ds, 0, 64));
    const size_t smem_bytes = ((ia.size() + 1023) & ~1023ull) + ((ib.size() + 1023) & ~1023ull) + 128 * 36 * 4 + 1024;
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    probe_kernel<<<1, 128, smem_bytes>>>(da, db, dd, dc, ds, p);
    CK(cudaDeviceSynchronize());
    std::vector<float> D((size_t)128 * N);
    long long cyc[4]; int st[2];
    CK(cudaMemcpy(D.data(), dd, D.size() * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(cyc, dc, sizeof(cyc), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(st, ds, sizeof(st), cudaMemcpyDeviceToHost));

    double max_err = 0; int bad = 0;
    for (int r = 0; r < M; ++r) {
        const int lane = (M == 128) ? r : (r % 16) + 32 * (r / 16);
        for (int n = 0; n < N; ++n) {
            double ref = 0;
            for (int k = 0; k < K; ++k) ref += (double)A[(size_t)r * K + k] * B[(size_t)n * K + k];
            const double err = fabs(ref - D[(size_t)lane * N + n]);
            if (err > max_err) max_err = err;
            if (err > 1e-3) ++bad;
        }
    }
    printf("test %d: M=%d N=%d K=%d a_fmt=%d b_fmt=%d lbo/sbo A %u/%u B %u/%u  barrier %s  max_err %.4g  bad %d  -> %s\n", test, M, N, K,
           a_fmt, b_fmt, p.a_lbo, p.a_sbo, p.b_lbo, p.b_sbo, st[0] ? "TIMEOUT" : "ok", max_err, bad,
           (!st[0] && bad == 0) ? "PASS" : "FAIL");
    printf("   cycles: issue %lld, issue+complete %lld (%.1f per MMA), readback+stage of %d columns %lld\n", cyc[0], cyc[1],
           (double)cyc[1] / (repeat * (K / 16)), N, cyc[2]);
    printf("   TMEM->reg->smem staging, 4 warps, 16 x 128 columns: %lld cycles (%.1f per 32-column load+stage per warp)\n", cyc[3], (double)cyc[3] / 64);
    return 0;
}
