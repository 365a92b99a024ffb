// A C++ host program that uses libmsda_b200.so the way a non-Python caller would: include/msda_b200.h, the
// CUDA runtime for memory and the stream, nothing else -- no torch, no Python.  It runs forward and backward
// on a small encoder-form problem (fp32 and bf16 values) and compares with the CPU oracle (oracle/msda_oracle.c,
// linked into THIS TEST ONLY).  Exit code 0 = parity.   Built and run by tests/test_msda_gpu.py.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "msda_b200.h"

extern "C" {
int msda_oracle_forward_f64(const double* value, const int64_t* shapes, const double* loc, const double* attn,
                            double* out, int N, int S, int M, int D, int L, int Lq, int P, int nthreads);
int msda_oracle_backward_f64(const double* value, const int64_t* shapes, const double* loc, const double* attn,
                             const double* grad_out, double* grad_value, double* grad_loc, double* grad_attn,
                             int N, int S, int M, int D, int L, int Lq, int P, int nthreads);
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 2; } } while (0)

static double frand(uint64_t& s) { s = s * 6364136223846793005ull + 1442695040888963407ull; return (double)(s >> 11) / 9007199254740992.0; }

template <typename T> static T* to_device(const std::vector<T>& h)
{
    T* d = nullptr;
    if (cudaMalloc(&d, h.size() * sizeof(T)) != cudaSuccess) return nullptr;
    cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
    return d;
}

static double rel_err(const std::vector<double>& got, const std::vector<double>& ref)
{
    double e = 0, m = 1e-30;
    for (size_t i = 0; i < ref.size(); ++i) { e = std::fmax(e, std::fabs(got[i] - ref[i])); m = std::fmax(m, std::fabs(ref[i])); }
    return e / m;
}

template <typename VT> static VT from_double(double v);
template <> float from_double<float>(double v) { return (float)v; }
template <> __nv_bfloat16 from_double<__nv_bfloat16>(double v) { return __float2bfloat16((float)v); }
static double to_double(float v) { return v; }
static double to_double(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename VT> static int run(int value_dtype, double tol, const char* name)
{
    const int64_t shapes_h[] = {11, 13, 6, 7, 3, 4};
    const int L = 3, P = 4, M = 8, D = 32, N = 2;
    int S = 0;
    for (int l = 0; l < L; ++l) S += (int)(shapes_h[2 * l] * shapes_h[2 * l + 1]);
    const int Lq = S;                                       // encoder form
    const msda_dims dims = {N, S, M, D, L, Lq, P};
    uint64_t seed = 7;
    const size_t nv = (size_t)N * S * M * D, nl = (size_t)N * Lq * M * L * P * 2, na = nl / 2, no = (size_t)N * Lq * M * D;
    std::vector<VT> value(nv), go(no);
    std::vector<float> loc(nl), attn(na);
    std::vector<double> value64(nv), go64(no), loc64(nl), attn64(na);
    for (size_t i = 0; i < nv; ++i) { value[i] = from_double<VT>(frand(seed) * 2 - 1); value64[i] = to_double(value[i]); }
    for (size_t i = 0; i < no; ++i) { go[i] = from_double<VT>(frand(seed) * 2 - 1); go64[i] = to_double(go[i]); }
    for (size_t i = 0; i < nl; ++i) { loc[i] = (float)(frand(seed) * 1.2 - 0.1); loc64[i] = loc[i]; }
    for (size_t r = 0; r < na / (L * P); ++r) {
        double sum = 0;
        for (int k = 0; k < L * P; ++k) { attn64[r * L * P + k] = frand(seed) + 1e-3; sum += attn64[r * L * P + k]; }
        for (int k = 0; k < L * P; ++k) { attn[r * L * P + k] = (float)(attn64[r * L * P + k] / sum); attn64[r * L * P + k] = attn[r * L * P + k]; }
    }
    std::vector<double> out_ref(no), gv_ref(nv), gl_ref(nl), ga_ref(na);
    msda_oracle_forward_f64(value64.data(), shapes_h, loc64.data(), attn64.data(), out_ref.data(), N, S, M, D, L, Lq, P, 4);
    msda_oracle_backward_f64(value64.data(), shapes_h, loc64.data(), attn64.data(), go64.data(), gv_ref.data(), gl_ref.data(),
                             ga_ref.data(), N, S, M, D, L, Lq, P, 4);

    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    std::vector<int64_t> shapes_v(shapes_h, shapes_h + 2 * L);
    int64_t* d_shapes = to_device(shapes_v);
    VT *d_value = to_device(value), *d_go = to_device(go), *d_out = nullptr, *d_gv = nullptr;
    float *d_loc = to_device(loc), *d_attn = to_device(attn), *d_gl = nullptr, *d_ga = nullptr;
    CK(cudaMalloc(&d_out, no * sizeof(VT)));
    CK(cudaMalloc(&d_gv, nv * sizeof(VT)));
    CK(cudaMalloc(&d_gl, nl * sizeof(float)));
    CK(cudaMalloc(&d_ga, na * sizeof(float)));
    const size_t ws_bytes = msda_b200_backward_workspace_bytes(value_dtype, &dims);
    void* d_ws = nullptr;
    if (ws_bytes) CK(cudaMalloc(&d_ws, ws_bytes));

    int rc = msda_b200_forward(d_value, value_dtype, d_shapes, nullptr, d_loc, d_attn, MSDA_F32, d_out, &dims, st);
    if (rc) { std::printf("%s forward: %s\n", name, msda_b200_error_string(rc)); return 3; }
    rc = msda_b200_backward(d_value, value_dtype, d_shapes, nullptr, d_loc, d_attn, MSDA_F32, d_go, d_gv, d_gl, d_ga, d_ws,
                            ws_bytes, &dims, st);
    if (rc) { std::printf("%s backward: %s\n", name, msda_b200_error_string(rc)); return 3; }
    CK(cudaStreamSynchronize(st));

    std::vector<VT> out(no), gv(nv);
    std::vector<float> gl(nl), ga(na);
    CK(cudaMemcpy(out.data(), d_out, no * sizeof(VT), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(gv.data(), d_gv, nv * sizeof(VT), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(gl.data(), d_gl, nl * sizeof(float), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(ga.data(), d_ga, na * sizeof(float), cudaMemcpyDeviceToHost));
    std::vector<double> o64(no), gv64(nv), gl64(nl), ga64(na);
    for (size_t i = 0; i < no; ++i) o64[i] = to_double(out[i]);
    for (size_t i = 0; i < nv; ++i) gv64[i] = to_double(gv[i]);
    for (size_t i = 0; i < nl; ++i) gl64[i] = gl[i];
    for (size_t i = 0; i < na; ++i) ga64[i] = ga[i];
    const double e_out = rel_err(o64, out_ref), e_gv = rel_err(gv64, gv_ref), e_ga = rel_err(ga64, ga_ref), e_gl = rel_err(gl64, gl_ref);
    std::printf("%s: rel err out %.3g grad_value %.3g grad_attn %.3g grad_loc %.3g (tol %.1g; grad_loc 10x: kinks)\n", name, e_out,
                e_gv, e_ga, e_gl, tol);
    cudaFree(d_shapes); cudaFree(d_value); cudaFree(d_go); cudaFree(d_out); cudaFree(d_gv); cudaFree(d_loc); cudaFree(d_attn);
    cudaFree(d_gl); cudaFree(d_ga); cudaFree(d_ws);
    cudaStreamDestroy(st);
    return (e_out <= tol && e_gv <= tol && e_ga <= tol && e_gl <= 10 * tol) ? 0 : 1;
}

int main()
{
    if (msda_b200_abi_version() != MSDA_B200_ABI_VERSION) { std::printf("ABI version mismatch\n"); return 4; }
    int rc = run<float>(MSDA_F32, 1e-5, "fp32");
    if (rc) return rc;
    return run<__nv_bfloat16>(MSDA_BF16, 1.2e-2, "bf16");
}
