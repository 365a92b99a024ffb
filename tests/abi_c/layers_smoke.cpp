// layers_smoke.cpp -- drives include/kinet_layers_b200.h from a plain C++ host program (CUDA runtime only:
// no torch, no Python) and checks the results against a double-precision restatement of the same arithmetic
// (x <- LayerNorm(x + r) of the reference's layers, deformable_transformer.py:294-296, and its gradient).
// Built and run by tests/test_layers_gpu.py::test_layer_abi_from_a_cpp_host_program.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "kinet_layers_b200.h"
#include "msda_b200.h"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 2; } } while (0)

static float frand() { return (float)std::rand() / RAND_MAX * 2.f - 1.f; }

template <typename T> static T* to_dev(const std::vector<T>& h)
{
    T* d = nullptr;
    cudaMalloc(&d, h.size() * sizeof(T));
    cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
    return d;
}

int main()
{
    const int64_t rows = 777, pos_rows = 259;      // pos broadcast over 3 "batch items"
    const int C = 288;
    std::srand(3);
    std::vector<float> x(rows * C), r(rows * C), pos(pos_rows * C), gam(C), bet(C), dy(rows * C), dyp(rows * C);
    for (auto& v : x) v = frand();
    for (auto& v : r) v = 0.5f * frand();
    for (auto& v : pos) v = 0.3f * frand();
    for (auto& v : gam) v = 1.f + 0.2f * frand();
    for (auto& v : bet) v = 0.2f * frand();
    for (auto& v : dy) v = frand();
    for (auto& v : dyp) v = frand();
    float *dx_ = to_dev(x), *dr_ = to_dev(r), *dpos = to_dev(pos), *dgam = to_dev(gam), *dbet = to_dev(bet);
    float *ddy = to_dev(dy), *ddyp = to_dev(dyp);
    float *y, *ypos, *stats, *gx, *gr, *dgb, *partial, *colsum_out;
    CK(cudaMalloc(&y, rows * C * 4)); CK(cudaMalloc(&ypos, rows * C * 4)); CK(cudaMalloc(&stats, rows * 8));
    CK(cudaMalloc(&gx, rows * C * 4)); CK(cudaMalloc(&gr, rows * C * 4)); CK(cudaMalloc(&dgb, 2 * C * 4));
    CK(cudaMalloc(&partial, kinet_b200_colsum_partial_floats(2 * C) * 4)); CK(cudaMalloc(&colsum_out, C * 4));
    cudaStream_t st;
    CK(cudaStreamCreate(&st));

    kinet_ln_fwd_args f = {dx_, MSDA_F32, dr_, MSDA_F32, dpos, MSDA_F32, pos_rows, dgam, dbet,
                           y, MSDA_F32, nullptr, ypos, MSDA_F32, stats, rows, C, 1e-5f};
    int rc = kinet_b200_add_layernorm_forward(&f, st);
    if (rc) { std::printf("forward rc=%d\n", rc); return 1; }
    kinet_ln_bwd_args b = {ddy, MSDA_F32, nullptr, ddyp, MSDA_F32, dx_, MSDA_F32, dr_, MSDA_F32, dgam, stats,
                           gx, gr, dgb, partial, rows, C};
    rc = kinet_b200_add_layernorm_backward(&b, st);
    if (rc) { std::printf("backward rc=%d\n", rc); return 1; }
    rc = kinet_b200_colsum(ddy, MSDA_F32, rows, C, colsum_out, MSDA_F32, partial, st);
    if (rc) { std::printf("colsum rc=%d\n", rc); return 1; }
    CK(cudaStreamSynchronize(st));

    std::vector<float> hy(rows * C), hyp(rows * C), hgx(rows * C), hgr(rows * C), hdgb(2 * C), hcs(C);
    CK(cudaMemcpy(hy.data(), y, rows * C * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(hyp.data(), ypos, rows * C * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(hgx.data(), gx, rows * C * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(hgr.data(), gr, rows * C * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(hdgb.data(), dgb, 2 * C * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(hcs.data(), colsum_out, C * 4, cudaMemcpyDeviceToHost));

    // double-precision restatement
    double e_y = 0, e_p = 0, e_g = 0, e_r = 0, m_g = 0;
    std::vector<double> dgamma(C, 0.0), dbeta(C, 0.0), cs(C, 0.0);
    for (int64_t i = 0; i < rows; ++i) {
        std::vector<double> z(C), xh(C), g(C);
        double mean = 0, var = 0;
        for (int c = 0; c < C; ++c) { z[c] = (double)x[i * C + c] + r[i * C + c]; mean += z[c]; }
        mean /= C;
        for (int c = 0; c < C; ++c) var += (z[c] - mean) * (z[c] - mean);
        const double rstd = 1.0 / std::sqrt(var / C + 1e-5);
        double c1 = 0, c2 = 0;
        for (int c = 0; c < C; ++c) {
            xh[c] = (z[c] - mean) * rstd;
            const double yy = xh[c] * gam[c] + bet[c];
            e_y = std::fmax(e_y, std::fabs(yy - hy[i * C + c]));
            e_p = std::fmax(e_p, std::fabs(yy + pos[(i % pos_rows) * C + c] - hyp[i * C + c]));
            g[c] = (double)dy[i * C + c] + dyp[i * C + c];
            dgamma[c] += g[c] * xh[c];
            dbeta[c] += g[c];
            cs[c] += dy[i * C + c];
            c1 += g[c] * gam[c];
            c2 += g[c] * gam[c] * xh[c];
        }
        c1 /= C; c2 /= C;
        for (int c = 0; c < C; ++c) {
            const double d = rstd * (g[c] * gam[c] - c1 - xh[c] * c2);
            m_g = std::fmax(m_g, std::fabs(d));
            e_g = std::fmax(e_g, std::fabs(d - hgx[i * C + c]));
            e_r = std::fmax(e_r, std::fabs(d - hgr[i * C + c]));
        }
    }
    double e_dg = 0, e_db = 0, e_cs = 0, m_dg = 0;
    for (int c = 0; c < C; ++c) {
        e_dg = std::fmax(e_dg, std::fabs(dgamma[c] - hdgb[c]));
        e_db = std::fmax(e_db, std::fabs(dbeta[c] - hdgb[C + c]));
        e_cs = std::fmax(e_cs, std::fabs(cs[c] - hcs[c]));
        m_dg = std::fmax(m_dg, std::fabs(dgamma[c]));
    }
    std::printf("layers_smoke: |y| %.2e |y+pos| %.2e |dx| %.2e (max %.2f) |dr| %.2e |dgamma| %.2e (max %.1f) |dbeta| %.2e |colsum| %.2e\n",
                e_y, e_p, e_g, m_g, e_r, e_dg, m_dg, e_db, e_cs);
    // argument errors come back as codes, never as exceptions or aborts
    kinet_ln_fwd_args bad = f;
    bad.C = 290;
    if (kinet_b200_add_layernorm_forward(&bad, st) != MSDA_E_SHAPE) { std::printf("expected MSDA_E_SHAPE\n"); return 1; }
    const bool ok = e_y < 5e-6 && e_p < 5e-6 && e_g < 2e-5 * std::fmax(1.0, m_g) && e_r < 2e-5 * std::fmax(1.0, m_g) &&
                    e_dg < 2e-5 * std::fmax(1.0, m_dg) && e_db < 5e-4 && e_cs < 5e-4;
    std::printf(ok ? "PASS\n" : "FAIL\n");
    return ok ? 0 : 1;
}
