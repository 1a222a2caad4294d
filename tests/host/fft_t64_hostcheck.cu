// Host validation of fft4096_t64.cuh: the header is compiled for the HOST (plain float arithmetic instead of the
// packed PTX) and the 64 lanes of one transform are run one after the other; the result is compared with a
// double-precision DFT.  Built and run by tests/test_fft_t64_host.py (no GPU needed).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <complex>
#include <vector>

#include "../../m_mimo_ofdm_with_nonlinear_pa_sim_b200/csrc/fft4096_t64.cuh"
using namespace mmimo;

static void transform(const std::vector<std::complex<float>>& in, std::vector<std::complex<float>>& out,
                      const std::vector<float2>& tab) {
    std::vector<pf2> buf(4096);
    static pf2 regs[64][64];
    for (int lane = 0; lane < 64; ++lane) {
        pf2 (&x)[64] = regs[lane];
        for (int n1 = 0; n1 < 64; ++n1) x[n1] = t64::qpk(in[lane + 64 * n1].real(), in[lane + 64 * n1].imag());
        t64::pass1_store(x, buf.data(), lane);
    }
    for (int lane = 0; lane < 64; ++lane) {
        pf2 (&x)[64] = regs[lane];
        t64::pass2_load(x, buf.data(), lane);
        const t64::TabA ta = t64::load_tab_a(tab.data(), lane);
        t64::pass2_compute<4>(x, tab.data(), ta, lane);
        for (int k1 = 0; k1 < 64; ++k1) {
            const float2 v = t64::qunpk(x[t64::rev64(k1)]);
            out[lane + 64 * k1] = {v.x, v.y};
        }
    }
}

int main(int argc, char** argv) {
    const int n_trials = argc > 1 ? atoi(argv[1]) : 3;
    std::vector<float2> tab(t64::TAB_FLOAT2);
    t64::build_table(tab.data());
    float tmax = 0.f, fmin = 1e9f;
    for (int i = 0; i < t64::TAB_A; ++i) {  // levels B and C only: |t| must stay small there
        const int plane = i < t64::TAB_B ? i / 1024 : (i - t64::TAB_B) / 256;
        const float2 v = tab[i];
        if (plane == 0) tmax = fmaxf(tmax, fmaxf(fabsf(v.x), fabsf(v.y)));
        if (plane == 1) { tmax = fmaxf(tmax, fabsf(v.x)); fmin = fminf(fmin, fabsf(v.y)); }
        if (plane == 2) fmin = fminf(fmin, fabsf(v.y));
    }
    printf("table: max |t| (levels B, C) = %.3f, min |f| = %.4f\n", tmax, fmin);
    const int N = 4096;
    std::vector<std::complex<double>> W(N);
    for (int k = 0; k < N; ++k) W[k] = std::polar(1.0, -2.0 * M_PI * k / N);
    double worst = 0.0;
    srand(1234);
    for (int trial = 0; trial < n_trials; ++trial) {
        std::vector<std::complex<float>> in(N), out(N);
        for (int n = 0; n < N; ++n) {
            float re = (float)rand() / RAND_MAX - 0.5f, im = (float)rand() / RAND_MAX - 0.5f;
            if (trial == 1 && ((n >> 6) >= 17 && (n >> 6) < 48)) re = im = 0.f;  // guard band empty, like the IFFT input
            if (trial == 2) { re = n == 65 ? 1.f : 0.f; im = 0.f; }                // one tone
            in[n] = {re, im};
        }
        transform(in, out, tab);
        double num = 0.0, den = 0.0;
        for (int k = 0; k < N; ++k) {
            std::complex<double> acc = 0.0;
            for (int n = 0; n < N; ++n) acc += std::complex<double>(in[n]) * W[(int)(((long long)n * k) % N)];
            num += std::norm(acc - std::complex<double>(out[k]));
            den += std::norm(acc);
        }
        const double rel = sqrt(num / den);
        printf("trial %d: relative L2 error %.3e\n", trial, rel);
        worst = fmax(worst, rel);
    }
    printf("worst %.3e\n", worst);
    return worst < 1.0e-6 ? 0 : 1;
}
