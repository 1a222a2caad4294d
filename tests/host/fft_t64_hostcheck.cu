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

// prune_in: level A of pass 1 in its guard-band form (registers 17..47 hold garbage on entry, as the frame kernel leaves
// them); prune_out: last level of pass 2 produces only the data bins (+ bin 1024 of lane 0)
static void transform(const std::vector<std::complex<float>>& in, std::vector<std::complex<float>>& out,
                      const std::vector<float2>& tab, bool prune_in = false, bool prune_out = false) {
    std::vector<pf2> buf(t64::XBUF);
    static pf2 regs[64][64];
    for (int lane = 0; lane < 64; ++lane) {
        pf2 (&x)[64] = regs[lane];
        for (int n1 = 0; n1 < 64; ++n1) x[n1] = t64::qpk(in[lane + 64 * n1].real(), in[lane + 64 * n1].imag());
        if (prune_in) {
            for (int n1 = 17; n1 < 48; ++n1) x[n1] = t64::qpk(123.f, -77.f);
            t64::p1_level_a_pruned(x);
        } else {
            t64::p1_level_a(x);
        }
        t64::p1_levels_bc(x);
        t64::exchange_store(x, buf.data(), lane);
    }
    for (int lane = 0; lane < 64; ++lane) {
        pf2 (&x)[64] = regs[lane];
        t64::pass2_load(x, buf.data(), lane);
        const t64::TabA ta = t64::load_tab_a(tab.data(), lane);
        t64::Raw3 ring[4];
        t64::p2_levels_ab<4>(x, ring, tab.data(), ta, lane);
        if (prune_out) t64::p2_level_c_all<4, true>(x, ring, tab.data(), lane);
        else t64::p2_level_c_all<4, false>(x, ring, tab.data(), lane);
        for (int k1 = 0; k1 < 64; ++k1) {
            const float2 v = t64::qunpk(x[t64::rev64(k1)]);
            out[lane + 64 * k1] = {v.x, v.y};
        }
    }
}

int main(int argc, char** argv) {
    const int n_trials = argc > 1 ? atoi(argv[1]) : 5;
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
            // guard band empty, like the frame kernel's first transform (position 1024 = register 16 of lane 0 is data)
            if ((trial == 1 || trial == 3) && ((n >> 6) >= 16 && (n >> 6) < 48) && n != 1024) re = im = 0.f;
            if (trial == 2) { re = n == 65 ? 1.f : 0.f; im = 0.f; }                // one tone
            in[n] = {re, im};
        }
        transform(in, out, tab, trial == 3, trial == 4);
        double num = 0.0, den = 0.0;
        for (int k = 0; k < N; ++k) {
            if (trial == 4 && !(k < 1024 || k >= 3072 || k == 1024)) continue;  // pruned outputs: data bins only
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
