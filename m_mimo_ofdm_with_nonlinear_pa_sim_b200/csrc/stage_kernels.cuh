// Stage-level (unfused) kernels behind the drop-in building-block classes: one kernel per reference
// method, any n_fft = 2^m (8..8192) and any even n_sub_carr < n_fft.  All arrays use the REFERENCE's
// orders: frequency-domain vectors in FFT-bin order (DC at 0), per-sub-carrier vectors in data order
// ("negative half then positive half", modulation.py:266-267).  These kernels are HBM-bound
// (K x n_fft complex64 in, the same out); the Monte-Carlo hot loop does not use them - it runs the
// fused frame kernel (frame_kernel.cuh).
#pragma once
#include "prep_kernels.cuh"

namespace mmimo {

// data-sub-carrier index of FFT bin n, or -1 (DC / guard band)
__host__ __device__ __forceinline__ int sc_of_bin(int n, int n_fft, int n_sc) {
    const int h = n_sc >> 1;
    if (n >= 1 && n <= h) return h + n - 1;
    if (n >= n_fft - h) return n - (n_fft - h);
    return -1;
}
__host__ __device__ __forceinline__ int bin_of_sc(int s, int n_fft, int n_sc) {
    const int h = n_sc >> 1;
    return s < h ? n_fft - h + s : s - h + 1;
}

// ---------------------------------------------------------------------------------------------
// bits (one byte per bit, MSB first) -> symbol values.  modulation.py:13-25 + utilities.py:54-67
// ---------------------------------------------------------------------------------------------
__global__ void k_bits_to_idx(const uint8_t* __restrict__ bits, uint8_t* __restrict__ idx, size_t n_sym, int bps) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_sym) return;
    uint32_t v = 0;
    for (int b = 0; b < bps; ++b) v = (v << 1) | (bits[i * bps + b] & 1u);
    idx[i] = (uint8_t)v;
}
// symbol values -> bits.  utilities.py:18-51
__global__ void k_idx_to_bits(const uint8_t* __restrict__ idx, uint8_t* __restrict__ bits, size_t n_sym, int bps) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_sym) return;
    const uint32_t v = idx[i];
    for (int b = 0; b < bps; ++b) bits[i * bps + b] = (uint8_t)((v >> (bps - 1 - b)) & 1u);
}
// symbol values -> constellation points (scale = alpha of a corrected constellation, modulation.py:167-176)
__global__ void k_idx_to_symbols(const uint8_t* __restrict__ idx, float2* __restrict__ sym, size_t n_sym, Qam q, float scale) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_sym) return;
    const float2 s = qam_symbol(q, idx[i]);
    sym[i] = make_float2(s.x * scale, s.y * scale);
}
// Modem.symbol_detection / demodulate (modulation.py:138-146, :63-77): nearest point of the scaled table
__global__ void k_detect(const float2* __restrict__ in, uint8_t* __restrict__ idx, float2* __restrict__ sym, size_t n_sym,
                         Qam q, float scale) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_sym) return;
    const float2 r = in[i];
    const float inv = 1.0f / scale;
    const uint32_t v = qam_detect(q, make_float2(r.x * inv, r.y * inv));
    if (idx) idx[i] = (uint8_t)v;
    if (sym) {
        const float2 s = qam_symbol(q, v);
        sym[i] = make_float2(s.x * scale, s.y * scale);
    }
}

// modulation.soft_decoding (modulation.py:29-59): exact log-likelihood ratio of every bit,
//   llr[i*bps + bps-1-b] = log( sum_{c: bit b of label(c) set} e^{-|y_i-c|^2/nv_i} / sum_{c: bit b clear} e^{...} ),
// +inf where the denominator underflows to zero (modulation.py:53-54), in double like the reference.
__global__ void k_soft_llr(const float2* __restrict__ in, const double* __restrict__ nv, size_t nv_n, double* __restrict__ llr,
                           size_t n_sym, Qam q, double scale, int bps) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_sym) return;
    const double yr = in[i].x, yi = in[i].y;
    const double inv_nv = 1.0 / nv[nv_n > 1 ? i : 0];
    double num[8], den[8];
#pragma unroll
    for (int b = 0; b < 8; ++b) num[b] = den[b] = 0.0;
    const uint32_t n_pts = 1u << bps;
    for (uint32_t v = 0; v < n_pts; ++v) {
        const float2 c = qam_symbol(q, v);
        const double dr = yr - scale * c.x, di = yi - scale * c.y;
        const double e = exp(-(dr * dr + di * di) * inv_nv);
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            if ((v >> b) & 1u) num[b] += e;
            else den[b] += e;
        }
    }
    for (int b = 0; b < bps; ++b)
        llr[i * bps + bps - 1 - b] = den[b] == 0.0 ? (double)INFINITY : log(num[b] / den[b]);
}

// ---------------------------------------------------------------------------------------------
// OfdmQamModem.modulate up to the IFFT input (modulation.py:346-382 + :264-267): symbol values of
// every user -> precode -> sum over users -> place on the n_fft grid, for every antenna.
//   idx   [U][n_sc]          symbol values
//   v     [K][U][n_sc]       precoder in AntennaArray.get_precoding_mat layout, or nullptr (no precoding)
//   extra [n_sc]             optional additive term (CncMuReceiver: the other user's symbols, corrector.py:322)
//   grid  [G][K][n_fft]      G = U when per_user (sum_usr_signals=False), else 1
// ---------------------------------------------------------------------------------------------
struct MapArgs {
    const uint8_t* __restrict__ idx; const float2* __restrict__ v; const float2* __restrict__ extra; float2* __restrict__ grid;
    int n_users, n_ant, n_fft, n_sc, per_user;
    Qam q; float scale;       // constellation scale
    float w_self, w_extra;    // weights of the mapped symbols and of `extra`
};
__device__ __forceinline__ float2 map_precode_bin(const MapArgs& a, int n, int k, int g) {
    const int s = sc_of_bin(n, a.n_fft, a.n_sc);
    float2 acc = make_float2(0.f, 0.f);
    if (s >= 0) {
        const int u0 = a.per_user ? g : 0, u1 = a.per_user ? g + 1 : a.n_users;
        for (int u = u0; u < u1; ++u) {
            float2 sym = qam_symbol(a.q, a.idx[(size_t)u * a.n_sc + s]);
            sym.x *= a.scale * a.w_self; sym.y *= a.scale * a.w_self;
            if (a.v) sym = cmul(sym, a.v[((size_t)k * a.n_users + u) * a.n_sc + s]);
            acc = cadd(acc, sym);
        }
        if (a.extra) {
            const float2 e = a.extra[s];
            acc.x = fmaf(a.w_extra, e.x, acc.x); acc.y = fmaf(a.w_extra, e.y, acc.y);
        }
    }
    return acc;
}
// four bins per thread (n, n + n_fft/4, ...): contiguous 8-byte accesses per warp, four independent loads in flight
__global__ void __launch_bounds__(256) k_map_precode(const MapArgs a) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x, q = a.n_fft >> 2;
    if (n >= q) return;
    const int k = blockIdx.y, g = blockIdx.z;
    float2 y[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) y[j] = map_precode_bin(a, n + j * q, k, g);
    float2* row = a.grid + ((size_t)g * a.n_ant + k) * a.n_fft + n;
#pragma unroll
    for (int j = 0; j < 4; ++j) row[j * q] = y[j];
}

// ---------------------------------------------------------------------------------------------
// Batched orthonormal FFT / IFFT, one CTA per row.  torch.fft.fft/ifft(norm="ortho") call sites
// utilities.py:329,339, modulation.py:270,290, corrector.py:93,98.
// n == 4096: the register/shared-memory radix-16 transform of fft4096.cuh.
// otherwise: radix-2 Stockham autosort in shared memory (n <= 8192).
// `skip` input samples are dropped in front of each input row (cyclic prefix removal, utilities.py:323);
// `cp` samples are prepended to each output row (modulation.py:273).
// ---------------------------------------------------------------------------------------------
#ifndef FFT_ROWS_ONEBUF
#define FFT_ROWS_ONEBUF 1
#endif
// one exchange buffer for both exchanges of a transform (a barrier in front of each store phase): 35 KB instead of 67 KB
// of shared memory per CTA, so that four CTAs instead of three are resident per SM (the kernel is HBM-latency-bound)
struct SyncCtaOneBuf : SyncCta {
    __device__ __forceinline__ void x_begin() const { __syncthreads(); }
};
constexpr int FFT_ROWS_SMEM = ((FFT_ROWS_ONEBUF ? 4096 : 8192) + 256) * 8;
template <int S>
__global__ void __launch_bounds__(FFT_THREADS, FFT_ROWS_ONEBUF ? 4 : 3) k_fft4096_rows(const float2* __restrict__ in, float2* __restrict__ out,
                                                              const float2* __restrict__ tw4096, const float2* __restrict__ tw256,
                                                              int skip, int cp) {
    extern __shared__ float2 smem[];
    pf2* bufA = reinterpret_cast<pf2*>(smem);
    pf2* bufB = FFT_ROWS_ONEBUF ? bufA : reinterpret_cast<pf2*>(smem + 4096);
    float2* tw2 = smem + (FFT_ROWS_ONEBUF ? 4096 : 8192);
    const int t = threadIdx.x;
    tw2[t] = tw256[t];
    Tw1 tw1;
    load_tw1(tw1, tw4096, t);
    __syncthreads();
    const float2* src = in + (size_t)blockIdx.x * (4096 + skip) + skip;
    float2* dst = out + (size_t)blockIdx.x * (4096 + cp) + cp;
    pf2 x[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] = pk2(src[t + 256 * j]);
#if FFT_ROWS_ONEBUF
    fft4096<S, false, false, SyncCtaOneBuf>(x, bufA, bufB, tw2, tw1, t);
#else
    fft4096<S, false, false>(x, bufA, bufB, tw2, tw1, t);
#endif
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        const float2 y = unpk(x[j]);
        const float2 o = make_float2(y.x * 0.015625f, y.y * 0.015625f);
        const int n = t + 256 * j;
        dst[n] = o;
        if (n >= 4096 - cp) dst[n - 4096] = o;  // cyclic prefix = last cp samples
    }
}

template <int S>
__global__ void __launch_bounds__(256) k_fft_generic_rows(const float2* __restrict__ in, float2* __restrict__ out, int n,
                                                          int log2n, int skip, int cp) {
    extern __shared__ float2 smem[];
    float2* a = smem;
    float2* b = smem + n;
    const float2* src = in + (size_t)blockIdx.x * (n + skip) + skip;
    float2* dst = out + (size_t)blockIdx.x * (n + cp) + cp;
    for (int i = threadIdx.x; i < n; i += blockDim.x) a[i] = src[i];
    __syncthreads();
    // Stockham autosort, decimation in frequency: sub-transform length nn halves, stride s doubles
    const int half = n >> 1;
    for (int st = 0, nn = n, s = 1; st < log2n; ++st, nn >>= 1, s <<= 1) {
        const int m = nn >> 1;
        for (int i = threadIdx.x; i < half; i += blockDim.x) {
            const int p = i / s, q = i - p * s;  // p in [0,m), q in [0,s)
            float sn, cs;
            sincospif((float)(S * 2 * p) / (float)nn, &sn, &cs);  // exp(S 2 pi i p / nn), argument exact in fp32
            const float2 c0 = a[q + s * p];
            const float2 c1 = a[q + s * (p + m)];
            b[q + s * (2 * p)] = cadd(c0, c1);
            b[q + s * (2 * p + 1)] = cmul(csub(c0, c1), make_float2(cs, sn));
        }
        __syncthreads();
        float2* tmp = a; a = b; b = tmp;
    }
    const float sc = rsqrtf((float)n);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float2 y = a[i];
        const float2 o = make_float2(y.x * sc, y.y * sc);
        dst[i] = o;
        if (i >= n - cp) dst[i - n] = o;
    }
}

// prepend the cyclic prefix to every row (modulation.py:273): out[r] = in[r][n-cp:] ++ in[r]
__global__ void k_add_cp(const float2* __restrict__ in, float2* __restrict__ out, int n, int cp) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
    if (i >= n + cp) return;
    out[(size_t)r * (n + cp) + i] = in[(size_t)r * n + (i < cp ? n - cp + i : i - cp)];
}

// ---------------------------------------------------------------------------------------------
// PA models on a flat array.  distortion.py:9-19, :102-113, :202-211
// ---------------------------------------------------------------------------------------------
template <int KIND>
__global__ void k_pa(const float2* __restrict__ in, float2* __restrict__ out, size_t n, float p0, float p1) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = pa_apply<KIND>(in[i], p0, sqrtf(p0), p1);
}
// the same on four samples per thread (two 16-byte loads in flight, 16-byte stores): n even, both pointers 16-byte aligned
template <int KIND>
__global__ void __launch_bounds__(256) k_pa4(const float4* __restrict__ in, float4* __restrict__ out, size_t n_pairs, float p0, float p1) {
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 2;
    const float sp = sqrtf(p0);
    if (i + 1 < n_pairs) {
        const float4 a = in[i], b = in[i + 1];
        const float2 a0 = pa_apply<KIND>(make_float2(a.x, a.y), p0, sp, p1), a1 = pa_apply<KIND>(make_float2(a.z, a.w), p0, sp, p1);
        const float2 b0 = pa_apply<KIND>(make_float2(b.x, b.y), p0, sp, p1), b1 = pa_apply<KIND>(make_float2(b.z, b.w), p0, sp, p1);
        out[i] = make_float4(a0.x, a0.y, a1.x, a1.y);
        out[i + 1] = make_float4(b0.x, b0.y, b1.x, b1.y);
    } else if (i < n_pairs) {
        const float4 a = in[i];
        const float2 a0 = pa_apply<KIND>(make_float2(a.x, a.y), p0, sp, p1), a1 = pa_apply<KIND>(make_float2(a.z, a.w), p0, sp, p1);
        out[i] = make_float4(a0.x, a0.y, a1.x, a1.y);
    }
}

// ---------------------------------------------------------------------------------------------
// Miso*Fd.propagate (channel.py:74-89): Y = X .* H, optionally summed over antennas.
// ---------------------------------------------------------------------------------------------
__global__ void k_propagate(const float2* __restrict__ x, const float2* __restrict__ h, float2* __restrict__ out, int n_ant,
                            int n, int sum) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (sum) {
        float2 acc = make_float2(0.f, 0.f);
        for (int k = 0; k < n_ant; ++k) acc = cfma(x[(size_t)k * n + i], h[(size_t)k * n + i], acc);
        out[i] = acc;
    } else {
        const int k = blockIdx.y;
        out[(size_t)k * n + i] = cmul(x[(size_t)k * n + i], h[(size_t)k * n + i]);
    }
}

// out = in + scale * noise   (Awgn.process, noise.py:64-83: the draw itself comes from the caller's numpy Generator)
__global__ void k_axpy_c(const float2* __restrict__ in, const float2* __restrict__ nz, float2* __restrict__ out, float sr,
                         size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = make_float2(fmaf(sr, nz[i].x, in[i].x), fmaf(sr, nz[i].y, in[i].y));
}

// utilities.count_mismatched_bits (utilities.py:94-104)
__global__ void k_count_mismatch(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, size_t n,
                                 unsigned long long* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int c = 0;
    for (; i < n; i += (size_t)gridDim.x * blockDim.x) c += (a[i] != b[i]);
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, (unsigned long long)c);
}

// ---------------------------------------------------------------------------------------------
// Channels on the full n_fft grid (channel.py:35-72 LOS, :116-167 two-path), per-antenna gains.
// ---------------------------------------------------------------------------------------------
struct ChanFullArgs {
    const double* tx_pos;   // [K][3]
    const double* tx_gain;  // [K] dB
    const double* freqs;    // [n_fft]
    double rx[3]; double rx_gain_db;
    int kind, skip_attenuation, n_fft;
    float2* h_out;          // [K][n_fft]
};
__global__ void k_channel_full(const ChanFullArgs a) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x, k = blockIdx.y;
    if (n >= a.n_fft) return;
    const double tz = a.tx_pos[k * 3 + 2];
    const double dx = a.tx_pos[k * 3 + 0] - a.rx[0], dy = a.tx_pos[k * 3 + 1] - a.rx[1], dz = tz - a.rx[2];
    const double gain = sqrt(pow(10.0, (a.tx_gain[k] + a.rx_gain_db) / 10.0));
    const double f = a.freqs[n];
    float2 h = path_coeff(sqrt(dx * dx + dy * dy + dz * dz), f, gain, a.skip_attenuation);
    if (a.kind == 1) {
        const double el = atan((tz + a.rx[2]) / sqrt(dx * dx + dy * dy));
        const double sn = sin(el);
        const float2 g = path_coeff(tz / sn + a.rx[2] / sn, f, gain, a.skip_attenuation);
        h.x -= g.x; h.y -= g.y;
    }
    a.h_out[(size_t)k * a.n_fft + n] = h;
}
// free-space attenuation matrix (channel.py:225-227) and its product with injected IID coefficients (:244-250)
__global__ void k_fspl_apply(const ChanFullArgs a, const float2* __restrict__ coeffs) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x, k = blockIdx.y;
    if (n >= a.n_fft) return;
    const double dx = a.tx_pos[k * 3 + 0] - a.rx[0], dy = a.tx_pos[k * 3 + 1] - a.rx[1], dz = a.tx_pos[k * 3 + 2] - a.rx[2];
    const double gain = sqrt(pow(10.0, (a.tx_gain[k] + a.rx_gain_db) / 10.0));
    const float att = (float)(gain * C_LIGHT / (4.0 * 3.14159265358979323846 * sqrt(dx * dx + dy * dy + dz * dz) * a.freqs[n]));
    const float2 c = coeffs ? coeffs[(size_t)k * a.n_fft + n] : make_float2(1.f, 0.f);
    a.h_out[(size_t)k * a.n_fft + n] = make_float2(c.x * att, c.y * att);
}
// MisoRandomPathsFd (channel.py:330-344), literally: exp(-2j f (tau_p + delta_k sin(phi_p / c))) summed over paths
struct RandPathArgs {
    const double* tx_pos; const double* freqs; const double* angles; const double* taus;
    int n_paths, n_fft; float2* h_out;
};
__global__ void k_channel_random_paths(const RandPathArgs a) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x, k = blockIdx.y;
    if (n >= a.n_fft) return;
    const double dx = a.tx_pos[k * 3 + 0] - a.tx_pos[0], dy = a.tx_pos[k * 3 + 1] - a.tx_pos[1], dz = a.tx_pos[k * 3 + 2] - a.tx_pos[2];
    const double delta = sqrt(dx * dx + dy * dy + dz * dz);
    const double f = a.freqs[n];
    double re = 0.0, im = 0.0;
    for (int p = 0; p < a.n_paths; ++p) {
        const double arg = -2.0 * f * (a.taus[p] + delta * sin(a.angles[p] / C_LIGHT));
        double s, c;
        sincos(arg, &s, &c);  // |arg| ~ 7e3 rad: double-precision range reduction
        re += c; im += s;
    }
    const double sc = 1.0 / sqrt((double)a.n_paths);
    a.h_out[(size_t)k * a.n_fft + n] = make_float2((float)(re * sc), (float)(im * sc));
}

// ---------------------------------------------------------------------------------------------
// AntennaArray.set_precoding_matrix (antenna_array.py:142-311).  h [U][K][n_fft] -> v [K][U][n_sc].
// One thread per data sub-carrier; antennas looped (coalesced across sub-carriers).
// ---------------------------------------------------------------------------------------------
enum PrecMode : int { PREC_PHASE = 0, PREC_MRT = 1, PREC_ZF = 2 };
constexpr int MAX_USERS = 8;
struct PrecArgs {
    const float2* h; float2* v;
    int n_users, n_ant, n_fft, n_sc, mode;
    int sep_carriers;  // user u owns the u-th block of n_sc / U sub-carriers (:275-305); output is [K][1][n_sc]
};
__global__ void k_precoding(const PrecArgs a) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= a.n_sc) return;
    const int bin = bin_of_sc(s, a.n_fft, a.n_sc);
    const int U = a.n_users, K = a.n_ant;
    const size_t us = (size_t)K * a.n_fft;
    if (a.sep_carriers) {
        const int u = s / (a.n_sc / U);
        const float2* hu = a.h + (size_t)u * us + bin;
        float nrm = 0.f;
        for (int k = 0; k < K; ++k) { const float2 h = hu[(size_t)k * a.n_fft]; nrm = fmaf(h.x, h.x, fmaf(h.y, h.y, nrm)); }
        const float inv = rsqrtf(nrm);
        for (int k = 0; k < K; ++k) {
            const float2 h = hu[(size_t)k * a.n_fft];
            float2 o;
            if (a.mode == PREC_MRT) o = make_float2(h.x * inv, -h.y * inv);
            else { const float m = rsqrtf(fmaf(h.x, h.x, h.y * h.y)); o = make_float2(h.x * m, -h.y * m); }
            a.v[(size_t)k * a.n_sc + s] = o;
        }
        return;
    }
    if (a.mode == PREC_PHASE) {  // exp(1j angle(conj h)), :178 / :259-264
        for (int u = 0; u < U; ++u)
            for (int k = 0; k < K; ++k) {
                const float2 h = a.h[(size_t)u * us + (size_t)k * a.n_fft + bin];
                const float m = rsqrtf(fmaf(h.x, h.x, h.y * h.y));
                a.v[((size_t)k * U + u) * a.n_sc + s] = make_float2(h.x * m, -h.y * m);
            }
        return;
    }
    if (a.mode == PREC_MRT) {  // conj(h_u) / sqrt(sum_u sum_k |h_u|^2), :170-171 / :202-220
        float nrm = 0.f;
        for (int u = 0; u < U; ++u)
            for (int k = 0; k < K; ++k) {
                const float2 h = a.h[(size_t)u * us + (size_t)k * a.n_fft + bin];
                nrm = fmaf(h.x, h.x, fmaf(h.y, h.y, nrm));
            }
        const float inv = 1.0f / sqrtf(nrm);
        for (int u = 0; u < U; ++u)
            for (int k = 0; k < K; ++k) {
                const float2 h = a.h[(size_t)u * us + (size_t)k * a.n_fft + bin];
                a.v[((size_t)k * U + u) * a.n_sc + s] = make_float2(h.x * inv, -h.y * inv);
            }
        return;
    }
    // zero forcing (:222-257): T = [h_1 .. h_U] (K x U); V = sqrt(K-U) conj(T) inv(T^T conj(T)); unit Frobenius norm.
    // Gram matrix and its inverse in double: the users' LOS channels can be close to collinear.
    double gr[MAX_USERS][MAX_USERS], gi[MAX_USERS][MAX_USERS], ir[MAX_USERS][MAX_USERS], ii[MAX_USERS][MAX_USERS];
    for (int p = 0; p < U; ++p)
        for (int q = 0; q < U; ++q) {
            double sr = 0.0, si = 0.0;
            for (int k = 0; k < K; ++k) {
                const float2 hp = a.h[(size_t)p * us + (size_t)k * a.n_fft + bin];
                const float2 hq = a.h[(size_t)q * us + (size_t)k * a.n_fft + bin];
                sr += (double)hp.x * hq.x + (double)hp.y * hq.y;  // h_p * conj(h_q)
                si += (double)hp.y * hq.x - (double)hp.x * hq.y;
            }
            gr[p][q] = sr; gi[p][q] = si;
            ir[p][q] = (p == q) ? 1.0 : 0.0; ii[p][q] = 0.0;
        }
    double gmax = 0.0;
    for (int p = 0; p < U; ++p) gmax = fmax(gmax, gr[p][p]);
    bool singular = false;
    double g0r[MAX_USERS][MAX_USERS], g0i[MAX_USERS][MAX_USERS];  // the Gram matrix itself, for the pseudo-inverse path
    for (int p = 0; p < U; ++p)
        for (int q = 0; q < U; ++q) { g0r[p][q] = gr[p][q]; g0i[p][q] = gi[p][q]; }
    for (int c = 0; c < U && !singular; ++c) {  // Gauss-Jordan with partial pivoting
        int piv = c; double best = gr[c][c] * gr[c][c] + gi[c][c] * gi[c][c];
        for (int r = c + 1; r < U; ++r) {
            const double m = gr[r][c] * gr[r][c] + gi[r][c] * gi[r][c];
            if (m > best) { best = m; piv = r; }
        }
        if (best <= 1e-30 * gmax * gmax) { singular = true; break; }  // np.linalg.inv would raise here (or return garbage)
        if (piv != c)
            for (int q = 0; q < U; ++q) {
                double t;
                t = gr[c][q]; gr[c][q] = gr[piv][q]; gr[piv][q] = t;
                t = gi[c][q]; gi[c][q] = gi[piv][q]; gi[piv][q] = t;
                t = ir[c][q]; ir[c][q] = ir[piv][q]; ir[piv][q] = t;
                t = ii[c][q]; ii[c][q] = ii[piv][q]; ii[piv][q] = t;
            }
        const double pr = gr[c][c] / best, pi = -gi[c][c] / best;  // 1 / pivot
        for (int q = 0; q < U; ++q) {
            double xr = gr[c][q], xi = gi[c][q];
            gr[c][q] = xr * pr - xi * pi; gi[c][q] = xr * pi + xi * pr;
            xr = ir[c][q]; xi = ii[c][q];
            ir[c][q] = xr * pr - xi * pi; ii[c][q] = xr * pi + xi * pr;
        }
        for (int r = 0; r < U; ++r) {
            if (r == c) continue;
            const double fr = gr[r][c], fi = gi[r][c];
            for (int q = 0; q < U; ++q) {
                gr[r][q] -= fr * gr[c][q] - fi * gi[c][q]; gi[r][q] -= fr * gi[c][q] + fi * gr[c][q];
                ir[r][q] -= fr * ir[c][q] - fi * ii[c][q]; ii[r][q] -= fr * ii[c][q] + fi * ir[c][q];
            }
        }
    }
    if (singular) {
        // antenna_array.py:243-251: the reference retries with np.linalg.pinv.  Pseudo-inverse of the Hermitian Gram matrix
        // by cyclic complex Jacobi rotations: G = W diag(w) W^H, pinv = W diag(1 / w_i for w_i > 1e-15 w_max) W^H.
        double wr[MAX_USERS][MAX_USERS], wi[MAX_USERS][MAX_USERS];
        for (int p = 0; p < U; ++p)
            for (int q = 0; q < U; ++q) { wr[p][q] = (p == q) ? 1.0 : 0.0; wi[p][q] = 0.0; }
        for (int sweep = 0; sweep < 24; ++sweep) {
            double off = 0.0;
            for (int p = 0; p < U; ++p)
                for (int q = p + 1; q < U; ++q) off += g0r[p][q] * g0r[p][q] + g0i[p][q] * g0i[p][q];
            if (off <= 1e-60 * gmax * gmax) break;
            for (int p = 0; p < U; ++p)
                for (int q = p + 1; q < U; ++q) {
                    const double m = sqrt(g0r[p][q] * g0r[p][q] + g0i[p][q] * g0i[p][q]);
                    if (m <= 1e-300) continue;
                    const double er = g0r[p][q] / m, ei = g0i[p][q] / m;          // e^{i phi} of the pivot element
                    const double tau = (g0r[q][q] - g0r[p][p]) / (2.0 * m);
                    const double tt = (tau >= 0.0 ? 1.0 : -1.0) / (fabs(tau) + sqrt(1.0 + tau * tau));
                    const double cs = 1.0 / sqrt(1.0 + tt * tt), sn = tt * cs;
                    // R = I except R[p][p] = c, R[p][q] = s e^{i phi}, R[q][p] = -s e^{-i phi}, R[q][q] = c;  G <- R^H G R
                    for (int k2 = 0; k2 < U; ++k2) {  // columns p, q of G R
                        const double apr = g0r[k2][p], api = g0i[k2][p], aqr = g0r[k2][q], aqi = g0i[k2][q];
                        g0r[k2][p] = cs * apr - sn * (aqr * er + aqi * ei);  g0i[k2][p] = cs * api - sn * (aqi * er - aqr * ei);
                        g0r[k2][q] = cs * aqr + sn * (apr * er - api * ei);  g0i[k2][q] = cs * aqi + sn * (api * er + apr * ei);
                        const double vpr = wr[k2][p], vpi = wi[k2][p], vqr = wr[k2][q], vqi = wi[k2][q];
                        wr[k2][p] = cs * vpr - sn * (vqr * er + vqi * ei);   wi[k2][p] = cs * vpi - sn * (vqi * er - vqr * ei);
                        wr[k2][q] = cs * vqr + sn * (vpr * er - vpi * ei);   wi[k2][q] = cs * vqi + sn * (vpi * er + vpr * ei);
                    }
                    for (int k2 = 0; k2 < U; ++k2) {  // rows p, q of R^H (G R)
                        const double apr = g0r[p][k2], api = g0i[p][k2], aqr = g0r[q][k2], aqi = g0i[q][k2];
                        g0r[p][k2] = cs * apr - sn * (aqr * er - aqi * ei);  g0i[p][k2] = cs * api - sn * (aqi * er + aqr * ei);
                        g0r[q][k2] = cs * aqr + sn * (apr * er + api * ei);  g0i[q][k2] = cs * aqi + sn * (api * er - apr * ei);
                    }
                }
        }
        double wmax = 0.0;
        for (int p = 0; p < U; ++p) wmax = fmax(wmax, g0r[p][p]);
        for (int p = 0; p < U; ++p)
            for (int q = 0; q < U; ++q) {
                double sr = 0.0, si = 0.0;
                for (int e = 0; e < U; ++e) {
                    if (!(g0r[e][e] > 1e-15 * wmax)) continue;
                    const double iw = 1.0 / g0r[e][e];   // W[p][e] conj(W[q][e]) / w_e
                    sr += (wr[p][e] * wr[q][e] + wi[p][e] * wi[q][e]) * iw;
                    si += (wi[p][e] * wr[q][e] - wr[p][e] * wi[q][e]) * iw;
                }
                ir[p][q] = sr; ii[p][q] = si;
            }
    }
    double fro = 0.0;
    for (int pass = 0; pass < 2; ++pass) {
        const double sc = pass ? 1.0 / sqrt(fro) : 0.0;
        for (int k = 0; k < K; ++k)
            for (int u = 0; u < U; ++u) {
                double vr = 0.0, vi = 0.0;
                for (int p = 0; p < U; ++p) {  // conj(h_p[k]) * Ginv[p][u]
                    const float2 hp = a.h[(size_t)p * us + (size_t)k * a.n_fft + bin];
                    vr += (double)hp.x * ir[p][u] + (double)hp.y * ii[p][u];
                    vi += (double)hp.x * ii[p][u] - (double)hp.y * ir[p][u];
                }
                if (!pass) fro += vr * vr + vi * vi;
                else a.v[((size_t)k * U + u) * a.n_sc + s] = make_float2((float)(vr * sc), (float)(vi * sc));
            }
    }
}

// ---------------------------------------------------------------------------------------------
// AGC / alpha bookkeeping: corrector.py:126-163,209-245 (+MU :348-403) and mp_model.py:290-329.
//   k_vk_pow : vk_pow[k] = sum_{u,s} |v[k][u][s]|^2 ; alpha[k] = calc_alpha(10 log10(ibo_lin n_sc / (vk_pow K)))
//   k_agc    : agc[bin(s)] = sum_k alpha_k h_k v_{k,usr}, hkvk likewise with alpha = 1; 1 elsewhere
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_vk_pow(const float2* __restrict__ v, double* __restrict__ vk_pow, float* __restrict__ alpha,
                                                int n_ant, int row_len, double ibo_lin, int n_sc, double alpha_fixed, int use_fixed) {
    __shared__ double red_s[8];
    const int k = blockIdx.x, t = threadIdx.x;
    double s = 0.0;
    for (int i = t; i < row_len; i += 256) {
        const float2 x = v[(size_t)k * row_len + i];
        s += (double)x.x * x.x + (double)x.y * x.y;
    }
    s = block_sum_256(s, red_s, t);
    if (t == 0) {
        vk_pow[k] = s;
        double al = alpha_fixed;
        if (!use_fixed) {
            const double g2 = ibo_lin * (double)n_sc / (s * (double)n_ant);
            const double g = sqrt(g2);
            al = 1.0 - exp(-g2) + (1.7724538509055160273 * g * 0.5) * erfc(g);
        }
        alpha[k] = (float)al;
    }
}
struct AgcArgs {
    const float2* h; const float2* v; const float* alpha;
    float2* agc_nfft; float2* hkvk_nfft;  // [n_fft] each
    double* scalers;                       // [2] mean |agc|^2, mean |hkvk|^2 over data sub-carriers (atomicAdd)
    int n_ant, n_users, usr, n_fft, n_sc;
};
__global__ void k_agc(const AgcArgs a) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    double p_agc = 0.0, p_hv = 0.0;
    if (n < a.n_fft) {
        const int s = sc_of_bin(n, a.n_fft, a.n_sc);
        float2 agc = make_float2(1.f, 0.f), hv = make_float2(1.f, 0.f);
        if (s >= 0) {
            agc = hv = make_float2(0.f, 0.f);
            for (int k = 0; k < a.n_ant; ++k) {
                const float2 m = cmul(a.h[(size_t)k * a.n_fft + n], a.v[((size_t)k * a.n_users + a.usr) * a.n_sc + s]);
                hv = cadd(hv, m);
                const float al = a.alpha[k];
                agc = make_float2(fmaf(al, m.x, agc.x), fmaf(al, m.y, agc.y));
            }
            p_agc = (double)agc.x * agc.x + (double)agc.y * agc.y;
            p_hv = (double)hv.x * hv.x + (double)hv.y * hv.y;
        }
        a.agc_nfft[n] = agc;
        a.hkvk_nfft[n] = hv;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        p_agc += __shfl_xor_sync(0xffffffffu, p_agc, o);
        p_hv += __shfl_xor_sync(0xffffffffu, p_hv, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&a.scalers[0], p_agc / a.n_sc);
        atomicAdd(&a.scalers[1], p_hv / a.n_sc);
    }
}

// ---------------------------------------------------------------------------------------------
// Receiver iteration pieces (corrector.py:71-110, :181-205, MU :305-343, :422-449)
// ---------------------------------------------------------------------------------------------
// take the data sub-carriers of a full-band vector (corrector.py:66,176)
__global__ void k_take_sc(const float2* __restrict__ fd, float2* __restrict__ sc, int n_fft, int n_sc) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n_sc) sc[s] = fd[bin_of_sc(s, n_fft, n_sc)];
}
// corr = rx - dist_est ; detect
__global__ void k_rx_correct_detect(const float2* __restrict__ rx_sc, const float2* __restrict__ dist_est, float2* __restrict__ corr_out,
                                    uint8_t* __restrict__ idx_out, int n_sc, Qam q, float scale) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_sc) return;
    float2 c = rx_sc[s];
    if (dist_est) c = csub(c, dist_est[s]);
    if (corr_out) corr_out[s] = c;
    const float inv = 1.0f / scale;
    idx_out[s] = (uint8_t)qam_detect(q, make_float2(c.x * inv, c.y * inv));
}
// dist_est[s] = (sum_k X[k][bin] H[k][bin]) / agc[bin] - symbol   (MCNC, K antennas; CNC: K = 1, h = nullptr, agc = alpha)
__global__ void k_rx_estimate(const float2* __restrict__ x, const float2* __restrict__ h, const float2* __restrict__ agc_nfft,
                              float inv_alpha, const uint8_t* __restrict__ idx, float2* __restrict__ dist_est, int n_ant, int n_fft,
                              int n_sc, Qam q, float scale) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_sc) return;
    const int bin = bin_of_sc(s, n_fft, n_sc);
    float2 acc = make_float2(0.f, 0.f);
    if (h) {
        for (int k = 0; k < n_ant; ++k) acc = cfma(x[(size_t)k * n_fft + bin], h[(size_t)k * n_fft + bin], acc);
        const float2 g = agc_nfft[bin];
        const float d = 1.0f / fmaf(g.x, g.x, g.y * g.y);
        acc = cmulc(acc, g);
        acc.x *= d; acc.y *= d;
    } else {
        acc = x[bin];
        acc.x *= inv_alpha; acc.y *= inv_alpha;
    }
    const float2 sy = qam_symbol(q, idx[s]);
    dist_est[s] = make_float2(acc.x - sy.x * scale, acc.y - sy.y * scale);
}

// ---------------------------------------------------------------------------------------------
// Welch power spectral density as the reference's scripts call scipy.signal.welch
// (main_multiuser/main_multiuser_mrt_precoding_radiation_pattern.py:371-381 and 59 more call sites):
//   welch(x, fs, nfft, nperseg, return_onesided=False) with scipy's defaults window="hann" (periodic), noverlap =
//   nperseg // 2, detrend="constant", scaling="density", average="mean".
// k_welch_segments: one CTA per segment: mean removal (double), window, zero padding to nfft -> rows [n_seg][nfft];
// the rows go through the batched stage FFT; k_welch_reduce averages |X|^2 over the segments per bin (double).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_welch_segments(const float2* __restrict__ x, float2* __restrict__ rows, int nperseg,
                                                        int step, int nfft) {
    __shared__ double red_s[16];
    const int seg = blockIdx.x, t = threadIdx.x;
    const float2* __restrict__ xs = x + (size_t)seg * step;
    double sr = 0.0, si = 0.0;
    for (int i = t; i < nperseg; i += 256) { sr += xs[i].x; si += xs[i].y; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { sr += __shfl_xor_sync(0xffffffffu, sr, o); si += __shfl_xor_sync(0xffffffffu, si, o); }
    if ((t & 31) == 0) { red_s[t >> 5] = sr; red_s[8 + (t >> 5)] = si; }
    __syncthreads();
    double mr = 0.0, mi = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) { mr += red_s[w]; mi += red_s[8 + w]; }
    mr /= nperseg; mi /= nperseg;
    float2* __restrict__ out = rows + (size_t)seg * nfft;
    for (int i = t; i < nfft; i += 256) {
        float2 v = make_float2(0.f, 0.f);
        if (i < nperseg) {
            const double w = 0.5 - 0.5 * cospi(2.0 * (double)i / (double)nperseg);  // scipy get_window("hann", N): periodic
            v = make_float2((float)(((double)xs[i].x - mr) * w), (float)(((double)xs[i].y - mi) * w));
        }
        out[i] = v;
    }
}
// acc[f] += sum_seg |X[seg][f]|^2 (the stage FFT is orthonormal; the caller's scale restores the plain DFT)
__global__ void __launch_bounds__(256) k_welch_reduce(const float2* __restrict__ rows, double* __restrict__ acc, int n_seg, int nfft) {
    const int f = blockIdx.x * 256 + threadIdx.x;
    if (f >= nfft) return;
    double s = 0.0;
    for (int g = 0; g < n_seg; ++g) {
        const float2 v = rows[(size_t)g * nfft + f];
        s += (double)v.x * v.x + (double)v.y * v.y;
    }
    acc[f] += s;
}
__global__ void k_scale_d(double* __restrict__ x, double s, int n) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i < n) x[i] *= s;
}

}  // namespace mmimo

static int stage_kernels_init();
