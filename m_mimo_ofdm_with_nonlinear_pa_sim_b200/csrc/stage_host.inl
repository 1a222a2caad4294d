// Stage-level C ABI (include/mmimo_b200.h "stage-level operations") - included by mmimo_b200.cu.
// Every entry point: H2D of the caller's arrays -> stage kernels on the context stream -> D2H -> synchronize.

static int stage_kernels_init() {
    cudaError_t e = cudaFuncSetAttribute(k_fft4096_rows<+1>, cudaFuncAttributeMaxDynamicSharedMemorySize, FFT_ROWS_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_fft4096_rows<-1>, cudaFuncAttributeMaxDynamicSharedMemorySize, FFT_ROWS_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_fft_generic_rows<+1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 8192 * 8);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_fft_generic_rows<-1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 8192 * 8);
    if (e != cudaSuccess) return fail("stage_kernels_init: %s", cudaGetErrorString(e));
    return 0;
}

// bump allocator over one context-owned device buffer, sized up front by each entry point
struct Arena {
    mmimo_ctx* c; size_t off = 0;
    int begin(size_t bytes) { off = 0; return c->s_a.reserve(bytes + 4096); }
    template <class T> T* get(size_t n) {
        T* p = reinterpret_cast<T*>(static_cast<char*>(c->s_a.p) + off);
        off += (n * sizeof(T) + 255) & ~(size_t)255;
        return p;
    }
    static size_t pad(size_t bytes) { return (bytes + 255) & ~(size_t)255; }
};

static int check_modem(const mmimo_modem_desc* m, int* bps, Qam* q) {
    if (!m) return fail("modem description is NULL");
    const int l2 = log2_side(m->constel_size);
    if (l2 < 0) return fail("Constellation size must be a power of some number, only square QAM supported. (got %d)", m->constel_size);
    if (m->n_fft < 8 || m->n_fft > 8192 || (m->n_fft & (m->n_fft - 1))) return fail("n_fft must be a power of two in 8..8192 (got %d)", m->n_fft);
    if (m->n_sub_carr < 2 || (m->n_sub_carr & 1) || m->n_sub_carr >= m->n_fft) return fail("n_sub_carr must be even and < n_fft (got %d)", m->n_sub_carr);
    if (m->cp_len < 0 || m->cp_len > m->n_fft) return fail("cp_len must be 0..n_fft (got %d)", m->cp_len);
    if (!(m->constel_scale > 0.0)) return fail("constel_scale must be > 0");
    *bps = 2 * l2;
    q->log2side = l2;
    return 0;
}

static int ilog2(int n) { int l = 0; while ((1 << l) < n) ++l; return l; }

// device-level batched transform (in != out)
static int fft_dev(mmimo_ctx* c, bool inverse, int n, int rows, int skip, int cp, const float2* in, float2* out) {
    if (rows <= 0) return 0;
    if (n == 4096) {
        if (inverse) k_fft4096_rows<+1><<<rows, FFT_THREADS, FFT_ROWS_SMEM, c->stream>>>(in, out, c->tw4096.as<float2>(), c->tw256.as<float2>(), skip, cp);
        else k_fft4096_rows<-1><<<rows, FFT_THREADS, FFT_ROWS_SMEM, c->stream>>>(in, out, c->tw4096.as<float2>(), c->tw256.as<float2>(), skip, cp);
    } else {
        if (inverse) k_fft_generic_rows<+1><<<rows, 256, 2 * n * 8, c->stream>>>(in, out, n, ilog2(n), skip, cp);
        else k_fft_generic_rows<-1><<<rows, 256, 2 * n * 8, c->stream>>>(in, out, n, ilog2(n), skip, cp);
    }
    c->launches++;
    CK(cudaGetLastError());
    return 0;
}

template <int KIND>
static void pa_launch(mmimo_ctx* c, size_t n, const float2* in, float2* out, float p0, float p1) {
    if (n % 2 == 0 && (((uintptr_t)in | (uintptr_t)out) & 15) == 0)
        k_pa4<KIND><<<nblk(n / 2, 512), 256, 0, c->stream>>>(reinterpret_cast<const float4*>(in), reinterpret_cast<float4*>(out), n / 2, p0, p1);
    else
        k_pa<KIND><<<nblk(n), 256, 0, c->stream>>>(in, out, n, p0, p1);
}
static int pa_dev(mmimo_ctx* c, const mmimo_pa_desc* pa, size_t n, const float2* in, float2* out) {
    const float p0 = (float)pa->p0, p1 = (float)pa->p1;
    switch (pa->kind) {
        case MMIMO_PA_NONE: CK(cudaMemcpyAsync(out, in, n * 8, cudaMemcpyDeviceToDevice, c->stream)); return 0;
        case MMIMO_PA_SOFT_LIMITER: pa_launch<PA_SOFT_LIMITER>(c, n, in, out, p0, p1); break;
        case MMIMO_PA_RAPP: pa_launch<PA_RAPP>(c, n, in, out, p0, p1); break;
        case MMIMO_PA_TOI: pa_launch<PA_TOI>(c, n, in, out, p0, p1); break;
        default: return fail("unknown pa kind %d", pa->kind);
    }
    c->launches++;
    return 0;
}

// symbol values -> [G][K][n_out] transmitted signal(s).  scratch: 2 buffers of G*K*n_fft complex
struct TxDev {
    const uint8_t* idx; const float2* v; const float2* extra; float w_self, w_extra;
    int n_ant, n_users, per_user;
};
static int array_transmit_dev(mmimo_ctx* c, const mmimo_modem_desc* m, Qam q, const mmimo_pa_desc* pa, const TxDev& t, int flags,
                              float2* grid, float2* td, float2* td_pa, float2* out, float2* out_clean) {
    const int G = t.per_user ? t.n_users : 1;
    const int rows = G * t.n_ant;
    const bool skip = (flags & MMIMO_TX_SKIP_DIST) || !pa || pa->kind == MMIMO_PA_NONE;
    const bool time_dom = flags & MMIMO_TX_TIME_DOMAIN;
    MapArgs ma{};
    ma.idx = t.idx; ma.v = t.v; ma.extra = t.extra; ma.grid = grid;
    ma.n_users = t.n_users; ma.n_ant = t.n_ant; ma.n_fft = m->n_fft; ma.n_sc = m->n_sub_carr; ma.per_user = t.per_user;
    ma.q = q; ma.scale = (float)m->constel_scale; ma.w_self = t.w_self; ma.w_extra = t.w_extra;
    k_map_precode<<<dim3(nblk(m->n_fft / 4), t.n_ant, G), 256, 0, c->stream>>>(ma);
    c->launches++;
    const size_t nfd = (size_t)rows * m->n_fft;
    if (skip) {
        // FFT(IFFT(x)) == x: the clean frequency-domain signal is the grid itself (transceiver.py:112-115)
        if (time_dom) { if (fft_dev(c, true, m->n_fft, rows, 0, m->cp_len, grid, out)) return 1; }
        else CK(cudaMemcpyAsync(out, grid, nfd * 8, cudaMemcpyDeviceToDevice, c->stream));
        if (out_clean) CK(cudaMemcpyAsync(out_clean, out, (size_t)rows * (m->n_fft + (time_dom ? m->cp_len : 0)) * 8, cudaMemcpyDeviceToDevice, c->stream));
        return 0;
    }
    if (fft_dev(c, true, m->n_fft, rows, 0, 0, grid, td)) return 1;
    if (pa_dev(c, pa, nfd, td, td_pa)) return 1;
    if (time_dom) {
        // the PA is memoryless: prefix of the distorted symbol == distorted prefix
        k_add_cp<<<dim3(nblk(m->n_fft + m->cp_len), rows), 256, 0, c->stream>>>(td_pa, out, m->n_fft, m->cp_len);
        c->launches++;
        if (out_clean) { k_add_cp<<<dim3(nblk(m->n_fft + m->cp_len), rows), 256, 0, c->stream>>>(td, out_clean, m->n_fft, m->cp_len); c->launches++; }
    } else {
        if (fft_dev(c, false, m->n_fft, rows, 0, 0, td_pa, out)) return 1;
        if (out_clean) CK(cudaMemcpyAsync(out_clean, grid, nfd * 8, cudaMemcpyDeviceToDevice, c->stream));
    }
    return 0;
}

#define STAGE_PROLOGUE(c)                                   \
    if (!(c)) return fail("NULL context");                  \
    CK(cudaSetDevice((c)->device));                         \
    Arena ar{(c)}

extern "C" int mmimo_fft(mmimo_ctx* c, int inverse, int n, int rows, int skip, int cp, const float* in, float* out) {
    STAGE_PROLOGUE(c);
    if (!in || !out) return fail("mmimo_fft: NULL buffer");
    if (n < 8 || n > 8192 || (n & (n - 1))) return fail("mmimo_fft: n must be a power of two in 8..8192 (got %d)", n);
    if (rows < 0 || skip < 0 || cp < 0 || cp > n) return fail("mmimo_fft: bad rows/skip/cp");
    const size_t ni = (size_t)rows * (n + skip), no = (size_t)rows * (n + cp);
    if (ar.begin(Arena::pad(ni * 8) + Arena::pad(no * 8))) return 1;
    float2* din = ar.get<float2>(ni);
    float2* dout = ar.get<float2>(no);
    CK(cudaMemcpyAsync(din, in, ni * 8, cudaMemcpyHostToDevice, c->stream));
    if (fft_dev(c, inverse != 0, n, rows, skip, cp, din, dout)) return 1;
    CK(cudaMemcpyAsync(out, dout, no * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int mmimo_welch_psd(mmimo_ctx* c, size_t n, const float* x, int nperseg, int nfft, double fs, double* freqs, double* psd) {
    STAGE_PROLOGUE(c);
    if (!x || !psd) return fail("mmimo_welch_psd: NULL buffer");
    if (nfft < 8 || nfft > 8192 || (nfft & (nfft - 1))) return fail("mmimo_welch_psd: nfft must be a power of two in 8..8192 (got %d)", nfft);
    if (nperseg < 2 || nperseg > nfft) return fail("mmimo_welch_psd: nperseg must be 2..nfft (got %d)", nperseg);
    if ((size_t)nperseg > n) return fail("mmimo_welch_psd: the signal is shorter than one segment (%zu < %d)", n, nperseg);
    if (!(fs > 0.0)) return fail("mmimo_welch_psd: fs must be positive");
    const int step = nperseg - nperseg / 2;                       // noverlap = nperseg // 2 (scipy default)
    const size_t n_seg = (n - (size_t)(nperseg / 2)) / (size_t)step;
    const size_t seg_chunk = std::min<size_t>(n_seg, std::max<size_t>(1, ((size_t)256 << 20) / ((size_t)nfft * 16)));
    if (ar.begin(Arena::pad(n * 8) + 2 * Arena::pad(seg_chunk * nfft * 8) + Arena::pad((size_t)nfft * 8))) return 1;
    float2* dx = ar.get<float2>(n);
    float2* rows = ar.get<float2>(seg_chunk * nfft);
    float2* spec = ar.get<float2>(seg_chunk * nfft);
    double* acc = ar.get<double>(nfft);
    CK(cudaMemcpyAsync(dx, x, n * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemsetAsync(acc, 0, (size_t)nfft * 8, c->stream));
    for (size_t s0 = 0; s0 < n_seg; s0 += seg_chunk) {
        const int ns = (int)std::min(seg_chunk, n_seg - s0);
        k_welch_segments<<<ns, 256, 0, c->stream>>>(dx + s0 * step, rows, nperseg, step, nfft);
        c->launches++;
        if (fft_dev(c, false, nfft, ns, 0, 0, rows, spec)) return 1;
        k_welch_reduce<<<(nfft + 255) / 256, 256, 0, c->stream>>>(spec, acc, ns, nfft);
        c->launches++;
    }
    // density scaling 1 / (fs sum w^2), mean over segments, and nfft to undo the orthonormal transform; for the periodic Hann
    // window sum_n w_n^2 = 3 N / 8 (N >= 3)
    double w2 = 0.0;
    for (int i = 0; i < nperseg; ++i) { const double w = 0.5 - 0.5 * cos(2.0 * M_PI * i / nperseg); w2 += w * w; }
    k_scale_d<<<(nfft + 255) / 256, 256, 0, c->stream>>>(acc, (double)nfft / (fs * w2 * (double)n_seg), nfft);
    c->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(psd, acc, (size_t)nfft * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    if (freqs)  // np.fft.fftfreq(nfft, 1 / fs)
        for (int i = 0; i < nfft; ++i) freqs[i] = (double)(i < (nfft + 1) / 2 ? i : i - nfft) * fs / nfft;
    return 0;
}

extern "C" int mmimo_pa_process(mmimo_ctx* c, const mmimo_pa_desc* pa, size_t n, const float* in, float* out) {
    STAGE_PROLOGUE(c);
    if (!pa || !in || !out) return fail("mmimo_pa_process: NULL argument");
    if (ar.begin(2 * Arena::pad(n * 8))) return 1;
    float2* din = ar.get<float2>(n);
    float2* dout = ar.get<float2>(n);
    CK(cudaMemcpyAsync(din, in, n * 8, cudaMemcpyHostToDevice, c->stream));
    if (pa_dev(c, pa, n, din, dout)) return 1;
    CK(cudaMemcpyAsync(out, dout, n * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int mmimo_qam_map(mmimo_ctx* c, const mmimo_modem_desc* m, size_t n_sym, const uint8_t* bits, float* symbols) {
    STAGE_PROLOGUE(c);
    int bps; Qam q;
    if (check_modem(m, &bps, &q)) return 1;
    if (!bits || !symbols) return fail("mmimo_qam_map: NULL buffer");
    if (ar.begin(Arena::pad(n_sym * bps) + Arena::pad(n_sym) + Arena::pad(n_sym * 8))) return 1;
    uint8_t* dbits = ar.get<uint8_t>(n_sym * bps);
    uint8_t* didx = ar.get<uint8_t>(n_sym);
    float2* dsym = ar.get<float2>(n_sym);
    CK(cudaMemcpyAsync(dbits, bits, n_sym * bps, cudaMemcpyHostToDevice, c->stream));
    k_bits_to_idx<<<nblk(n_sym), 256, 0, c->stream>>>(dbits, didx, n_sym, bps);
    k_idx_to_symbols<<<nblk(n_sym), 256, 0, c->stream>>>(didx, dsym, n_sym, q, (float)m->constel_scale);
    c->launches += 2;
    CK(cudaMemcpyAsync(symbols, dsym, n_sym * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int mmimo_qam_detect(mmimo_ctx* c, const mmimo_modem_desc* m, size_t n_sym, const float* in, float* symbols, uint8_t* bits) {
    STAGE_PROLOGUE(c);
    int bps; Qam q;
    if (check_modem(m, &bps, &q)) return 1;
    if (!in) return fail("mmimo_qam_detect: NULL input");
    if (ar.begin(2 * Arena::pad(n_sym * 8) + Arena::pad(n_sym) + Arena::pad(n_sym * bps))) return 1;
    float2* din = ar.get<float2>(n_sym);
    float2* dsym = ar.get<float2>(n_sym);
    uint8_t* didx = ar.get<uint8_t>(n_sym);
    uint8_t* dbits = ar.get<uint8_t>(n_sym * bps);
    CK(cudaMemcpyAsync(din, in, n_sym * 8, cudaMemcpyHostToDevice, c->stream));
    k_detect<<<nblk(n_sym), 256, 0, c->stream>>>(din, didx, symbols ? dsym : nullptr, n_sym, q, (float)m->constel_scale);
    c->launches++;
    if (symbols) CK(cudaMemcpyAsync(symbols, dsym, n_sym * 8, cudaMemcpyDeviceToHost, c->stream));
    if (bits) {
        k_idx_to_bits<<<nblk(n_sym), 256, 0, c->stream>>>(didx, dbits, n_sym, bps);
        c->launches++;
        CK(cudaMemcpyAsync(bits, dbits, n_sym * bps, cudaMemcpyDeviceToHost, c->stream));
    }
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int mmimo_qam_soft_llr(mmimo_ctx* c, const mmimo_modem_desc* m, size_t n_sym, const float* in, const double* noise_var,
                                  size_t n_noise_var, double* llr) {
    STAGE_PROLOGUE(c);
    int bps; Qam q;
    if (check_modem(m, &bps, &q)) return 1;
    if (!in || !noise_var || !llr) return fail("mmimo_qam_soft_llr: NULL buffer");
    if (n_noise_var != 1 && n_noise_var != n_sym) return fail("mmimo_qam_soft_llr: noise_var must hold 1 or n_sym values");
    if (n_sym == 0) return 0;
    if (ar.begin(Arena::pad(n_sym * 8) + Arena::pad(n_noise_var * 8) + Arena::pad(n_sym * bps * 8))) return 1;
    float2* din = ar.get<float2>(n_sym);
    double* dnv = ar.get<double>(n_noise_var);
    double* dllr = ar.get<double>(n_sym * bps);
    CK(cudaMemcpyAsync(din, in, n_sym * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(dnv, noise_var, n_noise_var * 8, cudaMemcpyHostToDevice, c->stream));
    k_soft_llr<<<nblk(n_sym), 256, 0, c->stream>>>(din, dnv, n_noise_var, dllr, n_sym, q, m->constel_scale, bps);
    c->launches++;
    CK(cudaMemcpyAsync(llr, dllr, n_sym * bps * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int mmimo_array_transmit(mmimo_ctx* c, const mmimo_modem_desc* m, const mmimo_pa_desc* pa, int n_ant, int n_users,
                                    const uint8_t* bits, const float* precoder, int flags, float* out, float* out_clean) {
    STAGE_PROLOGUE(c);
    int bps; Qam q;
    if (check_modem(m, &bps, &q)) return 1;
    if (n_ant < 1 || n_ant > 4096) return fail("mmimo_array_transmit: n_ant must be 1..4096 (got %d)", n_ant);
    if (n_users < 1 || n_users > MAX_USERS) return fail("mmimo_array_transmit: n_users must be 1..%d (got %d)", MAX_USERS, n_users);
    if (!bits || !out) return fail("mmimo_array_transmit: NULL buffer");
    const int per_user = (flags & MMIMO_TX_PER_USER) && n_users > 1;
    const int G = per_user ? n_users : 1;
    const size_t n_sym = (size_t)n_users * m->n_sub_carr;
    const size_t nfd = (size_t)G * n_ant * m->n_fft;
    const size_t n_out = (size_t)G * n_ant * (m->n_fft + ((flags & MMIMO_TX_TIME_DOMAIN) ? m->cp_len : 0));
    const size_t nv = (size_t)n_ant * n_users * m->n_sub_carr;
    if (ar.begin(Arena::pad(n_sym * bps) + Arena::pad(n_sym) + Arena::pad(nv * 8) + 3 * Arena::pad(nfd * 8) + 2 * Arena::pad(n_out * 8))) return 1;
    uint8_t* dbits = ar.get<uint8_t>(n_sym * bps);
    uint8_t* didx = ar.get<uint8_t>(n_sym);
    float2* dv = ar.get<float2>(nv);
    float2* grid = ar.get<float2>(nfd);
    float2* td = ar.get<float2>(nfd);
    float2* td_pa = ar.get<float2>(nfd);
    float2* dout = ar.get<float2>(n_out);
    float2* dclean = ar.get<float2>(n_out);
    CK(cudaMemcpyAsync(dbits, bits, n_sym * bps, cudaMemcpyHostToDevice, c->stream));
    if (precoder) CK(cudaMemcpyAsync(dv, precoder, nv * 8, cudaMemcpyHostToDevice, c->stream));
    k_bits_to_idx<<<nblk(n_sym), 256, 0, c->stream>>>(dbits, didx, n_sym, bps);
    c->launches++;
    TxDev t{didx, precoder ? dv : nullptr, nullptr, 1.f, 0.f, n_ant, n_users, per_user};
    if (array_transmit_dev(c, m, q, pa, t, flags, grid, td, td_pa, dout, out_clean ? dclean : nullptr)) return 1;
    CK(cudaMemcpyAsync(out, dout, n_out * 8, cudaMemcpyDeviceToHost, c->stream));
    if (out_clean) CK(cudaMemcpyAsync(out_clean, dclean, n_out * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int mmimo_ofdm_demodulate(mmimo_ctx* c, const mmimo_modem_desc* m, const float* td, float* symbols, uint8_t* bits) {
    STAGE_PROLOGUE(c);
    int bps; Qam q;
    if (check_modem(m, &bps, &q)) return 1;
    if (!td) return fail("mmimo_ofdm_demodulate: NULL input");
    const int n = m->n_fft, ns = m->n_sub_carr;
    if (ar.begin(Arena::pad((size_t)(n + m->cp_len) * 8) + Arena::pad((size_t)n * 8) + Arena::pad((size_t)ns * 8) + Arena::pad(ns) + Arena::pad((size_t)ns * bps))) return 1;
    float2* din = ar.get<float2>(n + m->cp_len);
    float2* fd = ar.get<float2>(n);
    float2* sc = ar.get<float2>(ns);
    uint8_t* didx = ar.get<uint8_t>(ns);
    uint8_t* dbits = ar.get<uint8_t>((size_t)ns * bps);
    CK(cudaMemcpyAsync(din, td, (size_t)(n + m->cp_len) * 8, cudaMemcpyHostToDevice, c->stream));
    if (fft_dev(c, false, n, 1, m->cp_len, 0, din, fd)) return 1;
    k_take_sc<<<nblk(ns), 256, 0, c->stream>>>(fd, sc, n, ns);
    c->launches++;
    if (symbols) CK(cudaMemcpyAsync(symbols, sc, (size_t)ns * 8, cudaMemcpyDeviceToHost, c->stream));
    if (bits) {
        k_detect<<<nblk(ns), 256, 0, c->stream>>>(sc, didx, nullptr, ns, q, (float)m->constel_scale);
        k_idx_to_bits<<<nblk(ns), 256, 0, c->stream>>>(didx, dbits, ns, bps);
        c->launches += 2;
        CK(cudaMemcpyAsync(bits, dbits, (size_t)ns * bps, cudaMemcpyDeviceToHost, c->stream));
    }
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int mmimo_propagate(mmimo_ctx* c, int n_ant, int n, const float* x, const float* h, int sum_signals, float* out) {
    STAGE_PROLOGUE(c);
    if (!x || !h || !out) return fail("mmimo_propagate: NULL buffer");
    if (n_ant < 1 || n < 1) return fail("mmimo_propagate: bad shape");
    const size_t nn = (size_t)n_ant * n, no = sum_signals ? (size_t)n : nn;
    if (ar.begin(2 * Arena::pad(nn * 8) + Arena::pad(no * 8))) return 1;
    float2* dx = ar.get<float2>(nn);
    float2* dh = ar.get<float2>(nn);
    float2* dout = ar.get<float2>(no);
    CK(cudaMemcpyAsync(dx, x, nn * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(dh, h, nn * 8, cudaMemcpyHostToDevice, c->stream));
    k_propagate<<<dim3(nblk(n), sum_signals ? 1 : n_ant), 256, 0, c->stream>>>(dx, dh, dout, n_ant, n, sum_signals);
    c->launches++;
    CK(cudaMemcpyAsync(out, dout, no * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int mmimo_awgn_add(mmimo_ctx* c, size_t n, const float* in, const float* unit_noise, double scale, float* out) {
    STAGE_PROLOGUE(c);
    if (!in || !unit_noise || !out) return fail("mmimo_awgn_add: NULL buffer");
    if (ar.begin(3 * Arena::pad(n * 8))) return 1;
    float2* di = ar.get<float2>(n);
    float2* dn = ar.get<float2>(n);
    float2* dout = ar.get<float2>(n);
    CK(cudaMemcpyAsync(di, in, n * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(dn, unit_noise, n * 8, cudaMemcpyHostToDevice, c->stream));
    k_axpy_c<<<nblk(n), 256, 0, c->stream>>>(di, dn, dout, (float)scale, n);
    c->launches++;
    CK(cudaMemcpyAsync(out, dout, n * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int mmimo_count_mismatched_bits(mmimo_ctx* c, size_t n, const uint8_t* a, const uint8_t* b, uint64_t* out) {
    STAGE_PROLOGUE(c);
    if (!a || !b || !out) return fail("mmimo_count_mismatched_bits: NULL buffer");
    if (ar.begin(2 * Arena::pad(n) + 256)) return 1;
    uint8_t* da = ar.get<uint8_t>(n);
    uint8_t* db = ar.get<uint8_t>(n);
    unsigned long long* dc = ar.get<unsigned long long>(1);
    CK(cudaMemcpyAsync(da, a, n, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(db, b, n, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemsetAsync(dc, 0, 8, c->stream));
    k_count_mismatch<<<std::max(1u, std::min(nblk(n), 1024u)), 256, 0, c->stream>>>(da, db, n, dc);
    c->launches++;
    unsigned long long h = 0;
    CK(cudaMemcpyAsync(&h, dc, 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    *out = h;
    return 0;
}

static int chan_common(mmimo_ctx* c, Arena& ar, int n_ant, int n_fft, const double* tx_pos, const double* tx_gain_db, const double* freqs,
                       ChanFullArgs* ca, size_t extra_bytes) {
    if (n_ant < 1 || n_fft < 1) return fail("channel: bad shape");
    if (!tx_pos || !freqs) return fail("channel: NULL geometry");
    if (ar.begin(Arena::pad((size_t)n_ant * 24) + Arena::pad((size_t)n_ant * 8) + Arena::pad((size_t)n_fft * 8) + Arena::pad((size_t)n_ant * n_fft * 8) + extra_bytes)) return 1;
    double* dpos = ar.get<double>((size_t)n_ant * 3);
    double* dgain = ar.get<double>(n_ant);
    double* dfr = ar.get<double>(n_fft);
    float2* dh = ar.get<float2>((size_t)n_ant * n_fft);
    CK(cudaMemcpyAsync(dpos, tx_pos, (size_t)n_ant * 24, cudaMemcpyHostToDevice, c->stream));
    if (tx_gain_db) CK(cudaMemcpyAsync(dgain, tx_gain_db, (size_t)n_ant * 8, cudaMemcpyHostToDevice, c->stream));
    else CK(cudaMemsetAsync(dgain, 0, (size_t)n_ant * 8, c->stream));
    CK(cudaMemcpyAsync(dfr, freqs, (size_t)n_fft * 8, cudaMemcpyHostToDevice, c->stream));
    ca->tx_pos = dpos; ca->tx_gain = dgain; ca->freqs = dfr; ca->n_fft = n_fft; ca->h_out = dh;
    return 0;
}

extern "C" int mmimo_channel_geometric(mmimo_ctx* c, int kind, int n_ant, int n_fft, const double* tx_pos, const double* tx_gain_db,
                                       double rx_gain_db, const double* rx_pos, const double* freqs, int skip_attenuation, float* h_out) {
    STAGE_PROLOGUE(c);
    if (kind != MMIMO_CHAN_LOS && kind != MMIMO_CHAN_TWO_PATH) return fail("mmimo_channel_geometric: kind must be LOS or TWO_PATH");
    if (!rx_pos || !h_out) return fail("mmimo_channel_geometric: NULL argument");
    ChanFullArgs ca{};
    if (chan_common(c, ar, n_ant, n_fft, tx_pos, tx_gain_db, freqs, &ca, 0)) return 1;
    ca.rx[0] = rx_pos[0]; ca.rx[1] = rx_pos[1]; ca.rx[2] = rx_pos[2];
    ca.rx_gain_db = rx_gain_db; ca.kind = kind == MMIMO_CHAN_TWO_PATH; ca.skip_attenuation = skip_attenuation;
    k_channel_full<<<dim3(nblk(n_fft), n_ant), 256, 0, c->stream>>>(ca);
    c->launches++;
    CK(cudaMemcpyAsync(h_out, ca.h_out, (size_t)n_ant * n_fft * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int mmimo_channel_fspl(mmimo_ctx* c, int n_ant, int n_fft, const double* tx_pos, const double* tx_gain_db, double rx_gain_db,
                                  const double* rx_pos, const double* freqs, const float* coeffs, float* h_out) {
    STAGE_PROLOGUE(c);
    if (!rx_pos || !h_out) return fail("mmimo_channel_fspl: NULL argument");
    ChanFullArgs ca{};
    const size_t nn = (size_t)n_ant * n_fft;
    if (chan_common(c, ar, n_ant, n_fft, tx_pos, tx_gain_db, freqs, &ca, Arena::pad(nn * 8))) return 1;
    ca.rx[0] = rx_pos[0]; ca.rx[1] = rx_pos[1]; ca.rx[2] = rx_pos[2];
    ca.rx_gain_db = rx_gain_db;
    float2* dco = nullptr;
    if (coeffs) {
        dco = ar.get<float2>(nn);
        CK(cudaMemcpyAsync(dco, coeffs, nn * 8, cudaMemcpyHostToDevice, c->stream));
    }
    k_fspl_apply<<<dim3(nblk(n_fft), n_ant), 256, 0, c->stream>>>(ca, dco);
    c->launches++;
    CK(cudaMemcpyAsync(h_out, ca.h_out, nn * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int mmimo_channel_random_paths(mmimo_ctx* c, int n_ant, int n_fft, const double* tx_pos, const double* freqs, int n_paths,
                                          const double* angles, const double* taus, float* h_out) {
    STAGE_PROLOGUE(c);
    if (!angles || !taus || !h_out || n_paths < 1) return fail("mmimo_channel_random_paths: bad argument");
    ChanFullArgs ca{};
    if (chan_common(c, ar, n_ant, n_fft, tx_pos, nullptr, freqs, &ca, 2 * Arena::pad((size_t)n_paths * 8))) return 1;
    double* da = ar.get<double>(n_paths);
    double* dt = ar.get<double>(n_paths);
    CK(cudaMemcpyAsync(da, angles, (size_t)n_paths * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(dt, taus, (size_t)n_paths * 8, cudaMemcpyHostToDevice, c->stream));
    RandPathArgs ra{ca.tx_pos, ca.freqs, da, dt, n_paths, n_fft, ca.h_out};
    k_channel_random_paths<<<dim3(nblk(n_fft), n_ant), 256, 0, c->stream>>>(ra);
    c->launches++;
    CK(cudaMemcpyAsync(h_out, ca.h_out, (size_t)n_ant * n_fft * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int mmimo_precoding(mmimo_ctx* c, int mode, int sep_carriers, int n_users, int n_ant, int n_fft, int n_sc, const float* h,
                               float* v) {
    STAGE_PROLOGUE(c);
    if (!h || !v) return fail("mmimo_precoding: NULL buffer");
    if (mode < 0 || mode > 2) return fail("mmimo_precoding: unknown mode %d", mode);
    if (n_users < 1 || n_users > MAX_USERS) return fail("mmimo_precoding: n_users must be 1..%d", MAX_USERS);
    if (mode == MMIMO_PREC_ZF && n_users >= n_ant) return fail("mmimo_precoding: zero forcing needs n_users < n_ant");
    if (sep_carriers && (mode == MMIMO_PREC_ZF || n_sc % n_users)) return fail("mmimo_precoding: sep_carriers needs MRT/phase mode and n_sub_carr divisible by n_users");
    if (n_ant < 1 || n_sc < 2 || (n_sc & 1) || n_sc >= n_fft) return fail("mmimo_precoding: bad shape");
    const size_t nh = (size_t)n_users * n_ant * n_fft, nv = (size_t)n_ant * (sep_carriers ? 1 : n_users) * n_sc;
    if (ar.begin(Arena::pad(nh * 8) + Arena::pad(nv * 8))) return 1;
    float2* dh = ar.get<float2>(nh);
    float2* dv = ar.get<float2>(nv);
    CK(cudaMemcpyAsync(dh, h, nh * 8, cudaMemcpyHostToDevice, c->stream));
    PrecArgs pa{dh, dv, n_users, n_ant, n_fft, n_sc, mode, sep_carriers};
    k_precoding<<<nblk(n_sc, 64), 64, 0, c->stream>>>(pa);
    c->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(v, dv, nv * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int mmimo_agc(mmimo_ctx* c, int n_ant, int n_users, int usr_idx, int n_fft, int n_sc, const float* h, const float* v,
                         double ibo_db, const double* alpha_fixed, float* agc_nfft, float* hkvk_nfft, double* scalers, double* vk_pow,
                         float* alpha) {
    STAGE_PROLOGUE(c);
    if (!h || !v) return fail("mmimo_agc: NULL buffer");
    if (n_ant < 1 || n_users < 1 || usr_idx < 0 || usr_idx >= n_users) return fail("mmimo_agc: bad shape");
    const size_t nh = (size_t)n_ant * n_fft, nv = (size_t)n_ant * n_users * n_sc;
    if (ar.begin(Arena::pad(nh * 8) + Arena::pad(nv * 8) + 2 * Arena::pad((size_t)n_fft * 8) + Arena::pad((size_t)n_ant * 8) + Arena::pad((size_t)n_ant * 4) + 256)) return 1;
    float2* dh = ar.get<float2>(nh);
    float2* dv = ar.get<float2>(nv);
    float2* dagc = ar.get<float2>(n_fft);
    float2* dhv = ar.get<float2>(n_fft);
    double* dpow = ar.get<double>(n_ant);
    float* dal = ar.get<float>(n_ant);
    double* dsc = ar.get<double>(2);
    CK(cudaMemcpyAsync(dh, h, nh * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(dv, v, nv * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemsetAsync(dsc, 0, 16, c->stream));
    k_vk_pow<<<n_ant, 256, 0, c->stream>>>(dv, dpow, dal, n_ant, n_users * n_sc, pow(10.0, ibo_db / 10.0), n_sc,
                                          alpha_fixed ? *alpha_fixed : 0.0, alpha_fixed != nullptr);
    AgcArgs aa{dh, dv, dal, dagc, dhv, dsc, n_ant, n_users, usr_idx, n_fft, n_sc};
    k_agc<<<nblk(n_fft), 256, 0, c->stream>>>(aa);
    c->launches += 2;
    if (agc_nfft) CK(cudaMemcpyAsync(agc_nfft, dagc, (size_t)n_fft * 8, cudaMemcpyDeviceToHost, c->stream));
    if (hkvk_nfft) CK(cudaMemcpyAsync(hkvk_nfft, dhv, (size_t)n_fft * 8, cudaMemcpyDeviceToHost, c->stream));
    if (scalers) CK(cudaMemcpyAsync(scalers, dsc, 16, cudaMemcpyDeviceToHost, c->stream));
    if (vk_pow) CK(cudaMemcpyAsync(vk_pow, dpow, (size_t)n_ant * 8, cudaMemcpyDeviceToHost, c->stream));
    if (alpha) CK(cudaMemcpyAsync(alpha, dal, (size_t)n_ant * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

// shared iteration loop of the four receivers (corrector.py): K antennas + channel (MCNC) or a single-antenna model (CNC)
struct RxLoop {
    const mmimo_modem_desc* m; const mmimo_pa_desc* pa; Qam q; int bps;
    int n_ant, n_users, usr_idx, n_iter;
    const float2* h; const float2* v; const float2* agc;   // MCNC (device)
    float inv_alpha;                                        // CNC
    const float2* extra; float w_self, w_extra;             // CncMu: other user's symbols
    uint8_t* idx_users;                                     // [U][n_sc] device: row usr_idx is rewritten every iteration
};
static int rx_loop_dev(mmimo_ctx* c, Arena& ar, const RxLoop& r, const float2* in_fd, uint8_t* det_bits_dev, float2* corr_dev) {
    const mmimo_modem_desc* m = r.m;
    const int n = m->n_fft, ns = m->n_sub_carr;
    const size_t nfd = (size_t)r.n_ant * n;
    float2* rx_sc = ar.get<float2>(ns);
    float2* dist = ar.get<float2>(ns);
    float2* grid = ar.get<float2>(nfd);
    float2* td = ar.get<float2>(nfd);
    float2* td_pa = ar.get<float2>(nfd);
    float2* x = ar.get<float2>(nfd);
    uint8_t* idx_row = r.idx_users + (size_t)r.usr_idx * ns;
    k_take_sc<<<nblk(ns), 256, 0, c->stream>>>(in_fd, rx_sc, n, ns);
    c->launches++;
    for (int it = 0; it <= r.n_iter; ++it) {
        k_rx_correct_detect<<<nblk(ns), 256, 0, c->stream>>>(rx_sc, it ? dist : nullptr, corr_dev ? corr_dev + (size_t)it * ns : nullptr,
                                                           idx_row, ns, r.q, (float)m->constel_scale);
        c->launches++;
        if (det_bits_dev) {
            k_idx_to_bits<<<nblk(ns), 256, 0, c->stream>>>(idx_row, det_bits_dev + (size_t)it * ns * r.bps, ns, r.bps);
            c->launches++;
        }
        if (it == r.n_iter) break;  // the reference regenerates once more and discards the result (corrector.py:198-205)
        TxDev t{r.idx_users, r.v, r.extra, r.w_self, r.w_extra, r.n_ant, r.n_users, 0};
        if (array_transmit_dev(c, m, r.q, r.pa, t, 0, grid, td, td_pa, x, nullptr)) return 1;
        k_rx_estimate<<<nblk(ns), 256, 0, c->stream>>>(x, r.h, r.agc, r.inv_alpha, idx_row, dist, r.n_ant, n, ns, r.q, (float)m->constel_scale);
        c->launches++;
    }
    CK(cudaGetLastError());
    return 0;
}
static size_t rx_loop_bytes(const mmimo_modem_desc* m, int n_ant) {
    return 2 * Arena::pad((size_t)m->n_sub_carr * 8) + 4 * Arena::pad((size_t)n_ant * m->n_fft * 8);
}

extern "C" int mmimo_cnc_receive(mmimo_ctx* c, const mmimo_modem_desc* m, const mmimo_pa_desc* pa, double alpha, int n_iter,
                                 const float* in_sig_fd, const float* other_usr_symbols, uint8_t* det_bits, float* corr) {
    STAGE_PROLOGUE(c);
    int bps; Qam q;
    if (check_modem(m, &bps, &q)) return 1;
    if (!pa || !in_sig_fd) return fail("mmimo_cnc_receive: NULL argument");
    if (n_iter < 0 || n_iter > 1024) return fail("mmimo_cnc_receive: n_iter must be 0..1024");
    if (!(alpha > 0.0)) return fail("mmimo_cnc_receive: alpha must be > 0");
    const int n = m->n_fft, ns = m->n_sub_carr;
    const size_t nb = (size_t)(n_iter + 1) * ns * bps, ncorr = (size_t)(n_iter + 1) * ns;
    if (ar.begin(rx_loop_bytes(m, 1) + Arena::pad((size_t)n * 8) + Arena::pad((size_t)ns * 8) + Arena::pad(ns) + Arena::pad(nb) + Arena::pad(ncorr * 8))) return 1;
    float2* din = ar.get<float2>(n);
    float2* dextra = ar.get<float2>(ns);
    uint8_t* didx = ar.get<uint8_t>(ns);
    uint8_t* dbits = ar.get<uint8_t>(nb);
    float2* dcorr = ar.get<float2>(ncorr);
    CK(cudaMemcpyAsync(din, in_sig_fd, (size_t)n * 8, cudaMemcpyHostToDevice, c->stream));
    if (other_usr_symbols) CK(cudaMemcpyAsync(dextra, other_usr_symbols, (size_t)ns * 8, cudaMemcpyHostToDevice, c->stream));
    RxLoop r{};
    r.m = m; r.pa = pa; r.q = q; r.bps = bps; r.n_ant = 1; r.n_users = 1; r.usr_idx = 0; r.n_iter = n_iter;
    r.inv_alpha = (float)(1.0 / alpha);
    const float w = other_usr_symbols ? 0.70710678118654752440f : 1.0f;  // corrector.py:322
    r.extra = other_usr_symbols ? dextra : nullptr; r.w_self = w; r.w_extra = w;
    r.idx_users = didx;
    if (rx_loop_dev(c, ar, r, din, det_bits ? dbits : nullptr, corr ? dcorr : nullptr)) return 1;
    if (det_bits) CK(cudaMemcpyAsync(det_bits, dbits, nb, cudaMemcpyDeviceToHost, c->stream));
    if (corr) CK(cudaMemcpyAsync(corr, dcorr, ncorr * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int mmimo_mcnc_receive(mmimo_ctx* c, const mmimo_modem_desc* m, const mmimo_pa_desc* pa, int n_ant, int n_users, int usr_idx,
                                  const float* h, const float* v, const float* agc_nfft, int n_iter, const float* in_sig_fd,
                                  const uint8_t* other_usr_bits, uint8_t* det_bits, float* corr) {
    STAGE_PROLOGUE(c);
    int bps; Qam q;
    if (check_modem(m, &bps, &q)) return 1;
    if (!h || !v || !agc_nfft || !in_sig_fd) return fail("mmimo_mcnc_receive: NULL argument");
    if (n_ant < 1 || n_users < 1 || n_users > MAX_USERS || usr_idx < 0 || usr_idx >= n_users) return fail("mmimo_mcnc_receive: bad shape");
    if (n_users > 1 && !other_usr_bits) return fail("mmimo_mcnc_receive: other_usr_bits is required with n_users > 1");
    if (n_iter < 0 || n_iter > 1024) return fail("mmimo_mcnc_receive: n_iter must be 0..1024");
    const int n = m->n_fft, ns = m->n_sub_carr;
    const size_t nh = (size_t)n_ant * n, nv = (size_t)n_ant * n_users * ns;
    const size_t nb = (size_t)(n_iter + 1) * ns * bps, ncorr = (size_t)(n_iter + 1) * ns;
    const size_t nob = (size_t)(n_users - 1) * ns * bps;
    if (ar.begin(rx_loop_bytes(m, n_ant) + Arena::pad(nh * 8) + Arena::pad(nv * 8) + 2 * Arena::pad((size_t)n * 8) + Arena::pad((size_t)n_users * ns) +
                 Arena::pad(nob + 1) + Arena::pad((size_t)n_users * ns) + Arena::pad(nb) + Arena::pad(ncorr * 8))) return 1;
    float2* dh = ar.get<float2>(nh);
    float2* dv = ar.get<float2>(nv);
    float2* dagc = ar.get<float2>(n);
    float2* din = ar.get<float2>(n);
    uint8_t* didx = ar.get<uint8_t>((size_t)n_users * ns);
    uint8_t* dob = ar.get<uint8_t>(nob + 1);
    uint8_t* doidx = ar.get<uint8_t>((size_t)n_users * ns);
    uint8_t* dbits = ar.get<uint8_t>(nb);
    float2* dcorr = ar.get<float2>(ncorr);
    CK(cudaMemcpyAsync(dh, h, nh * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(dv, v, nv * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(dagc, agc_nfft, (size_t)n * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(din, in_sig_fd, (size_t)n * 8, cudaMemcpyHostToDevice, c->stream));
    if (n_users > 1) {  // the other users' symbol values, in user order around usr_idx (corrector.py:437-440)
        CK(cudaMemcpyAsync(dob, other_usr_bits, nob, cudaMemcpyHostToDevice, c->stream));
        k_bits_to_idx<<<nblk((size_t)(n_users - 1) * ns), 256, 0, c->stream>>>(dob, doidx, (size_t)(n_users - 1) * ns, bps);
        c->launches++;
        for (int u = 0, o = 0; u < n_users; ++u) {
            if (u == usr_idx) continue;
            CK(cudaMemcpyAsync(didx + (size_t)u * ns, doidx + (size_t)o * ns, ns, cudaMemcpyDeviceToDevice, c->stream));
            ++o;
        }
    }
    RxLoop r{};
    r.m = m; r.pa = pa; r.q = q; r.bps = bps; r.n_ant = n_ant; r.n_users = n_users; r.usr_idx = usr_idx; r.n_iter = n_iter;
    r.h = dh; r.v = dv; r.agc = dagc; r.inv_alpha = 1.f; r.extra = nullptr; r.w_self = 1.f; r.w_extra = 0.f;
    r.idx_users = didx;
    if (rx_loop_dev(c, ar, r, din, det_bits ? dbits : nullptr, corr ? dcorr : nullptr)) return 1;
    if (det_bits) CK(cudaMemcpyAsync(det_bits, dbits, nb, cudaMemcpyDeviceToHost, c->stream));
    if (corr) CK(cudaMemcpyAsync(corr, dcorr, ncorr * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// mmimo_measure_stage_kernels (bench.py, "stages"): the unfused building-block kernels of one array transmit + propagate
// (k_map_precode -> k_fft4096_rows<+1> -> k_pa<soft limiter> -> k_fft4096_rows<-1> -> k_propagate) timed one by one with
// CUDA events on the context stream over `rows` device-resident rows of 4096 samples (rows * 32 KiB per buffer: larger than
// L2 for rows >= 8192), `reps` timed launches each after one warm-up.  ms_out[5] = average ms per launch in that order.
// These are HBM-bound kernels; a single stage call of the drop-in classes (64 rows, host buffers) is latency-bound instead.
// ---------------------------------------------------------------------------------------------
extern "C" int mmimo_measure_stage_kernels(mmimo_ctx* c, int rows, int reps, double* ms_out) {
    STAGE_PROLOGUE(c);
    if (!ms_out || rows < 1 || rows > 65535 || reps < 1) return fail("mmimo_measure_stage_kernels: rows 1..65535, reps >= 1");
    const int n = 4096, ns = 2048;
    const size_t nfd = (size_t)rows * n, nv = (size_t)rows * ns;
    if (ar.begin(4 * Arena::pad(nfd * 8) + 2 * Arena::pad(nv * 8) + Arena::pad(ns))) return 1;
    float2* grid = ar.get<float2>(nfd);
    float2* td = ar.get<float2>(nfd);
    float2* td_pa = ar.get<float2>(nfd);
    float2* fd = ar.get<float2>(nfd);
    float2* v = ar.get<float2>(nv);
    uint8_t* idx = ar.get<uint8_t>(ns);
    CK(cudaMemsetAsync(v, 0x3c, nv * 8, c->stream));  // 0x3c3c3c3c = 0.0115f
    CK(cudaMemsetAsync(idx, 0x15, ns, c->stream));
    MapArgs ma{};
    ma.idx = idx; ma.v = v; ma.extra = nullptr; ma.grid = grid;
    ma.n_users = 1; ma.n_ant = rows; ma.n_fft = n; ma.n_sc = ns; ma.per_user = 0;
    ma.q.log2side = 3; ma.scale = 1.f; ma.w_self = 1.f; ma.w_extra = 0.f;
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    int rc = 0;
    for (int kern = 0; kern < 5 && !rc; ++kern) {
        for (int rep = -1; rep < reps; ++rep) {  // rep -1: warm-up, then `reps` launches inside one event pair
            if (rep == 0) cudaEventRecord(a, c->stream);
            switch (kern) {
                case 0: k_map_precode<<<dim3(nblk(n / 4), rows, 1), 256, 0, c->stream>>>(ma); break;
                case 1: rc = fft_dev(c, true, n, rows, 0, 0, grid, td); break;
                case 2: pa_launch<PA_SOFT_LIMITER>(c, nfd, td, td_pa, 1e-4f, 0.f); break;
                case 3: rc = fft_dev(c, false, n, rows, 0, 0, td_pa, fd); break;
                case 4: k_propagate<<<dim3(nblk(n), rows), 256, 0, c->stream>>>(fd, grid, td, rows, n, 0); break;
            }
            if (kern != 1 && kern != 3) c->launches++;
            if (rc) break;
        }
        cudaEventRecord(b, c->stream);
        if (cudaEventSynchronize(b) != cudaSuccess || cudaGetLastError() != cudaSuccess) { rc = fail("mmimo_measure_stage_kernels: kernel %d failed", kern); break; }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, a, b);
        ms_out[kern] = (double)ms / reps;
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    return rc;
}
