// libmmimo_b200.so - C ABI (include/mmimo_b200.h) over the sm_100a link kernels.
// Host side only orchestrates: uploads, launches, downloads.  There is no CPU compute path.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/mmimo_b200.h"
#include "prep_kernels.cuh"
#include "mu_kernels.cuh"
#include "frame_kernel_r64.cuh"
#include "frame_kernel_t64.cuh"
#include "cnc_loop_t64.cuh"
#include "stage_kernels.cuh"

using namespace mmimo;

static thread_local std::string g_err;

static int fail(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return 1;
}

#define CK(call)                                                                                        \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess) return fail("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// growable device buffer owned by the context
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes) {
        if (bytes <= cap) return 0;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) return fail("cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
        cap = bytes;
        return 0;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct mmimo_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t own_stream = nullptr;
    bool configured = false;
    mmimo_link_config cfg{};
    std::vector<double> tx_pos_h, freqs_h;
    uint64_t launches = 0;
    int n_sm = 148;
    size_t scratch_budget = (size_t)12 << 30;
    // tables
    DevBuf tw4096, tw256, tw_t64, freq_slot, freq_bin, tx_pos;
    // frame-kernel form, read once from MMIMO_FRAME_KERNEL at mmimo_create: 0 = scaled-tangent radix-64 (default),
    // 1 = radix-64 with separate twiddle multiplies (round-1 form), 2 = three-pass radix-16; measurement knob only
    int frame_kernel = 0;
    // start offset (SM cycles) of antenna groups 2,3 of k_link_frame_t64 (FrameArgs::stagger); MMIMO_T64_STAGGER overrides
    int t64_stagger = 0;
    int cnc_split = CNC_SPLIT;  // CNC frames: pass 0 in the frame kernel, the receiver loop in k_cnc_loop_t64 (MMIMO_CNC_SPLIT=0/1)
    // closed-form realisation kernel for LOS + MRT (k_prep_los); MMIMO_PREP_ANALYTIC=0 keeps k_prep_realisation (A/B knob)
    int prep_analytic = 1;
    int ctas_per_sm_r16 = 2;
    // per-chunk scratch
    DevBuf h_true, h_loop, csi_nz, rxeq_scr, v, sscale, agc_inv, hkvk_inv, scal, alpha, rx_pos, tx_idx, noise, det, err, rx_eq, corr, stage;
    DevBuf cnc_ones, cnc_v, cnc_agc_inv;
    // Host inputs of the distorted frames (receiver positions, symbol values, noise: what a reference user hands over per
    // frame, mp_model.py:190-212) travel on their own stream into one of two staging slots, so that the copies of batch
    // n + 1 run under the frame kernel of batch n.  copied: H2D of the slot finished (copy stream); consumed: the slot has
    // been permuted into the working buffers (compute stream) and may be overwritten.
    struct InSlot { DevBuf pos, idx, nz; cudaEvent_t copied = nullptr, consumed = nullptr; bool in_use = false; };
    cudaStream_t copy_stream = nullptr;
    InSlot in_slot[2];
    unsigned slot_ctr = 0;
    // Pipelined realisations (MMIMO_PIPELINE_PREP=1, off by default; mmimo_link_enqueue with device-generated channels): the
    // per-realisation buffers exist twice; before every batch the members above are swapped with `alt` and the channel /
    // realisation kernels of the batch run on prep_stream while the previous batch's frame kernel is still on the compute
    // stream.  ready: the set has been filled (prep stream); released: the frame kernel that read the set has finished
    // (compute stream).  Measured (profiles/r02_pipelined_prep.txt): results identical, but the light k_gen_channel_bg
    // (64 threads, 56 registers, no shared memory, no barrier) is NOT scheduled beside resident frame-kernel CTAs although
    // registers (4096 free), shared memory (10.6 KB free) and barriers would fit - with equal stream priorities it wins the
    // race for the idle SMs and runs first, with a lower priority it waits for the frame kernel to end: 18.02-18.12 vs
    // 18.13-18.18 ms per step, i.e. nothing worth twice the channel memory.
    struct PrepSet {
        DevBuf h_true, h_loop, csi_nz, sscale, agc_inv, hkvk_inv, scal, alpha, rx_pos;
        cudaEvent_t ready = nullptr, released = nullptr;
        bool used = false;
    };
    PrepSet alt;
    cudaEvent_t cur_ready = nullptr, cur_released = nullptr;
    bool cur_used = false;
    cudaStream_t prep_stream = nullptr;
    int pipeline_prep = 0;
    DevBuf counters;  // [0..63] distorted iterations, [64] clean
    // multi-point batch in flight (mmimo_link_sweep): device tables of 10^(x/10) per point, 64 counters per point
    struct Sweep { const double* snr_lin = nullptr; const double* ibo_lin = nullptr; int fpp = 0; unsigned long long* counters = nullptr; } sweep;
    DevBuf sweep_tab, sweep_cnt;
    // stage-op scratch
    DevBuf s_a, s_b, s_c, s_d;
    // optional per-kernel CUDA-event timing (mmimo_profile)
    struct ProfRec { int kind; cudaEvent_t a, b; };
    bool prof = false;
    std::vector<ProfRec> prof_recs;
};

enum { PROF_CHAN = 0, PROF_PREP = 1, PROF_FRAME = 2, PROF_OTHER = 3 };
static void prof_begin(mmimo_ctx* c, int kind) {
    if (!c->prof) return;
    mmimo_ctx::ProfRec r;
    r.kind = kind;
    cudaEventCreate(&r.a);
    cudaEventCreate(&r.b);
    cudaEventRecord(r.a, c->stream);
    c->prof_recs.push_back(r);
}
static void prof_end(mmimo_ctx* c) {
    if (!c->prof || c->prof_recs.empty()) return;
    cudaEventRecord(c->prof_recs.back().b, c->stream);
}

template <class T>
__global__ void k_to_native(const T* __restrict__ src, T* __restrict__ dst, size_t rows) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * N_SC) return;
    const size_t r = i / N_SC;
    const int p = (int)(i % N_SC);
    dst[i] = src[r * N_SC + slot_to_sc(p)];
}
template <class T>
__global__ void k_from_native(const T* __restrict__ src, T* __restrict__ dst, size_t rows) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * N_SC) return;
    const size_t r = i / N_SC;
    const int s = (int)(i % N_SC);
    dst[i] = src[r * N_SC + sc_to_slot(s)];
}
// 1/z -> z for the AGC dump (agc_inv holds 1/(64 agc))
__global__ void k_agc_from_inv(const float2* __restrict__ inv, float2* __restrict__ out, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float2 z = inv[i];
    const double m = ((double)z.x * z.x + (double)z.y * z.y) * 64.0;
    out[i] = make_float2((float)(z.x / m), (float)(-z.y / m));
}
__global__ void k_scale(float2* __restrict__ x, float s, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { x[i].x *= s; x[i].y *= s; }
}
__global__ void k_fill_f(float* __restrict__ x, float v, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) x[i] = v;
}
__global__ void k_fill_c(float2* __restrict__ x, float2 v, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) x[i] = v;
}

static inline unsigned nblk(size_t n, int bs = 256) { return (unsigned)((n + bs - 1) / bs); }

extern "C" const char* mmimo_last_error(void) { return g_err.c_str(); }
extern "C" const char* mmimo_version(void) { return "mmimo_b200 0.1 (sm_100a)"; }

extern "C" int mmimo_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { fail("cudaGetDeviceCount failed"); return -1; }
    return n;
}

static int create_body(mmimo_ctx* c);
extern "C" void mmimo_destroy(mmimo_ctx* c);

extern "C" int mmimo_create(mmimo_ctx** out, int device) {
    if (!out) return fail("mmimo_create: out is NULL");
    *out = nullptr;
    int n = 0;
    CK(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return fail("mmimo_create: device %d not present (%d CUDA devices)", device, n);
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail("mmimo_create: device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
    mmimo_ctx* c = new mmimo_ctx();
    c->device = device;
    c->n_sm = prop.multiProcessorCount;
    if (const char* fk = getenv("MMIMO_FRAME_KERNEL")) c->frame_kernel = !strcmp(fk, "r64") ? 1 : !strcmp(fk, "r16") ? 2 : 0;
    if (const char* e = getenv("MMIMO_PREP_ANALYTIC")) c->prep_analytic = atoi(e) != 0;
    if (const char* e = getenv("MMIMO_PIPELINE_PREP")) c->pipeline_prep = atoi(e) != 0;
    if (const char* e = getenv("MMIMO_T64_STAGGER")) c->t64_stagger = std::max(0, atoi(e));
    if (const char* e = getenv("MMIMO_CNC_SPLIT")) c->cnc_split = atoi(e) != 0;
    if (const char* e = getenv("MMIMO_CTAS_PER_SM")) c->ctas_per_sm_r16 = std::min(8, std::max(1, atoi(e)));
    const int rc = create_body(c);
    if (rc) { mmimo_destroy(c); return rc; }
    *out = c;
    return 0;
}

// everything of mmimo_create that can fail after the context exists (the caller destroys it on failure)
static int create_body(mmimo_ctx* c) {
    // The compute stream gets the highest priority, the prep stream the lowest: when the frame kernel of batch n and the
    // background channel kernel of batch n + 1 become eligible together, the block scheduler must place the 148 frame CTAs
    // (one whole SM each) first and let the light channel blocks fill what is left - the other way round the frame kernel
    // would wait for the channel kernel to drain.  A caller-supplied stream (mmimo_set_stream) should be a high-priority one.
    int prio_least = 0, prio_greatest = 0;
    CK(cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest));
    CK(cudaStreamCreateWithPriority(&c->own_stream, cudaStreamNonBlocking, prio_greatest));
    c->stream = c->own_stream;
    CK(cudaStreamCreateWithPriority(&c->copy_stream, cudaStreamNonBlocking, prio_least));
    CK(cudaStreamCreateWithPriority(&c->prep_stream, cudaStreamNonBlocking, prio_least));
    for (cudaEvent_t* e : {&c->cur_ready, &c->cur_released, &c->alt.ready, &c->alt.released})
        CK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    for (auto& sl : c->in_slot) {
        CK(cudaEventCreateWithFlags(&sl.copied, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&sl.consumed, cudaEventDisableTiming));
    }
    // twiddle tables, computed in double (forward sign)
    std::vector<float2> t4(4096), t2(256);
    for (int i = 0; i < 4096; ++i) {
        const double a = -2.0 * M_PI * (double)i / 4096.0;
        t4[i] = make_float2((float)cos(a), (float)sin(a));
    }
    for (int k1 = 0; k1 < 16; ++k1)
        for (int n0 = 0; n0 < 16; ++n0) {
            const double a = -2.0 * M_PI * (double)(k1 * n0) / 256.0;
            t2[k1 * 16 + n0] = make_float2((float)cos(a), (float)sin(a));
        }
    std::vector<float2> tt(t64::TAB_FLOAT2);
    t64::build_table(tt.data());
    if (c->tw4096.reserve(4096 * 8) || c->tw256.reserve(256 * 8) || c->tw_t64.reserve(t64::TAB_FLOAT2 * 8) || c->counters.reserve(65 * 8)) return 1;
    CK(cudaMemcpy(c->tw4096.p, t4.data(), 4096 * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c->tw_t64.p, tt.data(), t64::TAB_FLOAT2 * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c->tw256.p, t2.data(), 256 * 8, cudaMemcpyHostToDevice));
    CK(cudaMemset(c->counters.p, 0, 65 * 8));
    CK(cudaFuncSetAttribute(k_link_frame<PA_NONE>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_link_frame<PA_SOFT_LIMITER>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_link_frame<PA_RAPP>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_link_frame<PA_TOI>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_link_frame_r64<PA_NONE>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME64_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_link_frame_r64<PA_SOFT_LIMITER>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME64_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_link_frame_r64<PA_RAPP>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME64_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_link_frame_r64<PA_TOI>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME64_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_link_frame_t64<PA_NONE>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_T64_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_link_frame_t64<PA_SOFT_LIMITER>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_T64_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_link_frame_t64<PA_RAPP>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_T64_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_link_frame_t64<PA_TOI>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_T64_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_cnc_loop_t64<PA_NONE>, cudaFuncAttributeMaxDynamicSharedMemorySize, CNC_LOOP_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_cnc_loop_t64<PA_SOFT_LIMITER>, cudaFuncAttributeMaxDynamicSharedMemorySize, CNC_LOOP_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_cnc_loop_t64<PA_RAPP>, cudaFuncAttributeMaxDynamicSharedMemorySize, CNC_LOOP_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_cnc_loop_t64<PA_TOI>, cudaFuncAttributeMaxDynamicSharedMemorySize, CNC_LOOP_SMEM_BYTES));
    CK(cudaFuncSetAttribute(k_link_frame_t64<PA_NONE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_T64_SMEM_BYTES_MU));
    CK(cudaFuncSetAttribute(k_link_frame_t64<PA_SOFT_LIMITER, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_T64_SMEM_BYTES_MU));
    CK(cudaFuncSetAttribute(k_link_frame_t64<PA_RAPP, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_T64_SMEM_BYTES_MU));
    CK(cudaFuncSetAttribute(k_link_frame_t64<PA_TOI, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_T64_SMEM_BYTES_MU));
    // the two-user and SDR kernels: per device, so set for every context (not once per process)
    CK(cudaFuncSetAttribute(k_link_frame<PA_NONE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_SMEM_BYTES_MU));
    CK(cudaFuncSetAttribute(k_link_frame<PA_SOFT_LIMITER, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_SMEM_BYTES_MU));
    CK(cudaFuncSetAttribute(k_link_frame<PA_RAPP, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_SMEM_BYTES_MU));
    CK(cudaFuncSetAttribute(k_link_frame<PA_TOI, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME_SMEM_BYTES_MU));
    // the background channel kernel runs beside resident frame-kernel CTAs: it must ask for the same (maximal) shared-memory
    // carve-out, or the SM would have to be reconfigured and the two kernels could not share it
    CK(cudaFuncSetAttribute(k_gen_channel_bg, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    if (stage_kernels_init()) return 1;
    return 0;
}

extern "C" void mmimo_destroy(mmimo_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->copy_stream) cudaStreamSynchronize(c->copy_stream);
    if (c->prep_stream) cudaStreamSynchronize(c->prep_stream);
    for (auto& r : c->prof_recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
    c->prof_recs.clear();
    for (auto& sl : c->in_slot) {
        sl.pos.release(); sl.idx.release(); sl.nz.release();
        if (sl.copied) cudaEventDestroy(sl.copied);
        if (sl.consumed) cudaEventDestroy(sl.consumed);
    }
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->prep_stream) { cudaStreamSynchronize(c->prep_stream); cudaStreamDestroy(c->prep_stream); }
    for (DevBuf* b : {&c->alt.h_true, &c->alt.h_loop, &c->alt.csi_nz, &c->alt.sscale, &c->alt.agc_inv, &c->alt.hkvk_inv, &c->alt.scal,
                      &c->alt.alpha, &c->alt.rx_pos})
        b->release();
    for (cudaEvent_t e : {c->cur_ready, c->cur_released, c->alt.ready, c->alt.released})
        if (e) cudaEventDestroy(e);
    DevBuf* all[] = {&c->tw4096, &c->tw256, &c->tw_t64, &c->sweep_tab, &c->sweep_cnt, &c->freq_slot, &c->freq_bin, &c->tx_pos, &c->h_true, &c->h_loop, &c->csi_nz, &c->rxeq_scr, &c->v, &c->sscale, &c->agc_inv,
                     &c->hkvk_inv, &c->scal, &c->alpha, &c->rx_pos, &c->tx_idx, &c->noise, &c->det, &c->err,
                     &c->rx_eq, &c->corr, &c->stage, &c->cnc_ones, &c->cnc_v, &c->cnc_agc_inv, &c->counters,
                     &c->s_a, &c->s_b, &c->s_c, &c->s_d};
    for (DevBuf* b : all) b->release();
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

extern "C" void* mmimo_stream(mmimo_ctx* c) { return c ? (void*)c->stream : nullptr; }
extern "C" int mmimo_set_stream(mmimo_ctx* c, void* stream) {
    if (!c) return fail("NULL context");
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    c->stream = stream ? (cudaStream_t)stream : c->own_stream;
    return 0;
}
extern "C" int mmimo_counters_to_device(mmimo_ctx* c, void* dst_dev, int clear) {
    if (!c || !dst_dev) return fail("NULL argument");
    CK(cudaSetDevice(c->device));
    CK(cudaMemcpyAsync(dst_dev, c->counters.p, 65 * 8, cudaMemcpyDeviceToDevice, c->stream));
    if (clear) CK(cudaMemsetAsync(c->counters.p, 0, 65 * 8, c->stream));
    return 0;
}
extern "C" uint64_t mmimo_launch_count(mmimo_ctx* c) { return c ? c->launches : 0; }
extern "C" int mmimo_synchronize(mmimo_ctx* c) {
    if (!c) return fail("NULL context");
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

static int log2_side(int constel) {
    switch (constel) {
        case 4: return 1;
        case 16: return 2;
        case 64: return 3;
        case 256: return 4;
    }
    return -1;
}

static double avg_symbol_power(int constel) {  // mean |c|^2 of square QAM with levels +-1, +-3, ...: 2 (M-1) / 3
    return 2.0 * (constel - 1) / 3.0;
}

extern "C" int mmimo_configure(mmimo_ctx* c, const mmimo_link_config* cfg) {
    if (!c || !cfg) return fail("mmimo_configure: NULL argument");
    if (cfg->n_fft != 4096 || cfg->n_sub_carr != 2048)
        return fail("mmimo_configure: only n_fft=4096 / n_sub_carr=2048 is built (got %d / %d)", cfg->n_fft, cfg->n_sub_carr);
    if (log2_side(cfg->constel_size) < 0)
        return fail("Constellation size must be a power of some number, only square QAM supported. (got %d)", cfg->constel_size);
    if (cfg->n_ant < 1 || cfg->n_ant > MAX_ANT) return fail("mmimo_configure: n_ant must be 1..%d (got %d)", MAX_ANT, cfg->n_ant);
    if (cfg->pa_kind < 0 || cfg->pa_kind > 3) return fail("mmimo_configure: unknown pa_kind %d", cfg->pa_kind);
    if (cfg->channel_kind < 0 || cfg->channel_kind > 3) return fail("mmimo_configure: unknown channel_kind %d", cfg->channel_kind);
    if (!cfg->tx_pos || !cfg->freqs) return fail("mmimo_configure: tx_pos and freqs are required");
    CK(cudaSetDevice(c->device));
    c->cfg = *cfg;
    c->tx_pos_h.assign(cfg->tx_pos, cfg->tx_pos + 3 * cfg->n_ant);
    c->freqs_h.assign(cfg->freqs, cfg->freqs + cfg->n_fft);
    c->cfg.tx_pos = c->tx_pos_h.data();
    c->cfg.freqs = c->freqs_h.data();
    std::vector<double> fs(N_SC);
    for (int p = 0; p < N_SC; ++p) fs[p] = c->freqs_h[slot_to_bin(p)];
    if (c->freq_slot.reserve(N_SC * 8) || c->freq_bin.reserve(4096 * 8) || c->tx_pos.reserve((size_t)cfg->n_ant * 24)) return 1;
    CK(cudaMemcpyAsync(c->freq_slot.p, fs.data(), N_SC * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->freq_bin.p, c->freqs_h.data(), 4096 * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->tx_pos.p, c->tx_pos_h.data(), (size_t)cfg->n_ant * 24, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    c->configured = true;
    return 0;
}

extern "C" int mmimo_set_ibo(mmimo_ctx* c, double ibo_db) {
    if (!c || !c->configured) return fail("mmimo_set_ibo: context not configured");
    c->cfg.ibo_db = ibo_db;
    return 0;
}

extern "C" int mmimo_profile(mmimo_ctx* c, int enable) {
    if (!c) return fail("NULL context");
    c->prof = enable != 0;
    return 0;
}

extern "C" int mmimo_profile_read(mmimo_ctx* c, double* ms4, uint64_t* n4) {
    if (!c) return fail("NULL context");
    CK(cudaStreamSynchronize(c->stream));
    for (int i = 0; i < 4; ++i) { if (ms4) ms4[i] = 0.0; if (n4) n4[i] = 0; }
    for (auto& r : c->prof_recs) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, r.a, r.b);
        if (ms4) ms4[r.kind] += ms;
        if (n4) n4[r.kind] += 1;
        cudaEventDestroy(r.a);
        cudaEventDestroy(r.b);
    }
    c->prof_recs.clear();
    return 0;
}

// FP32 FMA peak of this device, measured: 16 independent FFMA chains per thread
__global__ void __launch_bounds__(256) k_fma_peak(float* out, int iters, float a, float b) {
    float x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = (float)(threadIdx.x + i);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = fmaf(x[i], a, b);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += x[i];
    if (s == 123.456f) out[0] = s;
}

extern "C" int mmimo_measure_fp32_peak(mmimo_ctx* c, double* tflops) {
    if (!c || !tflops) return fail("NULL argument");
    CK(cudaSetDevice(c->device));
    if (c->s_d.reserve(64)) return 1;
    const int iters = 4096, grid = c->n_sm * 8;
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    double best = 0.0;
    for (int rep = 0; rep < 6; ++rep) {
        CK(cudaEventRecord(a, c->stream));
        k_fma_peak<<<grid, 256, 0, c->stream>>>(c->s_d.as<float>(), iters, 0.999f, 0.001f);
        CK(cudaEventRecord(b, c->stream));
        CK(cudaEventSynchronize(b));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, a, b));
        const double fl = 2.0 * 16 * 8 * (double)iters * 256.0 * grid;
        if (rep > 0) best = std::max(best, fl / (ms * 1e-3) * 1e-12);
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    c->launches += 6;
    *tflops = best;
    return 0;
}

#include "link_host.inl"
#include "mu_host.inl"
#include "stage_host.inl"

// ---------------------------------------------------------------------------------------------
// mmimo_welch_psd_noalloc (include/mmimo_b200.h): the Welch PSD of stage_host.inl for callers that may not allocate
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The three scratch arrays of mmimo_welch_psd (windowed segments, their spectra, the double accumulator) live in ONE
// file-scope device array of fixed size; every GPU that loads the library holds it until the process exits.
// All calls share it: calls that overlap on one GPU (different streams, different host threads) race.
// ---------------------------------------------------------------------------------------------
constexpr size_t NOALLOC_MAX_ELEMS = MMIMO_WELCH_NOALLOC_MAX_ELEMS;          // n_seg * nfft complex samples per call
constexpr size_t NOALLOC_ROWS_OFF = 0;
constexpr size_t NOALLOC_SPEC_OFF = NOALLOC_MAX_ELEMS * 8;
constexpr size_t NOALLOC_ACC_OFF = 2 * NOALLOC_MAX_ELEMS * 8;
constexpr size_t NOALLOC_BYTES = NOALLOC_ACC_OFF + 8192 * 8;                 // 8 MiB + 64 KiB
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

#include <mutex>
constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex g_noalloc_mu;
static unsigned char* g_noalloc_ptr[NOALLOC_MAX_DEVICES];  // device address of g_noalloc_scratch, per device (the name itself
                                                           // is only a host-side handle: never hand it to a kernel)

// once per device, with that device current: the array's device address and the transform's shared-memory opt-in
static int noalloc_setup(int dev, unsigned char** out) {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        unsigned char* p = nullptr;
        CK(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        if (!p) return fail("mmimo_welch_psd_noalloc: cudaGetSymbolAddress returned NULL");
        CK(cudaFuncSetAttribute(k_fft_generic_rows<-1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 8192 * 8));
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return 0;
}

extern "C" int mmimo_welch_psd_noalloc(size_t n, const float* x_dev, int nperseg, int nfft, double fs, double* psd_dev, void* stream) {
    if (!x_dev || !psd_dev) return fail("mmimo_welch_psd_noalloc: NULL buffer");
    if (nfft < 8 || nfft > 8192 || (nfft & (nfft - 1))) return fail("mmimo_welch_psd_noalloc: nfft must be a power of two in 8..8192 (got %d)", nfft);
    if (nperseg < 2 || nperseg > nfft) return fail("mmimo_welch_psd_noalloc: nperseg must be 2..nfft (got %d)", nperseg);
    if ((size_t)nperseg > n) return fail("mmimo_welch_psd_noalloc: the signal is shorter than one segment (%zu < %d)", n, nperseg);
    if (!(fs > 0.0)) return fail("mmimo_welch_psd_noalloc: fs must be positive");
    const int step = nperseg - nperseg / 2;
    const size_t n_seg = (n - (size_t)(nperseg / 2)) / (size_t)step;
    if (n_seg > NOALLOC_MAX_ELEMS / (size_t)nfft)
        return fail("mmimo_welch_psd_noalloc: %zu segments x nfft %d = %zu samples of scratch, the static array holds %zu "
                    "(MMIMO_WELCH_NOALLOC_MAX_ELEMS); use mmimo_welch_psd for longer signals",
                    n_seg, nfft, n_seg * (size_t)nfft, NOALLOC_MAX_ELEMS);
    int dev = -1;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return fail("mmimo_welch_psd_noalloc: device ordinal %d out of range", dev);
    unsigned char* base = nullptr;
    if (noalloc_setup(dev, &base)) return 1;
    cudaStream_t st = (cudaStream_t)stream;
    const float2* dx = reinterpret_cast<const float2*>(x_dev);
    float2* rows = reinterpret_cast<float2*>(base + NOALLOC_ROWS_OFF);
    float2* spec = reinterpret_cast<float2*>(base + NOALLOC_SPEC_OFF);
    double* acc = reinterpret_cast<double*>(base + NOALLOC_ACC_OFF);
    // the array keeps the previous call's contents: the accumulator must start at zero; rows / spec are fully overwritten
    CK(cudaMemsetAsync(acc, 0, (size_t)nfft * 8, st));
    k_welch_segments<<<(unsigned)n_seg, 256, 0, st>>>(dx, rows, nperseg, step, nfft);
    k_fft_generic_rows<-1><<<(unsigned)n_seg, 256, 2 * (size_t)nfft * 8, st>>>(rows, spec, nfft, ilog2(nfft), 0, 0);
    k_welch_reduce<<<(nfft + 255) / 256, 256, 0, st>>>(spec, acc, (int)n_seg, nfft);
    double w2 = 0.0;
    for (int i = 0; i < nperseg; ++i) { const double w = 0.5 - 0.5 * cos(2.0 * M_PI * i / nperseg); w2 += w * w; }
    k_scale_d<<<(nfft + 255) / 256, 256, 0, st>>>(acc, (double)nfft / (fs * w2 * (double)n_seg), nfft);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(psd_dev, acc, (size_t)nfft * 8, cudaMemcpyDeviceToDevice, st));
    return 0;
}
