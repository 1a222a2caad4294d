/* mmimo_b200.h - C ABI of libmmimo_b200.so: B200 (sm_100a) kernels for the Monte-Carlo link hot path of
 * MarcinWachowiak/m-mimo-ofdm-with-nonlinear-pa-sim.
 *
 * The reference has no FFI: its boundary for this path is the Python class surface of
 * modulation.py / transceiver.py / distortion.py / antenna_array.py / channel.py / noise.py /
 * corrector.py / mp_model.py.  Each entry point below names the reference method(s)
 * it replaces (paths relative to the reference root); the Python drop-in modules under
 * m_mimo_ofdm_with_nonlinear_pa_sim_b200/dropin/ bind them with ctypes.
 *
 * Conventions
 *   - plain C, no exceptions: every call returns 0 on success, non-zero on error;
 *     mmimo_last_error() returns the message of the last failure in the calling thread.
 *   - one context per process and GPU; calls on one context must not be issued concurrently.
 *   - all pointers in this header are HOST pointers unless the name ends in _dev; the caller owns
 *     every buffer, the library owns only its context scratch.
 *   - complex arrays are interleaved float (re, im) pairs ("c64"); per-sub-carrier host arrays use the
 *     reference's data-sub-carrier order ("negative half then positive half", modulation.py:266-267);
 *     full-band arrays use FFT-bin order (DC at index 0).
 *   - the FUSED entry points (mmimo_link_*, mmimo_mu_link_simulate, mmimo_sdr_simulate) take n_fft = 4096 with
 *     n_sub_carr = 2048 (every configuration in BASELINE.json); the stage-level entry points further down take any
 *     n_fft = 2^m from 8 to 8192 and any even n_sub_carr.
 */
#ifndef MMIMO_B200_H
#define MMIMO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mmimo_ctx mmimo_ctx;

enum { MMIMO_PA_NONE = 0, MMIMO_PA_SOFT_LIMITER = 1, MMIMO_PA_RAPP = 2, MMIMO_PA_TOI = 3 };
enum { MMIMO_CHAN_LOS = 0, MMIMO_CHAN_TWO_PATH = 1, MMIMO_CHAN_RAYLEIGH = 2, MMIMO_CHAN_INJECTED = 3 };
enum { MMIMO_RX_CNC = 0, MMIMO_RX_MCNC = 1 };
/* request flags: injected `noise` / `clean_noise` hold the raw standard-normal pairs of noise.py:66 and the device
 * applies the per-frame 0.5 * noise_std of noise.py:64 (which depends on the realisation's AGC vector) */
enum { MMIMO_REQ_NOISE_UNIT = 1 };

/* Static description of one link: the objects main_mp_miso_mcnc_ber_vs_ebn0.py:64-104 builds. */
typedef struct mmimo_link_config {
    int32_t constel_size;      /* OfdmQamModem(constel_size, ...)  modulation.py:296 : 4, 16, 64, 256 */
    int32_t n_fft;             /* 4096 */
    int32_t n_sub_carr;        /* 2048 */
    int32_t n_ant;             /* antenna_array.py:415 LinearArray(n_elements) ; 1..256 */
    int32_t pa_kind;           /* distortion.py:22 SoftLimiter / :116 Rapp / :215 ThirdOrderNonLin */
    int32_t channel_kind;      /* channel.py:11 MisoLosFd / :92 MisoTwoPathFd / :187 MisoRayleighFd */
    int32_t receiver_kind;     /* corrector.py:12 CncReceiver / :115 McncReceiver */
    int32_t skip_attenuation;  /* calc_channel_mat(skip_attenuation) channel.py:35 */
    double ibo_db;             /* SoftLimiter/Rapp ibo_db, ThirdOrderNonLin toi_db */
    double p_hardness;         /* Rapp p_hardness */
    double center_freq;        /* Transceiver.center_freq  transceiver.py:9 */
    double carrier_spacing;    /* Transceiver.carrier_spacing */
    double rx_pos[3];          /* nominal receiver position (cord_x, cord_y, cord_z) */
    double rx_loc_var;         /* Link(rx_loc_var) mp_model.py:32 */
    double ant_gain_db;        /* tx_ant_gain_db + rx_ant_gain_db (channel.py:65) */
    const double* tx_pos;      /* [n_ant][3] element coordinates (antenna_array.py:436-445) */
    const double* freqs;       /* [n_fft] sub-carrier frequencies in FFT-bin order exactly as channel.py:51-52
                                  produces them (float32-rounded values held in double) */
} mmimo_link_config;

/* One batch of Monte-Carlo frames = n_frames passes of the loop body mp_model.py:190-222. */
typedef struct mmimo_sim_request {
    int32_t n_frames;
    int32_t n_iter;            /* max(cnc_n_iter_lst): detections after 0..n_iter iterations are counted */
    int32_t incl_clean_run;    /* 1: also run the "no distortion" frames of mp_model.py:136-175; 2: run ONLY those */
    int32_t flags;             /* MMIMO_REQ_* */
    double snr_db;             /* Link.set_snr  mp_model.py:243 */
    double csi_eps;            /* Link(csi_epsylon): > 0 draws h_loop on the device (mp_model.py:264-284); ignored
                                  when h_loop is injected */
    uint64_t seed;             /* Philox key for everything not injected */
    uint64_t frame_id0;        /* global index of the first frame (counter): results do not depend on sharding */
    /* optional injected inputs (NULL = generate on the device) */
    const double* rx_pos;      /* [n_frames][3] receiver position per frame (mp_model.py:193-198) */
    const float* h_true;       /* c64 [n_frames][n_ant][n_sub_carr] channel on the data sub-carriers */
    const float* h_loop;       /* c64, same shape: channel knowledge with CSI error (mp_model.py:264-284); NULL = h_true */
    const uint8_t* tx_idx;     /* [n_frames][n_sub_carr] transmitted symbol values (bit groups, MSB first) */
    const float* noise;        /* c64 [n_frames][n_sub_carr] additive noise on the data sub-carriers (noise.py:83) */
    /* the same for the clean run (its own frames: mp_model.py:136-175 draws them first) */
    const double* clean_rx_pos;
    const float* clean_h_true;
    const float* clean_h_loop;
    const uint8_t* clean_tx_idx;
    const float* clean_noise;
    /* CSI-error draws (mp_model.py:272): c64 [n_frames][n_ant][n_sub_carr] unit-variance pairs exactly as
     * rng.standard_normal((n_sub_carr, 2)) yields them per antenna; the device applies
     * 0.5 * sqrt(2 * mean|h_k|^2) * csi_eps.  NULL = Philox.  Ignored unless csi_eps > 0 and h_loop == NULL. */
    const float* csi_noise;
    const float* clean_csi_noise;
} mmimo_sim_request;

typedef struct mmimo_sim_result {
    uint64_t* err;             /* [n_iter+1] bit errors per iteration (required) */
    uint64_t* clean_err;       /* [1] (required when incl_clean_run) */
    uint64_t bits_per_frame;   /* written by the library */
    /* optional per-frame dumps (NULL = skip) */
    uint8_t* det;              /* [n_frames][n_iter+1][n_sub_carr] detected symbol values */
    uint32_t* err_per_frame;   /* [n_frames][n_iter+1] */
    float* rx_eq;              /* c64 [n_frames][n_sub_carr] equalised received sub-carriers (mp_model.py:214) */
    float* corr;               /* c64 [n_frames][n_iter+1][n_sub_carr] slicer input per iteration (corrector.py:184) */
    float* h_true;             /* c64 [n_frames][n_ant][n_sub_carr] channel actually used */
    float* v;                  /* c64 [n_frames][n_ant][n_sub_carr] precoder (antenna_array.py:170-171) */
    float* agc;                /* c64 [n_frames][n_sub_carr]   sum_k alpha_k h_k v_k (mp_model.py:319-321) */
    float* pa_p0;              /* [n_frames] PA saturation power / cubic coefficient of the array */
    float* alpha;              /* [n_frames][n_ant] per-antenna Bussgang alpha (mp_model.py:315-317) */
    double* rx_pos;            /* [n_frames][3] */
    uint8_t* clean_det;        /* [n_frames][n_sub_carr] */
    uint32_t* clean_err_per_frame; /* [n_frames] */
} mmimo_sim_result;

/* context ------------------------------------------------------------------------------------ */
int mmimo_device_count(void);                   /* number of CUDA devices visible to this process, < 0 on error */
int mmimo_create(mmimo_ctx** out, int device);
void mmimo_destroy(mmimo_ctx* ctx);
const char* mmimo_last_error(void);
const char* mmimo_version(void);
void* mmimo_stream(mmimo_ctx* ctx);            /* cudaStream_t all work of this context is enqueued on */
int mmimo_synchronize(mmimo_ctx* ctx);
/* run all further work of this context on the caller's cudaStream_t (NULL = back to the context's own stream);
 * lets the host side order kernels against torch.distributed collectives without extra synchronisation */
int mmimo_set_stream(mmimo_ctx* ctx, void* cuda_stream);

/* Link.__init__ / update_distortion (mp_model.py:32-87, :230-241): (re)describe the link */
int mmimo_configure(mmimo_ctx* ctx, const mmimo_link_config* cfg);
int mmimo_set_ibo(mmimo_ctx* ctx, double ibo_db);

/* Link.simulate (mp_model.py:89-222) for a fixed number of frames.
 *   mmimo_link_simulate : H2D of injected inputs + kernels + D2H of results + synchronize.
 *   mmimo_link_enqueue  : the same work without the final D2H/sync; counters stay on the device and are
 *                         ACCUMULATED across calls until mmimo_link_fetch reads and clears them.
 * Buffer lifetime: mmimo_link_enqueue returns while the copies of the request's host arrays may still be in flight
 * (rx_pos / tx_idx / noise go through the context's copy stream into one of two staging slots, so that the copies of
 * batch n + 1 overlap the kernels of batch n; pinned memory makes them asynchronous).  The caller must leave every array
 * of the request untouched until mmimo_link_fetch or mmimo_synchronize has returned, or until TWO later enqueue calls
 * carrying the same kind of array have been issued and the stream they run on has passed them. */
int mmimo_link_simulate(mmimo_ctx* ctx, const mmimo_sim_request* req, mmimo_sim_result* res);
int mmimo_link_enqueue(mmimo_ctx* ctx, const mmimo_sim_request* req);
int mmimo_link_fetch(mmimo_ctx* ctx, int32_t n_iter, uint64_t* err, uint64_t* clean_err);
/* Multi-point batch: n_points * frames_per_point frames in one pass of the link pipeline; point p owns frames
 * [p * frames_per_point, (p + 1) * frames_per_point) of the batch (global frame ids base->frame_id0 + index) and runs at
 * snr_db[p] / ibo_db[p] (NULL = the value of `base` / of the configured link for every point).  Replaces the inner
 * (IBO, Eb/N0) double loop of main_mp_clipping_noise_cancellation/main_mp_miso_mcnc_constant_ber_req_ebn0_vs_ibo.py:116-150
 * and the point loops of ..._ber_vs_ebn0.py:108-140 / ..._ber_vs_ibo.py.  base->n_frames is ignored;
 * err is [n_points][n_iter+1]. */
int mmimo_link_sweep(mmimo_ctx* ctx, const mmimo_sim_request* base, int32_t n_points, int32_t frames_per_point,
                     const double* snr_db, const double* ibo_db, uint64_t* err);
/* stream-ordered copy of the 65 uint64 device counters ([0..63] bit errors after iteration i, [64] clean run)
 * into caller-owned DEVICE memory - the buffer handed to the NCCL all-reduce that replaces the shared
 * multiprocessing.Array accumulation of mp_model.py:172-173,221-222; clear != 0 zeroes the counters after the copy */
int mmimo_counters_to_device(mmimo_ctx* ctx, void* dst_dev, int clear);
/* Signal-to-distortion snapshots (main_beampatterns_plotting/sdr_mp.py:124-151): for n_frames channel realisations of the
 * configured link (receiver position re-drawn per snapshot like mp_model.py:193-198, MRT precoding, the configured PA/IBO)
 * desired[i]    = sum_s | sum_k alpha_k h_k X_clean_k |^2   and
 * distortion[i] = sum_s | sum_k h_k (X_k - alpha_k X_clean_k) |^2 , alpha_k the per-antenna Bussgang coefficient.
 * Optional injected inputs as in mmimo_sim_request (NULL = generated on the device). */
typedef struct mmimo_sdr_request {
    int32_t n_frames;
    int32_t reserved;
    uint64_t seed;
    uint64_t frame_id0;
    const double* rx_pos;      /* [n_frames][3] */
    const float* h_true;       /* c64 [n_frames][n_ant][n_sub_carr] */
    const uint8_t* tx_idx;     /* [n_frames][n_sub_carr] */
} mmimo_sdr_request;
int mmimo_sdr_simulate(mmimo_ctx* ctx, const mmimo_sdr_request* req, double* desired, double* distortion);

/* Two-user Monte-Carlo frames, fused (BASELINE config 5): the frame loop of
 * main_multiuser/main_multiuser_mcnc_ber_vs_ebn0.py:514-642 and multiuser_mp_prototype.py - per realisation one channel
 * per user (MisoLosFd / MisoTwoPathFd / MisoRayleighFd / injected), multi-user MRT or zero-forcing precoding
 * (antenna_array.py:202-257), array PA re-normalisation (:313-360), one array transmission carrying both users
 * (modulation.py:373), per user: propagate over its channel, AWGN, AGC, then McncMuReceiver.receive with knowledge of the
 * other user's bits (corrector.py:404-446) or one of the other receivers of the multi-user scripts (MMIMO_MU_RX_*).
 * Arrays with a user axis are ordered [n_frames][2][...]; err is [2][n_iter+1].  NULL optional inputs = generated on the
 * device (Philox keyed on (seed, 2*frame + user)). */
enum { MMIMO_MU_MRT = 0, MMIMO_MU_ZF = 1 };
/* per-user receiver of the two-user loop:
 *   MCNC_MU  McncMuReceiver (corrector.py:348-489): full two-user array model, other user's bits known
 *   MCNC_SU  single-user McncReceiver on the user's own channel, MRT (main_multiuser_mcnc_ber_vs_ebn0.py:563-572 with
 *            mr_precoding; with zf_precoding the script's single-user arrays use the phase-only branch
 *            antenna_array.py:175-178 - stage path only)
 *   CNC      single-antenna CncReceiver (corrector.py:12-112; main_multiuser_cnc_ber_vs_ebn0.py:124,600)
 *   CNC_MU   CncMuReceiver (corrector.py:248-345): single-antenna model fed with sqrt(2)/2 (own + other user's symbols) */
enum { MMIMO_MU_RX_MCNC_MU = 0, MMIMO_MU_RX_MCNC_SU = 1, MMIMO_MU_RX_CNC = 2, MMIMO_MU_RX_CNC_MU = 3 };
typedef struct mmimo_mu_request {
    int32_t n_frames;          /* channel realisations; each yields one frame per user */
    int32_t n_iter;
    int32_t precoding;         /* MMIMO_MU_MRT / MMIMO_MU_ZF */
    int32_t receiver;          /* MMIMO_MU_RX_*: which receiver runs per user (below) */
    double snr_db;
    uint64_t seed;
    uint64_t frame_id0;
    double usr_pos[2][3];      /* nominal position of each user (usr_pos_tup of the script); offsets +-rx_loc_var/2 on x and y */
    const double* rx_pos;      /* [n_frames][2][3] or NULL */
    const float* h;            /* c64 [n_frames][2][n_ant][n_sub_carr] or NULL */
    const uint8_t* tx_idx;     /* [n_frames][2][n_sub_carr] or NULL */
    const float* noise;        /* c64 [n_frames][2][n_sub_carr] or NULL */
    const float* rx_eq_in;     /* c64 [n_frames][2][n_sub_carr]: equalised received signal; skips the transmission (receiver-only use) */
} mmimo_mu_request;
typedef struct mmimo_mu_result {
    uint64_t* err;             /* [2][n_iter+1] (required) */
    uint64_t bits_per_frame;
    uint8_t* det;              /* [n_frames][2][n_iter+1][n_sub_carr] or NULL */
    float* rx_eq;              /* c64 [n_frames][2][n_sub_carr] or NULL */
    float* v;                  /* c64 [n_frames][2][n_ant][n_sub_carr] precoder per user or NULL (reference layout is [K,U,N_sc]) */
    float* agc;                /* c64 [n_frames][2][n_sub_carr] or NULL */
    float* pa_p0;              /* [n_frames] or NULL */
} mmimo_mu_result;
int mmimo_mu_link_simulate(mmimo_ctx* ctx, const mmimo_mu_request* req, mmimo_mu_result* res);

/* number of kernels launched by this context so far (bench.py "gpu_launches") */
uint64_t mmimo_launch_count(mmimo_ctx* ctx);


/* ---- stage-level (unfused) operations: one call per reference building-block method ------------------
 * Used by the drop-in classes (m_mimo_ofdm_with_nonlinear_pa_sim_b200/dropin/), where a script drives the
 * blocks one by one instead of through Link.simulate.  Any n_fft = 2^m (8..8192), any even n_sub_carr < n_fft.
 * Frequency-domain arrays: FFT-bin order; per-sub-carrier arrays: data order; bits: one byte (0/1) per bit,
 * MSB of each symbol first (utilities.py:18-67). */
typedef struct mmimo_modem_desc {
    int32_t constel_size;      /* QamModem(constel_size)  modulation.py:223 */
    int32_t n_fft;             /* OfdmQamModem.n_fft      modulation.py:296 */
    int32_t n_sub_carr;
    int32_t cp_len;
    double constel_scale;      /* 1, or alpha after Modem.correct_constellation (modulation.py:167-176) */
} mmimo_modem_desc;

typedef struct mmimo_pa_desc {
    int32_t kind;              /* MMIMO_PA_* */
    int32_t reserved;
    double p0;                 /* SoftLimiter/Rapp .sat_pow (distortion.py:37,131) ; ThirdOrderNonLin .cubic_dist_coeff (:228) */
    double p1;                 /* Rapp .p_hardness */
} mmimo_pa_desc;

enum { MMIMO_TX_SKIP_DIST = 1, MMIMO_TX_TIME_DOMAIN = 2, MMIMO_TX_PER_USER = 4 };
enum { MMIMO_PREC_PHASE = 0, MMIMO_PREC_MRT = 1, MMIMO_PREC_ZF = 2 };

/* utilities.to_freq_domain / to_time_domain (utilities.py:311-339), modulation._tx/_rx_ofdm_symbol transforms:
 * `rows` orthonormal transforms of length n; `skip` leading samples of every input row are dropped (cyclic prefix
 * removal), `cp` trailing output samples are repeated in front of every output row (cyclic prefix insertion). */
int mmimo_fft(mmimo_ctx* ctx, int inverse, int n, int rows, int skip, int cp, const float* in, float* out);
/* scipy.signal.welch as the reference's scripts call it - welch(x, fs=fs, nfft=nfft, nperseg=nperseg, return_onesided=False)
 * with scipy's defaults (periodic Hann window, 50 % overlap, constant detrend, density scaling, mean over segments):
 * main_multiuser/main_multiuser_mrt_precoding_radiation_pattern.py:371-381 (and the PSD blocks of the other beampattern / PSD
 * scripts).  x: n complex samples; psd [nfft] doubles in FFT-bin order, freqs [nfft] = fftfreq(nfft, 1 / fs) (may be NULL).
 * nfft a power of two in 8..8192, nperseg <= nfft. */
int mmimo_welch_psd(mmimo_ctx* ctx, size_t n, const float* x, int nperseg, int nfft, double fs, double* freqs, double* psd);
/* NO COUNTERPART IN THE REFERENCE - added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same computation as mmimo_welch_psd
 * (same kernels except that every nfft goes through the table-free transform), but the entry point allocates nothing and
 * needs no context: no cudaMalloc / cudaFree / allocating library routine is called.  Its scratch (windowed segments, their
 * spectra, the fp64 accumulator) is one file-scope `static __device__` array of 2 * MMIMO_WELCH_NOALLOC_MAX_ELEMS * 8 + 65536
 * bytes (8.06 MiB) that every GPU which loads the library holds until the process exits.
 *   x_dev   : n complex64 samples, DEVICE memory of the CURRENT device (cudaSetDevice is the caller's business);
 *   psd_dev : [nfft] doubles, DEVICE memory, FFT-bin order (frequencies: fftfreq(nfft, 1 / fs), not returned);
 *   stream  : cudaStream_t the work is enqueued on (NULL = default stream); the call does NOT synchronise.
 * Limit: n_seg * nfft <= MMIMO_WELCH_NOALLOC_MAX_ELEMS with n_seg = (n - nperseg / 2) / (nperseg - nperseg / 2); beyond it the
 * call fails (returns 1, mmimo_last_error() names the limit) and nothing is launched.  nfft a power of two in 8..8192,
 * nperseg <= nfft.  All calls share the one array, which keeps the last call's contents (the accumulator is cleared on the
 * stream at the start of every call): calls that OVERLAP on one GPU - two streams, two host threads - race; serialise them.
 * The first call on a device looks the array's address up (cudaGetSymbolAddress) and opts the transform into its shared
 * memory size; make that first call outside stream capture. */
#define MMIMO_WELCH_NOALLOC_MAX_ELEMS ((size_t)1 << 19)
int mmimo_welch_psd_noalloc(size_t n, const float* x_dev, int nperseg, int nfft, double fs, double* psd_dev, void* stream);
/* SoftLimiter/Rapp/ThirdOrderNonLin.process (distortion.py:91-98, :191-198, :281) on n complex samples */
int mmimo_pa_process(mmimo_ctx* ctx, const mmimo_pa_desc* pa, size_t n, const float* in, float* out);
/* modulation.modulate (modulation.py:13-25): bits -> n_sym constellation points */
int mmimo_qam_map(mmimo_ctx* ctx, const mmimo_modem_desc* m, size_t n_sym, const uint8_t* bits, float* symbols);
/* Modem.symbol_detection (:138-146) / demodulate (:63-77): nearest points (symbols, may be NULL) and their bits (may be NULL) */
int mmimo_qam_detect(mmimo_ctx* ctx, const mmimo_modem_desc* m, size_t n_sym, const float* in, float* symbols, uint8_t* bits);
/* modulation.soft_decoding / OfdmQamModem.soft_detection_llr (modulation.py:29-59, :407-417): exact bit LLRs
 * log(sum_{bit=1} e^{-|y-c|^2/nv} / sum_{bit=0} e^{...}), MSB of each symbol first, +inf where the denominator underflows;
 * noise_var holds 1 value (all symbols) or n_sym values; llr is [n_sym * log2(constel)] doubles */
int mmimo_qam_soft_llr(mmimo_ctx* ctx, const mmimo_modem_desc* m, size_t n_sym, const float* in, const double* noise_var,
                       size_t n_noise_var, double* llr);
/* AntennaArray.transmit (antenna_array.py:58-140) = Transceiver.transmit (transceiver.py:98-174) for n_ant antennas that
 * share one modem description and one PA (pa == NULL: no impairment).
 *   bits     [n_users][n_sub_carr * log2(constel)]
 *   precoder c64 [n_ant][n_users][n_sub_carr] (AntennaArray.get_precoding_mat layout) or NULL
 *   out      c64 [G][n_ant][n_out], G = n_users with MMIMO_TX_PER_USER else 1, n_out = n_fft or n_fft + cp_len (TIME_DOMAIN)
 *   out_clean same shape: the undistorted signal (return_both), or NULL */
int mmimo_array_transmit(mmimo_ctx* ctx, const mmimo_modem_desc* m, const mmimo_pa_desc* pa, int n_ant, int n_users,
                         const uint8_t* bits, const float* precoder, int flags, float* out, float* out_clean);
/* OfdmQamModem.demodulate (modulation.py:384-396): time-domain symbol with cyclic prefix -> data sub-carrier symbols
 * (may be NULL) and hard-decision bits (may be NULL) */
int mmimo_ofdm_demodulate(mmimo_ctx* ctx, const mmimo_modem_desc* m, const float* td, float* symbols, uint8_t* bits);
/* Miso*Fd.propagate (channel.py:74-89): x, h c64 [n_ant][n]; out [n] (sum_signals) or [n_ant][n] */
int mmimo_propagate(mmimo_ctx* ctx, int n_ant, int n, const float* x, const float* h, int sum_signals, float* out);
/* Awgn.process (noise.py:45-83): out = in + scale * unit_noise; the draw itself stays with the caller's numpy Generator */
int mmimo_awgn_add(mmimo_ctx* ctx, size_t n, const float* in, const float* unit_noise, double scale, float* out);
/* utilities.count_mismatched_bits (utilities.py:94-104) */
int mmimo_count_mismatched_bits(mmimo_ctx* ctx, size_t n, const uint8_t* a, const uint8_t* b, uint64_t* out);
/* MisoLosFd / MisoTwoPathFd.calc_channel_mat (channel.py:35-72, :116-167): kind = MMIMO_CHAN_LOS | MMIMO_CHAN_TWO_PATH;
 * h_out c64 [n_ant][n_fft] */
int mmimo_channel_geometric(mmimo_ctx* ctx, int kind, int n_ant, int n_fft, const double* tx_pos, const double* tx_gain_db,
                            double rx_gain_db, const double* rx_pos, const double* freqs, int skip_attenuation, float* h_out);
/* MisoRayleighFd (channel.py:225-250, :262-275): free-space attenuation matrix times the caller's IID coefficients
 * (coeffs == NULL: the attenuation matrix itself) */
int mmimo_channel_fspl(mmimo_ctx* ctx, int n_ant, int n_fft, const double* tx_pos, const double* tx_gain_db, double rx_gain_db,
                       const double* rx_pos, const double* freqs, const float* coeffs, float* h_out);
/* MisoRandomPathsFd (channel.py:330-344, :358-384) from the drawn departure angles and delays */
int mmimo_channel_random_paths(mmimo_ctx* ctx, int n_ant, int n_fft, const double* tx_pos, const double* freqs, int n_paths,
                               const double* angles, const double* taus, float* h_out);
/* AntennaArray.set_precoding_matrix (antenna_array.py:142-311): h c64 [n_users][n_ant][n_fft] -> v c64 [n_ant][n_users][n_sub_carr]
 * (sep_carriers: [n_ant][n_sub_carr], antenna_array.py:275-305) */
int mmimo_precoding(mmimo_ctx* ctx, int mode, int sep_carriers, int n_users, int n_ant, int n_fft, int n_sub_carr, const float* h,
                    float* v);
/* McncReceiver.__init__/update_agc (corrector.py:126-163, :209-245; MU :348-403, :453-489) and Link.recalculate_agc
 * (mp_model.py:290-329).  alpha_fixed != NULL: every antenna uses *alpha_fixed (third-order PA), else calc_alpha of the
 * per-antenna IBO.  Outputs (any may be NULL): agc_nfft / hkvk_nfft c64 [n_fft] (ones off the data sub-carriers),
 * scalers[2] = mean|sum_k alpha_k h_k v_k|^2, mean|sum_k h_k v_k|^2, vk_pow[n_ant], alpha[n_ant] */
int mmimo_agc(mmimo_ctx* ctx, int n_ant, int n_users, int usr_idx, int n_fft, int n_sub_carr, const float* h, const float* v,
              double ibo_db, const double* alpha_fixed, float* agc_nfft, float* hkvk_nfft, double* scalers, double* vk_pow,
              float* alpha);
/* CncReceiver.receive (corrector.py:52-112) / CncMuReceiver.receive (:288-345, other_usr_symbols != NULL).
 * in_sig_fd c64 [n_fft]; det_bits [n_iter+1][n_sub_carr*log2(constel)] and corr c64 [n_iter+1][n_sub_carr] (either may be NULL) */
int mmimo_cnc_receive(mmimo_ctx* ctx, const mmimo_modem_desc* m, const mmimo_pa_desc* pa, double alpha, int n_iter,
                      const float* in_sig_fd, const float* other_usr_symbols, uint8_t* det_bits, float* corr);
/* McncReceiver.receive (corrector.py:165-207) / McncMuReceiver.receive (:405-451, other_usr_bits [n_users-1][bits] in user order).
 * h c64 [n_ant][n_fft] (the receiver's channel knowledge), v [n_ant][n_users][n_sub_carr], agc_nfft c64 [n_fft] */
int mmimo_mcnc_receive(mmimo_ctx* ctx, const mmimo_modem_desc* m, const mmimo_pa_desc* pa, int n_ant, int n_users, int usr_idx,
                       const float* h, const float* v, const float* agc_nfft, int n_iter, const float* in_sig_fd,
                       const uint8_t* other_usr_bits, uint8_t* det_bits, float* corr);

/* measurement helpers (bench.py): per-kernel CUDA-event timing on the context's stream.
 * kinds: 0 channel generation, 1 realisation prep (precoder/AGC), 2 fused frame kernel, 3 other */
int mmimo_profile(mmimo_ctx* ctx, int enable);
int mmimo_profile_read(mmimo_ctx* ctx, double* ms4, uint64_t* n4);
/* measured FP32 FMA peak of the device in TFLOP/s (dependent-free FFMA chains on every SM) */
int mmimo_measure_fp32_peak(mmimo_ctx* ctx, double* tflops);
/* the unfused building-block kernels of Transceiver.transmit + propagate (transceiver.py:98-174, channel.py:84-89), timed one
 * by one on `rows` device-resident rows of 4096 samples, `reps` launches each: ms_out[5] = ms per launch of
 * map+precode, IFFT, soft limiter, FFT, propagate (elementwise form).  rows <= 65535 */
int mmimo_measure_stage_kernels(mmimo_ctx* ctx, int rows, int reps, double* ms_out);

#ifdef __cplusplus
}
#endif
#endif /* MMIMO_B200_H */
