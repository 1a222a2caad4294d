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
 *   - n_fft = 4096, n_sub_carr = 2048 in this version (every configuration in BASELINE.json).
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
    int32_t incl_clean_run;    /* also run the "no distortion" frames of mp_model.py:136-175 */
    int32_t reserved;
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
} mmimo_sim_result;

/* context ------------------------------------------------------------------------------------ */
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
 *                         ACCUMULATED across calls until mmimo_link_fetch reads and clears them. */
int mmimo_link_simulate(mmimo_ctx* ctx, const mmimo_sim_request* req, mmimo_sim_result* res);
int mmimo_link_enqueue(mmimo_ctx* ctx, const mmimo_sim_request* req);
int mmimo_link_fetch(mmimo_ctx* ctx, int32_t n_iter, uint64_t* err, uint64_t* clean_err);
/* stream-ordered copy of the 65 uint64 device counters ([0..63] bit errors after iteration i, [64] clean run)
 * into caller-owned DEVICE memory - the buffer handed to the NCCL all-reduce that replaces the shared
 * multiprocessing.Array accumulation of mp_model.py:172-173,221-222; clear != 0 zeroes the counters after the copy */
int mmimo_counters_to_device(mmimo_ctx* ctx, void* dst_dev, int clear);
/* number of kernels launched by this context so far (bench.py "gpu_launches") */
uint64_t mmimo_launch_count(mmimo_ctx* ctx);


/* measurement helpers (bench.py): per-kernel CUDA-event timing on the context's stream.
 * kinds: 0 channel generation, 1 realisation prep (precoder/AGC), 2 fused frame kernel, 3 other */
int mmimo_profile(mmimo_ctx* ctx, int enable);
int mmimo_profile_read(mmimo_ctx* ctx, double* ms4, uint64_t* n4);
/* measured FP32 FMA peak of the device in TFLOP/s (dependent-free FFMA chains on every SM) */
int mmimo_measure_fp32_peak(mmimo_ctx* ctx, double* tflops);

#ifdef __cplusplus
}
#endif
#endif /* MMIMO_B200_H */
