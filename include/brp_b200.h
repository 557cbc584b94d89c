/*
 * brp_b200.h -- C ABI of libbrp_b200.so: the Einstein@Home BRP per-template search loop
 * (resample -> R2C FFT + power spectrum -> harmonic sum -> toplist) on NVIDIA B200 (sm_100a).
 *
 * Two layers are exported, all `extern "C"`, plain pointers and sizes only:
 *
 *  1. The reference's own backend interface ("legacy" layer): the eleven functions every backend of
 *     VolunteerComputingHelp/boinc-app-eah-brp implements and demod_binary.c calls
 *     (cuda/app/demod_binary_cuda.h:42-52, cuda/app/demod_binary_hs_cuda.cuh:37-39; CPU flavour
 *     demod_binary_resamp_cpu.h:37-39, demod_binary_fft_fftw.h:38-40, demod_binary_hs_cpu.h:37-61).
 *     Same names, argument meaning, call order and error codes, one template per call, so the
 *     reference driver links against this library unchanged.
 *
 *  2. An additive batched layer (brp_search_*): many templates per launch, results kept on the
 *     device in a per-bin toplist table until brp_search_collect(); this is what is benchmarked.
 *
 * Every entry point returns 0 on success or one of the reference's RADPUL_* codes
 * (demod_binary.h:25-45), which the reference driver returns unchanged to the BOINC wrapper.
 * The library has no CPU fallback: without a CUDA device every compute entry point fails with
 * RADPUL_CUDA_DEVICE_FIND / RADPUL_CUDA_DEVICE_SET.
 */
#ifndef BRP_B200_H
#define BRP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- error codes: demod_binary.h:25-45 ---- */
#define RADPUL_EMEM 1
#define RADPUL_EFILE 2
#define RADPUL_EIO 3
#define RADPUL_EVAL 4
#define RADPUL_EMISC 5
#define RADPUL_CUDA_DEVICE_FIND 1001
#define RADPUL_CUDA_DEVICE_SET 1002
#define RADPUL_CUDA_DEVICE_PROP 1003
#define RADPUL_CUDA_DEVICE_VERSION 1004
#define RADPUL_CUDA_MEM_ALLOC_HOST 1005
#define RADPUL_CUDA_MEM_ALLOC_DEVICE 1006
#define RADPUL_CUDA_MEM_COPY_HOST_DEVICE 1007
#define RADPUL_CUDA_MEM_COPY_DEVICE_HOST 1008
#define RADPUL_CUDA_MEM_FREE_HOST 1009
#define RADPUL_CUDA_MEM_FREE_DEVICE 1010
#define RADPUL_CUDA_FFT_PLAN 1011
#define RADPUL_CUDA_FFT_EXEC 1012
#define RADPUL_CUDA_FFT_DESTROY 1013
#define RADPUL_CUDA_KERNEL_INVOKE 1015

/* ---- shared types ---- */

/* diptr.h:38-41 with USE_CUDA: device_ptr_t is CUdeviceptr (cuda/app/deviceptr.h:11), 8 bytes */
#ifndef BRP_B200_NO_LEGACY_TYPES
typedef unsigned long long brp_device_ptr_t;
typedef union ptr_union_float {
    float *host_ptr;
    brp_device_ptr_t device_ptr;
} DIfloatPtr;

/* structs.h:151-161 */
typedef struct {
    unsigned int nsamples;          /* padded length N = FFT length */
    unsigned int nsamples_unpadded; /* Nu */
    unsigned int fft_size;          /* N/2 + 1 */
    float tau;
    float Omega;
    float Psi0;
    float dt;
    float step_inv;
    float S0;
} RESAMP_PARAMS;
#endif

/* structs.h:121-132 (packed, 48 bytes): one toplist / checkpoint candidate */
typedef struct __attribute__((__packed__)) {
    double power;
    double P_b;
    double tau;
    double Psi;
    double fA;
    uint32_t n_harm;
    uint32_t f0;
} brp_cand;

#define BRP_N_CAND_5 100 /* demod_binary.c:83 */
#define BRP_N_CAND 500   /* demod_binary.c:84 */
#define BRP_LOG_PS_PAGE_SIZE 10 /* hs_common.h:36 */

/* =====================================================================================
 * 1. Legacy backend interface (one template per call)
 * ===================================================================================== */

/* cuda/app/demod_binary_cuda.h:42 ; called once, demod_binary.c:457.  cudDeviceIdGiven != 0 selects
 * *cudDeviceId, otherwise device 0 of the visible set is used and written back. */
int initialize_cuda(int cudDeviceIdGiven, int *cudDeviceId);

/* cuda/app/demod_binary_cuda.h:44-46 / demod_binary_resamp_cpu.h:37-39.
 * `input` is a HOST pointer to params->nsamples_unpadded floats (copied to the device once);
 * *output receives an opaque device handle (params->nsamples floats). The LUT arguments are
 * accepted for signature compatibility; the tables of erp_utilities.cpp:45-46 live in the library. */
int set_up_resampling(DIfloatPtr input, DIfloatPtr *output, const RESAMP_PARAMS *params, float *sinLUTsamples,
                      float *cosLUTsamples);
int run_resampling(DIfloatPtr input, DIfloatPtr output, const RESAMP_PARAMS *params);
int tear_down_resampling(DIfloatPtr output);

/* cuda/app/demod_binary_cuda.h:48-50 / demod_binary_fft_fftw.h:38-40.  nsamples must be 2^a or 3*2^a
 * (RADPUL_CUDA_FFT_PLAN otherwise).  *output: device handle of the power spectrum (fft_size floats). */
int set_up_fft(DIfloatPtr input, DIfloatPtr *output, uint32_t nsamples, unsigned int fft_size);
int run_fft(DIfloatPtr input, DIfloatPtr output, uint32_t nsamples, unsigned int fft_size, float norm_factor);
int tear_down_fft(DIfloatPtr output);

/* cuda/app/demod_binary_hs_cuda.cuh:37-39 / demod_binary_hs_cpu.h:37-61.  sumspec[1..4] and dirty[0..4]
 * are HOST arrays allocated here; after run_harmonic_summing sumspec[0] points at a host array of the
 * spectrum's length up to fundamental_idx_hi, dirty[k][page] is set for every page holding a value
 * > thresholds[k], and sumspec[k][j] (k = 0..4) holds the per-bin maximum for every such value; all other
 * entries are 0 and are never selected by the candidate scan (demod_binary.c:1350 tests "> threshold").
 * run_resampling and run_fft only enqueue work; run_harmonic_summing synchronises and reports their errors. */
int set_up_harmonic_summing(float **sumspec, int32_t **dirty, unsigned int *nr_pages_ptr, unsigned int fundamental_idx_hi,
                            unsigned int harmonic_idx_hi);
int run_harmonic_summing(float **sumspec, int32_t **dirty, unsigned int nr_pages, DIfloatPtr powerspectrum,
                         unsigned int window_2, unsigned int fundamental_idx_hi, unsigned int harmonic_idx_hi,
                         float *thresholds);
int tear_down_harmonic_summing(float **sumspec, int32_t **dirty);

/* cuda/app/demod_binary_cuda.h:52 ; demod_binary.c:1469 */
int shutdown_cuda(void);

/* =====================================================================================
 * 2. Host-side restatement of the driver arithmetic around the loop (no GPU needed)
 * ===================================================================================== */

typedef struct {
    uint32_t nsamples_unpadded;  /* Nu */
    uint32_t nsamples;           /* N  = (int)(padding*Nu + 0.5)            demod_binary.c:782  */
    uint32_t fft_size;           /* F  = (unsigned)(N*0.5 + 0.5) + 1        demod_binary.c:1092 */
    uint32_t window_2;           /* w2 = (unsigned)(window*0.5 + 0.5)       demod_binary.c:1095 */
    uint32_t fundamental_idx_hi; /* Fu                                      demod_binary.c:1097 */
    uint32_t harmonic_idx_hi;    /* H                                       demod_binary.c:1099 */
    float t_obs, dt, step_inv;   /*                                         demod_binary.c:1087-1089 */
} brp_geometry;

typedef struct {
    float P, tau, Psi0, Omega, S0; /* demod_binary.c:1208-1230 */
} brp_template;

int brp_derive_geometry(uint32_t nsamples_unpadded, double tsample_us, float padding, float f0, uint32_t window,
                        brp_geometry *g);
int brp_derive_template(double P, double tau, double Psi0, float step_inv, brp_template *t);
/* thr[k] = 0.5 * chi2inv_Q(prob, 2*2^k), prob = 1-(1-fA)^(1/F): demod_binary.c:1268-1282 without the
 * toplist feedback term (using the base threshold is exact). */
int brp_base_thresholds(float fA, uint32_t fft_size, float thr[5]);
double brp_chisq_Q(double x, int n_harm);
double brp_chisq_Qinv(double Q, int n_harm);

/* template bank text file, one "%lg %lg %lg\n" line per template (demod_binary.c:197,507-535,1180-1205).
 * brp_bank_read allocates *bank (3 doubles per template); free with brp_bank_free. */
int brp_bank_read(const char *path, double **bank, uint32_t *n_templates);
void brp_bank_free(double *bank);

/* work-unit reader: .bin4 (4-bit, high nibble first) or .binary (int8), header structs.h:74-105,
 * unpacking demod_binary.c:831-842.  Allocates *series (free with brp_series_free). */
int brp_wu_read(const char *path, float **series, uint32_t *nsamples, double *tsample_us);
void brp_series_free(float *series);

/* one above-threshold per-bin maximum, as collected from the device table */
typedef struct {
    float power;       /* max over all searched templates of the summed power at this bin */
    uint32_t bin;      /* fundamental frequency bin j (w2 <= j < Fu) */
    uint32_t tmpl;     /* global index of the earliest template attaining it */
    uint32_t harm_idx; /* k: n_harm = 2^k */
} brp_hit;

/* Build the reference's candidates_all[500] (state right after the template loop, demod_binary.c:1310-1397)
 * from per-bin maxima: per harmonic, the 100 strongest bins, sorted by descending power.
 * bank (3 doubles per template) supplies P_b/tau/Psi0; tmpl_idx_out (500 entries, may be NULL) receives
 * the template index of each slot (-1 = empty).  Hits from several GPUs may simply be concatenated. */
int brp_toplist_from_hits(const brp_hit *hits, uint32_t n_hits, const double *bank, uint32_t n_templates,
                          brp_cand cands_out[BRP_N_CAND], int32_t *tmpl_idx_out);
/* merge two candidate arrays that are in the state right after the loop (a = earlier part of the bank, e.g. a checkpoint) */
int brp_toplist_merge(const brp_cand a[BRP_N_CAND], const brp_cand b[BRP_N_CAND], brp_cand out[BRP_N_CAND]);
/* demod_binary.c:1502-1555: false-alarm statistic, power in sigma, sort by fA */
int brp_finalize_candidates(brp_cand cands[BRP_N_CAND], int32_t *tmpl_idx_io);
/* demod_binary.c:1632-1667: candidate lines + %DONE% into buf (returns bytes written, <0 on overflow).
 * Destroys cands like the reference does. */
int brp_format_candidates(brp_cand cands[BRP_N_CAND], const brp_geometry *g, double tsample_us, char *buf, int buflen);
/* same, written to `path` through a temporary file + rename (demod_binary.c:1558-1686, without the BOINC header) */
int brp_write_candidates(const char *path, brp_cand cands[BRP_N_CAND], const brp_geometry *g, double tsample_us);
/* checkpoint file: CP_Header + 500 CP_cand, structs.h:111-132, demod_binary.c:1743-1783 / :560-652 */
int brp_checkpoint_write(const char *path, uint32_t n_template, const char *originalfile, const brp_cand cands[BRP_N_CAND]);
int brp_checkpoint_read(const char *path, uint32_t *n_template, char originalfile[256], brp_cand cands[BRP_N_CAND]);

/* =====================================================================================
 * 3. Batched search (additive; replaces the body of the loop demod_binary.c:1180-1443)
 * ===================================================================================== */

typedef struct brp_search brp_search;

/* Allocate everything one work-unit needs on `device`: series, per-batch scratch for max_batch
 * templates in flight, FFT tables, the per-bin toplist table (5 x Fu x 8 bytes, zeroed). */
int brp_search_create(brp_search **out, int device, const brp_geometry *g, float fA, uint32_t max_batch);
int brp_search_destroy(brp_search *s);
/* dedispersed (whitened) time series: Nu floats from HOST memory (one H2D copy, cuda/app/demod_binary_cuda.cu:333).
 * The copy is enqueued, not awaited: host_series must stay valid and unchanged until brp_search_sync() or
 * brp_search_collect() has returned (with pageable memory the runtime has staged it by the time this call returns;
 * with pinned memory it has not). */
int brp_search_set_series(brp_search *s, const float *host_series);
/* same from DEVICE memory already resident on `device` (device-to-device copy) */
int brp_search_set_series_device(brp_search *s, const void *device_series);
/* zero the per-bin toplist table (start of a new work-unit or shard) */
int brp_search_reset(brp_search *s);
/* Enqueue the search of templates bank[0..n) (3 doubles each); their global indices are
 * first_global_index + i (used for "earliest template" tie-breaking across shards).
 * Asynchronous with respect to the host; brp_search_sync() or brp_search_collect() waits. */
int brp_search_run(brp_search *s, const double *bank, uint32_t n_templates, uint32_t first_global_index);
int brp_search_sync(brp_search *s);
/* Compact all non-empty entries of the device table into hits[0..*n_hits) (D2H).
 * Returns RADPUL_EMEM (and the needed count in *n_hits) if max_hits is too small. */
int brp_search_collect(brp_search *s, brp_hit *hits, uint32_t max_hits, uint32_t *n_hits);
/* device time of the kernels enqueued by the last brp_search_run, measured with CUDA events on the
 * library's stream (valid after brp_search_sync); per-stage times need profile != 0 at create time */
int brp_search_last_run_ms(brp_search *s, float *total_ms);
/* stream the library enqueues on (a cudaStream_t), for callers that time with their own events */
void *brp_search_stream(brp_search *s);
/* number of kernels launched by this handle so far */
uint64_t brp_search_kernel_launches(const brp_search *s);
/* per-stage device time of ONE batch of `batch` templates (templates taken from bank[0..batch)),
 * each stage bracketed by CUDA events: ms[0]=resample ms[1]=fft pass 1 ms[2]=fft pass 2+power ms[3]=harmonic sum */
int brp_search_profile_batch(brp_search *s, const double *bank, uint32_t batch, uint32_t repeats, float ms[4]);

/* ---- stage taps for parity tests (slot 0 of the batch scratch; results copied to HOST) ---- */
/* resample one template; any output pointer may be NULL.  resampled_out: N floats (mean-padded),
 * index_out: Nu int32 (entries >= n_steps untouched) */
int brp_stage_resample(brp_search *s, const brp_template *t, float *resampled_out, int32_t *index_out,
                       uint32_t *n_steps_out, float *mean_out);
/* For every template of bank[0..n): resample with the production kernel (timing model evaluated at every fourth sample,
 * the samples in between proven from the two ends) and with the timing model evaluated at every sample, and count the
 * positions where the two index arrays (or n_steps) differ.  Must be 0; runs wholly on the device. */
int brp_stage_resample_compare(brp_search *s, const double *bank, uint32_t n_templates, uint64_t *n_mismatch);
/* spectrum of the series resampled last by brp_stage_resample: ps_out[0..count), count <= F */
int brp_stage_spectrum(brp_search *s, float *ps_out, uint32_t count);
/* spectrum of an arbitrary HOST series of N floats (n_steps = N, no padding): FFT kernels alone */
int brp_stage_spectrum_of(brp_search *s, const float *host_series_n, float *ps_out, uint32_t count);
/* harmonic summing of a HOST spectrum (>= H floats) against thresholds thr[5]; returns the per-bin maxima
 * above threshold as hits (tmpl = tmpl_index) without touching the search's own table */
int brp_stage_harmonic_summing(brp_search *s, const float *host_ps, const float thr[5], uint32_t tmpl_index,
                               brp_hit *hits, uint32_t max_hits, uint32_t *n_hits);
/* device evaluation of the timing-model pieces, for bit-exactness tests:
 * sin_out[i] = LUT sine of x[i] (erp_utilities.cpp:176-209); idx_out[i] = (int)((double)f[i] + 0.5) */
int brp_stage_sin_lut(int device, const float *x, float *sin_out, uint32_t n);
int brp_stage_round_index(int device, const float *f, int32_t *idx_out, uint32_t n);

/* Whitening pre-step of a work-unit (demod_binary.c:858-1079, once per work-unit, before the template loop): FFT of the
 * zero-padded series, periodogram, exact running median over `window` bins (rngmed.c:48), rescaling to a white spectrum,
 * replacement of the zap-list bins by Gaussian noise (:982-1033), zeroing of window/2 edge bins, back-transform,
 * 1/sqrt(N).  `series` (HOST, nsamples_unpadded floats) is whitened in place.  zap_ranges: n_zap_ranges (fmin, fmax)
 * pairs in Hz in file order (brp_zaplist_read), may be NULL when n_zap_ranges == 0; tsample_us (header.tsample) is only
 * used to turn them into bin numbers (:996,1012-1013).  The noise is drawn on the host in the reference's order from
 * restatements of GSL's taus2 and ziggurat Gaussian (GSL is not part of the reference tree: parity with libgsl
 * unpinned).  device_ms (optional) receives the device time of the whole step. */
int brp_whiten_series_zap(int device, float *series, uint32_t nsamples_unpadded, float padding, uint32_t window,
                          double tsample_us, const double *zap_ranges, uint32_t n_zap_ranges, float *device_ms);
/* the same without a zap list (n_zap_ranges must be 0, RADPUL_EVAL otherwise) */
int brp_whiten_series(int device, float *series, uint32_t nsamples_unpadded, float padding, uint32_t window,
                      uint32_t n_zap_ranges, float *device_ms);
/* zap list file, one "%lg %lg" pair (fmin fmax, Hz) per line (demod_binary.c:982-1009): RADPUL_EFILE if it cannot be
 * opened, RADPUL_EIO on a line that does not parse.  Allocates *ranges (2 doubles per range; brp_zaplist_free). */
int brp_zaplist_read(const char *path, double **ranges, uint32_t *n_ranges);
void brp_zaplist_free(double *ranges);
/* the noise draws of the zap step alone (no GPU), in draw order: bins_out[i] and out[2i], out[2i+1] = bin, re, im of
 * draw i for i < max_bins (either may be NULL); *n_bins_out = total number of bins drawn.  seed_bits = bit pattern of
 * the first sample of the un-whitened series (:917), nsamples = PADDED length. */
int brp_zap_noise(uint32_t seed_bits, float padding, uint32_t nsamples, double tsample_us, const double *zap_ranges,
                  uint32_t n_zap_ranges, uint32_t *bins_out, float *out, uint32_t max_bins, uint32_t *n_bins_out);

/* =====================================================================================
 * 4. Resampling without any device allocation
 * ===================================================================================== */

/* No counterpart in the reference (every reference backend allocates its own scratch at set-up time).
 *
 * Resamples one template like run_resampling / brp_stage_resample, for callers that cannot allocate device
 * memory when they need it: a task admitted with a fixed device-memory budget (a failed mid-run allocation
 * costs a BOINC task a 15-minute back-off), a host that has given the whole card to its own arena, or a loop
 * that cannot take the device-wide synchronisation of cudaMalloc/cudaFree.  This call performs no cudaMalloc,
 * cudaFree or other allocating call.  The caller passes every buffer, all in DEVICE memory of the current
 * device: d_series (nsamples_unpadded floats, in), d_out (nsamples floats, 16-byte aligned, out: resampled
 * series with the mean padding), d_n_steps and d_mean (one element each, out).  `t` is a host struct
 * (brp_derive_template), `stream` a cudaStream_t (NULL = default stream); the work is enqueued, not awaited.
 *
 * The kernel's scratch is a fixed array of BRP_NOALLOC_SCRATCH_BYTES inside the library image, held by every
 * GPU that loads the library until the process exits; series longer than BRP_NOALLOC_MAX_SAMPLES are
 * rejected with RADPUL_EVAL and a message.  The first call on a GPU looks the array's address up once; issue
 * that first call before beginning any stream capture (the address dies with cudaDeviceReset: do not call again after one).  All calls on a GPU share the array: calls that may
 * overlap on the device (different streams or threads) race and must be ordered by the caller. */
#define BRP_NOALLOC_MAX_SAMPLES (1u << 24)
#define BRP_NOALLOC_SCRATCH_BYTES (8u * (BRP_NOALLOC_MAX_SAMPLES / 4096u) + 256u)
int brp_resample_noalloc(const void *d_series, uint32_t nsamples_unpadded, uint32_t nsamples, const brp_template *t,
                         float dt, float step_inv, void *d_out, uint32_t *d_n_steps, float *d_mean, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* BRP_B200_H */
