/* TEST INFRASTRUCTURE ONLY (oracle/): CPU restatement of the Einstein@Home BRP per-template search
 * loop, used solely as the checker for the CUDA path (tests/, __graft_entry__.smoke(), bench.py's
 * cpu_baseline leg).  The product library (boinc-app-eah-brp_b200/) never includes, links or loads
 * anything in this directory.
 *
 * Pinning: tests/test_oracle_vs_reference.py checks every function below against the reference's own
 * translation units compiled in place into oracle/_ref/ (resampling, harmonic summing and the driver
 * loop bit for bit; the spectrum against the shimmed reference run_fft and scipy's float64 rfft), and
 * tests/golden/ holds vectors produced by that reference build so the check also runs where
 * /root/reference is absent.  FFTW and GSL are not in the reference tree: the FFT arithmetic and the
 * chi-square quantiles are therefore "parity unpinned" third-party pieces, restated from their
 * published definitions (see fft_double.h, shim_gsl.c).
 */
#ifndef BRP_ORACLE_H
#define BRP_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define BRP_ORACLE_N_CAND_5 100
#define BRP_ORACLE_N_CAND 500
#define BRP_ORACLE_LOG_PAGE 10

/* layout of the reference CP_cand (structs.h:121-132), packed, 48 bytes */
typedef struct __attribute__((__packed__)) {
    double power, P_b, tau, Psi, fA;
    uint32_t n_harm, f0;
} brp_oracle_cand;

typedef struct {
    uint32_t nsamples_unpadded; /* Nu */
    uint32_t nsamples;          /* N = (int)(padding*Nu + 0.5), demod_binary.c:782 */
    uint32_t fft_size;          /* F, demod_binary.c:1092 */
    uint32_t window_2;          /* demod_binary.c:1095 */
    uint32_t fundamental_idx_hi;/* Fu, demod_binary.c:1097 */
    uint32_t harmonic_idx_hi;   /* H, demod_binary.c:1099 */
    float t_obs, dt, step_inv;  /* demod_binary.c:1087-1089 */
} brp_oracle_geometry;

typedef struct {
    float P, tau, Psi0, Omega, S0; /* demod_binary.c:1208-1230 */
} brp_oracle_template;

void brp_oracle_derive_geometry(uint32_t nsamples_unpadded, double tsample_us, float padding, float f0,
                                uint32_t window, brp_oracle_geometry *g);
void brp_oracle_derive_template(double P, double tau, double Psi0, float step_inv, brp_oracle_template *t);

float brp_oracle_sin_lut(float x);

/* demod_binary_resamp_cpu.c:80-136.  idx_out (n >= Nu entries) and del_t_out may be NULL. */
int brp_oracle_resample(const float *series, const brp_oracle_geometry *g, const brp_oracle_template *t, float *out,
                        int32_t *idx_out, uint32_t *n_steps_out, float *mean_out);

/* demod_binary_fft_fftw.c:88-113 with the double-precision DFT of fft_double.c */
int brp_oracle_power_spectrum(const float *resampled, uint32_t nsamples, float norm_factor, float *ps_out);

/* demod_binary_hs_cpu.c:62-86 + hs_common.c:33-171.  sumspec_out = 4 x Fu floats (2,4,8,16 harmonics),
 * dirty_out = 5 x nr_pages int32. */
int brp_oracle_harmonic_summing(const float *ps, const brp_oracle_geometry *g, const float thr[5], float *sumspec_out,
                                int32_t *dirty_out);

/* per-bin maxima without the threshold-dependent store rule: v_k[j] = max over the segment of S_k(i).
 * (what the candidate scan of demod_binary.c:1310-1397 can observe once values <= thr are ignored) */
int brp_oracle_harmonic_maxima(const float *ps, const brp_oracle_geometry *g, float *maxima_out /* 5 x Fu */);

double brp_oracle_chisq_Q(double x, int n_harm);     /* Q_{chi^2}(x; 2 n_harm) */
double brp_oracle_chisq_Qinv(double Q, int n_harm);
void brp_oracle_base_thresholds(float fA, uint32_t fft_size, float thr_out[5]); /* demod_binary.c:1268-1282 without the toplist term */

/* demod_binary.c:1310-1397 for one template; tmpl_idx_io (500 entries) tracks which template filled each slot */
void brp_oracle_toplist_update(brp_oracle_cand *cands, int32_t *tmpl_idx_io, const float *const sumspec[5],
                               const int32_t *const dirty[5], uint32_t nr_pages, const brp_oracle_geometry *g,
                               const float thrA[5], const brp_oracle_template *t, int32_t tmpl_idx);

/* demod_binary.c:1502-1555: fA, power/sigma, sort by fA */
void brp_oracle_finalize(brp_oracle_cand *cands, int32_t *tmpl_idx_io);

/* demod_binary.c:1632-1667: emitted lines ("%6.12f %6.12f %6.12f %6.12f %g %g %d\n") + %DONE%.
 * Destroys cands (as the reference does). Returns bytes written or -1. */
int brp_oracle_format_candidates(brp_oracle_cand *cands, const brp_oracle_geometry *g, double tsample_us, char *buf,
                                 int buflen);

/* whole loop demod_binary.c:1180-1443 over bank[3*T] = (P, tau, Psi0) doubles.
 * cands_out: 500 entries in the state they have right after the loop (before finalize). */
int brp_oracle_search(const float *series, const brp_oracle_geometry *g, float fA, const double *bank,
                      uint32_t n_templates, brp_oracle_cand *cands_out, int32_t *tmpl_idx_out);

/* whitening pre-step with an empty zap list, demod_binary.c:858-1079 (rngmed.c:48 restated as an exact sliding median):
 * series (nsamples_unpadded floats) is replaced by its whitened version. */
int brp_oracle_whiten(float *series, uint32_t nsamples_unpadded, float padding, uint32_t window);
/* the same with the zap-list noise replacement: n_zap (fmin, fmax) pairs in Hz, in file order.  GSL's taus2 and ziggurat
 * Gaussian are restated from their published descriptions (tables: tools/gen_ziggurat.py) -> parity with libgsl unpinned */
int brp_oracle_whiten_zap(float *series, uint32_t nsamples_unpadded, float padding, uint32_t window, double tsample_us,
                          const double *zap, uint32_t n_zap);
/* exact running median, semantics of rngmed.c: medians[i] = median(input[i .. i+bsize-1]); even bsize: (a+b)/2.0 */
int brp_oracle_rngmed(const float *input, uint32_t length, uint32_t bsize, float *medians);

#ifdef __cplusplus
}
#endif
#endif
