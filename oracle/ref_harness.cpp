/* TEST INFRASTRUCTURE ONLY (oracle/): thin extern "C" harness around the REFERENCE's own CPU
 * translation units, which oracle/Makefile compiles unmodified, in place, from /root/reference into
 * oracle/_ref/.  Nothing here restates reference arithmetic except brp_ref_resamp_indices(), which
 * re-evaluates the index expression of demod_binary_resamp_cpu.c:91-116 by calling the reference's
 * own sincosLUTLookup() (the reference never materialises the index array).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/--impl reference legs may load
 * the resulting library.  The product library never links or dlopens it.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "structs.h"
#include "diptr.h"
#include "demod_binary.h"
#include "erp_utilities.h"
#include "demod_binary_resamp_cpu.h"
#include "demod_binary_fft_fftw.h"
#include "demod_binary_hs_cpu.h"
#include "hs_common.h"
#include "rngmed.h"

extern "C" {

/* reference sincosLUTLookup (erp_utilities.cpp:176-209), sine branch */
float brp_ref_sin_lut(float x)
{
    float s = 0.0f, c = 0.0f;
    sincosLUTLookup(x, &s, &c);
    return s;
}

static void fill_params(RESAMP_PARAMS *p, uint32_t nsamples, uint32_t nsamples_unpadded, float tau, float Omega,
                        float Psi0, float dt, float step_inv, float S0)
{
    p->nsamples = nsamples;
    p->nsamples_unpadded = nsamples_unpadded;
    p->fft_size = (unsigned int)(nsamples * 0.5 + 0.5) + 1; /* demod_binary.c:1092 */
    p->tau = tau;
    p->Omega = Omega;
    p->Psi0 = Psi0;
    p->dt = dt;
    p->step_inv = step_inv;
    p->S0 = S0;
}

/* run the reference resampler once: set_up_resampling / run_resampling / tear_down_resampling
 * (demod_binary_resamp_cpu.c:39-147); out must hold nsamples floats */
int brp_ref_resample(const float *series, uint32_t nsamples, uint32_t nsamples_unpadded, float tau, float Omega,
                     float Psi0, float dt, float step_inv, float S0, float *out)
{
    RESAMP_PARAMS p;
    fill_params(&p, nsamples, nsamples_unpadded, tau, Omega, Psi0, dt, step_inv, S0);
    DIfloatPtr in, res;
    in.host_ptr = const_cast<float *>(series);
    res.host_ptr = NULL;
    int rc = set_up_resampling(in, &res, &p, NULL, NULL);
    if (rc) return rc;
    rc = run_resampling(in, res, &p);
    if (!rc) memcpy(out, res.host_ptr, sizeof(float) * nsamples);
    tear_down_resampling(res);
    return rc;
}

/* index array + n_steps exactly as demod_binary_resamp_cpu.c:91-116 computes them */
int brp_ref_resamp_indices(uint32_t nsamples_unpadded, float tau, float Omega, float Psi0, float dt, float step_inv,
                           float S0, int32_t *idx_out, uint32_t *n_steps_out, float *del_t_out)
{
    float *del_t = (float *)malloc(sizeof(float) * nsamples_unpadded);
    if (!del_t) return RADPUL_EMEM;
    unsigned int i;
    float i_f;
    for (i = 0, i_f = 0.0f; i < nsamples_unpadded; i++, i_f += 1.0f) {
        float t = i_f * dt;
        float sinValue = 0.0f, dummy = 0.0f;
        sincosLUTLookup(Omega * t + Psi0, &sinValue, &dummy);
        del_t[i] = tau * sinValue * step_inv - S0;
    }
    unsigned int n_steps = nsamples_unpadded - 1;
    while (n_steps - del_t[n_steps] >= nsamples_unpadded - 1) n_steps--;
    for (i = 0, i_f = 0.0f; i < n_steps; i++, i_f += 1.0f) idx_out[i] = (int)(i_f - del_t[i] + 0.5);
    *n_steps_out = n_steps;
    if (del_t_out) memcpy(del_t_out, del_t, sizeof(float) * nsamples_unpadded);
    free(del_t);
    return 0;
}

/* reference set_up_fft / run_fft / tear_down_fft (demod_binary_fft_fftw.c:46-125); the FFT behind
 * fftwf_execute is oracle/shim_fftw.c, NOT FFTW.  ps_out must hold fft_size floats. */
int brp_ref_fft_power(float *resampled, uint32_t nsamples, float norm_factor, float *ps_out)
{
    const unsigned int fft_size = (unsigned int)(nsamples * 0.5 + 0.5) + 1;
    DIfloatPtr in, ps;
    in.host_ptr = resampled;
    ps.host_ptr = NULL;
    int rc = set_up_fft(in, &ps, nsamples, fft_size);
    if (rc) return rc;
    rc = run_fft(in, ps, nsamples, fft_size, norm_factor);
    if (!rc) memcpy(ps_out, ps.host_ptr, sizeof(float) * fft_size);
    tear_down_fft(ps);
    return rc;
}

/* reference set_up/run/tear_down_harmonic_summing (demod_binary_hs_cpu.c:34-104, hs_common.c:33-171).
 * sumspec_out: 4 arrays (harmonics 2,4,8,16) of fundamental_idx_hi floats, concatenated;
 * dirty_out: 5 arrays of nr_pages int32, concatenated. */
int brp_ref_harmonic_summing(float *powerspectrum, uint32_t window_2, uint32_t fundamental_idx_hi,
                             uint32_t harmonic_idx_hi, float *thresholds, float *sumspec_out, int32_t *dirty_out,
                             uint32_t *nr_pages_out)
{
    float *sumspec[5] = { 0, 0, 0, 0, 0 };
    int32_t *dirty[5] = { 0, 0, 0, 0, 0 };
    unsigned int nr_pages = 0;
    int rc = set_up_harmonic_summing(sumspec, dirty, &nr_pages, fundamental_idx_hi, harmonic_idx_hi);
    if (rc) return rc;
    DIfloatPtr ps;
    ps.host_ptr = powerspectrum;
    rc = run_harmonic_summing(sumspec, dirty, nr_pages, ps, window_2, fundamental_idx_hi, harmonic_idx_hi, thresholds);
    if (!rc) {
        for (int k = 1; k < 5; k++)
            memcpy(sumspec_out + (size_t)(k - 1) * fundamental_idx_hi, sumspec[k], sizeof(float) * fundamental_idx_hi);
        for (int k = 0; k < 5; k++) memcpy(dirty_out + (size_t)k * nr_pages, dirty[k], sizeof(int32_t) * nr_pages);
        *nr_pages_out = nr_pages;
    }
    tear_down_harmonic_summing(sumspec, dirty);
    return rc;
}

/* reference running median (rngmed.c:48) */
int brp_ref_rngmed(const float *data, unsigned int n, unsigned int window, float *medians)
{
    rngmed(data, n, window, medians);
    return 0;
}

/* the reference driver itself (demod_binary.c:117), unmodified: same argv as the CLI */
int brp_ref_MAIN(int argc, char **argv) { return MAIN(argc, argv); }

} /* extern "C" */
