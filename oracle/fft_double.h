/* TEST INFRASTRUCTURE ONLY (oracle/): double-precision mixed-radix (2,3) FFT used as the
 * arithmetic behind the oracle.  Never linked into the product library.
 *
 * Contract (what FFTW's fftwf_plan_dft_r2c_1d documents and what
 * /root/reference/demod_binary_fft_fftw.c:70,95 relies on):
 *     X[k] = sum_{n<N} x[n] * exp(-2*pi*i*k*n/N),  k = 0..N/2   (unnormalised, forward)
 * N must be even and of the form 2^a * 3^b.
 */
#ifndef BRP_ORACLE_FFT_DOUBLE_H
#define BRP_ORACLE_FFT_DOUBLE_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct { double re, im; } dcplx;
typedef struct dfft_plan_s dfft_plan;

/* complex forward transform of length m (m = 2^a 3^b); in place, natural order */
dfft_plan *dfft_plan_create(size_t m);
void dfft_plan_destroy(dfft_plan *p);
void dfft_execute(const dfft_plan *p, dcplx *data);

/* real -> half-complex, n even; out has n/2+1 entries.  work may be NULL (allocated inside). */
typedef struct drfft_plan_s drfft_plan;
drfft_plan *drfft_plan_create(size_t n);
void drfft_plan_destroy(drfft_plan *p);
void drfft_execute_f32(const drfft_plan *p, const float *in, float *out_interleaved /* 2*(n/2+1) floats */);
void drfft_execute_f64(const drfft_plan *p, const float *in, dcplx *out /* n/2+1 */);
/* half-complex -> real (unnormalised backward, FFTW c2r contract); in has n/2+1 entries */
void drfft_execute_c2r_f32(const drfft_plan *p, const float *in_interleaved, float *out);

#ifdef __cplusplus
}
#endif
#endif
