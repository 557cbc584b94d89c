/* TEST INFRASTRUCTURE ONLY (oracle/): minimal stand-in for <fftw3.h>.
 *
 * FFTW 3 (single precision, libfftw3f; version unpinned by the reference, see
 * debian/control:6-7) is not installed in this image and cannot be installed.
 * The reference translation units only use the ten entry points below
 * (demod_binary_fft_fftw.c:53,66-70,95,120-122; demod_binary_resamp_cpu.c:62-64,144).
 * They are implemented in oracle/shim_fftw.c on top of a double-precision
 * mixed-radix (2,3) CPU transform, so what this header promises is the published
 * FFTW contract only: an unnormalised forward real-to-complex DFT in the
 * half-complex [re,im] x (n/2+1) layout.  It is NOT FFTW.
 */
#ifndef BRP_ORACLE_SHIM_FFTW3_H
#define BRP_ORACLE_SHIM_FFTW3_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef float fftwf_complex[2];
typedef struct brp_shim_plan_s *fftwf_plan;
#define FFTW_ESTIMATE (1U << 6)
#define FFTW_MEASURE (0U)
void *fftwf_malloc(size_t n);
float *fftwf_alloc_real(size_t n);
void fftwf_free(void *p);
int fftwf_import_system_wisdom(void);
int fftwf_import_wisdom_from_string(const char *s);
fftwf_plan fftwf_plan_dft_r2c_1d(int n, float *in, fftwf_complex *out, unsigned flags);
fftwf_plan fftwf_plan_dft_c2r_1d(int n, fftwf_complex *in, float *out, unsigned flags);
void fftwf_execute(const fftwf_plan p);
void fftwf_destroy_plan(fftwf_plan p);
#ifdef __cplusplus
}
#endif
#endif
