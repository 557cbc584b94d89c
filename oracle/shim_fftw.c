/* TEST INFRASTRUCTURE ONLY (oracle/): link-time stand-in for libfftw3f, backing the prototypes in
 * oracle/shim/fftw3.h with the double-precision transform in fft_double.c.  Used only to build
 * oracle/_ref/ (the reference's own CPU translation units compiled in place); never part of the
 * product.  Because FFTW is absent, the reference's FFT arithmetic itself is "parity unpinned":
 * what is pinned is the DFT definition, cross-checked against scipy's float64 rfft in tests/. */
#include "shim/fftw3.h"
#include "fft_double.h"
#include <stdlib.h>

struct brp_shim_plan_s {
    int n;
    int backward;
    float *rbuf;
    fftwf_complex *cbuf;
    drfft_plan *plan;
};

void *fftwf_malloc(size_t n)
{
    void *p = NULL;
    if (posix_memalign(&p, 64, n ? n : 64) != 0) return NULL;
    return p;
}
float *fftwf_alloc_real(size_t n) { return (float *)fftwf_malloc(n * sizeof(float)); }
void fftwf_free(void *p) { free(p); }
int fftwf_import_system_wisdom(void) { return 0; }
int fftwf_import_wisdom_from_string(const char *s) { (void)s; return 0; }

static fftwf_plan make_plan(int n, float *r, fftwf_complex *c, int backward)
{
    struct brp_shim_plan_s *p = (struct brp_shim_plan_s *)calloc(1, sizeof(*p));
    if (!p) return NULL;
    p->n = n;
    p->backward = backward;
    p->rbuf = r;
    p->cbuf = c;
    p->plan = drfft_plan_create((size_t)n);
    if (!p->plan) { free(p); return NULL; }
    return p;
}

fftwf_plan fftwf_plan_dft_r2c_1d(int n, float *in, fftwf_complex *out, unsigned flags)
{
    (void)flags;
    return make_plan(n, in, out, 0);
}

fftwf_plan fftwf_plan_dft_c2r_1d(int n, fftwf_complex *in, float *out, unsigned flags)
{
    (void)flags;
    return make_plan(n, out, in, 1);
}

void fftwf_execute(const fftwf_plan p)
{
    if (!p->backward) {
        /* in-place r2c (BRP_FFT_INPLACE) is safe: the transform copies its input first */
        drfft_execute_f32(p->plan, p->rbuf, (float *)p->cbuf);
    } else {
        drfft_execute_c2r_f32(p->plan, (const float *)p->cbuf, p->rbuf);
    }
}

void fftwf_destroy_plan(fftwf_plan p)
{
    if (!p) return;
    drfft_plan_destroy(p->plan);
    free(p);
}
