/* TEST INFRASTRUCTURE ONLY (oracle/): see fft_double.h.
 *
 * Double-precision complex FFT for lengths 2^a 3^b:
 *   - leaves (<= LEAF_MAX points): Stockham autosort, radices 4/2/3, one master twiddle table;
 *   - larger lengths: one four-step split m = n1*n2 (n1 a power of two <= 2048) so that every
 *     sub-transform runs out of cache; inter-stage twiddles W_m^p come from a two-level table
 *     (hi/lo) generated with long-double sin/cos.
 * Written for this repository; it is not FFTW and shares no code with it.
 */
#include "fft_double.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define LEAF_MAX 8192
#define LO_BITS 11
#define LO_SIZE (1u << LO_BITS)
#define COLBLK 8

struct dfft_plan_s {
    size_t m;
    int is_leaf;
    int nfac;
    int fac[40];
    dcplx *tw;        /* leaf: W_m^j, j < m */
    dcplx *leafwork;  /* leaf: ping-pong buffer, m entries */
    size_t n1, n2;
    dfft_plan *p1, *p2;
    dcplx *tw_hi, *tw_lo; /* split: W_m^{hi<<LO_BITS}, W_m^{lo} */
    dcplx *scratch;       /* split: m entries */
    dcplx *colbuf;        /* split: COLBLK*n1 entries */
};

static inline dcplx cmul(dcplx a, dcplx b)
{
    dcplx r = { a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re };
    return r;
}

static void fill_twiddle(dcplx *t, size_t count, size_t stride, size_t modulus)
{
    /* t[j] = exp(-2 pi i (j*stride) / modulus) */
    const long double twopi = 6.283185307179586476925286766559005768L;
    for (size_t j = 0; j < count; j++) {
        size_t p = (j * stride) % modulus;
        long double ang = -twopi * (long double)p / (long double)modulus;
        t[j].re = (double)cosl(ang);
        t[j].im = (double)sinl(ang);
    }
}

static int factorize(size_t m, int *fac)
{
    int n = 0;
    while (m % 4 == 0) { fac[n++] = 4; m /= 4; }
    while (m % 2 == 0) { fac[n++] = 2; m /= 2; }
    while (m % 3 == 0) { fac[n++] = 3; m /= 3; }
    return m == 1 ? n : -1;
}

dfft_plan *dfft_plan_create(size_t m)
{
    int tmp[40];
    if (m == 0 || factorize(m, tmp) < 0) return NULL;
    dfft_plan *p = (dfft_plan *)calloc(1, sizeof(*p));
    if (!p) return NULL;
    p->m = m;
    if (m <= LEAF_MAX) {
        p->is_leaf = 1;
        p->nfac = factorize(m, p->fac);
        p->tw = (dcplx *)malloc(sizeof(dcplx) * m);
        p->leafwork = (dcplx *)malloc(sizeof(dcplx) * m);
        fill_twiddle(p->tw, m, 1, m);
        return p;
    }
    size_t n1 = 2048;
    while (m % n1 != 0) n1 >>= 1;
    p->n1 = n1;
    p->n2 = m / n1;
    p->p1 = dfft_plan_create(p->n1);
    p->p2 = dfft_plan_create(p->n2);
    size_t nhi = (m >> LO_BITS) + 1;
    p->tw_hi = (dcplx *)malloc(sizeof(dcplx) * nhi);
    p->tw_lo = (dcplx *)malloc(sizeof(dcplx) * LO_SIZE);
    fill_twiddle(p->tw_hi, nhi, LO_SIZE, m);
    fill_twiddle(p->tw_lo, LO_SIZE, 1, m);
    p->scratch = (dcplx *)malloc(sizeof(dcplx) * m);
    p->colbuf = (dcplx *)malloc(sizeof(dcplx) * COLBLK * n1);
    if (!p->p1 || !p->p2) { dfft_plan_destroy(p); return NULL; }
    return p;
}

void dfft_plan_destroy(dfft_plan *p)
{
    if (!p) return;
    free(p->tw); free(p->leafwork); free(p->tw_hi); free(p->tw_lo); free(p->scratch); free(p->colbuf);
    dfft_plan_destroy(p->p1);
    dfft_plan_destroy(p->p2);
    free(p);
}

/* One Stockham pass of radix r: n = current sub-length, s = stride (= m/n). */
static void pass2(size_t n, size_t s, const dcplx *x, dcplx *y, const dcplx *tw)
{
    const size_t h = n / 2;
    for (size_t p = 0; p < h; p++) {
        const dcplx w = tw[p * s];
        for (size_t q = 0; q < s; q++) {
            dcplx a = x[q + s * p], b = x[q + s * (p + h)];
            dcplx d = { a.re - b.re, a.im - b.im };
            y[q + s * (2 * p)].re = a.re + b.re;
            y[q + s * (2 * p)].im = a.im + b.im;
            y[q + s * (2 * p + 1)] = cmul(d, w);
        }
    }
}

static void pass4(size_t n, size_t s, const dcplx *x, dcplx *y, const dcplx *tw)
{
    const size_t h = n / 4;
    for (size_t p = 0; p < h; p++) {
        const dcplx w1 = tw[p * s], w2 = tw[2 * p * s], w3 = tw[3 * p * s];
        for (size_t q = 0; q < s; q++) {
            dcplx a = x[q + s * p], b = x[q + s * (p + h)], c = x[q + s * (p + 2 * h)], d = x[q + s * (p + 3 * h)];
            dcplx apc = { a.re + c.re, a.im + c.im }, amc = { a.re - c.re, a.im - c.im };
            dcplx bpd = { b.re + d.re, b.im + d.im }, bmd = { b.re - d.re, b.im - d.im };
            /* -i * bmd */
            dcplx jbmd = { bmd.im, -bmd.re };
            dcplx t0 = { apc.re + bpd.re, apc.im + bpd.im };
            dcplx t1 = { amc.re + jbmd.re, amc.im + jbmd.im };
            dcplx t2 = { apc.re - bpd.re, apc.im - bpd.im };
            dcplx t3 = { amc.re - jbmd.re, amc.im - jbmd.im };
            y[q + s * (4 * p)] = t0;
            y[q + s * (4 * p + 1)] = cmul(t1, w1);
            y[q + s * (4 * p + 2)] = cmul(t2, w2);
            y[q + s * (4 * p + 3)] = cmul(t3, w3);
        }
    }
}

static void pass3(size_t n, size_t s, const dcplx *x, dcplx *y, const dcplx *tw)
{
    const size_t h = n / 3;
    const double c3 = -0.5, s3 = 0.86602540378443864676372317075294;
    for (size_t p = 0; p < h; p++) {
        const dcplx w1 = tw[p * s], w2 = tw[2 * p * s];
        for (size_t q = 0; q < s; q++) {
            dcplx a = x[q + s * p], b = x[q + s * (p + h)], c = x[q + s * (p + 2 * h)];
            dcplx bpc = { b.re + c.re, b.im + c.im }, bmc = { b.re - c.re, b.im - c.im };
            dcplx t0 = { a.re + bpc.re, a.im + bpc.im };
            dcplx m = { a.re + c3 * bpc.re, a.im + c3 * bpc.im };
            /* -i * s3 * bmc */
            dcplx r = { s3 * bmc.im, -s3 * bmc.re };
            dcplx t1 = { m.re + r.re, m.im + r.im };
            dcplx t2 = { m.re - r.re, m.im - r.im };
            y[q + s * (3 * p)] = t0;
            y[q + s * (3 * p + 1)] = cmul(t1, w1);
            y[q + s * (3 * p + 2)] = cmul(t2, w2);
        }
    }
}

static void leaf_execute(const dfft_plan *p, dcplx *data)
{
    dcplx *x = data, *y = p->leafwork;
    size_t n = p->m, s = 1;
    for (int f = 0; f < p->nfac; f++) {
        const int r = p->fac[f];
        if (r == 4) pass4(n, s, x, y, p->tw);
        else if (r == 2) pass2(n, s, x, y, p->tw);
        else pass3(n, s, x, y, p->tw);
        n /= (size_t)r;
        s *= (size_t)r;
        dcplx *t = x; x = y; y = t;
    }
    if (x != data) memcpy(data, x, sizeof(dcplx) * p->m);
}

void dfft_execute(const dfft_plan *p, dcplx *data)
{
    if (p->is_leaf) { leaf_execute(p, data); return; }
    const size_t n1 = p->n1, n2 = p->n2;
    const size_t lomask = LO_SIZE - 1;
    /* step 1+2: length-n1 transforms down the columns of the n1 x n2 row-major view, then twiddle */
    for (size_t b0 = 0; b0 < n2; b0 += COLBLK) {
        const size_t nb = (n2 - b0 < COLBLK) ? (n2 - b0) : COLBLK;
        for (size_t a = 0; a < n1; a++)
            for (size_t j = 0; j < nb; j++) p->colbuf[j * n1 + a] = data[a * n2 + b0 + j];
        for (size_t j = 0; j < nb; j++) {
            dcplx *col = p->colbuf + j * n1;
            dfft_execute(p->p1, col);
            const size_t b = b0 + j;
            for (size_t c = 0; c < n1; c++) {
                const size_t pw = b * c; /* < m */
                col[c] = cmul(col[c], cmul(p->tw_hi[pw >> LO_BITS], p->tw_lo[pw & lomask]));
            }
        }
        for (size_t c = 0; c < n1; c++)
            for (size_t j = 0; j < nb; j++) data[c * n2 + b0 + j] = p->colbuf[j * n1 + c];
    }
    /* step 3: length-n2 transforms along the rows */
    for (size_t c = 0; c < n1; c++) dfft_execute(p->p2, data + c * n2);
    /* step 4: X[c + n1*d] = Z[c][d]  (blocked transpose through scratch) */
    const size_t TB = 32;
    for (size_t c0 = 0; c0 < n1; c0 += TB)
        for (size_t d0 = 0; d0 < n2; d0 += TB) {
            const size_t cmax = (c0 + TB < n1) ? c0 + TB : n1, dmax = (d0 + TB < n2) ? d0 + TB : n2;
            for (size_t d = d0; d < dmax; d++)
                for (size_t c = c0; c < cmax; c++) p->scratch[d * n1 + c] = data[c * n2 + d];
        }
    memcpy(data, p->scratch, sizeof(dcplx) * p->m);
}

/* ---------------- real transforms ---------------- */

struct drfft_plan_s {
    size_t n, m;
    dfft_plan *cp;
    dcplx *z;            /* m entries */
    dcplx *tw_hi, *tw_lo; /* W_n^p two-level */
};

drfft_plan *drfft_plan_create(size_t n)
{
    if (n < 2 || (n & 1)) return NULL;
    drfft_plan *p = (drfft_plan *)calloc(1, sizeof(*p));
    if (!p) return NULL;
    p->n = n;
    p->m = n / 2;
    p->cp = dfft_plan_create(p->m);
    if (!p->cp) { free(p); return NULL; }
    p->z = (dcplx *)malloc(sizeof(dcplx) * p->m);
    size_t nhi = (n >> LO_BITS) + 1;
    p->tw_hi = (dcplx *)malloc(sizeof(dcplx) * nhi);
    p->tw_lo = (dcplx *)malloc(sizeof(dcplx) * LO_SIZE);
    fill_twiddle(p->tw_hi, nhi, LO_SIZE, n);
    fill_twiddle(p->tw_lo, LO_SIZE, 1, n);
    return p;
}

void drfft_plan_destroy(drfft_plan *p)
{
    if (!p) return;
    dfft_plan_destroy(p->cp);
    free(p->z); free(p->tw_hi); free(p->tw_lo);
    free(p);
}

static inline dcplx rtw(const drfft_plan *p, size_t k)
{
    return cmul(p->tw_hi[k >> LO_BITS], p->tw_lo[k & (LO_SIZE - 1)]);
}

static inline dcplx untangle(const drfft_plan *p, size_t k)
{
    /* X[k] = E + W_n^k O,  E = (Z[k]+conj Z[m-k])/2,  O = (Z[k]-conj Z[m-k])/(2i) */
    const size_t m = p->m;
    const dcplx a = p->z[k % m], b = p->z[(m - k) % m];
    const dcplx e = { 0.5 * (a.re + b.re), 0.5 * (a.im - b.im) };
    const dcplx d = { 0.5 * (a.re - b.re), 0.5 * (a.im + b.im) }; /* (Z[k]-conj Z[m-k])/2 */
    const dcplx o = { d.im, -d.re };                              /* divide by i */
    const dcplx wo = cmul(rtw(p, k), o);
    dcplx x = { e.re + wo.re, e.im + wo.im };
    return x;
}

static void rfft_core(const drfft_plan *p, const float *in)
{
    for (size_t j = 0; j < p->m; j++) {
        p->z[j].re = (double)in[2 * j];
        p->z[j].im = (double)in[2 * j + 1];
    }
    dfft_execute(p->cp, p->z);
}

void drfft_execute_f64(const drfft_plan *p, const float *in, dcplx *out)
{
    rfft_core(p, in);
    for (size_t k = 0; k <= p->m; k++) out[k] = untangle(p, k);
    out[0].im = 0.0;
    out[p->m].im = 0.0;
}

void drfft_execute_f32(const drfft_plan *p, const float *in, float *out)
{
    rfft_core(p, in);
    for (size_t k = 0; k <= p->m; k++) {
        const dcplx x = untangle(p, k);
        out[2 * k] = (float)x.re;
        out[2 * k + 1] = (float)x.im;
    }
    out[1] = 0.0f;
    out[2 * p->m + 1] = 0.0f;
}

void drfft_execute_c2r_f32(const drfft_plan *p, const float *in, float *out)
{
    /* Z'[k] = (X[k]+conj X[m-k]) + i W_n^{-k} (X[k]-conj X[m-k]);  z = conj(FFT(conj Z')) */
    const size_t m = p->m;
    for (size_t k = 0; k < m; k++) {
        const dcplx a = { in[2 * k], in[2 * k + 1] };
        const dcplx b = { in[2 * (m - k)], in[2 * (m - k) + 1] };
        const dcplx s = { a.re + b.re, a.im - b.im };
        const dcplx d = { a.re - b.re, a.im + b.im };
        dcplx w = rtw(p, k);
        w.im = -w.im;
        const dcplx wd = cmul(w, d);
        /* Z' = s + i*wd; store conj(Z') */
        p->z[k].re = s.re - wd.im;
        p->z[k].im = -(s.im + wd.re);
    }
    dfft_execute(p->cp, p->z);
    for (size_t j = 0; j < m; j++) {
        out[2 * j] = (float)p->z[j].re;
        out[2 * j + 1] = (float)(-p->z[j].im);
    }
}
