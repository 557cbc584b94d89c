/* TEST INFRASTRUCTURE ONLY (oracle/): see brp_oracle.h.
 *
 * CPU restatement of the reference algorithm, written for this repository.  Compiled as C++ with
 * "-O2 -ffp-contract=off" because the reference compiles its .c files with g++ (Makefile:80-111):
 * overload resolution of sin()/pow() on float arguments and the absence of FMA contraction are part
 * of the arithmetic being restated.  Every function cites the reference lines it follows.
 */
#include "brp_oracle.h"
#include "fft_double.h"
#include "../boinc-app-eah-brp_b200/csrc/ziggurat_tables.h" /* data only: regenerated GSL ziggurat tables */

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---- constants: erp_utilities.h:27-32, tables erp_utilities.cpp:45-46 (data, copied as data) ---- */
#define LUT_RES 64
#define LUT_RES_F (1.0f * LUT_RES)
#define LUT_RES_F_INV (1.0f / LUT_RES)
#define TWO_PI_F 6.283185f
#define TWO_PI_INV_F (1.0f / TWO_PI_F)

static const float kSin[LUT_RES + 1] = {
    0.000000f,  0.098017f,  0.195090f,  0.290285f,  0.382683f,  0.471397f,  0.555570f,  0.634393f,  0.707107f,
    0.773010f,  0.831470f,  0.881921f,  0.923880f,  0.956940f,  0.980785f,  0.995185f,  1.000000f,  0.995185f,
    0.980785f,  0.956940f,  0.923880f,  0.881921f,  0.831470f,  0.773010f,  0.707107f,  0.634393f,  0.555570f,
    0.471397f,  0.382683f,  0.290285f,  0.195091f,  0.098017f,  0.000000f,  -0.098017f, -0.195090f, -0.290284f,
    -0.382683f, -0.471397f, -0.555570f, -0.634393f, -0.707107f, -0.773010f, -0.831469f, -0.881921f, -0.923880f,
    -0.956940f, -0.980785f, -0.995185f, -1.000000f, -0.995185f, -0.980785f, -0.956940f, -0.923880f, -0.881921f,
    -0.831470f, -0.773011f, -0.707107f, -0.634394f, -0.555570f, -0.471397f, -0.382684f, -0.290285f, -0.195091f,
    -0.098017f, -0.000000f };
static const float kCos[LUT_RES + 1] = {
    1.000000f,  0.995185f,  0.980785f,  0.956940f,  0.923880f,  0.881921f,  0.831470f,  0.773010f,  0.707107f,
    0.634393f,  0.555570f,  0.471397f,  0.382683f,  0.290285f,  0.195090f,  0.098017f,  0.000000f,  -0.098017f,
    -0.195090f, -0.290285f, -0.382683f, -0.471397f, -0.555570f, -0.634393f, -0.707107f, -0.773010f, -0.831470f,
    -0.881921f, -0.923880f, -0.956940f, -0.980785f, -0.995185f, -1.000000f, -0.995185f, -0.980785f, -0.956940f,
    -0.923880f, -0.881921f, -0.831470f, -0.773011f, -0.707107f, -0.634393f, -0.555570f, -0.471397f, -0.382684f,
    -0.290285f, -0.195090f, -0.098017f, 0.000000f,  0.098017f,  0.195090f,  0.290285f,  0.382683f,  0.471397f,
    0.555570f,  0.634393f,  0.707107f,  0.773010f,  0.831470f,  0.881921f,  0.923879f,  0.956940f,  0.980785f,
    0.995185f,  1.000000f };

extern "C" {

/* demod_binary.c:782 (padding), :1087-1099 (derived sizes) */
void brp_oracle_derive_geometry(uint32_t nsamples_unpadded, double tsample_us, float padding, float f0,
                                uint32_t window, brp_oracle_geometry *g)
{
    uint32_t nsamples = nsamples_unpadded;
    nsamples = (int)(padding * nsamples + 0.5);
    g->nsamples_unpadded = nsamples_unpadded;
    g->nsamples = nsamples;
    g->t_obs = nsamples * tsample_us * 1.0e-6;
    g->dt = tsample_us * 1.0e-6;
    g->step_inv = 1.0 / g->dt;
    g->fft_size = (unsigned int)(nsamples * 0.5 + 0.5) + 1;
    g->window_2 = (unsigned int)(window * 0.5 + 0.5);
    const int fcap = (int)(g->fft_size - g->window_2);
    const int fu = (int)(f0 * g->t_obs + 0.5);
    const int hh = (int)(16.0 * f0 * g->t_obs + 0.5);
    g->fundamental_idx_hi = (unsigned int)(fcap < fu ? fcap : fu);
    g->harmonic_idx_hi = (unsigned int)(fcap < hh ? fcap : hh);
}

/* demod_binary.c:1208-1230 */
void brp_oracle_derive_template(double P_tmp, double tau_tmp, double Psi0_tmp, float step_inv, brp_oracle_template *t)
{
    float P = (float)P_tmp;
    float tau = (float)tau_tmp;
    float Psi0 = (float)Psi0_tmp;
    float Omega = 2.0 * M_PI / P;
    float S0 = tau * sin(Psi0) * step_inv; /* g++ resolves sin(float) to the float overload */
    t->P = P;
    t->tau = tau;
    t->Psi0 = Psi0;
    t->Omega = Omega;
    t->S0 = S0;
}

/* erp_utilities.cpp:176-209, sine branch only */
float brp_oracle_sin_lut(float x)
{
    float ipart;
    float xt = modff(TWO_PI_INV_F * x, &ipart);
    if (xt < 0.0f) xt += 1.0f;
    const int i0 = (int)(xt * LUT_RES_F + 0.5f);
    const float d = TWO_PI_F * (xt - LUT_RES_F_INV * i0);
    const float d2 = d * (0.5f * d);
    const float ts = kSin[i0], tc = kCos[i0];
    return ts + d * tc - d2 * ts;
}

/* demod_binary_resamp_cpu.c:80-136 */
int brp_oracle_resample(const float *series, const brp_oracle_geometry *g, const brp_oracle_template *t, float *out,
                        int32_t *idx_out, uint32_t *n_steps_out, float *mean_out)
{
    const uint32_t nu = g->nsamples_unpadded;
    float *del_t = (float *)malloc(sizeof(float) * nu);
    if (!del_t) return 1;
    float fi = 0.0f;
    for (uint32_t i = 0; i < nu; i++, fi += 1.0f) {
        const float tm = fi * g->dt;
        const float s = brp_oracle_sin_lut(t->Omega * tm + t->Psi0);
        del_t[i] = t->tau * s * g->step_inv - t->S0;
    }
    uint32_t n_steps = nu - 1;
    while (n_steps - del_t[n_steps] >= nu - 1) n_steps--; /* unsigned -> float, as in :108 */
    float mean = 0.0f;
    fi = 0.0f;
    uint32_t i = 0;
    for (; i < n_steps; i++, fi += 1.0f) {
        const int nearest = (int)(fi - del_t[i] + 0.5); /* float subtract, double add, truncate (:116) */
        if (idx_out) idx_out[i] = nearest;
        out[i] = series[nearest];
        mean += out[i];
    }
    mean /= fi;
    for (; i < g->nsamples; i++) out[i] = mean;
    if (n_steps_out) *n_steps_out = n_steps;
    if (mean_out) *mean_out = mean;
    free(del_t);
    return 0;
}

/* demod_binary_fft_fftw.c:88-113; the transform is the DFT definition in double precision */
int brp_oracle_power_spectrum(const float *resampled, uint32_t nsamples, float norm_factor, float *ps_out)
{
    /* one cached plan per process (test infrastructure: single-threaded by design) */
    static drfft_plan *cached = NULL;
    static uint32_t cached_n = 0;
    static float *spec = NULL;
    const size_t fft_size = (size_t)nsamples / 2 + 1;
    if (!cached || cached_n != nsamples) {
        drfft_plan_destroy(cached);
        free(spec);
        cached = drfft_plan_create(nsamples);
        cached_n = nsamples;
        spec = cached ? (float *)malloc(sizeof(float) * 2 * fft_size) : NULL;
        if (!cached || !spec) { drfft_plan_destroy(cached); cached = NULL; free(spec); spec = NULL; return 4; }
    }
    drfft_execute_f32(cached, resampled, spec); /* rounded to float like FFTW's float output */
    for (size_t k = 1; k < fft_size; k++) {
        const double re = spec[2 * k], im = spec[2 * k + 1];
        ps_out[k] = norm_factor * (re * re + im * im);
    }
    ps_out[0] = 0.0f;
    return 0;
}

/* hs_common.c:33-171 (with the wrapper's zeroing, demod_binary_hs_cpu.c:73-80) */
int brp_oracle_harmonic_summing(const float *ps, const brp_oracle_geometry *g, const float thr[5], float *sumspec_out,
                                int32_t *dirty_out)
{
    const uint32_t fu = g->fundamental_idx_hi, w2 = g->window_2, hi = g->harmonic_idx_hi;
    const uint32_t nr_pages = (fu >> BRP_ORACLE_LOG_PAGE) + 1;
    memset(sumspec_out, 0, sizeof(float) * 4 * (size_t)fu);
    memset(dirty_out, 0, sizeof(int32_t) * 5 * (size_t)nr_pages);
    float *ss[5] = { NULL, sumspec_out, sumspec_out + fu, sumspec_out + 2 * (size_t)fu, sumspec_out + 3 * (size_t)fu };
    int32_t *dirty[5];
    for (int k = 0; k < 5; k++) dirty[k] = dirty_out + (size_t)k * nr_pages;

    int prev_j[5] = { -1, -1, -1, -1, -1 };
    float prev_max[5] = { 0, 0, 0, 0, 0 };
    for (uint32_t i = w2; i < hi; i++) {
#define PH(l) ps[((uint32_t)(l) * i + 8u) >> 4]
        float sum = ps[i];
        if (sum > thr[0] && i < fu) dirty[0][i >> BRP_ORACLE_LOG_PAGE] = 1;
        float power = 0.0f; /* the reference leaves this stale when j >= Fu; never observable */
        for (int k = 1; k < 5; k++) {
            int j;
            if (k == 1) {
                j = (int)((8u * i + 8u) >> 4);
                sum += ps[j];
            } else if (k == 2) {
                j = (int)((4u * i + 8u) >> 4);
                sum += PH(12) + ps[j];
            } else if (k == 3) {
                j = (int)((2u * i + 8u) >> 4);
                sum += PH(14) + PH(10) + PH(6) + ps[j];
            } else {
                j = (int)((i + 8u) >> 4);
                sum += PH(15) + PH(13) + PH(11) + PH(9) + PH(7) + PH(5) + PH(3) + ps[j];
            }
            if (j != prev_j[k]) prev_max[k] = 0.0f;
            if ((uint32_t)j < fu) {
                power = (sum > prev_max[k]) ? sum : prev_max[k];
                if (power > thr[k]) {
                    ss[k][j] = power;
                    dirty[k][j >> BRP_ORACLE_LOG_PAGE] = 1;
                } else if (j != prev_j[k]) {
                    ss[k][j] = power;
                }
            }
            prev_j[k] = j;
            prev_max[k] = power;
        }
#undef PH
    }
    return 0;
}

int brp_oracle_harmonic_maxima(const float *ps, const brp_oracle_geometry *g, float *mx)
{
    const uint32_t fu = g->fundamental_idx_hi, w2 = g->window_2, hi = g->harmonic_idx_hi;
    memset(mx, 0, sizeof(float) * 5 * (size_t)fu);
    for (uint32_t i = w2; i < hi; i++) {
#define PH(l) ps[((uint32_t)(l) * i + 8u) >> 4]
        float s = ps[i];
        if (i < fu && s > mx[i]) mx[i] = s;
        uint32_t j = (8u * i + 8u) >> 4;
        s += ps[j];
        if (j < fu && s > mx[fu + j]) mx[fu + j] = s;
        j = (4u * i + 8u) >> 4;
        s += PH(12) + ps[j];
        if (j < fu && s > mx[2 * (size_t)fu + j]) mx[2 * (size_t)fu + j] = s;
        j = (2u * i + 8u) >> 4;
        s += PH(14) + PH(10) + PH(6) + ps[j];
        if (j < fu && s > mx[3 * (size_t)fu + j]) mx[3 * (size_t)fu + j] = s;
        j = (i + 8u) >> 4;
        s += PH(15) + PH(13) + PH(11) + PH(9) + PH(7) + PH(5) + PH(3) + ps[j];
        if (j < fu && s > mx[4 * (size_t)fu + j]) mx[4 * (size_t)fu + j] = s;
#undef PH
    }
    return 0;
}

/* chi-square upper tail for 2n degrees of freedom: Q(x) = exp(-x/2) sum_{k<n} (x/2)^k/k!
 * (stands in for gsl_cdf_chisq_Q / _Qinv, demod_binary.c:1281,1517-1545; GSL is not in the tree) */
static double log_chisq_Q(double x, int n)
{
    const double h = 0.5 * x;
    double term = 1.0, sum = 1.0;
    for (int k = 1; k < n; k++) { term *= h / k; sum += term; }
    return -h + log(sum);
}

double brp_oracle_chisq_Q(double x, int n)
{
    if (x <= 0.0) return 1.0;
    const double h = 0.5 * x;
    double term = 1.0, sum = 1.0;
    for (int k = 1; k < n; k++) { term *= h / k; sum += term; }
    return exp(-h) * sum;
}

double brp_oracle_chisq_Qinv(double Q, int n)
{
    if (Q >= 1.0) return 0.0;
    if (Q <= 0.0) return INFINITY;
    const double target = log(Q);
    double lo = 0.0, hi = 2.0 * n + 10.0;
    while (log_chisq_Q(hi, n) > target) hi *= 2.0;
    for (int it = 0; it < 200; it++) {
        const double mid = 0.5 * (lo + hi);
        if (log_chisq_Q(mid, n) > target) lo = mid; else hi = mid;
        if (hi - lo <= 4e-16 * hi) break;
    }
    return 0.5 * (lo + hi);
}

void brp_oracle_base_thresholds(float fA, uint32_t fft_size, float thr_out[5])
{
    for (unsigned int k = 0; k < 5; k++) {
        unsigned int N_h = (int)(pow(2.0, k) + 0.5);
        float prob = 1.0 - pow(1.0 - fA, 1.0 / fft_size);
        thr_out[k] = (float)(0.5 * brp_oracle_chisq_Qinv(prob, (int)N_h));
    }
}

/* demod_binary.c:1699-1710 */
static int by_power_desc(const void *a, const void *b)
{
    const brp_oracle_cand *x = (const brp_oracle_cand *)a, *y = (const brp_oracle_cand *)b;
    if (x->power < y->power) return 1;
    if (x->power > y->power) return -1;
    return 0;
}

/* demod_binary.c:1713-1740 */
static int by_ifa_desc(const void *a, const void *b)
{
    const brp_oracle_cand *x = (const brp_oracle_cand *)a, *y = (const brp_oracle_cand *)b;
    if (x->fA < y->fA) return 1;
    if (x->fA > y->fA) return -1;
    if (x->power < y->power) return 1;
    if (x->power > y->power) return -1;
    if (x->f0 < y->f0) return 1;
    if (x->f0 > y->f0) return -1;
    return 0;
}

/* the template index rides along in a parallel array; to keep it attached through qsort the sort is
 * done on (cand, idx) pairs with the same comparator (the comparator only looks at the cand part) */
typedef struct __attribute__((__packed__)) { brp_oracle_cand c; int32_t t; } cand_tagged;

static void sort_tagged(brp_oracle_cand *c, int32_t *t, int n, int (*cmp)(const void *, const void *))
{
    cand_tagged *tmp = (cand_tagged *)malloc(sizeof(cand_tagged) * n);
    for (int i = 0; i < n; i++) { tmp[i].c = c[i]; tmp[i].t = t ? t[i] : -1; }
    qsort(tmp, n, sizeof(cand_tagged), cmp);
    for (int i = 0; i < n; i++) { c[i] = tmp[i].c; if (t) t[i] = tmp[i].t; }
    free(tmp);
}

/* demod_binary.c:1310-1397 (screensaver binning :1383-1393 omitted: no effect on results) */
void brp_oracle_toplist_update(brp_oracle_cand *cands, int32_t *tidx, const float *const sumspec[5],
                               const int32_t *const dirty[5], uint32_t nr_pages, const brp_oracle_geometry *g,
                               const float thrA[5], const brp_oracle_template *t, int32_t tmpl_idx)
{
    const uint32_t fu = g->fundamental_idx_hi;
    for (unsigned int k = 0; k < 5; k++) {
        const int first = k * BRP_ORACLE_N_CAND_5, last = (k + 1) * BRP_ORACLE_N_CAND_5 - 1;
        const unsigned int N_h = (int)(pow(2.0, k) + 0.5);
        const float thr = thrA[k];
        uint32_t i = g->window_2;
        while (i < fu) {
            uint32_t page = i >> BRP_ORACLE_LOG_PAGE;
            while (page < nr_pages && dirty[k][page] == 0) { page++; i = page << BRP_ORACLE_LOG_PAGE; }
            if (i >= fu) break;
            uint32_t next = (page + 1) << BRP_ORACLE_LOG_PAGE;
            if (next > fu) next = fu;
            for (; i < next; i++) {
                const float power = sumspec[k][i];
                if (power > thr && power > cands[last].power) {
                    int store = last;
                    for (int c = first; c <= last; c++) {
                        if (cands[c].f0 == i) { store = (cands[c].power < power) ? c : -1; break; }
                    }
                    if (store >= first) {
                        cands[store].f0 = i;
                        cands[store].P_b = t->P;
                        cands[store].tau = t->tau;
                        cands[store].Psi = t->Psi0;
                        cands[store].power = power;
                        cands[store].n_harm = N_h;
                        if (tidx) tidx[store] = tmpl_idx;
                        sort_tagged(cands + first, tidx ? tidx + first : NULL, BRP_ORACLE_N_CAND_5, by_power_desc);
                    }
                }
            }
        }
    }
}

/* demod_binary.c:1502-1555 */
void brp_oracle_finalize(brp_oracle_cand *cands, int32_t *tidx)
{
    for (int i = 0; i < BRP_ORACLE_N_CAND; i++) {
        float sigma;
        int n = (int)cands[i].n_harm;
        switch (n) {
            case 1: sigma = 1.0; break;
            case 2: sigma = sqrt(2.0); break;
            case 4: sigma = 2.0; break;
            case 8: sigma = sqrt(8.0); break;
            case 16: sigma = 4.0; break;
            default: cands[i].fA = -10.0; continue;
        }
        const double fA = brp_oracle_chisq_Q(2.0 * cands[i].power, n);
        cands[i].fA = (fA > 0.0) ? -log10(fA) : 320.0;
        cands[i].power /= sigma;
    }
    sort_tagged(cands, tidx, BRP_ORACLE_N_CAND, by_ifa_desc);
}

/* demod_binary.c:1632-1667 */
int brp_oracle_format_candidates(brp_oracle_cand *cands, const brp_oracle_geometry *g, double tsample_us, char *buf,
                                 int buflen)
{
    int off = 0, counter = 0;
    while (counter < BRP_ORACLE_N_CAND_5 && cands[0].fA > 0.0) {
        const double t_obs = g->nsamples * tsample_us * 1.0e-6;
        const double res_factor = 1.0 / t_obs;
        int n = snprintf(buf + off, buflen - off, "%6.12f %6.12f %6.12f %6.12f %g %g %d\n", cands[0].f0 * res_factor,
                         cands[0].P_b, cands[0].tau, cands[0].Psi, cands[0].power, cands[0].fA, cands[0].n_harm);
        if (n < 0 || n >= buflen - off) return -1;
        off += n;
        counter++;
        const uint32_t f0 = cands[0].f0;
        for (int j = 0; j < BRP_ORACLE_N_CAND; j++)
            if (cands[j].f0 == f0) cands[j].fA = -10.0;
        qsort(cands, BRP_ORACLE_N_CAND, sizeof(brp_oracle_cand), by_ifa_desc);
    }
    int n = snprintf(buf + off, buflen - off, "%%DONE%%\n");
    if (n < 0 || n >= buflen - off) return -1;
    return off + n;
}

/* demod_binary.c:1180-1443 */
int brp_oracle_search(const float *series, const brp_oracle_geometry *g, float fA, const double *bank,
                      uint32_t n_templates, brp_oracle_cand *cands, int32_t *tidx)
{
    const uint32_t fu = g->fundamental_idx_hi;
    const uint32_t nr_pages = (fu >> BRP_ORACLE_LOG_PAGE) + 1;
    float *resampled = (float *)malloc(sizeof(float) * g->nsamples);
    float *ps = (float *)malloc(sizeof(float) * g->fft_size);
    float *ss = (float *)malloc(sizeof(float) * 4 * (size_t)fu);
    int32_t *dirty = (int32_t *)malloc(sizeof(int32_t) * 5 * (size_t)nr_pages);
    if (!resampled || !ps || !ss || !dirty) return 1;
    memset(cands, 0, sizeof(brp_oracle_cand) * BRP_ORACLE_N_CAND);
    if (tidx) for (int i = 0; i < BRP_ORACLE_N_CAND; i++) tidx[i] = -1;
    float base[5];
    brp_oracle_base_thresholds(fA, g->fft_size, base);
    int rc = 0;
    for (uint32_t it = 0; it < n_templates && !rc; it++) {
        brp_oracle_template t;
        brp_oracle_derive_template(bank[3 * it], bank[3 * it + 1], bank[3 * it + 2], g->step_inv, &t);
        rc = brp_oracle_resample(series, g, &t, resampled, NULL, NULL, NULL);
        if (rc) break;
        const float norm = 1.0 / g->nsamples;
        rc = brp_oracle_power_spectrum(resampled, g->nsamples, norm, ps);
        if (rc) break;
        float thrA[5];
        for (int k = 0; k < 5; k++) {
            const float p = cands[(k + 1) * BRP_ORACLE_N_CAND_5 - 1].power;
            thrA[k] = fmaxf(p, base[k]);
        }
        brp_oracle_harmonic_summing(ps, g, thrA, ss, dirty);
        const float *sp[5] = { ps, ss, ss + fu, ss + 2 * (size_t)fu, ss + 3 * (size_t)fu };
        const int32_t *dp[5];
        for (int k = 0; k < 5; k++) dp[k] = dirty + (size_t)k * nr_pages;
        brp_oracle_toplist_update(cands, tidx, sp, dp, nr_pages, g, thrA, &t, (int32_t)it);
    }
    free(resampled); free(ps); free(ss); free(dirty);
    return rc;
}

/* rngmed.c:48-341 computes the same numbers with a linked list and checkpoints; here: a sorted window */
int brp_oracle_rngmed(const float *input, uint32_t length, uint32_t bsize, float *medians)
{
    if (bsize == 0 || bsize > length) return 4;
    float *win = (float *)malloc(sizeof(float) * bsize);
    if (!win) return 1;
    memcpy(win, input, sizeof(float) * bsize);
    qsort(win, bsize, sizeof(float), [](const void *a, const void *b) {
        const float x = *(const float *)a, y = *(const float *)b;
        return (x > y) - (x < y);
    });
    const uint32_t mid = (bsize + (bsize & 1)) / 2 - 1; /* rngmed.c:104 */
    for (uint32_t i = 0;; i++) {
        medians[i] = (bsize & 1) ? win[mid] : (float)((win[mid] + win[mid + 1]) / 2.0); /* rngmed.c:177-179,327-329 */
        if (i + bsize >= length) break;
        /* replace input[i] by input[i+bsize] keeping the window sorted */
        const float oldv = input[i], newv = input[i + bsize];
        uint32_t lo = 0, hi = bsize;
        while (lo < hi) { const uint32_t m = (lo + hi) / 2; if (win[m] < oldv) lo = m + 1; else hi = m; }
        uint32_t pos = lo; /* first element == oldv */
        if (newv >= oldv) {
            while (pos + 1 < bsize && win[pos + 1] < newv) { win[pos] = win[pos + 1]; pos++; }
        } else {
            while (pos > 0 && win[pos - 1] > newv) { win[pos] = win[pos - 1]; pos--; }
        }
        win[pos] = newv;
    }
    free(win);
    return 0;
}

/* ---- zap-list noise: GSL taus2 + gsl_ran_gaussian_ziggurat restated (GSL is not under /root/reference: the generator
 * follows its published description, the tables are regenerated by tools/gen_ziggurat.py -> parity with libgsl unpinned;
 * pinned against oracle/_ref, whose GSL stand-in is oracle/shim_gsl.c) ---- */
struct Taus2 { unsigned long s1, s2, s3; };
static unsigned long taus2_get(Taus2 *r)
{
#define BRP_TAUS(s, a, b, c, d) ((((s) & (c)) << (d)) & 0xffffffffUL) ^ (((((s) << (a)) & 0xffffffffUL) ^ (s)) >> (b))
    r->s1 = BRP_TAUS(r->s1, 13, 19, 4294967294UL, 12);
    r->s2 = BRP_TAUS(r->s2, 2, 25, 4294967288UL, 4);
    r->s3 = BRP_TAUS(r->s3, 3, 11, 4294967280UL, 17);
#undef BRP_TAUS
    return r->s1 ^ r->s2 ^ r->s3;
}
static void taus2_set(Taus2 *r, unsigned long s)
{
    if (s == 0) s = 1;
    r->s1 = (69069UL * s) & 0xffffffffUL;  if (r->s1 < 2) r->s1 += 2UL;
    r->s2 = (69069UL * r->s1) & 0xffffffffUL; if (r->s2 < 8) r->s2 += 8UL;
    r->s3 = (69069UL * r->s2) & 0xffffffffUL; if (r->s3 < 16) r->s3 += 16UL;
    for (int i = 0; i < 6; i++) taus2_get(r);
}
static double zig_gaussian(Taus2 *r, double sigma)
{
    for (;;) {
        const unsigned long k = taus2_get(r);
        unsigned long i = k & 0xFF;
        const unsigned long j = (k >> 8) & 0xFFFFFF;
        const int sign = (i & 0x80) ? +1 : -1;
        i &= 0x7f;
        double x = j * brp_zig_wtab[i], y;
        if (j < brp_zig_ktab[i]) return sign * sigma * x;
        if (i < 127) {
            const double y0 = brp_zig_ytab[i], y1 = brp_zig_ytab[i + 1];
            y = y1 + (y0 - y1) * (taus2_get(r) / 4294967296.0);
        } else {
            const double U1 = 1.0 - taus2_get(r) / 4294967296.0;
            const double U2 = taus2_get(r) / 4294967296.0;
            x = BRP_ZIG_R - log(U1) / BRP_ZIG_R;
            y = exp(-BRP_ZIG_R * (x - 0.5 * BRP_ZIG_R)) * U2;
        }
        if (y < exp(-0.5 * x * x)) return sign * sigma * x;
    }
}

/* demod_binary.c:858-1079.  zap: n_zap pairs (fmin, fmax) in Hz in file order (:993-1023), may be NULL with n_zap = 0;
 * tsample_us is only used for the bin numbers of the zap ranges (:996,1012-1013). */
int brp_oracle_whiten_zap(float *series, uint32_t n_unpadded, float padding, uint32_t window, double tsample_us,
                          const double *zap, uint32_t n_zap)
{
    uint32_t nsamples = n_unpadded;
    nsamples = (int)(padding * nsamples + 0.5);
    const unsigned int window_2 = (int)(0.5 * window + 0.5);
    const unsigned int fft_size = (unsigned int)(0.5 * nsamples + 0.5) + 1;
    if (fft_size < window) return 4;
    drfft_plan *plan = drfft_plan_create(nsamples);
    if (!plan) return 4;
    float *time_series = (float *)calloc(nsamples, sizeof(float));
    float *fft = (float *)malloc(sizeof(float) * 2 * fft_size);
    float *ps = (float *)calloc(fft_size, sizeof(float));
    const int white_size = fft_size - window + 1;
    float *med = (float *)calloc(white_size, sizeof(float));
    if (!time_series || !fft || !ps || !med) return 1;
    for (uint32_t i = 0; i < n_unpadded; i++) time_series[i] = series[i];
    int32_t seed;
    memcpy(&seed, series, sizeof(seed)); /* :917 the first sample's bit pattern */
    drfft_execute_f32(plan, time_series, fft);
    for (unsigned int i = 1; i < fft_size; i++) {
        const double re = fft[2 * i], im = fft[2 * i + 1];
        ps[i] = re * re + im * im;
    }
    brp_oracle_rngmed(ps, fft_size, window, med);
    for (int i = 0; i < white_size; i++) {
        float factor = sqrt(M_LN2 / med[i]);
        fft[2 * (i + window_2)] *= factor;
        fft[2 * (i + window_2) + 1] *= factor;
    }
    if (n_zap) { /* :990-1023 */
        Taus2 r;
        taus2_set(&r, (unsigned long)seed);
        const double t_obs = nsamples * tsample_us * 1.0e-6;
        const double sigma = M_SQRT1_2 * sqrt(padding);
        for (uint32_t z = 0; z < n_zap; z++) {
            const unsigned int idx_min = (unsigned int)(zap[2 * z] * t_obs + 0.5);
            const unsigned int idx_max = (unsigned int)(zap[2 * z + 1] * t_obs + 0.5);
            for (unsigned int idx = idx_min; idx <= idx_max; idx++) {
                const float re = zig_gaussian(&r, sigma), im = zig_gaussian(&r, sigma);
                if (idx < fft_size) { fft[2 * idx] = re; fft[2 * idx + 1] = im; } /* the reference writes out of bounds beyond */
            }
        }
    }
    for (unsigned int i = 0; i < window_2; i++) {
        fft[2 * i] = fft[2 * i + 1] = 0.0f;
        fft[2 * (fft_size - i - 1)] = fft[2 * (fft_size - i - 1) + 1] = 0.0f;
    }
    drfft_execute_c2r_f32(plan, fft, time_series);
    const float norm_factor = 1.0 / sqrt((float)nsamples);
    for (uint32_t i = 0; i < n_unpadded; i++) series[i] = norm_factor * time_series[i];
    free(time_series); free(fft); free(ps); free(med);
    drfft_plan_destroy(plan);
    return 0;
}

int brp_oracle_whiten(float *series, uint32_t n_unpadded, float padding, uint32_t window)
{
    return brp_oracle_whiten_zap(series, n_unpadded, padding, window, 0.0, NULL, 0);
}

} /* extern "C" */
