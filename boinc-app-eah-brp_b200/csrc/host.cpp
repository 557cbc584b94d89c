// Host side of libbrp_b200.so: the driver arithmetic that surrounds the GPU loop, restated from
// demod_binary.c so that a caller of the batched API gets the reference's numbers without the reference
// driver: derived sizes (:782,1087-1099), per-template constants (:1208-1230), thresholds (:1268-1282),
// bank and work-unit readers (:507-535, :655-842), toplist (:1310-1397), final statistics and candidate
// file (:1502-1686), checkpoint (:1743-1783).  No GPU code here; compiled as C++ like the reference
// (its .c files are built with g++, Makefile:80-111), which matters for sin(float) below.
#include <errno.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

#include <algorithm>
#include <vector>

#include "../../include/brp_b200.h"

#define BRP_FN_LENGTH 256 /* structs.h:32 */
#define BRP_MICROSEC 1.0e-6 /* structs.h:34 */

extern "C" {

int brp_derive_geometry(uint32_t nsamples_unpadded, double tsample_us, float padding, float f0, uint32_t window,
                        brp_geometry *g)
{
    if (!g || nsamples_unpadded < 2) return RADPUL_EVAL;
    if (padding < 1.0f || padding > 10.0f) return RADPUL_EVAL; /* demod_binary.c:229-238 */
    uint32_t nsamples = nsamples_unpadded;
    nsamples = (int)(padding * nsamples + 0.5); /* float product, double add: demod_binary.c:782 */
    g->nsamples_unpadded = nsamples_unpadded;
    g->nsamples = nsamples;
    g->t_obs = nsamples * tsample_us * BRP_MICROSEC; /* of the PADDED series: the header count was overwritten at :782 */
    g->dt = tsample_us * BRP_MICROSEC;
    g->step_inv = 1.0 / g->dt;
    g->fft_size = (unsigned int)(nsamples * 0.5 + 0.5) + 1;
    g->window_2 = (unsigned int)(window * 0.5 + 0.5);
    if (g->fft_size < window) return RADPUL_EVAL; /* demod_binary.c:1102-1106 */
    const int cap = (int)(g->fft_size - g->window_2);
    const int fund = (int)(f0 * g->t_obs + 0.5);         /* float product */
    const int harm = (int)(16.0 * f0 * g->t_obs + 0.5);  /* double product */
    g->fundamental_idx_hi = (unsigned int)std::min(cap, fund);
    g->harmonic_idx_hi = (unsigned int)std::min(cap, harm);
    return 0;
}

int brp_derive_template(double P_in, double tau_in, double Psi0_in, float step_inv, brp_template *t)
{
    if (!t) return RADPUL_EVAL;
    const float P = (float)P_in, tau = (float)tau_in, Psi0 = (float)Psi0_in; /* demod_binary.c:1208-1210 */
    t->P = P;
    t->tau = tau;
    t->Psi0 = Psi0;
    t->Omega = (float)(2.0 * M_PI / P); /* :1218, double division */
    /* :1230 "tau * sin(Psi0) * step_inv": under g++ sin(float) is the float overload, so this is all-float */
    t->S0 = tau * sinf(Psi0) * step_inv;
    return 0;
}

/* Upper tail of chi^2 with 2n degrees of freedom: Q(x) = exp(-x/2) sum_{k<n} (x/2)^k / k!.
 * Stands in for gsl_cdf_chisq_Q / gsl_cdf_chisq_Qinv (demod_binary.c:1281,1517-1545); GSL is a third-party
 * dependency that is not part of the reference tree, the closed form is exact for even d.o.f. */
static double log_chisq_Q(double x, int n)
{
    const double h = 0.5 * x;
    double term = 1.0, sum = 1.0;
    for (int k = 1; k < n; k++) {
        term *= h / k;
        sum += term;
    }
    return -h + log(sum);
}

double brp_chisq_Q(double x, int n)
{
    if (x <= 0.0) return 1.0;
    const double h = 0.5 * x;
    double term = 1.0, sum = 1.0;
    for (int k = 1; k < n; k++) {
        term *= h / k;
        sum += term;
    }
    return exp(-h) * sum;
}

double brp_chisq_Qinv(double Q, int n)
{
    if (Q >= 1.0) return 0.0;
    if (Q <= 0.0) return INFINITY;
    const double target = log(Q);
    double lo = 0.0, hi = 2.0 * n + 10.0;
    while (log_chisq_Q(hi, n) > target) hi *= 2.0;
    for (int it = 0; it < 200 && hi - lo > 4e-16 * hi; it++) {
        const double mid = 0.5 * (lo + hi);
        (log_chisq_Q(mid, n) > target ? lo : hi) = mid;
    }
    return 0.5 * (lo + hi);
}

int brp_base_thresholds(float fA, uint32_t fft_size, float thr[5])
{
    if (!thr || fft_size == 0) return RADPUL_EVAL;
    for (unsigned int k = 0; k < 5; k++) {
        const unsigned int N_h = (int)(pow(2.0, k) + 0.5);
        const float prob = 1.0 - pow(1.0 - fA, 1.0 / fft_size); /* demod_binary.c:1274 */
        thr[k] = (float)(0.5 * brp_chisq_Qinv(prob, (int)N_h));  /* :1281 without the fmaxf(toplist) term */
    }
    return 0;
}

int brp_bank_read(const char *path, double **bank, uint32_t *n_templates)
{
    if (!path || !bank || !n_templates) return RADPUL_EVAL;
    FILE *f = fopen(path, "r");
    if (!f) return RADPUL_EIO;
    std::vector<double> v;
    char line[BRP_FN_LENGTH];
    /* same termination rule as demod_binary.c:1180 / :514: a last line without '\n' sets EOF and is dropped */
    while (fgets(line, BRP_FN_LENGTH, f) != NULL && !feof(f)) {
        double P, tau, psi;
        if (sscanf(line, "%lg %lg %lg\n", &P, &tau, &psi) != 3) {
            fclose(f);
            return RADPUL_EVAL; /* :521-523 */
        }
        v.push_back(P);
        v.push_back(tau);
        v.push_back(psi);
    }
    const bool bad = ferror(f);
    fclose(f);
    if (bad) return RADPUL_EIO;
    double *out = (double *)malloc(sizeof(double) * (v.empty() ? 1 : v.size()));
    if (!out) return RADPUL_EMEM;
    memcpy(out, v.data(), sizeof(double) * v.size());
    *bank = out;
    *n_templates = (uint32_t)(v.size() / 3);
    return 0;
}

void brp_bank_free(double *bank) { free(bank); }

/* packed header of a dedispersed time series, structs.h:74-105 */
struct __attribute__((__packed__)) brp_dd_header {
    double tsample, tobs, timestamp, fcenter, fchan, RA, DEC, gal_l, gal_b, AZstart, ZAstart, ASTstart, LSTstart, DM, scale;
    uint32_t filesize, datasize, nsamples;
    uint16_t smprec, nchan, nifs, lagformat, sum, level;
    char name[BRP_FN_LENGTH], originalfile[BRP_FN_LENGTH], proj_id[BRP_FN_LENGTH], observers[BRP_FN_LENGTH];
};

int brp_wu_read(const char *path, float **series, uint32_t *nsamples, double *tsample_us)
{
    if (!path || !series || !nsamples) return RADPUL_EVAL;
    int four_bit = 1; /* demod_binary.c:318-326: the extension selects the sample format */
    if (strstr(path, ".binary")) four_bit = 0;
    else if (!strstr(path, ".bin4")) return RADPUL_EFILE;
    gzFile in = gzopen(path, "rb");
    if (!in) return RADPUL_EIO;
    brp_dd_header h;
    if (gzread(in, &h, sizeof(h)) != (int)sizeof(h)) {
        gzclose(in);
        return RADPUL_EIO;
    }
    const uint32_t n = h.nsamples;
    const uint32_t nbytes = four_bit ? (uint32_t)(n * 0.5) : n; /* :779 */
    unsigned char *raw = (unsigned char *)calloc(nbytes ? nbytes : 1, 1);
    float *out = (float *)calloc(n ? n : 1, sizeof(float));
    if (!raw || !out) {
        free(raw);
        free(out);
        gzclose(in);
        return RADPUL_EMEM;
    }
    if (gzread(in, raw, nbytes) != (int)nbytes) {
        free(raw);
        free(out);
        gzclose(in);
        return RADPUL_EIO;
    }
    gzclose(in);
    for (uint32_t i = 0; i < nbytes; i++) {
        if (four_bit) { /* high nibble is the even sample, :835-836 */
            out[2 * i + 1] = (float)(raw[i] % 16) / h.scale;
            out[2 * i] = (float)(raw[i] >> 4) / h.scale;
        } else {
            out[i] = ((signed char *)raw)[i] / h.scale; /* :840 */
        }
    }
    free(raw);
    *series = out;
    *nsamples = n;
    if (tsample_us) *tsample_us = h.tsample;
    return 0;
}

void brp_series_free(float *series) { free(series); }

int brp_toplist_from_hits(const brp_hit *hits, uint32_t n_hits, const double *bank, uint32_t n_templates,
                          brp_cand cands_out[BRP_N_CAND], int32_t *tmpl_idx_out)
{
    if ((!hits && n_hits) || !cands_out) return RADPUL_EVAL;
    memset(cands_out, 0, sizeof(brp_cand) * BRP_N_CAND);
    if (tmpl_idx_out)
        for (int i = 0; i < BRP_N_CAND; i++) tmpl_idx_out[i] = -1;
    for (uint32_t k = 0; k < 5; k++) {
        std::vector<brp_hit> v;
        for (uint32_t i = 0; i < n_hits; i++)
            if (hits[i].harm_idx == k) v.push_back(hits[i]);
        /* one entry per bin: strongest power, earliest template on a tie (the strict '<' of demod_binary.c:1360);
         * needed when hit lists of several shards are concatenated */
        std::sort(v.begin(), v.end(), [](const brp_hit &a, const brp_hit &b) {
            if (a.bin != b.bin) return a.bin < b.bin;
            if (a.power != b.power) return a.power > b.power;
            return a.tmpl < b.tmpl;
        });
        v.erase(std::unique(v.begin(), v.end(), [](const brp_hit &a, const brp_hit &b) { return a.bin == b.bin; }), v.end());
        /* the 100 strongest bins, descending power (the order qsort leaves behind, :1377,1699-1710) */
        std::sort(v.begin(), v.end(), [](const brp_hit &a, const brp_hit &b) {
            if (a.power != b.power) return a.power > b.power;
            if (a.tmpl != b.tmpl) return a.tmpl < b.tmpl;
            return a.bin < b.bin;
        });
        const size_t keep = std::min<size_t>(v.size(), BRP_N_CAND_5);
        for (size_t i = 0; i < keep; i++) {
            brp_cand &c = cands_out[k * BRP_N_CAND_5 + i];
            if (v[i].tmpl >= n_templates) return RADPUL_EVAL;
            c.power = v[i].power;
            c.f0 = v[i].bin;
            c.n_harm = 1u << k;
            /* the driver stores the float-rounded template values (:1368-1370) */
            c.P_b = (float)bank[3 * (size_t)v[i].tmpl];
            c.tau = (float)bank[3 * (size_t)v[i].tmpl + 1];
            c.Psi = (float)bank[3 * (size_t)v[i].tmpl + 2];
            if (tmpl_idx_out) tmpl_idx_out[k * BRP_N_CAND_5 + i] = (int32_t)v[i].tmpl;
        }
    }
    return 0;
}

/* Merge two candidate arrays in the "right after the loop" state (e.g. a checkpoint and the hits of the templates searched
 * since): per harmonic, one entry per bin (the stronger; `a` wins ties, it is the earlier part of the bank), the 100
 * strongest kept, descending power.  Exact for resuming: a bin that missed the first part's top 100 cannot re-enter
 * through a weaker later value. */
int brp_toplist_merge(const brp_cand a[BRP_N_CAND], const brp_cand b[BRP_N_CAND], brp_cand out[BRP_N_CAND])
{
    if (!a || !b || !out) return RADPUL_EVAL;
    std::vector<brp_cand> res(BRP_N_CAND);
    memset(res.data(), 0, sizeof(brp_cand) * BRP_N_CAND);
    for (int k = 0; k < 5; k++) {
        std::vector<brp_cand> v;
        for (int i = 0; i < BRP_N_CAND_5; i++)
            if (a[k * BRP_N_CAND_5 + i].power > 0.0) v.push_back(a[k * BRP_N_CAND_5 + i]);
        const size_t na = v.size();
        for (int i = 0; i < BRP_N_CAND_5; i++) {
            const brp_cand &c = b[k * BRP_N_CAND_5 + i];
            if (!(c.power > 0.0)) continue;
            bool dup = false;
            for (size_t j = 0; j < na && !dup; j++)
                if (v[j].f0 == c.f0) {
                    if (c.power > v[j].power) v[j] = c;
                    dup = true;
                }
            if (!dup) v.push_back(c);
        }
        std::stable_sort(v.begin(), v.end(), [](const brp_cand &x, const brp_cand &y) { return x.power > y.power; });
        for (size_t i = 0; i < v.size() && i < BRP_N_CAND_5; i++) res[k * BRP_N_CAND_5 + i] = v[i];
    }
    memcpy(out, res.data(), sizeof(brp_cand) * BRP_N_CAND);
    return 0;
}

namespace {
struct Tagged {
    brp_cand c;
    int32_t t;
};
/* ordering of demod_binary.c:1713-1740: descending fA, then power, then bin */
bool ifa_before(const brp_cand &x, const brp_cand &y)
{
    if (x.fA != y.fA) return x.fA > y.fA;
    if (x.power != y.power) return x.power > y.power;
    return x.f0 > y.f0;
}
} // namespace

int brp_finalize_candidates(brp_cand cands[BRP_N_CAND], int32_t *tmpl_idx_io)
{
    if (!cands) return RADPUL_EVAL;
    std::vector<Tagged> v(BRP_N_CAND);
    for (int i = 0; i < BRP_N_CAND; i++) {
        brp_cand c = cands[i];
        /* 2*power ~ chi^2 with 2n d.o.f.; sigma_n = sqrt(n) held in a float like the reference (:1508-1512) */
        float sigma = 0.0f;
        switch (c.n_harm) {
            case 1: sigma = 1.0; break;
            case 2: sigma = sqrt(2.0); break;
            case 4: sigma = 2.0; break;
            case 8: sigma = sqrt(8.0); break;
            case 16: sigma = 4.0; break;
            default: break;
        }
        if (sigma > 0.0f) {
            const double fA = brp_chisq_Q(2.0 * c.power, (int)c.n_harm);
            c.fA = (fA > 0.0) ? -log10(fA) : 320.0;
            c.power /= sigma;
        } else {
            c.fA = -10.0; /* empty slot, :1551-1552 */
        }
        v[i].c = c;
        v[i].t = tmpl_idx_io ? tmpl_idx_io[i] : -1;
    }
    std::stable_sort(v.begin(), v.end(), [](const Tagged &a, const Tagged &b) { return ifa_before(a.c, b.c); });
    for (int i = 0; i < BRP_N_CAND; i++) {
        cands[i] = v[i].c;
        if (tmpl_idx_io) tmpl_idx_io[i] = v[i].t;
    }
    return 0;
}

int brp_format_candidates(brp_cand cands[BRP_N_CAND], const brp_geometry *g, double tsample_us, char *buf, int buflen)
{
    if (!cands || !g || !buf) return -1;
    int off = 0, emitted = 0;
    const double t_obs = g->nsamples * tsample_us * BRP_MICROSEC; /* :1637, double this time */
    const double res_factor = 1.0 / t_obs;
    std::vector<brp_cand> v(cands, cands + BRP_N_CAND);
    /* strongest remaining candidate first; every other entry at the same bin is then dismissed, which
     * maximises over the number of harmonics (:1632-1664) */
    while (emitted < BRP_N_CAND_5 && v[0].fA > 0.0) {
        const brp_cand &c = v[0];
        const int n = snprintf(buf + off, buflen - off, "%6.12f %6.12f %6.12f %6.12f %g %g %d\n", c.f0 * res_factor, c.P_b, c.tau,
                               c.Psi, c.power, c.fA, c.n_harm);
        if (n < 0 || n >= buflen - off) return -1;
        off += n;
        emitted++;
        const uint32_t bin = c.f0;
        for (brp_cand &o : v)
            if (o.f0 == bin) o.fA = -10.0;
        std::stable_sort(v.begin(), v.end(), ifa_before);
    }
    const int n = snprintf(buf + off, buflen - off, "%%DONE%%\n");
    if (n < 0 || n >= buflen - off) return -1;
    memcpy(cands, v.data(), sizeof(brp_cand) * BRP_N_CAND);
    return off + n;
}

int brp_write_candidates(const char *path, brp_cand cands[BRP_N_CAND], const brp_geometry *g, double tsample_us)
{
    if (!path) return RADPUL_EVAL;
    std::vector<char> buf(1 << 16);
    const int n = brp_format_candidates(cands, g, tsample_us, buf.data(), (int)buf.size());
    if (n < 0) return RADPUL_EIO;
    char tmp[BRP_FN_LENGTH + 8];
    if (snprintf(tmp, sizeof(tmp), "%s.tmp", path) >= (int)sizeof(tmp)) return RADPUL_EFILE;
    FILE *f = fopen(tmp, "w");
    if (!f) return RADPUL_EIO;
    const bool ok = fwrite(buf.data(), 1, (size_t)n, f) == (size_t)n;
    if (fclose(f) != 0 || !ok) return RADPUL_EIO;
    if (rename(tmp, path) != 0) return RADPUL_EFILE; /* atomic replace, :1681 */
    return 0;
}

/* checkpoint = CP_Header {uint32 n_template; char originalfile[256]} + 500 CP_cand (structs.h:111-132) */
int brp_checkpoint_write(const char *path, uint32_t n_template, const char *originalfile, const brp_cand cands[BRP_N_CAND])
{
    if (!path || !cands) return RADPUL_EVAL;
    char tmp[BRP_FN_LENGTH + 8];
    if (snprintf(tmp, sizeof(tmp), "%s.tmp", path) >= (int)sizeof(tmp)) return RADPUL_EFILE;
    FILE *f = fopen(tmp, "wb");
    if (!f) return RADPUL_EIO;
    char name[BRP_FN_LENGTH];
    memset(name, 0, sizeof(name));
    if (originalfile) strncpy(name, originalfile, sizeof(name) - 1);
    bool ok = fwrite(&n_template, sizeof(uint32_t), 1, f) == 1;
    ok = ok && fwrite(name, 1, sizeof(name), f) == sizeof(name);
    ok = ok && fwrite(cands, sizeof(brp_cand), BRP_N_CAND, f) == BRP_N_CAND;
    if (fclose(f) != 0 || !ok) return RADPUL_EIO;
    if (rename(tmp, path) != 0) return RADPUL_EFILE;
    return 0;
}

int brp_checkpoint_read(const char *path, uint32_t *n_template, char originalfile[256], brp_cand cands[BRP_N_CAND])
{
    if (!path || !n_template || !cands) return RADPUL_EVAL;
    FILE *f = fopen(path, "rb");
    if (!f) return RADPUL_EIO;
    char name[BRP_FN_LENGTH];
    bool ok = fread(n_template, sizeof(uint32_t), 1, f) == 1;
    ok = ok && fread(name, 1, sizeof(name), f) == sizeof(name);
    ok = ok && fread(cands, sizeof(brp_cand), BRP_N_CAND, f) == BRP_N_CAND;
    fclose(f);
    if (!ok) return RADPUL_EIO;
    if (originalfile) memcpy(originalfile, name, sizeof(name));
    return 0;
}

} /* extern "C" */
