// Internal interfaces between the translation units of libbrp_b200.so (not part of the C ABI).
#ifndef BRP_INTERNAL_H
#define BRP_INTERNAL_H

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/brp_b200.h"

// Per-template constants as the kernels consume them (16 bytes, one float4 load).
struct BrpTemplateDev {
    float tau, Omega, Psi0, S0;
};

// Device-side view of the work-unit geometry.
struct BrpGeomDev {
    uint32_t nu;        // unpadded samples
    uint32_t n;         // padded samples = FFT length
    uint32_t fft_size;  // n/2 + 1
    uint32_t w2, fu, h; // window_2, fundamental_idx_hi, harmonic_idx_hi
    float dt, step_inv;
};

// ---------------- resampling (resamp.cu) ----------------
static const int BRP_RESAMP_THREADS = 256;
static const int BRP_RESAMP_SAMPLES_PER_BLOCK = 4096;

inline uint32_t brp_resamp_blocks(uint32_t nu)
{
    return (nu + BRP_RESAMP_SAMPLES_PER_BLOCK - 1) / BRP_RESAMP_SAMPLES_PER_BLOCK;
}

// out: batch x out_stride floats; [0, n_steps) of each row is written (and [n_steps, n) with the mean when
// fill_padding).  idx_out (batch x nu int32) optional.  partials: batch x brp_resamp_blocks(nu) doubles;
// counters: batch uint32, must be zero on entry and are left zero on exit.
cudaError_t brp_launch_count_mismatch(const int32_t *a, const int32_t *b, uint32_t n, unsigned long long *count, cudaStream_t st);
cudaError_t brp_launch_resample(const float *series, const BrpTemplateDev *tmpl, int batch, const BrpGeomDev &g,
                                float *out, size_t out_stride, int32_t *idx_out, uint32_t *n_steps, float *mean,
                                double *partials, uint32_t *counters, bool fill_padding, cudaStream_t st,
                                bool every_sample = false /* skip the run detection: timing model at every sample */,
                                const BrpTemplateDev *tmpl_value = nullptr /* tmpl == NULL, batch 1: the template itself (host memory) */);
cudaError_t brp_launch_sin_lut(const float *x, float *out, uint32_t n, cudaStream_t st);
cudaError_t brp_launch_round_index(const float *f, int32_t *out, uint32_t n, cudaStream_t st);

// ---------------- FFT + power spectrum (fft.cu) ----------------
// tensor maps of the last launch, kept per plan (= per handle): a search launches every batch on the same buffers, so the three
// cuTensorMapEncodeTiled calls happen once per handle (and once more for a shorter last batch), not once per batch
struct BrpFftMapCache {
    const void *in, *work, *out;
    size_t in_stride, out_stride;
    uint32_t bound, count;
    int batch1, batch2, parts;
    bool valid1, valid2, ok2;
    unsigned char in_map[128], out_map[128], ps_map[128]; // copied with memcpy (the handle comes from calloc: no 64-byte alignment)
};

struct BrpFftPlan {
    uint32_t n;      // real length
    uint32_t m;      // complex length n/2
    uint32_t m1, m2; // m = m1*m2 : pass 1 transforms length m2 (stride m1), pass 2 length m1
    int radix3;      // 1 when m2 carries the factor 3
    float2 *tw_pass1;   // in-tile twiddles for length m2
    float2 *tw_pass2;   // in-tile twiddles for length m1
    float2 *tw_untangle_a; // W_n^{k2},      k2 < m2
    float2 *tw_untangle_b; // W_{2 m1}^{k1}, k1 < m1
    float2 *tw_inter_hi, *tw_inter_lo; // W_m^{p}: p = hi*1024 + lo
    // pass 1 as `parts` sub-transforms of 2048 rows (m2 = parts*2048, upper rows all padding): W_{m2}^{q p} for q = 1..parts-1,
    // p < 2048, and the second inter-pass factor for a last radix of 8, exp(-2 pi i col q / (8 m1))
    int parts;
    float2 *tw_parts, *tw_inter_parts;
    int smem_pass1, smem_pass2;
    int launches_per_batch;
    mutable BrpFftMapCache maps;
};

// nu_hint: number of samples that can be valid (Nu) if the caller knows it, else 0; it may move the split of m so that
// pass 1 can run as sub-transforms of 2048 rows (see fft.cu)
int brp_fft_plan_create(BrpFftPlan *p, uint32_t n, uint32_t nu_hint = 0);
void brp_fft_plan_destroy(BrpFftPlan *p);
// in: batch x in_stride floats (only [0, n_steps[b]) valid; the rest is taken as mean[b]);
// work: batch x m float2; ps: batch x ps_stride floats, bins [0, ps_count) written, ps[0] = 0.
cudaError_t brp_launch_fft_power(const BrpFftPlan &p, const float *in, size_t in_stride, const uint32_t *n_steps,
                                 const float *mean, float2 *work, float *ps, size_t ps_stride, uint32_t ps_count,
                                 float norm, int batch, cudaStream_t st, cudaEvent_t mid_event,
                                 uint32_t n_steps_bound = 0 /* upper bound of every n_steps[b] if the caller knows one (e.g. Nu) */);

cudaError_t brp_launch_fft_spectrum(const BrpFftPlan &p, const float *in, const uint32_t *n_steps, const float *mean,
                                    float2 *work, float2 *spectrum, uint32_t count, cudaStream_t st);
cudaError_t brp_launch_fft_conj(const BrpFftPlan &p, const float2 *in, const uint32_t *n_steps_full, const float *mean,
                                float2 *work, float *out, uint32_t pairs, float scale, cudaStream_t st);

// ---------------- harmonic summing + per-bin toplist (hs.cu) ----------------
struct BrpThresholds {
    float t[5];
};
// keys: 5 x fu uint64, key = float_bits(power) << 32 | (0xFFFFFFFF - template index); atomicMax semantics
cudaError_t brp_launch_harmonic_summing(const float *ps, size_t ps_stride, int batch, const BrpGeomDev &g,
                                        BrpThresholds thr, uint32_t first_template, unsigned long long *keys,
                                        cudaStream_t st);
// compact non-zero keys with bin >= w2 into hits (device), counter (device uint32, zeroed by caller)
cudaError_t brp_launch_collect(const unsigned long long *keys, uint32_t fu, uint32_t w2, brp_hit *hits,
                               uint32_t max_hits, uint32_t *counter, cudaStream_t st);

// One-template-per-call callers (the reference's backend interface): the same scan, but every non-empty entry is zeroed behind
// it (the table is empty again afterwards) and the hits go STRAIGHT to the caller's mapped pinned host memory, `hits` being its
// device-side address, with their number in *host_count (mapped pinned as well) -- no device-to-host copy to enqueue and await.
// ctl: three device words, all zero before the first call: slot counters of even and odd calls (`which` = 0 / 1; each call zeroes the
// other one for the next call, which therefore needs no memset) and the arrival count of the blocks.
cudaError_t brp_launch_collect_to_host(unsigned long long *keys, uint32_t fu, uint32_t w2, brp_hit *hits, uint32_t max_hits,
                                       uint32_t *ctl, int which, uint32_t *host_count, cudaStream_t st);

#endif
