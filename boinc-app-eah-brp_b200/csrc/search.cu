// Batched search handle (brp_search_*) and the per-stage taps used by the parity tests.
//
// One handle = one work-unit on one GPU.  All device memory is allocated at create time and laid out
// for a batch of `max_batch` templates in flight (batching does not fit the legacy
// one-template interface, so it is additive):
//
//   series     Nu floats                     shared by every template (stays L2-resident: 16.8 MB at 2^22)
//   templates  n x {tau, Omega, Psi0, S0}    whole bank, one H2D copy per brp_search_run
//   resampled  max_batch x Nu floats         rows hold [0, n_steps); the mean padding is never stored
//   n_steps, mean, partial sums, counters    per template of the batch
//   work       max_batch x M float2          FFT intermediate Y[k2][m1]
//   spectrum   max_batch x H' floats         only bins < H are ever produced
//   keys       5 x Fu uint64                 per-bin toplist table, lives until brp_search_collect
//
// A batch is four kernel launches on one stream (resample, FFT pass 1, FFT pass 2 + power, harmonic sum);
// nothing is copied back and the host never synchronises inside brp_search_run: the template constants are derived into
// one of two pinned staging buffers (an event per buffer says when its last upload has left it), so the upload of one
// call overlaps the kernels of the previous one.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "brp_internal.h"

#define CU_TRY(call, code)                                                                           \
    do {                                                                                             \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) {                                                                     \
            fprintf(stderr, "[brp_b200] %s failed: %s (%s:%d)\n", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return (code);                                                                           \
        }                                                                                            \
    } while (0)

// One lane = one stream with its own per-batch scratch.  Batches are dealt round-robin to the lanes so that kernels of
// different batches overlap on the GPU: the stages are bound by different resources (resampler: issue slots, FFT
// passes: L1/LSU wavefronts, harmonic sum: L2), so co-resident CTAs of different stages fill each other's idle units.
struct BrpLane {
    cudaStream_t st;
    float *resamp;
    uint32_t *nsteps;
    float *mean;
    double *partials;
    uint32_t *counters;
    float2 *work;
    float *ps;
    cudaEvent_t done;
};
#define BRP_MAX_LANES 4

struct brp_search {
    int device;
    cudaStream_t stream;
    BrpLane lane[BRP_MAX_LANES]; // lane[0] aliases the handle's own stream and scratch
    int n_lanes;
    cudaEvent_t ev_fork;
    brp_geometry g;
    BrpGeomDev gd;
    BrpThresholds thr;
    uint32_t max_batch;
    size_t rs_stride, ps_stride;
    uint32_t nblk;
    BrpFftPlan plan;
    float *d_series;
    BrpTemplateDev *d_tmpl, *h_tmpl[2]; // two pinned staging buffers, used alternately
    cudaEvent_t ev_tmpl[2];             // recorded behind the upload that read h_tmpl[k]
    int tmpl_slot;
    uint32_t tmpl_cap;
    float *d_resamp;
    uint32_t *d_nsteps;
    float *d_mean;
    double *d_partials;
    uint32_t *d_counters;
    float2 *d_work;
    float *d_ps;
    unsigned long long *d_keys;
    brp_hit *d_hits;
    uint32_t *d_hit_counter;
    uint32_t hit_cap;
    // lazily allocated buffers of the stage taps
    float *d_tap_series; // N floats
    int32_t *d_tap_idx;  // rs_stride int32
    float *d_tap_ps;     // F floats
    unsigned long long *d_tap_keys;
    cudaEvent_t ev_start, ev_stop;
    bool timed;
    uint64_t launches;
};

static int set_device(int device)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) {
        fprintf(stderr, "[brp_b200] no CUDA device available: this library has no CPU fallback\n");
        return RADPUL_CUDA_DEVICE_FIND;
    }
    if (device < 0 || device >= count) return RADPUL_CUDA_DEVICE_FIND;
    CU_TRY(cudaSetDevice(device), RADPUL_CUDA_DEVICE_SET);
    return 0;
}

extern "C" {

int brp_search_destroy(brp_search *s)
{
    if (!s) return 0;
    cudaSetDevice(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    brp_fft_plan_destroy(&s->plan);
    cudaFree(s->d_series); cudaFree(s->d_tmpl); cudaFreeHost(s->h_tmpl[0]); cudaFreeHost(s->h_tmpl[1]); cudaFree(s->d_resamp); cudaFree(s->d_nsteps);
    for (int k = 0; k < 2; k++)
        if (s->ev_tmpl[k]) cudaEventDestroy(s->ev_tmpl[k]);
    cudaFree(s->d_mean); cudaFree(s->d_partials); cudaFree(s->d_counters); cudaFree(s->d_work); cudaFree(s->d_ps);
    cudaFree(s->d_keys); cudaFree(s->d_hits); cudaFree(s->d_hit_counter); cudaFree(s->d_tap_series);
    cudaFree(s->d_tap_idx); cudaFree(s->d_tap_ps); cudaFree(s->d_tap_keys);
    for (int l = 1; l < s->n_lanes; l++) {
        BrpLane &ln = s->lane[l];
        if (ln.st) cudaStreamSynchronize(ln.st);
        cudaFree(ln.resamp); cudaFree(ln.nsteps); cudaFree(ln.mean); cudaFree(ln.partials); cudaFree(ln.counters);
        cudaFree(ln.work); cudaFree(ln.ps);
        if (ln.done) cudaEventDestroy(ln.done);
        if (ln.st) cudaStreamDestroy(ln.st);
    }
    if (s->lane[0].done) cudaEventDestroy(s->lane[0].done);
    if (s->ev_fork) cudaEventDestroy(s->ev_fork);
    if (s->ev_start) cudaEventDestroy(s->ev_start);
    if (s->ev_stop) cudaEventDestroy(s->ev_stop);
    if (s->stream) cudaStreamDestroy(s->stream);
    free(s);
    return 0;
}

int brp_search_create(brp_search **out, int device, const brp_geometry *g, float fA, uint32_t max_batch)
{
    if (!out || !g || max_batch == 0) return RADPUL_EVAL;
    *out = NULL;
    if (g->nsamples_unpadded < 2 || g->nsamples < g->nsamples_unpadded || g->harmonic_idx_hi > g->fft_size ||
        g->fundamental_idx_hi > g->fft_size || g->nsamples_unpadded > (1u << 24))
        return RADPUL_EVAL;
    int rc = set_device(device);
    if (rc) return rc;
    brp_search *s = (brp_search *)calloc(1, sizeof(brp_search));
    if (!s) return RADPUL_EMEM;
    s->device = device;
    s->g = *g;
    s->gd.nu = g->nsamples_unpadded; s->gd.n = g->nsamples; s->gd.fft_size = g->fft_size;
    s->gd.w2 = g->window_2; s->gd.fu = g->fundamental_idx_hi; s->gd.h = g->harmonic_idx_hi;
    s->gd.dt = g->dt; s->gd.step_inv = g->step_inv;
    s->max_batch = max_batch;
    rc = brp_base_thresholds(fA, g->fft_size, s->thr.t);
    if (rc) { free(s); return rc; }
    rc = brp_fft_plan_create(&s->plan, g->nsamples, g->nsamples_unpadded);
    if (rc) { free(s); return rc; }
    s->rs_stride = ((size_t)g->nsamples_unpadded + 3) & ~(size_t)3;
    // the harmonic-summing bulk copies round slices up to 16 bytes; the tensor stores of FFT pass 2 write the last row of
    // the stored part of the spectrum whole (a row is at most 6144 bins)
    s->ps_stride = (((size_t)g->harmonic_idx_hi + 6144 + 4) + 31) & ~(size_t)31;
    s->nblk = brp_resamp_blocks(g->nsamples_unpadded);
    s->hit_cap = 5u * (g->fundamental_idx_hi ? g->fundamental_idx_hi : 1u);
    const size_t B = max_batch;
#define ALLOC(ptr, bytes)                                              \
    if (cudaMalloc((void **)&(ptr), (bytes)) != cudaSuccess) {         \
        fprintf(stderr, "[brp_b200] cudaMalloc(%zu) failed\n", (size_t)(bytes)); \
        brp_search_destroy(s);                                         \
        return RADPUL_CUDA_MEM_ALLOC_DEVICE;                           \
    }
    ALLOC(s->d_series, sizeof(float) * s->rs_stride);
    ALLOC(s->d_resamp, sizeof(float) * s->rs_stride * B);
    ALLOC(s->d_nsteps, sizeof(uint32_t) * B);
    ALLOC(s->d_mean, sizeof(float) * B);
    ALLOC(s->d_partials, sizeof(double) * s->nblk * B);
    ALLOC(s->d_counters, sizeof(uint32_t) * B);
    ALLOC(s->d_work, sizeof(float2) * (size_t)s->plan.m * B);
    ALLOC(s->d_ps, sizeof(float) * s->ps_stride * B);
    ALLOC(s->d_keys, sizeof(unsigned long long) * s->hit_cap);
    ALLOC(s->d_hits, sizeof(brp_hit) * s->hit_cap);
    ALLOC(s->d_hit_counter, sizeof(uint32_t));
#undef ALLOC
    if (cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&s->ev_start) != cudaSuccess || cudaEventCreate(&s->ev_stop) != cudaSuccess) {
        brp_search_destroy(s);
        return RADPUL_CUDA_DEVICE_SET;
    }
    cudaMemsetAsync(s->d_counters, 0, sizeof(uint32_t) * B, s->stream);
    cudaMemsetAsync(s->d_keys, 0, sizeof(unsigned long long) * s->hit_cap, s->stream);
    cudaMemsetAsync(s->d_ps, 0, sizeof(float) * s->ps_stride * B, s->stream);
    // lanes: BRP_STREAMS (default 1) streams with private scratch; lane 0 is the handle's own
    {
        const char *env = getenv("BRP_STREAMS");
        int want = env ? atoi(env) : 1; // measured on B200: 2-4 lanes are ~10 % SLOWER than one (L2 thrash), so off by default
        if (want < 1) want = 1;
        if (want > BRP_MAX_LANES) want = BRP_MAX_LANES;
        s->n_lanes = 1;
        BrpLane &l0 = s->lane[0];
        l0.st = s->stream; l0.resamp = s->d_resamp; l0.nsteps = s->d_nsteps; l0.mean = s->d_mean; l0.partials = s->d_partials;
        l0.counters = s->d_counters; l0.work = s->d_work; l0.ps = s->d_ps;
        bool ok = cudaEventCreateWithFlags(&l0.done, cudaEventDisableTiming) == cudaSuccess &&
                  cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming) == cudaSuccess;
        for (int l = 1; ok && l < want; l++) {
            BrpLane &ln = s->lane[l];
            ok = cudaMalloc((void **)&ln.resamp, sizeof(float) * s->rs_stride * B) == cudaSuccess &&
                 cudaMalloc((void **)&ln.nsteps, sizeof(uint32_t) * B) == cudaSuccess &&
                 cudaMalloc((void **)&ln.mean, sizeof(float) * B) == cudaSuccess &&
                 cudaMalloc((void **)&ln.partials, sizeof(double) * s->nblk * B) == cudaSuccess &&
                 cudaMalloc((void **)&ln.counters, sizeof(uint32_t) * B) == cudaSuccess &&
                 cudaMalloc((void **)&ln.work, sizeof(float2) * (size_t)s->plan.m * B) == cudaSuccess &&
                 cudaMalloc((void **)&ln.ps, sizeof(float) * s->ps_stride * B) == cudaSuccess &&
                 cudaStreamCreateWithFlags(&ln.st, cudaStreamNonBlocking) == cudaSuccess &&
                 cudaEventCreateWithFlags(&ln.done, cudaEventDisableTiming) == cudaSuccess;
            s->n_lanes = l + 1; // so that destroy releases a partially built lane too
            if (ok) {
                cudaMemsetAsync(ln.counters, 0, sizeof(uint32_t) * B, s->stream);
                cudaMemsetAsync(ln.ps, 0, sizeof(float) * s->ps_stride * B, s->stream);
            }
        }
        if (!ok) {
            fprintf(stderr, "[brp_b200] could not set up %d lanes\n", want);
            brp_search_destroy(s);
            return RADPUL_CUDA_MEM_ALLOC_DEVICE;
        }
    }
    CU_TRY(cudaStreamSynchronize(s->stream), RADPUL_CUDA_KERNEL_INVOKE);
    *out = s;
    return 0;
}

int brp_search_set_series(brp_search *s, const float *host_series)
{
    if (!s || !host_series) return RADPUL_EVAL;
    CU_TRY(cudaSetDevice(s->device), RADPUL_CUDA_DEVICE_SET);
    CU_TRY(cudaMemcpyAsync(s->d_series, host_series, sizeof(float) * s->g.nsamples_unpadded, cudaMemcpyHostToDevice, s->stream),
           RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
    return 0;
}

int brp_search_set_series_device(brp_search *s, const void *device_series)
{
    if (!s || !device_series) return RADPUL_EVAL;
    CU_TRY(cudaSetDevice(s->device), RADPUL_CUDA_DEVICE_SET);
    CU_TRY(cudaMemcpyAsync(s->d_series, device_series, sizeof(float) * s->g.nsamples_unpadded, cudaMemcpyDeviceToDevice, s->stream),
           RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
    return 0;
}

int brp_search_reset(brp_search *s)
{
    if (!s) return RADPUL_EVAL;
    CU_TRY(cudaSetDevice(s->device), RADPUL_CUDA_DEVICE_SET);
    CU_TRY(cudaMemsetAsync(s->d_keys, 0, sizeof(unsigned long long) * s->hit_cap, s->stream), RADPUL_CUDA_KERNEL_INVOKE);
    return 0;
}

static int ensure_templates(brp_search *s, uint32_t n)
{
    if (n <= s->tmpl_cap) return 0;
    cudaStreamSynchronize(s->stream); // growing the buffers: rare (the first call, or a larger chunk than ever before)
    cudaFree(s->d_tmpl);
    cudaFreeHost(s->h_tmpl[0]);
    cudaFreeHost(s->h_tmpl[1]);
    s->d_tmpl = NULL; s->h_tmpl[0] = s->h_tmpl[1] = NULL; s->tmpl_cap = 0;
    const uint32_t cap = n < 1024 ? 1024 : n;
    CU_TRY(cudaMalloc((void **)&s->d_tmpl, sizeof(BrpTemplateDev) * cap), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    for (int k = 0; k < 2; k++) {
        CU_TRY(cudaMallocHost((void **)&s->h_tmpl[k], sizeof(BrpTemplateDev) * cap), RADPUL_CUDA_MEM_ALLOC_HOST);
        if (!s->ev_tmpl[k]) CU_TRY(cudaEventCreateWithFlags(&s->ev_tmpl[k], cudaEventDisableTiming), RADPUL_CUDA_DEVICE_SET);
    }
    s->tmpl_cap = cap;
    return 0;
}

// next pinned staging buffer, free to be overwritten (its last upload has completed; with two buffers that was two calls ago)
static BrpTemplateDev *staging_acquire(brp_search *s)
{
    s->tmpl_slot ^= 1;
    cudaEventSynchronize(s->ev_tmpl[s->tmpl_slot]); // returns at once for an event that was never recorded
    return s->h_tmpl[s->tmpl_slot];
}

// enqueue the upload of the first n staged templates (device buffer: stream order protects the kernels still reading it)
static int staging_upload(brp_search *s, uint32_t n)
{
    CU_TRY(cudaMemcpyAsync(s->d_tmpl, s->h_tmpl[s->tmpl_slot], sizeof(BrpTemplateDev) * n, cudaMemcpyHostToDevice, s->stream),
           RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
    cudaEventRecord(s->ev_tmpl[s->tmpl_slot], s->stream);
    return 0;
}

// upload bank[0..n) as device templates, without waiting for the work already enqueued
static int upload_templates(brp_search *s, const double *bank, uint32_t n)
{
    if (n == 0) return 0; // an empty bank: nothing to stage
    int rc = ensure_templates(s, n);
    if (rc) return rc;
    BrpTemplateDev *h = staging_acquire(s);
    for (uint32_t i = 0; i < n; i++) {
        brp_template t;
        brp_derive_template(bank[3 * (size_t)i], bank[3 * (size_t)i + 1], bank[3 * (size_t)i + 2], s->g.step_inv, &t);
        h[i].tau = t.tau; h[i].Omega = t.Omega; h[i].Psi0 = t.Psi0; h[i].S0 = t.S0;
    }
    return staging_upload(s, n);
}

static int run_batch(brp_search *s, const BrpLane &ln, uint32_t first, uint32_t nb, uint32_t first_global,
                     unsigned long long *keys, cudaEvent_t *ev /* optional 5 events */)
{
    const float norm = 1.0f / (float)s->g.nsamples; /* demod_binary.c:1255 */
    if (ev) cudaEventRecord(ev[0], ln.st);
    CU_TRY(brp_launch_resample(s->d_series, s->d_tmpl + first, (int)nb, s->gd, ln.resamp, s->rs_stride, NULL, ln.nsteps, ln.mean,
                               ln.partials, ln.counters, false, ln.st),
           RADPUL_CUDA_KERNEL_INVOKE);
    if (ev) cudaEventRecord(ev[1], ln.st);
    CU_TRY(brp_launch_fft_power(s->plan, ln.resamp, s->rs_stride, ln.nsteps, ln.mean, ln.work, ln.ps, s->ps_stride,
                                s->g.harmonic_idx_hi, norm, (int)nb, ln.st, ev ? ev[2] : NULL, s->g.nsamples_unpadded),
           RADPUL_CUDA_FFT_EXEC);
    if (ev) cudaEventRecord(ev[3], ln.st);
    CU_TRY(brp_launch_harmonic_summing(ln.ps, s->ps_stride, (int)nb, s->gd, s->thr, first_global + first, keys, ln.st),
           RADPUL_CUDA_KERNEL_INVOKE);
    if (ev) cudaEventRecord(ev[4], ln.st);
    s->launches += 4;
    return 0;
}

int brp_search_run(brp_search *s, const double *bank, uint32_t n_templates, uint32_t first_global_index)
{
    if (!s || (!bank && n_templates)) return RADPUL_EVAL;
    CU_TRY(cudaSetDevice(s->device), RADPUL_CUDA_DEVICE_SET);
    int rc = 0;
    if (s->n_lanes > 1) {
        rc = upload_templates(s, bank, n_templates);
        if (rc) return rc;
    } else if (n_templates) {
        rc = ensure_templates(s, n_templates);
        if (rc) return rc;
    }
    cudaEventRecord(s->ev_start, s->stream);
    // fork: the other lanes start after everything already queued on the handle's stream (series, templates, reset)
    cudaEventRecord(s->ev_fork, s->stream);
    for (int l = 1; l < s->n_lanes; l++) cudaStreamWaitEvent(s->lane[l].st, s->ev_fork, 0);
    // One lane (the default): the constants of a batch are derived and uploaded right before its kernels are enqueued, so the
    // GPU starts on the first batch while the host is still deriving the rest of the bank (6 662 templates take half a
    // millisecond of host time, which matters once the bank is split over eight GPUs).
    BrpTemplateDev *h = (s->n_lanes == 1 && n_templates) ? staging_acquire(s) : NULL;
    uint32_t k = 0;
    for (uint32_t first = 0; first < n_templates; first += s->max_batch, k++) {
        const uint32_t nb = n_templates - first < s->max_batch ? n_templates - first : s->max_batch;
        if (h) {
            for (uint32_t i = first; i < first + nb; i++) {
                brp_template t;
                brp_derive_template(bank[3 * (size_t)i], bank[3 * (size_t)i + 1], bank[3 * (size_t)i + 2], s->g.step_inv, &t);
                h[i].tau = t.tau; h[i].Omega = t.Omega; h[i].Psi0 = t.Psi0; h[i].S0 = t.S0;
            }
            CU_TRY(cudaMemcpyAsync(s->d_tmpl + first, h + first, sizeof(BrpTemplateDev) * nb, cudaMemcpyHostToDevice, s->stream),
                   RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
        }
        rc = run_batch(s, s->lane[k % s->n_lanes], first, nb, first_global_index, s->d_keys, NULL);
        if (rc) return rc;
    }
    if (h) cudaEventRecord(s->ev_tmpl[s->tmpl_slot], s->stream); // the staging buffer is free again once the last upload has left it
    // join: the handle's stream continues only after every lane has drained
    for (int l = 1; l < s->n_lanes; l++) {
        cudaEventRecord(s->lane[l].done, s->lane[l].st);
        cudaStreamWaitEvent(s->stream, s->lane[l].done, 0);
    }
    cudaEventRecord(s->ev_stop, s->stream);
    s->timed = true;
    return 0;
}

int brp_search_sync(brp_search *s)
{
    if (!s) return RADPUL_EVAL;
    CU_TRY(cudaSetDevice(s->device), RADPUL_CUDA_DEVICE_SET);
    CU_TRY(cudaStreamSynchronize(s->stream), RADPUL_CUDA_KERNEL_INVOKE);
    return 0;
}

static int collect_from(brp_search *s, const unsigned long long *keys, brp_hit *hits, uint32_t max_hits, uint32_t *n_hits)
{
    CU_TRY(cudaMemsetAsync(s->d_hit_counter, 0, sizeof(uint32_t), s->stream), RADPUL_CUDA_KERNEL_INVOKE);
    CU_TRY(brp_launch_collect(keys, s->g.fundamental_idx_hi, s->g.window_2, s->d_hits, s->hit_cap, s->d_hit_counter, s->stream),
           RADPUL_CUDA_KERNEL_INVOKE);
    s->launches += 1;
    uint32_t n = 0;
    CU_TRY(cudaMemcpyAsync(&n, s->d_hit_counter, sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream), RADPUL_CUDA_MEM_COPY_DEVICE_HOST);
    CU_TRY(cudaStreamSynchronize(s->stream), RADPUL_CUDA_KERNEL_INVOKE);
    *n_hits = n;
    if (n > max_hits) return RADPUL_EMEM;
    if (n) {
        CU_TRY(cudaMemcpyAsync(hits, s->d_hits, sizeof(brp_hit) * n, cudaMemcpyDeviceToHost, s->stream), RADPUL_CUDA_MEM_COPY_DEVICE_HOST);
        CU_TRY(cudaStreamSynchronize(s->stream), RADPUL_CUDA_KERNEL_INVOKE);
    }
    return 0;
}

int brp_search_collect(brp_search *s, brp_hit *hits, uint32_t max_hits, uint32_t *n_hits)
{
    if (!s || !n_hits || (!hits && max_hits)) return RADPUL_EVAL;
    CU_TRY(cudaSetDevice(s->device), RADPUL_CUDA_DEVICE_SET);
    return collect_from(s, s->d_keys, hits, max_hits, n_hits);
}

int brp_search_last_run_ms(brp_search *s, float *total_ms)
{
    if (!s || !total_ms || !s->timed) return RADPUL_EVAL;
    CU_TRY(cudaEventSynchronize(s->ev_stop), RADPUL_CUDA_KERNEL_INVOKE);
    CU_TRY(cudaEventElapsedTime(total_ms, s->ev_start, s->ev_stop), RADPUL_CUDA_KERNEL_INVOKE);
    return 0;
}

void *brp_search_stream(brp_search *s) { return s ? (void *)s->stream : NULL; }
uint64_t brp_search_kernel_launches(const brp_search *s) { return s ? s->launches : 0; }

int brp_search_profile_batch(brp_search *s, const double *bank, uint32_t batch, uint32_t repeats, float ms[4])
{
    if (!s || !bank || !ms || batch == 0 || batch > s->max_batch || repeats == 0) return RADPUL_EVAL;
    CU_TRY(cudaSetDevice(s->device), RADPUL_CUDA_DEVICE_SET);
    int rc = upload_templates(s, bank, batch);
    if (rc) return rc;
    if (!s->d_tap_keys) {
        CU_TRY(cudaMalloc((void **)&s->d_tap_keys, sizeof(unsigned long long) * s->hit_cap), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    }
    CU_TRY(cudaMemsetAsync(s->d_tap_keys, 0, sizeof(unsigned long long) * s->hit_cap, s->stream), RADPUL_CUDA_KERNEL_INVOKE);
    cudaEvent_t ev[5];
    for (int i = 0; i < 5; i++) CU_TRY(cudaEventCreate(&ev[i]), RADPUL_CUDA_DEVICE_SET);
    double acc[4] = {0, 0, 0, 0};
    rc = run_batch(s, s->lane[0], 0, batch, 0, s->d_tap_keys, NULL); // warm-up
    for (uint32_t r = 0; r < repeats && !rc; r++) {
        rc = run_batch(s, s->lane[0], 0, batch, 0, s->d_tap_keys, ev);
        if (rc) break;
        if (cudaStreamSynchronize(s->stream) != cudaSuccess) { rc = RADPUL_CUDA_KERNEL_INVOKE; break; }
        for (int i = 0; i < 4; i++) {
            float t = 0.f;
            cudaEventElapsedTime(&t, ev[i], ev[i + 1]);
            acc[i] += t;
        }
    }
    for (int i = 0; i < 5; i++) cudaEventDestroy(ev[i]);
    for (int i = 0; i < 4; i++) ms[i] = (float)(acc[i] / repeats);
    return rc;
}

// ------------------------------------------------------------------------------------------ stage taps

int brp_stage_resample(brp_search *s, const brp_template *t, float *resampled_out, int32_t *index_out,
                       uint32_t *n_steps_out, float *mean_out)
{
    if (!s || !t) return RADPUL_EVAL;
    CU_TRY(cudaSetDevice(s->device), RADPUL_CUDA_DEVICE_SET);
    int rc = ensure_templates(s, 1);
    if (rc) return rc;
    const size_t nfull = ((size_t)s->g.nsamples + 3) & ~(size_t)3;
    if (!s->d_tap_series) CU_TRY(cudaMalloc((void **)&s->d_tap_series, sizeof(float) * nfull), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    if (!s->d_tap_idx) CU_TRY(cudaMalloc((void **)&s->d_tap_idx, sizeof(int32_t) * s->rs_stride), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    {
        BrpTemplateDev *h = staging_acquire(s);
        h[0].tau = t->tau; h[0].Omega = t->Omega; h[0].Psi0 = t->Psi0; h[0].S0 = t->S0;
        rc = staging_upload(s, 1);
        if (rc) return rc;
    }
    CU_TRY(cudaMemsetAsync(s->d_tap_idx, 0xFF, sizeof(int32_t) * s->rs_stride, s->stream), RADPUL_CUDA_KERNEL_INVOKE);
    CU_TRY(brp_launch_resample(s->d_series, s->d_tmpl, 1, s->gd, s->d_tap_series, nfull, s->d_tap_idx, s->d_nsteps, s->d_mean,
                               s->d_partials, s->d_counters, true, s->stream),
           RADPUL_CUDA_KERNEL_INVOKE);
    s->launches += 2;
    // keep slot 0 of the batch scratch consistent so brp_stage_spectrum can follow
    CU_TRY(cudaMemcpyAsync(s->d_resamp, s->d_tap_series, sizeof(float) * s->g.nsamples_unpadded, cudaMemcpyDeviceToDevice, s->stream),
           RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
    if (resampled_out)
        CU_TRY(cudaMemcpyAsync(resampled_out, s->d_tap_series, sizeof(float) * s->g.nsamples, cudaMemcpyDeviceToHost, s->stream), RADPUL_CUDA_MEM_COPY_DEVICE_HOST);
    if (index_out)
        CU_TRY(cudaMemcpyAsync(index_out, s->d_tap_idx, sizeof(int32_t) * s->g.nsamples_unpadded, cudaMemcpyDeviceToHost, s->stream), RADPUL_CUDA_MEM_COPY_DEVICE_HOST);
    if (n_steps_out)
        CU_TRY(cudaMemcpyAsync(n_steps_out, s->d_nsteps, sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream), RADPUL_CUDA_MEM_COPY_DEVICE_HOST);
    if (mean_out)
        CU_TRY(cudaMemcpyAsync(mean_out, s->d_mean, sizeof(float), cudaMemcpyDeviceToHost, s->stream), RADPUL_CUDA_MEM_COPY_DEVICE_HOST);
    CU_TRY(cudaStreamSynchronize(s->stream), RADPUL_CUDA_KERNEL_INVOKE);
    return 0;
}

int brp_stage_resample_compare(brp_search *s, const double *bank, uint32_t n_templates, uint64_t *n_mismatch)
{
    if (!s || !bank || !n_mismatch) return RADPUL_EVAL;
    CU_TRY(cudaSetDevice(s->device), RADPUL_CUDA_DEVICE_SET);
    int rc = ensure_templates(s, 1);
    if (rc) return rc;
    const size_t nfull = ((size_t)s->g.nsamples + 3) & ~(size_t)3;
    const uint32_t nu = s->g.nsamples_unpadded;
    if (!s->d_tap_series) CU_TRY(cudaMalloc((void **)&s->d_tap_series, sizeof(float) * nfull), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    if (!s->d_tap_idx) CU_TRY(cudaMalloc((void **)&s->d_tap_idx, sizeof(int32_t) * s->rs_stride), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    int32_t *d_idx2 = NULL;
    unsigned long long *d_count = NULL;
    uint32_t *d_ns2 = NULL;
    CU_TRY(cudaMalloc((void **)&d_idx2, sizeof(int32_t) * s->rs_stride), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    if (cudaMalloc((void **)&d_count, 8) != cudaSuccess || cudaMalloc((void **)&d_ns2, 8) != cudaSuccess) {
        cudaFree(d_idx2);
        cudaFree(d_count);
        return RADPUL_CUDA_MEM_ALLOC_DEVICE;
    }
    cudaError_t e = cudaMemsetAsync(d_count, 0, 8, s->stream);
    for (uint32_t t = 0; t < n_templates && e == cudaSuccess; t++) {
        brp_template bt;
        brp_derive_template(bank[3 * t], bank[3 * t + 1], bank[3 * t + 2], s->g.step_inv, &bt);
        BrpTemplateDev *h = staging_acquire(s);
        h[0].tau = bt.tau; h[0].Omega = bt.Omega; h[0].Psi0 = bt.Psi0; h[0].S0 = bt.S0;
        e = staging_upload(s, 1) ? cudaErrorUnknown : cudaSuccess;
        if (e == cudaSuccess) e = cudaMemsetAsync(s->d_tap_idx, 0xFF, sizeof(int32_t) * nu, s->stream);
        if (e == cudaSuccess) e = cudaMemsetAsync(d_idx2, 0xFF, sizeof(int32_t) * nu, s->stream);
        if (e == cudaSuccess)
            e = brp_launch_resample(s->d_series, s->d_tmpl, 1, s->gd, s->d_tap_series, nfull, s->d_tap_idx, s->d_nsteps, s->d_mean,
                                    s->d_partials, s->d_counters, false, s->stream, false);
        if (e == cudaSuccess)
            e = brp_launch_resample(s->d_series, s->d_tmpl, 1, s->gd, s->d_tap_series, nfull, d_idx2, d_ns2, s->d_mean,
                                    s->d_partials, s->d_counters, false, s->stream, true);
        if (e == cudaSuccess) e = brp_launch_count_mismatch(s->d_tap_idx, d_idx2, nu, d_count, s->stream);
        if (e == cudaSuccess) e = brp_launch_count_mismatch((const int32_t *)s->d_nsteps, (const int32_t *)d_ns2, 1, d_count, s->stream);
        s->launches += 4;
    }
    unsigned long long h = 0;
    if (e == cudaSuccess) e = cudaMemcpyAsync(&h, d_count, 8, cudaMemcpyDeviceToHost, s->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
    cudaFree(d_idx2);
    cudaFree(d_count);
    cudaFree(d_ns2);
    if (e != cudaSuccess) return RADPUL_CUDA_KERNEL_INVOKE;
    *n_mismatch = h;
    return 0;
}

// n_steps_bound: how many samples of d_in can be valid (= how far the FFT's first pass may read)
static int spectrum_common(brp_search *s, const float *d_in, size_t in_stride, float *ps_out, uint32_t count, uint32_t n_steps_bound)
{
    if (count > s->g.fft_size) return RADPUL_EVAL;
    if (!s->d_tap_ps) CU_TRY(cudaMalloc((void **)&s->d_tap_ps, sizeof(float) * (s->g.fft_size + 32)), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    const float norm = 1.0f / (float)s->g.nsamples;
    CU_TRY(brp_launch_fft_power(s->plan, d_in, in_stride, s->d_nsteps, s->d_mean, s->d_work, s->d_tap_ps, s->g.fft_size + 32, count,
                                norm, 1, s->stream, NULL, n_steps_bound),
           RADPUL_CUDA_FFT_EXEC);
    s->launches += 2;
    CU_TRY(cudaMemcpyAsync(ps_out, s->d_tap_ps, sizeof(float) * count, cudaMemcpyDeviceToHost, s->stream), RADPUL_CUDA_MEM_COPY_DEVICE_HOST);
    CU_TRY(cudaStreamSynchronize(s->stream), RADPUL_CUDA_KERNEL_INVOKE);
    return 0;
}

int brp_stage_spectrum(brp_search *s, float *ps_out, uint32_t count)
{
    if (!s || !ps_out) return RADPUL_EVAL;
    CU_TRY(cudaSetDevice(s->device), RADPUL_CUDA_DEVICE_SET);
    // slot 0 holds the series resampled by the last brp_stage_resample, with its n_steps and mean
    return spectrum_common(s, s->d_resamp, s->rs_stride, ps_out, count, s->g.nsamples_unpadded);
}

int brp_stage_spectrum_of(brp_search *s, const float *host_series_n, float *ps_out, uint32_t count)
{
    if (!s || !host_series_n || !ps_out) return RADPUL_EVAL;
    CU_TRY(cudaSetDevice(s->device), RADPUL_CUDA_DEVICE_SET);
    const size_t nfull = ((size_t)s->g.nsamples + 3) & ~(size_t)3;
    if (!s->d_tap_series) CU_TRY(cudaMalloc((void **)&s->d_tap_series, sizeof(float) * nfull), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    CU_TRY(cudaMemcpyAsync(s->d_tap_series, host_series_n, sizeof(float) * s->g.nsamples, cudaMemcpyHostToDevice, s->stream), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
    const uint32_t full = s->g.nsamples;
    const float zero = 0.0f;
    CU_TRY(cudaMemcpyAsync(s->d_nsteps, &full, sizeof(uint32_t), cudaMemcpyHostToDevice, s->stream), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
    CU_TRY(cudaMemcpyAsync(s->d_mean, &zero, sizeof(float), cudaMemcpyHostToDevice, s->stream), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
    CU_TRY(cudaStreamSynchronize(s->stream), RADPUL_CUDA_KERNEL_INVOKE);
    return spectrum_common(s, s->d_tap_series, nfull, ps_out, count, s->g.nsamples);
}

int brp_stage_harmonic_summing(brp_search *s, const float *host_ps, const float thr[5], uint32_t tmpl_index, brp_hit *hits,
                               uint32_t max_hits, uint32_t *n_hits)
{
    if (!s || !host_ps || !thr || !n_hits) return RADPUL_EVAL;
    CU_TRY(cudaSetDevice(s->device), RADPUL_CUDA_DEVICE_SET);
    if (!s->d_tap_keys) CU_TRY(cudaMalloc((void **)&s->d_tap_keys, sizeof(unsigned long long) * s->hit_cap), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    CU_TRY(cudaMemsetAsync(s->d_tap_keys, 0, sizeof(unsigned long long) * s->hit_cap, s->stream), RADPUL_CUDA_KERNEL_INVOKE);
    CU_TRY(cudaMemcpyAsync(s->d_ps, host_ps, sizeof(float) * s->g.harmonic_idx_hi, cudaMemcpyHostToDevice, s->stream), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
    BrpThresholds t;
    for (int k = 0; k < 5; k++) t.t[k] = thr[k];
    CU_TRY(brp_launch_harmonic_summing(s->d_ps, s->ps_stride, 1, s->gd, t, tmpl_index, s->d_tap_keys, s->stream), RADPUL_CUDA_KERNEL_INVOKE);
    s->launches += 1;
    return collect_from(s, s->d_tap_keys, hits, max_hits, n_hits);
}

static int elementwise_tap(int device, const void *in, void *out, uint32_t n, bool sin_mode)
{
    if (!in || !out) return RADPUL_EVAL;
    int rc = set_device(device);
    if (rc) return rc;
    void *d_in = NULL, *d_out = NULL;
    CU_TRY(cudaMalloc(&d_in, 4 * (size_t)n + 4), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    if (cudaMalloc(&d_out, 4 * (size_t)n + 4) != cudaSuccess) { cudaFree(d_in); return RADPUL_CUDA_MEM_ALLOC_DEVICE; }
    cudaError_t e = cudaMemcpy(d_in, in, 4 * (size_t)n, cudaMemcpyHostToDevice);
    if (e == cudaSuccess)
        e = sin_mode ? brp_launch_sin_lut((const float *)d_in, (float *)d_out, n, 0)
                     : brp_launch_round_index((const float *)d_in, (int32_t *)d_out, n, 0);
    if (e == cudaSuccess) e = cudaMemcpy(out, d_out, 4 * (size_t)n, cudaMemcpyDeviceToHost);
    cudaFree(d_in);
    cudaFree(d_out);
    if (e != cudaSuccess) {
        fprintf(stderr, "[brp_b200] stage tap failed: %s\n", cudaGetErrorString(e));
        return RADPUL_CUDA_KERNEL_INVOKE;
    }
    return 0;
}

int brp_stage_sin_lut(int device, const float *x, float *sin_out, uint32_t n) { return elementwise_tap(device, x, sin_out, n, true); }
int brp_stage_round_index(int device, const float *f, int32_t *idx_out, uint32_t n) { return elementwise_tap(device, f, idx_out, n, false); }

} // extern "C"
