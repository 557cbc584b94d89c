// The reference's backend interface, one template per call, on top of the same kernels.
//
// These are the eleven symbols demod_binary.c calls when built with -DUSE_CUDA
// (cuda/app/demod_binary_cuda.h:42-52, cuda/app/demod_binary_hs_cuda.cuh:37-39); state is module-global
// exactly like in every reference backend (cuda/app/demod_binary_cuda.cu:37-59), i.e. one work-unit per
// process, single caller thread, set_up_* once in the order resampling -> fft -> harmonic summing, run_*
// once per template in that order.  This path exists to be a drop-in and a parity proof; throughput
// numbers come from the batched brp_search_* layer.
//
// What keeps the per-template cost down (the interface forces one host round trip per template, nothing forces three):
//   * run_resampling and run_fft only enqueue; the one synchronisation of a template is in run_harmonic_summing, which is
//     where the reference driver first looks at results (errors of the earlier kernels surface there);
//   * the handles passed between the stages are opaque, so the padding is not materialised: run_fft on the buffer of
//     run_resampling reads n_steps and the mean the resampler left on the device (pad-on-load, as in the batched path);
//   * nothing of the size of the spectrum crosses the bus or is cleared on the host per template: only the hit list comes
//     back (a few hundred bytes), sumspec[0..4] and dirty[] hold the values above threshold and zeros elsewhere -- all the
//     candidate scan reads (demod_binary.c:1350) -- and the entries written for one template are cleared one by one before
//     the next.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "brp_internal.h"

namespace {

struct LegacyState {
    bool initialised;
    int device;
    cudaStream_t stream;
    // resampling
    BrpGeomDev gd;
    float *d_series, *d_resamp;
    uint32_t *d_nsteps;
    float *d_mean;
    double *d_partials;
    uint32_t *d_counters;
    // fft
    BrpFftPlan plan;
    bool have_plan;
    float2 *d_work;
    float *d_ps;
    uint32_t *d_nsteps_full;
    float *d_mean_zero;
    // harmonic summing
    unsigned long long *d_keys;
    // mapped pinned host memory the collecting kernel writes straight into: [hit count (one brp_hit-sized slot)][hits ...];
    // d_hitbuf is the same memory as the device sees it.  d_ctl: the kernel's three control words (brp_launch_collect_to_host)
    brp_hit *d_hitbuf, *h_hitbuf;
    uint32_t *d_ctl;
    unsigned calls; // parity of the control words
    uint32_t hit_cap;
    float *h_spectrum; // handed out as sumspec[0]: the spectrum's values above threshold, 0 elsewhere
    uint32_t fu, ps_stride;
    uint32_t n_prev;   // hits of the previous template, still in h_hitbuf: their entries are cleared before the next fill
    bool resampled;    // d_resamp holds [0, n_steps) of the current template with d_nsteps / d_mean (padding not written)
};

LegacyState L;

#define LG_TRY(call, code)                                                                          \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess) {                                                                    \
            fprintf(stderr, "[brp_b200] %s failed: %s (%s:%d)\n", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return (code);                                                                          \
        }                                                                                           \
    } while (0)

inline float *as_dev(DIfloatPtr p) { return reinterpret_cast<float *>(p.device_ptr); }

} // namespace

extern "C" {

int initialize_cuda(int cudDeviceIdGiven, int *cudDeviceId)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) {
        fprintf(stderr, "[brp_b200] no CUDA device available: this backend has no CPU fallback\n");
        return RADPUL_CUDA_DEVICE_FIND;
    }
    int dev = 0;
    if (cudDeviceIdGiven && cudDeviceId) dev = *cudDeviceId;
    if (dev < 0 || dev >= count) return RADPUL_CUDA_DEVICE_FIND;
    LG_TRY(cudaSetDevice(dev), RADPUL_CUDA_DEVICE_SET);
    memset(&L, 0, sizeof(L));
    L.device = dev;
    LG_TRY(cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking), RADPUL_CUDA_DEVICE_SET);
    L.initialised = true;
    if (cudDeviceId) *cudDeviceId = dev;
    return 0;
}

int shutdown_cuda(void)
{
    if (!L.initialised) return 0;
    cudaSetDevice(L.device);
    cudaStreamSynchronize(L.stream);
    cudaStreamDestroy(L.stream);
    memset(&L, 0, sizeof(L));
    return 0;
}

int set_up_resampling(DIfloatPtr input, DIfloatPtr *output, const RESAMP_PARAMS *params, float *sinLUTsamples,
                      float *cosLUTsamples)
{
    (void)sinLUTsamples; // the tables of erp_utilities.cpp:45-46 are compiled into the library
    (void)cosLUTsamples;
    if (!L.initialised) {
        int rc = initialize_cuda(0, NULL);
        if (rc) return rc;
    }
    if (!input.host_ptr || !output || !params || params->nsamples_unpadded < 2 || params->nsamples < params->nsamples_unpadded)
        return RADPUL_EVAL;
    LG_TRY(cudaSetDevice(L.device), RADPUL_CUDA_DEVICE_SET);
    const uint32_t nu = params->nsamples_unpadded, n = params->nsamples;
    L.gd.nu = nu;
    L.gd.n = n;
    L.gd.fft_size = params->fft_size;
    const size_t nu4 = ((size_t)nu + 3) & ~(size_t)3, n4 = ((size_t)n + 3) & ~(size_t)3;
    LG_TRY(cudaMalloc((void **)&L.d_series, sizeof(float) * nu4), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    LG_TRY(cudaMalloc((void **)&L.d_resamp, sizeof(float) * n4), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    LG_TRY(cudaMalloc((void **)&L.d_nsteps, sizeof(uint32_t)), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    LG_TRY(cudaMalloc((void **)&L.d_mean, sizeof(float)), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    LG_TRY(cudaMalloc((void **)&L.d_partials, sizeof(double) * brp_resamp_blocks(nu)), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    LG_TRY(cudaMalloc((void **)&L.d_counters, sizeof(uint32_t)), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    LG_TRY(cudaMemset(L.d_counters, 0, sizeof(uint32_t)), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
    // the only bulk host->device transfer of a work-unit (cuda/app/demod_binary_cuda.cu:333)
    LG_TRY(cudaMemcpy(L.d_series, input.host_ptr, sizeof(float) * nu, cudaMemcpyHostToDevice), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
    output->device_ptr = (brp_device_ptr_t) reinterpret_cast<uintptr_t>(L.d_resamp);
    return 0;
}

int run_resampling(DIfloatPtr input, DIfloatPtr output, const RESAMP_PARAMS *params)
{
    (void)input; // the series was uploaded by set_up_resampling
    if (!L.initialised || !L.d_series || !params || as_dev(output) != L.d_resamp) return RADPUL_EVAL;
    LG_TRY(cudaSetDevice(L.device), RADPUL_CUDA_DEVICE_SET);
    BrpGeomDev g = L.gd;
    g.dt = params->dt;
    g.step_inv = params->step_inv;
    const BrpTemplateDev t = {params->tau, params->Omega, params->Psi0, params->S0};
    const size_t n4 = ((size_t)g.n + 3) & ~(size_t)3;
    // the template travels as a kernel argument: nothing to upload first
    LG_TRY(brp_launch_resample(L.d_series, NULL, 1, g, L.d_resamp, n4, NULL, L.d_nsteps, L.d_mean, L.d_partials, L.d_counters,
                               false, L.stream, false, &t),
           RADPUL_CUDA_KERNEL_INVOKE);
    L.resampled = true;
    return 0;
}

int tear_down_resampling(DIfloatPtr output)
{
    (void)output;
    if (!L.initialised) return 0;
    cudaSetDevice(L.device);
    cudaFree(L.d_series); cudaFree(L.d_resamp); cudaFree(L.d_nsteps); cudaFree(L.d_mean);
    cudaFree(L.d_partials); cudaFree(L.d_counters);
    L.d_series = L.d_resamp = NULL; L.d_nsteps = NULL; L.d_mean = NULL; L.d_partials = NULL; L.d_counters = NULL;
    return 0;
}

int set_up_fft(DIfloatPtr input, DIfloatPtr *output, uint32_t nsamples, unsigned int fft_size)
{
    (void)input;
    if (!L.initialised || !output) return RADPUL_EVAL;
    if (fft_size != nsamples / 2 + 1) return RADPUL_EVAL;
    LG_TRY(cudaSetDevice(L.device), RADPUL_CUDA_DEVICE_SET);
    int rc = brp_fft_plan_create(&L.plan, nsamples, L.gd.nu); // nu: lets pass 1 run as sub-transforms when the upper rows are padding
    if (rc) {
        fprintf(stderr, "[brp_b200] FFT length %u unsupported (2^a or 3*2^a, >= 512 / 1536)\n", nsamples);
        return rc;
    }
    L.have_plan = true;
    LG_TRY(cudaMalloc((void **)&L.d_work, sizeof(float2) * (size_t)L.plan.m), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    // rows of m2 bins, the last one whole, so that pass 2 may store through tensor copies (fft.cu, pass2_tma_map); the
    // harmonic-summing bulk copies read up to 16 bytes past the last bin they need
    L.ps_stride = (uint32_t)((((size_t)fft_size + L.plan.m2 - 1) / L.plan.m2 * L.plan.m2 + 64 + 31) & ~(size_t)31);
    LG_TRY(cudaMalloc((void **)&L.d_ps, sizeof(float) * L.ps_stride), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    LG_TRY(cudaMalloc((void **)&L.d_nsteps_full, sizeof(uint32_t)), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    LG_TRY(cudaMalloc((void **)&L.d_mean_zero, sizeof(float)), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    const float zero = 0.0f;
    LG_TRY(cudaMemcpy(L.d_nsteps_full, &nsamples, sizeof(uint32_t), cudaMemcpyHostToDevice), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
    LG_TRY(cudaMemcpy(L.d_mean_zero, &zero, sizeof(float), cudaMemcpyHostToDevice), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
    LG_TRY(cudaMemset(L.d_ps, 0, sizeof(float) * L.ps_stride), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
    output->device_ptr = (brp_device_ptr_t) reinterpret_cast<uintptr_t>(L.d_ps);
    return 0;
}

int run_fft(DIfloatPtr input, DIfloatPtr output, uint32_t nsamples, unsigned int fft_size, float norm_factor)
{
    if (!L.initialised || !L.have_plan || nsamples != L.plan.n || as_dev(output) != L.d_ps || !as_dev(input)) return RADPUL_EVAL;
    LG_TRY(cudaSetDevice(L.device), RADPUL_CUDA_DEVICE_SET);
    const size_t n4 = ((size_t)nsamples + 3) & ~(size_t)3;
    if (as_dev(input) == L.d_resamp && L.resampled) {
        // the buffer run_resampling filled: samples [0, n_steps) are there, everything beyond reads as the mean
        LG_TRY(brp_launch_fft_power(L.plan, L.d_resamp, n4, L.d_nsteps, L.d_mean, L.d_work, L.d_ps, L.ps_stride, fft_size, norm_factor, 1,
                                    L.stream, NULL, L.gd.nu),
               RADPUL_CUDA_FFT_EXEC);
    } else {
        // some other device buffer: a fully padded series, every sample taken from memory
        LG_TRY(brp_launch_fft_power(L.plan, as_dev(input), n4, L.d_nsteps_full, L.d_mean_zero, L.d_work, L.d_ps, L.ps_stride, fft_size,
                                    norm_factor, 1, L.stream, NULL),
               RADPUL_CUDA_FFT_EXEC);
    }
    return 0;
}

int tear_down_fft(DIfloatPtr output)
{
    (void)output;
    if (!L.initialised) return 0;
    cudaSetDevice(L.device);
    if (L.have_plan) brp_fft_plan_destroy(&L.plan);
    L.have_plan = false;
    cudaFree(L.d_work); cudaFree(L.d_ps); cudaFree(L.d_nsteps_full); cudaFree(L.d_mean_zero);
    L.d_work = NULL; L.d_ps = NULL; L.d_nsteps_full = NULL; L.d_mean_zero = NULL;
    return 0;
}

int set_up_harmonic_summing(float **sumspec, int32_t **dirty, unsigned int *nr_pages_ptr, unsigned int fundamental_idx_hi,
                            unsigned int harmonic_idx_hi)
{
    (void)harmonic_idx_hi;
    if (!L.initialised || !sumspec || !dirty || !nr_pages_ptr) return RADPUL_EVAL;
    LG_TRY(cudaSetDevice(L.device), RADPUL_CUDA_DEVICE_SET);
    const unsigned int nr_pages = (fundamental_idx_hi >> BRP_LOG_PS_PAGE_SIZE) + 1; // demod_binary_hs_cpu.c:37
    *nr_pages_ptr = nr_pages;
    for (int i = 1; i < 5; i++) {
        sumspec[i] = (float *)calloc(fundamental_idx_hi ? fundamental_idx_hi : 1, sizeof(float));
        if (!sumspec[i]) return RADPUL_EMEM;
    }
    for (int i = 0; i < 5; i++) {
        dirty[i] = (int32_t *)calloc(nr_pages, sizeof(int32_t));
        if (!dirty[i]) return RADPUL_EMEM;
    }
    L.fu = fundamental_idx_hi;
    L.hit_cap = 5u * (fundamental_idx_hi ? fundamental_idx_hi : 1u);
    LG_TRY(cudaMalloc((void **)&L.d_keys, sizeof(unsigned long long) * L.hit_cap), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    LG_TRY(cudaHostAlloc((void **)&L.h_hitbuf, sizeof(brp_hit) * ((size_t)L.hit_cap + 1), cudaHostAllocMapped), RADPUL_CUDA_MEM_ALLOC_HOST);
    LG_TRY(cudaHostGetDevicePointer((void **)&L.d_hitbuf, L.h_hitbuf, 0), RADPUL_CUDA_MEM_ALLOC_HOST);
    memset(L.h_hitbuf, 0, sizeof(brp_hit));
    LG_TRY(cudaMalloc((void **)&L.d_ctl, 4 * sizeof(uint32_t)), RADPUL_CUDA_MEM_ALLOC_DEVICE);
    LG_TRY(cudaMemset(L.d_ctl, 0, 4 * sizeof(uint32_t)), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
    L.calls = 0;
    // the per-bin table starts empty and is left empty by every run_harmonic_summing (the collecting kernel clears what it reads)
    LG_TRY(cudaMemset(L.d_keys, 0, sizeof(unsigned long long) * L.hit_cap), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
    L.h_spectrum = (float *)calloc(fundamental_idx_hi ? fundamental_idx_hi : 1, sizeof(float));
    if (!L.h_spectrum) return RADPUL_EMEM;
    L.n_prev = 0;
    return 0;
}

int run_harmonic_summing(float **sumspec, int32_t **dirty, unsigned int nr_pages, DIfloatPtr powerspectrum,
                         unsigned int window_2, unsigned int fundamental_idx_hi, unsigned int harmonic_idx_hi,
                         float *thresholds)
{
    if (!L.initialised || !L.d_keys || !sumspec || !dirty || !thresholds || fundamental_idx_hi != L.fu || !as_dev(powerspectrum))
        return RADPUL_EVAL;
    LG_TRY(cudaSetDevice(L.device), RADPUL_CUDA_DEVICE_SET);
    (void)nr_pages;
    // demod_binary_hs_cpu.c:73-83 zeroes the summed spectra and the page flags and aliases sumspec[0] to the spectrum; here
    // the arrays are all-zero except for what the previous template's hits put there, so those entries are cleared instead
    sumspec[0] = L.h_spectrum;
    for (uint32_t i = 0; i < L.n_prev; i++) {
        const brp_hit &h = L.h_hitbuf[1 + i];
        sumspec[h.harm_idx][h.bin] = 0.0f;
        dirty[h.harm_idx][h.bin >> BRP_LOG_PS_PAGE_SIZE] = 0;
    }
    L.n_prev = 0;
    BrpGeomDev g = L.gd;
    g.w2 = window_2;
    g.fu = fundamental_idx_hi;
    g.h = harmonic_idx_hi;
    BrpThresholds t;
    for (int k = 0; k < 5; k++) t.t[k] = thresholds[k];
    // three launches and no copy per template: the collecting kernel writes the hits and their number into mapped host memory
    LG_TRY(brp_launch_harmonic_summing(as_dev(powerspectrum), L.ps_stride, 1, g, t, 0, L.d_keys, L.stream), RADPUL_CUDA_KERNEL_INVOKE);
    LG_TRY(brp_launch_collect_to_host(L.d_keys, fundamental_idx_hi, 0, L.d_hitbuf + 1, L.hit_cap, L.d_ctl, (int)(L.calls++ & 1u),
                                      reinterpret_cast<uint32_t *>(L.d_hitbuf), L.stream),
           RADPUL_CUDA_KERNEL_INVOKE);
    LG_TRY(cudaStreamSynchronize(L.stream), RADPUL_CUDA_KERNEL_INVOKE); // the one synchronisation of a template
    const uint32_t n = *reinterpret_cast<const volatile uint32_t *>(L.h_hitbuf);
    for (uint32_t i = 0; i < n; i++) {
        const brp_hit &h = L.h_hitbuf[1 + i];
        sumspec[h.harm_idx][h.bin] = h.power; // harm_idx 0: the spectrum bin itself (P[i] > thr[0], hs_common.c:79)
        dirty[h.harm_idx][h.bin >> BRP_LOG_PS_PAGE_SIZE] = 1;
    }
    L.n_prev = n;
    L.resampled = false;
    return 0;
}

int tear_down_harmonic_summing(float **sumspec, int32_t **dirty)
{
    if (sumspec)
        for (int i = 1; i < 5; i++) { free(sumspec[i]); sumspec[i] = NULL; }
    if (dirty)
        for (int i = 0; i < 5; i++) { free(dirty[i]); dirty[i] = NULL; }
    if (!L.initialised) return 0;
    cudaSetDevice(L.device);
    cudaFree(L.d_keys); cudaFree(L.d_ctl); cudaFreeHost(L.h_hitbuf); free(L.h_spectrum);
    L.d_keys = NULL; L.d_ctl = NULL; L.d_hitbuf = NULL; L.h_hitbuf = NULL; L.h_spectrum = NULL;
    return 0;
}

} // extern "C"
