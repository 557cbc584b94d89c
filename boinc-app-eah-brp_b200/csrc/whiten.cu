// Whitening pre-step of a work-unit on the GPU: demod_binary.c:858-1079.
//
// Once per work-unit, before the template loop, the reference
//   1. zero-pads the series to N = padding*Nu samples and takes its real FFT            (:905-926)
//   2. forms the periodogram |X_k|^2 (DC excluded)                                      (:939)
//   3. takes a running median over `window` bins (rngmed.c:48)                          (:953)
//   4. scales X_{i+window/2} by sqrt(ln 2 / median_i)                                   (:964-969)
//   5. replaces zap-list bins by Gaussian noise (GSL taus2 + ziggurat)                  (:982-1033)
//   6. zeroes window/2 bins at either end, transforms back, multiplies by 1/sqrt(N)     (:1038-1069)
// Steps 1-4 and 6 are done here with the library's own FFT kernels (fft.cu) and an exact sliding median.  The noise of
// step 5 comes from a strictly sequential generator and is only a few thousand numbers: it is drawn on the host in the
// reference's order (zap.cpp) and scattered into the spectrum by zap_kernel between steps 4 and 6.
//
// The median is exact (same order statistics as rngmed.c, even windows average the two middle values in
// double like rngmed.c:179,329): see running_median_kernel.
#include <math.h>
#include <stdio.h>

#include <vector>

#include "brp_internal.h"

// zap.cpp: the replaced bins and their noise values, drawn in the reference's order
int brp_zap_draw(uint32_t seed_bits, float padding, uint32_t nsamples, double tsample_us, const double *zap, uint32_t n_zap,
                 uint32_t fft_size, std::vector<uint32_t> &bins, std::vector<float2> &values);

namespace {

constexpr int kMedThreads = 256;

__global__ void periodogram_kernel(const float2 *__restrict__ X, float *__restrict__ ps, uint32_t f)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= f) return;
    const float2 v = X[i];
    // gsl_pow_2(re) + gsl_pow_2(im) in double, rounded once (:939); the DC bin stays 0 (calloc + loop from 1)
    ps[i] = i == 0 ? 0.0f : (float)((double)v.x * (double)v.x + (double)v.y * (double)v.y);
}

// Running median.  A thread produces kMedRun consecutive outputs.  The first one selects the k-th smallest value of its
// window by bisection on the bit pattern of the non-negative floats (31 passes over the window).  Every further window
// differs from the previous one by one element out and one in, so the rank of the previous median moves by at most one:
// ONE pass that counts the values below / not above the previous median and finds its nearest neighbours on either side
// decides between "same value", "predecessor" and "successor".  An even window needs the (k+1)-th value too: one more pass
// relative to the new k-th value.  About 3.9 passes per output instead of 32, same order statistics.
// The window lives in shared memory; a thread's windows start kMedRun words apart, so each lane walks its window from a
// different offset (lane t starts at element t and wraps around): the 32 addresses of a warp then fall into 32 different banks.
constexpr int kMedRun = 16;

struct MedScan {
    int lt, le;        // values < pivot, values <= pivot
    uint32_t pred, succ; // largest value < pivot (0 if none), smallest value > pivot (+inf pattern if none)
};

__device__ __forceinline__ MedScan med_scan(const uint32_t *__restrict__ win, int n, int skew, uint32_t pivot)
{
    MedScan r{0, 0, 0u, 0x7F800000u};
    auto visit = [&](uint32_t v) {
        r.lt += (v < pivot);
        r.le += (v <= pivot);
        if (v < pivot) r.pred = max(r.pred, v);
        if (v > pivot) r.succ = min(r.succ, v);
    };
    for (int j = skew; j < n; ++j) visit(win[j]);
    for (int j = skew - 1; j >= 0; --j) visit(win[j]); // downwards: lane t reads word 17 t - 1 - step, again one bank per lane
    return r;
}

// k-th smallest (0-based) of win[0..n) for non-negative floats: bisection on the bit pattern
__device__ __forceinline__ uint32_t select_kth(const uint32_t *__restrict__ win, int n, int k, int skew)
{
    uint32_t x = 0;
    for (int bit = 30; bit >= 0; --bit) {
        const uint32_t cand = x | (1u << bit);
        int cnt = 0;
        for (int j = skew; j < n; ++j) cnt += (win[j] < cand);
        for (int j = skew - 1; j >= 0; --j) cnt += (win[j] < cand);
        if (cnt <= k) x = cand;
    }
    return x;
}

__global__ void __launch_bounds__(kMedThreads)
running_median_kernel(const float *__restrict__ ps, uint32_t length, uint32_t bsize, float *__restrict__ med, uint32_t n_out)
{
    extern __shared__ uint32_t sw[]; // kMedThreads*kMedRun + bsize - 1 values as bit patterns
    const uint32_t first = blockIdx.x * (kMedThreads * kMedRun);
    const uint32_t span = kMedThreads * kMedRun + bsize - 1;
    for (uint32_t j = threadIdx.x; j < span; j += kMedThreads) {
        const uint32_t src = first + j;
        sw[j] = src < length ? __float_as_uint(ps[src]) : 0x7F800000u; // +inf beyond the end (never selected)
    }
    __syncthreads();
    const int n = (int)bsize;
    const int mid = (int)((bsize + (bsize & 1u)) / 2u) - 1; // rngmed.c:104
    const int skew = min((int)(threadIdx.x & 31), n - 1);
    uint32_t a = 0;
    for (int r = 0; r < kMedRun; ++r) {
        const uint32_t i = first + threadIdx.x * kMedRun + r;
        if (i >= n_out) return;
        const uint32_t *win = sw + threadIdx.x * kMedRun + r;
        if (r == 0) {
            a = select_kth(win, n, mid, skew);
        } else {
            const MedScan s = med_scan(win, n, skew, a); // a = k-th value of the previous window
            if (mid < s.lt) a = s.pred;                  // rank moved up by one: the k-th value is the predecessor
            else if (mid >= s.le) a = s.succ;            // rank moved down by one (or the old median left the window)
        }
        if (bsize & 1u) {
            med[i] = __uint_as_float(a);
            continue;
        }
        // even window: the next order statistic is a itself if it occurs again above rank mid, else the smallest value > a
        const MedScan s2 = med_scan(win, n, skew, a);
        const uint32_t bb = (s2.le >= mid + 2) ? a : s2.succ;
        med[i] = (float)((double)(__uint_as_float(a) + __uint_as_float(bb)) / 2.0); // float add, double halving (rngmed.c:179)
    }
}

// steps 4 and the zeroing of step 6
__global__ void scale_and_edge_kernel(float2 *__restrict__ X, const float *__restrict__ med, uint32_t f, uint32_t window_2,
                                      uint32_t white_size)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= f) return;
    float2 v = X[k];
    if (k >= window_2 && k - window_2 < white_size) {
        const float factor = (float)sqrt(M_LN2 / (double)med[k - window_2]); // :966
        v.x *= factor;
        v.y *= factor;
    }
    if (k < window_2 || k + window_2 >= f) v = make_float2(0.f, 0.f); // :1038-1044
    X[k] = v;
}

// step 5: X[bin] = noise for the bins of the zap list (:1018-1019); bins inside the zeroed edges stay zero (:1038-1044 come
// later in the reference and win)
__global__ void zap_kernel(float2 *__restrict__ X, const uint32_t *__restrict__ bins, const float2 *__restrict__ values, uint32_t n,
                           uint32_t f, uint32_t window_2)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t k = bins[i];
    if (k < window_2 || k + window_2 >= f) return;
    X[k] = values[i];
}

// backward real transform of X[0..M] as one complex transform of length M:
//   Z'[k] = (X[k] + conj X[M-k]) + i W_N^{-k} (X[k] - conj X[M-k]),   x = conj(FFT(conj Z')) interleaved
__global__ void inverse_prep_kernel(const float2 *__restrict__ X, float2 *__restrict__ zc, uint32_t m, float two_over_n)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= m) return;
    const float2 a = X[k], b = X[m - k];
    const float2 s = make_float2(a.x + b.x, a.y - b.y);
    const float2 d = make_float2(a.x - b.x, a.y + b.y);
    float sn, cs;
    sincospif((float)k * two_over_n, &sn, &cs); // W_N^{-k} = exp(+2 pi i k / N)
    const float2 wd = make_float2(cs * d.x - sn * d.y, cs * d.y + sn * d.x);
    zc[k] = make_float2(s.x - wd.y, -(s.y + wd.x));
}

} // namespace

#define WH_TRY(call, code)                                                                                   \
    do {                                                                                                     \
        cudaError_t e_ = (call);                                                                             \
        if (e_ != cudaSuccess) {                                                                             \
            fprintf(stderr, "[brp_b200] %s failed: %s (%s:%d)\n", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            rc = (code);                                                                                     \
            goto done;                                                                                       \
        }                                                                                                    \
    } while (0)

extern "C" int brp_whiten_series_zap(int device, float *series, uint32_t nsamples_unpadded, float padding, uint32_t window,
                                     double tsample_us, const double *zap_ranges, uint32_t n_zap_ranges, float *device_ms)
{
    if (!series || nsamples_unpadded < 2 || padding < 1.0f || padding > 10.0f) return RADPUL_EVAL;
    if (n_zap_ranges != 0 && (!zap_ranges || !(tsample_us > 0.0))) return RADPUL_EVAL;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0 || device < 0 || device >= count) return RADPUL_CUDA_DEVICE_FIND;
    if (cudaSetDevice(device) != cudaSuccess) return RADPUL_CUDA_DEVICE_SET;

    const uint32_t nu = nsamples_unpadded;
    const uint32_t n = (uint32_t)(int)(padding * nu + 0.5);                 // demod_binary.c:782
    const uint32_t window_2 = (uint32_t)(int)(0.5 * window + 0.5);          // :869
    const uint32_t f = (uint32_t)(0.5 * n + 0.5) + 1;                       // :875
    if (f < window || window == 0) return RADPUL_EVAL;                       // :878-882
    if (window > 16384) return RADPUL_EVAL;                                  // shared-memory staging of the median window
    const uint32_t white_size = f - window + 1;
    const uint32_t m = n / 2;

    int rc = 0;
    BrpFftPlan plan;
    float *d_x = NULL, *d_ps = NULL, *d_med = NULL, *d_out = NULL, *d_mean = NULL;
    float2 *d_X = NULL, *d_zc = NULL, *d_work = NULL, *d_zap_val = NULL;
    uint32_t *d_ns = NULL, *d_zap_bin = NULL;
    std::vector<uint32_t> zap_bins;
    std::vector<float2> zap_vals;
    cudaEvent_t ev0 = NULL, ev1 = NULL;
    cudaStream_t st = 0;
    memset(&plan, 0, sizeof(plan));
    {
        uint32_t seed_bits; // the first sample's bit pattern, before anything is changed (:917)
        memcpy(&seed_bits, series, sizeof(seed_bits));
        rc = brp_zap_draw(seed_bits, padding, n, tsample_us, zap_ranges, n_zap_ranges, f, zap_bins, zap_vals);
        if (rc) return rc;
    }
    rc = brp_fft_plan_create(&plan, n);
    if (rc) return rc;
    {
        const uint32_t ns_host[2] = {nu, n};
        const float zero = 0.0f;
        const size_t nu4 = ((size_t)nu + 3) & ~(size_t)3;
        const int med_smem = (int)((kMedThreads * kMedRun + window - 1) * sizeof(uint32_t));
        WH_TRY(cudaMalloc((void **)&d_x, sizeof(float) * nu4), RADPUL_CUDA_MEM_ALLOC_DEVICE);
        WH_TRY(cudaMalloc((void **)&d_X, sizeof(float2) * ((size_t)m + 2)), RADPUL_CUDA_MEM_ALLOC_DEVICE);
        WH_TRY(cudaMalloc((void **)&d_zc, sizeof(float2) * (size_t)m), RADPUL_CUDA_MEM_ALLOC_DEVICE);
        WH_TRY(cudaMalloc((void **)&d_work, sizeof(float2) * (size_t)m), RADPUL_CUDA_MEM_ALLOC_DEVICE);
        WH_TRY(cudaMalloc((void **)&d_ps, sizeof(float) * (size_t)f), RADPUL_CUDA_MEM_ALLOC_DEVICE);
        WH_TRY(cudaMalloc((void **)&d_med, sizeof(float) * (size_t)white_size), RADPUL_CUDA_MEM_ALLOC_DEVICE);
        WH_TRY(cudaMalloc((void **)&d_out, sizeof(float) * (nu4 + 4)), RADPUL_CUDA_MEM_ALLOC_DEVICE);
        WH_TRY(cudaMalloc((void **)&d_ns, sizeof(uint32_t) * 2), RADPUL_CUDA_MEM_ALLOC_DEVICE);
        WH_TRY(cudaMalloc((void **)&d_mean, sizeof(float)), RADPUL_CUDA_MEM_ALLOC_DEVICE);
        WH_TRY(cudaEventCreate(&ev0), RADPUL_CUDA_DEVICE_SET);
        WH_TRY(cudaEventCreate(&ev1), RADPUL_CUDA_DEVICE_SET);
        WH_TRY(cudaMemcpy(d_x, series, sizeof(float) * nu, cudaMemcpyHostToDevice), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
        WH_TRY(cudaMemcpy(d_ns, ns_host, sizeof(ns_host), cudaMemcpyHostToDevice), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
        WH_TRY(cudaMemcpy(d_mean, &zero, sizeof(float), cudaMemcpyHostToDevice), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
        if (!zap_bins.empty()) {
            WH_TRY(cudaMalloc((void **)&d_zap_bin, sizeof(uint32_t) * zap_bins.size()), RADPUL_CUDA_MEM_ALLOC_DEVICE);
            WH_TRY(cudaMalloc((void **)&d_zap_val, sizeof(float2) * zap_vals.size()), RADPUL_CUDA_MEM_ALLOC_DEVICE);
            WH_TRY(cudaMemcpy(d_zap_bin, zap_bins.data(), sizeof(uint32_t) * zap_bins.size(), cudaMemcpyHostToDevice), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
            WH_TRY(cudaMemcpy(d_zap_val, zap_vals.data(), sizeof(float2) * zap_vals.size(), cudaMemcpyHostToDevice), RADPUL_CUDA_MEM_COPY_HOST_DEVICE);
        }
        if (med_smem > 48 * 1024)
            WH_TRY(cudaFuncSetAttribute((const void *)running_median_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, med_smem),
                   RADPUL_CUDA_KERNEL_INVOKE);
        cudaEventRecord(ev0, st);
        // 1. forward transform of the zero-padded series: samples >= Nu read as 0 (never materialised)
        WH_TRY(brp_launch_fft_spectrum(plan, d_x, d_ns, d_mean, d_work, d_X, m + 1, st), RADPUL_CUDA_FFT_EXEC);
        // 2.-4. periodogram, running median, rescale, zero the edges
        periodogram_kernel<<<(f + 255) / 256, 256, 0, st>>>(d_X, d_ps, f);
        running_median_kernel<<<(white_size + kMedThreads * kMedRun - 1) / (kMedThreads * kMedRun), kMedThreads, med_smem, st>>>(d_ps, f, window, d_med, white_size);
        scale_and_edge_kernel<<<(f + 255) / 256, 256, 0, st>>>(d_X, d_med, f, window_2, white_size);
        // 5. zap-list noise
        if (!zap_bins.empty())
            zap_kernel<<<((uint32_t)zap_bins.size() + 255) / 256, 256, 0, st>>>(d_X, d_zap_bin, d_zap_val, (uint32_t)zap_bins.size(), f, window_2);
        // 6. backward transform and renormalisation: 1/sqrt((float)N) held in a float (:1066)
        inverse_prep_kernel<<<(m + 255) / 256, 256, 0, st>>>(d_X, d_zc, m, 2.0f / (float)n);
        WH_TRY(cudaGetLastError(), RADPUL_CUDA_KERNEL_INVOKE);
        const float norm_factor = (float)(1.0 / sqrt((double)(float)n));
        WH_TRY(brp_launch_fft_conj(plan, d_zc, d_ns + 1, d_mean, d_work, d_out, (nu + 1) / 2, norm_factor, st), RADPUL_CUDA_FFT_EXEC);
        cudaEventRecord(ev1, st);
        WH_TRY(cudaMemcpy(series, d_out, sizeof(float) * nu, cudaMemcpyDeviceToHost), RADPUL_CUDA_MEM_COPY_DEVICE_HOST);
        if (device_ms) cudaEventElapsedTime(device_ms, ev0, ev1);
    }
done:
    cudaFree(d_x); cudaFree(d_X); cudaFree(d_zc); cudaFree(d_work); cudaFree(d_ps); cudaFree(d_med); cudaFree(d_out);
    cudaFree(d_ns); cudaFree(d_mean); cudaFree(d_zap_bin); cudaFree(d_zap_val);
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    brp_fft_plan_destroy(&plan);
    return rc;
}

// the original entry point: whitening without a zap list
extern "C" int brp_whiten_series(int device, float *series, uint32_t nsamples_unpadded, float padding, uint32_t window,
                                 uint32_t n_zap_ranges, float *device_ms)
{
    if (n_zap_ranges != 0) {
        fprintf(stderr, "[brp_b200] brp_whiten_series: pass the zap ranges to brp_whiten_series_zap\n");
        return RADPUL_EVAL;
    }
    return brp_whiten_series_zap(device, series, nsamples_unpadded, padding, window, 0.0, NULL, 0, device_ms);
}
