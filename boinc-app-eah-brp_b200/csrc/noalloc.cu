// brp_resample_noalloc: the resampling stage for callers that must not allocate device memory.
//
// Who needs this.  Every other entry point of this library sizes its scratch with cudaMalloc when a
// handle is created.  Some hosts cannot do that at the moment they want a series resampled:
//   * a volunteer-computing task that was admitted with a fixed device-memory budget -- a cudaMalloc
//     in mid-run can fail when another task has taken the rest of the card, and the BOINC wrapper
//     answers an allocation failure with a 15-minute back-off (erp_boinc_wrapper.cpp:560-570);
//   * an application that has handed the whole device to its own arena and gives this library
//     sub-buffers of it, so the driver allocator has nothing left to give;
//   * a latency-sensitive loop that cannot afford cudaMalloc/cudaFree, both of which synchronise the
//     whole device.
// For them the caller supplies every input and output buffer, and the kernel's scratch (the per-block
// partial sums of the time-series mean and the "blocks finished" counter) lives in a fixed array that
// is part of the library image.  The price: that array occupies BRP_NOALLOC_SCRATCH_BYTES on every GPU
// that loads the library until the process exits, the series length is capped at
// BRP_NOALLOC_MAX_SAMPLES, and there is exactly one such array per GPU (see "Sharing" below).
//
// Set-up.  In host code the name of a __device__ array is only a stand-in, not an address a kernel can
// use, so the first call on a GPU asks the runtime for the real address (cudaGetSymbolAddress, with
// that GPU current) and remembers it per device.  Make that first call before starting any stream
// capture.  No call here allocates: no cudaMalloc, cudaFree, or library routine that allocates.
//
// Sharing.  All calls on one GPU use the same scratch array, and it keeps whatever the previous call
// left in it.  The counter is cleared on the caller's stream at the start of every call; the partial
// sums are fully rewritten before they are read.  Two calls whose kernels can overlap on the same GPU
// (different streams, or different host threads) race on the array and give wrong means -- issue them
// on one stream, or order them with events.
//
// The arithmetic is the library's resampler (resamp.cu / brp_device.cuh; reference semantics
// demod_binary_resamp_cpu.c:80-136): n_steps and the gathered samples equal those of brp_stage_resample /
// run_resampling bit for bit; the mean is summed in double like theirs but over differently shaped
// partial sums, so it agrees to the last float ulp or one next to it (tests hold it to 1e-6).
// The remembered scratch address belongs to the device's current context: after cudaDeviceReset it is
// stale, and this entry point must not be used again in that process.
#include <stdio.h>

#include <mutex>

#include "brp_device.cuh"
#include "brp_internal.h"

#define NOALLOC_THREADS 256
#define NOALLOC_SAMPLES_PER_BLOCK 4096
#define NOALLOC_MAX_BLOCKS (BRP_NOALLOC_MAX_SAMPLES / NOALLOC_SAMPLES_PER_BLOCK)
#define NOALLOC_COUNTER_OFFSET (NOALLOC_MAX_BLOCKS * 8)

static_assert(BRP_NOALLOC_SCRATCH_BYTES >= NOALLOC_COUNTER_OFFSET + 256, "scratch array too small for its layout");

// [0, 8*MAX_BLOCKS): one double per block (partial sums); then a 32-bit counter of finished blocks.
static __device__ __align__(256) unsigned char g_noalloc_scratch[BRP_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_setup_lock;
void *g_scratch_addr[kMaxDevices]; // device address of g_noalloc_scratch, per device ordinal (0 = not set up yet)

struct SmemLut {
    const float2 *p;
    __device__ __forceinline__ float2 operator[](int i) const { return p[i]; }
};

__device__ __forceinline__ double warp_total(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

template <bool FAST>
__global__ void __launch_bounds__(NOALLOC_THREADS)
noalloc_resample_kernel(const float *__restrict__ series, BrpTemplateDev tp, uint32_t nu, float dt, float step_inv,
                        float *__restrict__ out, uint32_t *__restrict__ n_steps_out, float *__restrict__ mean_out,
                        double *__restrict__ partials, uint32_t *__restrict__ counter)
{
    __shared__ float2 s_tab[65];
    __shared__ int s_nsteps;
    __shared__ double s_warp[NOALLOC_THREADS / 32];
    __shared__ bool s_last;

    const uint32_t nblk = gridDim.x;
    const uint32_t block_lo = blockIdx.x * NOALLOC_SAMPLES_PER_BLOCK;
    if (threadIdx.x < 65) s_tab[threadIdx.x] = c_brp_sincos[threadIdx.x];
    if (threadIdx.x == 0) s_nsteps = -1;
    __syncthreads();
    const SmemLut tab{s_tab};

    // n_steps: last sample whose detector-time index still lies inside the series (demod_binary_resamp_cpu.c:105-109)
    const uint32_t reach = (uint32_t)(2.02f * fabsf(tp.tau) * step_inv) + 16u;
    const uint32_t win_lo = (nu - 1 > reach) ? (nu - 1 - reach) : 0u;
    uint32_t n_steps = nu;
    if (block_lo + NOALLOC_SAMPLES_PER_BLOCK > win_lo) {
        const float limit = (float)(nu - 1);
        int best = -1;
        for (uint32_t n = win_lo + threadIdx.x; n < nu; n += NOALLOC_THREADS) {
            const float nf = (float)n;
            const float dl = brp_del_t(nf, tp.tau, tp.Omega, tp.Psi0, tp.S0, dt, step_inv, tab);
            if (__fsub_rn(nf, dl) < limit) best = (int)n;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
        if ((threadIdx.x & 31) == 0 && best >= 0) atomicMax(&s_nsteps, best);
        __syncthreads();
        n_steps = s_nsteps >= 0 ? (uint32_t)s_nsteps : (win_lo > 0 ? win_lo - 1 : 0u);
        if (blockIdx.x == nblk - 1 && threadIdx.x == 0) *n_steps_out = n_steps;
    }

    double acc = 0.0;
#pragma unroll
    for (int it = 0; it < NOALLOC_SAMPLES_PER_BLOCK / (4 * NOALLOC_THREADS); ++it) {
        const uint32_t i0 = block_lo + (uint32_t)it * (4 * NOALLOC_THREADS) + 4u * threadIdx.x;
        if (i0 >= n_steps) continue;
        float v[4];
        const float base = (float)i0;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float fi = __fadd_rn(base, (float)e);
            const float dl = brp_del_t(fi, tp.tau, tp.Omega, tp.Psi0, tp.S0, dt, step_inv, tab);
            const float f = __fsub_rn(fi, dl);
            int k = FAST ? brp_round_index_fast(f) : brp_round_index_exact(f);
            k = min(k, (int)nu - 1);
            v[e] = __ldg(series + k);
        }
        if (i0 + 3 < n_steps) {
            *reinterpret_cast<float4 *>(out + i0) = make_float4(v[0], v[1], v[2], v[3]);
            acc += (double)((v[0] + v[1]) + (v[2] + v[3]));
        } else {
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (i0 + e < n_steps) {
                    out[i0 + e] = v[e];
                    acc += (double)v[e];
                }
        }
    }

    acc = warp_total(acc);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < NOALLOC_THREADS / 32; ++w) t += s_warp[w];
        partials[blockIdx.x] = t;
        __threadfence();
        s_last = (atomicAdd(counter, 1u) == nblk - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    double t = 0.0;
    for (uint32_t k = threadIdx.x; k < nblk; k += NOALLOC_THREADS) t += __ldcg(&partials[k]);
    t = warp_total(t);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
        double tot = 0.0;
#pragma unroll
        for (int w = 0; w < NOALLOC_THREADS / 32; ++w) tot += s_warp[w];
        *mean_out = (float)(tot / (double)__ldcg(n_steps_out));
    }
}

__global__ void noalloc_pad_kernel(float *__restrict__ out, const uint32_t *__restrict__ n_steps, const float *__restrict__ mean,
                                   uint32_t n)
{
    const float m = *mean;
    for (uint32_t i = *n_steps + blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = m;
}

// once per device: turn the host-side stand-in for g_noalloc_scratch into a device address
int scratch_for_current_device(void **addr)
{
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return RADPUL_CUDA_DEVICE_FIND;
    if (dev < 0 || dev >= kMaxDevices) return RADPUL_CUDA_DEVICE_FIND;
    std::lock_guard<std::mutex> guard(g_setup_lock);
    if (!g_scratch_addr[dev]) {
        void *p = NULL;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            fprintf(stderr, "[brp_b200] brp_resample_noalloc: cudaGetSymbolAddress failed on device %d: %s\n", dev, cudaGetErrorString(e));
            return RADPUL_CUDA_DEVICE_SET;
        }
        g_scratch_addr[dev] = p;
    }
    *addr = g_scratch_addr[dev];
    return 0;
}

} // namespace

extern "C" int brp_resample_noalloc(const void *d_series, uint32_t nsamples_unpadded, uint32_t nsamples, const brp_template *t,
                                    float dt, float step_inv, void *d_out, uint32_t *d_n_steps, float *d_mean, void *stream)
{
    if (!d_series || !t || !d_out || !d_n_steps || !d_mean || nsamples_unpadded < 2 || nsamples < nsamples_unpadded) return RADPUL_EVAL;
    if (reinterpret_cast<uintptr_t>(d_out) & 15u) return RADPUL_EVAL; // rows are written with 128-bit stores
    if (nsamples_unpadded > BRP_NOALLOC_MAX_SAMPLES) {
        fprintf(stderr, "[brp_b200] brp_resample_noalloc: series of %u samples exceeds the built-in scratch (at most %u samples); "
                        "use brp_search_create, which sizes its scratch with cudaMalloc\n", nsamples_unpadded, (unsigned)BRP_NOALLOC_MAX_SAMPLES);
        return RADPUL_EVAL;
    }
    void *scratch = NULL;
    const int rc = scratch_for_current_device(&scratch);
    if (rc) return rc;
    double *partials = reinterpret_cast<double *>(scratch);
    uint32_t *counter = reinterpret_cast<uint32_t *>(static_cast<unsigned char *>(scratch) + NOALLOC_COUNTER_OFFSET);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // the array still holds the previous call's state: the counter must start at zero
    if (cudaMemsetAsync(counter, 0, sizeof(uint32_t), st) != cudaSuccess) return RADPUL_CUDA_KERNEL_INVOKE;
    const BrpTemplateDev tp = {t->tau, t->Omega, t->Psi0, t->S0};
    const uint32_t nblk = (nsamples_unpadded + NOALLOC_SAMPLES_PER_BLOCK - 1) / NOALLOC_SAMPLES_PER_BLOCK;
    float *out = static_cast<float *>(d_out);
    const float *series = static_cast<const float *>(d_series);
    if (nsamples_unpadded <= (1u << 22))
        noalloc_resample_kernel<true><<<nblk, NOALLOC_THREADS, 0, st>>>(series, tp, nsamples_unpadded, dt, step_inv, out, d_n_steps, d_mean, partials, counter);
    else
        noalloc_resample_kernel<false><<<nblk, NOALLOC_THREADS, 0, st>>>(series, tp, nsamples_unpadded, dt, step_inv, out, d_n_steps, d_mean, partials, counter);
    if (cudaGetLastError() != cudaSuccess) return RADPUL_CUDA_KERNEL_INVOKE;
    noalloc_pad_kernel<<<64, 256, 0, st>>>(out, d_n_steps, d_mean, nsamples);
    if (cudaGetLastError() != cudaSuccess) return RADPUL_CUDA_KERNEL_INVOKE;
    return 0;
}
