// Kernel 1: resampling gather (north_star kernel 1).
//
// Replaces, per template, the reference's run_resampling (demod_binary_resamp_cpu.c:80-136; the
// reference CUDA backend needs six kernels and two blocking device->host copies for the same thing,
// cuda/app/demod_binary_cuda.cu:432-802):
//   * del_t[i] is evaluated in registers and never stored;
//   * n_steps (:105-109) is found by the few blocks that overlap the tail window, with a block-level
//     max reduction instead of a serial scan;
//   * out[i] = in[nearest(i)] for i < n_steps (:113-121), four consecutive samples per thread, one
//     128-bit store; the gather loads are nearly contiguous (index slope 1 +- 0.003) and the series
//     is shared by every template of the batch, so they are served from L1/L2;
//   * the time-series mean (:120,125) is accumulated in double: per-block partial sums are written to
//     scratch and reduced in a fixed order by whichever block of the template finishes last, so the
//     result does not depend on scheduling (the reference's serial float32 sum is documented, not
//     reproduced: SURVEY.md 9.4);
//   * the mean padding (:130-133) is not materialised in batched mode: the FFT's first pass substitutes
//     mean for i >= n_steps on load.  The legacy entry point asks for it to be written (fill_padding).
//
// grid = (ceil(Nu/4096), batch), block = 256.  HBM-bound: 4*Nu read (L2-resident across the batch) +
// 4*n_steps written per template.
#include "brp_device.cuh"
#include "brp_internal.h"

namespace {

struct ConstTable {
    __device__ __forceinline__ float2 operator[](int i) const { return c_brp_sincos[i]; }
};

struct SmemTable {
    const float2 *p;
    __device__ __forceinline__ float2 operator[](int i) const { return p[i]; }
};

template <bool FAST>
__device__ __forceinline__ int round_index(float f)
{
    return FAST ? brp_round_index_fast(f) : brp_round_index_exact(f);
}

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

template <bool FAST, bool WRITE_IDX>
__global__ void __launch_bounds__(BRP_RESAMP_THREADS)
resample_kernel(const float *__restrict__ series, const BrpTemplateDev *__restrict__ tmpl, BrpGeomDev g,
                float *__restrict__ out, size_t out_stride, int32_t *__restrict__ idx_out,
                uint32_t *__restrict__ n_steps_out, float *__restrict__ mean_out, double *__restrict__ partials,
                uint32_t *__restrict__ counters)
{
    __shared__ float2 s_tab[65];
    __shared__ int s_nsteps;
    __shared__ double s_warp[BRP_RESAMP_THREADS / 32];
    __shared__ bool s_last;

    const int b = blockIdx.y;
    const uint32_t nblk = gridDim.x;
    const uint32_t block_lo = blockIdx.x * BRP_RESAMP_SAMPLES_PER_BLOCK;
    const BrpTemplateDev tp = tmpl[b];
    const uint32_t nu = g.nu;

    if (threadIdx.x < 65) s_tab[threadIdx.x] = c_brp_sincos[threadIdx.x];
    if (threadIdx.x == 0) s_nsteps = -1;
    __syncthreads();
    const SmemTable tab{s_tab};

    // ---- n_steps = max{ n <= Nu-1 : (float)n - del_t[n] < (float)(Nu-1) }  (demod_binary_resamp_cpu.c:105-109)
    // |del_t| <= |tau|*step_inv*(|sin|max + 1) so only the last `reach` samples can fail the test.
    const uint32_t reach = (uint32_t)(2.02f * fabsf(tp.tau) * g.step_inv) + 16u;
    const uint32_t win_lo = (nu - 1 > reach) ? (nu - 1 - reach) : 0u;
    uint32_t n_steps = nu; // "everything in this block is below n_steps"
    const bool tail_block = block_lo + BRP_RESAMP_SAMPLES_PER_BLOCK > win_lo;
    if (tail_block) {
        const float limit = (float)(nu - 1);
        int best = -1;
        for (uint32_t n = win_lo + threadIdx.x; n < nu; n += BRP_RESAMP_THREADS) {
            const float nf = (float)n;
            const float dl = brp_del_t(nf, tp.tau, tp.Omega, tp.Psi0, tp.S0, g.dt, g.step_inv, tab);
            if (__fsub_rn(nf, dl) < limit) best = (int)n; // n increases along the loop
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
        if ((threadIdx.x & 31) == 0 && best >= 0) atomicMax(&s_nsteps, best);
        __syncthreads();
        // the window always holds a sample that passes; the fallback only guards degenerate parameters
        n_steps = s_nsteps >= 0 ? (uint32_t)s_nsteps : (win_lo > 0 ? win_lo - 1 : 0u);
        if (blockIdx.x == nblk - 1 && threadIdx.x == 0) n_steps_out[b] = n_steps;
    }

    // ---- gather: out[i] = series[(int)(i - del_t[i] + 0.5)] for i < n_steps  (demod_binary_resamp_cpu.c:113-121)
    float *const orow = out + (size_t)b * out_stride;
    double acc = 0.0;
    // the packed path needs |Omega*t + Psi0| / 2pi < 2^22 for every sample: true for any physical template
    const bool packed_ok = (fabsf(tp.Omega) * ((float)nu * g.dt) + fabsf(tp.Psi0)) * BRP_TWO_PI_INV < 4.0e6f;
#pragma unroll
    for (int it = 0; it < BRP_RESAMP_SAMPLES_PER_BLOCK / (4 * BRP_RESAMP_THREADS); ++it) {
        const uint32_t i0 = block_lo + (uint32_t)it * (4 * BRP_RESAMP_THREADS) + 4u * threadIdx.x;
        if (i0 >= n_steps) continue;
        float v[4];
        int id[4];
        const float base = (float)i0;
        if (FAST && packed_ok) {
            // two samples per instruction (FADD2/FMUL2/FFMA2), bit-identical to the scalar path
#pragma unroll
            for (int e = 0; e < 4; e += 2) {
                const float2 fi = brp_add2(brp_pair(base), make_float2((float)e, (float)(e + 1)));
                const float2 dl = brp_del_t2(fi, tp.tau, tp.Omega, tp.Psi0, tp.S0, g.dt, g.step_inv, tab);
                const int2 k = brp_round_index_fast2(brp_sub2(fi, dl));
                id[e] = min(k.x, (int)nu - 1);
                id[e + 1] = min(k.y, (int)nu - 1);
            }
        } else {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float fi = __fadd_rn(base, (float)e); // exact: integers below 2^24
                const float dl = brp_del_t(fi, tp.tau, tp.Omega, tp.Psi0, tp.S0, g.dt, g.step_inv, tab);
                const int k = round_index<FAST>(__fsub_rn(fi, dl));
                id[e] = min(k, (int)nu - 1); // the reference would read past the array here; keep the access in bounds
            }
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) v[e] = __ldg(series + id[e]);
        if (i0 + 3 < n_steps) {
            *reinterpret_cast<float4 *>(orow + i0) = make_float4(v[0], v[1], v[2], v[3]);
            acc += (double)((v[0] + v[1]) + (v[2] + v[3]));
            if (WRITE_IDX) *reinterpret_cast<int4 *>(idx_out + (size_t)b * nu + i0) = make_int4(id[0], id[1], id[2], id[3]);
        } else {
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (i0 + e < n_steps) {
                    orow[i0 + e] = v[e];
                    acc += (double)v[e];
                    if (WRITE_IDX) idx_out[(size_t)b * nu + i0 + e] = id[e];
                }
        }
    }

    // ---- mean: fixed-order block reduction, per-block partials, last block of the template finishes
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < BRP_RESAMP_THREADS / 32; ++w) t += s_warp[w];
        partials[(size_t)b * nblk + blockIdx.x] = t;
        __threadfence();
        const uint32_t done = atomicAdd(&counters[b], 1u);
        s_last = (done == nblk - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    double t = 0.0;
    for (uint32_t k = threadIdx.x; k < nblk; k += BRP_RESAMP_THREADS) t += __ldcg(&partials[(size_t)b * nblk + k]);
    t = warp_sum(t);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
        double tot = 0.0;
#pragma unroll
        for (int w = 0; w < BRP_RESAMP_THREADS / 32; ++w) tot += s_warp[w];
        const uint32_t ns = __ldcg(&n_steps_out[b]);
        mean_out[b] = (float)(tot / (double)ns); // mean /= (float)n_steps  (:125)
        counters[b] = 0u;                        // ready for the next launch
    }
}

__global__ void pad_kernel(float *__restrict__ out, size_t out_stride, const uint32_t *__restrict__ n_steps,
                           const float *__restrict__ mean, uint32_t n)
{
    const int b = blockIdx.y;
    const float m = mean[b];
    float *row = out + (size_t)b * out_stride;
    for (uint32_t i = n_steps[b] + blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) row[i] = m;
}

__global__ void sin_lut_kernel(const float *__restrict__ x, float *__restrict__ out, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = brp_sin_lut(x[i], ConstTable{});
}

__global__ void round_index_kernel(const float *__restrict__ f, int32_t *__restrict__ out, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float v = f[i];
    // same selection rule as the resampler: the FP32 path whenever the value is below 2^23
    out[i] = (v < 8388608.0f) ? brp_round_index_fast(v) : brp_round_index_exact(v);
}

} // namespace

cudaError_t brp_launch_resample(const float *series, const BrpTemplateDev *tmpl, int batch, const BrpGeomDev &g,
                                float *out, size_t out_stride, int32_t *idx_out, uint32_t *n_steps, float *mean,
                                double *partials, uint32_t *counters, bool fill_padding, cudaStream_t st)
{
    const dim3 grid(brp_resamp_blocks(g.nu), batch);
    // the FP32 index path needs every i - del_t below 2^23; |del_t| is far below 2^22 for any sane template
    const bool fast = g.nu <= (1u << 22);
    if (fast) {
        if (idx_out)
            resample_kernel<true, true><<<grid, BRP_RESAMP_THREADS, 0, st>>>(series, tmpl, g, out, out_stride, idx_out, n_steps, mean, partials, counters);
        else
            resample_kernel<true, false><<<grid, BRP_RESAMP_THREADS, 0, st>>>(series, tmpl, g, out, out_stride, idx_out, n_steps, mean, partials, counters);
    } else {
        if (idx_out)
            resample_kernel<false, true><<<grid, BRP_RESAMP_THREADS, 0, st>>>(series, tmpl, g, out, out_stride, idx_out, n_steps, mean, partials, counters);
        else
            resample_kernel<false, false><<<grid, BRP_RESAMP_THREADS, 0, st>>>(series, tmpl, g, out, out_stride, idx_out, n_steps, mean, partials, counters);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    if (fill_padding) {
        pad_kernel<<<dim3(64, batch), 256, 0, st>>>(out, out_stride, n_steps, mean, g.n);
        e = cudaGetLastError();
    }
    return e;
}

cudaError_t brp_launch_sin_lut(const float *x, float *out, uint32_t n, cudaStream_t st)
{
    sin_lut_kernel<<<(n + 255) / 256, 256, 0, st>>>(x, out, n);
    return cudaGetLastError();
}

cudaError_t brp_launch_round_index(const float *f, int32_t *out, uint32_t n, cudaStream_t st)
{
    round_index_kernel<<<(n + 255) / 256, 256, 0, st>>>(f, out, n);
    return cudaGetLastError();
}
