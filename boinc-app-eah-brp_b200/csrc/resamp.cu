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
// Run detection (MODE 2, the production path).  The shift idx(i) - i changes once per ~1000 samples only, so the
// timing model is evaluated at every 16th sample (one per lane and group of sixteen) and the fifteen samples in
// between are proven, not computed:
//   (1) G(i, x) = (int)(fl(i - x) + 0.5) is non-increasing in x (float subtraction and the rounding are monotone), so
//       lo <= del_t[i] <= hi  and  G(i, lo) == G(i, hi)  pin idx(i) without evaluating del_t[i];
//   (2) for integer i below 2^23 and i - x inside one binade [2^e, 2^(e+1)), fl(i - x) = i - c(x, e): i is an even
//       multiple of the binade's spacing, so round-to-nearest-even of i - x does not depend on i.  Hence
//       G(i, x) - i is the same for every i of the group and is computed once, at the group's first sample;
//   (3) del_t of the inner samples lies between the values at the two ends of the group, up to rounding: the computed
//       phase fl(fl(Omega*fl(i*dt)) + Psi0) is monotone in i, so is the expansion variable d while the table cell
//       does not change, and the table polynomial is monotone in d over a group (at the two cells where its
//       derivative changes sign the values may leave the interval of the two ends by at most (group phase span)^2 / 8,
//       which the precondition span < 3e-3 rad keeps below 1.2e-6; physical orbits span 1e-5 rad).  What is not monotone is the rounding of the seven
//       operations behind d: at most 3.7e-7 * |tau|*step_inv per value.  eta = 4e-6*|tau|*step_inv + 1e-6 is twice the sum
//       of two such roundings (7.4e-7) and the excursion (1.2e-6), all times |tau|*step_inv.
// A group falls back to evaluating all of its samples when its ends lie in different table cells, when the two
// bounds of (1) disagree (a shift change may lie inside), when i - x crosses a binade or is below 1.  A warp owns 512
// consecutive samples per step; when all 32 groups are proven with the same shift (about 60 % of the steps) the step is a
// plain displaced copy with one alignment case for the whole warp.  The tests
// compare the index arrays with the oracle bit for bit, and brp_stage_resample_compare() checks MODE 2 against MODE 1
// sample by sample on the device for whole banks.
//
// grid = (ceil(Nu/16384), batch), block = 256: a warp walks four adjacent 512-sample segments.  HBM traffic: 4*Nu read
// (L2-resident across the batch) + 4*n_steps written per template.
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

constexpr int kGroupLen = BRP_RESAMP_SAMPLES_PER_BLOCK / BRP_RESAMP_THREADS; // 16 samples per timing-model evaluation
constexpr int kGroups = BRP_RESAMP_THREADS;                                    // groups per chunk of 4096 samples
static_assert(kGroupLen == 16 && 128 % kGroupLen == 0, "the lane mapping below assumes 16-sample groups");
// A block walks kChunks chunks: the block-level epilogue (reduction, fence, atomic) is a serial chain of a few
// microseconds, and with one chunk per block that chain, not bandwidth or issue slots, set the kernel's duration.
// (One or two templates per launch -- the reference's one-template-at-a-time interface -- are the exception: 256 blocks
// per template leave most of the 148 SMs' slots empty, so those launches run with one chunk per block.)
constexpr int kChunksBatched = 4;

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

// MODE 0: index by the literal double arithmetic (series longer than 2^22); 1: FP32 index path, every sample
// evaluated; 2: FP32 index path with run detection.
#ifndef BRP_RESAMP_MIN_BLOCKS
#define BRP_RESAMP_MIN_BLOCKS 8
#endif
template <int MODE, bool WRITE_IDX, int kChunks = kChunksBatched>
__global__ void __launch_bounds__(BRP_RESAMP_THREADS, BRP_RESAMP_MIN_BLOCKS)
resample_kernel(const float *__restrict__ series, const BrpTemplateDev *__restrict__ tmpl, BrpGeomDev g,
                float *__restrict__ out, size_t out_stride, int32_t *__restrict__ idx_out,
                uint32_t *__restrict__ n_steps_out, float *__restrict__ mean_out, double *__restrict__ partials,
                uint32_t *__restrict__ counters, const BrpTemplateDev tmpl_value)
{
    constexpr bool FAST = MODE != 0;
    constexpr uint32_t kBlockSpan = (uint32_t)kChunks * BRP_RESAMP_SAMPLES_PER_BLOCK; // 8 warps x kChunks segments of 512
    __shared__ float2 s_tab[65];
    __shared__ int s_nsteps;
    __shared__ double s_warp[BRP_RESAMP_THREADS / 32];
    __shared__ unsigned int s_arrived;

    const int b = blockIdx.y;
    const uint32_t nblk = gridDim.x;
    const uint32_t block_lo = blockIdx.x * kBlockSpan;
    const BrpTemplateDev tp = tmpl ? tmpl[b] : tmpl_value; // one template handed over as a kernel argument: no upload before the launch
    const uint32_t nu = g.nu;

    if (threadIdx.x < 65) s_tab[threadIdx.x] = c_brp_sincos[threadIdx.x];
    if (threadIdx.x == 0) {
        s_nsteps = -1;
        s_arrived = 0u;
    }
    __syncthreads();
    const SmemTable tab{s_tab};

    // ---- n_steps = max{ n <= Nu-1 : (float)n - del_t[n] < (float)(Nu-1) }  (demod_binary_resamp_cpu.c:105-109)
    // |del_t| <= |tau|*step_inv*(|sin|max + 1) so only the last `reach` samples can fail the test.
    const uint32_t reach = (uint32_t)(2.02f * fabsf(tp.tau) * g.step_inv) + 16u;
    const uint32_t win_lo = (nu - 1 > reach) ? (nu - 1 - reach) : 0u;
    uint32_t n_steps = nu; // "everything in this block is below n_steps"
    const bool tail_block = block_lo + kBlockSpan > win_lo;
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
    int32_t *const irow = WRITE_IDX ? idx_out + (size_t)b * nu : nullptr;
    double acc = 0.0;
    // the packed path needs |Omega*t + Psi0| / 2pi < 2^22 for every sample: true for any physical template
    const bool packed_ok = (fabsf(tp.Omega) * ((float)nu * g.dt) + fabsf(tp.Psi0)) * BRP_TWO_PI_INV < 4.0e6f;
    // run detection additionally needs a group to span far less than one table cell (1/64 cycle)
    const float amp = fabsf(tp.tau) * g.step_inv;
    const bool runs_ok = MODE == 2 && packed_ok && fabsf(tp.Omega) * g.dt * (float)kGroupLen < 3.0e-3f && amp < 4.0e6f;
    const float eta = 4.0e-6f * amp + 1.0e-6f;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    // four samples i0..i0+3 by the timing model itself
    auto four_samples = [&](uint32_t i0, float (&v)[4], int (&id)[4]) {
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
    };
    auto emit = [&](uint32_t i0, const float (&v)[4], const int (&id)[4]) {
        if (i0 + 3 < n_steps) {
            *reinterpret_cast<float4 *>(orow + i0) = make_float4(v[0], v[1], v[2], v[3]);
            acc += (double)((v[0] + v[1]) + (v[2] + v[3]));
            if (WRITE_IDX) *reinterpret_cast<int4 *>(irow + i0) = make_int4(id[0], id[1], id[2], id[3]);
        } else {
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (i0 + e < n_steps) {
                    orow[i0 + e] = v[e];
                    acc += (double)v[e];
                    if (WRITE_IDX) irow[i0 + e] = id[e];
                }
        }
    };

    // A warp owns kChunks consecutive segments of 512 samples; in round r lane l handles the four samples at 128 r + 4 l.
    // One timing-model evaluation per group of kGroupLen samples (lane l: group l of the segment).  The right end of a
    // group is its neighbour's left end, fetched by shuffle; the right end of the segment's last group is lane 0's
    // evaluation for the NEXT segment, which is computed one iteration ahead.  No shared memory, no block barrier.
    const uint32_t warp_lo = block_lo + (uint32_t)warp * (kChunks * 32 * kGroupLen);
    float d_next = 0.0f;
    int cell_next = 0;
    if (MODE == 2 && runs_ok && warp_lo < n_steps)
        d_next = brp_del_t_cell((float)(warp_lo + (uint32_t)lane * kGroupLen), tp.tau, tp.Omega, tp.Psi0, tp.S0, g.dt, g.step_inv, tab, cell_next);
#pragma unroll 1
    for (int chunk = 0; chunk < kChunks; ++chunk) {
        const uint32_t seg_lo = warp_lo + (uint32_t)chunk * (32 * kGroupLen);
        if (seg_lo >= n_steps) break; // uniform within the warp
        if (MODE != 2 || !runs_ok) {
#pragma unroll
            for (int r = 0; r < kGroupLen / 4; ++r) {
                const uint32_t i0 = seg_lo + 128u * r + 4u * lane;
                if (i0 >= n_steps) continue;
                float v[4];
                int id[4];
                four_samples(i0, v, id);
                emit(i0, v, id);
            }
            continue;
        }
        if constexpr (MODE == 2) {
            const uint32_t g_lo = seg_lo + (uint32_t)lane * kGroupLen;
            const float d_own = d_next;
            const int cell = cell_next;
            d_next = brp_del_t_cell((float)(g_lo + 32 * kGroupLen), tp.tau, tp.Omega, tp.Psi0, tp.S0, g.dt, g.step_inv, tab, cell_next);
            float2 eb;
            eb.x = __shfl_down_sync(0xffffffffu, d_own, 1);
            eb.y = __int_as_float(__shfl_down_sync(0xffffffffu, cell, 1));
            const float d_first = __shfl_sync(0xffffffffu, d_next, 0);
            const int cell_first = __shfl_sync(0xffffffffu, cell_next, 0);
            if (lane == 31) {
                eb.x = d_first;
                eb.y = __int_as_float(cell_first);
            }
            const float base = (float)g_lo;
            const float lo = __fsub_rn(fminf(d_own, eb.x), eta), hi = __fadd_rn(fmaxf(d_own, eb.x), eta);
            const float a = __fsub_rn(base, hi);                                        // smallest i - x of the group
            const float z = __fsub_rn(__fadd_rn(base, (float)(kGroupLen - 1)), lo);     // largest
            const int ka = brp_round_index_fast(a), kb = brp_round_index_fast(__fsub_rn(base, lo));
            const bool proven = cell == __float_as_int(eb.y) && ka == kb && a >= 1.0f &&
                                ((__float_as_int(a) ^ __float_as_int(z)) & 0x7F800000) == 0;
            const int verdict = proven ? ka : -1; // index of the group's first sample, or "evaluate every sample"
            // Most segments (no shift change among their 512 samples) are a plain copy displaced by one warp-uniform shift:
            // pointers and the alignment case are then computed once per segment instead of once per four samples.
            const int sh = verdict - (int)g_lo, sh0 = __shfl_sync(0xffffffffu, sh, 0);
            const int k_seg = (int)seg_lo + sh0;
            const bool whole = seg_lo + 32u * kGroupLen <= n_steps && k_seg >= 0 && (uint32_t)k_seg + 32u * kGroupLen + 4u <= nu;
            if (__all_sync(0xffffffffu, verdict >= 0 && sh == sh0) && whole) {
                const float4 *s4 = reinterpret_cast<const float4 *>(series + (k_seg & ~3)) + lane;
                float4 *o4 = reinterpret_cast<float4 *>(orow + seg_lo) + lane;
                const int al = k_seg & 3;
#pragma unroll
                for (int r = 0; r < kGroupLen / 4; ++r) {
                    const float4 q0 = __ldg(s4 + 32 * r), q1 = __ldg(s4 + 32 * r + 1);
                    float4 o;
                    if (al == 0) o = q0;
                    else if (al == 1) o = make_float4(q0.y, q0.z, q0.w, q1.x);
                    else if (al == 2) o = make_float4(q0.z, q0.w, q1.x, q1.y);
                    else o = make_float4(q0.w, q1.x, q1.y, q1.z);
                    o4[32 * r] = o;
                    acc += (double)((o.x + o.y) + (o.z + o.w));
                    if (WRITE_IDX) {
                        const int k = k_seg + 128 * r + 4 * lane;
                        *reinterpret_cast<int4 *>(irow + seg_lo + 128u * r + 4u * lane) = make_int4(k, k + 1, k + 2, k + 3);
                    }
                }
                continue;
            }
#pragma unroll
            for (int r = 0; r < kGroupLen / 4; ++r) {
                const int src = r * (128 / kGroupLen) + lane / (kGroupLen / 4);
                const int first = __shfl_sync(0xffffffffu, verdict, src);
                const uint32_t i0 = seg_lo + 128u * r + 4u * lane;
                if (i0 >= n_steps) continue;
                float v[4];
                int id[4];
                if (first >= 0) {
                    const int k = first + 4 * (lane % (kGroupLen / 4));
#pragma unroll
                    for (int e = 0; e < 4; ++e) id[e] = min(k + e, (int)nu - 1);
                    // four consecutive input samples: two aligned 128-bit loads instead of four scalar ones
                    const int k4 = k & ~3;
                    if (k4 + 8 <= (int)nu) {
                        const float4 *p4 = reinterpret_cast<const float4 *>(series + k4);
                        const float4 q0 = __ldg(p4), q1 = __ldg(p4 + 1);
                        switch (k & 3) {
                        case 0: v[0] = q0.x; v[1] = q0.y; v[2] = q0.z; v[3] = q0.w; break;
                        case 1: v[0] = q0.y; v[1] = q0.z; v[2] = q0.w; v[3] = q1.x; break;
                        case 2: v[0] = q0.z; v[1] = q0.w; v[2] = q1.x; v[3] = q1.y; break;
                        default: v[0] = q0.w; v[1] = q1.x; v[2] = q1.y; v[3] = q1.z; break;
                        }
                    } else {
#pragma unroll
                        for (int e = 0; e < 4; ++e) v[e] = __ldg(series + id[e]);
                    }
                } else {
                    four_samples(i0, v, id);
                }
                emit(i0, v, id);
            }
        }
    } // chunks

    // ---- mean: fixed-order reduction.  Warp sums -> block partial (by whichever warp of the block finishes last) ->
    // template total (by whichever block finishes last).  The order of every sum is fixed by index, never by arrival, so
    // the mean does not depend on scheduling; no block-wide barrier: warps that are done leave.
    acc = warp_sum(acc);
    int last_warp = 0;
    if (lane == 0) {
        s_warp[warp] = acc;
        __threadfence(); // this warp's global stores (the series, n_steps) and its sum before the arrival is counted
        last_warp = atomicAdd(&s_arrived, 1u) == BRP_RESAMP_THREADS / 32 - 1;
    }
    if (!__shfl_sync(0xffffffffu, last_warp, 0)) return;
    int last_block = 0;
    if (lane == 0) {
        __threadfence_block();
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < BRP_RESAMP_THREADS / 32; ++w) t += ((volatile double *)s_warp)[w];
        partials[(size_t)b * nblk + blockIdx.x] = t;
        __threadfence();
        last_block = atomicAdd(&counters[b], 1u) == nblk - 1;
    }
    if (!__shfl_sync(0xffffffffu, last_block, 0)) return;
    __threadfence();
    double t = 0.0;
    for (uint32_t k = lane; k < nblk; k += 32) t += __ldcg(&partials[(size_t)b * nblk + k]);
    t = warp_sum(t);
    if (lane == 0) {
        const uint32_t ns = __ldcg(&n_steps_out[b]);
        mean_out[b] = (float)(t / (double)ns); // mean /= (float)n_steps  (:125)
        counters[b] = 0u;                      // ready for the next launch
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

__global__ void count_mismatch_kernel(const int32_t *__restrict__ a, const int32_t *__restrict__ b, uint32_t n,
                                      unsigned long long *__restrict__ count)
{
    unsigned int bad = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) bad += a[i] != b[i];
    bad = __reduce_add_sync(0xffffffffu, bad);
    if ((threadIdx.x & 31) == 0 && bad) atomicAdd(count, (unsigned long long)bad);
}

} // namespace

cudaError_t brp_launch_count_mismatch(const int32_t *a, const int32_t *b, uint32_t n, unsigned long long *count, cudaStream_t st)
{
    count_mismatch_kernel<<<296, 256, 0, st>>>(a, b, n, count);
    return cudaGetLastError();
}

cudaError_t brp_launch_resample(const float *series, const BrpTemplateDev *tmpl, int batch, const BrpGeomDev &g,
                                float *out, size_t out_stride, int32_t *idx_out, uint32_t *n_steps, float *mean,
                                double *partials, uint32_t *counters, bool fill_padding, cudaStream_t st, bool every_sample,
                                const BrpTemplateDev *tmpl_value)
{
    if (!tmpl && (!tmpl_value || batch != 1)) return cudaErrorInvalidValue;
    const BrpTemplateDev tv = tmpl_value ? *tmpl_value : BrpTemplateDev{};
    constexpr uint32_t kBlockSpan = (uint32_t)kChunksBatched * BRP_RESAMP_SAMPLES_PER_BLOCK;
    const dim3 grid((g.nu + kBlockSpan - 1) / kBlockSpan, batch);
    // the FP32 index path needs every i - del_t below 2^23; |del_t| is far below 2^22 for any sane template
    const bool fast = g.nu <= (1u << 22);
    if (fast && !every_sample && !idx_out && batch <= 2) {
        const dim3 fine(brp_resamp_blocks(g.nu), batch);
        resample_kernel<2, false, 1><<<fine, BRP_RESAMP_THREADS, 0, st>>>(series, tmpl, g, out, out_stride, idx_out, n_steps, mean, partials, counters, tv);
    } else if (fast && !every_sample) {
        if (idx_out)
            resample_kernel<2, true><<<grid, BRP_RESAMP_THREADS, 0, st>>>(series, tmpl, g, out, out_stride, idx_out, n_steps, mean, partials, counters, tv);
        else
            resample_kernel<2, false><<<grid, BRP_RESAMP_THREADS, 0, st>>>(series, tmpl, g, out, out_stride, idx_out, n_steps, mean, partials, counters, tv);
    } else if (fast) {
        if (idx_out)
            resample_kernel<1, true><<<grid, BRP_RESAMP_THREADS, 0, st>>>(series, tmpl, g, out, out_stride, idx_out, n_steps, mean, partials, counters, tv);
        else
            resample_kernel<1, false><<<grid, BRP_RESAMP_THREADS, 0, st>>>(series, tmpl, g, out, out_stride, idx_out, n_steps, mean, partials, counters, tv);
    } else {
        if (idx_out)
            resample_kernel<0, true><<<grid, BRP_RESAMP_THREADS, 0, st>>>(series, tmpl, g, out, out_stride, idx_out, n_steps, mean, partials, counters, tv);
        else
            resample_kernel<0, false><<<grid, BRP_RESAMP_THREADS, 0, st>>>(series, tmpl, g, out, out_stride, idx_out, n_steps, mean, partials, counters, tv);
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
