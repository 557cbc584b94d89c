// Kernel 3: fused harmonic summing + per-bin toplist (north_star kernel 3).
//
// Replaces run_harmonic_summing (demod_binary_hs_cpu.c:62-86 -> hs_common.c:33-171) AND the per-template
// candidate scan of the driver (demod_binary.c:1310-1397) for a whole batch of templates.
//
// For every top-harmonic bin i in [w2, H) the reference forms the running sums
//     S_0 = P[i]
//     S_1 = S_0 +  P[(8i+8)>>4]
//     S_2 = S_1 + (P[(12i+8)>>4] + P[(4i+8)>>4])
//     S_3 = S_2 + (((P[(14i+8)>>4] + P[(10i+8)>>4]) + P[(6i+8)>>4]) + P[(2i+8)>>4])
//     S_4 = S_3 + (((((((P[(15i+8)>>4] + P[13..]) + P[11..]) + P[9..]) + P[7..]) + P[5..]) + P[3..]) + P[(i+8)>>4])
// (same association, hs_common.c:107,125-126,145-148) and keeps, per harmonic k and fundamental bin
// j_k = (i*2^(4-k) + 8) >> 4, the maximum over the i that map to j_k.  Only maxima above the threshold are
// ever looked at by the candidate scan (demod_binary.c:1350), and "keep the maximum" commutes, so the
// segment maximum, the maximum over templates and the toplist's same-bin rule (:1356-1363, earliest
// template wins ties) collapse into ONE 64-bit atomicMax per above-threshold value on a device-resident
// table  key[k][j] = float_bits(power) << 32 | (0xFFFFFFFF - template index).
// Above-threshold values are rare (false-alarm rate ~1e-7 per bin): nothing is written in the common case.
//
// Data movement.  A CTA owns kSpan consecutive bins i.  Harmonic l of those bins lives in ONE contiguous
// slice of the spectrum, P[(l*i0+8)>>4 .. (l*i1+8)>>4] (about l*kSpan/16 floats), so the sixteen slices are
// brought into shared memory by sixteen 1-D bulk async copies (TMA, cp.async.bulk + mbarrier): no load
// instructions, no registers, no L1 wavefronts on the way in.  Consecutive lanes then read consecutive or
// identical shared-memory words (conflict-free, broadcast), and because a thread's bins are 256 apart,
// the slice index of its next bin is the previous one plus the constant 16*l: one index computation per
// harmonic and thread, immediate offsets afterwards.
// Algorithmic traffic 4*H bytes per template; the slices of different harmonics overlap in the spectrum,
// so L2 serves 8.5*H floats per template and HBM about H.
#include "brp_internal.h"

namespace {

constexpr int kHsThreads = 256;
constexpr int kHsPerThread = 4;                       // bins per thread, kHsThreads apart
constexpr int kSpan = kHsThreads * kHsPerThread;      // bins per CTA
// slice l holds at most l*kSpan/16 + 2 floats, rounded up to whole 16-byte units on both ends
__host__ __device__ constexpr int slice_cap(int l) { return ((l * kSpan / 16 + 2 + 3 + 3) / 4) * 4; }
__host__ __device__ constexpr int slice_off(int l)
{
    int o = 0;
    for (int m = 1; m < l; ++m) o += slice_cap(m);
    return o;
}
constexpr int kSmemFloats = slice_off(17);

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void offer(unsigned long long *slot, float power, uint32_t tkey)
{
    atomicMax(slot, ((unsigned long long)__float_as_uint(power) << 32) | tkey);
}

__global__ void __launch_bounds__(kHsThreads)
harmonic_summing_kernel(const float *__restrict__ ps, size_t ps_stride, BrpGeomDev g, BrpThresholds thr,
                        uint32_t first_template, unsigned long long *__restrict__ keys)
{
    extern __shared__ __align__(128) float sm[];
    __shared__ __align__(8) unsigned long long bar;

    const float *__restrict__ P = ps + (size_t)blockIdx.y * ps_stride;
    const uint32_t tkey = 0xFFFFFFFFu - (first_template + blockIdx.y);
    const uint32_t fu = g.fu;
    const uint32_t i0 = g.w2 + blockIdx.x * kSpan;           // first bin of this CTA
    const uint32_t i1 = min(i0 + kSpan, g.h) - 1;            // last bin of this CTA (i0 <= i1 by grid construction)

    // ---- stage the sixteen harmonic slices with bulk async copies
    const uint32_t bar_addr = smem_u32(&bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_addr));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x < 32) {
        const int l = threadIdx.x + 1;
        uint32_t bytes = 0, first = 0;
        if (l <= 16) {
            first = ((uint32_t)l * i0 + 8u) >> 4;
            const uint32_t last = ((uint32_t)l * i1 + 8u) >> 4;
            first &= ~3u;                                    // 16-byte aligned source
            bytes = (((last - first + 1u) + 3u) & ~3u) * 4u; // whole 16-byte units (the row is padded past H)
        }
        uint32_t total = bytes;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
        if (threadIdx.x == 0)
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_addr), "r"(total) : "memory");
        __syncwarp();
        if (l <= 16) {
            const uint32_t dst = smem_u32(sm + slice_off(l));
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                         "l"(P + first), "r"(bytes), "r"(bar_addr)
                         : "memory");
        }
    }

    // ---- per-thread slice indices of its first bin (the only index arithmetic of the kernel)
    const uint32_t i = i0 + threadIdx.x;
    uint32_t idx[17];
#pragma unroll
    for (int l = 1; l <= 16; ++l) {
        const uint32_t first = (((uint32_t)l * i0 + 8u) >> 4) & ~3u;
        idx[l] = slice_off(l) + ((((uint32_t)l * i + 8u) >> 4) - first);
    }
    const uint32_t j1 = (8u * i + 8u) >> 4, j2 = (4u * i + 8u) >> 4, j3 = (2u * i + 8u) >> 4, j4 = (i + 8u) >> 4;

    // wait for the copies (phase 0 of the barrier)
    {
        uint32_t done = 0;
        while (!done) {
            asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n selp.u32 %0, 1, 0, p;\n}"
                         : "=r"(done)
                         : "r"(bar_addr)
                         : "memory");
        }
    }

#pragma unroll
    for (int u = 0; u < kHsPerThread; ++u) {
        const uint32_t iu = i + u * kHsThreads;
        if (iu > i1) break;
        // bins kHsThreads apart: slice index advances by exactly l*kHsThreads/16
#define PH(l) sm[idx[l] + u * ((l) * kHsThreads / 16)]
        const float s0 = PH(16);
        const float s1 = s0 + PH(8);
        const float s2 = s1 + (PH(12) + PH(4));
        const float s3 = s2 + (((PH(14) + PH(10)) + PH(6)) + PH(2));
        const float s4 = s3 + (((((((PH(15) + PH(13)) + PH(11)) + PH(9)) + PH(7)) + PH(5)) + PH(3)) + PH(1));
#undef PH
        if ((s0 > thr.t[0]) | (s1 > thr.t[1]) | (s2 > thr.t[2]) | (s3 > thr.t[3]) | (s4 > thr.t[4])) {
            // rare path: fundamental bins of this i (they advance by kHsThreads/2^k per step)
            const uint32_t b0 = iu, b1 = j1 + u * (kHsThreads / 2), b2 = j2 + u * (kHsThreads / 4), b3 = j3 + u * (kHsThreads / 8),
                           b4 = j4 + u * (kHsThreads / 16);
            if (s0 > thr.t[0] && b0 < fu) offer(keys + b0, s0, tkey);
            if (s1 > thr.t[1] && b1 < fu) offer(keys + (size_t)fu + b1, s1, tkey);
            if (s2 > thr.t[2] && b2 < fu) offer(keys + 2 * (size_t)fu + b2, s2, tkey);
            if (s3 > thr.t[3] && b3 < fu) offer(keys + 3 * (size_t)fu + b3, s3, tkey);
            if (s4 > thr.t[4] && b4 < fu) offer(keys + 4 * (size_t)fu + b4, s4, tkey);
        }
    }
}

__global__ void collect_kernel(const unsigned long long *__restrict__ keys, uint32_t fu, uint32_t w2,
                               brp_hit *__restrict__ hits, uint32_t max_hits, uint32_t *__restrict__ counter)
{
    const uint32_t total = 5u * fu;
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
        const unsigned long long key = keys[e];
        if (key == 0ull) continue;
        const uint32_t k = e / fu, bin = e - k * fu;
        if (bin < w2) continue; // the candidate scan starts at window_2 (demod_binary.c:1323)
        const uint32_t slot = atomicAdd(counter, 1u);
        if (slot < max_hits) {
            brp_hit h;
            h.power = __uint_as_float((uint32_t)(key >> 32));
            h.bin = bin;
            h.tmpl = 0xFFFFFFFFu - (uint32_t)(key & 0xFFFFFFFFull);
            h.harm_idx = k;
            hits[slot] = h;
        }
    }
}

} // namespace

cudaError_t brp_launch_harmonic_summing(const float *ps, size_t ps_stride, int batch, const BrpGeomDev &g,
                                        BrpThresholds thr, uint32_t first_template, unsigned long long *keys,
                                        cudaStream_t st)
{
    if (g.h <= g.w2) return cudaSuccess;
    static bool configured[64]; // per device: the attribute belongs to the function on the current device
    const int smem = kSmemFloats * (int)sizeof(float);
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || !configured[dev]) {
        cudaError_t e = cudaFuncSetAttribute((const void *)harmonic_summing_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    const dim3 grid((g.h - g.w2 + kSpan - 1) / kSpan, batch);
    harmonic_summing_kernel<<<grid, kHsThreads, smem, st>>>(ps, ps_stride, g, thr, first_template, keys);
    return cudaGetLastError();
}

cudaError_t brp_launch_collect(const unsigned long long *keys, uint32_t fu, uint32_t w2, brp_hit *hits,
                               uint32_t max_hits, uint32_t *counter, cudaStream_t st)
{
    collect_kernel<<<296, 256, 0, st>>>(keys, fu, w2, hits, max_hits, counter);
    return cudaGetLastError();
}
