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
// Above-threshold values are rare (false-alarm rate ~1e-7 per bin), so the kernel is a pure streaming
// read of P[0..H): 4*H algorithmic bytes per template, nothing is written in the common case.
// A warp-level vote skips the atomic path entirely when no lane crossed a threshold.
#include "brp_internal.h"

namespace {

constexpr int kHsThreads = 256;
constexpr int kHsPerThread = 4;

__device__ __forceinline__ void offer(unsigned long long *slot, float power, uint32_t tkey)
{
    atomicMax(slot, ((unsigned long long)__float_as_uint(power) << 32) | tkey);
}

__global__ void __launch_bounds__(kHsThreads)
harmonic_summing_kernel(const float *__restrict__ ps, size_t ps_stride, BrpGeomDev g, BrpThresholds thr,
                        uint32_t first_template, unsigned long long *__restrict__ keys)
{
    const float *__restrict__ P = ps + (size_t)blockIdx.y * ps_stride;
    const uint32_t tkey = 0xFFFFFFFFu - (first_template + blockIdx.y);
    const uint32_t fu = g.fu;
    const uint32_t base = g.w2 + blockIdx.x * (kHsThreads * kHsPerThread) + threadIdx.x;
#pragma unroll
    for (int u = 0; u < kHsPerThread; ++u) {
        const uint32_t i = base + u * kHsThreads;
        const bool live = i < g.h;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f, s4 = 0.f;
        uint32_t j1 = 0, j2 = 0, j3 = 0, j4 = 0;
        if (live) {
#define PH(l) __ldg(P + (((uint32_t)(l) * i + 8u) >> 4))
            j1 = (8u * i + 8u) >> 4;
            j2 = (4u * i + 8u) >> 4;
            j3 = (2u * i + 8u) >> 4;
            j4 = (i + 8u) >> 4;
            s0 = __ldg(P + i);
            s1 = s0 + __ldg(P + j1);
            s2 = s1 + (PH(12) + __ldg(P + j2));
            s3 = s2 + (((PH(14) + PH(10)) + PH(6)) + __ldg(P + j3));
            s4 = s3 + (((((((PH(15) + PH(13)) + PH(11)) + PH(9)) + PH(7)) + PH(5)) + PH(3)) + __ldg(P + j4));
#undef PH
        }
        const bool h0 = live && s0 > thr.t[0] && i < fu;
        const bool h1 = live && s1 > thr.t[1] && j1 < fu;
        const bool h2 = live && s2 > thr.t[2] && j2 < fu;
        const bool h3 = live && s3 > thr.t[3] && j3 < fu;
        const bool h4 = live && s4 > thr.t[4] && j4 < fu;
        if (__any_sync(0xffffffffu, h0 | h1 | h2 | h3 | h4)) {
            if (h0) offer(keys + i, s0, tkey);
            if (h1) offer(keys + (size_t)fu + j1, s1, tkey);
            if (h2) offer(keys + 2 * (size_t)fu + j2, s2, tkey);
            if (h3) offer(keys + 3 * (size_t)fu + j3, s3, tkey);
            if (h4) offer(keys + 4 * (size_t)fu + j4, s4, tkey);
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
    const uint32_t per_block = kHsThreads * kHsPerThread;
    const dim3 grid((g.h - g.w2 + per_block - 1) / per_block, batch);
    harmonic_summing_kernel<<<grid, kHsThreads, 0, st>>>(ps, ps_stride, g, thr, first_template, keys);
    return cudaGetLastError();
}

cudaError_t brp_launch_collect(const unsigned long long *keys, uint32_t fu, uint32_t w2, brp_hit *hits,
                               uint32_t max_hits, uint32_t *counter, cudaStream_t st)
{
    collect_kernel<<<296, 256, 0, st>>>(keys, fu, w2, hits, max_hits, counter);
    return cudaGetLastError();
}
