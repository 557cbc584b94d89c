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
// Data movement.  Work is cut into tiles of kSpan = 1024 consecutive bins i that START AT MULTIPLES OF kSpan
// (the first tile of a template is partly below w2 and masked).  Harmonic l of a tile lives in ONE contiguous
// slice of the spectrum, P[(l*i0+8)>>4 .. (l*i1+8)>>4] (l*kSpan/16 + 1 floats), so the sixteen slices are
// brought into shared memory by sixteen 1-D bulk async copies (TMA, cp.async.bulk + mbarrier): no load
// instructions, no registers, no L1 wavefronts on the way in.  CTAs are persistent and walk the tiles of the
// whole batch with a two-deep ring; a producer warp issues the copies of tile t+1 while four consumer warps sum
// tile t (eight bins per thread), stages handed over through full/empty mbarriers without any block-wide barrier.
// Measured on B200 (tools/hs_sweep.sh): 2 stages x 1024 bins x 4 consumer warps, three CTAs per SM, is the fastest
// shape (77 us per 16 templates); deeper rings or wider CTAs lose more to the lower CTA count than they gain in
// prefetch distance.  What bounds the kernel is shared-memory bandwidth: 16 x 4 bytes read per bin by the consumers
// (80 B/clk/SM) plus the 8.5*H floats per template written by the bulk copies (43 B/clk/SM) against 128 B/clk/SM;
// halving the bytes copied (probe: odd harmonics not fetched) shortens it by only 7 %.
// Because i0 is a multiple of 1024, (l*i0+8)>>4 = l*i0/16 exactly, and the position of bin i0+tid+128*r inside
// slice l is ((l*tid+8)>>4) + 8*l*r: a per-thread constant plus a compile-time offset.  The hot loop therefore
// has no index arithmetic at all; consecutive lanes read consecutive or identical shared-memory words
// (conflict-free, broadcast).
// Algorithmic traffic 4*H bytes per template; the slices of different harmonics overlap in the spectrum,
// so L2 serves 8.5*H floats per template and HBM about H.
#include "brp_internal.h"

namespace {

#ifndef BRP_HS_THREADS
#define BRP_HS_THREADS 128
#endif
constexpr int kHsThreads = BRP_HS_THREADS;  // consumer threads (one more warp issues the copies)
#ifndef BRP_HS_PER_THREAD
#define BRP_HS_PER_THREAD 8
#endif
#ifndef BRP_HS_STAGES
#define BRP_HS_STAGES 2
#endif
constexpr int kHsPerThread = BRP_HS_PER_THREAD;       // bins per thread, kHsThreads apart
constexpr int kStages = BRP_HS_STAGES;                // depth of the shared-memory ring (kStages - 1 tiles in flight)
constexpr int kSpan = kHsThreads * kHsPerThread;      // bins per tile; tiles start at multiples of kSpan
// slice l of an aligned tile holds l*kSpan/16 + 1 floats, rounded up to whole 16-byte units
__host__ __device__ constexpr int slice_cap(int l) { return ((l * kSpan / 16 + 1 + 3) / 4) * 4; }
__host__ __device__ constexpr int slice_off(int l)
{
    int o = 0;
    for (int m = 1; m < l; ++m) o += slice_cap(m);
    return o;
}
constexpr int kSmemFloats = slice_off(17);

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// rare path, kept out of line so that the hot loop stays small: offer every sum of bin i that beats its threshold
__device__ __noinline__ void record_hits(uint32_t i, float s0, float s1, float s2, float s3, float s4, const BrpThresholds &thr,
                                         uint32_t fu, uint32_t tkey, unsigned long long *__restrict__ keys)
{
    const float s[5] = {s0, s1, s2, s3, s4};
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const uint32_t j = (i * (16u >> k) + 8u) >> 4; // fundamental bin of level k (hs_common.c:84,107,125,145)
        if (s[k] > thr.t[k] && j < fu)
            atomicMax(keys + (size_t)k * fu + j, ((unsigned long long)__float_as_uint(s[k]) << 32) | tkey);
    }
}

// stage the sixteen harmonic slices of tile (bx, template b) into `dst` with bulk async copies (called by warp 0)
__device__ __forceinline__ void issue_tile(const float *__restrict__ ps, size_t ps_stride, const BrpGeomDev &g, uint32_t tile0,
                                           uint32_t tile, uint32_t tiles_per_template, float *dst, uint32_t bar_addr)
{
    const uint32_t b = tile / tiles_per_template, bx = tile - b * tiles_per_template;
    const float *__restrict__ P = ps + (size_t)b * ps_stride;
    const uint32_t i0 = (tile0 + bx) * kSpan, i1 = min(i0 + kSpan, g.h) - 1; // i0 is a multiple of kSpan
    const int l = (threadIdx.x & 31) + 1;
    uint32_t bytes = 0, first = 0;
    if (l <= 16) {
        first = (((uint32_t)l * i0 + 8u) >> 4) & ~3u;                    // 16-byte aligned source
        const uint32_t last = ((uint32_t)l * i1 + 8u) >> 4;
        bytes = (((last - first + 1u) + 3u) & ~3u) * 4u;                 // whole 16-byte units (rows are padded past H)
    }
    uint32_t total = bytes;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
    if ((threadIdx.x & 31) == 0)
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_addr), "r"(total) : "memory");
    __syncwarp();
    if (l <= 16)
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         smem_u32(dst + slice_off(l))),
                     "l"(P + first), "r"(bytes), "r"(bar_addr)
                     : "memory");
}

template <int IMM>
__device__ __forceinline__ float lds_at(uint32_t a)
{
    float v;
    asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v) : "r"(a), "n"(IMM));
    return v;
}

// the five sums of bin iu = i + U*256 of the tile in ring stage ST, and the rare path if any beats its threshold
template <int ST, int U>
__device__ __forceinline__ void sum_bin(const uint32_t (&addr)[17], uint32_t i, const BrpGeomDev &g, const BrpThresholds &thr,
                                        uint32_t tkey, unsigned long long *__restrict__ keys)
{
    const uint32_t iu = i + U * kHsThreads;
#define PH(l) lds_at<4 * (ST * kSmemFloats + U * ((l) * kHsThreads / 16))>(addr[l])
    const float s0 = PH(16);
    const float s1 = s0 + PH(8);
    const float s2 = s1 + (PH(12) + PH(4));
    const float s3 = s2 + (((PH(14) + PH(10)) + PH(6)) + PH(2));
    const float s4 = s3 + (((((((PH(15) + PH(13)) + PH(11)) + PH(9)) + PH(7)) + PH(5)) + PH(3)) + PH(1));
#undef PH
    const bool any = (s0 > thr.t[0]) | (s1 > thr.t[1]) | (s2 > thr.t[2]) | (s3 > thr.t[3]) | (s4 > thr.t[4]);
    // bins outside [w2, H) only exist in the first and last tile of a template (their slots may hold stale data)
    if (any && iu >= g.w2 && iu < g.h) record_hits(iu, s0, s1, s2, s3, s4, thr, g.fu, tkey, keys);
}

__device__ __forceinline__ void mbar_wait(uint32_t addr, uint32_t parity)
{
    uint32_t done = 0;
    while (!done) {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done)
                     : "r"(addr), "r"(parity)
                     : "memory");
    }
}

// Two bins per instruction: the sums of bins iu = i + U*128 and i + (U+1)*128 ride in the two halves of packed FP32x2
// registers (FADD2).  Each half is an ordinary IEEE addition in the reference's association, so the sums are bit-identical
// to sum_bin's; the kernel is bound by issue slots (63 % busy, FMA pipe 20 %), and this takes 7.5 of the 15 additions per bin
// out of the instruction stream.
#ifndef BRP_HS_PACKED
#define BRP_HS_PACKED 1
#endif
template <int ST, int U>
__device__ __forceinline__ void sum_bin_pair(const uint32_t (&addr)[17], uint32_t i, const BrpGeomDev &g, const BrpThresholds &thr,
                                             uint32_t tkey, unsigned long long *__restrict__ keys)
{
#define PH(l) lds_at<4 * (ST * kSmemFloats + U * ((l) * kHsThreads / 16))>(addr[l])
#define PH1(l) lds_at<4 * (ST * kSmemFloats + (U + 1) * ((l) * kHsThreads / 16))>(addr[l])
#define P2(l) make_float2(PH(l), PH1(l))
    const float2 s0 = P2(16);
    const float2 s1 = __fadd2_rn(s0, P2(8));
    const float2 s2 = __fadd2_rn(s1, __fadd2_rn(P2(12), P2(4)));
    const float2 s3 = __fadd2_rn(s2, __fadd2_rn(__fadd2_rn(__fadd2_rn(P2(14), P2(10)), P2(6)), P2(2)));
    const float2 s4 = __fadd2_rn(
        s3, __fadd2_rn(__fadd2_rn(__fadd2_rn(__fadd2_rn(__fadd2_rn(__fadd2_rn(__fadd2_rn(P2(15), P2(13)), P2(11)), P2(9)), P2(7)), P2(5)), P2(3)), P2(1)));
#undef P2
#undef PH1
#undef PH
    const uint32_t iu = i + U * kHsThreads, iv = iu + kHsThreads;
    const bool any0 = (s0.x > thr.t[0]) | (s1.x > thr.t[1]) | (s2.x > thr.t[2]) | (s3.x > thr.t[3]) | (s4.x > thr.t[4]);
    const bool any1 = (s0.y > thr.t[0]) | (s1.y > thr.t[1]) | (s2.y > thr.t[2]) | (s3.y > thr.t[3]) | (s4.y > thr.t[4]);
    if (any0 && iu >= g.w2 && iu < g.h) record_hits(iu, s0.x, s1.x, s2.x, s3.x, s4.x, thr, g.fu, tkey, keys);
    if (any1 && iv >= g.w2 && iv < g.h) record_hits(iv, s0.y, s1.y, s2.y, s3.y, s4.y, thr, g.fu, tkey, keys);
}

template <int ST, int U>
__device__ __forceinline__ void sum_bins(const uint32_t (&addr)[17], uint32_t i, const BrpGeomDev &g, const BrpThresholds &thr,
                                         uint32_t tkey, unsigned long long *__restrict__ keys)
{
    if constexpr (BRP_HS_PACKED != 0 && U + 1 < kHsPerThread) {
        sum_bin_pair<ST, U>(addr, i, g, thr, tkey, keys);
        sum_bins<ST, U + 2>(addr, i, g, thr, tkey, keys);
    } else if constexpr (U < kHsPerThread) {
        sum_bin<ST, U>(addr, i, g, thr, tkey, keys);
        sum_bins<ST, U + 1>(addr, i, g, thr, tkey, keys);
    }
}

// consumer side of one tile: wait until ring stage ST is full, sum the thread's bins, hand the stage back
template <int ST>
__device__ __forceinline__ void sum_tile(const BrpGeomDev &g, const BrpThresholds &thr, uint32_t first_template, uint32_t tile0,
                                         uint32_t tiles_per_template, unsigned long long *__restrict__ keys, uint32_t full0,
                                         uint32_t empty0, const uint32_t (&addr)[17], uint32_t tile, uint32_t n)
{
    const uint32_t b = tile / tiles_per_template, bx = tile - b * tiles_per_template;
    const uint32_t tkey = 0xFFFFFFFFu - (first_template + b);
    const uint32_t i = (tile0 + bx) * kSpan + threadIdx.x;
    mbar_wait(full0 + 8 * ST, (n / kStages) & 1u); // use number n / kStages of this stage
    sum_bins<ST, 0>(addr, i, g, thr, tkey, keys);
    __syncwarp();
    if ((threadIdx.x & 31) == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(empty0 + 8 * ST) : "memory");
}

// Persistent CTAs, warp-specialised: the last warp only issues the bulk copies (tile n into ring stage n % kStages, as
// soon as the consumer warps have handed that stage back), the other warps only sum.  Stages are passed through full/empty
// mbarriers, so there is no block-wide barrier in the loop and a consumer warp never waits for a slower sibling.
__global__ void __launch_bounds__(kHsThreads + 32)
harmonic_summing_kernel(const float *__restrict__ ps, size_t ps_stride, BrpGeomDev g, BrpThresholds thr,
                        uint32_t first_template, uint32_t tile0, uint32_t tiles_per_template, uint32_t total_tiles,
                        unsigned long long *__restrict__ keys)
{
    extern __shared__ __align__(128) float sm[];
    __shared__ __align__(8) unsigned long long bar[2 * kStages];

    const uint32_t full0 = smem_u32(&bar[0]), empty0 = smem_u32(&bar[kStages]);
    if (threadIdx.x == 0) {
        for (int q = 0; q < kStages; ++q) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(full0 + 8 * q));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(empty0 + 8 * q), "r"(kHsThreads / 32));
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t tile = blockIdx.x;

    if (threadIdx.x >= kHsThreads) { // producer warp
        for (uint32_t n = 0; tile < total_tiles; ++n, tile += gridDim.x) {
            const uint32_t st = n % kStages;
            if (n >= kStages) mbar_wait(empty0 + 8 * st, (n / kStages - 1u) & 1u);
            issue_tile(ps, ps_stride, g, tile0, tile, tiles_per_template, sm + st * kSmemFloats, full0 + 8 * st);
        }
        return;
    }

    // Tiles are aligned to kSpan, so bin i0 + t + 128 u of harmonic l sits at slice offset ((l t + 8) >> 4) + 8 l u:
    // the first term is a per-thread constant for the whole kernel (kept as a shared-memory byte address), the second
    // and the ring stage are immediates of the load instruction (the tile loop is unrolled over the ring stages).
    uint32_t addr[17];
#pragma unroll
    for (int l = 1; l <= 16; ++l) addr[l] = smem_u32(sm) + 4u * (slice_off(l) + (((uint32_t)l * threadIdx.x + 8u) >> 4));

#define BRP_HS_STEP(ST)                                                                                                         \
    if constexpr (ST < kStages) {                                                                                               \
        if (tile >= total_tiles) break;                                                                                         \
        sum_tile<ST>(g, thr, first_template, tile0, tiles_per_template, keys, full0, empty0, addr, tile, n + ST);               \
        tile += gridDim.x;                                                                                                      \
    }
    for (uint32_t n = 0;; n += kStages) {
        BRP_HS_STEP(0)
        BRP_HS_STEP(1)
        BRP_HS_STEP(2)
        BRP_HS_STEP(3)
        BRP_HS_STEP(4)
        BRP_HS_STEP(5)
    }
#undef BRP_HS_STEP
}

// TO_HOST (one-template-per-call callers): the table is left empty behind the scan (saves a memset of the whole table per call),
// `hits` is mapped host memory, and the block that finishes last publishes the hit count there and resets the control words
template <bool TO_HOST>
__global__ void collect_kernel(unsigned long long *__restrict__ keys, uint32_t fu, uint32_t w2,
                               brp_hit *__restrict__ hits, uint32_t max_hits, uint32_t *__restrict__ counter,
                               uint32_t *__restrict__ next_counter, uint32_t *__restrict__ arrived, uint32_t *__restrict__ host_count)
{
    const uint32_t total = 5u * fu;
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
        const unsigned long long key = keys[e];
        if (key == 0ull) continue;
        if (TO_HOST) keys[e] = 0ull;
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
    if constexpr (TO_HOST) {
        __syncthreads(); // this block's slot requests are all issued
        if (threadIdx.x == 0) {
            __threadfence();
            if (atomicAdd(arrived, 1u) == gridDim.x - 1) {
                __threadfence();
                const uint32_t n = atomicAdd(counter, 0u);
                *host_count = n < max_hits ? n : max_hits;
                *next_counter = 0u;
                *arrived = 0u;
            }
        }
    }
}

} // namespace

cudaError_t brp_launch_harmonic_summing(const float *ps, size_t ps_stride, int batch, const BrpGeomDev &g,
                                        BrpThresholds thr, uint32_t first_template, unsigned long long *keys,
                                        cudaStream_t st)
{
    if (g.h <= g.w2 || batch <= 0) return cudaSuccess;
    static int ctas_per_device[64]; // persistent grid size, per device (0 = not configured yet)
    const int smem = kStages * kSmemFloats * (int)sizeof(float);
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (!ctas_per_device[dev]) {
        cudaError_t e = cudaFuncSetAttribute((const void *)harmonic_summing_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        cudaFuncSetAttribute((const void *)harmonic_summing_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        int sms = 0, per_sm = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, harmonic_summing_kernel, kHsThreads + 32, smem);
        if (e != cudaSuccess) return e;
        ctas_per_device[dev] = (sms > 0 ? sms : 148) * (per_sm > 0 ? per_sm : 1); // one resident wave
    }
    // aligned tiles covering [w2, H): tile index range [w2 / kSpan, ceil(H / kSpan))
    const uint32_t tile0 = g.w2 / kSpan;
    const uint32_t tiles_per_template = (g.h + kSpan - 1) / kSpan - tile0;
    const uint32_t total = tiles_per_template * (uint32_t)batch;
    const uint32_t grid = total < (uint32_t)ctas_per_device[dev] ? total : (uint32_t)ctas_per_device[dev];
    harmonic_summing_kernel<<<grid, kHsThreads + 32, smem, st>>>(ps, ps_stride, g, thr, first_template, tile0, tiles_per_template, total, keys);
    return cudaGetLastError();
}

cudaError_t brp_launch_collect(const unsigned long long *keys, uint32_t fu, uint32_t w2, brp_hit *hits,
                               uint32_t max_hits, uint32_t *counter, cudaStream_t st)
{
    collect_kernel<false><<<296, 256, 0, st>>>(const_cast<unsigned long long *>(keys), fu, w2, hits, max_hits, counter, nullptr, nullptr, nullptr);
    return cudaGetLastError();
}

cudaError_t brp_launch_collect_to_host(unsigned long long *keys, uint32_t fu, uint32_t w2, brp_hit *hits, uint32_t max_hits,
                                       uint32_t *ctl, int which, uint32_t *host_count, cudaStream_t st)
{
    collect_kernel<true><<<296, 256, 0, st>>>(keys, fu, w2, hits, max_hits, ctl + (which & 1), ctl + ((which & 1) ^ 1), ctl + 2, host_count);
    return cudaGetLastError();
}
