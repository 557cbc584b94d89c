// Kernel 2: two-pass shared-memory-staged R2C FFT with fused power-spectrum epilogue (north_star kernel 2).
//
// Contract (what the reference gets from fftwf_plan_dft_r2c_1d + the loop in demod_binary_fft_fftw.c:88-113):
//     P[k] = norm * |sum_{n<N} x[n] exp(-2 pi i k n / N)|^2   for 1 <= k < count,   P[0] = 0
// where x is the resampled series mean-padded to N samples (demod_binary_resamp_cpu.c:130-133).
// No cuFFT: the transform is written here.
//
// The N reals are viewed as M = N/2 complex points z[m] = x[2m] + i x[2m+1] and M = M1*M2 is split
// four-step style, so the data makes exactly two trips through HBM:
//
//   pass 1 (fft_pass1_kernel): one tile = ROWS consecutive columns m1 of the M2 x M1 view; for each column
//       a length-M2 transform over m2 (input stride M1), the padding rule applied on load
//       (i >= n_steps -> mean, so the pad is never materialised), times the twiddle W_M^{m1*k2}, written
//       transposed as Y[k2][m1].
//   pass 2 (fft_pass2_kernel): one tile = T2 rows k2 and their T2 mirror rows M2-k2 of Y; a length-M1
//       transform along each (contiguous) row gives Z[k2 + M2*k1]; row k2 and row M2-k2 hold exactly
//       the pairs (Z[k], Z[M-k]) the real-input "untangle" needs, so the epilogue forms
//       X[k] = E + W_N^k O, X[M-k] = conj(E - W_N^k O) and stores norm*|X|^2 for bins < count only.
//
// In a tile every thread keeps a radix-16 (or 32/8/4/2/3) butterfly in registers; stages exchange data through shared
// memory (decimation in frequency, in place, digit-reversed result).  Shared-memory positions are padded and lanes
// permuted per tile shape (Layout<>, chosen with tools/bank_conflicts.py and tools/pass1_tma_layout.py) so the
// power-of-two strides of the butterflies and the digit-reversed reads of the epilogue stay free of bank conflicts.
//
// Global traffic.  A pass-1 tile is a few columns wide (32-byte runs 8 KB apart) and pass 2 writes 16-byte runs: moved
// by the lanes, every run costs an L1 tag lookup, and those lookups -- not HBM, not issue slots -- bounded both kernels.
// For the production shapes the tiles therefore travel by TMA tensor copies (fft_pass1_tma_kernel, fft_pass2_tma_kernel:
// dense tile images in shared memory, cp.async.bulk.tensor boxes of 256 rows); every other shape uses the plain kernels
// (fft_pass1_kernel, fft_pass2_kernel), whose first stage reads straight from global memory and whose last stage writes
// straight back.  Both families do the same arithmetic (tests/test_gpu_fft.py compares them bit for bit).
//
// HBM traffic per template: 4N (x) + 8M (Y out) + 8M (Y in) + 4*count (P) = 12N + 4*count bytes,
// against the algorithmic 4N + 4F of SURVEY.md 8(d).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include <cuda.h> // CUtensorMap (types only; the encoder is looked up at run time, libcuda is not linked)

#include "brp_internal.h"

namespace {

#ifndef BRP_BIG_TILE_L
#define BRP_BIG_TILE_L 0 // tile length that gets a 16384-element tile (8 rows, 512 threads, one CTA per SM); 0 = none.
// Measured with 2048 on B200: 64-byte global runs but only one CTA per SM -> pass 1 takes 178 us instead of 164 us.
#endif

// ------------------------------------------------------------------ compile-time plan of one tile transform
__host__ __device__ constexpr int cfg_radix(int L, int s)
{
    // two radix-32 stages for the production pass-2 length (one shared-memory exchange instead of two)
    if (L == 1024) return s < 2 ? 32 : 0;
    if (L == 512) return s == 0 ? 32 : (s == 1 ? 16 : 0);
    int rem = L;
    if (rem % 3 == 0) {
        if (s == 0) return 3;
        rem /= 3;
        --s;
    }
    while (rem > 1) {
        const int r = rem >= 16 ? 16 : rem;
        if (s == 0) return r;
        rem /= r;
        --s;
    }
    return 0;
}
__host__ __device__ constexpr int cfg_stages(int L)
{
    int s = 0;
    while (cfg_radix(L, s) != 0) ++s;
    return s;
}
// sub-transform length entering stage s
__host__ __device__ constexpr int cfg_n(int L, int s)
{
    int n = L;
    for (int i = 0; i < s; ++i) n /= cfg_radix(L, i);
    return n;
}
__host__ __device__ constexpr int cfg_stride(int L, int s) { return cfg_n(L, s) / cfg_radix(L, s); }
// offset of stage s in the tile twiddle table: stage s holds (R-1)*S entries, index (q-1)*S + p
__host__ __device__ constexpr int cfg_tw_offset(int L, int s)
{
    int off = 0;
    for (int i = 0; i < s; ++i) off += (cfg_radix(L, i) - 1) * cfg_stride(L, i);
    return off;
}
__host__ __device__ constexpr int cfg_tw_size(int L) { return cfg_tw_offset(L, cfg_stages(L)); }

// Shared-memory layout of a tile: element `pos` of row `row` lives at row*PITCH + pos + pos/16 + C8*(pos/256)
// (float2 units).  The padding constants, the pitch residue and the lane permutation of the row-fast stages were
// chosen per tile shape with tools/bank_conflicts.py, which replays every shared-memory access of the transform:
// with these values all stages of the production shapes are conflict-free (pass 1: 2048x4, pass 2: 1024x8 incl.
// the digit-reversed reads of the untangle epilogue).
#ifndef BRP_FFT_P1_LAYOUT
#define BRP_FFT_P1_LAYOUT 2 // 1: the round-1 layout of the tensor-copy pass 1 (2-way conflicts in the dense stage-0 read), for A/B runs
#endif
template <int L, int ROWS, int PASS>
struct Layout {
    // the radix-32 plans (L = 1024: 32x32, L = 512: 32x16) pad every 32 (resp. 16) positions
    static constexpr int C4 = (L == 1024 || (L == 512 && PASS == 1)) ? 0 : 1;
    static constexpr int C5 = (L == 1024) ? 1 : 0;
    static constexpr int C8 = ((PASS == 2 && (L == 2048 || L == 4096)) || PASS == 3) ? 1 : 0; // PASS 3: pass 1 through TMA, see below
    static constexpr int EXTRA = PASS == 3 ? (BRP_FFT_P1_LAYOUT == 1 ? 2 : 4) : PASS == 2 ? (L == 2048 ? (ROWS == 8 ? 2 : 4) : L == 1024 ? 4 : L == 512 ? 2 : (L == 4096 || L == 16) ? 0 : 1)
                                           : (L == 1024 ? 2 : ROWS == 4 ? (L % 3 == 0 ? 1 : 2) : ROWS == 2 ? (L == 4096 ? 8 : 4) : 1);
    // permute the low four bits of the butterfly index in the first (global -> shared) stage of pass 1
    // (the tensor-copy pass 1, PASS 3, needs none: with its pitch of 12 mod 16 plain row-fast lanes are conflict-free both in the dense
    //  read and in the padded write of stage 0; the matching lane map of its last stage is in the kernel)
    static constexpr bool ROT0 = (PASS == 1 || (PASS == 3 && BRP_FFT_P1_LAYOUT == 1)) && L != 1024 && (ROWS == 4 || ROWS == 8 || (ROWS == 2 && L != 4096));
    __host__ __device__ static constexpr int phys(int pos) { return pos + C4 * (pos >> 4) + C5 * (pos >> 5) + C8 * (pos >> 8); }
    static constexpr int PITCH = phys(L) + EXTRA;
    static constexpr int BYTES = ROWS * PITCH * 8;
};
__host__ __device__ constexpr int cfg_rows(int L)
{
    const int tile = (L % 3 == 0) ? 12288 : (L == BRP_BIG_TILE_L ? 16384 : 8192);
    return tile / L < 2 ? 2 : tile / L;
}

#ifndef BRP_FFT_WIDE_L
#define BRP_FFT_WIDE_L 2048 // tile length that runs with 512 threads (one butterfly per thread and stage), two CTAs per SM
#endif
__host__ __device__ constexpr int cfg_threads(int L) { return (cfg_rows(L) * L >= 16384 || L == BRP_FFT_WIDE_L) ? 512 : 256; }
__host__ __device__ constexpr int cfg_ctas(int L) { return cfg_rows(L) * L >= 16384 ? 1 : 2; }

// ------------------------------------------------------------------ complex helpers
// Blackwell packed FP32x2 arithmetic (FADD2 / FFMA2): one instruction per complex add or subtract
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return __ffma2_rn(b, make_float2(-1.0f, -1.0f), a); }
__device__ __forceinline__ float2 cmul(float2 a, float2 b)
{
    return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}

// multiply by exp(-2 pi i e / 16), e = 0..7
template <int E>
__device__ __forceinline__ float2 mul_w16(float2 d)
{
    constexpr float C1 = 0.92387953251128673848f, S1 = 0.38268343236508978178f, H = 0.70710678118654752440f;
    if constexpr (E == 0) return d;
    else if constexpr (E == 1) return make_float2(fmaf(d.x, C1, d.y * S1), fmaf(d.y, C1, -d.x * S1));
    else if constexpr (E == 2) return make_float2((d.x + d.y) * H, (d.y - d.x) * H);
    else if constexpr (E == 3) return make_float2(fmaf(d.x, S1, d.y * C1), fmaf(d.y, S1, -d.x * C1));
    else if constexpr (E == 4) return make_float2(d.y, -d.x);
    else if constexpr (E == 5) return make_float2(fmaf(d.y, C1, -d.x * S1), -fmaf(d.x, C1, d.y * S1));
    else if constexpr (E == 6) return make_float2((d.y - d.x) * H, -(d.x + d.y) * H);
    else return make_float2(fmaf(d.y, S1, -d.x * C1), -fmaf(d.x, S1, d.y * C1));
}

// multiply by exp(-2 pi i e / 32), e = 0..15
template <int E>
__device__ __forceinline__ float2 mul_w32(float2 d)
{
    if constexpr (E % 2 == 0) return mul_w16<E / 2>(d);
    else {
        // cos/sin of (2 pi E / 32) for odd E, from the two base angles pi/16 and 3 pi/16
        constexpr float CA = 0.98078528040323044913f, SA = 0.19509032201612826785f;
        constexpr float CB = 0.83146961230254523708f, SB = 0.55557023301960222474f;
        constexpr float C = E == 1 ? CA : E == 3 ? CB : E == 5 ? SB : E == 7 ? SA : E == 9 ? -SA : E == 11 ? -SB : E == 13 ? -CB : -CA;
        constexpr float S = E == 1 ? SA : E == 3 ? SB : E == 5 ? CB : E == 7 ? CA : E == 9 ? CA : E == 11 ? CB : E == 13 ? SB : SA;
        return make_float2(fmaf(d.x, C, d.y * S), fmaf(d.y, C, -d.x * S)); // d * (C - i S)
    }
}

// in-register DFT of R = 2^a points (R <= 32), decimation in frequency; output q ends up in v[bitrev(q)]
template <int R>
struct Dft {
    template <int J>
    static __device__ __forceinline__ void level(float2 *v)
    {
        if constexpr (J < R / 2) {
            const float2 a = v[J], b = v[J + R / 2];
            v[J] = cadd(a, b);
            v[J + R / 2] = mul_w32<J *(32 / R)>(csub(a, b));
            level<J + 1>(v);
        }
    }
    static __device__ __forceinline__ void run(float2 *v)
    {
        level<0>(v);
        Dft<R / 2>::run(v);
        Dft<R / 2>::run(v + R / 2);
    }
};
template <>
struct Dft<1> {
    static __device__ __forceinline__ void run(float2 *) {}
};
// 3-point DFT, natural output order
template <>
struct Dft<3> {
    static __device__ __forceinline__ void run(float2 *v)
    {
        constexpr float S3 = 0.86602540378443864676f;
        const float2 s = cadd(v[1], v[2]), d = csub(v[1], v[2]);
        const float2 m = make_float2(fmaf(-0.5f, s.x, v[0].x), fmaf(-0.5f, s.y, v[0].y));
        const float2 r = make_float2(S3 * d.y, -S3 * d.x); // -i * S3 * d
        v[0] = cadd(v[0], s);
        v[1] = cadd(m, r);
        v[2] = csub(m, r);
    }
};

template <int R>
__host__ __device__ constexpr int out_slot(int q)
{
    if (R == 3) return q;
    int r = 0;
    for (int b = 1, t = R >> 1; b < R; b <<= 1, t >>= 1)
        if (q & b) r |= t;
    return r;
}

// frequency index of the value that ends at in-place position pos: pos = sum q_s S_s  ->  k = sum q_s (L / n_s)
template <int L>
__device__ __forceinline__ int freq_of_pos(int pos)
{
    int k = 0;
#pragma unroll
    for (int s = 0; s < cfg_stages(L); ++s) {
        const int q = (pos / cfg_stride(L, s)) % cfg_radix(L, s);
        k += q * (L / cfg_n(L, s));
    }
    return k;
}
template <int L>
__device__ __forceinline__ int pos_of_freq(int k)
{
    int pos = 0;
#pragma unroll
    for (int s = 0; s < cfg_stages(L); ++s) {
        const int q = (k / (L / cfg_n(L, s))) % cfg_radix(L, s);
        pos += q * cfg_stride(L, s);
    }
    return pos;
}

enum { MAP_ROW_FAST = 0, MAP_BF_FAST = 1 };

// One DIF stage over the whole tile.  The I/O objects are used as
//     auto c = io.begin(row, base);   io.get(c, off);   io.put(c, off, q, value);
// where base = blk*n + p is the butterfly's first position and off = j*S is a compile-time constant after
// unrolling, so all per-element address arithmetic folds into immediates.
// Who synchronises: the whole CTA (the default), or one group of cfg_threads(L) threads of a wider CTA that works on its own tile
// (named barrier `id`; the persistent pass-1 kernel runs two such groups per CTA).
struct CtaSync {
    __device__ __forceinline__ int tid() const { return (int)threadIdx.x; }
    __device__ __forceinline__ void sync() const { __syncthreads(); }
};
template <int NT>
struct GroupSync {
    int id, t;
    __device__ __forceinline__ int tid() const { return t; }
    __device__ __forceinline__ void sync() const { asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(NT) : "memory"); }
};

// TW_SMEM: `tw` points into shared memory (a copy of the table made by the kernel), read with plain loads
template <int L, int ROWS, int STAGE, int MAP, bool ROT, bool SYNC_AFTER_LOAD = false, bool TREE = false, bool TW_SMEM = false, typename Load,
          typename Store, typename Sync = CtaSync>
__device__ __forceinline__ void tile_stage(const float2 *__restrict__ tw, const Load &load, const Store &store, const Sync &sy = Sync())
{
    constexpr int R = cfg_radix(L, STAGE);
    constexpr int n = cfg_n(L, STAGE);
    constexpr int S = cfg_stride(L, STAGE);
    constexpr int BF = L / R; // butterflies per row
    constexpr int ITEMS = ROWS * BF;
    constexpr bool LAST = (STAGE == cfg_stages(L) - 1);
    constexpr int TWO = cfg_tw_offset(L, STAGE);
    constexpr int NT = cfg_threads(L);
    static_assert(ITEMS % NT == 0, "tile must be a whole number of butterflies per thread");
#pragma unroll
    for (int it = 0; it < ITEMS / NT; ++it) {
        const int item = sy.tid() + it * NT;
        int row, bf;
        if (MAP == MAP_ROW_FAST) {
            row = item % ROWS;
            bf = item / ROWS;
            // same set of butterflies, different lane order: spreads a half-warp's shared-memory stores over all banks
            if (ROT && BF >= 16) bf = (bf & ~15) | ((bf & 1) << 3) | ((bf >> 1) & 7);
        } else {
            row = item / BF;
            bf = item % BF;
        }
        const int blk = bf / S, p = bf % S;
        const int base = blk * n + p;
        float2 v[R];
        const auto lc = load.begin(row, base);
#pragma unroll
        for (int j = 0; j < R; ++j) v[j] = load.get(lc, j * S);
        if constexpr (SYNC_AFTER_LOAD) { // the store region overlaps the load region: everybody reads before anybody writes
            static_assert(ITEMS == NT, "one butterfly per thread");
            sy.sync();
        }
        Dft<R>::run(v);
        const auto sc = store.begin(row, base);
        const float2 *__restrict__ twp = tw + TWO + p;
        if constexpr (!LAST && ((MAP == MAP_ROW_FAST && ROWS > 1) || TREE)) {
            // Row-fast lanes share p between the ROWS rows, so table loads would fetch every twiddle ROWS times
            // (measured: 0.17 L1 wavefronts per element in pass 1).  Load W_n^p once and square/multiply up to
            // W_n^{p q}: the product tree is at most log2(R) deep, i.e. a few ulp.
            float2 w[R];
            w[1] = TW_SMEM ? *twp : __ldg(twp);
#pragma unroll
            for (int q = 2; q < R; ++q) {
                int hb = 1;
                while (hb * 2 <= q) hb *= 2;
                w[q] = (q == hb) ? cmul(w[hb / 2], w[hb / 2]) : cmul(w[hb], w[q - hb]);
            }
#pragma unroll
            for (int q = 0; q < R; ++q) {
                float2 o = v[out_slot<R>(q)];
                if (q > 0) o = cmul(o, w[q]);
                store.put(sc, q * S, q, o);
            }
        } else {
#pragma unroll
            for (int q = 0; q < R; ++q) {
                float2 o = v[out_slot<R>(q)];
                if (!LAST && q > 0) o = cmul(o, TW_SMEM ? twp[(q - 1) * S] : __ldg(twp + (q - 1) * S));
                store.put(sc, q * S, q, o);
            }
        }
    }
}

// Shared-memory tile.  phys(base + off) == phys(base) + phys(off) for every (base, off) a stage produces:
// off = j*S with S a power of two, base = blk*R*S + p with p < S, so neither the /16 nor the /256 padding
// term carries from the p part into the j part (S | 16 or 16 | S, S | 256 or 256 | S).
template <int L, int ROWS, int PASS>
struct SmemIO {
    using Lay = Layout<L, ROWS, PASS>;
    float2 *sm;
    __device__ __forceinline__ float2 *begin(int row, int base) const { return sm + row * Lay::PITCH + Lay::phys(base); }
    __device__ __forceinline__ float2 get(float2 *c, int off) const { return c[Lay::phys(off)]; }
    __device__ __forceinline__ void put(float2 *c, int off, int, float2 v) const { c[Lay::phys(off)] = v; }
};

// middle stages (shared -> shared), recursively unrolled over the stage index
template <int L, int ROWS, int PASS, int STAGE, int END, typename Sync = CtaSync, bool TW_SMEM = false>
__device__ __forceinline__ void tile_middle(const float2 *tw, float2 *sm, const Sync &sy = Sync())
{
    if constexpr (STAGE < END) {
        SmemIO<L, ROWS, PASS> io{sm};
        tile_stage<L, ROWS, STAGE, MAP_BF_FAST, false, false, false, TW_SMEM>(tw, io, io, sy); // (a power tree here is slower: measured 257 vs 250 us)
        sy.sync();
        tile_middle<L, ROWS, PASS, STAGE + 1, END, Sync, TW_SMEM>(tw, sm, sy);
    }
}

// ------------------------------------------------------------------ pass 1
// first stage: column `col` of the M2 x M1 view, element m2 = pos  ->  z = (x[2c], x[2c+1]), c = pos*M1 + col
template <int L>
struct Pass1Load {
    const float *x;
    uint32_t m1, col0, n_steps;
    float mean;
    struct Ctx {
        const float *p;
        uint32_t i0;
        bool valid, fast;
    };
    __device__ __forceinline__ Ctx begin(int row, int base) const
    {
        constexpr int R = cfg_radix(L, 0), S = cfg_stride(L, 0);
        Ctx c;
        const uint32_t col = col0 + row;
        c.valid = col < m1;
        c.i0 = 2u * ((uint32_t)base * m1 + col);
        c.p = x + c.i0;
        // whole butterfly below n_steps: no per-element padding test
        c.fast = c.valid && (c.i0 + 2u * (uint32_t)((R - 1) * S) * m1 + 1u < n_steps);
        return c;
    }
    __device__ __forceinline__ float2 get(const Ctx &c, int off) const
    {
        const uint32_t d = 2u * (uint32_t)off * m1;
        if (c.fast) return *reinterpret_cast<const float2 *>(c.p + d);
        if (!c.valid) return make_float2(0.f, 0.f);
        // mean padding on load (demod_binary_resamp_cpu.c:130-133): samples >= n_steps take the series mean
        const uint32_t i = c.i0 + d;
        if (i + 1 < n_steps) return *reinterpret_cast<const float2 *>(c.p + d);
        return make_float2(i < n_steps ? c.p[d] : mean, mean);
    }
};

// last stage: output k2 = kb + q*(L/R) of column col, times the inter-pass twiddle
//     W_M^{col*k2} = W_M^{col*kb} * W_{M1*R}^{col*q}
// first factor: one exact-argument sincospi per butterfly (col*kb < M); second factor: table tb[col][q]
// generated in double precision.  Stored transposed: Y[k2][col].
template <int L>
struct Pass1Store {
    float2 *y;
    const float2 *tb;
    uint32_t m1, col0;
    float inv_m;
    static constexpr int RL = cfg_radix(L, cfg_stages(L) - 1);
    struct Ctx {
        float2 *p;
        const float2 *t;
        float2 a;
        bool valid;
    };
    __device__ __forceinline__ Ctx begin(int row, int base) const
    {
        Ctx c;
        const uint32_t col = col0 + row;
        c.valid = col < m1;
        const uint32_t kb = (uint32_t)freq_of_pos<L>(base);
        c.p = y + (size_t)kb * m1 + col;
        c.t = tb + (size_t)col * RL;
        float sn, cs;
        sincospif(-2.0f * (float)(col * kb) * inv_m, &sn, &cs);
        c.a = make_float2(cs, sn);
        return c;
    }
    __device__ __forceinline__ void put(const Ctx &c, int, int q, float2 v) const
    {
        if (!c.valid) return;
        const float2 w = (q == 0) ? c.a : cmul(c.a, __ldg(c.t + q));
        c.p[(size_t)q * (L / RL) * m1] = cmul(v, w);
    }
};

template <int L, int ROWS>
__global__ void __launch_bounds__(cfg_threads(L), cfg_ctas(L))
fft_pass1_kernel(const float *__restrict__ in, size_t in_stride, const uint32_t *__restrict__ n_steps_arr,
                 const float *__restrict__ mean_arr, float2 *__restrict__ work, uint32_t m1, uint32_t m,
                 const float2 *__restrict__ tw, const float2 *__restrict__ tb)
{
    extern __shared__ __align__(128) float2 sm[];
    constexpr int NS = cfg_stages(L);
    const int b = blockIdx.y;
    const uint32_t col0 = blockIdx.x * ROWS;
    const Pass1Load<L> gload{in + (size_t)b * in_stride, m1, col0, n_steps_arr[b], mean_arr[b]};
    const Pass1Store<L> gstore{work + (size_t)b * m, tb, m1, col0, 1.0f / (float)m};

    if constexpr (NS == 1) {
        tile_stage<L, ROWS, 0, MAP_ROW_FAST, false>(tw, gload, gstore);
    } else {
        SmemIO<L, ROWS, 1> io{sm};
        tile_stage<L, ROWS, 0, MAP_ROW_FAST, Layout<L, ROWS, 1>::ROT0>(tw, gload, io);
        __syncthreads();
        tile_middle<L, ROWS, 1, 1, NS - 1>(tw, sm);
        tile_stage<L, ROWS, NS - 1, MAP_ROW_FAST, false>(tw, io, gstore);
    }
}

// ------------------------------------------------------------------ pass 1 with TMA tensor copies (production shape)
// A pass-1 tile is 2048 rows of 4 complex columns: 32-byte runs, 8 KB apart.  Moved by per-lane loads and stores every
// run costs one L1 tag lookup, and those lookups (19 M of the kernel's 28 M L1 wavefronts per launch) set the speed of
// the plain kernel above.  Here the tile is fetched by eight cp.async.bulk.tensor boxes (256 rows x 32 bytes) that land
// DENSE in shared memory ([row][column]) without passing through the LSU, stage 0 reads it from there, and the last
// stage puts its results back in the same dense form for eight tensor stores.  The dense image and the padded
// exchange layout share the same shared memory, so stage 0 and the last stage read everything before anybody writes.
// Lane mappings and padding chosen with tools/pass1_tma_layout.py: every shared-memory access of the kernel is conflict-free.
// Arithmetic is the same as in fft_pass1_kernel (same butterflies, same twiddle construction): results are bit-identical.
constexpr int kTmaL = 2048, kTmaRows = 4, kTmaBox = 256;
#ifndef BRP_FFT_PASS2_TREE
#define BRP_FFT_PASS2_TREE 1
#endif
#ifndef BRP_FFT_L2_PREFETCH
#define BRP_FFT_L2_PREFETCH 148
#endif

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// PART < 0: plain length-2048 pass.  PART = q in 0..parts-1: the tile is sub-transform q of a length parts*2048 pass whose rows
// beyond the first 2048 are all mean padding (see fft_pass1_tma_kernel): the value handed to the 2048-point stages is the
// output q of the radix-`parts` first stage, y_q[p] = (x_p + c sum_{j>0} W_parts^{jq}) W_{m2}^{q p} with c = (mean, mean), i.e.
// x + (parts-1) c for q = 0 and (x - c) W^{qp} otherwise.
struct TmaDenseLoad {
    const float2 *dense;
    uint32_t m1, col0, n_steps;
    float mean;
    int part, parts;
    const float2 *twq; // W_{m2}^{q p}, p < 2048 (part > 0)
    bool twq_smem = false; // twq is the kernel's shared-memory copy of the row q = 1; q = 2 squares it
    struct Ctx {
        const float2 *p;
        const float2 *t;
        uint32_t i0;
        bool fast;
    };
    __device__ __forceinline__ Ctx begin(int row, int base) const
    {
        constexpr int R = cfg_radix(kTmaL, 0), S = cfg_stride(kTmaL, 0);
        Ctx c;
        c.p = dense + base * kTmaRows + row;
        c.t = twq + base;
        c.i0 = 2u * ((uint32_t)base * m1 + col0 + row);
        c.fast = c.i0 + 2u * (uint32_t)((R - 1) * S) * m1 + 1u < n_steps;
        return c;
    }
    __device__ __forceinline__ float2 get(const Ctx &c, int off) const
    {
        float2 v = c.p[off * kTmaRows]; // rows that were not fetched hold junk: never used, see below
        if (!c.fast) {
            const uint32_t i = c.i0 + 2u * (uint32_t)off * m1;
            if (i + 1 >= n_steps) v = make_float2(i < n_steps ? v.x : mean, mean); // mean padding, demod_binary_resamp_cpu.c:130-133
        }
        if (part < 0) return v;
        const float c2 = mean + mean; // parts = 3: the same operations as Dft<3> on (v, c, c)
        if (part == 0) return parts == 3 ? cadd(v, make_float2(c2, c2)) : cadd(v, make_float2(mean, mean));
        const float2 d = parts == 3 ? make_float2(fmaf(-0.5f, c2, v.x), fmaf(-0.5f, c2, v.y)) : make_float2(v.x - mean, v.y - mean);
        if (twq_smem) {
            const float2 w = c.t[off];
            return cmul(d, part == 2 ? cmul(w, w) : w);
        }
        return cmul(d, __ldg(c.t + off));
    }
};

// PARTS = 2, 3: pass-1 length PARTS x 2048 when only the first 2048 rows can hold samples (n_steps <= Nu <= N/PARTS: the
// reference's production padding -P 3.0, and -P 2.0).  The radix-PARTS first stage then has one variable and PARTS-1 constant
// inputs, its output blocks are independent 2048-point transforms of values derived from the same 2048 loaded rows, and
// output k2 = PARTS k' + q of block q lands in row PARTS k' + q.  One CTA per (column tile, q): same tile engine as the 2048
// case, the first stage folded into the load (for PARTS = 3 the same butterfly values as the generic radix-3 stage on these
// inputs; the twiddles W^{qp} are read from a table), a 4-D tensor map for the strided rows.
template <int PARTS>
__global__ void __launch_bounds__(512, 2)
fft_pass1_tma_kernel(const __grid_constant__ CUtensorMap in_map, const __grid_constant__ CUtensorMap out_map,
                     const uint32_t *__restrict__ n_steps_arr, const float *__restrict__ mean_arr, uint32_t m1, uint32_t m,
                     const float2 *__restrict__ tw, const float2 *__restrict__ tb, const float2 *__restrict__ tw_parts)
{
    const int part = PARTS > 1 ? (int)blockIdx.z : -1;
    constexpr int L = kTmaL, ROWS = kTmaRows, NS = cfg_stages(L), RL = cfg_radix(L, NS - 1);
    static_assert(NS == 3 && RL == 8 && cfg_threads(L) == 512, "written for 2048 = 16 x 16 x 8 with 512 threads");
    extern __shared__ __align__(128) float2 sm[];
    __shared__ __align__(8) unsigned long long bar;
    const int b = blockIdx.y;
    const uint32_t col0 = blockIdx.x * ROWS;
    const uint32_t n_steps = n_steps_arr[b];
    const float mean = mean_arr[b];
    // rows (positions m2) that hold at least one sample below n_steps; a row is 2*m1 samples
    const uint32_t rows_needed = min((uint32_t)L, (n_steps + 2u * m1 - 1u) / (2u * m1));
    const uint32_t boxes = (rows_needed + kTmaBox - 1) / kTmaBox;
    const uint32_t bar_a = smem_addr(&bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(boxes * kTmaBox * ROWS * 8u) : "memory");
        for (uint32_t k = 0; k < boxes; ++k)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                             smem_addr(sm + k * kTmaBox * ROWS)),
                         "l"(&in_map), "r"((int)(2u * col0)), "r"((int)(k * kTmaBox)), "r"(b), "r"(bar_a)
                         : "memory");
#if BRP_FFT_L2_PREFETCH > 0
        // pull the input tile of a CTA that starts about half a resident wave later into L2 (measured: pass 1 2.4 % faster at a
        // distance of 148 tiles, nothing at 296, slower beyond)
        {
            const uint32_t ahead = (blockIdx.y * gridDim.x + blockIdx.x) + BRP_FFT_L2_PREFETCH;
            const uint32_t by = ahead / gridDim.x, bx = ahead - by * gridDim.x;
            if (by < gridDim.y) {
                const uint32_t ns2 = n_steps_arr[by];
                const uint32_t need2 = min((uint32_t)L, (ns2 + 2u * m1 - 1u) / (2u * m1));
                for (uint32_t k = 0; k * kTmaBox < need2; ++k)
                    asm volatile("cp.async.bulk.prefetch.tensor.3d.L2.global.tile [%0, {%1, %2, %3}];" ::"l"(&in_map),
                                 "r"((int)(2u * bx * ROWS)), "r"((int)(k * kTmaBox)), "r"((int)by)
                                 : "memory");
            }
        }
#endif
    }
    __syncthreads();
    {
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n selp.u32 %0, 1, 0, p;\n}"
                         : "=r"(done)
                         : "r"(bar_a)
                         : "memory");
    }

    SmemIO<L, ROWS, 3> io{sm};
    const TmaDenseLoad dload{sm, m1, col0, n_steps, mean, part, PARTS, part > 0 ? tw_parts + (part - 1) * kTmaL : tw_parts};
    tile_stage<L, ROWS, 0, MAP_ROW_FAST, Layout<L, ROWS, 3>::ROT0, true>(tw, dload, io);
    __syncthreads();
    tile_middle<L, ROWS, 3, 1, NS - 1>(tw, sm);

    // last stage: radix 8 over consecutive positions.  Lanes take butterflies in transposed order (high nibble of the butterfly
    // number = x % 16, so the four butterflies of a half-warp produce outputs k2 = 0..3 mod 4: the dense stores hit 16 different
    // bank pairs) with the low nibble skewed by 2 t, t = (0, 1, 1, 2)[x % 4] (so that their padded READ positions differ mod 4 as
    // well: with the row pitch of 12 mod 16 that makes the read conflict-free too).
    constexpr int PER = ROWS * (L / RL) / 512; // butterflies per thread (2)
    float2 v[PER][RL];
    int rows_[PER], kb_[PER];
#pragma unroll
    for (int it = 0; it < PER; ++it) {
        const int item = threadIdx.x + it * 512;
        const int row = item % ROWS, x = item / ROWS;
        const int a4 = x & 3;
        const int bf = BRP_FFT_P1_LAYOUT == 1 ? 16 * (x % 16) + x / 16 : 16 * (x % 16) + ((x / 16 + 2 * (a4 - (a4 >> 1))) & 15);
        const int base = bf * RL;
        rows_[it] = row;
        kb_[it] = freq_of_pos<L>(base);
        const float2 *c = io.begin(row, base);
#pragma unroll
        for (int j = 0; j < RL; ++j) v[it][j] = io.get(const_cast<float2 *>(c), j);
    }
    __syncthreads(); // the dense image overwrites the padded tile
    const float inv_m = 1.0f / (float)m;
#pragma unroll
    for (int it = 0; it < PER; ++it) {
        Dft<RL>::run(v[it]);
        const uint32_t col = col0 + rows_[it];
        float sn, cs;
        const uint32_t k2b = PARTS > 1 ? (uint32_t)PARTS * (uint32_t)kb_[it] + (uint32_t)part : (uint32_t)kb_[it];
        sincospif(-2.0f * (float)(col * k2b) * inv_m, &sn, &cs);
        const float2 a = make_float2(cs, sn);
        const float2 *__restrict__ t = tb + (size_t)col * RL;
        float2 *d = sm + kb_[it] * ROWS + rows_[it];
#pragma unroll
        for (int q = 0; q < RL; ++q) {
            const float2 w = (q == 0) ? a : cmul(a, __ldg(t + q));
            d[q * (L / RL) * ROWS] = cmul(v[it][out_slot<RL>(q)], w);
        }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < L / kTmaBox; ++k) {
            if constexpr (PARTS > 1)
                asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%1, %2, %3, %4}], [%5];" ::"l"(&out_map),
                             "r"((int)(2u * col0)), "r"(part), "r"(k * kTmaBox), "r"(b), "r"(smem_addr(sm + k * kTmaBox * ROWS))
                             : "memory");
            else
                asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];" ::"l"(&out_map),
                             "r"((int)(2u * col0)), "r"(k * kTmaBox), "r"(b), "r"(smem_addr(sm + k * kTmaBox * ROWS))
                             : "memory");
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
}


// ------------------------------------------------------------------ pass 1, persistent (three tile buffers per SM)
// In fft_pass1_tma_kernel a CTA loads its tile, waits for it (a fifth of all warp samples: profiles/README.md), transforms it
// and leaves; two CTAs per SM cover each other's waits only in part, and a third does not fit the register file.  Here ONE
// CTA of 1024 threads per SM stays for the whole launch and owns THREE tile buffers.  Its two halves of 512 threads work like
// the two CTAs did -- same tile engine, same arithmetic, synchronising among themselves with named barriers -- on the tiles
// n = 0, 1, 2, ... of the CTA in turn (even n: half 0, odd n: half 1; tile n lives in buffer n mod 3), and the tensor load of
// tile n + 3 is issued into a buffer as soon as the stores of tile n have left it: a tile is under way for about half a tile time
// before anybody waits for it.  Per buffer one mbarrier ("full"), completed by the tensor loads of each use in turn.  There is
// no L2 prefetch here: the tensor load itself is the prefetch (measured: cp.async.bulk.prefetch.tensor of the tile two or three
// loads ahead, issued with each load, makes the kernel 50-70 % slower; waiting for the stores only inside stage 0 of the next tile
// 13 % slower -- what counts is how early the load of the other half's next tile goes out).
#ifndef BRP_FFT_P1_PERSIST
#define BRP_FFT_P1_PERSIST 1
#endif
#ifndef BRP_FFT_P1_TWQ_SMEM
#define BRP_FFT_P1_TWQ_SMEM 1 // 0: the persistent pass 1 with PARTS > 1 reads W^{qp} from global memory (A/B runs)
#endif
#ifndef BRP_FFT_P1_TB_SMEM
#define BRP_FFT_P1_TB_SMEM 1 // 0: the last stage of the persistent pass 1 reads tb[col][q] from global memory (A/B runs)
#endif
#ifndef BRP_FFT_P1_TW_SMEM
#define BRP_FFT_P1_TW_SMEM 1 // 0: the persistent pass 1 reads its in-tile twiddles from global memory in every tile (A/B runs)
#endif
#ifndef BRP_FFT_P1_BOXWISE
#define BRP_FFT_P1_BOXWISE 1 // 0: reload a buffer only after ALL stores of the tile have read it (A/B runs)
#endif

template <int PARTS>
__global__ void __launch_bounds__(1024, 1)
fft_pass1_persist_kernel(const __grid_constant__ CUtensorMap in_map, const __grid_constant__ CUtensorMap out_map,
                         const uint32_t *__restrict__ n_steps_arr, const float *__restrict__ mean_arr, uint32_t m1, uint32_t m,
                         const float2 *__restrict__ tw, const float2 *__restrict__ tb, const float2 *__restrict__ tw_parts, uint32_t tiles_x,
                         uint32_t total_tiles)
{
    constexpr int L = kTmaL, ROWS = kTmaRows, NS = cfg_stages(L), RL = cfg_radix(L, NS - 1);
    constexpr int BUF = Layout<L, ROWS, 3>::BYTES; // 70 016 bytes: dense image and padded exchange layout share it
    static_assert(BUF % 128 == 0, "tile buffers must keep the 128-byte alignment of tensor copies");
    extern __shared__ __align__(128) unsigned char smraw[];
    __shared__ __align__(8) unsigned long long full[3];
    // second factors of the inter-pass twiddle for the tile's four columns (tb[col][0..RL)): 256 contiguous bytes per tile, fetched by a
    // bulk copy together with the tile instead of seven global loads per thread in the last stage
    __shared__ __align__(128) float2 tbs[3][BRP_FFT_P1_TB_SMEM ? ROWS * RL : 1];
    const int half = threadIdx.x >> 9, t = threadIdx.x & 511;
    const GroupSync<512> sy{1 + half, t};
    // tiles of this CTA: g(n) = blockIdx.x + n * gridDim.x  ->  (column tile bx, part, template b), bx fastest
    const uint32_t my_tiles = total_tiles > blockIdx.x ? (total_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0u;
    auto decode = [&](uint32_t n, uint32_t &bx, int &part, uint32_t &b) {
        const uint32_t g = blockIdx.x + n * gridDim.x, rest = g / tiles_x;
        bx = g - rest * tiles_x;
        part = PARTS > 1 ? (int)(rest % PARTS) : -1;
        b = PARTS > 1 ? rest / PARTS : rest;
    };
    // tensor load of tile n into buffer n % 3 (one thread)
    auto issue_load = [&](uint32_t n) {
        uint32_t bx, b;
        int part;
        decode(n, bx, part, b);
        const uint32_t ns = n_steps_arr[b];
        const uint32_t rows_needed = min((uint32_t)L, (ns + 2u * m1 - 1u) / (2u * m1));
        const uint32_t boxes = (rows_needed + kTmaBox - 1) / kTmaBox;
        const uint32_t bar_a = smem_addr(&full[n % 3]);
        float2 *dst = reinterpret_cast<float2 *>(smraw + (size_t)(n % 3) * BUF);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a),
                     "r"(boxes * kTmaBox * ROWS * 8u + (BRP_FFT_P1_TB_SMEM ? ROWS * RL * 8u : 0u))
                     : "memory");
        if (BRP_FFT_P1_TB_SMEM)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(&tbs[n % 3][0])),
                         "l"(tb + (size_t)bx * ROWS * RL), "r"(ROWS * RL * 8), "r"(bar_a)
                         : "memory");
        for (uint32_t k = 0; k < boxes; ++k)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                             smem_addr(dst + k * kTmaBox * ROWS)),
                         "l"(&in_map), "r"((int)(2u * bx * ROWS)), "r"((int)(k * kTmaBox)), "r"((int)b), "r"(bar_a)
                         : "memory");
    };
    // The in-tile twiddles the kernel reads are thread constants: W^p of stage 0's product tree (128 values) and the table of
    // stage 1 (15 x 8).  Fetched from global memory in every tile they miss the few KB of L1 the three tile buffers leave, and the first
    // of them -- behind the wait for the tile, at the head of the dependency chain -- was the largest single stall of the kernel (10 %
    // of the warp samples on that one load).  A 2 KB copy in shared memory serves them instead.
    static_assert(NS == 3 && ROWS * (L / cfg_radix(L, 0)) == 512 && !Layout<L, ROWS, 3>::ROT0, "stage 0: one butterfly per thread, plain row-fast lanes");
    constexpr int TW0 = cfg_stride(L, 0), TW1 = (cfg_radix(L, 1) - 1) * cfg_stride(L, 1);
    __shared__ float2 tws[BRP_FFT_P1_TW_SMEM ? TW0 + TW1 : 1];
    if (BRP_FFT_P1_TW_SMEM && threadIdx.x < TW0 + TW1)
        tws[threadIdx.x] = __ldg(tw + (threadIdx.x < TW0 ? cfg_tw_offset(L, 0) + (int)threadIdx.x : cfg_tw_offset(L, 1) + (int)threadIdx.x - TW0));
    // pointers such that the stages' own offsets (cfg_tw_offset + ...) land in the copy
    const float2 *const tw_s0 = BRP_FFT_P1_TW_SMEM ? tws - cfg_tw_offset(L, 0) : tw;
    const float2 *const tw_s1 = BRP_FFT_P1_TW_SMEM ? tws + TW0 - cfg_tw_offset(L, 1) : tw;
    // PARTS > 1: the twiddles W^{qp} of the folded first stage (sixteen per thread and tile) come from a shared-memory copy of the row
    // q = 1 as well (16 KB, all that is left beside the three buffers); the row q = 2 is its square
    constexpr bool TWQ = BRP_FFT_P1_TWQ_SMEM != 0 && PARTS > 1;
    __shared__ float2 twq1[TWQ ? L : 1];
    if (TWQ)
        for (int i = threadIdx.x; i < L; i += 1024) twq1[i] = __ldg(tw_parts + i);
    if (threadIdx.x == 0) {
        for (int k = 0; k < 3; ++k) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&full[k])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        for (uint32_t n = 0; n < 3 && n < my_tiles; ++n) issue_load(n);
    }
    __syncthreads();

    for (uint32_t n = (uint32_t)half; n < my_tiles; n += 2) {
        uint32_t bx, b;
        int part;
        decode(n, bx, part, b);
        float2 *sm = reinterpret_cast<float2 *>(smraw + (size_t)(n % 3) * BUF);
        const uint32_t col0 = bx * ROWS;
        const uint32_t n_steps = n_steps_arr[b];
        const float mean = mean_arr[b];
        {
            const uint32_t bar_a = smem_addr(&full[n % 3]), parity = (n / 3) & 1u;
            uint32_t done = 0;
            while (!done)
                asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                             : "=r"(done)
                             : "r"(bar_a), "r"(parity)
                             : "memory");
        }
        SmemIO<L, ROWS, 3> io{sm};
        const TmaDenseLoad dload{sm, m1, col0, n_steps, mean, part, PARTS, TWQ ? twq1 : (part > 0 ? tw_parts + (part - 1) * kTmaL : tw_parts), TWQ};
        tile_stage<L, ROWS, 0, MAP_ROW_FAST, Layout<L, ROWS, 3>::ROT0, true, false, BRP_FFT_P1_TW_SMEM != 0>(tw_s0, dload, io, sy);
        sy.sync();
        tile_middle<L, ROWS, 3, 1, NS - 1, GroupSync<512>, BRP_FFT_P1_TW_SMEM != 0>(tw_s1, sm, sy);

        constexpr int PER = ROWS * (L / RL) / 512;
        float2 v[PER][RL];
        int rows_[PER], kb_[PER];
#pragma unroll
        for (int it = 0; it < PER; ++it) {
            const int item = t + it * 512;
            const int row = item % ROWS, x = item / ROWS;
            const int a4 = x & 3;
            const int bf = BRP_FFT_P1_LAYOUT == 1 ? 16 * (x % 16) + x / 16 : 16 * (x % 16) + ((x / 16 + 2 * (a4 - (a4 >> 1))) & 15);
            const int base = bf * RL;
            rows_[it] = row;
            kb_[it] = freq_of_pos<L>(base);
            const float2 *c = io.begin(row, base);
#pragma unroll
            for (int j = 0; j < RL; ++j) v[it][j] = io.get(const_cast<float2 *>(c), j);
        }
        sy.sync(); // the dense image overwrites the padded tile
        const float inv_m = 1.0f / (float)m;
#pragma unroll
        for (int it = 0; it < PER; ++it) {
            Dft<RL>::run(v[it]);
            const uint32_t col = col0 + rows_[it];
            float sn, cs;
            const uint32_t k2b = PARTS > 1 ? (uint32_t)PARTS * (uint32_t)kb_[it] + (uint32_t)part : (uint32_t)kb_[it];
            sincospif(-2.0f * (float)(col * k2b) * inv_m, &sn, &cs);
            const float2 a = make_float2(cs, sn);
            const float2 *__restrict__ tq = BRP_FFT_P1_TB_SMEM ? &tbs[n % 3][rows_[it] * RL] : tb + (size_t)col * RL;
            float2 *d = sm + kb_[it] * ROWS + rows_[it];
#pragma unroll
            for (int q = 0; q < RL; ++q) {
                const float2 w = (q == 0) ? a : cmul(a, BRP_FFT_P1_TB_SMEM ? tq[q] : __ldg(tq + q));
                d[q * (L / RL) * ROWS] = cmul(v[it][out_slot<RL>(q)], w);
            }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        sy.sync();
        if (t == 0) {
            // one bulk group per box, so that each box of the buffer can be reloaded as soon as ITS store has read it
#pragma unroll
            for (int k = 0; k < L / kTmaBox; ++k) {
                if constexpr (PARTS > 1)
                    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%1, %2, %3, %4}], [%5];" ::"l"(&out_map),
                                 "r"((int)(2u * col0)), "r"(part), "r"(k * kTmaBox), "r"((int)b), "r"(smem_addr(sm + k * kTmaBox * ROWS))
                                 : "memory");
                else
                    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];" ::"l"(&out_map),
                                 "r"((int)(2u * col0)), "r"(k * kTmaBox), "r"((int)b), "r"(smem_addr(sm + k * kTmaBox * ROWS))
                                 : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            // Reload the buffer with tile n + 3 AT ONCE, box by box behind the stores: this thread's warp then enters its next tile
            // a microsecond late, but the tile it loads -- the other half's next one -- gets its full lead.  (Measured: deferring
            // the wait to a point of the next tile where it costs nothing makes the loads late: 515-558 us instead of 455 us per 64
            // templates.)
            uint32_t boxes3 = 0, bx3 = 0, b3 = 0;
            if (n + 3 < my_tiles) {
                int part3;
                decode(n + 3, bx3, part3, b3);
                const uint32_t need3 = min((uint32_t)L, (n_steps_arr[b3] + 2u * m1 - 1u) / (2u * m1));
                boxes3 = (need3 + kTmaBox - 1) / kTmaBox;
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(&full[n % 3])),
                             "r"(boxes3 * kTmaBox * ROWS * 8u + (BRP_FFT_P1_TB_SMEM ? ROWS * RL * 8u : 0u))
                             : "memory");
                if (BRP_FFT_P1_TB_SMEM) // everybody of this half is past its reads of the slot (the barrier above)
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                     smem_addr(&tbs[n % 3][0])),
                                 "l"(tb + (size_t)bx3 * ROWS * RL), "r"(ROWS * RL * 8), "r"(smem_addr(&full[n % 3]))
                                 : "memory");
            }
            static_assert(L / kTmaBox == 8, "eight boxes per tile");
#define BRP_RELOAD_BOX(k)                                                                                                               \
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(BRP_FFT_P1_BOXWISE ? 7 - (k) : 0) : "memory"); /* store k has read its box */ \
    if ((uint32_t)(k) < boxes3)                                                                                                         \
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"( \
                         smem_addr(sm + (k) * kTmaBox * ROWS)),                                                                         \
                     "l"(&in_map), "r"((int)(2u * bx3 * ROWS)), "r"((int)((k) * kTmaBox)), "r"((int)b3), "r"(smem_addr(&full[n % 3]))    \
                     : "memory");
            BRP_RELOAD_BOX(0) BRP_RELOAD_BOX(1) BRP_RELOAD_BOX(2) BRP_RELOAD_BOX(3)
            BRP_RELOAD_BOX(4) BRP_RELOAD_BOX(5) BRP_RELOAD_BOX(6) BRP_RELOAD_BOX(7)
#undef BRP_RELOAD_BOX
        }
    }
}

// ------------------------------------------------------------------ pass 2 + untangle + power
// tile rows: r < T2 -> k2 = a + r ; r >= T2 -> k2 = (m2 - (a + r - T2)) mod m2 (the mirror row).
// Only rows k2 <= m2/2 (and their mirrors) are owned by a block; the grid covers k2 = 0..m2/2.
template <int L, int ROWS>
struct Pass2Load {
    const float2 *y;
    uint32_t a, m2;
    __device__ __forceinline__ const float2 *begin(int row, int base) const
    {
        constexpr int T2 = ROWS / 2;
        const uint32_t f = a + (uint32_t)(row < T2 ? row : row - T2);
        if (2 * f > m2) return nullptr;
        const uint32_t k2 = row < T2 ? f : (m2 - f) % m2;
        return y + (size_t)k2 * L + base;
    }
    __device__ __forceinline__ float2 get(const float2 *c, int off) const { return c ? c[off] : make_float2(0.f, 0.f); }
};

// MODE 0: power spectrum norm*|X|^2 of the real-input transform (the per-template hot path)
// MODE 1: the real-input transform itself, X[0..M] as float2 (whitening pre-step, forward direction)
// MODE 2: the plain complex transform Z[0..M) written as scale*conj(Z) pairs, i.e. out[2k] = s Re Z[k], out[2k+1] = -s Im Z[k]
//         (backward real transform = conj(FFT(conj(.))), whitening pre-step; `ps_count` = number of float2 pairs wanted)
enum { PASS2_POWER = 0, PASS2_SPECTRUM = 1, PASS2_CONJ = 2 };

template <int L, int ROWS, int MODE, bool TMA_OUT>
__device__ __forceinline__ void pass2_body(const float2 *__restrict__ work, float *__restrict__ ps, size_t ps_stride, uint32_t ps_count,
                                           uint32_t m2, uint32_t m, float norm, const float2 *__restrict__ tw,
                                           const float2 *__restrict__ tw_a, const float2 *__restrict__ tw_b, float2 *sm,
                                           const CUtensorMap *out_map)
{
    constexpr int NS = cfg_stages(L);
    constexpr int T2 = ROWS / 2;
    constexpr uint32_t M1 = L;
    const int b = blockIdx.y;
    const uint32_t a = blockIdx.x * T2;
    float *P = ps + (size_t)b * ps_stride;
    SmemIO<L, ROWS, 2> io{sm};
    const Pass2Load<L, ROWS> gload{work + (size_t)b * m, a, m2};

    // twiddles W^{pq} of this stage by a power tree from W^p: the table loads were a seventh of the kernel's L1 traffic
    // (measured: 203 -> 191 us per 32 templates)
    // (only for the power spectrum of the search: the whitening modes keep the double-rounded table twiddles, the tree's
    // few-ulp twiddle error doubles the error of the whitened series)
    tile_stage<L, ROWS, 0, MAP_BF_FAST, false, false, (BRP_FFT_PASS2_TREE != 0 && MODE == PASS2_POWER)>(tw, gload, io);
    __syncthreads();
    tile_middle<L, ROWS, 2, 1, NS>(tw, sm);

    if constexpr (MODE == PASS2_CONJ) {
        // every row of the tile is an output row here; rows are owned once: k2 <= m2/2 as "A" rows, the rest as mirrors
        float2 *out = reinterpret_cast<float2 *>(ps) + (size_t)b * ps_stride;
        for (int item = threadIdx.x; item < ROWS * (int)M1; item += cfg_threads(L)) {
            const int r = item % ROWS;
            const uint32_t k1 = item / ROWS;
            const uint32_t f = a + (uint32_t)(r < T2 ? r : r - T2);
            if (2 * f > m2) continue;
            if (r >= T2 && (f == 0 || 2 * f == m2)) continue; // self-mirrored rows already written as "A" rows
            const uint32_t k2 = r < T2 ? f : m2 - f;
            const uint32_t k = k2 + m2 * k1;
            if (k >= ps_count) continue;
            const float2 z = *io.begin(r, pos_of_freq<L>((int)k1));
            out[k] = make_float2(norm * z.x, -norm * z.y);
        }
        return;
    }
    // ---- untangle + power.  Pair (k, M-k): Z[k] in row r at k1, Z[M-k] in row r+T2 at k1' = M1-1-k1
    // (k1' = (M1-k1) mod M1 on the self-mirrored row k2 = 0).
    //   E = (Z[k] + conj Z[M-k])/2,  O = (Z[k] - conj Z[M-k])/(2i)
    //   X[k] = E + W_N^k O,  X[M-k] = conj(E - W_N^k O)            (N = 2M)
    // Interior tiles (no self-mirrored row): thread t owns row r = t % T2 and k1 = t / T2 + KSTEP it.  KSTEP is a multiple of
    // L / R_last, so only the last digit of k1 changes from one iteration to the next: the digit-reversed position (and its
    // padded address, the padding units being multiples of R_last) advances by the constant CPOS, the mirror operand recedes
    // by it, and the whole index computation leaves the loop.
    constexpr int NT2 = cfg_threads(L), RL2 = cfg_radix(L, NS - 1), KSTEP = NT2 / T2, UNIT = L / RL2;
    constexpr bool AFFINE = MODE == PASS2_POWER && NT2 % T2 == 0 && KSTEP % UNIT == 0 && (T2 * L) % NT2 == 0;
    if constexpr (AFFINE) {
        if (a >= 1 && 2 * (a + T2 - 1) < m2) {
            using Lay = Layout<L, ROWS, 2>;
            constexpr int CPOS = KSTEP / UNIT, ITERS = T2 * L / NT2;
            static_assert((Lay::C4 == 0 || 16 % RL2 == 0) && (Lay::C5 == 0 || 32 % RL2 == 0) && (Lay::C8 == 0 || 256 % RL2 == 0),
                          "padding units must be multiples of the last radix");
            const int r = threadIdx.x % T2, k10 = threadIdx.x / T2;
            const uint32_t k2 = a + r;
            const float2 *pa = io.begin(r, pos_of_freq<L>(k10));
            const float2 *pb = io.begin(r + T2, pos_of_freq<L>((int)L - 1 - k10));
            const float2 wa = __ldg(&tw_a[k2]);
            const float2 *__restrict__ twb = tw_b + k10;
            const uint32_t k0 = k2 + m2 * (uint32_t)k10;
            float *Pk = P + k0, *Pm = P + (m - k0);
            const uint32_t step = (uint32_t)KSTEP * m2;
            const float qnorm = 0.25f * norm;
            if constexpr (TMA_OUT) {
                static_assert(L == 1024 && ROWS == 8 && ITERS == 16 && CPOS == 2 && KSTEP == 64, "staging image below is written for this shape");
                // The 16-byte runs of this tile's outputs (four consecutive k2 per k1) cost one L1 tag lookup each when stored by
                // the lanes.  The powers of the rows k2 = a..a+3 are therefore staged densely in shared memory, [k1][4] (over the
                // tile, after everybody has read its operands), and four tensor stores (256 rows x 16 bytes) write them out; rows
                // at or beyond ps_count / m2 are clipped by the tensor map.  The mirror rows m2-a-3..m2-a start one float past a
                // 16-byte boundary, which a tensor copy cannot address, so they keep the per-lane stores.
                float2 za[16], zb[16];
#pragma unroll
                for (int it = 0; it < 16; ++it) {
                    za[it] = pa[2 * it];
                    zb[it] = pb[-2 * it];
                }
                __syncthreads();
                float *stage_a = reinterpret_cast<float *>(sm); // 1024 x 4 floats
#pragma unroll
                for (int it = 0; it < 16; ++it) {
                    // 2E, 2O and the factor 1/4 moved into norm: scaling by powers of two is exact, the powers are unchanged
                    const float2 e = make_float2(za[it].x + zb[it].x, za[it].y - zb[it].y);
                    const float2 o = make_float2(za[it].y + zb[it].y, zb[it].x - za[it].x);
                    const float2 w = cmul(wa, __ldg(twb + 64 * it));
                    const float2 wo = cmul(w, o);
                    const float2 xp = cadd(e, wo), xm = csub(e, wo);
                    stage_a[4 * (k10 + 64 * it) + r] = qnorm * fmaf(xp.x, xp.x, xp.y * xp.y);
                    const uint32_t k = k0 + (uint32_t)it * step;
                    if (m - k < ps_count) Pm[-(ptrdiff_t)((size_t)it * step)] = qnorm * fmaf(xm.x, xm.x, xm.y * xm.y);
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncthreads();
                if (threadIdx.x == 0) {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];" ::"l"(out_map),
                                     "r"((int)a), "r"(256 * j), "r"(b), "r"(smem_addr(stage_a + 1024 * j))
                                     : "memory");
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                }
                return;
            }
#pragma unroll
            for (int it = 0; it < ITERS; ++it) {
                const float2 za = pa[CPOS * it], zb = pb[-CPOS * it];
                const float2 e = make_float2(za.x + zb.x, za.y - zb.y); // 2E, 2O: the factor 1/4 sits in qnorm (exact)
                const float2 o = make_float2(za.y + zb.y, zb.x - za.x);
                const float2 w = cmul(wa, __ldg(twb + KSTEP * it));
                const float2 wo = cmul(w, o);
                const float2 xp = cadd(e, wo), xm = csub(e, wo);
                const uint32_t k = k0 + (uint32_t)it * step;
                if (k < ps_count) Pk[(size_t)it * step] = qnorm * fmaf(xp.x, xp.x, xp.y * xp.y);
                if (m - k < ps_count) Pm[-(ptrdiff_t)((size_t)it * step)] = qnorm * fmaf(xm.x, xm.x, xm.y * xm.y);
            }
            return;
        }
    }
#pragma unroll 4
    for (int item = threadIdx.x; item < T2 * (int)M1; item += cfg_threads(L)) {
        const int r = item % T2;
        const uint32_t k1 = item / T2;
        const uint32_t k2 = a + r;
        if (2 * k2 > m2) continue;
        uint32_t k1m;
        if (k2 == 0) {
            if (k1 > M1 / 2) continue; // pairs (k1, M1-k1) once
            k1m = (M1 - k1) % M1;
        } else {
            if (2 * k2 == m2 && k1 >= M1 / 2) continue; // row m2/2 mirrors onto itself
            k1m = M1 - 1 - k1;
        }
        const float2 za = *io.begin(r, pos_of_freq<L>((int)k1));
        const float2 zb = *io.begin(r + T2, pos_of_freq<L>((int)k1m));
        const float2 e = make_float2(0.5f * (za.x + zb.x), 0.5f * (za.y - zb.y));
        const float2 d = make_float2(0.5f * (za.x - zb.x), 0.5f * (za.y + zb.y));
        const float2 o = make_float2(d.y, -d.x);
        const float2 w = cmul(__ldg(&tw_a[k2]), __ldg(&tw_b[k1]));
        const float2 wo = cmul(w, o);
        const uint32_t k = k2 + m2 * k1;
        const uint32_t km = m - k;
        const float2 xp = cadd(e, wo), xm = csub(e, wo);
        if constexpr (MODE == PASS2_SPECTRUM) {
            float2 *X = reinterpret_cast<float2 *>(ps) + (size_t)b * ps_stride;
            if (k < ps_count) X[k] = xp;
            if (km < ps_count && km != k) X[km] = make_float2(xm.x, -xm.y);
        } else {
            if (k < ps_count) P[k] = (k == 0) ? 0.0f : norm * fmaf(xp.x, xp.x, xp.y * xp.y);
            if (km < ps_count && km != k) P[km] = norm * fmaf(xm.x, xm.x, xm.y * xm.y);
        }
    }
}

template <int L, int ROWS, int MODE>
__global__ void __launch_bounds__(cfg_threads(L), cfg_ctas(L))
fft_pass2_kernel(const float2 *__restrict__ work, float *__restrict__ ps, size_t ps_stride, uint32_t ps_count,
                 uint32_t m2, uint32_t m, float norm, const float2 *__restrict__ tw,
                 const float2 *__restrict__ tw_a, const float2 *__restrict__ tw_b)
{
    extern __shared__ __align__(128) float2 sm[];
    pass2_body<L, ROWS, MODE, false>(work, ps, ps_stride, ps_count, m2, m, norm, tw, tw_a, tw_b, sm, nullptr);
}

// production shape of the per-template hot path: powers leave through tensor stores (see pass2_body)
__global__ void __launch_bounds__(256, 2)
fft_pass2_tma_kernel(const __grid_constant__ CUtensorMap out_map, const float2 *__restrict__ work, float *__restrict__ ps,
                     size_t ps_stride, uint32_t ps_count, uint32_t m2, uint32_t m, float norm, const float2 *__restrict__ tw,
                     const float2 *__restrict__ tw_a, const float2 *__restrict__ tw_b)
{
    extern __shared__ __align__(128) float2 sm[];
    pass2_body<1024, 8, PASS2_POWER, true>(work, ps, ps_stride, ps_count, m2, m, norm, tw, tw_a, tw_b, sm, &out_map);
}

// ------------------------------------------------------------------ host side
typedef void (*pass1_fn)(const float *, size_t, const uint32_t *, const float *, float2 *, uint32_t, uint32_t, const float2 *,
                         const float2 *);
typedef void (*pass2_fn)(const float2 *, float *, size_t, uint32_t, uint32_t, uint32_t, float, const float2 *, const float2 *,
                         const float2 *);

struct TileInfo {
    int L, rows, threads, smem1, smem2, tw_size, last_radix;
    pass1_fn p1;
    pass2_fn p2, p2_spectrum, p2_conj;
};

template <int L>
TileInfo make_info()
{
    constexpr int ROWS = cfg_rows(L);
    TileInfo t;
    t.L = L;
    t.rows = ROWS;
    t.threads = cfg_threads(L);
    t.smem1 = Layout<L, ROWS, 1>::BYTES;
    t.smem2 = Layout<L, ROWS, 2>::BYTES;
    t.tw_size = cfg_tw_size(L);
    t.last_radix = cfg_radix(L, cfg_stages(L) - 1);
    t.p1 = fft_pass1_kernel<L, ROWS>;
    t.p2 = fft_pass2_kernel<L, ROWS, PASS2_POWER>;
    t.p2_spectrum = fft_pass2_kernel<L, ROWS, PASS2_SPECTRUM>;
    t.p2_conj = fft_pass2_kernel<L, ROWS, PASS2_CONJ>;
    return t;
}

const TileInfo *find_tile(uint32_t L)
{
    static const TileInfo tiles[] = {
        make_info<16>(),   make_info<32>(),   make_info<64>(),   make_info<128>(),  make_info<256>(),
        make_info<512>(),  make_info<1024>(), make_info<2048>(), make_info<4096>(), make_info<48>(),
        make_info<96>(),   make_info<192>(),  make_info<384>(),  make_info<768>(),  make_info<1536>(),
        make_info<3072>(), make_info<6144>(),
    };
    for (const TileInfo &t : tiles)
        if ((uint32_t)t.L == L) return &t;
    return nullptr;
}

// host generator for the in-tile twiddles of length L (double precision, rounded once)
template <int L>
void gen_tile_tw(std::vector<float2> &out)
{
    out.resize(cfg_tw_size(L) > 0 ? cfg_tw_size(L) : 1);
    for (int s = 0; s + 1 < cfg_stages(L); ++s) {
        const int R = cfg_radix(L, s), n = cfg_n(L, s), S = cfg_stride(L, s), off = cfg_tw_offset(L, s);
        for (int q = 1; q < R; ++q)
            for (int p = 0; p < S; ++p) {
                const double ang = -2.0 * M_PI * (double)((long long)p * q % n) / (double)n;
                out[off + (q - 1) * S + p] = make_float2((float)cos(ang), (float)sin(ang));
            }
    }
}

void gen_tile_tw_dispatch(uint32_t L, std::vector<float2> &out)
{
    switch (L) {
#define C(X) case X: gen_tile_tw<X>(out); break;
        C(16) C(32) C(64) C(128) C(256) C(512) C(1024) C(2048) C(4096)
        C(48) C(96) C(192) C(384) C(768) C(1536) C(3072) C(6144)
#undef C
        default: out.clear();
    }
}

float2 *upload(const std::vector<float2> &v)
{
    float2 *d = nullptr;
    if (cudaMalloc(&d, sizeof(float2) * v.size()) != cudaSuccess) return nullptr;
    if (cudaMemcpy(d, v.data(), sizeof(float2) * v.size(), cudaMemcpyHostToDevice) != cudaSuccess) {
        cudaFree(d);
        return nullptr;
    }
    return d;
}

} // namespace

int brp_fft_plan_create(BrpFftPlan *p, uint32_t n, uint32_t nu_hint)
{
    memset(p, 0, sizeof(*p));
    if (n < 512 || (n & 1)) return RADPUL_CUDA_FFT_PLAN;
    uint32_t m = n / 2, odd = m;
    int a = 0;
    while ((odd & 1) == 0) { odd >>= 1; ++a; }
    if (odd != 1 && odd != 3) return RADPUL_CUDA_FFT_PLAN; // lengths 2^a and 3*2^a only (like the OpenCL backend)
    // pass 2 length m1 = 2^floor(a/2); pass 1 length m2 = m / m1 carries the factor 3 if there is one
    uint32_t m1 = 1u << (a / 2), m2 = m / m1;
    // m = 2^22 (n = 2^23) splits 2048 x 2048 by the rule above; when only the first half of the series can hold samples
    // (padding 2) 1024 x 4096 is better: pass 1 runs as two 2048-row sub-transforms on the tensor-copy engine and pass 2 gets
    // its fastest shape
    if (odd == 1 && m2 == 2048 && m1 == 2048 && nu_hint != 0 && nu_hint <= 2u * 1024u * 2048u) {
        m1 = 1024;
        m2 = 4096;
    }
    // shorter series (n = 2^20, 2^21 and three times that): a 2048-row pass 1 here too, so that it runs on the persistent
    // tensor-copy kernel, and a shorter pass 2 (measured at 2^20: pass 1 117 instead of 153 us per 64 templates, pass 2 unchanged;
    // whole search 180 k instead of 162 k templates/s; 2^21: 96.8 k instead of 86.1 k).  With the factor 3 the 6144 rows run as
    // three 2048-row parts, which needs all samples inside the first 2048 rows.
    if ((a == 19 || a == 20) && (odd == 1 || (nu_hint != 0 && nu_hint <= 2u * (m / 6144u) * 2048u))) {
        m2 = 2048u * odd;
        m1 = m / m2;
    }
    if (const char *force = getenv("BRP_FFT_M2")) { // measurement aid: another split of the same length
        const uint32_t f = (uint32_t)atoi(force);
        if (f >= 16 && m % f == 0 && find_tile(f) && find_tile(m / f)) {
            m2 = f;
            m1 = m / f;
        }
    }
    const TileInfo *t1 = find_tile(m2), *t2 = find_tile(m1);
    if (!t1 || !t2) return RADPUL_CUDA_FFT_PLAN;
    p->n = n; p->m = m; p->m1 = m1; p->m2 = m2; p->radix3 = (odd == 3);
    p->smem_pass1 = t1->smem1; p->smem_pass2 = t2->smem2;
    p->launches_per_batch = 2;

    std::vector<float2> tw;
    gen_tile_tw_dispatch(m2, tw);
    p->tw_pass1 = upload(tw);
    gen_tile_tw_dispatch(m1, tw);
    p->tw_pass2 = upload(tw);
    // second factor of the inter-pass twiddle: tb[col][q] = exp(-2 pi i col q / (m1 * R_last))
    tw.resize((size_t)m1 * t1->last_radix);
    for (uint32_t col = 0; col < m1; ++col)
        for (int q = 0; q < t1->last_radix; ++q) {
            const double ang = -2.0 * M_PI * (double)((unsigned long long)col * q) / ((double)m1 * t1->last_radix);
            tw[(size_t)col * t1->last_radix + q] = make_float2((float)cos(ang), (float)sin(ang));
        }
    p->tw_inter_hi = upload(tw);
    tw.resize(m2);
    for (uint32_t k = 0; k < m2; ++k) {
        const double ang = -2.0 * M_PI * (double)k / (double)n;
        tw[k] = make_float2((float)cos(ang), (float)sin(ang));
    }
    p->tw_untangle_a = upload(tw);
    tw.resize(m1);
    for (uint32_t k = 0; k < m1; ++k) {
        const double ang = -2.0 * M_PI * (double)k / (2.0 * (double)m1);
        tw[k] = make_float2((float)cos(ang), (float)sin(ang));
    }
    p->tw_untangle_b = upload(tw);
    p->parts = (m2 % 2048 == 0 && m2 / 2048 >= 1 && m2 / 2048 <= 3) ? (int)(m2 / 2048) : 0;
    if (p->parts > 1) {
        tw.resize((size_t)(p->parts - 1) * 2048);
        for (int q = 1; q < p->parts; ++q)
            for (uint32_t pos = 0; pos < 2048; ++pos) {
                const double ang = -2.0 * M_PI * (double)((unsigned long long)q * pos % m2) / (double)m2;
                tw[(size_t)(q - 1) * 2048 + pos] = make_float2((float)cos(ang), (float)sin(ang));
            }
        std::vector<float2> tile2048; // in-tile twiddles of the 2048-point stages, appended behind the W^{qp} rows
        gen_tile_tw<2048>(tile2048);
        tw.insert(tw.end(), tile2048.begin(), tile2048.end());
        p->tw_parts = upload(tw);
        tw.resize((size_t)m1 * 8);
        for (uint32_t col = 0; col < m1; ++col)
            for (int q = 0; q < 8; ++q) {
                const double ang = -2.0 * M_PI * (double)((unsigned long long)col * q) / ((double)m1 * 8.0);
                tw[(size_t)col * 8 + q] = make_float2((float)cos(ang), (float)sin(ang));
            }
        p->tw_inter_parts = upload(tw);
        if (!p->tw_parts || !p->tw_inter_parts) p->parts = 0; // fall back to the plain kernel
    }
    if (!p->tw_pass1 || !p->tw_pass2 || !p->tw_untangle_a || !p->tw_untangle_b || !p->tw_inter_hi) {
        brp_fft_plan_destroy(p);
        return RADPUL_CUDA_MEM_ALLOC_DEVICE;
    }
    // (measured: forcing the maximum shared-memory carve-out slows both passes by 15-35 %: the twiddle tables and the
    //  short global runs live off the L1 that is left over, so the driver's default split is kept)
    if (cudaFuncSetAttribute((const void *)t1->p1, cudaFuncAttributeMaxDynamicSharedMemorySize, t1->smem1) != cudaSuccess ||
        cudaFuncSetAttribute((const void *)t2->p2, cudaFuncAttributeMaxDynamicSharedMemorySize, t2->smem2) != cudaSuccess ||
        cudaFuncSetAttribute((const void *)t2->p2_spectrum, cudaFuncAttributeMaxDynamicSharedMemorySize, t2->smem2) != cudaSuccess ||
        cudaFuncSetAttribute((const void *)t2->p2_conj, cudaFuncAttributeMaxDynamicSharedMemorySize, t2->smem2) != cudaSuccess) {
        brp_fft_plan_destroy(p);
        return RADPUL_CUDA_FFT_PLAN;
    }
    return 0;
}

void brp_fft_plan_destroy(BrpFftPlan *p)
{
    cudaFree(p->tw_pass1);
    cudaFree(p->tw_pass2);
    cudaFree(p->tw_untangle_a);
    cudaFree(p->tw_untangle_b);
    cudaFree(p->tw_inter_hi);
    cudaFree(p->tw_inter_lo);
    cudaFree(p->tw_parts);
    cudaFree(p->tw_inter_parts);
    memset(p, 0, sizeof(*p));
}

// ---- host side of the TMA pass 1: tensor maps over the input series and the intermediate array
typedef CUresult (*tensor_map_encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                         const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                         CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static tensor_map_encode_fn tensor_map_encoder()
{
    // looked up once (a function-local static is initialised exactly once even when several host threads, one per device,
    // enqueue their first batch at the same time)
    static const tensor_map_encode_fn fn = []() -> tensor_map_encode_fn {
        const char *off = getenv("BRP_FFT_TMA"); // BRP_FFT_TMA=0 selects the plain kernel (A/B measurements)
        if (off && off[0] == '0') return nullptr;
        void *sym = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            return (tensor_map_encode_fn)sym;
        return nullptr;
    }();
    return fn;
}

// [batch][m2 rows][2*m1 floats] with the given batch stride (in floats); box = 256 rows x 8 floats
static bool encode_tile_map(tensor_map_encode_fn enc, CUtensorMap *map, const void *base, uint32_t m1, uint32_t m2, size_t batch_stride,
                            int batch)
{
    if (((uintptr_t)base & 15) || (batch > 1 && (batch_stride & 3))) return false;
    const cuuint64_t dims[3] = {2ull * m1, m2, (cuuint64_t)batch};
    const cuuint64_t strides[2] = {8ull * m1, batch > 1 ? 4ull * batch_stride : 8ull * m1 * m2};
    const cuuint32_t box[3] = {2 * kTmaRows, kTmaBox, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// returns 0 (use the plain kernel) or the number of 2048-row sub-transforms pass 1 runs as (1, 2 or 3)
static int pass1_tma_maps(const BrpFftPlan &p, const float *in, size_t in_stride, const float2 *work, int batch, uint32_t n_steps_bound,
                          CUtensorMap *in_map, CUtensorMap *out_map)
{
    if (p.m1 % kTmaRows != 0 || p.parts < 1) return 0;
    // more than one part: only valid when no sample can sit beyond row 2048 (a row is 2*m1 samples)
    if (p.parts > 1 && (n_steps_bound == 0 || n_steps_bound > 2u * p.m1 * (uint32_t)kTmaL)) return 0;
    // the kernel fetches whole boxes of kTmaBox rows up to the one that holds sample n_steps - 1: unless the samples a
    // template can have fill a whole number of boxes, the last box would read past the end of the caller's buffer
    if ((n_steps_bound ? n_steps_bound : p.n) % ((uint32_t)kTmaBox * 2u * p.m1) != 0) return 0;
    tensor_map_encode_fn enc = tensor_map_encoder();
    if (!enc) return 0;
    static bool configured_on[64]; // function attributes are per device
    int dev = 0;
    cudaGetDevice(&dev);
    bool &configured = configured_on[(dev >= 0 && dev < 64) ? dev : 0];
    if (!configured) {
        const int bytes = Layout<kTmaL, kTmaRows, 3>::BYTES;
        if (cudaFuncSetAttribute((const void *)fft_pass1_tma_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess ||
            cudaFuncSetAttribute((const void *)fft_pass1_tma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess ||
            cudaFuncSetAttribute((const void *)fft_pass1_tma_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess)
            return 0;
        configured = true;
    }
    if (!encode_tile_map(enc, in_map, in, p.m1, p.m2, in_stride, batch)) return 0;
    if (p.parts == 1) return encode_tile_map(enc, out_map, work, p.m1, p.m2, 2 * (size_t)p.m, batch) ? 1 : 0;
    // rows parts*k' + q of the intermediate as [batch][k'][q][2*m1 floats]
    if ((uintptr_t)work & 15) return 0;
    const cuuint64_t dims[4] = {2ull * p.m1, (cuuint64_t)p.parts, (cuuint64_t)kTmaL, (cuuint64_t)batch};
    const cuuint64_t strides[3] = {8ull * p.m1, 8ull * p.m1 * p.parts, 8ull * p.m};
    const cuuint32_t box[4] = {2 * kTmaRows, 1, kTmaBox, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    return enc(out_map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float2 *>(work), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS
               ? p.parts
               : 0;
}

// Power spectrum as [batch][rows k1][m2 floats]; only the rows that hold a bin below `count` exist in the map, so the
// tensor stores of pass 2 clip everything else.  The last of those rows is written whole: the caller's buffer must hold
// m2 * ceil(count / m2) floats per template (out_stride says how many it holds).
static bool pass2_tma_map(const BrpFftPlan &p, const float *out, size_t out_stride, uint32_t count, int batch, CUtensorMap *map)
{
    // (count = m + 1, the whole spectrum with its Nyquist bin as the reference's run_fft wants it: bin m is the mirror of bin 0 and
    //  is written by the edge tile's own stores, all m1 rows go through the map)
    if (p.m1 != 1024 || p.m2 < 16 || count == 0 || count > p.m + 1) return false;
    const uint32_t rows = ((count < p.m ? count : p.m) + p.m2 - 1) / p.m2;
    if ((size_t)rows * p.m2 > out_stride || (out_stride & 3) || ((uintptr_t)out & 15)) return false;
    tensor_map_encode_fn enc = tensor_map_encoder();
    if (!enc) return false;
    static bool configured_on[64]; // function attributes are per device
    int dev = 0;
    cudaGetDevice(&dev);
    bool &configured = configured_on[(dev >= 0 && dev < 64) ? dev : 0];
    if (!configured) {
        if (cudaFuncSetAttribute((const void *)fft_pass2_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 Layout<1024, 8, 2>::BYTES) != cudaSuccess)
            return false;
        configured = true;
    }
    const cuuint64_t dims[3] = {p.m2, rows, (cuuint64_t)batch};
    const cuuint64_t strides[2] = {4ull * p.m2, 4ull * out_stride};
    const cuuint32_t box[3] = {4, 256, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float *>(out), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static cudaError_t launch_fft(const BrpFftPlan &p, int mode, const float *in, size_t in_stride, const uint32_t *n_steps,
                              const float *mean, float2 *work, float *out, size_t out_stride, uint32_t out_count, float norm,
                              int batch, cudaStream_t st, cudaEvent_t mid_event, uint32_t n_steps_bound = 0)
{
    const TileInfo *t1 = find_tile(p.m2), *t2 = find_tile(p.m1);
    const dim3 g1((p.m1 + t1->rows - 1) / t1->rows, batch);
    static_assert(sizeof(CUtensorMap) == 128, "BrpFftMapCache stores tensor maps as 128 bytes");
    CUtensorMap in_map, out_map;
    BrpFftMapCache &mc = p.maps;
    int parts;
    if (mc.valid1 && mc.in == in && mc.in_stride == in_stride && mc.work == work && mc.batch1 == batch && mc.bound == n_steps_bound) {
        parts = mc.parts;
        memcpy(&in_map, mc.in_map, 128);
        memcpy(&out_map, mc.out_map, 128);
    } else {
        parts = pass1_tma_maps(p, in, in_stride, work, batch, n_steps_bound, &in_map, &out_map);
        mc.in = in; mc.in_stride = in_stride; mc.work = work; mc.batch1 = batch; mc.bound = n_steps_bound; mc.parts = parts;
        memcpy(mc.in_map, &in_map, 128);
        memcpy(mc.out_map, &out_map, 128);
        mc.valid1 = true;
    }
    const int tma_smem = Layout<kTmaL, kTmaRows, 3>::BYTES;
    const dim3 gp(p.m1 / kTmaRows, batch, parts > 0 ? parts : 1);
    // persistent form (one CTA of 1024 threads per SM, three tile buffers): BRP_FFT_PERSIST=0 selects the one-tile-per-CTA kernels
    static const bool persist_off = BRP_FFT_P1_PERSIST == 0 || (getenv("BRP_FFT_PERSIST") && getenv("BRP_FFT_PERSIST")[0] == '0');
    bool launched_persistent = false;
    if (parts >= 1 && !persist_off) {
        static int sms_on[64]; // resident CTAs = SMs, per device; 0 = not configured yet, -1 = not usable
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev < 0 || dev >= 64) dev = 0;
        const int psmem = 3 * tma_smem;
        if (sms_on[dev] == 0) {
            int sms = 0;
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            const bool ok = sms > 0 &&
                            cudaFuncSetAttribute((const void *)fft_pass1_persist_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, psmem) == cudaSuccess &&
                            cudaFuncSetAttribute((const void *)fft_pass1_persist_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, psmem) == cudaSuccess &&
                            cudaFuncSetAttribute((const void *)fft_pass1_persist_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, psmem) == cudaSuccess;
            sms_on[dev] = ok ? sms : -1;
            cudaGetLastError();
        }
        if (sms_on[dev] > 0) {
            const uint32_t tiles_x = p.m1 / kTmaRows, total = tiles_x * (uint32_t)batch * (uint32_t)parts;
            const uint32_t grid = total < (uint32_t)sms_on[dev] ? total : (uint32_t)sms_on[dev];
            if (parts == 1)
                fft_pass1_persist_kernel<1><<<grid, 1024, psmem, st>>>(in_map, out_map, n_steps, mean, p.m1, p.m, p.tw_pass1, p.tw_inter_hi, nullptr,
                                                                     tiles_x, total);
            else if (parts == 2)
                fft_pass1_persist_kernel<2><<<grid, 1024, psmem, st>>>(in_map, out_map, n_steps, mean, p.m1, p.m, p.tw_parts + 1 * kTmaL,
                                                                     p.tw_inter_parts, p.tw_parts, tiles_x, total);
            else
                fft_pass1_persist_kernel<3><<<grid, 1024, psmem, st>>>(in_map, out_map, n_steps, mean, p.m1, p.m, p.tw_parts + 2 * kTmaL,
                                                                     p.tw_inter_parts, p.tw_parts, tiles_x, total);
            launched_persistent = true;
        }
    }
    if (launched_persistent) {
        // done above
    } else if (parts == 1)
        fft_pass1_tma_kernel<1><<<gp, 512, tma_smem, st>>>(in_map, out_map, n_steps, mean, p.m1, p.m, p.tw_pass1, p.tw_inter_hi, nullptr);
    else if (parts == 2) // tw_parts: (parts-1) rows of W^{qp}, then the in-tile twiddles of the 2048-point stages
        fft_pass1_tma_kernel<2><<<gp, 512, tma_smem, st>>>(in_map, out_map, n_steps, mean, p.m1, p.m, p.tw_parts + 1 * kTmaL,
                                                         p.tw_inter_parts, p.tw_parts);
    else if (parts == 3)
        fft_pass1_tma_kernel<3><<<gp, 512, tma_smem, st>>>(in_map, out_map, n_steps, mean, p.m1, p.m, p.tw_parts + 2 * kTmaL,
                                                         p.tw_inter_parts, p.tw_parts);
    else
        t1->p1<<<g1, t1->threads, t1->smem1, st>>>(in, in_stride, n_steps, mean, work, p.m1, p.m, p.tw_pass1, p.tw_inter_hi);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    if (mid_event) cudaEventRecord(mid_event, st);
    const int T2 = t2->rows / 2;
    const dim3 g2((p.m2 / 2 + 1 + T2 - 1) / T2, batch); // rows k2 = 0 .. m2/2
    CUtensorMap ps_map;
    bool tma2 = false;
    if (mode == PASS2_POWER) {
        if (mc.valid2 && mc.out == out && mc.out_stride == out_stride && mc.count == out_count && mc.batch2 == batch) {
            tma2 = mc.ok2;
            memcpy(&ps_map, mc.ps_map, 128);
        } else {
            tma2 = pass2_tma_map(p, out, out_stride, out_count, batch, &ps_map);
            mc.out = out; mc.out_stride = out_stride; mc.count = out_count; mc.batch2 = batch; mc.ok2 = tma2;
            memcpy(mc.ps_map, &ps_map, 128);
            mc.valid2 = true;
        }
    }
    if (tma2) {
        fft_pass2_tma_kernel<<<g2, 256, Layout<1024, 8, 2>::BYTES, st>>>(ps_map, work, out, out_stride, out_count, p.m2, p.m, norm,
                                                                       p.tw_pass2, p.tw_untangle_a, p.tw_untangle_b);
        return cudaGetLastError();
    }
    const pass2_fn k2 = mode == PASS2_POWER ? t2->p2 : (mode == PASS2_SPECTRUM ? t2->p2_spectrum : t2->p2_conj);
    k2<<<g2, t2->threads, t2->smem2, st>>>(work, out, out_stride, out_count, p.m2, p.m, norm, p.tw_pass2, p.tw_untangle_a,
                                         p.tw_untangle_b);
    return cudaGetLastError();
}

cudaError_t brp_launch_fft_power(const BrpFftPlan &p, const float *in, size_t in_stride, const uint32_t *n_steps,
                                 const float *mean, float2 *work, float *ps, size_t ps_stride, uint32_t ps_count,
                                 float norm, int batch, cudaStream_t st, cudaEvent_t mid_event, uint32_t n_steps_bound)
{
    return launch_fft(p, PASS2_POWER, in, in_stride, n_steps, mean, work, ps, ps_stride, ps_count, norm, batch, st, mid_event,
                      n_steps_bound);
}

// X[0..count) of the real-input transform of `in` (n reals; samples >= *n_steps read as *mean), count <= n/2 + 1
cudaError_t brp_launch_fft_spectrum(const BrpFftPlan &p, const float *in, const uint32_t *n_steps, const float *mean,
                                    float2 *work, float2 *spectrum, uint32_t count, cudaStream_t st)
{
    return launch_fft(p, PASS2_SPECTRUM, in, 0, n_steps, mean, work, reinterpret_cast<float *>(spectrum), 0, count, 1.0f, 1, st, NULL);
}

// out[2k] = scale * Re Z[k], out[2k+1] = -scale * Im Z[k] for k < pairs, Z = complex transform of the n/2 complex points in `in`
cudaError_t brp_launch_fft_conj(const BrpFftPlan &p, const float2 *in, const uint32_t *n_steps_full, const float *mean,
                                float2 *work, float *out, uint32_t pairs, float scale, cudaStream_t st)
{
    return launch_fft(p, PASS2_CONJ, reinterpret_cast<const float *>(in), 0, n_steps_full, mean, work, out, 0, pairs, scale, 1, st, NULL);
}
