// Probe: how fast can an FFT pass-1 tile (2048 rows x 4 complex columns = 2048 runs of 32 bytes, 8 KB apart) be GATHERED from the
// L2-resident input series with a per-row displacement (the resampling shift), i.e. what does it cost to fold the resampler
// into pass 1's load?  The displaced runs are not 16-byte aligned, so tensor copies cannot fetch them.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/gather_probe tools/gather_probe.cu && tools/bin/gather_probe
//
// Methods (512 threads, 2 CTAs per SM, persistent over all tiles of a 16-template volume):
//   0  cp.async 4 bytes (LDGSTS.32), 8 lanes per row: exact run, dense [row][8 floats] image
//   1  cp.async 16 bytes (LDGSTS.128, .ca), 3 lanes per row: aligned 48-byte window that contains the run
//   2  same with .cg (L2 only)
//   3  LDG.32 8 lanes per row -> registers -> STS.32
//   4  cp.async.bulk 48 bytes per row (one bulk copy per row, mbarrier completion)
//   5  no gather at all (loop + barriers only): the floor of the harness
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x)                                                                       \
    do {                                                                            \
        cudaError_t e_ = (x);                                                       \
        if (e_ != cudaSuccess) {                                                    \
            printf("%s: %s (line %d)\n", #x, cudaGetErrorString(e_), __LINE__);      \
            exit(1);                                                                \
        }                                                                           \
    } while (0)

constexpr int ROWS = 2048, ROWLEN = 2048 /* samples per row */, NU = ROWS * ROWLEN, BATCH = 16, TILES = 256;

__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int METHOD>
__global__ void __launch_bounds__(512, 2) probe(const float *__restrict__ series, const int *__restrict__ shift /* [BATCH][ROWS] */,
                                                int n_tiles, float *sink)
{
    extern __shared__ __align__(128) float sm[]; // 2048 rows x 12 floats
    __shared__ __align__(8) unsigned long long bar;
    const uint32_t bar_a = s32(&bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    __syncthreads();
    uint32_t phase = 0;
    float acc = 0.f;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int b = tile / TILES, tx = tile - b * TILES;
        const int *sh = shift + b * ROWS;
        if (METHOD == 0 || METHOD == 3) {
            const int rl = lane >> 3, e = lane & 7;
#pragma unroll 8
            for (int it = 0; it < 32; ++it) {
                const int row = warp * 128 + it * 4 + rl;
                int k = row * ROWLEN + 8 * tx + e - __ldg(sh + row);
                k = min(max(k, 0), NU - 1);
                if (METHOD == 0)
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s32(sm + row * 8 + e)), "l"(series + k) : "memory");
                else
                    sm[row * 8 + e] = __ldg(series + k);
            }
            if (METHOD == 0) asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
        } else if (METHOD == 1 || METHOD == 2) {
            // 30 lanes: 10 rows x 3 pieces of 16 bytes
            const int rl = lane / 3, piece = lane - 3 * rl;
            for (int it = 0; it < 13; ++it) {
                const int row = warp * 128 + it * 10 + rl;
                if (lane < 30 && row < warp * 128 + 128) {
                    int k = row * ROWLEN + 8 * tx - __ldg(sh + row);
                    k = min(max(k, 0), NU - 16) & ~3;
                    if (METHOD == 1)
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(s32(sm + row * 12 + 4 * piece)), "l"(series + k + 4 * piece) : "memory");
                    else
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s32(sm + row * 12 + 4 * piece)), "l"(series + k + 4 * piece) : "memory");
                }
            }
            asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
        } else if (METHOD == 4) {
            if (threadIdx.x == 0)
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(ROWS * 48) : "memory");
            __syncthreads();
#pragma unroll
            for (int it = 0; it < 4; ++it) {
                const int row = threadIdx.x + 512 * it;
                int k = row * ROWLEN + 8 * tx - __ldg(sh + row);
                k = min(max(k, 0), NU - 16) & ~3;
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 48, [%2];" ::"r"(s32(sm + row * 12)),
                             "l"(series + k), "r"(bar_a)
                             : "memory");
            }
            asm volatile("{\n.reg .pred p;\nWAIT:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE;\nbra WAIT;\nDONE:\n}" ::"r"(bar_a),
                         "r"(phase)
                         : "memory");
            phase ^= 1;
        }
        __syncthreads();
        acc += sm[threadIdx.x * 12] + sm[threadIdx.x * 12 + 7]; // touch
        __syncthreads();
    }
    if (acc == 123.456f) sink[0] = acc;
}

template <int METHOD>
static void run(const float *series, const int *shift, float *sink, const char *what)
{
    const int smem = ROWS * 12 * 4;
    CK(cudaFuncSetAttribute(probe<METHOD>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int n_tiles = TILES * BATCH, grid = 148 * 2;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int i = 0; i < 3; i++) probe<METHOD><<<grid, 512, smem>>>(series, shift, n_tiles, sink);
    CK(cudaDeviceSynchronize());
    const int reps = 20;
    CK(cudaEventRecord(e0));
    for (int i = 0; i < reps; i++) probe<METHOD><<<grid, 512, smem>>>(series, shift, n_tiles, sink);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double us = 1e3 * ms / reps;
    printf("method %d  %-58s %7.1f us per 16-template volume = %5.2f us per template, %6.0f GB/s of useful bytes\n", METHOD, what, us, us / BATCH,
           (double)NU * 4 * BATCH / us * 1e-3);
}

int main()
{
    float *series, *sink;
    int *shift;
    CK(cudaMalloc(&series, sizeof(float) * NU));
    CK(cudaMalloc(&sink, 4));
    CK(cudaMalloc(&shift, sizeof(int) * BATCH * ROWS));
    CK(cudaMemset(series, 0, sizeof(float) * NU));
    int *h = (int *)malloc(sizeof(int) * BATCH * ROWS);
    for (int b = 0; b < BATCH; b++)
        for (int r = 0; r < ROWS; r++) h[b * ROWS + r] = (int)(4000.0 * __builtin_sin(0.004 * r + 0.7 * b)) ; // like a 0.26 s orbit amplitude
    CK(cudaMemcpy(shift, h, sizeof(int) * BATCH * ROWS, cudaMemcpyHostToDevice));
    run<5>(series, shift, sink, "no gather (harness floor)");
    run<0>(series, shift, sink, "cp.async 4 B, 8 lanes per row");
    run<1>(series, shift, sink, "cp.async.ca 16 B, 48-byte aligned window, 3 lanes per row");
    run<2>(series, shift, sink, "cp.async.cg 16 B, 48-byte aligned window, 3 lanes per row");
    run<3>(series, shift, sink, "LDG.32 -> STS.32, 8 lanes per row");
    run<4>(series, shift, sink, "cp.async.bulk 48 B per row");
    return 0;
}
