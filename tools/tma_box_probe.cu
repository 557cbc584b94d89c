// Probe: how fast can TMA *tensor* copies move an FFT pass-1 tile (2048 rows x W complex columns, rows 8 KB apart)
// between global and shared memory?  Decides whether pass 1 should stage its tiles with cp.async.bulk.tensor instead
// of per-lane global loads/stores (which are bound by L1 wavefronts: one per 128-byte line touched).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tma_box_probe tools/tma_box_probe.cu && /tmp/tma_box_probe
//
// For W in {2, 4, 8} columns: persistent CTAs copy every tile of a 16-template volume A -> smem -> B.
#include <cuda.h>
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

constexpr int ROWS = 2048;   // pass-1 FFT length
constexpr int COLS = 1024;   // complex elements per row (row pitch 8 KB)
constexpr int BATCH = 16;
constexpr int BOX_ROWS = 256;

__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int W, bool STORE>
__global__ void __launch_bounds__(256) probe(const __grid_constant__ CUtensorMap src, const __grid_constant__ CUtensorMap dst,
                                             int tiles_per_template, int n_tiles, float *sink, int store_shift)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) unsigned long long bar;
    const uint32_t bar_a = s32(&bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    __syncthreads();
    uint32_t phase = 0;
    float acc = 0.f;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int b = tile / tiles_per_template, tx = tile - b * tiles_per_template;
        const int c0 = tx * W * 2;            // inner coordinate in floats
        const int r0 = b * ROWS;              // outer coordinate (rows of all templates stacked)
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(ROWS * W * 8) : "memory");
#pragma unroll
            for (int k = 0; k < ROWS / BOX_ROWS; ++k)
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                                 s32(smem + (size_t)k * BOX_ROWS * W * 8)),
                             "l"(&src), "r"(c0), "r"(r0 + k * BOX_ROWS), "r"(bar_a)
                             : "memory");
        }
        // everybody waits for the tile
        asm volatile("{\n.reg .pred p;\nWAIT:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE;\nbra WAIT;\nDONE:\n}" ::"r"(bar_a),
                     "r"(phase)
                     : "memory");
        phase ^= 1;
        acc += reinterpret_cast<const float *>(smem)[threadIdx.x];  // touch
        if (STORE) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (threadIdx.x == 0) {
#pragma unroll
                for (int k = 0; k < ROWS / BOX_ROWS; ++k)
                    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(&dst), "r"(c0 + store_shift),
                                 "r"(r0 + k * BOX_ROWS), "r"(s32(smem + (size_t)k * BOX_ROWS * W * 8))
                                 : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
        }
        __syncthreads();
    }
    if (acc == 123.456f) sink[0] = acc;
}

typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                             const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static CUtensorMap make_map(EncodeFn enc, void *base, int w)
{
    CUtensorMap m;
    cuuint64_t dims[2] = {(cuuint64_t)COLS * 2, (cuuint64_t)ROWS * BATCH};
    cuuint64_t strides[1] = {(cuuint64_t)COLS * 8};
    cuuint32_t box[2] = {(cuuint32_t)w * 2, BOX_ROWS};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        printf("cuTensorMapEncodeTiled failed: %d\n", (int)r);
        exit(1);
    }
    return m;
}

template <int W, bool STORE>
static void run(EncodeFn enc, float *a, float *b, float *sink, int ctas_per_sm, int store_shift = 0)
{
    CUtensorMap ma = make_map(enc, a, W), mb = make_map(enc, b, W);
    const int smem = ROWS * W * 8;
    CK(cudaFuncSetAttribute(probe<W, STORE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int tiles_per_template = COLS / W, n_tiles = tiles_per_template * BATCH;
    const int grid = 148 * ctas_per_sm;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int i = 0; i < 3; i++) probe<W, STORE><<<grid, 256, smem>>>(ma, mb, tiles_per_template, n_tiles, sink, store_shift);
    CK(cudaDeviceSynchronize());
    const int reps = 20;
    CK(cudaEventRecord(e0));
    for (int i = 0; i < reps; i++) probe<W, STORE><<<grid, 256, smem>>>(ma, mb, tiles_per_template, n_tiles, sink, store_shift);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double us = 1e3 * ms / reps, bytes = (double)ROWS * COLS * 8 * BATCH * (STORE ? 2 : 1);
    printf("W=%2d columns (%3d-byte runs) %s, %d CTA/SM, store shifted by %d floats: %7.1f us per 16-template volume, %6.0f GB/s (%s)\n", W,
           W * 8, STORE ? "load+store" : "load only ", ctas_per_sm, store_shift, us, bytes / us * 1e-3, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    EncodeFn enc = (EncodeFn)fn;
    const size_t bytes = (size_t)ROWS * COLS * 8 * BATCH;
    float *a, *b, *sink;
    CK(cudaMalloc(&a, bytes));
    CK(cudaMalloc(&b, bytes));
    CK(cudaMalloc(&sink, 4));
    CK(cudaMemset(a, 0, bytes));
    run<4, false>(enc, a, b, sink, 2);
    run<4, false>(enc, a, b, sink, 3);
    run<4, true>(enc, a, b, sink, 2);
    run<4, true>(enc, a, b, sink, 3);
    run<2, false>(enc, a, b, sink, 4);
    run<2, true>(enc, a, b, sink, 4);
    run<2, true>(enc, a, b, sink, 4, 1);   // is a store that starts 4 bytes past a 16-byte boundary legal?
    run<8, false>(enc, a, b, sink, 1);
    run<8, true>(enc, a, b, sink, 1);
    return 0;
}
