// Device-side restatement of the reference's timing model, bit-exact with its CPU build.
//
//   sincosLUTLookup   erp_utilities.cpp:176-209   (tables :45-46, constants erp_utilities.h:27-32)
//   del_t[i]          demod_binary_resamp_cpu.c:91-102
//   nearest index     demod_binary_resamp_cpu.c:116
//
// The reference CPU build rounds every operation separately (SSE, no FMA; Debian forces
// -ffp-contract=off for this file), so every floating-point operation below is an explicit
// round-to-nearest intrinsic.  An FMA is used only where the product is exact (multiplication by a
// power of two), in which case fused and unfused results are identical.  Conversions (F2I, I2F, FRND,
// F2F.F64) run at a quarter of the FP32 rate on sm_100, so they are replaced by exact add-magic
// sequences; tests/test_gpu_timing_model.py checks the results bit for bit against the oracle.
#ifndef BRP_DEVICE_CUH
#define BRP_DEVICE_CUH

#include <cuda_runtime.h>
#include <stdint.h>

#define BRP_TWO_PI 6.283185f                 // ERP_TWO_PI (erp_utilities.h:31), NOT 2*pi rounded
#define BRP_TWO_PI_INV (1.0f / 6.283185f)    // ERP_TWO_PI_INV, folded in IEEE single like the host compiler does

// {sin, cos} sample pairs: the 6-decimal literals of erp_utilities.cpp:45-46, copied as data
// (they are not exact sines: e.g. sin[35] = -0.290284 but sin[29] = 0.290285).
static __constant__ float2 c_brp_sincos[65] = {
    {0.000000f, 1.000000f},   {0.098017f, 0.995185f},   {0.195090f, 0.980785f},   {0.290285f, 0.956940f},
    {0.382683f, 0.923880f},   {0.471397f, 0.881921f},   {0.555570f, 0.831470f},   {0.634393f, 0.773010f},
    {0.707107f, 0.707107f},   {0.773010f, 0.634393f},   {0.831470f, 0.555570f},   {0.881921f, 0.471397f},
    {0.923880f, 0.382683f},   {0.956940f, 0.290285f},   {0.980785f, 0.195090f},   {0.995185f, 0.098017f},
    {1.000000f, 0.000000f},   {0.995185f, -0.098017f},  {0.980785f, -0.195090f},  {0.956940f, -0.290285f},
    {0.923880f, -0.382683f},  {0.881921f, -0.471397f},  {0.831470f, -0.555570f},  {0.773010f, -0.634393f},
    {0.707107f, -0.707107f},  {0.634393f, -0.773010f},  {0.555570f, -0.831470f},  {0.471397f, -0.881921f},
    {0.382683f, -0.923880f},  {0.290285f, -0.956940f},  {0.195091f, -0.980785f},  {0.098017f, -0.995185f},
    {0.000000f, -1.000000f},  {-0.098017f, -0.995185f}, {-0.195090f, -0.980785f}, {-0.290284f, -0.956940f},
    {-0.382683f, -0.923880f}, {-0.471397f, -0.881921f}, {-0.555570f, -0.831470f}, {-0.634393f, -0.773011f},
    {-0.707107f, -0.707107f}, {-0.773010f, -0.634393f}, {-0.831469f, -0.555570f}, {-0.881921f, -0.471397f},
    {-0.923880f, -0.382684f}, {-0.956940f, -0.290285f}, {-0.980785f, -0.195090f}, {-0.995185f, -0.098017f},
    {-1.000000f, 0.000000f},  {-0.995185f, 0.098017f},  {-0.980785f, 0.195090f},  {-0.956940f, 0.290285f},
    {-0.923880f, 0.382683f},  {-0.881921f, 0.471397f},  {-0.831470f, 0.555570f},  {-0.773011f, 0.634393f},
    {-0.707107f, 0.707107f},  {-0.634394f, 0.773010f},  {-0.555570f, 0.831470f},  {-0.471397f, 0.881921f},
    {-0.382684f, 0.923879f},  {-0.290285f, 0.956940f},  {-0.195091f, 0.980785f},  {-0.098017f, 0.995185f},
    {-0.000000f, 1.000000f}};

// sin(x) by table + second-order Taylor step, erp_utilities.cpp:184-205:
//   xt = modff(x/2pi); if (xt < 0) xt += 1;  i0 = (int)(xt*64 + 0.5f);  d = 2pi*(xt - i0/64);  d2 = d*(0.5f*d)
//   sin = ts + d*tc - d2*ts
// `cell` receives the table index i0 (0..64) the value was expanded around.
template <typename Table>
__device__ __forceinline__ float brp_sin_lut_cell(float x, const Table &tab, int &cell)
{
    const float y = __fmul_rn(BRP_TWO_PI_INV, x);
    float fr;
    if (fabsf(y) < 4194304.0f) {
        // y - rint(y), exact.  With xt = fr < 0 ? fr + 1 : fr this equals the reference's modff + wrap for
        // either sign of y (both are the exact fractional part folded into [0,1); the only rounding is the
        // reference's own "+ 1.0f" on a negative fraction, which is reproduced).
        const float r = __fsub_rn(__fadd_rn(y, 12582912.0f), 12582912.0f);
        fr = __fsub_rn(y, r);
    } else {
        fr = __fsub_rn(y, truncf(y)); // modff for huge arguments (never reached by a physical template)
    }
    const float xt = (fr < 0.0f) ? __fadd_rn(fr, 1.0f) : fr;
    // (int)(xt*64 + 0.5f): the product is exact, so one FMA rounds like the reference's mul + add;
    // floor of a value in [0.5, 64.5] by a round-down add of 2^23 (mantissa then holds the integer)
    const float z = __fmaf_rn(xt, 64.0f, 0.5f);
    const float zf = __fadd_rd(z, 8388608.0f);
    const int i0 = __float_as_int(zf) & 0xFF;
    const float i0f = __fsub_rn(zf, 8388608.0f);
    // d = 2pi * (xt - (1/64)*i0): (1/64)*i0 is exact, so the FMA equals the reference's separate ops
    const float d = __fmul_rn(BRP_TWO_PI, __fmaf_rn(i0f, -0.015625f, xt));
    const float d2 = __fmul_rn(d, __fmul_rn(0.5f, d));
    const float2 sc = tab[i0];
    cell = i0;
    return __fsub_rn(__fadd_rn(sc.x, __fmul_rn(d, sc.y)), __fmul_rn(d2, sc.x));
}

template <typename Table>
__device__ __forceinline__ float brp_sin_lut(float x, const Table &tab)
{
    int cell;
    return brp_sin_lut_cell(x, tab, cell);
}

// del_t for sample index i_f (a float that holds an exact integer), demod_binary_resamp_cpu.c:93-101
template <typename Table>
__device__ __forceinline__ float brp_del_t(float i_f, float tau, float Omega, float Psi0, float S0, float dt,
                                           float step_inv, const Table &tab)
{
    const float t = __fmul_rn(i_f, dt);
    const float s = brp_sin_lut(__fadd_rn(__fmul_rn(Omega, t), Psi0), tab);
    return __fsub_rn(__fmul_rn(__fmul_rn(tau, s), step_inv), S0);
}

// same, also reporting the table cell (used by the run detection of the resampler)
template <typename Table>
__device__ __forceinline__ float brp_del_t_cell(float i_f, float tau, float Omega, float Psi0, float S0, float dt,
                                                float step_inv, const Table &tab, int &cell)
{
    const float t = __fmul_rn(i_f, dt);
    const float s = brp_sin_lut_cell(__fadd_rn(__fmul_rn(Omega, t), Psi0), tab, cell);
    return __fsub_rn(__fmul_rn(__fmul_rn(tau, s), step_inv), S0);
}

// (int)((double)f + 0.5) for f < 2^23, without leaving FP32 (demod_binary_resamp_cpu.c:116).
// For f >= 0 this is floor(f) + (frac >= 0.5); arguments in (-1.5, 0) give 0 like the C truncation
// (more negative ones would index before the array in the reference and are clamped to 0 here).
__device__ __forceinline__ int brp_round_index_fast(float f)
{
    const float fp = fmaxf(f, 0.0f);
    const float s = __fadd_rd(fp, 8388608.0f);  // 2^23 + floor(fp), exact
    const int ip = __float_as_int(s) & 0x7FFFFF;
    const float fl = __fsub_rn(s, 8388608.0f);
    const float frac = __fsub_rn(fp, fl);       // exact
    return ip + (frac >= 0.5f ? 1 : 0);
}

// general form (any magnitude), the literal arithmetic of the reference
__device__ __forceinline__ int brp_round_index_exact(float f)
{
    const int v = __double2int_rz(__dadd_rn((double)f, 0.5));
    return v < 0 ? 0 : v;
}

// ---- two samples per instruction: Blackwell packed FP32x2 arithmetic (FADD2 / FMUL2 / FFMA2, every rounding mode).
// Each lane of a packed operation rounds exactly like the scalar intrinsic above, so the pair version is bit-identical
// to two scalar evaluations; a - b is written as fma(b, -1, a), which is one exact product and one rounding.
// Only valid while |x / 2pi| < 2^22 for both lanes (the caller checks the template's largest argument once).
// The CUDA builtins __fmul2_rn/__fadd2_rn are NOT safe here: nvcc contracts a packed multiply followed by a packed add
// into FFMA2 even with -fmad=false (seen in SASS; ~4 % of the resampling indices changed).  PTX instructions that carry
// an explicit rounding modifier are never contracted by ptxas, so the packed operations are spelled in PTX.
#define BRP_PK2(name, op)                                                                                              \
    __device__ __forceinline__ float2 name(float2 a, float2 b)                                                         \
    {                                                                                                                  \
        float2 r;                                                                                                      \
        asm("{\n .reg .b64 ra, rb, rc;\n mov.b64 ra, {%2, %3};\n mov.b64 rb, {%4, %5};\n " op " rc, ra, rb;\n"         \
            " mov.b64 {%0, %1}, rc;\n}"                                                                                \
            : "=f"(r.x), "=f"(r.y)                                                                                     \
            : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));                                                                 \
        return r;                                                                                                      \
    }
// ptxas nevertheless contracts "mul.rn.f32x2 ; add.rn.f32x2" into FFMA2 (CUDA 12.9, also with -fmad=false), which the
// scalar forms never do.  Hence brp_add2_of_product(): an addition whose operand is a fresh packed product is done with
// two scalar __fadd_rn, which are protected; everything else stays packed.  The index arrays of
// tests/test_gpu_resample.py are the check: bit-exact against the oracle.
BRP_PK2(brp_mul2, "mul.rn.f32x2")
BRP_PK2(brp_add2, "add.rn.f32x2")
BRP_PK2(brp_add2_rd, "add.rm.f32x2")
#undef BRP_PK2
__device__ __forceinline__ float2 brp_fma2(float2 a, float2 b, float2 c)
{
    float2 r;
    asm("{\n .reg .b64 ra, rb, rc, rd;\n mov.b64 ra, {%2, %3};\n mov.b64 rb, {%4, %5};\n mov.b64 rc, {%6, %7};\n"
        " fma.rn.f32x2 rd, ra, rb, rc;\n mov.b64 {%0, %1}, rd;\n}"
        : "=f"(r.x), "=f"(r.y)
        : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
    return r;
}
__device__ __forceinline__ float2 brp_add2_of_product(float2 prod, float2 c)
{
    return make_float2(__fadd_rn(prod.x, c.x), __fadd_rn(prod.y, c.y));
}
__device__ __forceinline__ float2 brp_pair(float v) { return make_float2(v, v); }
__device__ __forceinline__ float2 brp_sub2(float2 a, float2 b) { return brp_fma2(b, make_float2(-1.0f, -1.0f), a); }

template <typename Table>
__device__ __forceinline__ float2 brp_sin_lut2(float2 x, const Table &tab)
{
    const float2 y = brp_mul2(brp_pair(BRP_TWO_PI_INV), x);
    const float2 r = brp_add2(brp_add2(y, brp_pair(12582912.0f)), brp_pair(-12582912.0f));
    const float2 fr = brp_sub2(y, r);
    const float2 wrap = make_float2(fr.x < 0.0f ? 1.0f : 0.0f, fr.y < 0.0f ? 1.0f : 0.0f); // fr + 0 is exact
    const float2 xt = brp_add2(fr, wrap);
    const float2 z = brp_fma2(xt, brp_pair(64.0f), brp_pair(0.5f));
    const float2 zf = brp_add2_rd(z, brp_pair(8388608.0f));
    const int i0 = __float_as_int(zf.x) & 0xFF, i1 = __float_as_int(zf.y) & 0xFF;
    const float2 i0f = brp_add2(zf, brp_pair(-8388608.0f));
    const float2 d = brp_mul2(brp_pair(BRP_TWO_PI), brp_fma2(i0f, brp_pair(-0.015625f), xt));
    const float2 d2 = brp_mul2(d, brp_mul2(brp_pair(0.5f), d));
    const float2 sc0 = tab[i0], sc1 = tab[i1];
    const float2 ts = make_float2(sc0.x, sc1.x), tc = make_float2(sc0.y, sc1.y);
    return brp_sub2(brp_add2_of_product(brp_mul2(d, tc), ts), brp_mul2(d2, ts));
}

template <typename Table>
__device__ __forceinline__ float2 brp_del_t2(float2 i_f, float tau, float Omega, float Psi0, float S0, float dt,
                                             float step_inv, const Table &tab)
{
    const float2 t = brp_mul2(i_f, brp_pair(dt));
    const float2 s = brp_sin_lut2(brp_add2_of_product(brp_mul2(brp_pair(Omega), t), brp_pair(Psi0)), tab);
    return brp_sub2(brp_mul2(brp_mul2(brp_pair(tau), s), brp_pair(step_inv)), brp_pair(S0));
}

// pair version of brp_round_index_fast
__device__ __forceinline__ int2 brp_round_index_fast2(float2 f)
{
    const float2 fp = make_float2(fmaxf(f.x, 0.0f), fmaxf(f.y, 0.0f));
    const float2 s = brp_add2_rd(fp, brp_pair(8388608.0f));
    const float2 fl = brp_add2(s, brp_pair(-8388608.0f));
    const float2 frac = brp_sub2(fp, fl);
    return make_int2((__float_as_int(s.x) & 0x7FFFFF) + (frac.x >= 0.5f ? 1 : 0),
                     (__float_as_int(s.y) & 0x7FFFFF) + (frac.y >= 0.5f ? 1 : 0));
}

#endif
