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
template <typename Table>
__device__ __forceinline__ float brp_sin_lut(float x, const Table &tab)
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
    return __fsub_rn(__fadd_rn(sc.x, __fmul_rn(d, sc.y)), __fmul_rn(d2, sc.x));
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

#endif
