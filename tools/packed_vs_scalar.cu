// Stand-alone check (B200): packed FP32x2 timing model vs the scalar one, bit for bit, for every sample of one template.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o /tmp/pvs tools/packed_vs_scalar.cu && /tmp/pvs   -> "bad=0"
// It is how the FFMA2 contraction of packed mul+add (see brp_device.cuh) was found.
#include <cstdio>
#include <cmath>
#include "../boinc-app-eah-brp_b200/csrc/brp_device.cuh"
struct CT { __device__ float2 operator[](int i) const { return c_brp_sincos[i]; } };
__global__ void k(float tau, float Om, float Psi, float S0, float dt, float sinv, int n, int *bad) {
  int i = 2*(blockIdx.x*blockDim.x+threadIdx.x); if (i+1>=n) return;
  float2 fi = make_float2((float)i,(float)(i+1));
  float2 d2 = brp_del_t2(fi,tau,Om,Psi,S0,dt,sinv,CT{});
  float a = brp_del_t(fi.x,tau,Om,Psi,S0,dt,sinv,CT{}), b = brp_del_t(fi.y,tau,Om,Psi,S0,dt,sinv,CT{});
  if (__float_as_int(a)!=__float_as_int(d2.x) || __float_as_int(b)!=__float_as_int(d2.y)) {
    int s = atomicAdd(bad,1);
    if (s<5) {
      // intermediate comparison for lane x
      float x = __fadd_rn(__fmul_rn(Om, __fmul_rn(fi.x,dt)),Psi);
      float2 xx = __fadd2_rn(__fmul2_rn(brp_pair(Om), __fmul2_rn(fi, brp_pair(dt))), brp_pair(Psi));
      float s1 = brp_sin_lut(x,CT{}); float2 s2 = brp_sin_lut2(xx,CT{});
      printf("i=%d del %.9g/%.9g %.9g/%.9g  x %.9g/%.9g sin %.9g/%.9g\n", i, a,d2.x,b,d2.y, x, xx.x, s1, s2.x);
    }
  }
}
int main(){ int *bad; cudaMallocManaged(&bad,4); *bad=0;
  double P=1536.21668,tau=0.261931251,psi=0.52613317; float dt=65.47619047619051e-6f; float sinv=1.0/ (double)dt;
  float Pf=P,tf=tau,pf=psi; float Om=2.0*M_PI/Pf; float S0=tf*sinf(pf)*sinv;
  k<<<256,256>>>(tf,Om,pf,S0,dt,sinv,131072,bad); cudaDeviceSynchronize(); printf("bad=%d %s\n",*bad,cudaGetErrorString(cudaGetLastError())); }
