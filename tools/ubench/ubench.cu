// Micro-benchmarks that size the instruction diets of K1/K2 on B200:
// issue rate of FFMA vs FFMA2 (f32x2), DFMA/DADD, DSETP+SEL min, F2F, MUFU,
// and the accuracy of rsqrt.approx.ftz.f64 (MUFU.RSQ64H).
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>
#include <math.h>

#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("%s: %s\n",#x,cudaGetErrorString(e)); return 1;}}while(0)

__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c){
  unsigned long long ra=*reinterpret_cast<unsigned long long*>(&a), rb=*reinterpret_cast<unsigned long long*>(&b), rc=*reinterpret_cast<unsigned long long*>(&c), rd;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
  return *reinterpret_cast<float2*>(&rd);
}
constexpr int ITER = 4096;
// 8 independent chains per thread
__global__ void k_ffma(float *o, float a, float b){
  float x[8]; for(int i=0;i<8;i++) x[i]=threadIdx.x+i;
  for(int it=0; it<ITER; ++it){
#pragma unroll
    for(int i=0;i<8;i++) x[i]=__fmaf_rn(x[i],a,b);
  }
  float s=0; for(int i=0;i<8;i++) s+=x[i]; o[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
__global__ void k_ffma2(float *o, float a, float b){
  float2 x[8]; for(int i=0;i<8;i++) x[i]=make_float2(threadIdx.x+i, i);
  float2 A=make_float2(a,a), B=make_float2(b,b);
  for(int it=0; it<ITER; ++it){
#pragma unroll
    for(int i=0;i<8;i++) x[i]=fma2(x[i],A,B);
  }
  float s=0; for(int i=0;i<8;i++) s+=x[i].x+x[i].y; o[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
__global__ void k_dfma(double *o, double a, double b){
  double x[8]; for(int i=0;i<8;i++) x[i]=threadIdx.x+i;
  for(int it=0; it<ITER; ++it){
#pragma unroll
    for(int i=0;i<8;i++) x[i]=__fma_rn(x[i],a,b);
  }
  double s=0; for(int i=0;i<8;i++) s+=x[i]; o[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
__global__ void k_dadd(double *o, double a){
  double x[8]; for(int i=0;i<8;i++) x[i]=threadIdx.x+i;
  for(int it=0; it<ITER; ++it){
#pragma unroll
    for(int i=0;i<8;i++) x[i]=__dadd_rn(x[i],a);
  }
  double s=0; for(int i=0;i<8;i++) s+=x[i]; o[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
// min via DSETP + 2 SEL against a changing value
__global__ void k_dmin(double *o, double a){
  double x[8], y[8]; for(int i=0;i<8;i++){ x[i]=1e300; y[i]=threadIdx.x*1e3+i; }
  for(int it=0; it<ITER; ++it){
#pragma unroll
    for(int i=0;i<8;i++){ y[i]=__dadd_rn(y[i],a); x[i]=(y[i]<x[i])?y[i]:x[i]; }
  }
  double s=0; for(int i=0;i<8;i++) s+=x[i]; o[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
__global__ void k_f2f(float *o, double a){
  double x[8]; for(int i=0;i<8;i++) x[i]=threadIdx.x+i;
  float acc[8]; for(int i=0;i<8;i++) acc[i]=0;
  for(int it=0; it<ITER; ++it){
#pragma unroll
    for(int i=0;i<8;i++){ float f=(float)x[i]; x[i]=(double)f * a; }
  }
  float s=0; for(int i=0;i<8;i++) s+=(float)x[i]; o[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
__global__ void k_mufu(float *o){
  float x[8]; for(int i=0;i<8;i++) x[i]=threadIdx.x+i+1.5f;
  for(int it=0; it<ITER; ++it){
#pragma unroll
    for(int i=0;i<8;i++){ float r; asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x[i])); x[i]=r+2.0f; }
  }
  float s=0; for(int i=0;i<8;i++) s+=x[i]; o[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
__global__ void k_fsel(float *o, float lo, float hi){
  float x[8]; for(int i=0;i<8;i++) x[i]=threadIdx.x+i+1.5f;
  for(int it=0; it<ITER; ++it){
#pragma unroll
    for(int i=0;i<8;i++){ float v=x[i]*1.0001f; v=(v<lo)?lo:v; v=(v>hi)?hi:v; x[i]=v; }
  }
  float s=0; for(int i=0;i<8;i++) s+=x[i]; o[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
// accuracy of rsqrt.approx.ftz.f64
__global__ void k_rsq_acc(double *maxerr, double *maxerr_newton, uint64_t seed, int n){
  uint64_t s = seed + (blockIdx.x*blockDim.x+threadIdx.x)*0x9E3779B97F4A7C15ull;
  double me=0, mn=0;
  for(int i=0;i<n;i++){
    s ^= s<<13; s ^= s>>7; s ^= s<<17;
    // mantissa random, exponent in [-20, 40]
    int e = (int)((s>>52) % 61) - 20;
    double m = 1.0 + (double)(s & 0xfffffffffffffull) / 4503599627370496.0;
    double x = ldexp(m, e);
    double y; asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double ex = 1.0/sqrt(x);
    double err = fabs(y-ex)/ex; if(err>me) me=err;
    double g = x*y; double ee = fma(-g,g,x); double d = fma(ee, 0.5*y, g);
    double sx = sqrt(x);
    double e2 = fabs(d-sx)/sx; if(e2>mn) mn=e2;
  }
  // block max via atomics on bits (positive doubles order as uint64)
  atomicMax((unsigned long long*)maxerr, (unsigned long long)__double_as_longlong(me));
  atomicMax((unsigned long long*)maxerr_newton, (unsigned long long)__double_as_longlong(mn));
}

template<typename F> float timeit(F f){
  cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms,a,b); return ms;
}
int main(){
  int sms=0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  int clk=0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  const int blocks = sms*8, threads=256;
  void *o; CK(cudaMalloc(&o, (size_t)blocks*threads*8));
  double nthr = (double)blocks*threads;
  double warp_instr = nthr/32*ITER*8;
  auto rep=[&](const char*name, float ms, double instr_per_iter){
    double wi = warp_instr*instr_per_iter; // warp instrs of interest
    double per_smsp_clk = wi/(sms*4)/(ms*1e-3*clk*1e3);
    printf("%-10s %.3f ms  %.3f warp-instr/clk/SMSP (counting %g instr/iter)\n", name, ms, per_smsp_clk, instr_per_iter);
  };
  rep("FFMA", timeit([&]{k_ffma<<<blocks,threads>>>((float*)o,1.0001f,0.5f);}),1);
  rep("FFMA2", timeit([&]{k_ffma2<<<blocks,threads>>>((float*)o,1.0001f,0.5f);}),1);
  rep("DFMA", timeit([&]{k_dfma<<<blocks,threads>>>((double*)o,1.0001,0.5);}),1);
  rep("DADD", timeit([&]{k_dadd<<<blocks,threads>>>((double*)o,0.5);}),1);
  rep("DADD+DMIN", timeit([&]{k_dmin<<<blocks,threads>>>((double*)o,-0.5);}),4);
  rep("F2Fx2+DMUL", timeit([&]{k_f2f<<<blocks,threads>>>((float*)o,1.0001);}),3);
  rep("MUFU+FADD", timeit([&]{k_mufu<<<blocks,threads>>>((float*)o);}),2);
  rep("FMUL+clamp", timeit([&]{k_fsel<<<blocks,threads>>>((float*)o,1.f,1e30f);}),5);
  double *d; CK(cudaMalloc(&d,16)); CK(cudaMemset(d,0,16));
  k_rsq_acc<<<sms*4,256>>>(d,d+1,12345,20000);
  double h[2]; CK(cudaMemcpy(h,d,16,cudaMemcpyDeviceToHost));
  printf("rsqrt.approx.ftz.f64 max rel err = %.3e (2^%.2f); after one Newton step sqrt err = %.3e (2^%.2f)\n", h[0], log2(h[0]), h[1], log2(h[1]));
  printf("sms=%d clk=%d kHz\n", sms, clk);
  return 0;
}
