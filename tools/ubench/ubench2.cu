// Round-2 micro-benchmarks / exhaustive checks behind the packed-f32 limiter math
// (k_pass.cu):
//   * issue rate of FFMA / FFMA2 / FMUL2 / FADD2 and of 3-input FMNMX
//   * does ptxas contract mul.rn.f32x2 + add.rn.f32x2 into FFMA2?  (numerical
//     check against separately rounded scalar ops)
//   * float32(|(1/3) * (double)q|)  ==  |Markstein q/3|  for EVERY float q
//     (numpy: abs((1./3.) * q) with q float32 promotes to float64, the result is
//     rounded to float32 at the end -- hfun/raster.py:1403-1405, 1432-1435)
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>
#include <math.h>

#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("%s: %s\n",#x,cudaGetErrorString(e)); return 1;}}while(0)

typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b){ u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(u64 v, float &a, float &b){ asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c){ u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b){ u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 add2(u64 a, u64 b){ u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

constexpr int ITER = 4096;
__global__ void k_ffma(float *o, float a, float b){
  float x[8]; for(int i=0;i<8;i++) x[i]=threadIdx.x+i;
  for(int it=0; it<ITER; ++it){
#pragma unroll
    for(int i=0;i<8;i++) x[i]=__fmaf_rn(x[i],a,b);
  }
  float s=0; for(int i=0;i<8;i++) s+=x[i]; o[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
__global__ void k_ffma2(float *o, float a, float b){
  u64 x[8]; for(int i=0;i<8;i++) x[i]=pk(threadIdx.x+i, i);
  u64 A=pk(a,a), B=pk(b,b);
  for(int it=0; it<ITER; ++it){
#pragma unroll
    for(int i=0;i<8;i++) x[i]=fma2(x[i],A,B);
  }
  float s=0; for(int i=0;i<8;i++){ float p,q; upk(x[i],p,q); s+=p+q; } o[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
__global__ void k_fmul2(float *o, float a){
  u64 x[8]; for(int i=0;i<8;i++) x[i]=pk(threadIdx.x+i, i+1);
  u64 A=pk(a,a);
  for(int it=0; it<ITER; ++it){
#pragma unroll
    for(int i=0;i<8;i++) x[i]=mul2(x[i],A);
  }
  float s=0; for(int i=0;i<8;i++){ float p,q; upk(x[i],p,q); s+=p+q; } o[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
__global__ void k_fadd2(float *o, float a){
  u64 x[8]; for(int i=0;i<8;i++) x[i]=pk(threadIdx.x+i, i+1);
  u64 A=pk(a,a);
  for(int it=0; it<ITER; ++it){
#pragma unroll
    for(int i=0;i<8;i++) x[i]=add2(x[i],A);
  }
  float s=0; for(int i=0;i<8;i++){ float p,q; upk(x[i],p,q); s+=p+q; } o[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
// mixed: FFMA2 and FMNMX alternating (different pipes)
__global__ void k_mix(float *o, float a, float b, float lo){
  u64 x[8]; for(int i=0;i<8;i++) x[i]=pk(threadIdx.x+i, i+1);
  float m[8]; for(int i=0;i<8;i++) m[i]=i;
  u64 A=pk(a,a), B=pk(b,b);
  for(int it=0; it<ITER; ++it){
#pragma unroll
    for(int i=0;i<8;i++){ x[i]=fma2(x[i],A,B); m[i]=fmaxf(m[i]*1.0f, lo); }
  }
  float s=0; for(int i=0;i<8;i++){ float p,q; upk(x[i],p,q); s+=p+q+m[i]; } o[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
__global__ void k_min3(float *o, float a, float b){
  float x[8]; for(int i=0;i<8;i++) x[i]=threadIdx.x+i;
  for(int it=0; it<ITER; ++it){
#pragma unroll
    for(int i=0;i<8;i++){ float r; asm volatile("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(x[i]), "f"(a), "f"(b)); x[i]=r+1.0f; }
  }
  float s=0; for(int i=0;i<8;i++) s+=x[i]; o[blockIdx.x*blockDim.x+threadIdx.x]=s;
}

// ---- contraction check ----------------------------------------------------
__global__ void k_contract(unsigned *bad, uint64_t seed, int n){
  uint64_t s = seed + (blockIdx.x*blockDim.x+threadIdx.x)*0x9E3779B97F4A7C15ull;
  unsigned nb=0;
  for(int i=0;i<n;i++){
    float v[6];
    for(int j=0;j<6;j++){ s ^= s<<13; s ^= s>>7; s ^= s<<17; v[j] = 1.0f + (float)(s & 0x7fffff) * (1.0f/8388608.0f); }
    u64 p = mul2(pk(v[0],v[1]), pk(v[2],v[3]));
    u64 r = add2(p, pk(v[4],-v[5]));
    float r0,r1; upk(r,r0,r1);
    float e0 = __fadd_rn(__fmul_rn(v[0],v[2]), v[4]);
    float e1 = __fadd_rn(__fmul_rn(v[1],v[3]), -v[5]);
    nb += (r0!=e0) + (r1!=e1);
  }
  if(nb) atomicAdd(bad, nb);
}

// ---- exhaustive q/3 --------------------------------------------------------
__global__ void k_third(unsigned long long *bad, unsigned *first_bad){
  const float y = 1.0f/3.0f;   // RN(1/3)
  for(uint64_t i = blockIdx.x*(uint64_t)blockDim.x+threadIdx.x; i < (1ull<<32); i += (uint64_t)gridDim.x*blockDim.x){
    const float q = __uint_as_float((unsigned)i);
    const float ref = (float)fabs(__dmul_rn(1.0/3.0, (double)q));
    const float h0 = __fmul_rn(q, y);
    const float rem = __fmaf_rn(-h0, 3.0f, q);
    const float h = fabsf(__fmaf_rn(rem, y, h0));
    const bool same = (__float_as_uint(h)==__float_as_uint(ref)) || (h!=h && ref!=ref);
    if(!same){ atomicAdd(bad, 1ull); atomicMin(first_bad, (unsigned)i); }
  }
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
    double wi = warp_instr*instr_per_iter;
    double per_smsp_clk = wi/(sms*4)/(ms*1e-3*clk*1e3);
    printf("%-10s %.3f ms  %.3f warp-instr/clk/SMSP (counting %g instr/iter)\n", name, ms, per_smsp_clk, instr_per_iter);
  };
  rep("FFMA", timeit([&]{k_ffma<<<blocks,threads>>>((float*)o,1.0001f,0.5f);}),1);
  rep("FFMA2", timeit([&]{k_ffma2<<<blocks,threads>>>((float*)o,1.0001f,0.5f);}),1);
  rep("FMUL2", timeit([&]{k_fmul2<<<blocks,threads>>>((float*)o,1.0001f);}),1);
  rep("FADD2", timeit([&]{k_fadd2<<<blocks,threads>>>((float*)o,0.5f);}),1);
  rep("FFMA2+FMUL+FMNMX", timeit([&]{k_mix<<<blocks,threads>>>((float*)o,1.0001f,0.5f,1.0f);}),3);
  rep("MIN3+FADD", timeit([&]{k_min3<<<blocks,threads>>>((float*)o,1e30f,1e31f);}),2);
  unsigned *d; CK(cudaMalloc(&d,32)); CK(cudaMemset(d,0,32));
  k_contract<<<sms*4,256>>>(d,777,2000);
  unsigned hb; CK(cudaMemcpy(&hb,d,4,cudaMemcpyDeviceToHost));
  printf("mul.rn.f32x2 + add.rn.f32x2 vs separately rounded scalar: %u mismatches of %llu (0 = not contracted)\n", hb, (unsigned long long)sms*4*256*2000*2);
  unsigned long long *b2; unsigned *fb; CK(cudaMalloc(&b2,8)); CK(cudaMalloc(&fb,4));
  CK(cudaMemset(b2,0,8)); CK(cudaMemset(fb,0xff,4));
  k_third<<<sms*16,256>>>(b2,fb);
  unsigned long long nb; unsigned f; CK(cudaMemcpy(&nb,b2,8,cudaMemcpyDeviceToHost)); CK(cudaMemcpy(&f,fb,4,cudaMemcpyDeviceToHost));
  printf("Markstein q/3 vs float32(|(1/3)*(double)q|): %llu mismatches over all 2^32 floats (first bits 0x%08x)\n", nb, f);
  printf("sms=%d clk=%d kHz\n", sms, clk);
  return 0;
}
