#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b){u64 d; asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(a), "f"(b)); return d;}
__device__ __forceinline__ void upk(u64 d, float&a, float&b){asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(d));}
__global__ void k(float* o, const float* in){
  float a=in[threadIdx.x], b=in[threadIdx.x+32], c=in[64], d=in[65];
  u64 x=pk(a,b), y=pk(c,c), z=pk(d,d);
  u64 r3; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r3) : "l"(x), "l"(y));
  u64 r5; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r5) : "l"(r3), "l"(z)); r3=r5;
  float p,q; upk(r3,p,q); o[threadIdx.x]=p; o[threadIdx.x+32]=q;
}
__global__ void k2(float* o, const float* in){
  float a=in[threadIdx.x], b=in[threadIdx.x+32], c=in[64], d=in[65];
  float2 r = __fmul2_rn(make_float2(a,b), make_float2(c,c));
  r = __fadd2_rn(r, make_float2(d,d));
  o[threadIdx.x]=r.x; o[threadIdx.x+32]=r.y;
}
