#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b){u64 d; asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(a), "f"(b)); return d;}
__device__ __forceinline__ void upk(u64 d, float&a, float&b){asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(d));}
__global__ void k(float* o, const float* in){
  float a=in[threadIdx.x], b=in[threadIdx.x+32], c=in[64], d=in[65];
  u64 x=pk(a,b), y=pk(c,c), z=pk(d,d), r;
  asm("fma.rm.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(x), "l"(y), "l"(z));
  u64 r2; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r2) : "l"(r), "l"(y));
  u64 r3; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r3) : "l"(r2), "l"(x));
#ifdef SAT
  u64 r4; asm("fma.rn.sat.f32x2 %0, %1, %2, %3;" : "=l"(r4) : "l"(r3), "l"(y), "l"(z)); r3=r4;
#endif
#ifdef SUB
  u64 r5; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r5) : "l"(r3), "l"(y)); r3=r5;
#endif
  float p,q; upk(r3,p,q); o[threadIdx.x]=p; o[threadIdx.x+32]=q;
}
