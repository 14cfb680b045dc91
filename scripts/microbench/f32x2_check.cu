// Does the packed path round like the scalar one?  A2(M2(a,b), M2(c,d)) through the sm_100 intrinsics and through inline PTX
// (mul.rn.f32x2 / add.rn.f32x2) against __fadd_rn(__fmul_rn(a,b), __fmul_rn(c,d)), bit for bit, on random operands.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o f32x2_check f32x2_check.cu && ./f32x2_check
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ float2 mul2_asm(float2 a, float2 b) {
  unsigned long long r, ua, ub;
  ua = ((unsigned long long)__float_as_uint(a.y) << 32) | __float_as_uint(a.x);
  ub = ((unsigned long long)__float_as_uint(b.y) << 32) | __float_as_uint(b.x);
  asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(ua), "l"(ub));
  return make_float2(__uint_as_float((unsigned)r), __uint_as_float((unsigned)(r >> 32)));
}
__device__ __forceinline__ float2 add2_asm(float2 a, float2 b) {
  unsigned long long r, ua, ub;
  ua = ((unsigned long long)__float_as_uint(a.y) << 32) | __float_as_uint(a.x);
  ub = ((unsigned long long)__float_as_uint(b.y) << 32) | __float_as_uint(b.x);
  asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(ua), "l"(ub));
  return make_float2(__uint_as_float((unsigned)r), __uint_as_float((unsigned)(r >> 32)));
}
// sum of two packed values as fma(a, 1, b): exact (a * 1 is a), and an FMA cannot absorb the multiply that produced a or b
__device__ __forceinline__ float2 add2_fma(float2 a, float2 b, float one) {
  unsigned long long r, ua, ub, uo;
  ua = ((unsigned long long)__float_as_uint(a.y) << 32) | __float_as_uint(a.x);
  ub = ((unsigned long long)__float_as_uint(b.y) << 32) | __float_as_uint(b.x);
  uo = ((unsigned long long)__float_as_uint(one) << 32) | __float_as_uint(one);
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(ua), "l"(uo), "l"(ub));
  return make_float2(__uint_as_float((unsigned)r), __uint_as_float((unsigned)(r >> 32)));
}
__device__ float rnd(uint32_t& s) { s = s * 1664525u + 1013904223u; return ((int)(s >> 8) - (1 << 23)) * (1.0f / (1 << 20)); }
__global__ void k(unsigned long long* bad, float one) {
  uint32_t s = blockIdx.x * 1315423911u + threadIdx.x * 2654435761u + 12345u;
  unsigned long long b_intr = 0, b_asm = 0, b_chain = 0;
  for (int it = 0; it < 4096; ++it) {
    const float a = rnd(s), b = rnd(s), c = rnd(s), d = rnd(s), e = rnd(s), f = rnd(s);
    const float ref0 = __fadd_rn(__fmul_rn(a, b), __fmul_rn(c, d)), ref1 = __fadd_rn(__fmul_rn(e, b), __fmul_rn(f, d));
    const float2 p = make_float2(a, e), q = make_float2(c, f);
    const float2 r1 = __fadd2_rn(__fmul2_rn(p, make_float2(b, b)), __fmul2_rn(q, make_float2(d, d)));
    const float2 r2 = add2_fma(__fmul2_rn(p, make_float2(b, b)), __fmul2_rn(q, make_float2(d, d)), one);
    // the three-term chain of the kernel: (x0*A + x1*B) + x2*C
    const float ref2 = __fadd_rn(__fadd_rn(__fmul_rn(a, b), __fmul_rn(c, d)), __fmul_rn(e, f));
    const float2 r3 = add2_fma(add2_fma(__fmul2_rn(make_float2(a, a), make_float2(b, b)), __fmul2_rn(make_float2(c, c), make_float2(d, d)), one),
                               __fmul2_rn(make_float2(e, e), make_float2(f, f)), one);
    b_intr += (__float_as_uint(r1.x) != __float_as_uint(ref0)) + (__float_as_uint(r1.y) != __float_as_uint(ref1));
    b_asm += (__float_as_uint(r2.x) != __float_as_uint(ref0)) + (__float_as_uint(r2.y) != __float_as_uint(ref1));
    b_chain += (__float_as_uint(r3.x) != __float_as_uint(ref2)) + (__float_as_uint(r3.y) != __float_as_uint(ref2));
  }
  atomicAdd(bad + 0, b_intr); atomicAdd(bad + 1, b_asm); atomicAdd(bad + 2, b_chain);
}
int main() {
  unsigned long long* d; unsigned long long h[3] = {0, 0, 0};
  cudaMalloc(&d, 24); cudaMemset(d, 0, 24);
  k<<<256, 128>>>(d, 1.0f);
  cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost);
  printf("mismatches against the scalar reference out of %d results each: intrinsics mul2+add2 %llu, mul2 + fma(a,1,b) %llu, three-term chain with fma(a,1,b) %llu\n",
         256 * 128 * 4096 * 2, h[0], h[1], h[2]);
  return 0;
}
