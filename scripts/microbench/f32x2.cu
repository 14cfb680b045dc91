// Micro-benchmark: throughput of scalar FMUL / FADD against the packed sm_100 forms (__fmul2_rn / __fadd2_rn: two IEEE-rounded
// FP32 operations per lane and instruction).  Each thread runs 8 independent dependency chains; prints G lane-operations / s.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o f32x2 f32x2.cu && ./f32x2
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float* out, int iters, float a, float b) {
  float2 x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = make_float2(1.f + threadIdx.x * 1e-6f + i, 2.f + i);
  const float2 a2 = make_float2(a, a), b2 = make_float2(b, b);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) { x[i].x = __fmul_rn(x[i].x, a); x[i].y = __fmul_rn(x[i].y, a); }
      if (MODE == 1) x[i] = __fmul2_rn(x[i], a2);
      if (MODE == 2) { x[i].x = __fadd_rn(x[i].x, b); x[i].y = __fadd_rn(x[i].y, b); }
      if (MODE == 3) x[i] = __fadd2_rn(x[i], b2);
      if (MODE == 4) { x[i].x = __fadd_rn(__fmul_rn(x[i].x, a), b); x[i].y = __fadd_rn(__fmul_rn(x[i].y, a), b); }
      if (MODE == 5) x[i] = __fadd2_rn(__fmul2_rn(x[i], a2), b2);
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += x[i].x + x[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
void run(const char* name, int ops_per_elem, int warps_per_sm) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int threads = 128, blocks = sms * warps_per_sm / 4, iters = 20000;
  float* out;
  cudaMalloc(&out, (size_t)blocks * threads * 4);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<blocks, threads>>>(out, 100, 1.0000001f, 1e-9f);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<MODE><<<blocks, threads>>>(out, iters, 1.0000001f, 1e-9f);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  const double ops = (double)blocks * threads * iters * 16.0 * ops_per_elem;
  printf("%-22s warps/SM %2d: %8.1f G lane-ops/s  (%.3f ms)  = %.1f lane-ops / clk / SM at 1.965 GHz\n", name, warps_per_sm, ops / ms / 1e6, ms,
         ops / (ms * 1e-3) / sms / 1.965e9);
  cudaFree(out);
}
int main() {
  for (int w : {4, 8, 16}) {
    run<0>("FMUL scalar", 1, w);
    run<1>("FMUL2 packed", 1, w);
    run<2>("FADD scalar", 1, w);
    run<3>("FADD2 packed", 1, w);
    run<4>("FMUL+FADD scalar", 2, w);
    run<5>("FMUL2+FADD2 packed", 2, w);
  }
  return 0;
}
