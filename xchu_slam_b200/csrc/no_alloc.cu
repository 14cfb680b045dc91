// no_alloc.cu -- ndtb_fitness_noalloc: getFitnessScore for callers that cannot call cudaMalloc (code under CUDA graph capture,
// tight serving loops, harnesses that forbid it).  It has NO counterpart in the reference;
// it changes no existing code path.
//
// What it computes is what ndtb_fitness computes (pcl::Registration::getFitnessScore: the mean, in double, of the float squared
// distances <= max_range from every transformed source point to its exact nearest target point), by the exhaustive variant of the
// query kernel (k_nn_far in voxel_build.cu: the target streamed through shared memory in tiles, slices combined with an integer
// atomicMin on the non-negative float bits) -- the cell grid of the ordinary path is what needs allocations.  Its scratch (one
// float per source point + {sum, count}) lives in a file-scope __device__ array of kNoAllocMaxSource floats: every GPU that loads
// the library holds those 1 MiB + 256 B until exit.  Every call shares the array and it keeps the last call's contents: each call
// re-arms what must start clean (the per-point minima, the sums), and two calls that overlap on the GPU (different streams,
// different host threads) RACE on it -- callers serialise.
//
// Compiled with -fmad=false like voxel_build.cu: the float expressions below are the ordinary path's, bit for bit.
#include <cfloat>
#include <cstdio>
#include <mutex>

#include <cuda_runtime.h>

#include "../../include/ndt_b200/ndt_c_api.h"

namespace {

constexpr size_t kNoAllocMaxSource = NDTB_NOALLOC_MAX_SOURCE;   // source points per call (ndt_c_api.h)
constexpr size_t kHeadBytes = 256;                              // {sum, count} as two doubles, padded to the array's alignment
constexpr int kThreads = 128, kTile = 1024, kMaxSlices = 64;
constexpr int kMaxDevices = 64;

// THE scratch.  In host code this name is a host-side stand-in, not a device address: the host asks the runtime for the device
// address once per device (scratch_address below) and hands that to the kernels.
static __device__ __align__(256) unsigned char g_scratch[kHeadBytes + kNoAllocMaxSource * sizeof(float)];

struct Mat16 { float m[16]; };   // the transform travels by value in the kernel arguments: no device copy of it is needed

__global__ void k_na_arm(float* __restrict__ d2, size_t n, double* __restrict__ head) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) d2[i] = FLT_MAX;
  if (i == 0) { head[0] = 0.0; head[1] = 0.0; }
}

// every query against one slice (blockIdx.y) of the target
__global__ void __launch_bounds__(kThreads) k_na_nearest(const float4* __restrict__ tgt, size_t n_tgt, const float4* __restrict__ src, size_t n_src,
                                                         Mat16 T, float* __restrict__ d2) {
  __shared__ float4 tile[kTile];
  const size_t i = (size_t)blockIdx.x * kThreads + threadIdx.x;
  const bool live = i < n_src;
  float qx = 0.f, qy = 0.f, qz = 0.f;
  if (live) {
    const float4 p = src[i];
    // pcl::transformPointCloud by the transform (float, left to right) -- the expression of k_nn_query / k_nn_far
    qx = ((T.m[0] * p.x + T.m[4] * p.y) + T.m[8] * p.z) + T.m[12];
    qy = ((T.m[1] * p.x + T.m[5] * p.y) + T.m[9] * p.z) + T.m[13];
    qz = ((T.m[2] * p.x + T.m[6] * p.y) + T.m[10] * p.z) + T.m[14];
  }
  float best = FLT_MAX;
  const size_t slice = (n_tgt + gridDim.y - 1) / gridDim.y;
  const size_t lo = (size_t)blockIdx.y * slice, hi = lo + slice < n_tgt ? lo + slice : n_tgt;
  for (size_t base = lo; base < hi; base += kTile) {
    const int m = (int)(hi - base < (size_t)kTile ? hi - base : (size_t)kTile);
    __syncthreads();
    for (int t = threadIdx.x; t < m; t += kThreads) tile[t] = tgt[base + t];
    __syncthreads();
#pragma unroll 8
    for (int t = 0; t < m; ++t) {
      const float4 tp = tile[t];
      const float ex = qx - tp.x, ey = qy - tp.y, ez = qz - tp.z;
      const float d = (ex * ex + ey * ey) + ez * ez;
      // (a target point with a non-finite coordinate gives NaN here and never wins the comparison)
      if (d < best) best = d;
    }
  }
  // squared distances are >= 0: their bit patterns order like the floats, so the slices combine with an integer atomicMin
  if (live) atomicMin(reinterpret_cast<int*>(d2) + i, __float_as_int(best));
}

// {sum, count} over d2 <= max_range, one block, fixed order (thread-strided partial sums, then lanes, then warps), then the mean
__global__ void __launch_bounds__(1024) k_na_mean(const float* __restrict__ d2, size_t n, double max_range, double* __restrict__ head,
                                                  double* __restrict__ out) {
  __shared__ double ss[32], sc[32];
  double s = 0.0, c = 0.0;
  for (size_t i = threadIdx.x; i < n; i += 1024) {
    const float d = d2[i];
    if (d != FLT_MAX && (double)d <= max_range) { s += (double)d; c += 1.0; }   // FLT_MAX: the armed value, i.e. no target point at all
  }
  for (int o = 16; o > 0; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); c += __shfl_xor_sync(0xffffffffu, c, o); }
  if ((threadIdx.x & 31) == 0) { ss[threadIdx.x >> 5] = s; sc[threadIdx.x >> 5] = c; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double ts = 0.0, tc = 0.0;
    for (int w = 0; w < 32; ++w) { ts += ss[w]; tc += sc[w]; }
    head[0] = ts; head[1] = tc;
    *out = tc > 0.0 ? ts / tc : 1.7976931348623157e308;   // std::numeric_limits<double>::max () when no point is in range
  }
}

thread_local char t_err[256] = "";
int fail(const char* what, cudaError_t ce = cudaSuccess) {
  if (ce != cudaSuccess) std::snprintf(t_err, sizeof(t_err), "ndtb_fitness_noalloc: %s: %s", what, cudaGetErrorString(ce));
  else std::snprintf(t_err, sizeof(t_err), "ndtb_fitness_noalloc: %s", what);
  return 1;
}

// The device address of g_scratch on `device` (which the caller has made current): asked for once per device.
std::mutex g_mu;
unsigned char* g_addr[kMaxDevices] = {nullptr};
cudaError_t scratch_address(int device, unsigned char** out) {
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_addr[device]) {
    void* p = nullptr;
    const cudaError_t ce = cudaGetSymbolAddress(&p, g_scratch);
    if (ce != cudaSuccess) return ce;
    g_addr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_addr[device];
  return cudaSuccess;
}

}  // namespace

extern "C" {

const char* ndtb_noalloc_last_error(void) { return t_err; }

int ndtb_fitness_noalloc(int device, const void* source_dev, size_t n_source, const void* target_dev, size_t n_target, const float T[16],
                         double max_range, double* out_dev, void* cuda_stream) {
  if (device < 0 || device >= kMaxDevices) return fail("device ordinal out of range");
  if (!T || !out_dev) return fail("null transform or output pointer");
  if (n_source > kNoAllocMaxSource) {
    std::snprintf(t_err, sizeof(t_err), "ndtb_fitness_noalloc: %zu source points exceed the scratch array's %zu (NDTB_NOALLOC_MAX_SOURCE)", n_source,
                  kNoAllocMaxSource);
    return 1;
  }
  if ((n_source > 0 && !source_dev) || (n_target > 0 && !target_dev)) return fail("null cloud pointer");
  cudaError_t ce = cudaSetDevice(device);
  if (ce != cudaSuccess) return fail("cudaSetDevice", ce);
  unsigned char* base = nullptr;
  ce = scratch_address(device, &base);
  if (ce != cudaSuccess) return fail("cudaGetSymbolAddress", ce);
  double* head = reinterpret_cast<double*>(base);
  float* d2 = reinterpret_cast<float*>(base + kHeadBytes);
  cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
  const float4* src = static_cast<const float4*>(source_dev);
  const float4* tgt = static_cast<const float4*>(target_dev);
  const unsigned qb = (unsigned)((n_source + kThreads - 1) / kThreads);
  k_na_arm<<<qb > 0 ? qb : 1, kThreads, 0, s>>>(d2, n_source, head);
  if (n_source > 0 && n_target > 0) {
    Mat16 M;
    for (int k = 0; k < 16; ++k) M.m[k] = T[k];
    unsigned slices = (unsigned)((n_target + 4095) / 4096);
    if (slices > (unsigned)kMaxSlices) slices = kMaxSlices;
    if (slices < 1) slices = 1;
    k_na_nearest<<<dim3(qb, slices), kThreads, 0, s>>>(tgt, n_target, src, n_source, M, d2);
  }
  k_na_mean<<<1, 1024, 0, s>>>(d2, n_source, max_range, head, out_dev);
  ce = cudaGetLastError();
  if (ce != cudaSuccess) return fail("kernel launch", ce);
  return 0;
}

}  // extern "C"
