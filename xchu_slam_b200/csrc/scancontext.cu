// scancontext.cu -- ScanContext descriptors, column-shift distance and loop detection on the device
// (C ABI: include/ndt_b200/sc_c_api.h).  Reference: xchu_mapping/include/scancontext/
// Scancontext.cpp (= SC) and Scancontext.h:83-103.  Every reduction runs in the reference's order (sequential, ascending
// index, as the oracle and the stand-in Eigen evaluate norm / dot / mean) and this file is compiled with -fmad=false, so the
// doubles are the oracle's bit for bit; the per-bin maximum of makeScancontext is order-independent.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/ndt_b200/sc_c_api.h"

namespace {

constexpr int kMaxRing = 64, kMaxSector = 128, kMaxCand = 16;
constexpr double kPi = 3.14159265358979323846;   // M_PI

struct DevParams {
  int R, S, cand;
  double max_radius, lidar_height, search_ratio, dist_thres;
};

// ---- makeScancontext (SC:151-196) ------------------------------------------------------------------------------
// SC:23-37.  (180 / M_PI) is a double constant; atan takes the float quotient widened to double; the result narrows to float.
__device__ __forceinline__ float xy2theta(float x, float y) {
  const double k = 180 / kPi;
  if ((x >= 0) & (y >= 0)) return (float)(k * atan((double)(y / x)));
  if ((x < 0) & (y >= 0)) return (float)(180 - (k * atan((double)(y / (-x)))));
  if ((x < 0) & (y < 0)) return (float)(180 + (k * atan((double)(y / x))));
  if ((x >= 0) & (y < 0)) return (float)(360 - (k * atan((double)((-y) / x))));
  return 0.f;
}
// order-preserving float <-> int
__device__ __forceinline__ int enc(float f) { const int b = __float_as_int(f); return b >= 0 ? b : b ^ 0x7fffffff; }
__device__ __forceinline__ float dec(int e) { return __int_as_float(e >= 0 ? e : e ^ 0x7fffffff); }

// one block per scan: per-bin maximum of z + lidar_height, then the two keys
__global__ void __launch_bounds__(256) k_sc_make(DevParams P, const float4* __restrict__ pts, const long long* __restrict__ offs, int first_id,
                                                  double* __restrict__ desc, double* __restrict__ ring, double* __restrict__ sector,
                                                  float* __restrict__ invkey) {
  extern __shared__ double sm_d[];                 // R*S doubles, the first R*S ints double as the bins
  int* bins = reinterpret_cast<int*>(sm_d + P.R * P.S);
  const int R = P.R, S = P.S, tid = threadIdx.x;
  const int scan = blockIdx.x;
  const int no_point = enc(-1000.0f);              // NO_POINT (SC:158): a bin is raised only by z > -1000
  for (int k = tid; k < R * S; k += blockDim.x) bins[k] = no_point;
  __syncthreads();
  const long long b = offs[scan], e = offs[scan + 1];
  for (long long i = b + tid; i < e; i += blockDim.x) {
    const float4 p = pts[i];
    const float pz = (float)((double)p.z + P.lidar_height);                       // SC:168
    const float azim_range = (float)sqrt((double)(p.x * p.x + p.y * p.y));       // SC:171
    const float azim_angle = xy2theta(p.x, p.y);
    if ((double)azim_range > P.max_radius) continue;                              // SC:175
    int ring_idx = (int)ceil(((double)azim_range / P.max_radius) * R);
    ring_idx = max(min(R, ring_idx), 1);
    int sctor_idx = (int)ceil(((double)azim_angle / 360.0) * S);
    sctor_idx = max(min(S, sctor_idx), 1);
    atomicMax(&bins[(sctor_idx - 1) * R + (ring_idx - 1)], enc(pz));              // SC:182-183
  }
  __syncthreads();
  for (int k = tid; k < R * S; k += blockDim.x) {
    const int v = bins[k];
    sm_d[k] = v == no_point ? 0.0 : (double)dec(v);                               // SC:187-190
  }
  __syncthreads();
  const size_t id = (size_t)first_id + scan;
  double* d = desc + id * R * S;
  for (int k = tid; k < R * S; k += blockDim.x) d[k] = sm_d[k];
  if (tid < R) {                                                                   // SC:199-211 row mean
    double s = sm_d[tid];
    for (int c = 1; c < S; ++c) s += sm_d[c * R + tid];
    const double m = s / (double)S;
    ring[id * R + tid] = m;
    invkey[id * R + tid] = (float)m;                                               // eig2stdvec, SC:64-68
  }
  if (tid >= 64 && tid - 64 < S) {                                                 // SC:214-228 column mean
    const int c = tid - 64;
    double s = sm_d[c * R];
    for (int r = 1; r < R; ++r) s += sm_d[c * R + r];
    sector[id * S + c] = s / (double)R;
  }
}

// ---- distanceBtnScanContext (SC:118-148), cooperatively by one block of 128 threads -----------------------------------------
struct DistShared {
  double v1[kMaxSector], v2[kMaxSector];     // sector keys
  double n1[kMaxSector], n2[kMaxSector];     // column norms
  double t[kMaxSector];                      // per-shift norms, then per-column similarities
  int skip[kMaxSector];
  double best;
  int best_shift, align;
};
__device__ __forceinline__ int src_col(int c, int shift, int S) { return ((c - shift) % S + S) % S; }   // column c of circshift(m, shift)

// a, b: R x S column-major descriptors in global memory; sk1 / sk2: their sector keys.  Result in sh.best / sh.best_shift.
__device__ void block_distance(const DevParams& P, const double* __restrict__ a, const double* __restrict__ b,
                               const double* __restrict__ sk1, const double* __restrict__ sk2, DistShared& sh) {
  const int R = P.R, S = P.S, tid = threadIdx.x;
  if (tid < S) {
    sh.v1[tid] = sk1[tid];
    sh.v2[tid] = sk2[tid];
    const double* ca = a + (size_t)tid * R;
    const double* cb = b + (size_t)tid * R;
    double na = ca[0] * ca[0], nb = cb[0] * cb[0];
    for (int r = 1; r < R; ++r) { na += ca[r] * ca[r]; nb += cb[r] * cb[r]; }
    sh.n1[tid] = sqrt(na);
    sh.n2[tid] = sqrt(nb);
  }
  __syncthreads();
  if (tid < S) {                                  // fastAlignUsingVkey (SC:96-115): thread = shift
    double d0 = sh.v1[0] - sh.v2[src_col(0, tid, S)];
    double ss = d0 * d0;
    for (int c = 1; c < S; ++c) { const double d = sh.v1[c] - sh.v2[src_col(c, tid, S)]; ss += d * d; }
    sh.t[tid] = sqrt(ss);
  }
  __syncthreads();
  if (tid == 0) {
    int argmin = 0;
    double mn = 10000000;
    for (int s = 0; s < S; ++s)
      if (sh.t[s] < mn) { argmin = s; mn = sh.t[s]; }
    sh.align = argmin;
    sh.best = 10000000;
    sh.best_shift = 0;
  }
  __syncthreads();
  const int radius = (int)round(0.5 * P.search_ratio * S);     // SC:126
  // the search space {align, align +- 1 .. radius} mod S in ascending order (SC:127-133): walk all shifts, keep members
  for (int shift = 0; shift < S; ++shift) {
    int dd = shift - sh.align;
    dd = ((dd % S) + S) % S;
    const bool member = dd <= radius || S - dd <= radius;
    if (!member) continue;                       // block-uniform
    if (tid < S) {                               // distDirectSC (SC:71-93) against circshift(b, shift): thread = column
      const int cs = src_col(tid, shift, S);
      const double* ca = a + (size_t)tid * R;
      const double* cb = b + (size_t)cs * R;
      double dot = ca[0] * cb[0];
      for (int r = 1; r < R; ++r) dot += ca[r] * cb[r];
      const double n1 = sh.n1[tid], n2 = sh.n2[cs];
      const bool sk = (n1 == 0) | (n2 == 0);
      sh.skip[tid] = sk ? 1 : 0;
      sh.t[tid] = sk ? 0.0 : dot / (n1 * n2);
    }
    __syncthreads();
    if (tid == 0) {
      int num_eff = 0;
      double sum = 0;
      for (int c = 0; c < S; ++c)
        if (!sh.skip[c]) { sum = sum + sh.t[c]; num_eff = num_eff + 1; }
      const double cur = 1.0 - sum / num_eff;
      if (cur < sh.best) { sh.best = cur; sh.best_shift = shift; }
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(128) k_sc_distance(DevParams P, const double* __restrict__ desc, const double* __restrict__ sector,
                                                      const int* __restrict__ id1, const int* __restrict__ id2, double* __restrict__ dist,
                                                      int* __restrict__ shift) {
  __shared__ DistShared sh;
  const int i = blockIdx.x;
  const size_t a = (size_t)id1[i], b = (size_t)id2[i];
  block_distance(P, desc + a * P.R * P.S, desc + b * P.R * P.S, sector + a * P.S, sector + b * P.S, sh);
  if (threadIdx.x == 0) { dist[i] = sh.best; shift[i] = sh.best_shift; }
}

// ---- detectLoopClosureID (SC:331-423) for one query per block -----------------------------------------------------------------
// squared L2 between float ring keys as nanoflann's L2_Adaptor::evalMetric accumulates it (nanoflann.hpp:383-408)
__device__ __forceinline__ float key_dist2(const float* __restrict__ a, const float* __restrict__ b, int dim) {
  float result = 0.f;
  int k = 0;
  for (; k + 4 <= dim; k += 4) {
    const float d0 = a[k] - b[k], d1 = a[k + 1] - b[k + 1], d2 = a[k + 2] - b[k + 2], d3 = a[k + 3] - b[k + 3];
    result += d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
  }
  for (; k < dim; ++k) { const float d = a[k] - b[k]; result += d * d; }
  return result;
}

__global__ void __launch_bounds__(128) k_sc_detect(DevParams P, const double* __restrict__ desc, const double* __restrict__ sector,
                                                    const float* __restrict__ invkey, const int* __restrict__ query_id,
                                                    const int* __restrict__ tree_size, int* __restrict__ loop_id, float* __restrict__ yaw,
                                                    double* __restrict__ min_dist, int* __restrict__ nn_idx) {
  __shared__ DistShared sh;
  __shared__ float q[kMaxRing];
  __shared__ float wd[4];
  __shared__ int wi[4];
  __shared__ int cand[kMaxCand];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int qi = blockIdx.x;
  const int n = tree_size[qi];
  const size_t qid = (size_t)query_id[qi];
  if (n <= 0) {                                                                    // SC:343-347
    if (tid == 0) { loop_id[qi] = -1; yaw[qi] = 0.f; min_dist[qi] = 10000000; nn_idx[qi] = 0; }
    return;
  }
  if (tid < P.R) q[tid] = invkey[qid * P.R + tid];
  __syncthreads();
  // the k nearest ring keys in ascending (distance, index) order: k rounds of a block-wide arg-min over the keys not taken yet
  const int k = min(P.cand, n);
  for (int round_ = 0; round_ < k; ++round_) {
    float bd = INFINITY;
    int bi = 0x7fffffff;
    for (int i = tid; i < n; i += blockDim.x) {
      bool taken = false;
      for (int c = 0; c < round_; ++c) taken |= cand[c] == i;
      if (taken) continue;
      const float d = key_dist2(q, invkey + (size_t)i * P.R, P.R);
      if (d < bd || (d == bd && i < bi)) { bd = d; bi = i; }
    }
    for (int o = 16; o; o >>= 1) {
      const float od = __shfl_down_sync(0xffffffffu, bd, o);
      const int oi = __shfl_down_sync(0xffffffffu, bi, o);
      if (od < bd || (od == bd && oi < bi)) { bd = od; bi = oi; }
    }
    if (lane == 0) { wd[warp] = bd; wi[warp] = bi; }
    __syncthreads();
    if (tid == 0) {
      for (int w = 1; w < 4; ++w)
        if (wd[w] < wd[0] || (wd[w] == wd[0] && wi[w] < wi[0])) { wd[0] = wd[w]; wi[0] = wi[w]; }
      cand[round_] = wi[0] == 0x7fffffff ? 0 : wi[0];
    }
    __syncthreads();
  }
  // fewer keys than candidates: the reference's result slots keep their zero initialisation (SC:372-373) -> descriptor 0 again
  if (tid == 0)
    for (int c = k; c < P.cand; ++c) cand[c] = 0;
  __syncthreads();
  double best = 10000000;
  int best_align = 0, best_idx = 0;
  for (int c = 0; c < P.cand; ++c) {                                               // SC:383-397
    const size_t ci = (size_t)cand[c];
    block_distance(P, desc + qid * P.R * P.S, desc + ci * P.R * P.S, sector + qid * P.S, sector + ci * P.S, sh);
    if (sh.best < best) { best = sh.best; best_align = sh.best_shift; best_idx = (int)ci; }
    __syncthreads();
  }
  if (tid == 0) {
    loop_id[qi] = best < P.dist_thres ? best_idx : -1;                             // SC:403-406
    const double unit = 360.0 / double(P.S);                                       // PC_UNIT_SECTORANGLE
    const float deg = (float)(best_align * unit);                                  // float argument of deg2rad (SC:17-20, 419)
    yaw[qi] = (float)((double)deg * kPi / 180.0);
    min_dist[qi] = best;
    nn_idx[qi] = best_idx;
  }
}

}  // namespace

struct ndtb_sc {
  int device = 0;
  ndtb_sc_params p;
  std::string err;
  cudaStream_t stream = nullptr;
  size_t n = 0, cap = 0;
  double *desc = nullptr, *ring = nullptr, *sector = nullptr;
  float* invkey = nullptr;
  float4* pts = nullptr;
  size_t pts_cap = 0;
  float4* h_pts = nullptr;       // pinned staging
  size_t h_pts_cap = 0;
  long long* offs = nullptr;
  size_t offs_cap = 0;
  void* scratch = nullptr;       // ids in / results out of distance and detect
  size_t scratch_cap = 0;
  int fail(const std::string& m) { err = m; return 1; }
  DevParams dev() const { return DevParams{p.num_ring, p.num_sector, p.num_candidates, p.max_radius, p.lidar_height, p.search_ratio, p.dist_thres}; }
  int reserve(size_t need) {
    if (need <= cap) return 0;
    const size_t nc = need + need / 2 + 64;
    const size_t RS = (size_t)p.num_ring * p.num_sector;
    double *d = nullptr, *r = nullptr, *s = nullptr;
    float* k = nullptr;
    if (cudaMalloc(&d, nc * RS * 8) != cudaSuccess || cudaMalloc(&r, nc * p.num_ring * 8) != cudaSuccess ||
        cudaMalloc(&s, nc * p.num_sector * 8) != cudaSuccess || cudaMalloc(&k, nc * p.num_ring * 4) != cudaSuccess) {
      cudaFree(d); cudaFree(r); cudaFree(s); cudaFree(k);
      return fail("descriptor store allocation failed");
    }
    if (n) {
      cudaMemcpyAsync(d, desc, n * RS * 8, cudaMemcpyDeviceToDevice, stream);
      cudaMemcpyAsync(r, ring, n * p.num_ring * 8, cudaMemcpyDeviceToDevice, stream);
      cudaMemcpyAsync(s, sector, n * p.num_sector * 8, cudaMemcpyDeviceToDevice, stream);
      cudaMemcpyAsync(k, invkey, n * p.num_ring * 4, cudaMemcpyDeviceToDevice, stream);
    }
    if (cudaStreamSynchronize(stream) != cudaSuccess) return fail("descriptor store copy failed");
    cudaFree(desc); cudaFree(ring); cudaFree(sector); cudaFree(invkey);
    desc = d; ring = r; sector = s; invkey = k;
    cap = nc;
    return 0;
  }
  int ensure_scratch(size_t bytes) {
    if (bytes <= scratch_cap) return 0;
    cudaFree(scratch);
    scratch = nullptr;
    scratch_cap = bytes + bytes / 2 + 256;
    if (cudaMalloc(&scratch, scratch_cap) != cudaSuccess) { scratch_cap = 0; return fail("scratch allocation failed"); }
    return 0;
  }
};

static std::string g_sc_create_error;

extern "C" {

void ndtb_sc_default_params(ndtb_sc_params* p) {
  p->num_ring = 20; p->num_sector = 60; p->max_radius = 80.0; p->lidar_height = 2.0; p->search_ratio = 0.1;
  p->num_exclude_recent = 30; p->num_candidates = 3; p->tree_making_period = 30; p->dist_thres = 0.2;
}

int ndtb_sc_create(int device, const ndtb_sc_params* p, ndtb_sc** out) {
  if (!out) return 1;
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) { g_sc_create_error = "no CUDA device available (this library has no CPU fallback)"; return 1; }
  if (device < 0 || device >= count) { g_sc_create_error = "invalid device ordinal"; return 1; }
  ndtb_sc* s = new ndtb_sc();
  s->device = device;
  if (p) s->p = *p; else ndtb_sc_default_params(&s->p);
  if (s->p.num_ring < 1 || s->p.num_ring > kMaxRing || s->p.num_sector < 1 || s->p.num_sector > kMaxSector || s->p.num_candidates < 1 ||
      s->p.num_candidates > kMaxCand || !(s->p.max_radius > 0)) {
    g_sc_create_error = "parameter out of range (num_ring <= 64, num_sector <= 128, num_candidates <= 16)";
    delete s;
    return 1;
  }
  if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) != cudaSuccess) {
    g_sc_create_error = "stream creation failed";
    delete s;
    return 1;
  }
  *out = s;
  return 0;
}
int ndtb_sc_destroy(ndtb_sc* s) {
  if (!s) return 0;
  cudaSetDevice(s->device);
  cudaFree(s->desc); cudaFree(s->ring); cudaFree(s->sector); cudaFree(s->invkey); cudaFree(s->pts); cudaFree(s->offs); cudaFree(s->scratch);
  cudaFreeHost(s->h_pts);
  if (s->stream) cudaStreamDestroy(s->stream);
  delete s;
  return 0;
}
const char* ndtb_sc_last_error(const ndtb_sc* s) { return s ? s->err.c_str() : g_sc_create_error.c_str(); }
int ndtb_sc_set_dist_thres(ndtb_sc* s, double thres) { s->p.dist_thres = thres; return 0; }
int ndtb_sc_size(const ndtb_sc* s) { return (int)s->n; }
int ndtb_sc_clear(ndtb_sc* s) { s->n = 0; return 0; }

int ndtb_sc_add_scans(ndtb_sc* s, int n_scans, const void* const* pts, const size_t* n_points, size_t stride_bytes, int* first_id) {
  if (cudaSetDevice(s->device) != cudaSuccess) return s->fail("cudaSetDevice failed");
  if (first_id) *first_id = (int)s->n;
  if (n_scans <= 0) return 0;
  if (stride_bytes < 12) return s->fail("stride_bytes must be >= 12");
  size_t total = 0;
  std::vector<long long> offs((size_t)n_scans + 1);
  for (int i = 0; i < n_scans; ++i) { offs[i] = (long long)total; total += n_points[i]; }
  offs[n_scans] = (long long)total;
  if (s->reserve(s->n + n_scans)) return 1;
  if (cudaStreamSynchronize(s->stream) != cudaSuccess) return s->fail("sync failed");     // staging may still be in flight
  const bool direct = stride_bytes == 16;     // 16-byte records go up as they are (the kernel reads x y z only)
  if (!direct && total > s->h_pts_cap) {
    cudaFreeHost(s->h_pts);
    s->h_pts = nullptr;
    s->h_pts_cap = total + total / 4 + 1024;
    if (cudaMallocHost(&s->h_pts, s->h_pts_cap * sizeof(float4)) != cudaSuccess) { s->h_pts_cap = 0; return s->fail("pinned staging allocation failed"); }
  }
  if (total > s->pts_cap) {
    cudaFree(s->pts);
    s->pts = nullptr;
    s->pts_cap = total + total / 4 + 1024;
    if (cudaMalloc(&s->pts, s->pts_cap * sizeof(float4)) != cudaSuccess) { s->pts_cap = 0; return s->fail("point buffer allocation failed"); }
  }
  if ((size_t)n_scans + 1 > s->offs_cap) {
    cudaFree(s->offs);
    s->offs = nullptr;
    s->offs_cap = (size_t)n_scans + 65;
    if (cudaMalloc(&s->offs, s->offs_cap * sizeof(long long)) != cudaSuccess) { s->offs_cap = 0; return s->fail("offset buffer allocation failed"); }
  }
  if (direct) {
    for (int i = 0; i < n_scans; ++i)
      if (n_points[i] && cudaMemcpyAsync(s->pts + offs[i], pts[i], n_points[i] * sizeof(float4), cudaMemcpyHostToDevice, s->stream) != cudaSuccess)
        return s->fail("upload failed");
  } else {
    for (int i = 0; i < n_scans; ++i) {
      const char* src = static_cast<const char*>(pts[i]);
      float4* dst = s->h_pts + offs[i];
      for (size_t j = 0; j < n_points[i]; ++j) {
        const float* f = reinterpret_cast<const float*>(src + j * stride_bytes);
        dst[j] = make_float4(f[0], f[1], f[2], 0.f);
      }
    }
    if (total && cudaMemcpyAsync(s->pts, s->h_pts, total * sizeof(float4), cudaMemcpyHostToDevice, s->stream) != cudaSuccess) return s->fail("upload failed");
  }
  if (cudaMemcpyAsync(s->offs, offs.data(), offs.size() * sizeof(long long), cudaMemcpyHostToDevice, s->stream) != cudaSuccess) return s->fail("upload failed");
  const size_t RS = (size_t)s->p.num_ring * s->p.num_sector;
  const size_t smem = RS * sizeof(double) + RS * sizeof(int);
  k_sc_make<<<n_scans, 256, smem, s->stream>>>(s->dev(), s->pts, s->offs, (int)s->n, s->desc, s->ring, s->sector, s->invkey);
  if (cudaGetLastError() != cudaSuccess || cudaStreamSynchronize(s->stream) != cudaSuccess) return s->fail("descriptor kernel failed");
  s->n += n_scans;
  return 0;
}

int ndtb_sc_get(ndtb_sc* s, int id, double* desc, double* ring, double* sector, float* invkey) {
  if (cudaSetDevice(s->device) != cudaSuccess) return s->fail("cudaSetDevice failed");
  if (id < 0 || (size_t)id >= s->n) return s->fail("descriptor id out of range");
  const size_t R = s->p.num_ring, S = s->p.num_sector;
  cudaError_t ce = cudaSuccess;
  if (desc && ce == cudaSuccess) ce = cudaMemcpyAsync(desc, s->desc + (size_t)id * R * S, R * S * 8, cudaMemcpyDeviceToHost, s->stream);
  if (ring && ce == cudaSuccess) ce = cudaMemcpyAsync(ring, s->ring + (size_t)id * R, R * 8, cudaMemcpyDeviceToHost, s->stream);
  if (sector && ce == cudaSuccess) ce = cudaMemcpyAsync(sector, s->sector + (size_t)id * S, S * 8, cudaMemcpyDeviceToHost, s->stream);
  if (invkey && ce == cudaSuccess) ce = cudaMemcpyAsync(invkey, s->invkey + (size_t)id * R, R * 4, cudaMemcpyDeviceToHost, s->stream);
  if (ce != cudaSuccess || cudaStreamSynchronize(s->stream) != cudaSuccess) return s->fail("copy failed");
  return 0;
}

int ndtb_sc_distance(ndtb_sc* s, int n_pairs, const int* id1, const int* id2, double* dist, int* shift) {
  if (cudaSetDevice(s->device) != cudaSuccess) return s->fail("cudaSetDevice failed");
  if (n_pairs <= 0) return 0;
  for (int i = 0; i < n_pairs; ++i)
    if (id1[i] < 0 || id2[i] < 0 || (size_t)id1[i] >= s->n || (size_t)id2[i] >= s->n) return s->fail("descriptor id out of range");
  const size_t np = (size_t)n_pairs;
  if (s->ensure_scratch(np * (4 + 4 + 8 + 4) + 64)) return 1;
  char* base = static_cast<char*>(s->scratch);
  double* d_dist = reinterpret_cast<double*>(base);
  int* d_id1 = reinterpret_cast<int*>(base + np * 8);
  int* d_id2 = d_id1 + np;
  int* d_shift = d_id2 + np;
  if (cudaMemcpyAsync(d_id1, id1, np * 4, cudaMemcpyHostToDevice, s->stream) != cudaSuccess ||
      cudaMemcpyAsync(d_id2, id2, np * 4, cudaMemcpyHostToDevice, s->stream) != cudaSuccess) return s->fail("upload failed");
  k_sc_distance<<<n_pairs, 128, 0, s->stream>>>(s->dev(), s->desc, s->sector, d_id1, d_id2, d_dist, d_shift);
  if (cudaGetLastError() != cudaSuccess) return s->fail("distance kernel launch failed");
  if (cudaMemcpyAsync(dist, d_dist, np * 8, cudaMemcpyDeviceToHost, s->stream) != cudaSuccess ||
      cudaMemcpyAsync(shift, d_shift, np * 4, cudaMemcpyDeviceToHost, s->stream) != cudaSuccess ||
      cudaStreamSynchronize(s->stream) != cudaSuccess) return s->fail("distance kernel failed");
  return 0;
}

int ndtb_sc_detect(ndtb_sc* s, int n_queries, const int* query_id, const int* tree_size, int* loop_id, float* yaw_diff_rad,
                   double* min_dist, int* nn_idx) {
  if (cudaSetDevice(s->device) != cudaSuccess) return s->fail("cudaSetDevice failed");
  if (n_queries <= 0) return 0;
  for (int i = 0; i < n_queries; ++i) {
    if (query_id[i] < 0 || (size_t)query_id[i] >= s->n) return s->fail("query id out of range");
    if (tree_size[i] > 0 && (size_t)tree_size[i] > s->n) return s->fail("tree size exceeds the number of descriptors");
  }
  const size_t nq = (size_t)n_queries;
  if (s->ensure_scratch(nq * (8 + 4 * 5) + 64)) return 1;
  char* base = static_cast<char*>(s->scratch);
  double* d_min = reinterpret_cast<double*>(base);
  int* d_q = reinterpret_cast<int*>(base + nq * 8);
  int* d_ts = d_q + nq;
  int* d_loop = d_ts + nq;
  int* d_nn = d_loop + nq;
  float* d_yaw = reinterpret_cast<float*>(d_nn + nq);
  if (cudaMemcpyAsync(d_q, query_id, nq * 4, cudaMemcpyHostToDevice, s->stream) != cudaSuccess ||
      cudaMemcpyAsync(d_ts, tree_size, nq * 4, cudaMemcpyHostToDevice, s->stream) != cudaSuccess) return s->fail("upload failed");
  k_sc_detect<<<n_queries, 128, 0, s->stream>>>(s->dev(), s->desc, s->sector, s->invkey, d_q, d_ts, d_loop, d_yaw, d_min, d_nn);
  if (cudaGetLastError() != cudaSuccess) return s->fail("detect kernel launch failed");
  cudaError_t ce = cudaMemcpyAsync(loop_id, d_loop, nq * 4, cudaMemcpyDeviceToHost, s->stream);
  if (ce == cudaSuccess) ce = cudaMemcpyAsync(yaw_diff_rad, d_yaw, nq * 4, cudaMemcpyDeviceToHost, s->stream);
  if (ce == cudaSuccess && min_dist) ce = cudaMemcpyAsync(min_dist, d_min, nq * 8, cudaMemcpyDeviceToHost, s->stream);
  if (ce == cudaSuccess && nn_idx) ce = cudaMemcpyAsync(nn_idx, d_nn, nq * 4, cudaMemcpyDeviceToHost, s->stream);
  if (ce != cudaSuccess || cudaStreamSynchronize(s->stream) != cudaSuccess) return s->fail("detect kernel failed");
  return 0;
}

}  // extern "C"
