// voxel_build.cu -- target VoxelGridCovariance build on the GPU (K1/K2 of SURVEY.md 2.3).
//
// Replaces pclomp::VoxelGridCovariance::applyFilter (reference
// xchu_mapping/include/pclomp/voxel_grid_covariance_omp_impl.hpp:48-370, VGC_I below):
//   pass 1 (VGC_I:209-264, std::map insert per point)  ->  packed-key radix sort + run detection,
//          then ONE thread per voxel walks its points in input order (the sort is stable), forming
//          exactly the sequential f64 / f32 sums the reference forms, so the single-pass covariance
//          (which cancels catastrophically, VGC_I:329) sees the same rounded sums;
//   pass 2 (VGC_I:282-367) -> per-voxel finalise: mean, covariance, symmetric 3x3 eigen
//          decomposition, eigenvalue inflation, inverse, validity flags, written to a 64 B record
//          and an open-addressing hash table keyed by the absolute cell coordinate.
// Appending points (incremental rebuild on localmap refresh) runs the same kernels on the new
// points only and continues the stored running sums of the voxels they touch; the result is
// bit-identical to a full rebuild over the concatenated cloud.
#include <cub/cub.cuh>

#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstring>

#include "ndt_internal.h"

namespace ndtb {

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return e_; } while (0)

struct BuildWorkspace {
  size_t cap = 0;
  unsigned long long *keys_in = nullptr, *keys_out = nullptr;
  unsigned int *idx_in = nullptr, *idx_out = nullptr;
  int *flags = nullptr, *scan = nullptr, *run_start = nullptr, *run_vidx = nullptr, *run_slot = nullptr,
      *run_isnew = nullptr, *new_prefix = nullptr;
  void* cub_temp = nullptr;
  size_t cub_bytes = 0;
  float* mm = nullptr;      // device: 6 floats (min xyz, max xyz) + [6] = non-finite points seen (int)
  int* counts = nullptr;    // device: [0] = num_runs, [1] = n_new, [2] = largest voxel population, [3] = voxels above the warp kernel's capacity
  float* h_mm = nullptr;    // pinned
  int* h_counts = nullptr;  // pinned
  // counting-sort build (table_build_counting): exclusive scan over the cells of the dense index
  unsigned long long* cellscan = nullptr;
  size_t cellscan_cap = 0;
  void* scan_temp = nullptr;
  size_t scan_temp_bytes = 0;
  bool mm_ready = false;          // mm holds the identity of min / max (k_minmax_init ran since the last read)
  float4* finite_pts = nullptr;   // the cloud without its non-finite points (only when it has any)
  size_t finite_cap = 0;
};

BuildWorkspace* build_ws_create() {
  BuildWorkspace* ws = new BuildWorkspace();
  if (cudaMalloc(&ws->mm, 8 * sizeof(float)) != cudaSuccess || cudaMalloc(&ws->counts, 4 * sizeof(int)) != cudaSuccess ||
      cudaMallocHost(&ws->h_mm, 8 * sizeof(float)) != cudaSuccess || cudaMallocHost(&ws->h_counts, 4 * sizeof(int)) != cudaSuccess) {
    delete ws;
    return nullptr;
  }
  return ws;
}
static void ws_free_arrays(BuildWorkspace* ws) {
  cudaFree(ws->keys_in); cudaFree(ws->keys_out); cudaFree(ws->idx_in); cudaFree(ws->idx_out);
  cudaFree(ws->flags); cudaFree(ws->scan); cudaFree(ws->run_start); cudaFree(ws->run_vidx); cudaFree(ws->run_slot);
  cudaFree(ws->run_isnew); cudaFree(ws->new_prefix); cudaFree(ws->cub_temp);
  ws->keys_in = ws->keys_out = nullptr; ws->idx_in = ws->idx_out = nullptr;
  ws->flags = ws->scan = ws->run_start = ws->run_vidx = ws->run_slot = ws->run_isnew = ws->new_prefix = nullptr;
  ws->cub_temp = nullptr; ws->cub_bytes = 0; ws->cap = 0;
}
void build_ws_destroy(BuildWorkspace* ws) {
  if (!ws) return;
  ws_free_arrays(ws);
  cudaFree(ws->mm); cudaFree(ws->counts); cudaFreeHost(ws->h_mm); cudaFreeHost(ws->h_counts);
  cudaFree(ws->cellscan); cudaFree(ws->scan_temp); cudaFree(ws->finite_pts);
  delete ws;
}
static cudaError_t ws_reserve(BuildWorkspace* ws, size_t n) {
  if (n <= ws->cap) return cudaSuccess;
  ws_free_arrays(ws);
  const size_t cap = n + n / 4 + 1024;
  CK(cudaMalloc(&ws->keys_in, cap * 8)); CK(cudaMalloc(&ws->keys_out, cap * 8));
  CK(cudaMalloc(&ws->idx_in, cap * 4)); CK(cudaMalloc(&ws->idx_out, cap * 4));
  CK(cudaMalloc(&ws->flags, cap * 4)); CK(cudaMalloc(&ws->scan, cap * 4));
  CK(cudaMalloc(&ws->run_start, (cap + 1) * 4)); CK(cudaMalloc(&ws->run_vidx, cap * 4)); CK(cudaMalloc(&ws->run_slot, cap * 4));
  CK(cudaMalloc(&ws->run_isnew, cap * 4)); CK(cudaMalloc(&ws->new_prefix, cap * 4));
  size_t b1 = 0, b2 = 0, b3 = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, b1, ws->keys_in, ws->keys_out, ws->idx_in, ws->idx_out, (int)cap, 0, 64);
  cub::DeviceRadixSort::SortPairs(nullptr, b3, reinterpret_cast<unsigned int*>(ws->keys_in), reinterpret_cast<unsigned int*>(ws->keys_out),
                                  ws->idx_in, ws->idx_out, (int)cap, 0, 32);
  if (b3 > b1) b1 = b3;
  cub::DeviceScan::InclusiveSum(nullptr, b2, ws->flags, ws->scan, (int)cap);
  ws->cub_bytes = (b1 > b2 ? b1 : b2) + 256;
  CK(cudaMalloc(&ws->cub_temp, ws->cub_bytes));
  ws->cap = cap;
  return cudaSuccess;
}

// ---------------------------------------------------------------------------------------
// table storage
// ---------------------------------------------------------------------------------------
void table_free(TableStorage& ts) {
  VoxelTable& t = ts.t;
  cudaFree(t.slots); cudaFree(t.rec); cudaFree(t.frec); cudaFree(t.icov64); cudaFree(t.centroid); cudaFree(t.vkey); cudaFree(t.vn);
  cudaFree(t.vsum); cudaFree(t.vsxx); cudaFree(t.vcsum); cudaFree(ts.dense_buf);
  ts.dense_buf = nullptr; ts.dense_cap = 0; t.dense = nullptr;
  t.slots = nullptr; t.rec = nullptr; t.frec = nullptr; t.icov64 = nullptr; t.centroid = nullptr; t.vkey = nullptr; t.vn = nullptr;
  t.vsum = nullptr; t.vsxx = nullptr; t.vcsum = nullptr;
  ts.vox_cap = ts.slot_cap = 0; ts.n_vox = 0; ts.valid = false; ts.have_bbox = false;
}

__global__ void k_fill_slots(HashSlot* slots, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { HashSlot s; s.key = kEmptyKey; s.vidx = -1; s.nr_points = 0; slots[i] = s; }
}
__global__ void k_rehash(HashSlot* slots, uint32_t mask, const unsigned long long* vkey, const VoxelRec* rec, int n_vox) {
  int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n_vox) return;
  const unsigned long long key = vkey[v];
  uint32_t h = hash_key(key) & mask;
  while (true) {
    unsigned long long prev = atomicCAS(&slots[h].key, kEmptyKey, key);
    if (prev == kEmptyKey || prev == key) { slots[h].vidx = v; slots[h].nr_points = rec[v].nr_points; return; }
    h = (h + 1) & mask;
  }
}

template <typename T>
static cudaError_t grow(T*& p, size_t old_n, size_t new_n, cudaStream_t s) {
  T* q = nullptr;
  CK(cudaMalloc(&q, new_n * sizeof(T)));
  if (p && old_n) CK(cudaMemcpyAsync(q, p, old_n * sizeof(T), cudaMemcpyDeviceToDevice, s));
  CK(cudaStreamSynchronize(s));
  cudaFree(p);
  p = q;
  return cudaSuccess;
}

cudaError_t table_reserve(TableStorage& ts, size_t need, cudaStream_t s, long long* launches) {
  VoxelTable& t = ts.t;
  if (need > ts.vox_cap) {
    const size_t nc = need + need / 2 + 4096;
    const size_t used = (size_t)ts.n_vox;
    CK(grow(t.rec, used, nc, s)); CK(grow(t.frec, used, nc, s)); CK(grow(t.icov64, used * 9, nc * 9, s)); CK(grow(t.centroid, used, nc, s)); CK(grow(t.vkey, used, nc, s)); CK(grow(t.vn, used, nc, s));
    CK(grow(t.vsum, used * 3, nc * 3, s)); CK(grow(t.vsxx, used * 6, nc * 6, s)); CK(grow(t.vcsum, used * 3, nc * 3, s));
    ts.vox_cap = nc;
  }
  size_t want = 1024;
  while (want < need * 4) want <<= 1;   // load factor <= 0.25
  if (want > ts.slot_cap) {
    cudaFree(t.slots);
    t.slots = nullptr;
    CK(cudaMalloc(&t.slots, want * sizeof(HashSlot)));
    ts.slot_cap = want;
    t.mask = (uint32_t)(want - 1);
    k_fill_slots<<<(unsigned)((want + 255) / 256), 256, 0, s>>>(t.slots, want);
    if (launches) ++*launches;
    if (ts.n_vox > 0) {
      k_rehash<<<(ts.n_vox + 255) / 256, 256, 0, s>>>(t.slots, t.mask, t.vkey, t.rec, ts.n_vox);
      if (launches) ++*launches;
    }
    CK(cudaGetLastError());
  }
  return cudaSuccess;
}

// Empties the hash slots and takes the voxels that go away out of the dense cell index (their cells back to -1), in one launch:
// table_build_dense refills the dense array only when its box changes, so a target REPLACED by a cloud with the same bounding
// cells must not find the old cloud's codes in cells the new cloud leaves empty.
__global__ void k_clear_table(VoxelTable t, size_t n_slots, int n_vox_old) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < (size_t)n_vox_old) {
    int a, b, c;
    unpack_key(t.vkey[i], a, b, c);
    t.dense[((size_t)(c - t.dlo[2]) * t.dn[1] + (b - t.dlo[1])) * t.dn[0] + (a - t.dlo[0])] = -1;
  }
  if (i < n_slots) { HashSlot s; s.key = kEmptyKey; s.vidx = -1; s.nr_points = 0; t.slots[i] = s; }
}
cudaError_t table_clear(TableStorage& ts, cudaStream_t s, long long* launches) {
  // the dense index is live (built over every voxel at the end of the last append): un-scatter those voxels; otherwise its
  // contents are not known to be clean, and the next dense build refills it (dn[0] = -1 never equals a real box)
  const int unscatter = ts.t.dense != nullptr && ts.n_vox > 0 ? ts.n_vox : 0;
  if (!unscatter) ts.t.dn[0] = -1;
  const size_t threads = ts.slot_cap > (size_t)unscatter ? ts.slot_cap : (size_t)unscatter;
  if (threads) {
    k_clear_table<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(ts.t, ts.slot_cap, unscatter);
    if (launches) ++*launches;
    CK(cudaGetLastError());
  }
  ts.t.dense = nullptr;
  ts.n_vox = 0;
  ts.valid = false;
  ts.have_bbox = false;
  return cudaSuccess;
}

// ---------------------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------------------
__global__ void k_minmax(const float4* __restrict__ pts, size_t n, float* __restrict__ out6, int first) {
  // single-block-per-grid-stride reduction; launched with a modest grid, finished with atomics on
  // order-preserving integer images of the floats.
  float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  int bad = 0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float4 p = pts[i];
    // non-finite points take no part in the grid (the reference skips them for non-dense clouds: the pcl_isfinite test of
    // VGC_I's first pass and of pcl::getMinMax3D); they are counted so that the caller can compact the cloud
    if (!(isfinite(p.x) && isfinite(p.y) && isfinite(p.z))) { ++bad; continue; }
    mn[0] = fminf(mn[0], p.x); mn[1] = fminf(mn[1], p.y); mn[2] = fminf(mn[2], p.z);
    mx[0] = fmaxf(mx[0], p.x); mx[1] = fmaxf(mx[1], p.y); mx[2] = fmaxf(mx[2], p.z);
  }
  if (__any_sync(0xffffffffu, bad != 0)) {
    for (int o = 16; o > 0; o >>= 1) bad += __shfl_xor_sync(0xffffffffu, bad, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(reinterpret_cast<int*>(out6) + 6, bad);
  }
  for (int a = 0; a < 3; ++a)
    for (int o = 16; o > 0; o >>= 1) {
      mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
      mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
    }
  // warps of the block through shared memory, then ONE set of six atomics per block (a million-point cloud used to end in
  // ~47k atomics on six addresses, which took longer than reading the points)
  __shared__ float s_mn[8][3], s_mx[8][3];
  const int warp = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0)
    for (int a = 0; a < 3; ++a) { s_mn[warp][a] = mn[a]; s_mx[warp][a] = mx[a]; }
  __syncthreads();
  if (threadIdx.x < 3) {
    const int a = threadIdx.x, nw = (blockDim.x + 31) >> 5;
    float lo = s_mn[0][a], hi = s_mx[0][a];
    for (int w = 1; w < nw; ++w) { lo = fminf(lo, s_mn[w][a]); hi = fmaxf(hi, s_mx[w][a]); }
    auto enc = [](float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7FFFFFFF; };
    int* o = reinterpret_cast<int*>(out6);
    atomicMin(&o[a], enc(lo));
    atomicMax(&o[3 + a], enc(hi));
  }
}
// grid for k_minmax (256 threads): enough blocks for the points, at most 4 per SM
static inline unsigned minmax_grid(size_t n) {
  const size_t b = (n + 1023) / 1024;   // >= 4 points per thread
  return (unsigned)(b < 1 ? 1 : (b > 148 * 4 ? 148 * 4 : b));
}
__global__ void k_minmax_init(float* out6) {
  int* o = reinterpret_cast<int*>(out6);
  if (threadIdx.x < 3) o[threadIdx.x] = 0x7FFFFFFF;
  else if (threadIdx.x < 6) o[threadIdx.x] = (int)0x80000000;
  else if (threadIdx.x == 6) o[6] = 0;
}
// Bounding box of the finite points of pts[0 .. n) on the host (blocks the stream once); *n_bad = non-finite points.
// The accumulator is re-armed behind the read, off the critical path, and the order-preserving integer images are decoded here.
static cudaError_t minmax_host(BuildWorkspace* ws, const float4* pts, size_t n, cudaStream_t s, float mn[3], float mx[3], int* n_bad,
                               long long& L) {
  if (!ws->mm_ready) { k_minmax_init<<<1, 32, 0, s>>>(ws->mm); ++L; }
  k_minmax<<<minmax_grid(n), 256, 0, s>>>(pts, n, ws->mm, 1);
  ++L;
  CK(cudaMemcpyAsync(ws->h_mm, ws->mm, 7 * sizeof(float), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  k_minmax_init<<<1, 32, 0, s>>>(ws->mm);
  ++L;
  ws->mm_ready = true;
  const int* h = reinterpret_cast<const int*>(ws->h_mm);
  for (int a = 0; a < 6; ++a) {
    int i = h[a];
    i = i >= 0 ? i : i ^ 0x7FFFFFFF;
    float f;
    memcpy(&f, &i, 4);
    (a < 3 ? mn[a] : mx[a - 3]) = f;
  }
  if (n_bad) *n_bad = h[6];
  return cudaGetLastError();
}

// Build-time cell coordinate: floor(x * inverse_leaf_size) (VGC_I:218-220; the subtraction of
// float(min_b) there is exact for |cell| < 2^24, so the absolute cell is the same integer).
__device__ __forceinline__ int build_cell(float x, float inv_leaf) { return (int)floorf(x * inv_leaf); }

// KeyT: unsigned int when bi + bj + bk <= 32 (sort_key32 below), else unsigned long long: the same integers in
// the same (stable) order, two thirds of the bytes per radix pass.
template <typename KeyT>
__global__ void k_keys(const float4* __restrict__ pts, size_t n, float inv_leaf, int i0, int j0, int k0, int bi, int bj,
                       KeyT* __restrict__ keys, unsigned int* __restrict__ idx) {
  size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const float4 p = pts[t];
  const unsigned long long i = (unsigned long long)(build_cell(p.x, inv_leaf) - i0);
  const unsigned long long j = (unsigned long long)(build_cell(p.y, inv_leaf) - j0);
  const unsigned long long k = (unsigned long long)(build_cell(p.z, inv_leaf) - k0);
  keys[t] = (KeyT)((k << (bi + bj)) | (j << bi) | i);   // ascending == reference linear index order (k, j, i)
  idx[t] = (unsigned int)t;
}
template <typename KeyT>
__global__ void k_flags(const KeyT* __restrict__ keys, size_t n, int* __restrict__ flags) {
  size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  flags[t] = (t == 0 || keys[t] != keys[t - 1]) ? 1 : 0;
}
__global__ void k_run_starts(const int* __restrict__ flags, const int* __restrict__ scan, size_t n, int* __restrict__ run_start,
                             int* __restrict__ counts) {
  size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  if (flags[t]) run_start[scan[t] - 1] = (int)t;
  if (t == n - 1) { run_start[scan[t]] = (int)n; counts[0] = scan[t]; }
}

__device__ __forceinline__ int probe_find(const HashSlot* slots, uint32_t mask, unsigned long long key, int* slot_out) {
  uint32_t h = hash_key(key) & mask;
  while (true) {
    const unsigned long long k = slots[h].key;
    if (k == key) { *slot_out = (int)h; return slots[h].vidx; }
    if (k == kEmptyKey) { *slot_out = -1; return -1; }
    h = (h + 1) & mask;
  }
}

__global__ void k_run_lookup(const float4* __restrict__ pts, const unsigned int* __restrict__ sorted_idx,
                             const int* __restrict__ run_start, int n_runs, VoxelTable t, int* __restrict__ run_vidx,
                             int* __restrict__ run_slot, int* __restrict__ run_isnew) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_runs) return;
  const float4 p = pts[sorted_idx[run_start[r]]];
  const unsigned long long key = pack_key(build_cell(p.x, t.inv_leaf), build_cell(p.y, t.inv_leaf), build_cell(p.z, t.inv_leaf));
  int slot;
  const int v = probe_find(t.slots, t.mask, key, &slot);
  run_vidx[r] = v;
  run_slot[r] = slot;
  run_isnew[r] = v < 0 ? 1 : 0;
}

// ---- per-voxel finalisation (VGC_I:282-367) ---------------------------------------------
struct Mat3 { double a[3][3]; };

__device__ __forceinline__ double cof3(const Mat3& m, int i, int j) {
  const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
  return m.a[i1][j1] * m.a[i2][j2] - m.a[i1][j2] * m.a[i2][j1];
}
// general 3x3 inverse by cofactors (what Matrix3d::inverse() evaluates, VGC_I:355, 359)
__device__ void inverse3(const Mat3& m, Mat3& r) {
  const double c00 = cof3(m, 0, 0), c10 = cof3(m, 1, 0), c20 = cof3(m, 2, 0);
  const double det = (c00 * m.a[0][0] + c10 * m.a[1][0]) + c20 * m.a[2][0];
  const double invdet = 1.0 / det;
  r.a[0][0] = c00 * invdet; r.a[0][1] = c10 * invdet; r.a[0][2] = c20 * invdet;
  for (int j = 1; j < 3; ++j)
    for (int i = 0; i < 3; ++i) r.a[j][i] = cof3(m, i, j) * invdet;
}
__device__ void mul3(const Mat3& x, const Mat3& y, Mat3& r) {
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) r.a[i][j] = (x.a[i][0] * y.a[0][j] + x.a[i][1] * y.a[1][j]) + x.a[i][2] * y.a[2][j];
}

// Symmetric 3x3 eigen-decomposition of the lower triangle of `in` by cyclic Jacobi rotations
// (each rotation is a closed-form 2x2 diagonalisation; 3-5 sweeps reach full double accuracy).
// A one-shot trigonometric closed form was rejected: the parity bar on icov is 1e-10 relative and
// the trigonometric eigenvectors lose that for the near-degenerate (planar / linear) voxels that
// dominate LiDAR maps.  Eigenvalues ascending, eigenvectors in the columns of V.
__device__ void eigh3(const Mat3& in, double w[3], Mat3& V) {
  double A[3][3];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j <= i; ++j) A[i][j] = A[j][i] = in.a[i][j];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) V.a[i][j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 32; ++sweep) {
    const double off = A[0][1] * A[0][1] + A[0][2] * A[0][2] + A[1][2] * A[1][2];
    const double dia = A[0][0] * A[0][0] + A[1][1] * A[1][1] + A[2][2] * A[2][2];
    if (off == 0.0 || off <= 1e-34 * dia) break;
#pragma unroll
    for (int pq = 0; pq < 3; ++pq) {
      const int p = pq == 2 ? 1 : 0, q = pq == 0 ? 1 : 2;
      const double apq = A[p][q];
      if (apq == 0.0) continue;
      const double theta = (A[q][q] - A[p][p]) / (2.0 * apq);
      const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
      const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
      for (int k = 0; k < 3; ++k) { const double x = A[k][p], y = A[k][q]; A[k][p] = c * x - s * y; A[k][q] = s * x + c * y; }
      for (int k = 0; k < 3; ++k) { const double x = A[p][k], y = A[q][k]; A[p][k] = c * x - s * y; A[q][k] = s * x + c * y; }
      for (int k = 0; k < 3; ++k) { const double x = V.a[k][p], y = V.a[k][q]; V.a[k][p] = c * x - s * y; V.a[k][q] = s * x + c * y; }
    }
  }
  // ascending order (3-element sorting network on the diagonal)
  double d[3] = {A[0][0], A[1][1], A[2][2]};
  int o[3] = {0, 1, 2};
  if (d[o[1]] < d[o[0]]) { int x = o[0]; o[0] = o[1]; o[1] = x; }
  if (d[o[2]] < d[o[1]]) { int x = o[1]; o[1] = o[2]; o[2] = x; }
  if (d[o[1]] < d[o[0]]) { int x = o[0]; o[0] = o[1]; o[1] = x; }
  Mat3 Vs;
  for (int k = 0; k < 3; ++k) {
    w[k] = d[o[k]];
    for (int i = 0; i < 3; ++i) Vs.a[i][k] = V.a[i][o[k]];
  }
  V = Vs;
}

__device__ void finalize_voxel(const VoxelTable& t, int v, int n, const double sum[3], const double sxx[6], const float csum[3],
                               int slot) {
  VoxelRec rec;
  const float nf = (float)n;
  float4 c;
  c.x = csum[0] / nf; c.y = csum[1] / nf; c.z = csum[2] / nf; c.w = __int_as_float(n);
  double mean[3];
  for (int a = 0; a < 3; ++a) mean[a] = sum[a] / (double)n;
  for (int a = 0; a < 3; ++a) rec.mean[a] = mean[a];
  for (int a = 0; a < 9; ++a) rec.icov[a] = 0.f;
  double ic64[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  int nr = n;
  if (n >= t.min_points) {
    Mat3 cov;
    const int sidx[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b)
        cov.a[a][b] = (sxx[sidx[a][b]] - 2 * (sum[a] * mean[b])) / (double)n + mean[a] * mean[b];  // VGC_I:329
    const double f = ((double)n - 1.0) / (double)n;                                                 // VGC_I:330
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b) cov.a[a][b] *= f;
    double w[3];
    Mat3 V;
    eigh3(cov, w, V);
    if (w[0] < 0 || w[1] < 0 || w[2] <= 0) {
      nr = -1;  // VGC_I:337-341: icov_ stays Zero
    } else {
      const double min_ev = t.eig_mult * w[2];
      if (w[0] < min_ev) {  // VGC_I:346-356
        w[0] = min_ev;
        if (w[1] < min_ev) w[1] = min_ev;
        Mat3 D, VD, Vi;
        for (int a = 0; a < 3; ++a)
          for (int b = 0; b < 3; ++b) D.a[a][b] = (a == b) ? w[a] : 0.0;
        mul3(V, D, VD);
        inverse3(V, Vi);
        mul3(VD, Vi, cov);
      }
      Mat3 icov;
      inverse3(cov, icov);
      double mxc = -INFINITY, mnc = INFINITY;
      for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) { mxc = fmax(mxc, icov.a[a][b]); mnc = fmin(mnc, icov.a[a][b]); rec.icov[a * 3 + b] = (float)icov.a[a][b]; ic64[a * 3 + b] = icov.a[a][b]; }
      // fmax/fmin ignore NaN like Eigen's maxCoeff does not; the +-inf test is what VGC_I:360-364 checks
      if (mxc == INFINITY || mnc == -INFINITY) nr = -1;
    }
  }
  rec.nr_points = nr;
  t.rec[v] = rec;
  FastRec fr;
  for (int a = 0; a < 3; ++a) {
    fr.mh[a] = (float)mean[a];
    fr.ml[a] = (float)(mean[a] - (double)fr.mh[a]);
  }
  fr.c[0] = rec.icov[0]; fr.c[1] = rec.icov[1]; fr.c[2] = rec.icov[2]; fr.c[3] = rec.icov[4]; fr.c[4] = rec.icov[5]; fr.c[5] = rec.icov[8];
  t.frec[v] = fr;
  for (int a = 0; a < 9; ++a) t.icov64[(size_t)v * 9 + a] = ic64[a];
  t.centroid[v] = c;
  t.slots[slot].nr_points = nr;
}

// FRESH: the table was empty before this append, so every run is a new voxel and run r is voxel r -- no k_run_lookup, no
// prefix scan, no count to read back (run_vidx / run_slot / new_prefix are not read).
template <bool FRESH>
__global__ void k_run_merge(const float4* __restrict__ pts, const unsigned int* __restrict__ sorted_idx,
                            const int* __restrict__ run_start, const int* __restrict__ run_vidx,
                            const int* __restrict__ run_slot, const int* __restrict__ new_prefix, int n_runs, int n_vox_old,
                            VoxelTable t, int* __restrict__ counts) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_runs) return;
  const int s0 = run_start[r], s1 = run_start[r + 1];
  int v = FRESH ? -1 : run_vidx[r];
  int slot = FRESH ? 0 : run_slot[r];
  double sum[3] = {0, 0, 0};
  double sxx[6] = {1.0, 0, 0, 1.0, 0, 1.0};  // Leaf::cov_ starts at Identity (VGC_H:101) and is accumulated onto
  float cs[3] = {0.f, 0.f, 0.f};
  int n = 0;
  if (v >= 0) {
    n = t.vn[v];
    for (int a = 0; a < 3; ++a) { sum[a] = t.vsum[(size_t)v * 3 + a]; cs[a] = t.vcsum[(size_t)v * 3 + a]; }
    for (int a = 0; a < 6; ++a) sxx[a] = t.vsxx[(size_t)v * 6 + a];
  } else {
    v = FRESH ? r : n_vox_old + new_prefix[r];
    const float4 p = pts[sorted_idx[s0]];
    const unsigned long long key = pack_key(build_cell(p.x, t.inv_leaf), build_cell(p.y, t.inv_leaf), build_cell(p.z, t.inv_leaf));
    t.vkey[v] = key;
    uint32_t h = hash_key(key) & t.mask;
    while (true) {
      const unsigned long long prev = atomicCAS(&t.slots[h].key, kEmptyKey, key);
      if (prev == kEmptyKey) break;
      h = (h + 1) & t.mask;
    }
    t.slots[h].vidx = v;
    slot = (int)h;
    if (!FRESH && r == n_runs - 1) counts[1] = new_prefix[r] + 1;
  }
  if (!FRESH && r == n_runs - 1 && run_vidx[r] >= 0) counts[1] = new_prefix[r];
  for (int s = s0; s < s1; ++s) {  // input order within the voxel (stable sort) == reference order
    const float4 p = pts[sorted_idx[s]];
    const double x = (double)p.x, y = (double)p.y, z = (double)p.z;
    sum[0] += x; sum[1] += y; sum[2] += z;
    sxx[0] += x * x; sxx[1] += x * y; sxx[2] += x * z; sxx[3] += y * y; sxx[4] += y * z; sxx[5] += z * z;
    cs[0] += p.x; cs[1] += p.y; cs[2] += p.z;
  }
  n += s1 - s0;
  t.vn[v] = n;
  for (int a = 0; a < 3; ++a) { t.vsum[(size_t)v * 3 + a] = sum[a]; t.vcsum[(size_t)v * 3 + a] = cs[a]; }
  for (int a = 0; a < 6; ++a) t.vsxx[(size_t)v * 6 + a] = sxx[a];
  finalize_voxel(t, v, n, sum, sxx, cs, slot);
}

// ---- dense cell index (VoxelTable::dense) -------------------------------------------------
__global__ void k_dense_fill(int* __restrict__ d, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) d[i] = -1;
}
__global__ void k_dense_scatter(VoxelTable t, int n_vox) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n_vox) return;
  int i, j, k;
  unpack_key(t.vkey[v], i, j, k);
  const int nr = t.rec[v].nr_points;
  int code = -1;
  if (nr >= t.min_points) code = v;
  else if (nr == -1 && t.vn[v] >= t.min_points) code = -2 - v;
  t.dense[((size_t)(k - t.dlo[2]) * t.dn[1] + (j - t.dlo[1])) * t.dn[0] + (i - t.dlo[0])] = code;
}
// (Re)build the dense index over [min_b - 2, max_b + 2] when that box holds <= dense_max_cells cells.
// Every voxel is re-scattered (V is small next to the point count); the -1 fill only runs when the box changed.
static cudaError_t table_build_dense(TableStorage& ts, cudaStream_t s, long long* launches) {
  VoxelTable& t = ts.t;
  t.dense = nullptr;
  if (!ts.valid || ts.n_vox <= 0 || ts.dense_max_cells <= 0) return cudaSuccess;
  int lo[3], dn[3];
  long long cells = 1;
  for (int a = 0; a < 3; ++a) {
    lo[a] = ts.min_b[a] - 2;
    dn[a] = ts.max_b[a] - ts.min_b[a] + 5;
    cells *= dn[a];
    if (cells > ts.dense_max_cells) return cudaSuccess;
  }
  bool refill = false;
  if ((size_t)cells > ts.dense_cap) {
    cudaFree(ts.dense_buf);
    ts.dense_buf = nullptr;
    ts.dense_cap = 0;
    const size_t cap = (size_t)cells + (size_t)cells / 8;
    CK(cudaMalloc(&ts.dense_buf, cap * sizeof(int)));
    ts.dense_cap = cap;
    refill = true;
  }
  for (int a = 0; a < 3; ++a)
    if (lo[a] != t.dlo[a] || dn[a] != t.dn[a]) refill = true;
  for (int a = 0; a < 3; ++a) { t.dlo[a] = lo[a]; t.dn[a] = dn[a]; }
  t.dense = ts.dense_buf;
  if (refill) {
    k_dense_fill<<<148 * 8, 256, 0, s>>>(t.dense, (size_t)cells);
    if (launches) ++*launches;
  }
  k_dense_scatter<<<(ts.n_vox + 255) / 256, 256, 0, s>>>(t, ts.n_vox);
  if (launches) ++*launches;
  return cudaGetLastError();
}

static int bits_for(long long range) {
  int b = 1;
  while ((1ll << b) < range) ++b;
  return b;
}

// Cell keys of at most 32 bits (any submap: 21 bits at 1M points / 1 m) are generated, sorted and compared as 32-bit words in
// the lower halves of the same workspace arrays: two thirds of the bytes per radix pass (64 builds: 5.3 -> 4.7 ms, GPU suite
// green with it: profiles/r02).  NDTB_SORT_KEY32=0 forces 64-bit keys.
static bool sort_key32(int bits) {
  static const bool on = [] { const char* e = getenv("NDTB_SORT_KEY32"); return !(e && atoi(e) == 0); }();
  return on && bits <= 32;
}
// keys_in / idx_in -> keys_out / idx_out (stable), then flags[t] = key t starts a run
static cudaError_t sort_and_flag(BuildWorkspace* ws, size_t n, int bits, cudaStream_t s) {
  size_t tb = ws->cub_bytes;
  const unsigned nb = (unsigned)((n + 255) / 256);
  if (sort_key32(bits)) {
    unsigned int *ki = reinterpret_cast<unsigned int*>(ws->keys_in), *ko = reinterpret_cast<unsigned int*>(ws->keys_out);
    CK(cub::DeviceRadixSort::SortPairs(ws->cub_temp, tb, ki, ko, ws->idx_in, ws->idx_out, (int)n, 0, bits, s));
    k_flags<unsigned int><<<nb, 256, 0, s>>>(ko, n, ws->flags);
  } else {
    CK(cub::DeviceRadixSort::SortPairs(ws->cub_temp, tb, ws->keys_in, ws->keys_out, ws->idx_in, ws->idx_out, (int)n, 0, bits, s));
    k_flags<unsigned long long><<<nb, 256, 0, s>>>(ws->keys_out, n, ws->flags);
  }
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Counting-sort build of an EMPTY table over the dense cell grid (the full rebuild of a target: every setInputTarget of the
// batch path and of the keyframe-count refresh strategy).  The radix sort of (64-bit key, index) pairs -- two thirds of the
// old build's device time -- is replaced by: cell histogram (atomics on the dense array, which is L2-resident) -> exclusive
// scan over the cells (cell -> first slot AND cell -> voxel number, one packed 64-bit sum; voxels come out in ascending
// (k, j, i) order, the reference's leaf order) -> scatter of the point indices.  The scatter leaves the indices of a voxel in
// arbitrary order; the sequential f64 / f32 sums of the reference (VGC_I:233-243) need INPUT order, so the merge kernel
// sorts each voxel's handful of indices (a warp per voxel, bitonic network in shared memory), gathers the points a warp-wide
// tile at a time and lets twelve lanes carry the twelve running sums (3 + 6 doubles, 3 floats) in exactly the reference's
// order of additions.  Finalisation (eigen-decomposition, inverse, hash insert, dense code) is a thread per voxel.
// ---------------------------------------------------------------------------------------
constexpr int kWarpVoxelCap = 512;    // indices a warp sorts in its shared-memory slice; larger voxels go to k_cs_merge_big
constexpr int kBigVoxelCap = 8192;    // larger than this: fall back to the radix-sort build (nothing to sort per voxel there)

struct CellScanIn {
  __host__ __device__ __forceinline__ unsigned long long operator()(const int& c) const {
    return (unsigned long long)(unsigned)c | (c > 0 ? (1ull << 32) : 0ull);
  }
};

// cell of every point (dense-grid linear index; -1: outside the box, i.e. not finite) and the per-cell populations
__global__ void k_cs_hist(const float4* __restrict__ pts, size_t n, float inv_leaf, int lo0, int lo1, int lo2, int dn0, int dn1, int dn2,
                          int* __restrict__ cell_of, int* __restrict__ dense) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const float4 p = pts[t];
  const int a = build_cell(p.x, inv_leaf) - lo0, b = build_cell(p.y, inv_leaf) - lo1, c = build_cell(p.z, inv_leaf) - lo2;
  int cell = -1;
  if ((unsigned)a < (unsigned)dn0 && (unsigned)b < (unsigned)dn1 && (unsigned)c < (unsigned)dn2 && isfinite(p.x) && isfinite(p.y) && isfinite(p.z)) {
    cell = (c * dn1 + b) * dn0 + a;
    atomicAdd(dense + cell, 1);
  }
  cell_of[t] = cell;
}
// per occupied cell: voxel number, first slot, population; the dense entry becomes the scatter cursor (occupied) or -1 (final)
__global__ void k_cs_cells(int* __restrict__ dense, const unsigned long long* __restrict__ scan, size_t cells, int* __restrict__ v_start,
                           int* __restrict__ v_cnt, int* __restrict__ v_cell, int* __restrict__ counts) {
  const size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = c < cells;       // (no early return: the warp-wide reductions below need every lane)
  const int cnt = live ? dense[c] : 0;
  const unsigned long long e = live ? scan[c] : 0ull;
  const int start = (int)(unsigned)(e & 0xFFFFFFFFull), v = (int)(unsigned)(e >> 32);
  int mx = 0;
  if (cnt > 0) {
    v_start[v] = start; v_cnt[v] = cnt; v_cell[v] = (int)c;
    dense[c] = start;
    mx = cnt;
  } else if (live) {
    dense[c] = -1;
  }
  if (live && c == cells - 1) counts[0] = v + (cnt > 0 ? 1 : 0);
  // largest population / voxels beyond the warp kernel's capacity: one atomic per warp that has something to report
  const int wmx = __reduce_max_sync(0xffffffffu, mx);
  const unsigned big = __ballot_sync(0xffffffffu, cnt > kWarpVoxelCap);
  if ((threadIdx.x & 31) == 0) {
    if (wmx > kWarpVoxelCap / 2) atomicMax(counts + 2, wmx);
    if (big) atomicAdd(counts + 3, __popc(big));
  }
}
__global__ void k_cs_scatter(const int* __restrict__ cell_of, size_t n, int* __restrict__ dense, unsigned int* __restrict__ sorted_idx) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int cell = cell_of[t];
  if (cell < 0) return;
  sorted_idx[atomicAdd(dense + cell, 1)] = (unsigned int)t;
}

// the twelve running sums of a voxel: lanes 0-2 sum x (f64), 3-8 Identity + sum xx^T (f64; xx xy xz yy yz zz), 9-11 f32 centroid sum
struct VoxelSums {
  double d;   // lanes 0..8
  float f;    // lanes 9..11
};
__device__ __forceinline__ void sums_init(VoxelSums& a, int lane) {
  a.d = (lane == 3 || lane == 6 || lane == 8) ? 1.0 : 0.0;   // Leaf::cov_ starts at Identity (VGC_H:101)
  a.f = 0.f;
}
// The lane's two factors out of (x, y, z, 1), chosen once: the product of two floats is exact in double, so x * 1.0 is x and
// every lane performs exactly the reference's addition for its component (mean_ += pt3d; cov_ += pt3d * pt3d^T; centroid += pt).
__device__ __forceinline__ void sums_roles(int lane, int& ia, int& ib) {
  const int ra[12] = {0, 1, 2, 0, 0, 0, 1, 1, 2, 0, 1, 2}, rb[12] = {3, 3, 3, 0, 1, 2, 1, 2, 2, 3, 3, 3};
  ia = 0; ib = 3;
#pragma unroll
  for (int k = 0; k < 12; ++k)
    if (lane == k) { ia = ra[k]; ib = rb[k]; }
}
__device__ __forceinline__ float pick4(const float4 p, int i) { return i == 0 ? p.x : (i == 1 ? p.y : (i == 2 ? p.z : 1.f)); }
// Both chains advance in every lane (a lane keeps the one its role names): no divergent branch inside the per-point loop.
__device__ __forceinline__ void sums_add(VoxelSums& a, int ia, int ib, const float4 p) {
  const float fa = pick4(p, ia), fb = pick4(p, ib);
  a.d = __dadd_rn(a.d, __dmul_rn((double)fa, (double)fb));
  a.f = __fadd_rn(a.f, fa);
}
__device__ __forceinline__ void sums_store(const VoxelSums& a, int lane, const VoxelTable& t, int v) {
  if (lane < 3) t.vsum[(size_t)v * 3 + lane] = a.d;
  else if (lane < 9) t.vsxx[(size_t)v * 6 + (lane - 3)] = a.d;
  else if (lane < 12) t.vcsum[(size_t)v * 3 + (lane - 9)] = a.f;
}
__device__ __forceinline__ unsigned long long cell_to_key(int cell, const VoxelTable& t) {
  const int a = cell % t.dn[0], r = cell / t.dn[0];
  const int b = r % t.dn[1], c = r / t.dn[1];
  return pack_key(a + t.dlo[0], b + t.dlo[1], c + t.dlo[2]);
}

// Bitonic sort of R * 32 keys held in registers, key (r * 32 + lane) in v[r], ascending.  Exchanges at a distance of 32 or more
// stay inside a lane (register to register, direction known at compile time); closer ones are one shuffle + min / max.
template <int R>
__device__ __forceinline__ void warp_bitonic(unsigned int (&v)[R], int lane) {
  constexpr int M = R * 32;
#pragma unroll
  for (int size = 2; size <= M; size <<= 1) {
#pragma unroll
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      if (stride >= 32) {
        const int rs = stride >> 5;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          if ((r & rs) == 0) {
            const bool up = ((r << 5) & size) == 0;          // size >= 64: the bit is one of r's
            const unsigned int a = v[r], b = v[r | rs];
            const unsigned int mn = min(a, b), mx = max(a, b);
            v[r] = up ? mn : mx;
            v[r | rs] = up ? mx : mn;
          }
        }
      } else {
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const unsigned int other = __shfl_xor_sync(0xffffffffu, v[r], stride);
          const bool up = ((((r << 5) + lane) & size) == 0);
          const bool lower = (lane & stride) == 0;
          const unsigned int mn = min(v[r], other), mx = max(v[r], other);
          v[r] = (lower == up) ? mn : mx;
        }
      }
    }
  }
}

constexpr int kSumRow = 33;     // doubles per row of a warp's staging tile: rows land in different banks for the role-indexed reads
struct MergeTile {
  double d[4][kSumRow];         // x, y, z of 32 points widened to f64 (once per point, not once per sum), and a row of ones
  float f[3][kSumRow];          // x, y, z as they are, for the f32 centroid sums
};

// One voxel of up to R * 32 points by one warp: indices sorted in registers, points fetched a warp-wide tile at a time (the next
// tile in flight behind the current one's additions), twelve lanes carrying the twelve running sums in the reference's order.
template <int R>
__device__ __forceinline__ void merge_voxel(const float4* __restrict__ pts, unsigned int* __restrict__ sorted_idx, int start, int n, int lane,
                                            MergeTile& tile, int ia, int ib, int fr, double& sum_d, float& sum_f) {
  unsigned int v[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int k = (r << 5) + lane;
    v[r] = k < n ? sorted_idx[start + k] : 0xFFFFFFFFu;
  }
  warp_bitonic<R>(v, lane);
  float4 nxt = make_float4(0.f, 0.f, 0.f, 0.f);
  if (lane < n) nxt = __ldg(pts + v[0]);
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int base = r << 5;
    if (base < n) {                                   // warp-uniform
      const float4 p = nxt;
      if (r + 1 < R && base + 32 + lane < n) nxt = __ldg(pts + v[r + 1 < R ? r + 1 : r]);
      if (base + lane < n) sorted_idx[start + base + lane] = v[r];     // later readers of the voxel's points (appends, dumps) want input order
      tile.d[0][lane] = (double)p.x; tile.d[1][lane] = (double)p.y; tile.d[2][lane] = (double)p.z;
      tile.f[0][lane] = p.x; tile.f[1][lane] = p.y; tile.f[2][lane] = p.z;
      __syncwarp();
      const int cnt = min(32, n - base);
      if (lane < 12) {
        const double* da = tile.d[ia];
        const double* db = tile.d[ib];
        const float* fa = tile.f[fr];
#pragma unroll 4
        for (int k = 0; k < cnt; ++k) {
          sum_d = __dadd_rn(sum_d, __dmul_rn(da[k], db[k]));   // the product of two widened floats is exact; x * 1.0 is x
          sum_f = __fadd_rn(sum_f, fa[k]);
        }
      }
      __syncwarp();
    }
  }
}

// warp per voxel (population <= kWarpVoxelCap)
__global__ void __launch_bounds__(128) k_cs_merge(const float4* __restrict__ pts, unsigned int* __restrict__ sorted_idx, const int* __restrict__ v_start,
                                                  const int* __restrict__ v_cnt, int n_vox, VoxelTable t, int* __restrict__ big_list,
                                                  int* __restrict__ big_count) {
  __shared__ MergeTile s_tile[4];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int v = blockIdx.x * 4 + warp;
  if (v >= n_vox) return;
  const int start = v_start[v], n = v_cnt[v];
  if (n > kWarpVoxelCap) {
    if (lane == 0) big_list[atomicAdd(big_count, 1)] = v;
    return;
  }
  MergeTile& tile = s_tile[warp];
  tile.d[3][lane] = 1.0;
  // lanes 0-2: sum x (x_a * 1); 3-8: Identity + sum xx^T (xx xy xz yy yz zz); 9-11: f32 centroid sum (their f64 chain is idle work)
  int ia, ib;
  sums_roles(lane, ia, ib);
  const int fr = (lane >= 9 && lane < 12) ? lane - 9 : 0;
  VoxelSums acc;
  sums_init(acc, lane);
  if (n <= 32) merge_voxel<1>(pts, sorted_idx, start, n, lane, tile, ia, ib, fr, acc.d, acc.f);
  else if (n <= 64) merge_voxel<2>(pts, sorted_idx, start, n, lane, tile, ia, ib, fr, acc.d, acc.f);
  else if (n <= 128) merge_voxel<4>(pts, sorted_idx, start, n, lane, tile, ia, ib, fr, acc.d, acc.f);
  else if (n <= 256) merge_voxel<8>(pts, sorted_idx, start, n, lane, tile, ia, ib, fr, acc.d, acc.f);
  else merge_voxel<16>(pts, sorted_idx, start, n, lane, tile, ia, ib, fr, acc.d, acc.f);
  sums_store(acc, lane, t, v);
}
// block per voxel with kWarpVoxelCap < population <= kBigVoxelCap: rank-by-counting sort into `scratch`, then the same sums
__global__ void __launch_bounds__(256) k_cs_merge_big(const float4* __restrict__ pts, unsigned int* __restrict__ sorted_idx, unsigned int* __restrict__ scratch,
                                                      const int* __restrict__ v_start, const int* __restrict__ v_cnt, const int* __restrict__ big_list,
                                                      const int* __restrict__ big_count, VoxelTable t) {
  __shared__ unsigned int s_tile[1024];
  __shared__ float4 s_pts[256];
  const int n_big = *big_count;
  for (int b = blockIdx.x; b < n_big; b += gridDim.x) {
    const int v = big_list[b];
    const int start = v_start[v], n = v_cnt[v];
    const unsigned int* seg = sorted_idx + start;
    unsigned int* out = scratch + start;
    // rank of every element = number of smaller elements (indices are distinct)
    for (int e0 = 0; e0 < n; e0 += 256) {
      const int e = e0 + threadIdx.x;
      const unsigned int mine = e < n ? seg[e] : 0u;
      int rank = 0;
      for (int j0 = 0; j0 < n; j0 += 1024) {
        __syncthreads();
        for (int j = threadIdx.x; j < 1024; j += 256) s_tile[j] = j0 + j < n ? seg[j0 + j] : 0xFFFFFFFFu;
        __syncthreads();
        const int lim = min(1024, n - j0);
        for (int j = 0; j < lim; ++j) rank += s_tile[j] < mine ? 1 : 0;
      }
      if (e < n) out[rank] = mine;
    }
    __syncthreads();
    VoxelSums acc;
    sums_init(acc, threadIdx.x);
    int ia, ib;
    sums_roles(threadIdx.x, ia, ib);
    for (int base = 0; base < n; base += 256) {
      __syncthreads();
      if (base + threadIdx.x < n) s_pts[threadIdx.x] = __ldg(pts + out[base + threadIdx.x]);
      __syncthreads();
      const int cnt = min(256, n - base);
      if (threadIdx.x < 12)
        for (int k = 0; k < cnt; ++k) sums_add(acc, ia, ib, s_pts[k]);
    }
    if (threadIdx.x < 12) sums_store(acc, threadIdx.x, t, v);
    __syncthreads();
    for (int k = threadIdx.x; k < n; k += 256) sorted_idx[start + k] = out[k];
    __syncthreads();
  }
}
// thread per voxel: hash insert, finalisation (VGC_I:282-367), dense code
__global__ void __launch_bounds__(128) k_cs_finalize(VoxelTable t, int n_vox, const int* __restrict__ v_cell, const int* __restrict__ v_cnt) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n_vox) return;
  const unsigned long long key = cell_to_key(v_cell[v], t);
  t.vkey[v] = key;
  t.vn[v] = v_cnt[v];
  uint32_t h = hash_key(key) & t.mask;
  while (true) {
    const unsigned long long prev = atomicCAS(&t.slots[h].key, kEmptyKey, key);
    if (prev == kEmptyKey) break;
    h = (h + 1) & t.mask;
  }
  t.slots[h].vidx = v;
  const int n = v_cnt[v];
  double sum[3], sxx[6];
  float cs[3];
  for (int a = 0; a < 3; ++a) { sum[a] = t.vsum[(size_t)v * 3 + a]; cs[a] = t.vcsum[(size_t)v * 3 + a]; }
  for (int a = 0; a < 6; ++a) sxx[a] = t.vsxx[(size_t)v * 6 + a];
  finalize_voxel(t, v, n, sum, sxx, cs, (int)h);
  const int nr = t.rec[v].nr_points;
  int code = -1;
  if (nr >= t.min_points) code = v;
  else if (nr == -1 && n >= t.min_points) code = -2 - v;
  t.dense[v_cell[v]] = code;
}

static bool counting_build_enabled() {
  static const bool on = [] { const char* e = getenv("NDTB_BUILD_RADIX"); return !(e && atoi(e) > 0); }();   // NDTB_BUILD_RADIX=1: the radix-sort build only
  return on;
}

// Fresh build over the dense grid.  Preconditions: ts holds no voxel, ts.min_b / max_b / valid are set, hash slots are empty.
// *done = false: the caller builds with the radix-sort path instead (box too large for a dense index, or a voxel too populous
// to sort in shared memory) -- the dense array is then marked for a refill.
static cudaError_t table_build_counting(TableStorage& ts, BuildWorkspace* ws, const float4* pts, size_t n, cudaStream_t s, long long& L,
                                        long long* launches, bool* done) {
  *done = false;
  VoxelTable& t = ts.t;
  if (!counting_build_enabled() || ts.dense_max_cells <= 0 || n >= (size_t)1 << 31) return cudaSuccess;
  int lo[3], dn[3];
  long long cells = 1;
  for (int a = 0; a < 3; ++a) {
    lo[a] = ts.min_b[a] - 2;
    dn[a] = ts.max_b[a] - ts.min_b[a] + 5;
    cells *= dn[a];
    if (cells > ts.dense_max_cells) return cudaSuccess;
  }
  if (cells > 32ll * (long long)n + (1ll << 21) || cells >= (1ll << 31)) return cudaSuccess;   // a scan over mostly empty cells would cost more than the sort
  if ((size_t)cells > ts.dense_cap) {
    cudaFree(ts.dense_buf);
    ts.dense_buf = nullptr;
    ts.dense_cap = 0;
    const size_t cap = (size_t)cells + (size_t)cells / 8;
    CK(cudaMalloc(&ts.dense_buf, cap * sizeof(int)));
    ts.dense_cap = cap;
  }
  if ((size_t)cells > ws->cellscan_cap) {
    cudaFree(ws->cellscan);
    ws->cellscan = nullptr;
    ws->cellscan_cap = 0;
    const size_t cap = (size_t)cells + (size_t)cells / 4;
    CK(cudaMalloc(&ws->cellscan, cap * sizeof(unsigned long long)));
    ws->cellscan_cap = cap;
  }
  cub::TransformInputIterator<unsigned long long, CellScanIn, const int*> scan_in(ts.dense_buf, CellScanIn());
  size_t need = 0;
  CK(cub::DeviceScan::ExclusiveSum(nullptr, need, scan_in, ws->cellscan, (int)cells, s));
  if (need > ws->scan_temp_bytes) {
    cudaFree(ws->scan_temp);
    ws->scan_temp = nullptr;
    ws->scan_temp_bytes = 0;
    CK(cudaMalloc(&ws->scan_temp, need + need / 2 + 256));
    ws->scan_temp_bytes = need + need / 2 + 256;
  }
  for (int a = 0; a < 3; ++a) { t.dlo[a] = lo[a]; t.dn[a] = dn[a]; }
  t.dense = ts.dense_buf;
  int* cell_of = reinterpret_cast<int*>(ws->idx_in);
  int *v_start = ws->run_start, *v_cnt = ws->run_vidx, *v_cell = ws->run_slot, *big_list = ws->run_isnew;
  CK(cudaMemsetAsync(t.dense, 0, (size_t)cells * sizeof(int), s));
  CK(cudaMemsetAsync(ws->counts, 0, 4 * sizeof(int), s));
  const unsigned nb = (unsigned)((n + 255) / 256);
  k_cs_hist<<<nb, 256, 0, s>>>(pts, n, t.inv_leaf, lo[0], lo[1], lo[2], dn[0], dn[1], dn[2], cell_of, t.dense);
  size_t tb = ws->scan_temp_bytes;
  CK(cub::DeviceScan::ExclusiveSum(ws->scan_temp, tb, scan_in, ws->cellscan, (int)cells, s));
  k_cs_cells<<<(unsigned)((cells + 255) / 256), 256, 0, s>>>(t.dense, ws->cellscan, (size_t)cells, v_start, v_cnt, v_cell, ws->counts);
  CK(cudaMemcpyAsync(ws->h_counts, ws->counts, 4 * sizeof(int), cudaMemcpyDeviceToHost, s));
  k_cs_scatter<<<nb, 256, 0, s>>>(cell_of, n, t.dense, ws->idx_out);
  L += 8;
  CK(cudaStreamSynchronize(s));
  const int n_vox = ws->h_counts[0], max_pop = ws->h_counts[2], n_big = ws->h_counts[3];
  if (max_pop > kBigVoxelCap) {     // the dense array holds scatter cursors now: the radix path refills it
    t.dense = nullptr;
    t.dn[0] = -1;
    return cudaSuccess;
  }
  CK(table_reserve(ts, (size_t)n_vox, s, launches));
  if (n_vox > 0) {
    // 6 KB of static shared memory per 4-voxel block, 16 blocks per SM: ask for the large carve-out
    static const cudaError_t carve = cudaFuncSetAttribute(k_cs_merge, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    (void)carve;
    CK(cudaMemsetAsync(ws->counts + 1, 0, sizeof(int), s));     // [1] = big voxels listed by k_cs_merge
    k_cs_merge<<<(unsigned)((n_vox + 3) / 4), 128, 0, s>>>(pts, ws->idx_out, v_start, v_cnt, n_vox, t, big_list, ws->counts + 1);
    L += 2;
    if (n_big > 0) {
      k_cs_merge_big<<<(unsigned)(n_big < 296 ? n_big : 296), 256, 0, s>>>(pts, ws->idx_out, ws->idx_in, v_start, v_cnt, big_list, ws->counts + 1, t);
      ++L;
    }
    k_cs_finalize<<<(unsigned)((n_vox + 127) / 128), 128, 0, s>>>(t, n_vox, v_cell, v_cnt);
    ++L;
  }
  ts.n_vox = n_vox;
  *done = true;
  return cudaGetLastError();
}

static __global__ void k_finite_flags(const float4* __restrict__ pts, size_t n, int* __restrict__ flags) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = pts[i];
  flags[i] = (isfinite(p.x) && isfinite(p.y) && isfinite(p.z)) ? 1 : 0;
}

cudaError_t table_append(TableStorage& ts, BuildWorkspace* ws, const float4* pts, size_t n, cudaStream_t s, long long* launches) {
  if (n == 0) return cudaSuccess;
  VoxelTable& t = ts.t;
  CK(ws_reserve(ws, n));
  long long dummy = 0;
  long long& L = launches ? *launches : dummy;
  // bounding box of the new points
  float nmn[3], nmx[3];
  int n_bad = 0;
  CK(minmax_host(ws, pts, n, s, nmn, nmx, &n_bad, L));
  if (n_bad > 0) {
    // non-finite points are dropped before keying, in input order, as the reference's first pass skips them (VGC_I:209-216)
    if ((size_t)n_bad >= n) return cudaSuccess;
    if (n > ws->finite_cap) {
      cudaFree(ws->finite_pts);
      ws->finite_pts = nullptr;
      ws->finite_cap = 0;
      CK(cudaMalloc(&ws->finite_pts, (n + n / 4) * sizeof(float4)));
      ws->finite_cap = n + n / 4;
    }
    k_finite_flags<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(pts, n, ws->flags);
    size_t need = 0;
    CK(cub::DeviceSelect::Flagged(nullptr, need, pts, ws->flags, ws->finite_pts, ws->counts, (int)n, s));
    if (need > ws->cub_bytes) return cudaErrorMemoryAllocation;
    size_t tb2 = ws->cub_bytes;
    CK(cub::DeviceSelect::Flagged(ws->cub_temp, tb2, pts, ws->flags, ws->finite_pts, ws->counts, (int)n, s));
    L += 3;
    pts = ws->finite_pts;
    n -= (size_t)n_bad;
  }
  for (int a = 0; a < 3; ++a) {
    ts.mn[a] = ts.have_bbox ? fminf(ts.mn[a], nmn[a]) : nmn[a];
    ts.mx[a] = ts.have_bbox ? fmaxf(ts.mx[a], nmx[a]) : nmx[a];
  }
  ts.have_bbox = true;
  // overflow guard of the reference, on the bounding box of the WHOLE target (VGC_I:75-84)
  const float inv = t.inv_leaf;
  const long long dx = (long long)((ts.mx[0] - ts.mn[0]) * inv) + 1, dy = (long long)((ts.mx[1] - ts.mn[1]) * inv) + 1,
                  dz = (long long)((ts.mx[2] - ts.mn[2]) * inv) + 1;
  bool ok = dx * dy * dz <= 2147483647ll;
  for (int a = 0; a < 3; ++a) {
    ts.min_b[a] = (int)floorf(ts.mn[a] * inv);   // VGC_I:87-92
    ts.max_b[a] = (int)floorf(ts.mx[a] * inv);
    if (!coord_ok(ts.min_b[a]) || !coord_ok(ts.max_b[a])) ok = false;
    if (!(ts.mn[a] == ts.mn[a]) || !(ts.mx[a] == ts.mx[a])) ok = false;
  }
  if (!ok) {  // "Leaf size is too small for the input dataset": the reference leaves an EMPTY grid
    ts.valid = false;
    ts.n_vox = 0;
    t.dense = nullptr;
    t.dn[0] = -1;   // the dense array still holds the voxels dropped here: refill it at the next build
    if (ts.slot_cap) { k_fill_slots<<<(unsigned)((ts.slot_cap + 255) / 256), 256, 0, s>>>(t.slots, ts.slot_cap); ++L; }
    return cudaGetLastError();
  }
  ts.valid = true;
  if (ts.n_vox == 0) {
    bool done = false;
    CK(table_build_counting(ts, ws, pts, n, s, L, launches, &done));
    if (done) return cudaSuccess;
  }
  int c0[3], c1[3];
  for (int a = 0; a < 3; ++a) { c0[a] = (int)floorf(nmn[a] * inv); c1[a] = (int)floorf(nmx[a] * inv); }
  const int bi = bits_for((long long)c1[0] - c0[0] + 1), bj = bits_for((long long)c1[1] - c0[1] + 1), bk = bits_for((long long)c1[2] - c0[2] + 1);
  const unsigned nb = (unsigned)((n + 255) / 256);
  if (sort_key32(bi + bj + bk))
    k_keys<unsigned int><<<nb, 256, 0, s>>>(pts, n, inv, c0[0], c0[1], c0[2], bi, bj, reinterpret_cast<unsigned int*>(ws->keys_in), ws->idx_in);
  else
    k_keys<unsigned long long><<<nb, 256, 0, s>>>(pts, n, inv, c0[0], c0[1], c0[2], bi, bj, ws->keys_in, ws->idx_in);
  ++L;
  CK(sort_and_flag(ws, n, bi + bj + bk, s));
  L += (bi + bj + bk + 7) / 8 + 2;  // CUB onesweep: histogram + scan + one pass per 8 key bits (library kernels)
  size_t tb = ws->cub_bytes;
  CK(cub::DeviceScan::InclusiveSum(ws->cub_temp, tb, ws->flags, ws->scan, (int)n, s));
  k_run_starts<<<nb, 256, 0, s>>>(ws->flags, ws->scan, n, ws->run_start, ws->counts);
  L += 4;
  CK(cudaMemcpyAsync(ws->h_counts, ws->counts, sizeof(int), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  const int n_runs = ws->h_counts[0];
  CK(table_reserve(ts, (size_t)ts.n_vox + n_runs, s, launches));
  const unsigned rb = (unsigned)((n_runs + 127) / 128);
  // a build into an empty table (the radix path's fresh build: boxes too large for the dense grid) skips the run lookup, the
  // prefix scan and the third blocking read of the chain: every run is a new voxel, run r = voxel r (k_run_merge<true>)
  if (ts.n_vox == 0) {
    k_run_merge<true><<<rb, 128, 0, s>>>(pts, ws->idx_out, ws->run_start, nullptr, nullptr, nullptr, n_runs, 0, t, ws->counts);
    L += 1;
    ts.n_vox = n_runs;
  } else {
    k_run_lookup<<<rb, 128, 0, s>>>(pts, ws->idx_out, ws->run_start, n_runs, t, ws->run_vidx, ws->run_slot, ws->run_isnew);
    tb = ws->cub_bytes;
    CK(cub::DeviceScan::ExclusiveSum(ws->cub_temp, tb, ws->run_isnew, ws->new_prefix, n_runs, s));
    k_run_merge<false><<<rb, 128, 0, s>>>(pts, ws->idx_out, ws->run_start, ws->run_vidx, ws->run_slot, ws->new_prefix, n_runs, ts.n_vox, t,
                                          ws->counts);
    L += 4;
    CK(cudaMemcpyAsync(ws->h_counts + 1, ws->counts + 1, sizeof(int), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    ts.n_vox += ws->h_counts[1];
  }
  CK(table_build_dense(ts, s, launches));
  return cudaGetLastError();
}


// ---------------------------------------------------------------------------------------
// pcl::VoxelGrid<PointXYZI>::applyFilter with downsample_all_data_ = true (the default), as odom_node uses it for
// downSizeFilterLocalmap (reference src/odom_node.cpp:98, 334-335; PCL itself is not vendored -- SURVEY.md appendix B):
// one output point per occupied leaf = the centroid of x, y, z AND intensity over the leaf's points, float sums divided
// by float(count); output in ascending leaf index i + j*dx + k*dx*dy, i.e. ascending (k, j, i).  PCL orders the points
// of a leaf with std::sort on the leaf index alone (unspecified among equals); here, as in the oracle, they are summed
// in INPUT order (what a stable sort gives).  Leaf of a point: floor(x * inverse_leaf_size) - min_b, the same
// expression the target build uses.  Overflow guard (dx*dy*dz > INT_MAX): PCL warns and returns the input unchanged.
// in.w carries the intensity.  Blocks the stream twice (bounding box, leaf count).
// ---------------------------------------------------------------------------------------
__global__ void k_vg_centroid(const float4* __restrict__ pts, const unsigned int* __restrict__ sorted_idx, const int* __restrict__ run_start,
                              int n_runs, float4* __restrict__ out) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_runs) return;
  const int b = run_start[r], e = run_start[r + 1];
  float4 s = pts[sorted_idx[b]];
  for (int q = b + 1; q < e; ++q) {
    const float4 p = pts[sorted_idx[q]];
    s.x = __fadd_rn(s.x, p.x); s.y = __fadd_rn(s.y, p.y); s.z = __fadd_rn(s.z, p.z); s.w = __fadd_rn(s.w, p.w);
  }
  const float cnt = (float)(e - b);
  out[r] = make_float4(__fdiv_rn(s.x, cnt), __fdiv_rn(s.y, cnt), __fdiv_rn(s.z, cnt), __fdiv_rn(s.w, cnt));
}

cudaError_t voxelgrid_downsample(BuildWorkspace* ws, const float4* in, size_t n, float leaf, float4* out, size_t* n_out, cudaStream_t s,
                                 long long* launches) {
  *n_out = 0;
  if (n == 0) return cudaSuccess;
  CK(ws_reserve(ws, n));
  long long dummy = 0;
  long long& L = launches ? *launches : dummy;
  float mn[3], mx[3];
  CK(minmax_host(ws, in, n, s, mn, mx, nullptr, L));
  const float inv = 1.0f / leaf;     // pcl::VoxelGrid::setLeafSize
  const long long dx = (long long)((mx[0] - mn[0]) * inv) + 1, dy = (long long)((mx[1] - mn[1]) * inv) + 1, dz = (long long)((mx[2] - mn[2]) * inv) + 1;
  int c0[3], c1[3];
  bool ok = dx * dy * dz <= 2147483647ll;
  for (int a = 0; a < 3; ++a) {
    c0[a] = (int)floorf(mn[a] * inv);
    c1[a] = (int)floorf(mx[a] * inv);
    if (!coord_ok(c0[a]) || !coord_ok(c1[a]) || !(mn[a] == mn[a]) || !(mx[a] == mx[a])) ok = false;
  }
  if (!ok) {   // "Leaf size is too small for the input dataset. Integer indices would overflow.": output = input
    CK(cudaMemcpyAsync(out, in, n * sizeof(float4), cudaMemcpyDeviceToDevice, s));
    *n_out = n;
    return cudaSuccess;
  }
  const int bi = bits_for((long long)c1[0] - c0[0] + 1), bj = bits_for((long long)c1[1] - c0[1] + 1), bk = bits_for((long long)c1[2] - c0[2] + 1);
  const unsigned nb = (unsigned)((n + 255) / 256);
  if (sort_key32(bi + bj + bk))
    k_keys<unsigned int><<<nb, 256, 0, s>>>(in, n, inv, c0[0], c0[1], c0[2], bi, bj, reinterpret_cast<unsigned int*>(ws->keys_in), ws->idx_in);
  else
    k_keys<unsigned long long><<<nb, 256, 0, s>>>(in, n, inv, c0[0], c0[1], c0[2], bi, bj, ws->keys_in, ws->idx_in);
  CK(sort_and_flag(ws, n, bi + bj + bk, s));
  size_t tb = ws->cub_bytes;
  CK(cub::DeviceScan::InclusiveSum(ws->cub_temp, tb, ws->flags, ws->scan, (int)n, s));
  k_run_starts<<<nb, 256, 0, s>>>(ws->flags, ws->scan, n, ws->run_start, ws->counts);
  L += 5 + (bi + bj + bk + 7) / 8 + 2;
  CK(cudaMemcpyAsync(ws->h_counts, ws->counts, sizeof(int), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  const int n_runs = ws->h_counts[0];
  k_vg_centroid<<<(unsigned)((n_runs + 127) / 128), 128, 0, s>>>(in, ws->idx_out, ws->run_start, n_runs, out);
  ++L;
  *n_out = (size_t)n_runs;
  return cudaGetLastError();
}


// ---------------------------------------------------------------------------------------
// Exact 1-NN over the raw target for pcl::Registration::getFitnessScore (K6): target points
// sorted by cell, cell -> [start, end) in a hash table, ring search with the same stopping rule
// as a kd-tree search would give (exact minimum of the float squared distances).
// ---------------------------------------------------------------------------------------
struct __align__(16) CellSlot { unsigned long long key; int start; int end; };
struct NNGridStorage {
  float4* sorted = nullptr;
  size_t pts_cap = 0, n = 0;
  CellSlot* slots = nullptr;
  size_t slot_cap = 0;
  uint32_t mask = 0;
  float cell = 1.f;
  int* far_idx = nullptr;      // queries the ring search could not settle ...
  size_t far_cap = 0;
  int* far_count = nullptr;    // ... and how many there are
};
NNGridStorage* nn_create() { return new NNGridStorage(); }
void nn_destroy(NNGridStorage* g) {
  if (!g) return;
  cudaFree(g->sorted); cudaFree(g->slots); cudaFree(g->far_idx); cudaFree(g->far_count);
  delete g;
}
template <typename KeyT>
__global__ void k_nn_keys(const float4* __restrict__ pts, size_t n, float cell, int i0, int j0, int k0, int bi, int bj,
                          KeyT* __restrict__ keys, unsigned int* __restrict__ idx) {
  size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const float4 p = pts[t];
  const unsigned long long i = (unsigned long long)((int)floorf(p.x / cell) - i0), j = (unsigned long long)((int)floorf(p.y / cell) - j0),
                           k = (unsigned long long)((int)floorf(p.z / cell) - k0);
  keys[t] = (KeyT)((k << (bi + bj)) | (j << bi) | i);
  idx[t] = (unsigned int)t;
}
__global__ void k_nn_gather(const float4* __restrict__ pts, const unsigned int* __restrict__ idx, size_t n, float4* __restrict__ out) {
  size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) out[t] = pts[idx[t]];
}
__global__ void k_nn_fill(CellSlot* slots, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { CellSlot s; s.key = kEmptyKey; s.start = 0; s.end = 0; slots[i] = s; }
}
__global__ void k_nn_insert(const float4* __restrict__ sorted, const int* __restrict__ run_start, const int* __restrict__ n_runs_dev, float cell,
                            CellSlot* __restrict__ slots, uint32_t mask) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= *n_runs_dev) return;
  const float4 p = sorted[run_start[r]];
  const unsigned long long key = pack_key((int)floorf(p.x / cell), (int)floorf(p.y / cell), (int)floorf(p.z / cell));
  uint32_t h = hash_key(key) & mask;
  while (true) {
    const unsigned long long prev = atomicCAS(&slots[h].key, kEmptyKey, key);
    if (prev == kEmptyKey) break;
    h = (h + 1) & mask;
  }
  slots[h].start = run_start[r];
  slots[h].end = run_start[r + 1];
}
// bbox: the float bounding box of pts when the caller has it (min xyz, max xyz; the voxel table keeps it) -- the build
// then never blocks the stream; nullptr: computed here (one synchronisation).
cudaError_t nn_build(NNGridStorage* g, BuildWorkspace* ws, const float4* pts, size_t n, float cell, const float* bbox, cudaStream_t s,
                     long long* launches) {
  long long dummy = 0;
  long long& L = launches ? *launches : dummy;
  g->n = n;
  g->cell = cell;
  if (n == 0) return cudaSuccess;
  CK(ws_reserve(ws, n));
  if (n > g->pts_cap) {
    cudaFree(g->sorted);
    g->sorted = nullptr;
    CK(cudaMalloc(&g->sorted, (n + n / 2) * sizeof(float4)));
    g->pts_cap = n + n / 2;
  }
  float bb[6];
  if (bbox) {
    for (int a = 0; a < 6; ++a) bb[a] = bbox[a];
  } else {
    CK(minmax_host(ws, pts, n, s, bb, bb + 3, nullptr, L));
  }
  int c0[3], c1[3];
  for (int a = 0; a < 3; ++a) {
    c0[a] = (int)floorf(bb[a] / cell); c1[a] = (int)floorf(bb[3 + a] / cell);
    if (!coord_ok(c0[a]) || !coord_ok(c1[a])) return cudaErrorInvalidValue;
  }
  const int bi = bits_for((long long)c1[0] - c0[0] + 1), bj = bits_for((long long)c1[1] - c0[1] + 1), bk = bits_for((long long)c1[2] - c0[2] + 1);
  const unsigned nb = (unsigned)((n + 255) / 256);
  if (sort_key32(bi + bj + bk))
    k_nn_keys<unsigned int><<<nb, 256, 0, s>>>(pts, n, cell, c0[0], c0[1], c0[2], bi, bj, reinterpret_cast<unsigned int*>(ws->keys_in), ws->idx_in);
  else
    k_nn_keys<unsigned long long><<<nb, 256, 0, s>>>(pts, n, cell, c0[0], c0[1], c0[2], bi, bj, ws->keys_in, ws->idx_in);
  CK(sort_and_flag(ws, n, bi + bj + bk, s));
  k_nn_gather<<<nb, 256, 0, s>>>(pts, ws->idx_out, n, g->sorted);
  size_t tb = ws->cub_bytes;
  CK(cub::DeviceScan::InclusiveSum(ws->cub_temp, tb, ws->flags, ws->scan, (int)n, s));
  k_run_starts<<<nb, 256, 0, s>>>(ws->flags, ws->scan, n, ws->run_start, ws->counts);
  // occupied cells <= points: a table of >= 2n slots keeps the load factor <= 0.5 without asking the device how many there are
  size_t want = 1024;
  while (want < n * 2) want <<= 1;
  if (want > g->slot_cap) {
    cudaFree(g->slots);
    g->slots = nullptr;
    CK(cudaMalloc(&g->slots, want * sizeof(CellSlot)));
    g->slot_cap = want;
  }
  g->mask = (uint32_t)(want - 1);
  k_nn_fill<<<(unsigned)((want + 255) / 256), 256, 0, s>>>(g->slots, want);
  k_nn_insert<<<(unsigned)((n + 127) / 128), 128, 0, s>>>(g->sorted, ws->run_start, ws->counts, cell, g->slots, g->mask);
  L += 8 + (bi + bj + bk + 7) / 8;
  return cudaGetLastError();
}

// Ring search over the cell grid.  A query is settled once its best squared distance is <= ((ring-1)*cell)^2, the least
// any point of a farther ring can be away.  Queries still open after `max_ring` rings (the scan reaches far beyond a
// small localmap) go to k_nn_far instead of walking ever larger, empty shells.
// Eight lanes share a query: the cells of a shell are dealt round-robin to the lanes, the minimum is taken across the group
// after every ring (the minimum of the same float expressions whatever the order: the exact 1-NN distance).  One thread per
// query walked up to 26 + 98 cells of dependent hash probes and point loops alone and left most SMs without a block.
constexpr int kNNGroup = 8;
__global__ void __launch_bounds__(128) k_nn_query(const float4* __restrict__ sorted, size_t n_tgt, const CellSlot* __restrict__ slots, uint32_t mask,
                                                  float cell, const float4* __restrict__ src, size_t n, const float* __restrict__ T16,
                                                  int max_ring, float* __restrict__ d2_out, int* __restrict__ far_idx,
                                                  int* __restrict__ far_count) {
  const size_t gt = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t i = gt / kNNGroup;
  const int g = (int)(gt % kNNGroup);
  const bool live = i < n;                       // (no early return: the group shuffles below need every lane of the group)
  const unsigned gmask = 0xFFu << ((threadIdx.x & 31) & ~(kNNGroup - 1));
  const float4 p = live ? src[i] : make_float4(0.f, 0.f, 0.f, 0.f);
  // pcl::transformPointCloud by final_transformation_ (float, left to right)
  const float qx = ((T16[0] * p.x + T16[4] * p.y) + T16[8] * p.z) + T16[12];
  const float qy = ((T16[1] * p.x + T16[5] * p.y) + T16[9] * p.z) + T16[13];
  const float qz = ((T16[2] * p.x + T16[6] * p.y) + T16[10] * p.z) + T16[14];
  const int ci = (int)floorf(qx / cell), cj = (int)floorf(qy / cell), ck = (int)floorf(qz / cell);
  float best = FLT_MAX;
  bool certain = false;
  for (int ring = 0; ring <= max_ring + 1; ++ring) {
    if (ring >= 1) {
      const float lim = (float)(ring - 1) * cell;   // every point in rings >= `ring` is at least this far
      if (best <= lim * lim) { certain = true; break; }
    }
    if (ring > max_ring) break;
    const int side = 2 * ring + 1, cells = side * side * side;
    if (live)
      for (int c = g; c < cells; c += kNNGroup) {
        const int dx = c % side - ring, dy = (c / side) % side - ring, dz = c / (side * side) - ring;
        if (max(abs(dx), max(abs(dy), abs(dz))) != ring) continue;
        const int a = ci + dx, b = cj + dy, cc = ck + dz;
        if (!coord_ok(a) || !coord_ok(b) || !coord_ok(cc)) continue;
        const unsigned long long key = pack_key(a, b, cc);
        uint32_t h = hash_key(key) & mask;
        int s0 = 0, s1 = 0;
        while (true) {
          const CellSlot sl = slots[h];
          if (sl.key == key) { s0 = sl.start; s1 = sl.end; break; }
          if (sl.key == kEmptyKey) break;
          h = (h + 1) & mask;
        }
        for (int t = s0; t < s1; ++t) {
          const float4 tp = sorted[t];
          const float ex = qx - tp.x, ey = qy - tp.y, ez = qz - tp.z;
          const float d = (ex * ex + ey * ey) + ez * ez;
          if (d < best) best = d;
        }
      }
    // the group's minimum (groups are aligned 8-lane segments of a warp and shuffle under their own mask: `best`, hence every
    // loop exit, is uniform inside a group, while the groups of a warp leave the loop at different rings)
#pragma unroll
    for (int o = kNNGroup / 2; o > 0; o >>= 1) best = fminf(best, __shfl_xor_sync(gmask, best, o));
  }
  if (live && g == 0) {
    d2_out[i] = best;
    if (!certain) far_idx[atomicAdd(far_count, 1)] = (int)i;
  }
}
// The open queries against EVERY target point: one thread per query, the target streamed through shared memory in
// tiles (the minimum of the same float expression, so the result is the exact 1-NN distance whatever the order).
constexpr int kFarThreads = 128, kFarTile = 1024;
__global__ void __launch_bounds__(kFarThreads)
k_nn_far(const float4* __restrict__ sorted, size_t n_tgt, const float4* __restrict__ src, const float* __restrict__ T16,
         const int* __restrict__ far_idx, const int* __restrict__ far_count, float* __restrict__ d2_out) {
  __shared__ float4 tile[kFarTile];
  const int n_far = *far_count;
  if ((int)(blockIdx.x * kFarThreads) >= n_far) return;
  const int f = blockIdx.x * kFarThreads + threadIdx.x;
  const bool live = f < n_far;
  const int i = live ? far_idx[f] : 0;
  float qx = 0.f, qy = 0.f, qz = 0.f;
  if (live) {
    const float4 p = src[i];
    qx = ((T16[0] * p.x + T16[4] * p.y) + T16[8] * p.z) + T16[12];
    qy = ((T16[1] * p.x + T16[5] * p.y) + T16[9] * p.z) + T16[13];
    qz = ((T16[2] * p.x + T16[6] * p.y) + T16[10] * p.z) + T16[14];
  }
  float best = FLT_MAX;
  // blockIdx.y takes one slice of the target, so that no thread walks the whole of it
  const size_t slice = (n_tgt + gridDim.y - 1) / gridDim.y;
  const size_t lo = (size_t)blockIdx.y * slice, hi = min(n_tgt, lo + slice);
  for (size_t base = lo; base < hi; base += kFarTile) {
    const int m = (int)min((size_t)kFarTile, hi - base);
    __syncthreads();
    for (int t = threadIdx.x; t < m; t += kFarThreads) tile[t] = sorted[base + t];
    __syncthreads();
#pragma unroll 8
    for (int t = 0; t < m; ++t) {
      const float4 tp = tile[t];
      const float ex = qx - tp.x, ey = qy - tp.y, ez = qz - tp.z;
      const float d = (ex * ex + ey * ey) + ez * ez;
      best = fminf(best, d);
    }
  }
  // squared distances are >= 0: their bit patterns order like the floats, so the slices combine with an integer atomicMin
  // (the entry already holds the ring search's best, itself a distance to a real target point)
  if (live) atomicMin(reinterpret_cast<int*>(d2_out) + i, __float_as_int(best));
}
cudaError_t nn_query(NNGridStorage* g, const float4* src, size_t n, const float* T16_dev, float* d2_out, cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  if (n > g->far_cap) {
    cudaFree(g->far_idx);
    g->far_idx = nullptr;
    CK(cudaMalloc(&g->far_idx, (n + n / 4) * sizeof(int)));
    g->far_cap = n + n / 4;
  }
  if (!g->far_count) CK(cudaMalloc(&g->far_count, sizeof(int)));
  CK(cudaMemsetAsync(g->far_count, 0, sizeof(int), s));
  // small target: hand a query to the exhaustive pass early (its cost is n_far * n_tgt); large target: keep walking rings
  const int max_ring = g->n <= (size_t)1 << 17 ? 2 : 8;
  k_nn_query<<<(unsigned)((n * kNNGroup + 127) / 128), 128, 0, s>>>(g->sorted, g->n, g->slots, g->mask, g->cell, src, n, T16_dev, max_ring, d2_out,
                                                         g->far_idx, g->far_count);
  unsigned slices = (unsigned)((g->n + 2047) / 2048);
  if (slices > 64) slices = 64;
  if (slices < 1) slices = 1;
  k_nn_far<<<dim3((unsigned)((n + kFarThreads - 1) / kFarThreads), slices), kFarThreads, 0, s>>>(g->sorted, g->n, src, T16_dev, g->far_idx,
                                                                                                 g->far_count, d2_out);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Morton order of a batch's sources (see ndt_internal.h)
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned int spread3(unsigned int v) {   // 6 bits -> every third bit
  v &= 0x3Fu;
  v = (v | (v << 8)) & 0x0000F00Fu;   // only the low 6 bits are populated: the masks of the 10-bit version, truncated
  v = (v | (v << 4)) & 0x000C30C3u;
  v = (v | (v << 2)) & 0x00249249u;
  return v;
}
__global__ void k_source_keys(const SrcDesc* __restrict__ desc, int n_src, size_t total, unsigned int* __restrict__ keys,
                              float4* __restrict__ vals) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  int lo = 0, hi = n_src - 1;              // the last source whose first point is <= t
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if ((size_t)desc[mid].first <= t) lo = mid; else hi = mid - 1;
  }
  const SrcDesc d = desc[lo];
  const float4 p = d.p[t - d.first];
  const int a = __float2int_rd(p.x * 0.5f), b = __float2int_rd(p.y * 0.5f), c = __float2int_rd(p.z * 0.5f);   // 2 m cells, wrapped at 64
  const unsigned int m = spread3((unsigned int)a) | (spread3((unsigned int)b) << 1) | (spread3((unsigned int)c) << 2);
  keys[t] = ((unsigned int)lo << 18) | m;
  vals[t] = p;
}
void source_order_free(SourceOrderWs& w) {
  cudaFree(w.desc); cudaFree(w.k_in); cudaFree(w.k_out); cudaFree(w.v_in); cudaFree(w.temp);
  w = SourceOrderWs();
}
cudaError_t source_order_sort(SourceOrderWs& w, const SrcDesc* desc_host, int n_src, size_t total, float4* out, cudaStream_t s,
                              long long* launches) {
  if (n_src <= 0 || total == 0) return cudaSuccess;
  if ((size_t)n_src > w.desc_cap) {
    cudaFree(w.desc); w.desc = nullptr; w.desc_cap = 0;
    CK(cudaMalloc(&w.desc, ((size_t)n_src + 64) * sizeof(SrcDesc)));
    w.desc_cap = (size_t)n_src + 64;
  }
  if (total > w.cap) {
    cudaFree(w.k_in); cudaFree(w.k_out); cudaFree(w.v_in);
    w.k_in = w.k_out = nullptr; w.v_in = nullptr; w.cap = 0;
    const size_t nc = total + total / 8;
    CK(cudaMalloc(&w.k_in, nc * sizeof(unsigned int)));
    CK(cudaMalloc(&w.k_out, nc * sizeof(unsigned int)));
    CK(cudaMalloc(&w.v_in, nc * sizeof(float4)));
    w.cap = nc;
  }
  int bits = 18;
  while (bits < 32 && (1u << (bits - 18)) < (unsigned int)n_src) ++bits;
  size_t need = 0;
  CK(cub::DeviceRadixSort::SortPairs(nullptr, need, w.k_in, w.k_out, w.v_in, out, (int)total, 0, bits, s));
  if (need > w.temp_bytes) {
    cudaFree(w.temp); w.temp = nullptr; w.temp_bytes = 0;
    CK(cudaMalloc(&w.temp, need + need / 4 + 256));
    w.temp_bytes = need + need / 4 + 256;
  }
  CK(cudaMemcpyAsync(w.desc, desc_host, (size_t)n_src * sizeof(SrcDesc), cudaMemcpyHostToDevice, s));
  k_source_keys<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(w.desc, n_src, total, w.k_in, w.v_in);
  size_t tb = w.temp_bytes;
  CK(cub::DeviceRadixSort::SortPairs(w.temp, tb, w.k_in, w.k_out, w.v_in, out, (int)total, 0, bits, s));
  if (launches) *launches += 3 + (bits + 7) / 8;
  return cudaGetLastError();
}

// getFitnessScore of one registration: sum and count of the squared distances <= max_range, one block, fixed order (thread-
// strided partial sums, then lanes, then warps) -- the host reads 16 bytes instead of one float per source point.
__global__ void __launch_bounds__(1024) k_d2_reduce(const float* __restrict__ d2, size_t n, double max_range, double* __restrict__ out) {
  __shared__ double ss[32], sc[32];
  double s = 0.0, c = 0.0;
  for (size_t i = threadIdx.x; i < n; i += 1024) {
    const float d = d2[i];
    if ((double)d <= max_range) { s += (double)d; c += 1.0; }
  }
  for (int o = 16; o > 0; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); c += __shfl_xor_sync(0xffffffffu, c, o); }
  if ((threadIdx.x & 31) == 0) { ss[threadIdx.x >> 5] = s; sc[threadIdx.x >> 5] = c; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double ts = 0.0, tc = 0.0;
    for (int w = 0; w < 32; ++w) { ts += ss[w]; tc += sc[w]; }
    out[0] = ts; out[1] = tc;
  }
}
cudaError_t d2_reduce(const float* d2, size_t n, double max_range, double* out2_dev, cudaStream_t s) {
  k_d2_reduce<<<1, 1024, 0, s>>>(d2, n, max_range, out2_dev);
  return cudaGetLastError();
}

// Batched getFitnessScore (loop-closure candidates are accepted on it, reference src/pgo_node.cpp:460-476): every
// registration of a batch that shares one target, in one launch.  blockIdx.y walks `items` (ids into `states`); the
// transform is the registration's final one.  A query the ring search cannot settle scans the whole target itself (a scan
// registered against a submap lies inside it, so this is rare).  Per item: sum and count of the squared distances
// <= max_range, reduced in a fixed order (lane, then warp, then block partials added by k_fitness_finish).
__global__ void __launch_bounds__(128)
k_nn_fitness_batch(const float4* __restrict__ sorted, size_t n_tgt, const CellSlot* __restrict__ slots, uint32_t mask, float cell,
                   const AlignState* __restrict__ states, const int* __restrict__ items, double max_range, int blocks_per_item,
                   double* __restrict__ partial /* [n_items_total][blocks_per_item][2] */) {
  const int item = items[blockIdx.y];
  const AlignState& st = states[item];
  const float* T16 = st.final_T;
  const int n = st.n_src;
  double s = 0.0, c = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += blocks_per_item * blockDim.x) {
    const float4 p = st.src[i];
    const float qx = ((T16[0] * p.x + T16[4] * p.y) + T16[8] * p.z) + T16[12];
    const float qy = ((T16[1] * p.x + T16[5] * p.y) + T16[9] * p.z) + T16[13];
    const float qz = ((T16[2] * p.x + T16[6] * p.y) + T16[10] * p.z) + T16[14];
    const int ci = (int)floorf(qx / cell), cj = (int)floorf(qy / cell), ck = (int)floorf(qz / cell);
    float best = FLT_MAX;
    bool certain = false;
    for (int ring = 0; ring <= 9; ++ring) {
      if (ring >= 1) {
        const float lim = (float)(ring - 1) * cell;
        if (best <= lim * lim) { certain = true; break; }
      }
      if (ring > 8) break;
      for (int dz = -ring; dz <= ring; ++dz)
        for (int dy = -ring; dy <= ring; ++dy)
          for (int dx = -ring; dx <= ring; ++dx) {
            if (max(abs(dx), max(abs(dy), abs(dz))) != ring) continue;
            const int a = ci + dx, b = cj + dy, cc = ck + dz;
            if (!coord_ok(a) || !coord_ok(b) || !coord_ok(cc)) continue;
            const unsigned long long key = pack_key(a, b, cc);
            uint32_t h = hash_key(key) & mask;
            int s0 = 0, s1 = 0;
            while (true) {
              const CellSlot sl = slots[h];
              if (sl.key == key) { s0 = sl.start; s1 = sl.end; break; }
              if (sl.key == kEmptyKey) break;
              h = (h + 1) & mask;
            }
            for (int t = s0; t < s1; ++t) {
              const float4 tp = sorted[t];
              const float ex = qx - tp.x, ey = qy - tp.y, ez = qz - tp.z;
              const float d = (ex * ex + ey * ey) + ez * ez;
              if (d < best) best = d;
            }
          }
    }
    if (!certain)
      for (size_t t = 0; t < n_tgt; ++t) {
        const float4 tp = sorted[t];
        const float ex = qx - tp.x, ey = qy - tp.y, ez = qz - tp.z;
        const float d = (ex * ex + ey * ey) + ez * ez;
        if (d < best) best = d;
      }
    if ((double)best <= max_range) { s += (double)best; c += 1.0; }
  }
  __shared__ double sh[2][4];
  for (int o = 16; o > 0; o >>= 1) { s += __shfl_down_sync(0xffffffffu, s, o); c += __shfl_down_sync(0xffffffffu, c, o); }
  if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = s; sh[1][threadIdx.x >> 5] = c; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double* o = partial + ((size_t)item * blocks_per_item + blockIdx.x) * 2;
    o[0] = ((sh[0][0] + sh[0][1]) + sh[0][2]) + sh[0][3];
    o[1] = ((sh[1][0] + sh[1][1]) + sh[1][2]) + sh[1][3];
  }
}
__global__ void k_fitness_finish(const double* __restrict__ partial, int n_items, int blocks_per_item, double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_items) return;
  double s = 0.0, c = 0.0;
  for (int b = 0; b < blocks_per_item; ++b) { s += partial[((size_t)i * blocks_per_item + b) * 2]; c += partial[((size_t)i * blocks_per_item + b) * 2 + 1]; }
  out[i] = c > 0.0 ? s / c : 1.7976931348623157e308;   // std::numeric_limits<double>::max () when no point is in range
}
cudaError_t nn_fitness_batch(NNGridStorage* g, const AlignState* states, const int* items_dev, int n_group, double max_range,
                             int blocks_per_item, double* partial, cudaStream_t s) {
  if (n_group <= 0) return cudaSuccess;
  k_nn_fitness_batch<<<dim3((unsigned)blocks_per_item, (unsigned)n_group), 128, 0, s>>>(g->sorted, g->n, g->slots, g->mask, g->cell, states,
                                                                                    items_dev, max_range, blocks_per_item, partial);
  return cudaGetLastError();
}
cudaError_t fitness_finish(const double* partial, int n_items, int blocks_per_item, double* out, cudaStream_t s) {
  k_fitness_finish<<<(n_items + 127) / 128, 128, 0, s>>>(partial, n_items, blocks_per_item, out);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// filter_node's front end on the device (reference src/filter_node.cpp:236-273):
// pcl::removeNaNFromPointCloud + range gate 1 < r_xy < 60 (:236-247), pcl::VoxelGrid 0.5 m (:249-251, the kernels of
// voxelgrid_downsample above), pcl::StatisticalOutlierRemoval k = 30, 1 sigma (:254-262).  The third-party pieces are
// restated in oracle/thirdparty_math.h; everything here is float / double arithmetic in that order, so the result
// is the same cloud bit for bit.
// ---------------------------------------------------------------------------------------
__global__ void k_gate_flags(const float4* __restrict__ pts, size_t n, double r_min, double r_max, int* __restrict__ flags) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = pts[i];
  const bool finite = isfinite(p.x) && isfinite(p.y) && isfinite(p.z);
  const double r = sqrt((double)p.x * (double)p.x + (double)p.y * (double)p.y);   // std::sqrt(pow(x, 2.0) + pow(y, 2.0))
  flags[i] = (finite && r_min < r && r < r_max) ? 1 : 0;
}
__global__ void k_sor_flags(const float* __restrict__ dist, size_t n, double threshold, int* __restrict__ flags) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flags[i] = ((double)dist[i] > threshold) ? 0 : 1;
}

constexpr int kKnnMax = 64;   // mean_k + 1 <= kKnnMax
// ascending array of the k smallest squared distances seen so far
__device__ __forceinline__ void knn_insert(float* best, int& cnt, int k, float d) {
  if (cnt == k && !(d < best[k - 1])) return;
  int pos = cnt < k ? cnt : k - 1;
  while (pos > 0 && best[pos - 1] > d) { best[pos] = best[pos - 1]; --pos; }
  best[pos] = d;
  if (cnt < k) ++cnt;
}
// dist_sum (double) += sqrt(nn_dists[k]), k = 1..mean_k; distances[i] = float(dist_sum / mean_k)
__device__ __forceinline__ float knn_mean(const float* best, int mean_k) {
  double s = 0.0;
  for (int k = 1; k < mean_k + 1; ++k) s += sqrt((double)best[k]);
  return (float)(s / (double)mean_k);
}
// Ring search over the cell grid, one thread per point of the CELL-SORTED copy (position s; its index in the cloud is
// order[s]).  Settled once the k-th best is <= ((ring-1)*cell)^2; points still open after max_ring rings go to k_knn_sweep.
__global__ void k_knn_mean(const float4* __restrict__ sorted, const unsigned int* __restrict__ order, size_t n, const CellSlot* __restrict__ slots,
                           uint32_t mask, float cell, int mean_k, int max_ring, float* __restrict__ dist, int* __restrict__ far_idx,
                           int* __restrict__ far_count) {
  const size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const float4 q = sorted[s];
  const int ci = (int)floorf(q.x / cell), cj = (int)floorf(q.y / cell), ck = (int)floorf(q.z / cell);
  float best[kKnnMax];
  int cnt = 0;
  const int k = mean_k + 1;
  bool certain = false;
  for (int ring = 0; ring <= max_ring + 1; ++ring) {
    if (ring >= 1 && cnt == k) {
      const float lim = (float)(ring - 1) * cell;
      if (best[k - 1] <= lim * lim) { certain = true; break; }
    }
    if (ring > max_ring) break;
    for (int dz = -ring; dz <= ring; ++dz)
      for (int dy = -ring; dy <= ring; ++dy)
        for (int dx = -ring; dx <= ring; ++dx) {
          if (max(abs(dx), max(abs(dy), abs(dz))) != ring) continue;
          const int a = ci + dx, b = cj + dy, c = ck + dz;
          if (!coord_ok(a) || !coord_ok(b) || !coord_ok(c)) continue;
          const unsigned long long key = pack_key(a, b, c);
          uint32_t h = hash_key(key) & mask;
          int s0 = 0, s1 = 0;
          while (true) {
            const CellSlot sl = slots[h];
            if (sl.key == key) { s0 = sl.start; s1 = sl.end; break; }
            if (sl.key == kEmptyKey) break;
            h = (h + 1) & mask;
          }
          for (int t = s0; t < s1; ++t) {
            const float4 tp = sorted[t];
            const float ex = q.x - tp.x, ey = q.y - tp.y, ez = q.z - tp.z;
            knn_insert(best, cnt, k, (ex * ex + ey * ey) + ez * ez);
          }
        }
  }
  if (certain) dist[order[s]] = knn_mean(best, mean_k);
  else far_idx[atomicAdd(far_count, 1)] = (int)s;
}
// Exhaustive k-NN for the listed positions (far_idx == nullptr: every position): the cell-sorted copy is swept OUTWARD from
// the query's own position, so that its spatial neighbours come first, the k-th best tightens within a few dozen
// candidates and almost every later candidate is rejected by one comparison.  Exact: every point is visited.
__global__ void __launch_bounds__(128)
k_knn_sweep(const float4* __restrict__ sorted, const unsigned int* __restrict__ order, size_t n, int mean_k, const int* __restrict__ far_idx,
            const int* __restrict__ far_count, float* __restrict__ dist) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  const int n_list = far_idx ? *far_count : (int)n;
  if (f >= n_list) return;
  const int s = far_idx ? far_idx[f] : f;
  const float4 q = sorted[s];
  float best[kKnnMax];
  int cnt = 0;
  const int k = mean_k + 1;
  const int nn = (int)n;
  float kth = FLT_MAX;            // register copy of best[k-1] once k candidates are held: the common case is one compare
  // four candidates from each side per round, loads first: the sweep is bound by the latency of its loads otherwise
  for (int o = 0; o < nn; o += 4) {
    float dh[4], dl[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int jh = s + o + u, jl = s - 1 - o - u;
      dh[u] = FLT_MAX; dl[u] = FLT_MAX;
      if (jh < nn) {
        const float4 tp = __ldg(sorted + jh);
        const float ex = q.x - tp.x, ey = q.y - tp.y, ez = q.z - tp.z;
        dh[u] = (ex * ex + ey * ey) + ez * ez;
      }
      if (jl >= 0) {
        const float4 tp = __ldg(sorted + jl);
        const float ex = q.x - tp.x, ey = q.y - tp.y, ez = q.z - tp.z;
        dl[u] = (ex * ex + ey * ey) + ez * ez;
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (s + o + u < nn && (cnt < k || dh[u] < kth)) { knn_insert(best, cnt, k, dh[u]); if (cnt == k) kth = best[k - 1]; }
      if (s - 1 - o - u >= 0 && (cnt < k || dl[u] < kth)) { knn_insert(best, cnt, k, dl[u]); if (cnt == k) kth = best[k - 1]; }
    }
    if (s + o + 4 >= nn && s - 1 - o - 4 < 0) break;
  }
  dist[order[s]] = knn_mean(best, mean_k);
}

struct FilterWorkspace {
  float4 *a = nullptr, *b = nullptr;   // gated cloud, downsampled cloud
  int* flags = nullptr;
  float* dist = nullptr;
  size_t cap = 0;
  int* d_count = nullptr;
  int* h_count = nullptr;               // pinned
  float* h_dist = nullptr;              // pinned, cap floats
  NNGridStorage* nn = nullptr;
};
FilterWorkspace* filter_ws_create() {
  FilterWorkspace* f = new FilterWorkspace();
  f->nn = nn_create();
  if (cudaMalloc(&f->d_count, sizeof(int)) != cudaSuccess || cudaMallocHost(&f->h_count, sizeof(int)) != cudaSuccess) { delete f; return nullptr; }
  return f;
}
void filter_ws_destroy(FilterWorkspace* f) {
  if (!f) return;
  cudaFree(f->a); cudaFree(f->b); cudaFree(f->flags); cudaFree(f->dist); cudaFree(f->d_count);
  cudaFreeHost(f->h_count); cudaFreeHost(f->h_dist);
  nn_destroy(f->nn);
  delete f;
}
static cudaError_t filter_ws_reserve(FilterWorkspace* f, size_t n) {
  if (n <= f->cap) return cudaSuccess;
  cudaFree(f->a); cudaFree(f->b); cudaFree(f->flags); cudaFree(f->dist); cudaFreeHost(f->h_dist);
  f->a = f->b = nullptr; f->flags = nullptr; f->dist = nullptr; f->h_dist = nullptr; f->cap = 0;
  const size_t cap = n + n / 4 + 1024;
  CK(cudaMalloc(&f->a, cap * sizeof(float4))); CK(cudaMalloc(&f->b, cap * sizeof(float4)));
  CK(cudaMalloc(&f->flags, cap * sizeof(int))); CK(cudaMalloc(&f->dist, cap * sizeof(float)));
  CK(cudaMallocHost(&f->h_dist, cap * sizeof(float)));
  f->cap = cap;
  return cudaSuccess;
}
// order-preserving compaction of the flagged records; *n_out on the host (blocks the stream once)
static cudaError_t compact_flagged(FilterWorkspace* f, BuildWorkspace* ws, const float4* in, size_t n, float4* out, size_t* n_out,
                                   cudaStream_t s) {
  size_t need = 0;
  CK(cub::DeviceSelect::Flagged(nullptr, need, in, f->flags, out, f->d_count, (int)n, s));
  if (need > ws->cub_bytes) return cudaErrorMemoryAllocation;   // ws_reserve(n) sized the scratch for a 64-bit pair sort of n: always larger
  size_t tb = ws->cub_bytes;
  CK(cub::DeviceSelect::Flagged(ws->cub_temp, tb, in, f->flags, out, f->d_count, (int)n, s));
  CK(cudaMemcpyAsync(f->h_count, f->d_count, sizeof(int), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  *n_out = (size_t)*f->h_count;
  return cudaSuccess;
}

// raw: n device records (x y z intensity).  out: room for n records.  counts[0..2] = points after the gate, after the
// voxel downsample, after the outlier removal (= records written to out).
cudaError_t filter_scan(FilterWorkspace* f, BuildWorkspace* ws, const float4* raw, size_t n, double r_min, double r_max, float leaf, int mean_k,
                        double std_mul, float4* out, size_t counts[3], cudaStream_t s, long long* launches) {
  counts[0] = counts[1] = counts[2] = 0;
  if (n == 0) return cudaSuccess;
  if (mean_k + 1 > kKnnMax) return cudaErrorInvalidValue;   // mean_k <= 0: no outlier removal
  long long dummy = 0;
  long long& L = launches ? *launches : dummy;
  CK(filter_ws_reserve(f, n));
  CK(ws_reserve(ws, n));
  const unsigned nb = (unsigned)((n + 255) / 256);
  k_gate_flags<<<nb, 256, 0, s>>>(raw, n, r_min, r_max, f->flags);
  size_t n_gate = 0;
  CK(compact_flagged(f, ws, raw, n, f->a, &n_gate, s));
  L += 4;
  counts[0] = n_gate;
  if (n_gate == 0) return cudaSuccess;
  size_t n_ds = 0;
  if (leaf > 0.f) CK(voxelgrid_downsample(ws, f->a, n_gate, leaf, f->b, &n_ds, s, launches));
  else { CK(cudaMemcpyAsync(f->b, f->a, n_gate * sizeof(float4), cudaMemcpyDeviceToDevice, s)); n_ds = n_gate; }
  counts[1] = n_ds;
  if (mean_k <= 0 || n_ds < (size_t)mean_k + 1) {   // no outlier removal asked for, or too few points for the neighbour statistics: returned unchanged (thirdparty_math.h DECISION)
    CK(cudaMemcpyAsync(out, f->b, n_ds * sizeof(float4), cudaMemcpyDeviceToDevice, s));
    counts[2] = n_ds;
    return cudaSuccess;
  }
  // exact k-NN over a cell grid of the downsampled cloud (cells of two leaves: a handful of points each)
  const float cell = leaf > 0.f ? 2.f * leaf : 1.f;
  CK(nn_build(f->nn, ws, f->b, n_ds, cell, nullptr, s, launches));
  NNGridStorage* g = f->nn;
  if (n_ds > g->far_cap) {
    cudaFree(g->far_idx);
    g->far_idx = nullptr;
    CK(cudaMalloc(&g->far_idx, (n_ds + n_ds / 4) * sizeof(int)));
    g->far_cap = n_ds + n_ds / 4;
  }
  if (!g->far_count) CK(cudaMalloc(&g->far_count, sizeof(int)));
  CK(cudaMemsetAsync(g->far_count, 0, sizeof(int), s));
  const unsigned qb = (unsigned)((n_ds + 127) / 128);
  // ws->idx_out still holds the permutation of nn_build's sort: position in the cell-sorted copy -> index in the cloud
  if (n_ds <= 32768) {
    // a downsampled 64-beam scan is sparse (the 31st neighbour of most points lies several cells away): sweeping the whole
    // cloud per point beats walking mostly empty shells of cells
    k_knn_sweep<<<qb, 128, 0, s>>>(g->sorted, ws->idx_out, n_ds, mean_k, nullptr, nullptr, f->dist);
    L += 1;
  } else {
    k_knn_mean<<<qb, 128, 0, s>>>(g->sorted, ws->idx_out, n_ds, g->slots, g->mask, cell, mean_k, 2, f->dist, g->far_idx, g->far_count);
    k_knn_sweep<<<qb, 128, 0, s>>>(g->sorted, ws->idx_out, n_ds, mean_k, g->far_idx, g->far_count, f->dist);
    L += 2;
  }
  CK(cudaMemcpyAsync(f->h_dist, f->dist, n_ds * sizeof(float), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  // mean and standard deviation of the distance vector, sequential in point order (statistical_outlier_removal.hpp)
  double sum = 0, sq_sum = 0;
  for (size_t i = 0; i < n_ds; ++i) {
    sum += f->h_dist[i];
    sq_sum += f->h_dist[i] * f->h_dist[i];   // a FLOAT product added to a double, as in PCL
  }
  const double nv = (double)n_ds;
  const double mean = sum / nv;
  const double variance = (sq_sum - sum * sum / nv) / (nv - 1);
  const double threshold = mean + std_mul * sqrt(variance);
  k_sor_flags<<<(unsigned)((n_ds + 255) / 256), 256, 0, s>>>(f->dist, n_ds, threshold, f->flags);
  size_t n_keep = 0;
  CK(compact_flagged(f, ws, f->b, n_ds, out, &n_keep, s));
  L += 4;
  counts[2] = n_keep;
  return cudaGetLastError();
}

}  // namespace ndtb
