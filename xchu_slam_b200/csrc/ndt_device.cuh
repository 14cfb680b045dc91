// ndt_device.cuh -- device helpers shared by the derivative kernels: point transform, query cell,
// hash probe and the neighbour search of every NeighborSearchMethod.
//
// Everything here that decides WHICH voxels a point sees must be bit-identical to the reference
// arithmetic (neighbour sets are a bit-exact parity gate), so it is written with explicit
// round-to-nearest intrinsics (__fmul_rn / __fadd_rn / __fsub_rn are never contracted into FMAs):
// the header gives the same bits whether the including file is compiled with -fmad=false
// (ndt_core.cu) or with FMA contraction enabled (ndt_fast.cu).
#pragma once
#include "ndt_internal.h"

namespace ndtb {

static __constant__ int c_off7[7][3] = {{0, 0, 0}, {1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};  // VGC_I:423-430
// pcl::getAllNeighborCellIndices(): 13 "half" neighbours followed by their negations (VGC_I:413)
static __constant__ int c_off26[26][3] = {
    {-1, -1, -1}, {-1, 0, -1}, {-1, 1, -1}, {0, -1, -1}, {0, 0, -1}, {0, 1, -1}, {1, -1, -1}, {1, 0, -1}, {1, 1, -1},
    {-1, -1, 0}, {0, -1, 0}, {1, -1, 0}, {-1, 0, 0},
    {1, 1, 1}, {1, 0, 1}, {1, -1, 1}, {0, 1, 1}, {0, 0, 1}, {0, -1, 1}, {-1, 1, 1}, {-1, 0, 1}, {-1, -1, 1},
    {1, 1, 0}, {0, 1, 0}, {-1, 1, 0}, {1, 0, 0}};

// ---- hash probe: one 128-bit load per probed slot -----------------------------------------
__device__ __forceinline__ HashSlot load_slot(const HashSlot* p) {
  const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
  HashSlot s;
  s.key = ((unsigned long long)v.y << 32) | v.x;
  s.vidx = (int)v.z;
  s.nr_points = (int)v.w;
  return s;
}
__device__ __forceinline__ bool probe(const VoxelTable& t, int i, int j, int k, HashSlot& out) {
  if (!coord_ok(i) || !coord_ok(j) || !coord_ok(k)) return false;
  const unsigned long long key = pack_key(i, j, k);
  uint32_t h = hash_key(key) & t.mask;
  while (true) {
    const HashSlot s = load_slot(t.slots + h);
    if (s.key == key) { out = s; return true; }
    if (s.key == kEmptyKey) return false;
    h = (h + 1) & t.mask;
  }
}

// Query-time cell: floor(x / leaf_size) (VGC_I:379-381) -- an IEEE float DIVISION, unlike the build key.
__device__ __forceinline__ int query_cell(float x, float leaf) { return (int)floorf(__fdiv_rn(x, leaf)); }

// pcl::transformPointCloud (NDT_I:93, 822, 866): x' = ((m00 x + m01 y) + m02 z) + m03, float, no FMA
__device__ __forceinline__ void xform(const float* T /*row-major 3x4*/, float x, float y, float z, float& ox, float& oy, float& oz) {
  ox = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[0], x), __fmul_rn(T[1], y)), __fmul_rn(T[2], z)), T[3]);
  oy = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[4], x), __fmul_rn(T[5], y)), __fmul_rn(T[6], z)), T[7]);
  oz = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[8], x), __fmul_rn(T[9], y)), __fmul_rn(T[10], z)), T[11]);
}

// ---- cell lookup: dense index when the target has one, hash probe otherwise ----------------------
// Returns the dense code of cell (i, j, k): v >= 0 usable voxel, -2 - v present-but-invalid voxel with
// >= min_points points, -1 nothing a search may return.
__device__ __forceinline__ int cell_code_hash(const VoxelTable& t, int i, int j, int k) {
  HashSlot s;
  if (!probe(t, i, j, k, s)) return -1;
  if (s.nr_points >= t.min_points) return s.vidx;
  if (s.nr_points == -1 && __float_as_int(__ldg(t.centroid + s.vidx).w) >= t.min_points) return -2 - s.vidx;
  return -1;
}
__device__ __forceinline__ int cell_code(const VoxelTable& t, int i, int j, int k) {
  if (t.dense != nullptr) {
    const int a = i - t.dlo[0], b = j - t.dlo[1], c = k - t.dlo[2];
    if (i < t.dlo[0] || j < t.dlo[1] || k < t.dlo[2] || a >= t.dn[0] || b >= t.dn[1] || c >= t.dn[2]) return -1;
    return __ldg(t.dense + ((size_t)c * t.dn[1] + b) * t.dn[0] + a);
  }
  return cell_code_hash(t, i, j, k);
}

// DIRECT7 (VGC_I:419-433): codes of the seven cells in the reference's visiting order
// 0, +x, -x, +y, -y, +z, -z; v[o] >= 0 <=> the reference would return that leaf.
__device__ __forceinline__ void neighbors7(const VoxelTable& t, float xt, float yt, float zt, int* v) {
#pragma unroll
  for (int o = 0; o < 7; ++o) v[o] = -1;
  if (t.slots == nullptr) return;   // empty grid (no target, or the overflow guard of VGC_I:79-84 tripped)
  const int ci = query_cell(xt, t.leaf), cj = query_cell(yt, t.leaf), ck = query_cell(zt, t.leaf);
  if (t.dense != nullptr) {
    // the index is padded by 2 cells: a home cell within 1 cell of the box has all 7 cells in range,
    // any other home cell has no occupied neighbour at all
    if (ci <= t.dlo[0] || cj <= t.dlo[1] || ck <= t.dlo[2] || ci >= t.dlo[0] + t.dn[0] - 1 || cj >= t.dlo[1] + t.dn[1] - 1 ||
        ck >= t.dlo[2] + t.dn[2] - 1)
      return;
    const int sx = t.dn[0], sxy = t.dn[0] * t.dn[1];
    const int* __restrict__ c = t.dense + ((size_t)(ck - t.dlo[2]) * t.dn[1] + (cj - t.dlo[1])) * t.dn[0] + (ci - t.dlo[0]);
    v[0] = __ldg(c); v[1] = __ldg(c + 1); v[2] = __ldg(c - 1); v[3] = __ldg(c + sx); v[4] = __ldg(c - sx);
    v[5] = __ldg(c + sxy); v[6] = __ldg(c - sxy);
    return;
  }
#pragma unroll
  for (int o = 0; o < 7; ++o) {
    const int code = cell_code_hash(t, ci + c_off7[o][0], cj + c_off7[o][1], ck + c_off7[o][2]);
    v[o] = code >= 0 ? code : -1;
  }
}

// Neighbour voxels of a transformed point in the reference's visiting order.
// DIRECT1/7/26: VGC_I:373-442 (found && nr_points >= min_points).
// KDTREE: VGC_H:470-499 over the centroid cloud (every voxel that had >= min_points inserted points,
// flagged-invalid ones included), squared float distance < float(r*r), ascending distance.
__device__ inline int find_neighbors(const VoxelTable& t, int method, float x, float y, float z, int* vout) {
  if (t.slots == nullptr) return 0;
  int cnt = 0;
  if (method == 2) {
    int v[7];
    neighbors7(t, x, y, z, v);
#pragma unroll
    for (int o = 0; o < 7; ++o)
      if (v[o] >= 0) vout[cnt++] = v[o];
    return cnt;
  }
  const int ci = query_cell(x, t.leaf), cj = query_cell(y, t.leaf), ck = query_cell(z, t.leaf);
  if (method == 3) {
    const int code = cell_code(t, ci, cj, ck);
    if (code >= 0) vout[cnt++] = code;
  } else if (method == 1) {
    for (int o = 0; o < 26; ++o) {
      const int code = cell_code(t, ci + c_off26[o][0], cj + c_off26[o][1], ck + c_off26[o][2]);
      if (code >= 0) vout[cnt++] = code;
    }
  } else {
    const double rr = (double)t.leaf * (double)t.leaf;  // radius = resolution_ (NDT_I:233), r*r in double then float
    const float r2 = (float)rr;
    float d2s[27];
    unsigned long long keys[27];
    for (int dz = -1; dz <= 1; ++dz)
      for (int dy = -1; dy <= 1; ++dy)
        for (int dx = -1; dx <= 1; ++dx) {
          const int code = cell_code(t, ci + dx, cj + dy, ck + dz);
          if (code == -1) continue;
          const int vidx = code >= 0 ? code : -2 - code;
          const float4 c = __ldg(t.centroid + vidx);
          const float ex = __fsub_rn(x, c.x), ey = __fsub_rn(y, c.y), ez = __fsub_rn(z, c.z);
          const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)), __fmul_rn(ez, ez));
          if (!(d2 < r2)) continue;
          // insertion sort by (d2, cell key): ascending distance, ties by ascending linear index
          const unsigned long long key = pack_key(ci + dx, cj + dy, ck + dz);
          int pos = cnt;
          while (pos > 0 && (d2s[pos - 1] > d2 || (d2s[pos - 1] == d2 && keys[pos - 1] > key))) {
            d2s[pos] = d2s[pos - 1]; keys[pos] = keys[pos - 1]; vout[pos] = vout[pos - 1]; --pos;
          }
          d2s[pos] = d2; keys[pos] = key; vout[pos] = vidx;
          ++cnt;
        }
  }
  return cnt;
}

constexpr int kPartialStride = 44;  // 43 accumulators + pair count: one per-block partial of a derivative launch

}  // namespace ndtb
