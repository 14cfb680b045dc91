// ndt_internal.h -- device data layout and kernel launch interfaces shared by the .cu files.
//
// HBM layout:
//   points      float4[n]            x, y, z, (unused)            16 B / point, coalesced
//   HashSlot    16 B                 {packed cell key, voxel index, nr_points}; open addressing,
//                                    linear probing, load factor <= 0.25, one 128-bit load / probe
//   VoxelRec    64 B, 64 B-aligned   {mean f64[3], icov f32[9], nr_points}: what one (point, voxel)
//                                    pair of the derivative kernel reads -- four 128-bit loads
//   VoxelAcc    per voxel running sums (n, sum x f64[3], sum xx^T f64[6], f32 centroid sum[3]) kept so
//                                    that an incremental append continues the SAME sequential sums
//                                    a full rebuild would form (reference VGC_I:233-243)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ndtb {

constexpr int kCoordBias = 1 << 20;               // cell coordinates must lie in (-2^20, 2^20)
constexpr uint64_t kEmptyKey = 0xFFFFFFFFFFFFFFFFull;
constexpr int kAccN = 43;                         // score + g[6] + H[36]

__host__ __device__ inline uint64_t pack_key(int i, int j, int k) {
  return ((uint64_t)(uint32_t)(k + kCoordBias) << 42) | ((uint64_t)(uint32_t)(j + kCoordBias) << 21) |
         (uint64_t)(uint32_t)(i + kCoordBias);
}
__host__ __device__ inline bool coord_ok(int c) { return c > -kCoordBias && c < kCoordBias; }
__host__ __device__ inline void unpack_key(uint64_t key, int& i, int& j, int& k) {
  i = (int)(key & 0x1FFFFF) - kCoordBias;
  j = (int)((key >> 21) & 0x1FFFFF) - kCoordBias;
  k = (int)((key >> 42) & 0x1FFFFF) - kCoordBias;
}
// Spatial hash of the three biased 21-bit cell fields: three 32-bit multiplies and one xor-shift.
__host__ __device__ inline uint32_t hash_fields(uint32_t fi, uint32_t fj, uint32_t fk) {
  uint32_t h = fi * 0x9E3779B1u + fj * 0x85EBCA77u + fk * 0xC2B2AE3Du;
  h ^= h >> 15;
  return h;
}
__host__ __device__ inline uint32_t hash_key(uint64_t k) {
  return hash_fields((uint32_t)(k & 0x1FFFFF), (uint32_t)((k >> 21) & 0x1FFFFF), (uint32_t)((k >> 42) & 0x1FFFFF));
}

struct __align__(16) HashSlot {
  unsigned long long key;
  int vidx;
  int nr_points;   // reference Leaf::nr_points after pass 2 (-1 = flagged invalid)
};

struct __align__(64) VoxelRec {
  double mean[3];  // Leaf::mean_
  float icov[9];   // Leaf::icov_ narrowed to float (NDT_I:502), row-major
  int nr_points;
};

// 48 B record of the fast derivative kernel: exactly the 48 B per (point, voxel) pair of the
// algorithmic-bytes model (mean 24 B + symmetric inverse covariance 24 B), three 128-bit loads.
// The f64 mean is carried as an unevaluated float sum hi + lo (48 significant bits) so that
// x - mean is formed in two float subtractions to the same float the reference's double
// subtraction rounds to (NDT_I:257-260, 500).
// (A 64-byte, 64-B aligned record -- one 256-bit + one 128-bit load per pair -- was measured 3 % slower on B200:
// profiles/r02/r02_c3_variants.txt; the 48-byte record stays.)
struct __align__(16) FastRec {
  float mh[3];     // (float)mean
  float ml[3];     // (float)(mean - (double)mh)
  float c[6];      // icov (float) upper triangle: 00 01 02 11 12 22
};
static_assert(sizeof(FastRec) == 48, "FastRec layout");

struct VoxelTable {          // all pointers device
  HashSlot* slots;
  uint32_t mask;             // capacity - 1 (capacity is a power of two)
  VoxelRec* rec;
  FastRec* frec;             // same voxels, layout of the fast kernel
  // Dense cell -> voxel index over the target's bounding box padded by 2 cells (nullptr when the box
  // exceeds dense_max_cells: the hash table is then the only index).  Code per cell: v >= 0 voxel v
  // usable by the DIRECT searches (nr_points >= min_points); -2 - v: voxel v holds >= min_points
  // points but is flagged invalid (seen only by the radius search, VGC_H:485-497); -1 otherwise.
  int* dense;
  int dlo[3];                // cell coordinate of dense[0] (= min_b - 2)
  int dn[3];                 // dense dimensions in cells (x fastest)
  double* icov64;            // [v][9] Leaf::icov_ in f64, read only by the FP64 computeHessian / calculateScore kernels
  float4* centroid;          // x, y, z (f32, VGC_I:289), w = bit pattern of inserted point count
  unsigned long long* vkey;  // packed absolute cell of voxel v
  int* vn;                   // inserted point count
  double* vsum;              // [v][3]  sum x
  double* vsxx;              // [v][6]  Identity + sum xx^T (xx xy xz yy yz zz): Leaf::cov_ starts at I (VGC_H:101)
  float* vcsum;              // [v][3]  f32 running centroid sum
  float leaf;                // resolution
  float inv_leaf;            // 1.0f / leaf (pcl::VoxelGrid::setLeafSize)
  int min_points;
  double eig_mult;
};

// Everything one computeDerivatives launch needs to know about one registration.
struct EvalReq {
  float T[12];               // row-major 3x4 float transform applied to the source
  float jang[8][3];          // NDT_I:329-336
  float hang[15][3];         // NDT_I:362-396
  double jang_d[8][3];       // f64 twins for computeHessian (NDT_I:319-326, 340-359)
  double hang_d[15][3];
  double d1, d2, d3;         // gauss constants, NDT_I:80-87
  int mode;                  // 0 idle, 1 derivatives + hessian, 2 derivatives only, 3 computeHessian (f64),
                             // 4 calculateScore
  int search_method;
};

enum Phase { PH_INIT = 0, PH_MT_FIRST = 1, PH_MT_LOOP = 2, PH_MT_HESS = 3, PH_DONE = 4 };

struct AlignState {
  EvalReq req;
  // problem
  int n_src, target, first_block, n_blocks;
  const float4* src;         // this registration's source points (pool slice or a borrowed device buffer)
  double step_size, trans_eps;
  int max_iter;
  float resolution;
  // Newton state (NDT_I:73-164)
  double p[6], score, g[6], H[36], dir[6], dir_norm;
  int nr_iterations, converged, done, phase, n_evals;
  float final_T[16];         // column-major, final_transformation_
  double trans_probability;
  // More-Thuente state (NDT_I:760-916)
  double phi_0, d_phi_0, a_l, f_l, g_l, a_u, f_u, g_u, a_t, x_t[6];
  int step_iterations, open_interval, interval_converged;
  long long pairs;           // (point, voxel) pairs of the last evaluation (diagnostics)
  long long pairs_total;     // ... summed over all evaluations of this registration
  // What decides whether a registration run with the factorised kernel is replayed with the term-by-term one (ndt_api.cu
  // run_items): `borderline` = a discrete decision of the Newton loop (the |step| < eps convergence test NDT_I:152-157, the
  // descent-direction flip NDT_I:774-784) was taken within a relative 1e-3 of its threshold, so a 1e-7 difference in score /
  // gradient could have taken it the other way; n_rise = line-search steps that grew again after the step had dropped below
  // step_size (the iteration is not contracting); last_norm = |delta_p| of the last Newton solve; prev_at = previous step.
  int borderline, n_rise;
  int abort_iters;           // > 0 (first pass of a hybrid batch): leave the loop unconverged once nr_iterations exceeds it
  double last_norm, prev_at, min_at;
  float norm_hist[32];       // |delta_p| of the Newton solve of iteration k (diagnostics: ndtb_batch_get_step_history)
};

constexpr int kTraceRec = 52;  // doubles per trace record (see ndtb_get_trace)

// ---- launch wrappers (definitions in the .cu files) --------------------------------------
struct BuildWorkspace;  // opaque; owns CUB temp storage and sort buffers
BuildWorkspace* build_ws_create();
void build_ws_destroy(BuildWorkspace* ws);

// Table storage management
struct TableStorage {
  VoxelTable t{};
  size_t vox_cap = 0;        // capacity of per-voxel arrays
  size_t slot_cap = 0;       // hash capacity (power of two)
  int n_vox = 0;             // voxels in use (host mirror)
  bool valid = false;        // false = empty grid (no target / overflow guard VGC_I:79-84)
  int min_b[3] = {0, 0, 0}, max_b[3] = {0, 0, 0};
  float mn[3], mx[3];        // float bounding box of all inserted points
  bool have_bbox = false;
  long long dense_max_cells = 0;   // 0: no dense index
  size_t dense_cap = 0;            // ints allocated for t.dense
  int* dense_buf = nullptr;        // the allocation (t.dense is nullptr while the box is too large)
};
cudaError_t table_reserve(TableStorage& ts, size_t n_vox_needed, cudaStream_t s, long long* launches);
void table_free(TableStorage& ts);
cudaError_t table_clear(TableStorage& ts, cudaStream_t s, long long* launches);
// Merge n points (device float4, input order preserved) into the table: new voxels appended,
// touched voxels re-finalised.  Equivalent to a full rebuild over (old points ++ new points).
cudaError_t table_append(TableStorage& ts, BuildWorkspace* ws, const float4* pts, size_t n, cudaStream_t s,
                         long long* launches);

cudaError_t launch_derivatives(const AlignState* states, const VoxelTable* tables, const float4* src_pool,
                               const int2* blockmap, int n_blocks, int chunk, double* partials, int method,
                               bool with_gradient_only, cudaStream_t s);
// Fast variant of launch_derivatives: same outputs within the 1e-6 parity tolerance, per-point
// factorised math with FMAs (ndt_fast.cu).  Writes the same 44-double per-block partials.
cudaError_t launch_derivatives_fast(const AlignState* states, const VoxelTable* tables, const float4* src_pool,
                                    const int2* blockmap, int n_blocks, int chunk, double* partials, int method,
                                    bool with_gradient_only, cudaStream_t s);
cudaError_t launch_hessian_f64(const AlignState* states, const VoxelTable* tables, const float4* src_pool,
                               const int2* blockmap, int n_blocks, int chunk, double* partials, cudaStream_t s);
cudaError_t launch_newton(AlignState* states, int n_items, const double* partials, double* trace, int trace_cap,
                          int* n_active, cudaStream_t s, unsigned long long cond_handle = 0, int eval_cap = 0);
cudaError_t launch_neighbors(const VoxelTable* table, const float* xyz, size_t n, int method, int* cnt, int* ijk,
                             cudaStream_t s);
cudaError_t launch_transform(const float4* src, size_t n, const float* T16_colmajor_dev, float4* out, cudaStream_t s);
// One keyframe of a loop-closure submap: n points at src go to out[offset ..) through the row-major 3x4 float transform T.
struct SubmapSeg { const float4* src; long long offset; size_t n; float T[12]; };
cudaError_t launch_submap_gather(const SubmapSeg* segs_dev, int n_seg, size_t max_n, float4* out, cudaStream_t s);
cudaError_t launch_prepare_eval(AlignState* st, const float* T16_dev, const double* p6_dev, int mode, cudaStream_t s);
cudaError_t launch_pull(void* dst_dev, const void* src_mapped_dev, size_t bytes, cudaStream_t s);
cudaError_t launch_replan(AlignState* states, int n_items, int2* blockmap, int* plan_mapped_host, int target_blocks, cudaStream_t s);
// states[idx[k]] = src_mapped[idx[k]] for k < n (selected registration states pulled out of mapped host memory)
// host evaluation of the float pose chain the device uses (NDT_H:210-229)
void pose_to_matrix_host(const double p[6], float T[16]);
cudaError_t launch_pull_items(AlignState* states, const AlignState* src_mapped_dev, const int* idx_mapped_dev, int n, cudaStream_t s);
cudaError_t launch_set_int(int* dst, int v, cudaStream_t s);
cudaError_t launch_widen_xyz(const float* packed_dev, float4* dst, size_t n, cudaStream_t s);
cudaError_t launch_publish(const int* src_dev, int* dst_mapped_host, cudaStream_t s);
cudaError_t launch_prefix_equal(const float4* a, const float4* b, size_t n, int* flag, cudaStream_t s);

// pcl::VoxelGrid centroid downsample (odom_node's downSizeFilterLocalmap); out needs room for n points
cudaError_t voxelgrid_downsample(BuildWorkspace* ws, const float4* in, size_t n, float leaf, float4* out, size_t* n_out,
                                 cudaStream_t s, long long* launches);

// filter_node's front end (range gate, pcl::VoxelGrid, pcl::StatisticalOutlierRemoval) on device records
struct FilterWorkspace;
FilterWorkspace* filter_ws_create();
void filter_ws_destroy(FilterWorkspace* f);
cudaError_t filter_scan(FilterWorkspace* f, BuildWorkspace* ws, const float4* raw, size_t n, double r_min, double r_max, float leaf,
                        int mean_k, double std_mul, float4* out, size_t counts[3], cudaStream_t s, long long* launches);

// fitness (exact 1-NN over the raw target, grid of key-sorted points)
struct NNGridStorage;
NNGridStorage* nn_create();
void nn_destroy(NNGridStorage* g);
cudaError_t nn_build(NNGridStorage* g, BuildWorkspace* ws, const float4* pts, size_t n, float cell, const float* bbox,
                     cudaStream_t s, long long* launches);
// Morton order of the sources of a batch (ndt_api.cu run_items): every point of every source -> (source ordinal, 18-bit Morton
// code of its 2 m cell) -> one radix sort of (key, point) pairs -> `out`, where source k occupies [first_k, first_k + n_k) as in
// the input numbering, its points now in Morton order.  Grow-only workspace.
struct SrcDesc { const float4* p; unsigned int first, n; };
struct SourceOrderWs {
  SrcDesc* desc = nullptr; size_t desc_cap = 0;
  unsigned int *k_in = nullptr, *k_out = nullptr; float4* v_in = nullptr; size_t cap = 0;
  void* temp = nullptr; size_t temp_bytes = 0;
};
void source_order_free(SourceOrderWs& w);
cudaError_t source_order_sort(SourceOrderWs& w, const SrcDesc* desc_host, int n_src, size_t total, float4* out, cudaStream_t s,
                              long long* launches);
cudaError_t d2_reduce(const float* d2, size_t n, double max_range, double* out2_dev, cudaStream_t s);   // {sum, count} of d2 <= max_range
cudaError_t nn_query(NNGridStorage* g, const float4* src, size_t n, const float* T16_dev, float* d2_out,
                     cudaStream_t s);
// getFitnessScore of every registration of a batch that shares the grid's target (items_dev: ids into states)
cudaError_t nn_fitness_batch(NNGridStorage* g, const AlignState* states, const int* items_dev, int n_group, double max_range,
                             int blocks_per_item, double* partial, cudaStream_t s);
cudaError_t fitness_finish(const double* partial, int n_items, int blocks_per_item, double* out, cudaStream_t s);
constexpr int kPhaseStart = 10, kPhaseEvalOnly = 11;   // AlignState::phase values used by the host

}  // namespace ndtb
