// ndt_api.cu -- host side of libndt_b200.so: the registration engine (targets, sources, batched
// Newton loop driver) and the extern "C" ABI declared in include/ndt_b200/ndt_c_api.h.
// Host logic mirrors pcl::Registration::align + pclomp::NormalDistributionsTransform
// (reference xchu_mapping/include/pclomp/ndt_omp.h:110-278, ndt_omp_impl.hpp:73-164).
#include <algorithm>
#include <atomic>
#include <functional>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/ndt_b200/ndt_c_api.h"
#include "ndt_internal.h"

using namespace ndtb;

namespace {

thread_local std::string g_create_error;

#define CKE(x)                                                                                       \
  do {                                                                                               \
    cudaError_t e_ = (x);                                                                            \
    if (e_ != cudaSuccess) { return fail(std::string(#x) + ": " + cudaGetErrorString(e_)); }         \
  } while (0)

struct Target {
  float4* pts = nullptr;       // owned allocation (cap points)
  const float4* use = nullptr; // what the build reads: pts, or a borrowed device buffer
  size_t n = 0, cap = 0;
  TableStorage ts;
  bool built = false;
  cudaEvent_t uploaded = nullptr;   // recorded on the copy stream after the H2D of this target
  bool pending_upload = false;
};
struct Source { long long offset; size_t n; const float4* borrowed; };

// Grow-only pinned + mapped host buffer: the host fills it, a kernel pulls it (ndt_core.cu launch_pull).
struct PinnedBuf {
  void* h = nullptr;
  void* d = nullptr;    // device alias of h
  size_t cap = 0;
  cudaError_t reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (h) cudaFreeHost(h);
    h = d = nullptr;
    cap = 0;
    const size_t nc = ((bytes + bytes / 4 + 4095) / 4096) * 4096;
    cudaError_t e = cudaHostAlloc(&h, nc, cudaHostAllocMapped);
    if (e != cudaSuccess) return e;
    e = cudaHostGetDevicePointer(&d, h, 0);
    if (e != cudaSuccess) return e;
    cap = nc;
    return cudaSuccess;
  }
  void release() { if (h) cudaFreeHost(h); h = d = nullptr; cap = 0; }
};

// Eigen 3.3 MatrixBase::eulerAngles(0,1,2) on the float rotation block (NDT_I:102)
static void euler_012(const float T[16], float out[3]) {
  auto m = [&](int r, int c) { return T[c * 4 + r]; };
  const float pi = (float)M_PI;
  float r0 = std::atan2(m(1, 2), m(2, 2));
  const float c2 = std::sqrt(m(0, 0) * m(0, 0) + m(0, 1) * m(0, 1));
  float r1;
  if (r0 > 0.f) { r0 -= pi; r1 = std::atan2(-m(0, 2), -c2); }
  else r1 = std::atan2(-m(0, 2), c2);
  const float s1 = std::sin(r0), c1 = std::cos(r0);
  const float r2 = std::atan2(s1 * m(2, 0) - c1 * m(1, 0), c1 * m(1, 1) - s1 * m(2, 1));
  out[0] = -r0; out[1] = -r1; out[2] = -r2;
}

struct Engine {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = true;
  // Concurrent table builds of a batch: a few host threads, each with its own stream and workspace, so the small
  // per-target kernels (one thread per voxel, radix-sort passes) of different submaps overlap on the device.
  static constexpr int kBuildLanes = 16;    // upper bound; build_lanes() of them are used
  cudaStream_t bstream[kBuildLanes] = {};
  BuildWorkspace* bws[kBuildLanes] = {};
  cudaEvent_t bdone[kBuildLanes] = {};
  // A table build is ~25 short launches and two blocking reads (bounding box, voxel count): one lane leaves the GPU idle most
  // of the time, so several host threads build different targets on their own streams.
  static int build_lanes() {
    static const int n = [] {
      int v = 12;   // 4: 8.3 ms, 8: 6.9 ms, 12: 6.5 ms, 16: 6.6 ms per 64 one-million-point submaps (B200)
      if (const char* ev = getenv("NDTB_BUILD_LANES")) v = atoi(ev);   // tuning knob
      return v < 1 ? 1 : (v > kBuildLanes ? kBuildLanes : v);
    }();
    return n;
  }
  cudaEvent_t bstart = nullptr;
  cudaStream_t copy_stream = nullptr;      // batch uploads from host memory run here, overlapped with table builds
  struct PackedStage { void* p = nullptr; size_t cap = 0; };
  PackedStage stage_main, stage_copy;      // device landing zones of packed (12-byte) uploads, one per stream
  cudaEvent_t sources_uploaded = nullptr;
  bool sources_pending = false;
  std::string err;
  ndtb_params params;
  BuildWorkspace* ws = nullptr;
  NNGridStorage* nn = nullptr;
  bool nn_valid = false;
  // getFitnessScore's NN grid of a NEW target is built on a side stream under the next align (the caller has asked for a fitness
  // score before: odom_node does after every align, ODOM:266-280) instead of inside the fitness call
  bool fitness_used = false, nn_pending = false;
  cudaStream_t nn_stream = nullptr;
  cudaEvent_t ev_nn_dep = nullptr, ev_nn_ready = nullptr;
  BuildWorkspace* ws_nn = nullptr;
  std::vector<Target> targets;
  int n_targets_live = 0;
  VoxelTable* d_tables = nullptr;
  size_t d_tables_cap = 0;
  float4* src_pool = nullptr;
  size_t src_cap = 0, src_used = 0;
  std::vector<Source> sources;
  AlignState* d_states = nullptr;
  size_t states_cap = 0;
  PinnedBuf st_states, st_bmap, st_tables;   // staging of the run set-up
  AlignState* h_states = nullptr;            // = st_states.h
  int2* d_blockmap = nullptr;
  size_t blockmap_cap = 0;
  double* d_partials = nullptr;
  size_t partials_cap = 0;
  double* d_trace = nullptr;
  size_t trace_doubles = 0;
  int trace_cap = 0;
  int* d_active = nullptr;
  int* h_active = nullptr;       // pinned + mapped: the device publishes the unfinished-registration count into it
  int* h_active_dev = nullptr;
  float* d_T16 = nullptr;
  double* d_p6 = nullptr;
  float4* d_output = nullptr;
  size_t output_cap = 0;
  float* d_d2 = nullptr;
  size_t d2_cap = 0;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  std::vector<cudaEvent_t> deriv_events;
  ndtb_batch_stats stats{};
  double last_build_ms = 0, last_align_ms = 0;
  // single-handle state (pcl::Registration members)
  ndtb_result last{};
  bool have_result = false;
  std::vector<double> h_trace;
  int h_trace_n = 0;
  // odom_node's localmap on the device (ndtb_localmap_*): localmap, tmp_map, the cloud appended last
  struct DevCloud { float4* p = nullptr; size_t n = 0, cap = 0; };
  DevCloud lm, lm_tmp, lm_last;
  float4* lm_scratch = nullptr;
  size_t lm_scratch_cap = 0;
  size_t lm_snapshot = 0;
  bool lm_snapshot_valid = false;
  unsigned long long lm_gen = 1;     // bumped whenever the localmap's contents stop being an extension of what they were
  unsigned long long tgt_gen = 0;    // localmap generation the target's points were taken from (0: not from the localmap)
  size_t tgt_from_lm = 0;            // leading localmap points already merged into the target
  // the keyframe-count window (ndtb_window_*): per id one transformed + downsampled keyframe cloud
  std::vector<DevCloud> win;           // index = id; p == nullptr once dropped
  std::vector<DevCloud> win_free;      // buffers of dropped clouds, reused
  float4* win_cat = nullptr;           // concatenation scratch
  size_t win_cat_cap = 0;
  std::vector<int> last_item_target;   // targets of the items of the last run (ndtb_batch_fitness)
  double* d_fit_partial = nullptr;
  size_t fit_partial_cap = 0;
  double* d_fit_out = nullptr;
  size_t fit_out_cap = 0;
  int* d_fit_items = nullptr;
  size_t fit_items_cap = 0;
  FilterWorkspace* flt = nullptr;    // filter_node front end (ndtb_filter_scan)
  float4* flt_raw = nullptr;
  size_t flt_raw_cap = 0;
  // loop-closure keyframe store (ndtb_batch_add_keyframe / ndtb_batch_add_submap)
  struct KeyFrame { float4* p = nullptr; size_t n = 0; };
  std::vector<KeyFrame> kfs;
  SubmapSeg* d_segs = nullptr;
  size_t segs_cap = 0;
  float4* kf_cat = nullptr;          // stacked, not yet downsampled submap
  size_t kf_cat_cap = 0;
  float4* lane_cat[16] = {};         // the same per build lane (ndtb_batch_add_submaps)
  size_t lane_cat_cap[16] = {};
  SubmapSeg* lane_segs[16] = {};
  NNGridStorage* lane_nn[16] = {};   // one NN grid per build lane (ndtb_batch_fitness over distinct targets)
  size_t lane_segs_cap[16] = {};
  double* d_fit2 = nullptr;          // {sum, count} of getFitnessScore on the device
  PinnedBuf st_d2;                   // host landing zone of the per-point squared NN distances (getFitnessScore)
  PinnedBuf st_src;                  // host staging of strided sources (x y z intensity packed to 16 B)

  int fail(const std::string& m) { err = m; return 1; }

  // Which derivative kernel runs.  DIRECT26 skips the point's own voxel; Newton on that objective is not
  // contractive and float-level differences between the factorised and the term-by-term arithmetic grow
  // from iterate to iterate (centimetres after ~10 iterations), so that mode always gets the strict kernel.
  // strict_math = -1 (default): the strict kernel everywhere (a single handle is latency-bound, the term-by-term kernel costs
  // nothing measurable there; a batch pays 1.9x the hybrid mode's time for results that are the reference's by construction).
  bool single = false;
  bool use_strict() const { return kernel_mode() == KM_STRICT; }

  int init(int dev) {
    device = dev;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) return fail("no CUDA device available (this library has no CPU fallback)");
    if (dev < 0 || dev >= count) return fail("invalid device ordinal");
    CKE(cudaSetDevice(dev));
    CKE(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
    if (sm_count < 1) sm_count = 1;
    CKE(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    CKE(cudaStreamCreateWithFlags(&copy_stream, cudaStreamNonBlocking));
    CKE(cudaEventCreateWithFlags(&sources_uploaded, cudaEventDisableTiming));
    ws = build_ws_create();
    if (!ws) return fail("workspace allocation failed");
    nn = nn_create();
    CKE(cudaMalloc(&d_active, sizeof(int)));
    CKE(cudaHostAlloc(&h_active, 4 * sizeof(int), cudaHostAllocMapped));   // [0] unfinished registrations, [1..2] k_replan's plan
    CKE(cudaHostGetDevicePointer(&h_active_dev, h_active, 0));
    CKE(cudaMalloc(&d_T16, 16 * sizeof(float)));
    CKE(cudaMalloc(&d_p6, 6 * sizeof(double)));
    for (auto& e : ev) CKE(cudaEventCreate(&e));
    ndtb_default_params(&params);
    return 0;
  }
  void destroy() {
    cudaSetDevice(device);
    for (auto& t : targets) { cudaFree(t.pts); table_free(t.ts); if (t.uploaded) cudaEventDestroy(t.uploaded); }
    for (int g = 0; g < kBuildLanes; ++g) {
      if (bws[g]) build_ws_destroy(bws[g]);
      if (bstream[g]) cudaStreamDestroy(bstream[g]);
      if (bdone[g]) cudaEventDestroy(bdone[g]);
    }
    if (bstart) cudaEventDestroy(bstart);
    cudaFree(stage_main.p); cudaFree(stage_copy.p);
    source_order_free(so_ws); cudaFree(src_sorted);
    if (copy_stream) cudaStreamDestroy(copy_stream);
    if (sources_uploaded) cudaEventDestroy(sources_uploaded);
    targets.clear();
    build_ws_destroy(ws);
    if (ws_nn) build_ws_destroy(ws_nn);
    if (nn_stream) cudaStreamDestroy(nn_stream);
    if (ev_nn_dep) cudaEventDestroy(ev_nn_dep);
    if (ev_nn_ready) cudaEventDestroy(ev_nn_ready);
    nn_destroy(nn);
    st_states.release(); st_bmap.release(); st_tables.release(); st_idx.release();
    cudaFree(d_tables); cudaFree(src_pool); cudaFree(d_states); cudaFree(d_blockmap); cudaFree(d_partials);
    cudaFree(d_trace); cudaFree(d_active); cudaFreeHost(h_active); cudaFree(d_T16); cudaFree(d_p6);
    cudaFree(d_output); cudaFree(d_d2); cudaFree(d_fit2);
    destroy_align_graph();
    cudaFree(ag.d_state); cudaFree(ag.d_bmap); cudaFree(ag.d_partials); cudaFree(ag.d_trace); cudaFree(ag.d_active);
    ag.h_in.release(); ag.h_out.release();
    cudaFree(lm.p); cudaFree(lm_tmp.p); cudaFree(lm_last.p); cudaFree(lm_scratch);
    for (auto& c : win) cudaFree(c.p);
    for (auto& c : win_free) cudaFree(c.p);
    cudaFree(win_cat);
    st_src.release(); st_d2.release();
    filter_ws_destroy(flt); cudaFree(flt_raw);
    cudaFree(d_fit_partial); cudaFree(d_fit_out); cudaFree(d_fit_items);
    for (auto& k : kfs) cudaFree(k.p);
    kfs.clear();
    cudaFree(d_segs); cudaFree(kf_cat);
    for (int g = 0; g < 16; ++g) { cudaFree(lane_cat[g]); cudaFree(lane_segs[g]); nn_destroy(lane_nn[g]); }
    for (auto& e : ev) if (e) cudaEventDestroy(e);
    for (auto& e : deriv_events) cudaEventDestroy(e);
    if (stream && own_stream) cudaStreamDestroy(stream);
  }

  // copy n records (xyz = first three floats) into a float4 device array
  int upload_points(float4* dst, const void* pts, size_t n, size_t stride, bool on_device, cudaStream_t stream_) {
    cudaStream_t stream = stream_;
    if (n == 0) return 0;
    if (stride < 12) return fail("stride_bytes must be >= 12");
    const cudaMemcpyKind kind = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    if (stride == 16) {
      CKE(cudaMemcpyAsync(dst, pts, n * 16, kind, stream));
    } else if (stride > 16) {
      CKE(cudaMemcpy2DAsync(dst, 16, pts, stride, 16, n, kind, stream));
    } else if (stride == 12) {
      // packed xyz: one contiguous copy (3/4 of the bytes of a 16-byte upload) into a staging buffer, widened by a kernel
      // behind it on the same stream -- a 2-D copy of a million 12-byte rows runs at a fraction of the link rate
      PackedStage& ps = (stream == copy_stream) ? stage_copy : stage_main;
      if (n * 12 > ps.cap) {
        CKE(cudaStreamSynchronize(stream));      // the buffer may still feed a widen kernel
        cudaFree(ps.p);
        ps.p = nullptr; ps.cap = 0;
        const size_t nc = n * 12 + n * 3 + 4096;
        CKE(cudaMalloc(&ps.p, nc));
        ps.cap = nc;
      }
      CKE(cudaMemcpyAsync(ps.p, pts, n * 12, kind, stream));
      CKE(launch_widen_xyz(static_cast<const float*>(ps.p), dst, n, stream));
    } else {  // 13..15 bytes: widen on the way
      CKE(cudaMemsetAsync(dst, 0, n * 16, stream));
      CKE(cudaMemcpy2DAsync(dst, 16, pts, stride, 12, n, kind, stream));
    }
    return 0;
  }

  int reserve_target(Target& t, size_t n) {
    if (n <= t.cap) return 0;
    float4* q = nullptr;
    const size_t nc = n + n / 4 + 1024;
    CKE(cudaMalloc(&q, nc * sizeof(float4)));
    if (t.pts && t.n) CKE(cudaMemcpyAsync(q, t.pts, t.n * sizeof(float4), cudaMemcpyDeviceToDevice, stream));
    CKE(cudaStreamSynchronize(stream));
    cudaFree(t.pts);
    t.pts = q;
    t.cap = nc;
    return 0;
  }

  void table_config(Target& t) {
    t.ts.t.leaf = params.resolution;
    t.ts.t.inv_leaf = 1.0f / params.resolution;   // pcl::VoxelGrid::setLeafSize: inverse_leaf_size_ = 1 / leaf (float)
    t.ts.t.min_points = params.min_points;
    t.ts.t.eig_mult = params.eig_mult;
    t.ts.dense_max_cells = params.dense_max_cells;
  }

  // full (re)build of a target's voxel table from its resident points: init(), NDT_H:271-278
  int build_target(Target& t) {
    table_config(t);
    if (t.pending_upload) { CKE(cudaStreamWaitEvent(stream, t.uploaded, 0)); t.pending_upload = false; }
    t.built = false;
    tables_dirty = true;
    CKE(table_clear(t.ts, stream, &stats.kernel_launches));
    CKE(table_append(t.ts, ws, t.use ? t.use : t.pts, t.n, stream, &stats.kernel_launches));   // n == 0: an empty grid, as the reference's empty leaves_
    t.built = true;
    return 0;
  }

  // Builds every target that needs it.  One target: on the engine's stream.  Several: kBuildLanes host threads, each
  // driving its own stream / workspace (table_append synchronises its stream three times per target; a thread per
  // lane keeps those waits from serialising the lanes).  The engine's stream then waits for every lane.
  int ensure_lanes(int lanes) {
    if (!bstart) CKE(cudaEventCreateWithFlags(&bstart, cudaEventDisableTiming));
    for (int g = 0; g < lanes; ++g) {
      if (!bstream[g]) CKE(cudaStreamCreateWithFlags(&bstream[g], cudaStreamNonBlocking));
      if (!bdone[g]) CKE(cudaEventCreateWithFlags(&bdone[g], cudaEventDisableTiming));
      if (!bws[g]) { bws[g] = build_ws_create(); if (!bws[g]) return fail("workspace allocation failed"); }
    }
    return 0;
  }
  // `overlap` (host-only work of the caller, e.g. filling the registration states) runs on the calling thread while the other
  // lanes are already building; targets are handed out dynamically, so the calling thread joins in late at no cost.
  int build_pending_targets(const std::function<void()>& overlap = std::function<void()>()) {
    std::vector<Target*> todo;
    // (a target without points -- an empty submap -- is "built" into an empty grid, so that a slot reused after
    // ndtb_batch_clear never presents the previous cloud's voxel table)
    for (int k = 0; k < (single ? (int)targets.size() : n_targets_live); ++k)
      if (!targets[k].built) todo.push_back(&targets[k]);
    if (todo.size() <= 1) {
      if (overlap) overlap();
      return todo.empty() ? 0 : build_target(*todo[0]);
    }
    const int lanes = (int)std::min<size_t>((size_t)build_lanes(), todo.size());
    if (ensure_lanes(lanes)) return 1;
    CKE(cudaEventRecord(bstart, stream));
    std::string errs[kBuildLanes];
    long long launches[kBuildLanes] = {};
    std::atomic<size_t> next{0};
    auto work = [&](int g) {
      if (cudaSetDevice(device) != cudaSuccess) { errs[g] = "cudaSetDevice failed"; return; }
      cudaStream_t bs = bstream[g];
      cudaError_t ce = cudaStreamWaitEvent(bs, bstart, 0);
      for (size_t k = next.fetch_add(1); k < todo.size() && ce == cudaSuccess; k = next.fetch_add(1)) {
        Target& t = *todo[k];
        table_config(t);
        if (t.pending_upload) { ce = cudaStreamWaitEvent(bs, t.uploaded, 0); t.pending_upload = false; }
        if (ce == cudaSuccess) ce = table_clear(t.ts, bs, &launches[g]);
        if (ce == cudaSuccess) ce = table_append(t.ts, bws[g], t.use ? t.use : t.pts, t.n, bs, &launches[g]);
        t.built = (ce == cudaSuccess);
        if (ce != cudaSuccess) t.ts.valid = false;
      }
      if (ce == cudaSuccess) ce = cudaEventRecord(bdone[g], bs);
      if (ce != cudaSuccess) errs[g] = std::string("table build: ") + cudaGetErrorString(ce);
    };
    std::vector<std::thread> th;
    for (int g = 1; g < lanes; ++g) th.emplace_back(work, g);
    if (overlap) overlap();
    work(0);
    for (auto& x : th) x.join();
    for (int g = 0; g < lanes; ++g) {
      if (!errs[g].empty()) return fail(errs[g]);
      stats.kernel_launches += launches[g];
      CKE(cudaStreamWaitEvent(stream, bdone[g], 0));
    }
    return 0;
  }

  int sync_tables() {
    if (targets.size() > d_tables_cap) {
      cudaFree(d_tables);
      d_tables = nullptr;
      d_tables_cap = targets.size() + 8;
      CKE(cudaMalloc(&d_tables, ((d_tables_cap * sizeof(VoxelTable) + 15) / 16) * 16));
    }
    if (targets.empty()) return 0;
    CKE(cudaStreamSynchronize(stream));   // the staging buffer may still be read by the previous pull
    CKE(st_tables.reserve(targets.size() * sizeof(VoxelTable)));
    VoxelTable* h = static_cast<VoxelTable*>(st_tables.h);
    for (size_t i = 0; i < targets.size(); ++i) h[i] = targets[i].ts.t;
    CKE(launch_pull(d_tables, st_tables.d, targets.size() * sizeof(VoxelTable), stream));
    return 0;
  }

  // where: 0 host memory, 1 device memory (copied), 2 device memory borrowed (float4 records, no copy; the
  // caller keeps it alive until the batch is cleared).  async: host uploads go to the copy stream.
  // room for n more points behind src_used in the source pool
  int reserve_sources(size_t n) {
    if (src_used + n <= src_cap) return 0;
    const size_t nc = (src_used + n) + (src_used + n) / 2 + 4096;
    float4* q = nullptr;
    CKE(cudaMalloc(&q, nc * sizeof(float4)));
    CKE(cudaStreamSynchronize(copy_stream));
    if (src_pool && src_used) CKE(cudaMemcpyAsync(q, src_pool, src_used * sizeof(float4), cudaMemcpyDeviceToDevice, stream));
    CKE(cudaStreamSynchronize(stream));
    cudaFree(src_pool);
    src_pool = q;
    src_cap = nc;
    return 0;
  }
  int add_source(const void* pts, size_t n, size_t stride, int where, bool async, int* id) {
    if (where == 2) {
      if (stride != 16) return fail("a borrowed device buffer must hold 16-byte records");
      sources.push_back(Source{-1, n, static_cast<const float4*>(pts)});
      if (id) *id = (int)sources.size() - 1;
      return 0;
    }
    if (reserve_sources(n)) return 1;
    const bool on_copy = async && where == 0;
    if (upload_points(src_pool + src_used, pts, n, stride, where == 1, on_copy ? copy_stream : stream)) return 1;
    if (on_copy) sources_pending = true;
    sources.push_back(Source{(long long)src_used, n, nullptr});
    if (id) *id = (int)sources.size() - 1;
    src_used += n;
    return 0;
  }
  const float4* source_ptr(const Source& s) const { return s.borrowed ? s.borrowed : src_pool + s.offset; }

  template <typename T>
  int ensure(T*& p, size_t& cap, size_t need) {
    if (need <= cap) return 0;
    cudaFree(p);
    p = nullptr;
    cap = need + need / 4 + 16;
    CKE(cudaMalloc(&p, ((cap * sizeof(T) + 15) / 16) * 16));
    return 0;
  }

  void fill_state(AlignState& st, int target, const Source& src, const float* guess) {
    std::memset(&st, 0, sizeof(st));
    st.n_src = (int)src.n;
    st.target = target;
    st.src = source_ptr(src);
    st.step_size = params.step_size;
    st.trans_eps = params.trans_eps;
    st.max_iter = params.max_iter;
    st.resolution = params.resolution;
    // gauss constants, recomputed at every computeTransformation (NDT_I:80-87)
    const double c1 = 10 * (1 - params.outlier_ratio);
    const double c2 = params.outlier_ratio / std::pow((double)params.resolution, 3);
    const double d3 = -std::log(c2);
    const double d1 = -std::log(c1 + c2) - d3;
    const double d2 = -2 * std::log((-std::log(c1 * std::exp(-0.5) + c2) - d3) / d1);
    st.req.d1 = d1; st.req.d2 = d2; st.req.d3 = d3;
    st.req.search_method = params.search_method;
    st.req.mode = 0;
    // final_transformation_ = guess (Identity when guess == Identity: same bits either way, NDT_I:89-94)
    std::memcpy(st.final_T, guess, 16 * sizeof(float));
    float e[3];
    euler_012(st.final_T, e);                                    // NDT_I:96-104
    st.p[0] = st.final_T[12]; st.p[1] = st.final_T[13]; st.p[2] = st.final_T[14];
    st.p[3] = e[0]; st.p[4] = e[1]; st.p[5] = e[2];
    st.phase = kPhaseStart;
  }

  // Which derivative kernel(s) a run uses.  The default (strict_math = -1) is the term-by-term kernel for handles AND batches:
  // its per-pair float terms are the reference's bit for bit, so every result is the reference arithmetic's by construction.
  // KM_HYBRID (a batch with strict_math = -2, opt-in): every registration runs with the factorised kernel; those whose Newton
  // iteration gives a 1e-7 arithmetic difference room to grow -- more than params.replay_iters iterations, not converged, a
  // tiny last step, or a discrete decision of the loop taken at its threshold (AlignState::borderline) -- are then REPLAYED from
  // their initial guess with the term-by-term kernel, in the same call.  1.9x the strict mode's throughput; on held-out worlds
  // 5 of 20480 registrations still end outside the 1e-4 m / 1e-5 rad gate (profiles/r02/r02_c11_hybrid_seeds.txt): the
  // trajectories that diverge do so from any perturbation, however small, and no loop diagnostic separates all of them.
  enum { KM_FAST = 0, KM_STRICT = 1, KM_HYBRID = 2 };
  int kernel_mode() const {
    if (params.strict_math > 0 || params.search_method == NDTB_DIRECT26) return KM_STRICT;
    if (params.strict_math == 0) return KM_FAST;
    if (params.strict_math == -2) return single ? KM_STRICT : KM_HYBRID;
    return KM_STRICT;
  }
  int sm_count = 148;                 // cudaDevAttrMultiProcessorCount of `device` (init)
  int target_blocks() const { return sm_count * 8; }
  PinnedBuf st_idx;                   // replay pass: ids of the registrations run again
  // Morton-ordered copy of a large batch's sources, read by the derivative kernels only (the pool / borrowed buffers keep input
  // order for downloads); run_src[k] = where source k's ordered points start, nullptr = not ordered in this run
  SourceOrderWs so_ws;
  float4* src_sorted = nullptr;
  size_t src_sorted_cap = 0;
  std::vector<const float4*> run_src;
  const float4* run_source_ptr(int source_id) const {
    return (source_id < (int)run_src.size() && run_src[source_id]) ? run_src[source_id] : source_ptr(sources[source_id]);
  }
  std::vector<unsigned char> replay_flag;   // per item of the last run: 1 = replayed with the strict kernel

  // ---- a single registration as ONE CUDA graph launch (the loop NDT_I:114-159) -----------------
  // ndtb_align used to enqueue derivatives -> newton pairs and poll the device every fourth pair: ~0.23 ms of launch and
  // synchronisation latency around ~60 us of kernel time.  The graph: pull the registration state out of mapped host memory ->
  // newton (first request) -> WHILE conditional node { derivatives; newton (advances computeTransformation / the line search,
  // then tells the node whether another evaluation is wanted) } -> state back into mapped host memory.  One launch, one
  // synchronisation, convergence tested on the device.  The derivative grid is fixed (room for `blocks` 256-point blocks; blocks
  // past the source's end find nothing to do), buffers are the graph's own, so neither the source size nor the test hooks
  // disturb it; it is rebuilt when the kernel selection, the table descriptor's address or the capacity changes.
  struct AlignGraph {
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    int blocks = 0, method = -1, trace_cap = 0, eval_cap = 0;
    bool strict = false, mt_live = false;
    const VoxelTable* tables = nullptr;
    AlignState* d_state = nullptr;      // one registration
    int2* d_bmap = nullptr;
    double* d_partials = nullptr;
    double* d_trace = nullptr;
    size_t trace_doubles = 0;
    int* d_active = nullptr;
    PinnedBuf h_in, h_out;              // the state as the host fills it / as the device leaves it
    bool failed = false;                // graph construction failed once: stay on the eager path
  } ag;
  bool use_align_graph() const {
    static const bool off = [] { const char* e = getenv("NDTB_NO_ALIGN_GRAPH"); return e && atoi(e) > 0; }();
    return single && !off && !ag.failed && !getenv("NDTB_CHUNK");
  }
  void destroy_align_graph() {
    if (ag.exec) cudaGraphExecDestroy(ag.exec);
    if (ag.graph) cudaGraphDestroy(ag.graph);
    ag.exec = nullptr; ag.graph = nullptr;
  }
  int build_align_graph(int blocks_needed, bool strict, bool mt_live, int eval_cap, int tcap) {
    destroy_align_graph();
    const int chunk = 256;
    int blocks = 128;
    while (blocks < blocks_needed) blocks *= 2;
    if (!ag.d_state) CKE(cudaMalloc(&ag.d_state, sizeof(AlignState) + 16));
    if (!ag.d_active) CKE(cudaMalloc(&ag.d_active, sizeof(int)));
    if (blocks > ag.blocks || !ag.d_bmap) {
      cudaFree(ag.d_bmap); cudaFree(ag.d_partials);
      ag.d_bmap = nullptr; ag.d_partials = nullptr;
      CKE(cudaMalloc(&ag.d_bmap, (size_t)blocks * sizeof(int2)));
      CKE(cudaMalloc(&ag.d_partials, (size_t)blocks * 44 * sizeof(double)));
    }
    std::vector<int2> bm((size_t)blocks);
    for (int b = 0; b < blocks; ++b) bm[b] = make_int2(0, b * chunk);
    CKE(cudaMemcpyAsync(ag.d_bmap, bm.data(), bm.size() * sizeof(int2), cudaMemcpyHostToDevice, stream));
    CKE(cudaStreamSynchronize(stream));
    const size_t need_trace = (size_t)tcap * kTraceRec;
    if (need_trace > ag.trace_doubles) {
      cudaFree(ag.d_trace);
      ag.d_trace = nullptr;
      CKE(cudaMalloc(&ag.d_trace, need_trace * sizeof(double)));
      ag.trace_doubles = need_trace;
    }
    CKE(ag.h_in.reserve(sizeof(AlignState)));
    CKE(ag.h_out.reserve(sizeof(AlignState)));
    ag.blocks = blocks; ag.method = params.search_method; ag.strict = strict; ag.mt_live = mt_live;
    ag.trace_cap = tcap; ag.eval_cap = eval_cap; ag.tables = d_tables;
    // capture: prologue on the engine's stream, the loop body into the conditional node's graph on a second stream
    cudaStream_t s2 = nullptr;
    CKE(cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
    auto bail = [&](const std::string& m) {
      cudaGraph_t junk = nullptr;
      cudaStreamEndCapture(stream, &junk);
      if (junk) cudaGraphDestroy(junk);
      cudaStreamDestroy(s2);
      cudaGetLastError();
      ag.failed = true;
      return fail("align graph: " + m);
    };
    if (cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) { cudaStreamDestroy(s2); ag.failed = true; return fail("align graph: begin capture failed"); }
    if (launch_pull(ag.d_state, ag.h_in.d, sizeof(AlignState), stream) != cudaSuccess) return bail("prologue");
    if (launch_set_int(ag.d_active, 1, stream) != cudaSuccess) return bail("prologue");
    if (launch_newton(ag.d_state, 1, ag.d_partials, ag.d_trace, tcap, ag.d_active, stream) != cudaSuccess) return bail("prologue");
    cudaStreamCaptureStatus st_;
    cudaGraph_t g = nullptr;
    const cudaGraphNode_t* deps = nullptr;
    size_t n_deps = 0;
    if (cudaStreamGetCaptureInfo(stream, &st_, nullptr, &g, &deps, &n_deps) != cudaSuccess || st_ != cudaStreamCaptureStatusActive) return bail("capture info");
    cudaGraphConditionalHandle cond;
    if (cudaGraphConditionalHandleCreate(&cond, g, 1, cudaGraphCondAssignDefault) != cudaSuccess) return bail("conditional handle");
    cudaGraphNodeParams cp = {};
    cp.type = cudaGraphNodeTypeConditional;
    cp.conditional.handle = cond;
    cp.conditional.type = cudaGraphCondTypeWhile;
    cp.conditional.size = 1;
    cudaGraphNode_t cnode;
    if (cudaGraphAddNode(&cnode, g, deps, n_deps, &cp) != cudaSuccess) return bail("conditional node");
    cudaGraph_t body = cp.conditional.phGraph_out[0];
    if (cudaStreamUpdateCaptureDependencies(stream, &cnode, 1, cudaStreamSetCaptureDependencies) != cudaSuccess) return bail("capture dependencies");
    if (cudaStreamBeginCaptureToGraph(s2, body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal) != cudaSuccess) return bail("body capture");
    cudaError_t ce = (strict ? launch_derivatives : launch_derivatives_fast)(ag.d_state, d_tables, src_pool, ag.d_bmap, blocks, chunk, ag.d_partials,
                                                                             params.search_method, mt_live, s2);
    if (ce == cudaSuccess && mt_live) ce = launch_hessian_f64(ag.d_state, d_tables, src_pool, ag.d_bmap, blocks, chunk, ag.d_partials, s2);
    if (ce == cudaSuccess) ce = launch_newton(ag.d_state, 1, ag.d_partials, ag.d_trace, tcap, ag.d_active, s2, (unsigned long long)cond, eval_cap);
    cudaGraph_t body_out = nullptr;
    if (cudaStreamEndCapture(s2, &body_out) != cudaSuccess || ce != cudaSuccess) return bail("body");
    if (launch_pull(ag.h_out.d, ag.d_state, sizeof(AlignState), stream) != cudaSuccess) return bail("epilogue");
    if (cudaStreamEndCapture(stream, &ag.graph) != cudaSuccess) { cudaStreamDestroy(s2); cudaGetLastError(); ag.failed = true; return fail("align graph: end capture failed"); }
    cudaStreamDestroy(s2);
    if (cudaGraphInstantiate(&ag.exec, ag.graph, 0) != cudaSuccess) { cudaGetLastError(); destroy_align_graph(); ag.failed = true; return fail("align graph: instantiate failed"); }
    return 0;
  }
  // ndtb_align through the graph.  Returns 0 ok, 1 error, 2 = not applicable (the caller takes the eager path).
  int align_graph(const float* guess, ndtb_result* out) {
    if (!use_align_graph()) return 2;
    const Source& src = sources[0];
    const int mode = kernel_mode();
    if (mode == KM_HYBRID) return 2;
    const bool strict = mode == KM_STRICT;
    const bool mt_live = !((params.step_size - params.trans_eps / 2) > 0);
    const int eval_cap = 1 + (params.max_iter + 2) * (mt_live ? 12 : 1);
    const int tcap = eval_cap + 1;
    const int nb = (int)((src.n + 255) / 256);
    if (sync_tables_if_dirty()) return 1;
    if (!ag.exec || ag.blocks < nb || ag.method != params.search_method || ag.strict != strict || ag.mt_live != mt_live || ag.tables != d_tables ||
        ag.eval_cap != eval_cap || ag.trace_cap != tcap) {
      // the kernels' shared-memory attributes are set by their first ordinary launch: the first align of a configuration
      // runs eagerly, the graph is built behind it
      if (!ag_warm) { ag_warm = true; return 2; }
      if (build_align_graph(nb, strict, mt_live, eval_cap, tcap)) { err.clear(); return 2; }
    }
    AlignState& st = *static_cast<AlignState*>(ag.h_in.h);
    fill_state(st, 0, src, guess);
    st.first_block = 0;
    st.n_blocks = nb;
    stats = ndtb_batch_stats{};
    CKE(cudaEventRecord(ev[2], stream));
    CKE(cudaGraphLaunch(ag.exec, stream));
    CKE(cudaEventRecord(ev[3], stream));
    CKE(cudaStreamSynchronize(stream));
    float ms = 0;
    CKE(cudaEventElapsedTime(&ms, ev[2], ev[3]));
    last_align_ms = ms;
    const AlignState& r = *static_cast<const AlignState*>(ag.h_out.h);
    state_to_result(r, *out);
    stats.points_evaluated = (long long)r.n_src * r.n_evals;
    stats.pairs_evaluated = r.pairs_total;
    stats.deriv_launches = r.n_evals;
    stats.kernel_launches = 4 + 2ll * r.n_evals;
    trace_cap = tcap;
    trace_on_device = true;
    return 0;
  }
  bool ag_warm = false;
  bool trace_on_device = false;     // the last align's trace still sits in ag.d_trace (copied on demand, ndtb_get_trace)
  bool tables_dirty = true;         // the table descriptor on the device is older than targets[0].ts.t
  int sync_tables_if_dirty() {
    if (!tables_dirty && d_tables) return 0;
    if (sync_tables()) return 1;
    tables_dirty = false;
    return 0;
  }

  int plan_chunk(size_t total_pts) const {
    // small blocks keep the tail iterations (few registrations still active) balanced over the SMs
    int chunk = (int)((total_pts / (size_t)target_blocks() + 255) / 256 * 256);
    if (chunk < 256) chunk = 256;
    if (chunk > 1024) chunk = 1024;   // 1024 amortises the per-block set-up and reduction best (profiles/r01_v9_kernel_variants.txt)
    if (const char* ev = getenv("NDTB_CHUNK")) { const int c = atoi(ev); if (c >= 128) chunk = c; }   // tuning knob
    return chunk;
  }

  // The evaluation loop of one pass: derivatives -> newton pairs until the device reports no unfinished registration.
  int newton_loop(int n_items, bool strict, bool want_trace, bool replan, int n_blocks, int chunk, size_t& ev_used, size_t ev_first) {
    const bool mt_live = !((params.step_size - params.trans_eps / 2) > 0);   // NDT_I:807: the inner loop runs only then
    const int evals_per_iter = mt_live ? 12 : 1;
    const int max_evals = 1 + (params.max_iter + 2) * evals_per_iter;
    (void)ev_first;
    // PH_START: set up the first evaluation request
    CKE(launch_newton(d_states, n_items, d_partials, want_trace ? d_trace : nullptr, trace_cap, d_active, stream));
    stats.kernel_launches++;
    int poll = 4;
    if (const char* ev = getenv("NDTB_POLL")) { const int v = atoi(ev); if (v >= 1) poll = v; }   // tuning knob
    for (int it = 0; it < max_evals; ++it) {
      if (deriv_events.size() < ev_used + 2) {
        cudaEvent_t a, b;
        CKE(cudaEventCreate(&a)); CKE(cudaEventCreate(&b));
        deriv_events.push_back(a); deriv_events.push_back(b);
      }
      CKE(cudaEventRecord(deriv_events[ev_used], stream));
      CKE((strict ? launch_derivatives : launch_derivatives_fast)(d_states, d_tables, src_pool, d_blockmap, n_blocks, chunk,
                                                                  d_partials, params.search_method, mt_live, stream));
      CKE(cudaEventRecord(deriv_events[ev_used + 1], stream));
      ev_used += 2;
      stats.kernel_launches += mt_live ? 2 : 1;
      if (mt_live) {
        CKE(launch_hessian_f64(d_states, d_tables, src_pool, d_blockmap, n_blocks, chunk, d_partials, stream));
        stats.kernel_launches++;
      }
      CKE(launch_newton(d_states, n_items, d_partials, want_trace ? d_trace : nullptr, trace_cap, d_active, stream));
      stats.kernel_launches++;
      if ((it + 1) % poll == 0 || it + 1 == max_evals) {
        // published by a one-thread kernel straight into mapped host memory: a D2H memcpy would queue on a copy engine
        // behind another batch's multi-megabyte uploads and stall this loop for their whole duration
        if (replan) { CKE(launch_replan(d_states, n_items, d_blockmap, h_active_dev + 1, target_blocks(), stream)); stats.kernel_launches++; }
        CKE(launch_publish(d_active, h_active_dev, stream));
        stats.kernel_launches++;
        CKE(cudaStreamSynchronize(stream));
        if (*(volatile int*)h_active <= 0) break;
        if (replan) { n_blocks = ((volatile int*)h_active)[1]; chunk = ((volatile int*)h_active)[2]; }
      }
    }
    return 0;
  }

  static void state_to_result(const AlignState& st, ndtb_result& r) {
    std::memcpy(r.T, st.final_T, sizeof(r.T));
    r.iterations = st.nr_iterations;
    r.converged = st.done ? st.converged : 0;
    r.score = st.score;
    r.trans_probability = st.trans_probability;
    r.n_evals = st.n_evals;
  }

  // Run n_items registrations to completion on the device.  want_trace: per-evaluation records.
  int run_items(int n_items, const int* item_target, const int* item_source, const float* guesses, ndtb_result* results,
                bool want_trace) {
    last_item_target.clear();
    replay_flag.clear();
    if (n_items <= 0) return 0;
    CKE(cudaSetDevice(device));
    last_item_target.assign(item_target, item_target + n_items);
    stats = ndtb_batch_stats{};
    const int mode = kernel_mode();
    // work decomposition: one block per `chunk` source points of one registration
    size_t total_pts = 0;
    for (int i = 0; i < n_items; ++i) {
      if (item_target[i] < 0 || item_target[i] >= (int)targets.size()) return fail("item target id out of range");
      if (item_source[i] < 0 || item_source[i] >= (int)sources.size()) return fail("item source id out of range");
      total_pts += sources[item_source[i]].n;
    }
    int chunk = plan_chunk(total_pts);
    size_t total_blocks = 0;
    for (int i = 0; i < n_items; ++i) total_blocks += (sources[item_source[i]].n + chunk - 1) / chunk;
    CKE(cudaStreamSynchronize(stream));     // the staging buffers may still be read by the previous run's pulls
    CKE(st_states.reserve((size_t)n_items * sizeof(AlignState)));
    // (the replay pass plans its own, smaller, set of blocks at 256 points each at worst: room for that too)
    size_t blocks256 = 0;
    for (int i = 0; i < n_items; ++i) blocks256 += (sources[item_source[i]].n + 255) / 256;
    CKE(st_bmap.reserve(std::max(total_blocks, blocks256) * sizeof(int2) + 16));
    h_states = static_cast<AlignState*>(st_states.h);
    int2* bmap = static_cast<int2*>(st_bmap.h);
    size_t nb_fill = 0;
    // the registration states and the block map are filled on this thread WHILE the other lanes build the voxel tables
    // k_replan never needs more blocks than the start (chunk at its cap) or target_blocks() + one per registration (smaller chunks)
    const size_t blocks_cap = std::max(std::max(total_blocks, blocks256), (size_t)(target_blocks() + n_items)) + 16;
    // (a block size forced through NDTB_CHUNK may exceed the planner's cap, which is what blocks_cap assumes: no re-planning then)
    const bool replan = n_items > 1 && !getenv("NDTB_NO_REPLAN") && !getenv("NDTB_CHUNK");
    if (ensure(d_states, states_cap, (size_t)n_items)) return 1;
    if (ensure(d_blockmap, blockmap_cap, blocks_cap)) return 1;
    if (ensure(d_partials, partials_cap, blocks_cap * 44)) return 1;
    cudaError_t pull_err = cudaSuccess;
    // hybrid batches: the replay threshold, known before the first pass so that it can stop registrations that cross it
    int k_iters = params.replay_iters;
    if (const char* ev = getenv("NDTB_REPLAY_ITERS")) k_iters = atoi(ev);   // tuning knob
    const int abort_iters = (mode == KM_HYBRID && k_iters > 0 && !getenv("NDTB_REPLAY_NO_ABORT")) ? k_iters : 0;
    // Large batches: the derivative kernels read a Morton-ordered copy of every source (neighbouring threads then gather
    // neighbouring voxels: -6 % kernel time on filter_node's order, and no +42 % cliff for an arbitrarily ordered scan).  One
    // sort of all sources on this stream, under the table builds that run on the lanes' streams.  The sums only change in the
    // grouping of their FP64 additions (1e-16); small batches and single handles skip it (test_batch_matches_single's bit
    // equality is about those).  NDTB_SOURCE_ORDER=0 turns it off.
    run_src.clear();
    {
      static const bool off = [] { const char* ev = getenv("NDTB_SOURCE_ORDER"); return ev && atoi(ev) == 0; }();
      size_t all_pts = 0;
      for (const Source& sc : sources) all_pts += sc.n;
      if (!off && !single && all_pts >= ((size_t)1 << 20) && all_pts < ((size_t)1 << 31) && sources.size() <= (size_t)1 << 14) {
        std::vector<SrcDesc> desc(sources.size());
        size_t first = 0;
        for (size_t k = 0; k < sources.size(); ++k) {
          desc[k].p = source_ptr(sources[k]);
          desc[k].first = (unsigned int)first;
          desc[k].n = (unsigned int)sources[k].n;
          first += sources[k].n;
        }
        if (all_pts > src_sorted_cap) {
          cudaFree(src_sorted);
          src_sorted = nullptr; src_sorted_cap = 0;
          CKE(cudaMalloc(&src_sorted, (all_pts + all_pts / 8) * sizeof(float4)));
          src_sorted_cap = all_pts + all_pts / 8;
        }
        if (sources_pending) { CKE(cudaStreamWaitEvent(stream, sources_uploaded, 0)); sources_pending = false; }
        CKE(source_order_sort(so_ws, desc.data(), (int)desc.size(), all_pts, src_sorted, stream, &stats.kernel_launches));
        run_src.resize(sources.size());
        for (size_t k = 0; k < sources.size(); ++k) run_src[k] = sources[k].n ? src_sorted + desc[k].first : nullptr;
      }
    }
    auto fill = [&]() {
      for (int i = 0; i < n_items; ++i) {
        AlignState& st = h_states[i];
        fill_state(st, item_target[i], sources[item_source[i]], guesses + (size_t)i * 16);
        st.src = run_source_ptr(item_source[i]);
        st.abort_iters = abort_iters;
        st.first_block = (int)nb_fill;
        st.n_blocks = (st.n_src + chunk - 1) / chunk;
        for (int b = 0; b < st.n_blocks; ++b) bmap[nb_fill++] = make_int2(i, b * chunk);
      }
      // ... and pulled into HBM on the engine's stream, also under the builds (which run on the lanes' own streams)
      pull_err = launch_pull(d_states, st_states.d, (size_t)n_items * sizeof(AlignState), stream);
      if (pull_err == cudaSuccess) pull_err = launch_pull(d_blockmap, st_bmap.d, nb_fill * sizeof(int2), stream);
    };
    CKE(cudaEventRecord(ev[0], stream));
    if (build_pending_targets(fill)) return 1;
    CKE(pull_err);
    CKE(cudaEventRecord(ev[1], stream));
    if (sources_pending) { CKE(cudaStreamWaitEvent(stream, sources_uploaded, 0)); sources_pending = false; }
    if (sync_tables()) return 1;
    const bool mt_live = !((params.step_size - params.trans_eps / 2) > 0);
    const int max_evals = 1 + (params.max_iter + 2) * (mt_live ? 12 : 1);
    trace_cap = want_trace ? max_evals + 1 : 0;
    if (want_trace && ensure(d_trace, trace_doubles, (size_t)n_items * trace_cap * kTraceRec)) return 1;
    CKE(launch_set_int(d_active, n_items, stream));
    stats.kernel_launches += 3;
    CKE(cudaEventRecord(ev[2], stream));
    size_t ev_used = 0;
    if (newton_loop(n_items, mode == KM_STRICT, want_trace, replan, (int)nb_fill, chunk, ev_used, 0)) return 1;
    const size_t ev_pass1 = ev_used;
    CKE(cudaMemcpyAsync(h_states, d_states, (size_t)n_items * sizeof(AlignState), cudaMemcpyDeviceToHost, stream));
    CKE(cudaStreamSynchronize(stream));
    for (int i = 0; i < n_items; ++i) {
      const AlignState& st = h_states[i];
      state_to_result(st, results[i]);
      stats.points_evaluated += (long long)st.n_src * st.n_evals;
      stats.pairs_evaluated += st.pairs_total;
    }
    // ---- replay pass of a hybrid batch --------------------------------------------------------------------------
    int n_replay = 0;
    if (mode == KM_HYBRID) {
      replay_flag.assign((size_t)n_items, 0);
      CKE(st_idx.reserve((size_t)n_items * sizeof(int)));
      int* idx = static_cast<int*>(st_idx.h);
      double min_norm = params.replay_min_norm >= 0 ? params.replay_min_norm : 0.05 * params.trans_eps;
      if (const char* ev = getenv("NDTB_REPLAY_MIN_NORM")) min_norm = atof(ev);   // tuning knob
      int k_rise = 1 << 30;
      if (const char* ev = getenv("NDTB_REPLAY_RISE")) k_rise = atoi(ev);     // tuning knob: replay when n_rise >= this
      size_t pts = 0;
      for (int i = 0; i < n_items; ++i) {
        const AlignState& st = h_states[i];
        // last Newton step far below the step clamp (eps / 2): the final move is eps / 2 along delta_p / |delta_p|, a direction
        // that a difference of ~1e-6 in delta_p (one voxel flip of one point) turns by 1e-6 / |delta_p|
        const bool tiny_last = st.last_norm < min_norm;
        // fewer source points than unknowns: every point adds rank <= 3 to H, the 6x6 solve is (nearly) rank-deficient and
        // amplifies the factorised kernel's 1e-7 differences by the condition number (the one-point source of the edge tests)
        const bool underdetermined = st.n_src < 6;
        const bool again = !st.done || !st.converged || st.borderline || tiny_last || underdetermined ||
                           (k_iters > 0 && st.nr_iterations > k_iters) || st.n_rise >= k_rise;
        if (!again || st.n_src <= 0) continue;
        idx[n_replay++] = i;
        replay_flag[i] = 1;
        pts += (size_t)st.n_src;
      }
      if (n_replay > 0) {
        chunk = plan_chunk(pts);
        nb_fill = 0;
        for (int k = 0; k < n_replay; ++k) {
          const int i = idx[k];
          AlignState& st = h_states[i];
          fill_state(st, item_target[i], sources[item_source[i]], guesses + (size_t)i * 16);
          st.src = run_source_ptr(item_source[i]);
          st.first_block = (int)nb_fill;
          st.n_blocks = (st.n_src + chunk - 1) / chunk;
          for (int b = 0; b < st.n_blocks; ++b) bmap[nb_fill++] = make_int2(i, b * chunk);
        }
        CKE(launch_pull_items(d_states, static_cast<const AlignState*>(st_states.d), static_cast<const int*>(st_idx.d), n_replay, stream));
        CKE(launch_pull(d_blockmap, st_bmap.d, nb_fill * sizeof(int2), stream));
        CKE(launch_set_int(d_active, n_replay, stream));
        stats.kernel_launches += 3;
        if (newton_loop(n_items, true, false, replan, (int)nb_fill, chunk, ev_used, ev_pass1)) return 1;
        CKE(cudaMemcpyAsync(h_states, d_states, (size_t)n_items * sizeof(AlignState), cudaMemcpyDeviceToHost, stream));
      }
    }
    CKE(cudaEventRecord(ev[3], stream));
    if (want_trace) {
      h_trace.resize((size_t)n_items * trace_cap * kTraceRec);
      CKE(cudaMemcpyAsync(h_trace.data(), d_trace, h_trace.size() * sizeof(double), cudaMemcpyDeviceToHost, stream));
    }
    CKE(cudaStreamSynchronize(stream));
    float ms = 0;
    CKE(cudaEventElapsedTime(&ms, ev[0], ev[1])); stats.build_ms = ms;
    CKE(cudaEventElapsedTime(&ms, ev[0], ev[3])); stats.total_ms = ms;
    CKE(cudaEventElapsedTime(&ms, ev[2], ev[3])); last_align_ms = ms;
    stats.deriv_ms = 0;
    for (size_t k = 0; k + 1 < ev_pass1; k += 2) { CKE(cudaEventElapsedTime(&ms, deriv_events[k], deriv_events[k + 1])); stats.deriv_ms += ms; }
    stats.deriv_launches = (long long)(ev_pass1 / 2);
    for (size_t k = ev_pass1; k + 1 < ev_used; k += 2) { CKE(cudaEventElapsedTime(&ms, deriv_events[k], deriv_events[k + 1])); stats.replay_deriv_ms += ms; }
    stats.replay_deriv_launches = (long long)((ev_used - ev_pass1) / 2);
    stats.replayed = n_replay;
    if (n_replay > 0) {
      CKE(cudaEventElapsedTime(&ms, deriv_events[ev_pass1], ev[3])); stats.replay_ms = ms;
      const int* idx = static_cast<const int*>(st_idx.h);
      for (int k = 0; k < n_replay; ++k) {
        const AlignState& st = h_states[idx[k]];
        state_to_result(st, results[idx[k]]);
        stats.replay_points += (long long)st.n_src * st.n_evals;
        stats.replay_pairs += st.pairs_total;
      }
    }
    return 0;
  }
};

}  // namespace

struct ndtb_handle { Engine e; };
struct ndtb_batch { Engine e; };

extern "C" {

void ndtb_default_params(ndtb_params* p) {
  p->resolution = 1.0f; p->step_size = 0.1; p->trans_eps = 0.1; p->max_iter = 35; p->outlier_ratio = 0.55;
  p->search_method = NDTB_DIRECT7; p->min_points = 6; p->eig_mult = 0.01; p->strict_math = -1; p->dense_max_cells = 1ll << 26;
  p->replay_iters = 11; p->replay_min_norm = -1.0;
}

static int create_engine(int device, Engine& e) { return e.init(device); }

int ndtb_create(int device, ndtb_handle** out) {
  if (!out) return 1;
  ndtb_handle* h = new ndtb_handle();
  if (create_engine(device, h->e)) { g_create_error = h->e.err; delete h; *out = nullptr; return 1; }
  h->e.targets.resize(1);
  h->e.single = true;
  *out = h;
  return 0;
}
int ndtb_destroy(ndtb_handle* h) {
  if (!h) return 0;
  h->e.destroy();
  delete h;
  return 0;
}
const char* ndtb_last_error(const ndtb_handle* h) { return h ? h->e.err.c_str() : g_create_error.c_str(); }

static int check_params(Engine& e, const ndtb_params* p) {
  if (!p) return e.fail("null params");
  if (!(p->resolution > 0)) return e.fail("resolution must be > 0");
  if (p->search_method < 0 || p->search_method > 3) return e.fail("unknown search method");
  return 0;
}

int ndtb_set_params(ndtb_handle* h, const ndtb_params* p) {
  Engine& e = h->e;
  if (check_params(e, p)) return 1;
  cudaSetDevice(e.device);
  const bool rebuild = p->resolution != e.params.resolution || p->min_points != e.params.min_points || p->eig_mult != e.params.eig_mult ||
                       p->dense_max_cells != e.params.dense_max_cells;
  e.params = *p;
  Target& t = e.targets[0];
  if (rebuild && t.n > 0) {   // setResolution -> init() (NDT_H:127-137)
    t.ts.have_bbox = false;
    if (e.build_target(t)) return 1;
    if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  }
  return 0;
}
int ndtb_get_params(const ndtb_handle* h, ndtb_params* p) { *p = h->e.params; return 0; }

int ndtb_set_target(ndtb_handle* h, const void* pts, size_t n, size_t stride_bytes) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  Target& t = e.targets[0];
  if (e.reserve_target(t, n)) return 1;
  t.n = 0;
  cudaEventRecord(e.ev[0], e.stream);
  if (e.upload_points(t.pts, pts, n, stride_bytes, false, e.stream)) return 1;
  t.use = nullptr;
  t.n = n;
  e.nn_valid = false;
  e.tgt_gen = 0;
  t.ts.have_bbox = false;
  if (e.build_target(t)) return 1;
  cudaEventRecord(e.ev[1], e.stream);
  if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  float ms = 0;
  cudaEventElapsedTime(&ms, e.ev[0], e.ev[1]);
  e.last_build_ms = ms;
  return 0;
}

int ndtb_update_target(ndtb_handle* h, const void* pts, size_t n, size_t stride_bytes) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  Target& t = e.targets[0];
  if (!t.built || t.n == 0) return ndtb_set_target(h, pts, n, stride_bytes);
  if (n == 0) return 0;
  if (e.reserve_target(t, t.n + n)) return 1;
  cudaEventRecord(e.ev[0], e.stream);
  if (e.upload_points(t.pts + t.n, pts, n, stride_bytes, false, e.stream)) return 1;
  e.table_config(t);
  e.tables_dirty = true;
  cudaError_t ce = table_append(t.ts, e.ws, t.pts + t.n, n, e.stream, &e.stats.kernel_launches);
  if (ce != cudaSuccess) return e.fail(std::string("table_append: ") + cudaGetErrorString(ce));
  t.n += n;
  e.nn_valid = false;
  e.tgt_gen = 0;
  cudaEventRecord(e.ev[1], e.stream);
  if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  float ms = 0;
  cudaEventElapsedTime(&ms, e.ev[0], e.ev[1]);
  e.last_build_ms = ms;
  return 0;
}

int ndtb_set_target_auto(ndtb_handle* h, const void* pts, size_t n, size_t stride_bytes, int* was_incremental) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  Target& t = e.targets[0];
  if (was_incremental) *was_incremental = 0;
  if (!t.built || t.n == 0 || n < t.n || n - t.n > t.n || !t.ts.valid) return ndtb_set_target(h, pts, n, stride_bytes);
  // upload behind the current points, compare the shared prefix on the device
  const size_t old_n = t.n;
  if (e.reserve_target(t, old_n + n)) return 1;
  cudaEventRecord(e.ev[0], e.stream);
  float4* fresh = t.pts + old_n;
  if (e.upload_points(fresh, pts, n, stride_bytes, false, e.stream)) return 1;
  *e.h_active = 1;
  cudaMemcpyAsync(e.d_active, e.h_active, sizeof(int), cudaMemcpyHostToDevice, e.stream);
  cudaError_t ce = launch_prefix_equal(t.pts, fresh, old_n, e.d_active, e.stream);
  if (ce != cudaSuccess) return e.fail("prefix compare failed");
  cudaMemcpyAsync(e.h_active, e.d_active, sizeof(int), cudaMemcpyDeviceToHost, e.stream);
  if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  if (*e.h_active == 1) {
    // the tail [old_n, n) of the fresh copy already sits at t.pts + old_n + old_n; move it down
    const size_t tail = n - old_n;
    if (tail > 0) {
      // tail <= old_n (the guard above): [2 old_n, 2 old_n + tail) -> [old_n, old_n + tail) never overlap
      cudaMemcpyAsync(t.pts + old_n, fresh + old_n, tail * sizeof(float4), cudaMemcpyDeviceToDevice, e.stream);
      e.table_config(t);
      e.tables_dirty = true;
      ce = table_append(t.ts, e.ws, t.pts + old_n, tail, e.stream, &e.stats.kernel_launches);
      if (ce != cudaSuccess) return e.fail(std::string("table_append: ") + cudaGetErrorString(ce));
      t.n = n;
      e.nn_valid = false;
      e.tgt_gen = 0;
    }
    if (was_incremental) *was_incremental = 1;
  } else {
    // not an extension of the old cloud: move the fresh copy down to the front.  Source [old_n, old_n + n) and destination
    // [0, n) overlap whenever n > old_n, and an overlapping cudaMemcpy is undefined: ascending chunks of at most old_n points
    // (the shift) on the one stream are each disjoint from their own source, and a chunk only overwrites source ranges that
    // earlier chunks have already consumed.
    for (size_t off = 0; off < n; off += old_n) {
      const size_t len = std::min(old_n, n - off);
      if (cudaMemcpyAsync(t.pts + off, fresh + off, len * sizeof(float4), cudaMemcpyDeviceToDevice, e.stream) != cudaSuccess)
        return e.fail("device copy failed");
    }
    t.n = n;
    e.nn_valid = false;
    e.tgt_gen = 0;
    t.ts.have_bbox = false;
    if (e.build_target(t)) return 1;
  }
  cudaEventRecord(e.ev[1], e.stream);
  if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  float ms = 0;
  cudaEventElapsedTime(&ms, e.ev[0], e.ev[1]);
  e.last_build_ms = ms;
  return 0;
}

int ndtb_set_source(ndtb_handle* h, const void* pts, size_t n, size_t stride_bytes) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  e.sources.clear();
  e.src_used = 0;
  e.have_result = false;
  if (stride_bytes >= 20 && n > 0) {
    // pcl::PointXYZI-like records: pack x y z intensity into the pinned staging buffer on the host (one pass over
    // 32-byte records) and upload 16 B / point in one copy -- cheaper than a strided pageable copy, and the
    // intensity travels in .w for the localmap entry points (no kernel of the NDT path reads .w)
    if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");   // staging may still be in flight
    if (e.st_src.reserve(n * sizeof(float4)) != cudaSuccess) return e.fail("pinned staging allocation failed");
    float4* dst = static_cast<float4*>(e.st_src.h);
    const char* src = static_cast<const char*>(pts);
    for (size_t i = 0; i < n; ++i) {
      const float* f = reinterpret_cast<const float*>(src + i * stride_bytes);
      dst[i] = make_float4(f[0], f[1], f[2], f[4]);
    }
    return e.add_source(e.st_src.h, n, 16, 0, false, nullptr);
  }
  return e.add_source(pts, n, stride_bytes, 0, false, nullptr);
}

static cudaError_t nn_build_for(Engine& e, BuildWorkspace* ws, cudaStream_t s) {
  Target& t = e.targets[0];
  const float cell = e.params.resolution > 0.25f ? e.params.resolution : 0.25f;
  // the voxel table already holds the float bounding box of the target's points: the NN grid build never blocks
  float bb[6];
  for (int a = 0; a < 3; ++a) { bb[a] = t.ts.mn[a]; bb[3 + a] = t.ts.mx[a]; }
  return nn_build(e.nn, ws, t.use ? t.use : t.pts, t.n, cell, t.built && t.ts.have_bbox ? bb : nullptr, s, &e.stats.kernel_launches);
}
// Enqueues the NN-grid build of a changed target on the side stream, behind everything the main stream holds so far.
static void nn_prebuild(Engine& e) {
  static const bool off = [] { const char* ev = getenv("NDTB_NN_PREBUILD"); return ev && atoi(ev) == 0; }();
  Target& t = e.targets[0];
  if (off || !e.single || !e.fitness_used || e.nn_valid || t.n == 0 || !t.built || !t.ts.have_bbox) return;
  if (!e.nn_stream) {
    if (cudaStreamCreateWithFlags(&e.nn_stream, cudaStreamNonBlocking) != cudaSuccess) { e.nn_stream = nullptr; return; }
    cudaEventCreateWithFlags(&e.ev_nn_dep, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&e.ev_nn_ready, cudaEventDisableTiming);
    e.ws_nn = build_ws_create();
  }
  if (!e.ws_nn || !e.ev_nn_dep || !e.ev_nn_ready) return;
  cudaEventRecord(e.ev_nn_dep, e.stream);
  cudaStreamWaitEvent(e.nn_stream, e.ev_nn_dep, 0);
  if (nn_build_for(e, e.ws_nn, e.nn_stream) != cudaSuccess) { cudaGetLastError(); return; }   // fitness builds it the ordinary way
  cudaEventRecord(e.ev_nn_ready, e.nn_stream);
  e.nn_valid = true;
  e.nn_pending = true;
}

int ndtb_align(ndtb_handle* h, const float guess[16], ndtb_result* out) {
  Engine& e = h->e;
  if (e.sources.empty()) return e.fail("no source cloud set");
  if (!e.targets[0].built) return e.fail("no target cloud set");
  const int zero = 0;
  if (cudaSetDevice(e.device) != cudaSuccess) return e.fail("cudaSetDevice failed");
  nn_prebuild(e);
  const int via_graph = e.align_graph(guess, &e.last);
  if (via_graph == 1) return 1;
  if (via_graph == 2) {
    e.trace_on_device = false;
    if (e.run_items(1, &zero, &zero, guess, &e.last, true)) return 1;
  }
  // whatever the main stream does next (the fitness query, a target refresh that rewrites the points) comes after the grid build
  if (e.nn_pending) { cudaStreamWaitEvent(e.stream, e.ev_nn_ready, 0); e.nn_pending = false; }
  e.have_result = true;
  e.h_trace_n = e.last.n_evals;
  if (out) *out = e.last;
  return 0;
}

int ndtb_get_output(ndtb_handle* h, void* pts, size_t stride_bytes) {
  Engine& e = h->e;
  if (!e.have_result) return e.fail("no align result");
  cudaSetDevice(e.device);
  const size_t n = e.sources[0].n;
  if (n > e.output_cap) { cudaFree(e.d_output); e.d_output = nullptr; if (cudaMalloc(&e.d_output, n * sizeof(float4)) != cudaSuccess) return e.fail("alloc"); e.output_cap = n; }
  cudaMemcpyAsync(e.d_T16, e.last.T, 64, cudaMemcpyHostToDevice, e.stream);
  if (launch_transform(e.src_pool + e.sources[0].offset, n, e.d_T16, e.d_output, e.stream) != cudaSuccess) return e.fail("transform failed");
  if (cudaMemcpy2DAsync(pts, stride_bytes, e.d_output, 16, 12, n, cudaMemcpyDeviceToHost, e.stream) != cudaSuccess) return e.fail("copy failed");
  if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  return 0;
}

int ndtb_fitness(ndtb_handle* h, double max_range, const float* T, double* out) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  Target& t = e.targets[0];
  if (e.sources.empty() || t.n == 0) return e.fail("fitness needs a source and a target");
  if (!T && !e.have_result) return e.fail("no align result");
  e.fitness_used = true;
  if (!e.nn_valid) {
    cudaError_t ce = nn_build_for(e, e.ws, e.stream);
    if (ce != cudaSuccess) return e.fail(std::string("nn_build: ") + cudaGetErrorString(ce));
    e.nn_valid = true;
  }
  const size_t n = e.sources[0].n;
  if (n > e.d2_cap) { cudaFree(e.d_d2); e.d_d2 = nullptr; if (cudaMalloc(&e.d_d2, n * sizeof(float)) != cudaSuccess) return e.fail("alloc"); e.d2_cap = n; }
  cudaMemcpyAsync(e.d_T16, T ? T : e.last.T, 64, cudaMemcpyHostToDevice, e.stream);
  if (nn_query(e.nn, e.src_pool + e.sources[0].offset, n, e.d_T16, e.d_d2, e.stream) != cudaSuccess) return e.fail("nn_query failed");
  // pcl::Registration::getFitnessScore: the mean (double) of the float squared distances <= max_range.  Summed on the device in
  // a fixed order (the reference adds in point order: same terms, grouping differs at the 1e-16 level); the host reads {sum, n}.
  if (e.st_d2.reserve(2 * sizeof(double)) != cudaSuccess) return e.fail("pinned staging allocation failed");
  if (!e.d_fit2 && cudaMalloc(&e.d_fit2, 2 * sizeof(double)) != cudaSuccess) return e.fail("alloc");
  if (d2_reduce(e.d_d2, n, max_range, e.d_fit2, e.stream) != cudaSuccess) return e.fail("fitness reduction failed");
  cudaMemcpyAsync(e.st_d2.h, e.d_fit2, 2 * sizeof(double), cudaMemcpyDeviceToHost, e.stream);
  if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  const double* r = static_cast<const double*>(e.st_d2.h);
  *out = r[1] > 0 ? r[0] / r[1] : 1.7976931348623157e308;
  return 0;
}

// shared by eval_derivatives / compute_hessian / calculate_score: one evaluation in `mode`
static int single_eval(Engine& e, const float T[16], const double p[6], int mode, AlignState* out_state) {
  cudaSetDevice(e.device);
  if (e.sources.empty()) return e.fail("no source cloud set");
  Target& t = e.targets[0];
  if (!t.built) return e.fail("no target cloud set");
  if (e.sync_tables()) return 1;
  const int chunk = 256;
  AlignState st;
  static const float I16[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  e.fill_state(st, 0, e.sources[0], I16);
  st.first_block = 0;
  st.n_blocks = (st.n_src + chunk - 1) / chunk;
  std::vector<int2> bmap;
  for (int b = 0; b < st.n_blocks; ++b) bmap.push_back(make_int2(0, b * chunk));
  if (e.ensure(e.d_states, e.states_cap, (size_t)1)) return 1;
  if (e.ensure(e.d_blockmap, e.blockmap_cap, bmap.size())) return 1;
  if (e.ensure(e.d_partials, e.partials_cap, bmap.size() * 44)) return 1;
  cudaMemcpyAsync(e.d_states, &st, sizeof(st), cudaMemcpyHostToDevice, e.stream);
  cudaMemcpyAsync(e.d_blockmap, bmap.data(), bmap.size() * sizeof(int2), cudaMemcpyHostToDevice, e.stream);
  cudaMemcpyAsync(e.d_T16, T, 64, cudaMemcpyHostToDevice, e.stream);
  cudaMemcpyAsync(e.d_p6, p, 48, cudaMemcpyHostToDevice, e.stream);
  *e.h_active = 1;
  cudaMemcpyAsync(e.d_active, e.h_active, sizeof(int), cudaMemcpyHostToDevice, e.stream);
  if (launch_prepare_eval(e.d_states, e.d_T16, e.d_p6, mode, e.stream) != cudaSuccess) return e.fail("prepare failed");
  cudaError_t ce;
  if (mode == 1 || mode == 2) ce = (e.use_strict() ? launch_derivatives : launch_derivatives_fast)(e.d_states, e.d_tables, e.src_pool, e.d_blockmap, st.n_blocks, chunk, e.d_partials, e.params.search_method, true, e.stream);
  else ce = launch_hessian_f64(e.d_states, e.d_tables, e.src_pool, e.d_blockmap, st.n_blocks, chunk, e.d_partials, e.stream);
  if (ce != cudaSuccess) return e.fail(std::string("eval launch: ") + cudaGetErrorString(ce));
  if (launch_newton(e.d_states, 1, e.d_partials, nullptr, 0, e.d_active, e.stream) != cudaSuccess) return e.fail("reduce failed");
  cudaMemcpyAsync(out_state, e.d_states, sizeof(AlignState), cudaMemcpyDeviceToHost, e.stream);
  ce = cudaStreamSynchronize(e.stream);
  if (ce != cudaSuccess) return e.fail(std::string("eval: ") + cudaGetErrorString(ce));
  return 0;
}

int ndtb_eval_derivatives(ndtb_handle* h, const float T[16], const double p[6], int compute_hessian, double* score, double g[6],
                          double H[36]) {
  AlignState st;
  if (single_eval(h->e, T, p, compute_hessian ? 1 : 2, &st)) return 1;
  if (score) *score = st.score;
  if (g) std::memcpy(g, st.g, 48);
  if (H) std::memcpy(H, st.H, 288);
  h->e.stats.pairs_evaluated = st.pairs;
  return 0;
}
int ndtb_compute_hessian(ndtb_handle* h, const float T[16], const double p[6], double H[36]) {
  AlignState st;
  if (single_eval(h->e, T, p, 3, &st)) return 1;
  std::memcpy(H, st.H, 288);
  return 0;
}
int ndtb_calculate_score(ndtb_handle* h, const float T[16], double* out) {
  AlignState st;
  const double p0[6] = {0, 0, 0, 0, 0, 0};
  if (single_eval(h->e, T, p0, 4, &st)) return 1;
  *out = st.score / (double)h->e.sources[0].n;   // NDT_I:951
  return 0;
}

int ndtb_num_voxels(ndtb_handle* h, size_t* n_all) { *n_all = h->e.targets[0].ts.valid ? (size_t)h->e.targets[0].ts.n_vox : 0; return 0; }

int ndtb_dump_voxels(ndtb_handle* h, int* ijk, int* nr_points, int* n_raw, double* mean, float* icov, float* centroid) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  Target& t = e.targets[0];
  const size_t n = t.ts.valid ? (size_t)t.ts.n_vox : 0;
  if (n == 0) return 0;
  std::vector<unsigned long long> keys(n);
  std::vector<VoxelRec> rec(n);
  std::vector<float4> cen(n);
  std::vector<int> vn(n);
  cudaMemcpy(keys.data(), t.ts.t.vkey, n * 8, cudaMemcpyDeviceToHost);
  cudaMemcpy(rec.data(), t.ts.t.rec, n * sizeof(VoxelRec), cudaMemcpyDeviceToHost);
  cudaMemcpy(cen.data(), t.ts.t.centroid, n * sizeof(float4), cudaMemcpyDeviceToHost);
  if (cudaMemcpy(vn.data(), t.ts.t.vn, n * 4, cudaMemcpyDeviceToHost) != cudaSuccess) return e.fail("dump copy failed");
  for (size_t v = 0; v < n; ++v) {
    int a, b, c;
    unpack_key(keys[v], a, b, c);
    if (ijk) { ijk[v * 3] = a; ijk[v * 3 + 1] = b; ijk[v * 3 + 2] = c; }
    if (nr_points) nr_points[v] = rec[v].nr_points;
    if (n_raw) n_raw[v] = vn[v];
    if (mean) for (int k = 0; k < 3; ++k) mean[v * 3 + k] = rec[v].mean[k];
    if (icov) for (int k = 0; k < 9; ++k) icov[v * 9 + k] = rec[v].icov[k];
    if (centroid) { centroid[v * 3] = cen[v].x; centroid[v * 3 + 1] = cen[v].y; centroid[v * 3 + 2] = cen[v].z; }
  }
  return 0;
}

int ndtb_dump_icov64(ndtb_handle* h, double* icov) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  Target& t = e.targets[0];
  const size_t n = t.ts.valid ? (size_t)t.ts.n_vox : 0;
  if (n == 0) return 0;
  if (cudaMemcpy(icov, t.ts.t.icov64, n * 9 * sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess) return e.fail("dump copy failed");
  return 0;
}

void ndtb_pose_to_matrix(const double p[6], float T[16]) { pose_to_matrix_host(p, T); }

int ndtb_host_alloc(size_t bytes, void** out) {
  *out = nullptr;
  return cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable) == cudaSuccess ? 0 : 1;
}
int ndtb_host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? 0 : 1; }

int ndtb_neighbors(ndtb_handle* h, const float* xyz, size_t n, int method, int* cnt, int* ijk) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  if (!e.targets[0].built) return e.fail("no target cloud set");
  if (e.sync_tables()) return 1;
  float* d_xyz = nullptr; int* d_cnt = nullptr; int* d_ijk = nullptr;
  if (cudaMalloc(&d_xyz, n * 12) != cudaSuccess || cudaMalloc(&d_cnt, n * 4) != cudaSuccess || cudaMalloc(&d_ijk, n * 27 * 3 * 4) != cudaSuccess)
    return e.fail("alloc failed");
  cudaMemcpyAsync(d_xyz, xyz, n * 12, cudaMemcpyHostToDevice, e.stream);
  cudaMemsetAsync(d_ijk, 0, n * 27 * 3 * 4, e.stream);
  cudaError_t ce = launch_neighbors(e.d_tables, d_xyz, n, method, d_cnt, d_ijk, e.stream);
  cudaMemcpyAsync(cnt, d_cnt, n * 4, cudaMemcpyDeviceToHost, e.stream);
  cudaMemcpyAsync(ijk, d_ijk, n * 27 * 3 * 4, cudaMemcpyDeviceToHost, e.stream);
  cudaError_t ce2 = cudaStreamSynchronize(e.stream);
  cudaFree(d_xyz); cudaFree(d_cnt); cudaFree(d_ijk);
  if (ce != cudaSuccess || ce2 != cudaSuccess) return e.fail("neighbors kernel failed");
  return 0;
}

int ndtb_get_trace(ndtb_handle* h, double* out, int max_recs, int* n) {
  Engine& e = h->e;
  if (!e.have_result) return e.fail("no align result");
  if (e.trace_on_device) {     // the align ran as a graph: its trace is fetched when somebody asks for it
    cudaSetDevice(e.device);
    const int have = e.h_trace_n < e.trace_cap ? e.h_trace_n : e.trace_cap;
    e.h_trace.resize((size_t)(have > 0 ? have : 1) * kTraceRec);
    if (have > 0 && cudaMemcpy(e.h_trace.data(), e.ag.d_trace, (size_t)have * kTraceRec * sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess)
      return e.fail("trace copy failed");
    e.trace_on_device = false;
  }
  const int avail = e.h_trace_n < e.trace_cap ? e.h_trace_n : e.trace_cap;
  if (n) *n = avail;
  const int m = avail < max_recs ? avail : max_recs;
  if (out && m > 0) std::memcpy(out, e.h_trace.data(), (size_t)m * kTraceRec * sizeof(double));
  return 0;
}

int ndtb_last_timing(ndtb_handle* h, double* build_ms, double* align_ms) {
  if (build_ms) *build_ms = h->e.last_build_ms;
  if (align_ms) *align_ms = h->e.last_align_ms;
  return 0;
}

// ---- odom_node's localmap on the device (reference src/odom_node.cpp:218-233, 289-290, 329-356) ---------
static int lm_reserve(Engine& e, Engine::DevCloud& c, size_t need) {
  if (need <= c.cap) return 0;
  const size_t nc = need + need / 2 + 4096;
  float4* q = nullptr;
  if (cudaMalloc(&q, nc * sizeof(float4)) != cudaSuccess) return e.fail("localmap allocation failed");
  if (c.p && c.n) cudaMemcpyAsync(q, c.p, c.n * sizeof(float4), cudaMemcpyDeviceToDevice, e.stream);
  if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  cudaFree(c.p);
  c.p = q;
  c.cap = nc;
  return 0;
}

int ndtb_localmap_clear(ndtb_handle* h) {
  Engine& e = h->e;
  e.lm.n = e.lm_tmp.n = e.lm_last.n = 0;
  e.lm_snapshot = 0;
  e.lm_snapshot_valid = false;
  ++e.lm_gen;
  return 0;
}

int ndtb_localmap_append_source(ndtb_handle* h, const float T[16], float leaf, int also_tmp, size_t* n_added) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  if (n_added) *n_added = 0;
  if (e.sources.empty()) return e.fail("ndtb_localmap_append_source needs a source (ndtb_set_source)");
  const size_t n = e.sources[0].n;
  if (n == 0) return 0;
  if (lm_reserve(e, e.lm_last, n)) return 1;
  if (n > e.lm_scratch_cap) {
    cudaFree(e.lm_scratch);
    e.lm_scratch = nullptr;
    if (cudaMalloc(&e.lm_scratch, (n + n / 4) * sizeof(float4)) != cudaSuccess) return e.fail("alloc");
    e.lm_scratch_cap = n + n / 4;
  }
  if (cudaMemcpyAsync(e.d_T16, T, 64, cudaMemcpyHostToDevice, e.stream) != cudaSuccess) return e.fail("device copy failed");
  const bool ds = leaf > 0.f;
  if (launch_transform(e.source_ptr(e.sources[0]), n, e.d_T16, ds ? e.lm_scratch : e.lm_last.p, e.stream) != cudaSuccess)
    return e.fail("transform failed");
  ++e.stats.kernel_launches;
  size_t m = n;
  if (ds) {
    cudaError_t ce = voxelgrid_downsample(e.ws, e.lm_scratch, n, leaf, e.lm_last.p, &m, e.stream, &e.stats.kernel_launches);
    if (ce != cudaSuccess) return e.fail(std::string("voxelgrid_downsample: ") + cudaGetErrorString(ce));
  }
  e.lm_last.n = m;
  if (lm_reserve(e, e.lm, e.lm.n + m)) return 1;
  if (cudaMemcpyAsync(e.lm.p + e.lm.n, e.lm_last.p, m * sizeof(float4), cudaMemcpyDeviceToDevice, e.stream) != cudaSuccess) return e.fail("device copy failed");
  e.lm.n += m;
  if (also_tmp) {
    if (lm_reserve(e, e.lm_tmp, e.lm_tmp.n + m)) return 1;
    if (cudaMemcpyAsync(e.lm_tmp.p + e.lm_tmp.n, e.lm_last.p, m * sizeof(float4), cudaMemcpyDeviceToDevice, e.stream) != cudaSuccess) return e.fail("device copy failed");
    e.lm_tmp.n += m;
  }
  if (n_added) *n_added = m;
  return cudaGetLastError() == cudaSuccess ? 0 : e.fail("localmap append failed");
}

int ndtb_localmap_snapshot(ndtb_handle* h) {
  h->e.lm_snapshot = h->e.lm.n;
  h->e.lm_snapshot_valid = true;
  return 0;
}

int ndtb_localmap_set_target(ndtb_handle* h, int force_full, int* was_incremental) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  if (was_incremental) *was_incremental = 0;
  if (!e.lm_snapshot_valid) return e.fail("ndtb_localmap_set_target needs a snapshot taken after the last roll / clear");
  Target& t = e.targets[0];
  const size_t n = e.lm_snapshot;
  const bool inc = !force_full && t.built && t.ts.valid && e.tgt_gen == e.lm_gen && e.tgt_from_lm <= n && t.n == e.tgt_from_lm && t.n > 0;
  cudaEventRecord(e.ev[0], e.stream);
  if (inc) {
    const size_t tail = n - e.tgt_from_lm;
    if (tail > 0) {
      if (e.reserve_target(t, n)) return 1;
      if (cudaMemcpyAsync(t.pts + t.n, e.lm.p + e.tgt_from_lm, tail * sizeof(float4), cudaMemcpyDeviceToDevice, e.stream) != cudaSuccess) return e.fail("device copy failed");
      e.table_config(t);
      e.tables_dirty = true;
      cudaError_t ce = table_append(t.ts, e.ws, t.pts + t.n, tail, e.stream, &e.stats.kernel_launches);
      if (ce != cudaSuccess) return e.fail(std::string("table_append: ") + cudaGetErrorString(ce));
      t.n = n;
      e.nn_valid = false;
    }
    if (was_incremental) *was_incremental = 1;
  } else {
    if (e.reserve_target(t, n)) return 1;
    t.n = 0;
    if (n) cudaMemcpyAsync(t.pts, e.lm.p, n * sizeof(float4), cudaMemcpyDeviceToDevice, e.stream);
    t.use = nullptr;
    t.n = n;
    e.nn_valid = false;
    t.ts.have_bbox = false;
    if (e.build_target(t)) return 1;
  }
  e.tgt_gen = e.lm_gen;
  e.tgt_from_lm = n;
  cudaEventRecord(e.ev[1], e.stream);
  if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  float ms = 0;
  cudaEventElapsedTime(&ms, e.ev[0], e.ev[1]);
  e.last_build_ms = ms;
  return 0;
}

int ndtb_localmap_roll(ndtb_handle* h) {
  Engine& e = h->e;
  std::swap(e.lm, e.lm_tmp);
  e.lm_tmp.n = 0;
  ++e.lm_gen;
  e.lm_snapshot = 0;
  e.lm_snapshot_valid = false;
  return 0;
}

int ndtb_localmap_sizes(ndtb_handle* h, size_t sizes[3]) {
  sizes[0] = h->e.lm.n; sizes[1] = h->e.lm_tmp.n; sizes[2] = h->e.lm_snapshot_valid ? h->e.lm_snapshot : 0;
  return 0;
}

int ndtb_localmap_download(ndtb_handle* h, int which, void* pts, size_t cap, size_t stride_bytes, size_t* n) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  const Engine::DevCloud* c = which == 0 ? &e.lm : which == 1 ? &e.lm_tmp : which == 2 ? &e.lm_last : nullptr;
  if (!c) return e.fail("which must be 0 (localmap), 1 (tmp_map) or 2 (last appended cloud)");
  if (stride_bytes < 12) return e.fail("stride_bytes must be >= 12");
  const size_t m = c->n < cap ? c->n : cap;
  if (n) *n = m;
  if (m == 0) return 0;
  if (stride_bytes == 16) {
    if (cudaMemcpyAsync(pts, c->p, m * 16, cudaMemcpyDeviceToHost, e.stream) != cudaSuccess) return e.fail("device copy failed");
  } else {
    if (cudaMemcpy2DAsync(pts, stride_bytes, c->p, 16, 12, m, cudaMemcpyDeviceToHost, e.stream) != cudaSuccess) return e.fail("device copy failed");
    if (stride_bytes >= 20)
      if (cudaMemcpy2DAsync(static_cast<char*>(pts) + 16, stride_bytes, reinterpret_cast<const char*>(c->p) + 12, 16, 4, m, cudaMemcpyDeviceToHost, e.stream) != cudaSuccess) return e.fail("device copy failed");
  }
  if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  return 0;
}

// ---- keyframe-count window (reference src/odom_node.cpp:737-802) ----------------------------------------
static int copy_cloud_to_host(Engine& e, const float4* d, size_t m, void* pts, size_t stride_bytes);
int ndtb_window_add_source(ndtb_handle* h, const float T[16], float leaf, int* id, size_t* n_points) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  if (n_points) *n_points = 0;
  if (e.sources.empty()) return e.fail("ndtb_window_add_source needs a source (ndtb_set_source)");
  const size_t n = e.sources[0].n;
  Engine::DevCloud c;
  // a buffer of a dropped cloud that is large enough, else a fresh one
  for (size_t k = 0; k < e.win_free.size(); ++k)
    if (e.win_free[k].cap >= n) { c = e.win_free[k]; e.win_free.erase(e.win_free.begin() + (long)k); break; }
  c.n = 0;
  if (c.cap < n || !c.p) {
    cudaFree(c.p);
    c.p = nullptr;
    c.cap = n + n / 4 + 256;
    if (cudaMalloc(&c.p, c.cap * sizeof(float4)) != cudaSuccess) return e.fail("window cloud allocation failed");
  }
  size_t m = n;
  if (n > 0) {
    const bool ds = leaf > 0.f;
    if (ds && n > e.lm_scratch_cap) {
      cudaFree(e.lm_scratch);
      e.lm_scratch = nullptr;
      if (cudaMalloc(&e.lm_scratch, (n + n / 4) * sizeof(float4)) != cudaSuccess) return e.fail("alloc");
      e.lm_scratch_cap = n + n / 4;
    }
    if (cudaMemcpyAsync(e.d_T16, T, 64, cudaMemcpyHostToDevice, e.stream) != cudaSuccess) return e.fail("device copy failed");
    if (launch_transform(e.source_ptr(e.sources[0]), n, e.d_T16, ds ? e.lm_scratch : c.p, e.stream) != cudaSuccess) return e.fail("transform failed");
    ++e.stats.kernel_launches;
    if (ds) {
      cudaError_t ce = voxelgrid_downsample(e.ws, e.lm_scratch, n, leaf, c.p, &m, e.stream, &e.stats.kernel_launches);
      if (ce != cudaSuccess) return e.fail(std::string("voxelgrid_downsample: ") + cudaGetErrorString(ce));
    }
  }
  c.n = m;
  e.win.push_back(c);
  if (id) *id = (int)e.win.size() - 1;
  if (n_points) *n_points = m;
  return 0;
}
int ndtb_window_drop(ndtb_handle* h, int id) {
  Engine& e = h->e;
  if (id < 0 || id >= (int)e.win.size() || !e.win[id].p) return e.fail("ndtb_window_drop: no such window cloud");
  e.win_free.push_back(e.win[id]);
  e.win[id] = Engine::DevCloud();
  return 0;
}
int ndtb_window_clear(ndtb_handle* h) {
  Engine& e = h->e;
  for (auto& c : e.win)
    if (c.p) e.win_free.push_back(c);
  e.win.clear();
  return 0;
}
int ndtb_window_set_target(ndtb_handle* h, const int* ids, int n_ids, float leaf, size_t* n_target) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  if (n_target) *n_target = 0;
  size_t total = 0;
  for (int k = 0; k < n_ids; ++k) {
    if (ids[k] < 0 || ids[k] >= (int)e.win.size() || !e.win[ids[k]].p) return e.fail("ndtb_window_set_target: no such window cloud");
    total += e.win[ids[k]].n;
  }
  Target& t = e.targets[0];
  cudaEventRecord(e.ev[0], e.stream);
  const bool ds = leaf > 0.f && total > 0;
  if (ds && e.ensure(e.win_cat, e.win_cat_cap, total)) return 1;
  if (e.reserve_target(t, total)) return 1;
  t.n = 0;
  float4* dst = ds ? e.win_cat : t.pts;
  size_t off = 0;
  for (int k = 0; k < n_ids; ++k) {
    const Engine::DevCloud& c = e.win[ids[k]];
    if (c.n && cudaMemcpyAsync(dst + off, c.p, c.n * sizeof(float4), cudaMemcpyDeviceToDevice, e.stream) != cudaSuccess) return e.fail("device copy failed");
    off += c.n;
  }
  size_t m = total;
  if (ds) {
    cudaError_t ce = voxelgrid_downsample(e.ws, e.win_cat, total, leaf, t.pts, &m, e.stream, &e.stats.kernel_launches);
    if (ce != cudaSuccess) return e.fail(std::string("voxelgrid_downsample: ") + cudaGetErrorString(ce));
  }
  t.use = nullptr;
  t.n = m;
  e.nn_valid = false;
  e.tgt_gen = 0;
  t.ts.have_bbox = false;
  if (e.build_target(t)) return 1;
  cudaEventRecord(e.ev[1], e.stream);
  if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  float ms = 0;
  cudaEventElapsedTime(&ms, e.ev[0], e.ev[1]);
  e.last_build_ms = ms;
  if (n_target) *n_target = m;
  return 0;
}
int ndtb_get_target(ndtb_handle* h, void* pts, size_t cap, size_t stride_bytes, size_t* n) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  if (stride_bytes < 12) return e.fail("stride_bytes must be >= 12");
  const Target& t = e.targets[0];
  const size_t m = t.n < cap ? t.n : cap;
  if (n) *n = m;
  if (m == 0) return 0;
  if (copy_cloud_to_host(e, t.use ? t.use : t.pts, m, pts, stride_bytes)) return 1;
  if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  return 0;
}

// ---- filter_node's front end (reference src/filter_node.cpp:236-262) -----------------------------------
void ndtb_default_filter_params(ndtb_filter_params* p) { p->r_min = 1.0; p->r_max = 60.0; p->leaf = 0.5f; p->mean_k = 30; p->stddev_mul = 1.0; }

static int copy_cloud_to_host(Engine& e, const float4* d, size_t m, void* pts, size_t stride_bytes) {
  if (stride_bytes == 16) {
    if (cudaMemcpyAsync(pts, d, m * 16, cudaMemcpyDeviceToHost, e.stream) != cudaSuccess) return e.fail("device copy failed");
  } else {
    if (cudaMemcpy2DAsync(pts, stride_bytes, d, 16, 12, m, cudaMemcpyDeviceToHost, e.stream) != cudaSuccess) return e.fail("device copy failed");
    if (stride_bytes >= 20)
      if (cudaMemcpy2DAsync(static_cast<char*>(pts) + 16, stride_bytes, reinterpret_cast<const char*>(d) + 12, 16, 4, m, cudaMemcpyDeviceToHost, e.stream) != cudaSuccess) return e.fail("device copy failed");
  }
  return 0;
}

int ndtb_filter_scan(ndtb_handle* h, const void* pts, size_t n, size_t stride_bytes, const ndtb_filter_params* p, size_t counts[3]) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  ndtb_filter_params fp;
  if (p) fp = *p; else ndtb_default_filter_params(&fp);
  if (fp.mean_k > 63) return e.fail("mean_k must be <= 63");
  size_t c[3] = {0, 0, 0};
  // the raw scan goes through the same upload path as a source, then is filtered in place of it
  if (ndtb_set_source(h, pts, n, stride_bytes)) return 1;
  if (n > 0) {
    if (!e.flt) { e.flt = filter_ws_create(); if (!e.flt) return e.fail("filter workspace allocation failed"); }
    if (n > e.flt_raw_cap) {
      cudaFree(e.flt_raw);
      e.flt_raw = nullptr;
      if (cudaMalloc(&e.flt_raw, (n + n / 4) * sizeof(float4)) != cudaSuccess) return e.fail("alloc");
      e.flt_raw_cap = n + n / 4;
    }
    float4* src = e.src_pool + e.sources[0].offset;
    if (cudaMemcpyAsync(e.flt_raw, src, n * sizeof(float4), cudaMemcpyDeviceToDevice, e.stream) != cudaSuccess) return e.fail("device copy failed");
    cudaError_t ce = filter_scan(e.flt, e.ws, e.flt_raw, n, fp.r_min, fp.r_max, fp.leaf, fp.mean_k, fp.stddev_mul, src, c, e.stream,
                                 &e.stats.kernel_launches);
    if (ce != cudaSuccess) return e.fail(std::string("filter_scan: ") + cudaGetErrorString(ce));
    e.sources[0].n = c[2];
    e.src_used = c[2];
  }
  if (counts) { counts[0] = c[0]; counts[1] = c[1]; counts[2] = c[2]; }
  return 0;
}

int ndtb_get_source(ndtb_handle* h, void* pts, size_t cap, size_t stride_bytes, size_t* n) {
  Engine& e = h->e;
  cudaSetDevice(e.device);
  if (e.sources.empty()) return e.fail("no source");
  if (stride_bytes < 12) return e.fail("stride_bytes must be >= 12");
  const size_t m = e.sources[0].n < cap ? e.sources[0].n : cap;
  if (n) *n = m;
  if (m == 0) return 0;
  if (copy_cloud_to_host(e, e.source_ptr(e.sources[0]), m, pts, stride_bytes)) return 1;
  if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  return 0;
}

// ---- batch -------------------------------------------------------------------------------
int ndtb_batch_create(int device, ndtb_batch** out) {
  if (!out) return 1;
  ndtb_batch* b = new ndtb_batch();
  if (create_engine(device, b->e)) { g_create_error = b->e.err; delete b; *out = nullptr; return 1; }
  *out = b;
  return 0;
}
int ndtb_batch_destroy(ndtb_batch* b) {
  if (!b) return 0;
  b->e.destroy();
  delete b;
  return 0;
}
const char* ndtb_batch_last_error(const ndtb_batch* b) { return b ? b->e.err.c_str() : g_create_error.c_str(); }
int ndtb_batch_set_params(ndtb_batch* b, const ndtb_params* p) {
  if (check_params(b->e, p)) return 1;
  b->e.params = *p;
  for (auto& t : b->e.targets) { t.built = false; t.ts.have_bbox = false; }
  return 0;
}
int ndtb_batch_clear(ndtb_batch* b) {
  Engine& e = b->e;
  cudaSetDevice(e.device);
  // keep the allocations (grow-only pools); forget the contents
  for (auto& t : e.targets) { t.n = 0; t.built = false; t.ts.have_bbox = false; t.use = nullptr; t.pending_upload = false; }
  e.n_targets_live = 0;
  e.sources_pending = false;
  e.sources.clear();
  e.src_used = 0;
  return 0;
}
int ndtb_batch_add_target(ndtb_batch* b, const void* pts, size_t n, size_t stride_bytes, int on_device, int* id) {
  Engine& e = b->e;
  cudaSetDevice(e.device);
  if (e.n_targets_live == (int)e.targets.size()) e.targets.emplace_back();
  Target& t = e.targets[e.n_targets_live];
  t.n = 0;
  t.use = nullptr;
  t.pending_upload = false;
  if (on_device == 2) {   // borrow: the table build reads the caller's device buffer in place
    if (stride_bytes != 16) return e.fail("a borrowed device buffer must hold 16-byte records");
    t.use = static_cast<const float4*>(pts);
  } else {
    if (e.reserve_target(t, n)) return 1;
    if (on_device == 0) {   // host memory: upload on the copy stream; the build of THIS target waits for it
      if (!t.uploaded && cudaEventCreateWithFlags(&t.uploaded, cudaEventDisableTiming) != cudaSuccess) return e.fail("event creation failed");
      if (e.upload_points(t.pts, pts, n, stride_bytes, false, e.copy_stream)) return 1;
      if (cudaEventRecord(t.uploaded, e.copy_stream) != cudaSuccess) return e.fail("event record failed");
      t.pending_upload = true;
    } else if (e.upload_points(t.pts, pts, n, stride_bytes, true, e.stream)) return 1;
  }
  t.n = n;
  t.built = false;
  t.ts.have_bbox = false;
  if (id) *id = e.n_targets_live;
  e.n_targets_live++;
  return 0;
}
int ndtb_batch_add_source(ndtb_batch* b, const void* pts, size_t n, size_t stride_bytes, int on_device, int* id) {
  Engine& e = b->e;
  cudaSetDevice(e.device);
  if (e.add_source(pts, n, stride_bytes, on_device, true, id)) return 1;
  if (on_device == 0 && cudaEventRecord(e.sources_uploaded, e.copy_stream) != cudaSuccess) return e.fail("event record failed");
  return 0;
}
int ndtb_batch_add_targets(ndtb_batch* b, int n, const void* const* pts, const size_t* ns, size_t stride_bytes, int on_device, int* first_id) {
  if (first_id) *first_id = b->e.n_targets_live;
  for (int i = 0; i < n; ++i)
    if (ndtb_batch_add_target(b, pts[i], ns[i], stride_bytes, on_device, nullptr)) return 1;
  return 0;
}
int ndtb_batch_add_sources(ndtb_batch* b, int n, const void* const* pts, const size_t* ns, size_t stride_bytes, int on_device, int* first_id) {
  if (first_id) *first_id = (int)b->e.sources.size();
  for (int i = 0; i < n; ++i)
    if (ndtb_batch_add_source(b, pts[i], ns[i], stride_bytes, on_device, nullptr)) return 1;
  return 0;
}
int ndtb_batch_wait_uploads(ndtb_batch* b) {
  Engine& e = b->e;
  cudaSetDevice(e.device);
  if (cudaStreamSynchronize(e.copy_stream) != cudaSuccess) return e.fail("upload failed");
  return 0;
}
int ndtb_batch_run(ndtb_batch* b, int n_items, const int* item_target, const int* item_source, const float* guesses,
                   ndtb_result* results) {
  Engine& e = b->e;
  for (int i = 0; i < n_items; ++i)
    if (item_target[i] < 0 || item_target[i] >= e.n_targets_live) return e.fail("item target id out of range");
  return e.run_items(n_items, item_target, item_source, guesses, results, false);
}
int ndtb_batch_fitness(ndtb_batch* b, double max_range, double* out) {
  Engine& e = b->e;
  const int n = (int)e.last_item_target.size();
  if (n == 0) return e.fail("ndtb_batch_fitness needs a completed ndtb_batch_run");
  if (cudaSetDevice(e.device) != cudaSuccess) return e.fail("cudaSetDevice failed");
  // blocks per registration: 8 when many registrations share a target (the C5 batch), up to 160 when each target has one
  // (loop-closure candidates), so that a launch always fills the device
  int g_max = 1;
  {
    std::vector<int> cnt(e.targets.size(), 0);
    for (int i = 0; i < n; ++i) g_max = std::max(g_max, ++cnt[e.last_item_target[i]]);
  }
  const int bpi = std::min(160, std::max(8, (e.target_blocks() + g_max - 1) / g_max));
  if (e.ensure(e.d_fit_partial, e.fit_partial_cap, (size_t)n * bpi * 2)) return 1;
  if (e.ensure(e.d_fit_out, e.fit_out_cap, (size_t)n)) return 1;
  if (e.ensure(e.d_fit_items, e.fit_items_cap, (size_t)n)) return 1;
  // registrations grouped by target: one NN grid build and one launch per target
  std::vector<int> order(n);
  for (int i = 0; i < n; ++i) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](int a, int c) { return e.last_item_target[a] < e.last_item_target[c]; });
  if (cudaMemcpyAsync(e.d_fit_items, order.data(), (size_t)n * sizeof(int), cudaMemcpyHostToDevice, e.stream) != cudaSuccess)
    return e.fail("upload failed");
  if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");   // `order` is pageable: landed before it goes away
  // distinct targets are independent (each item owns its slice of the partials): their grid builds and fitness launches go round
  // the build lanes' streams, one NN grid per lane, so that the ~12 small launches of a build overlap those of the other targets
  int n_groups = 0;
  for (int g0 = 0; g0 < n; ++n_groups) {
    const int tid = e.last_item_target[order[g0]];
    while (g0 < n && e.last_item_target[order[g0]] == tid) ++g0;
  }
  const int lanes = n_groups > 1 ? std::min(e.build_lanes(), n_groups) : 0;
  if (lanes) {
    if (e.ensure_lanes(lanes)) return 1;
    for (int g = 0; g < lanes; ++g)
      if (!e.lane_nn[g]) e.lane_nn[g] = nn_create();
    if (cudaEventRecord(e.bstart, e.stream) != cudaSuccess) return e.fail("event record failed");
    for (int g = 0; g < lanes; ++g)
      if (cudaStreamWaitEvent(e.bstream[g], e.bstart, 0) != cudaSuccess) return e.fail("stream wait failed");
  }
  int group = 0;
  for (int g0 = 0; g0 < n; ++group) {
    const int tid = e.last_item_target[order[g0]];
    int g1 = g0;
    while (g1 < n && e.last_item_target[order[g1]] == tid) ++g1;
    Target& t = e.targets[tid];
    if (t.n == 0 || !t.built) return e.fail("ndtb_batch_fitness: a target of the last run is gone (ndtb_batch_clear was called)");
    const float cell = e.params.resolution > 0.25f ? e.params.resolution : 0.25f;
    float bb[6];
    for (int a = 0; a < 3; ++a) { bb[a] = t.ts.mn[a]; bb[3 + a] = t.ts.mx[a]; }
    const int lane = lanes ? group % lanes : 0;
    NNGridStorage* nn = lanes ? e.lane_nn[lane] : e.nn;
    BuildWorkspace* ws = lanes ? e.bws[lane] : e.ws;
    cudaStream_t st = lanes ? e.bstream[lane] : e.stream;
    cudaError_t ce = nn_build(nn, ws, t.use ? t.use : t.pts, t.n, cell, t.ts.have_bbox ? bb : nullptr, st, &e.stats.kernel_launches);
    if (ce == cudaSuccess) ce = nn_fitness_batch(nn, e.d_states, e.d_fit_items + g0, g1 - g0, max_range, bpi, e.d_fit_partial, st);
    if (ce != cudaSuccess) return e.fail(std::string("batch fitness: ") + cudaGetErrorString(ce));
    ++e.stats.kernel_launches;
    g0 = g1;
  }
  for (int g = 0; g < lanes; ++g)
    if (cudaEventRecord(e.bdone[g], e.bstream[g]) != cudaSuccess || cudaStreamWaitEvent(e.stream, e.bdone[g], 0) != cudaSuccess)
      return e.fail("stream wait failed");
  e.nn_valid = false;
  if (fitness_finish(e.d_fit_partial, n, bpi, e.d_fit_out, e.stream) != cudaSuccess) return e.fail("fitness_finish failed");
  if (cudaMemcpyAsync(out, e.d_fit_out, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, e.stream) != cudaSuccess) return e.fail("copy failed");
  if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  return 0;
}
// ---- loop-closure submaps (reference src/pgo_node.cpp:404-476, 530-554, 568-590) -------------------------
int ndtb_batch_add_keyframe(ndtb_batch* b, const void* pts, size_t n, size_t stride_bytes, int* kf_id) {
  Engine& e = b->e;
  if (cudaSetDevice(e.device) != cudaSuccess) return e.fail("cudaSetDevice failed");
  if (n > 0 && stride_bytes < 12) return e.fail("stride_bytes must be >= 12");
  Engine::KeyFrame k;
  k.n = n;
  if (n > 0) {
    if (cudaMalloc(&k.p, n * sizeof(float4)) != cudaSuccess) return e.fail("keyframe allocation failed");
    // x y z intensity packed on the host as ndtb_set_source does, one 16 B / point upload
    if (cudaStreamSynchronize(e.stream) != cudaSuccess) { cudaFree(k.p); return e.fail("sync failed"); }
    if (e.st_src.reserve(n * sizeof(float4)) != cudaSuccess) { cudaFree(k.p); return e.fail("pinned staging allocation failed"); }
    float4* dst = static_cast<float4*>(e.st_src.h);
    const char* src = static_cast<const char*>(pts);
    for (size_t i = 0; i < n; ++i) {
      const float* f = reinterpret_cast<const float*>(src + i * stride_bytes);
      dst[i] = make_float4(f[0], f[1], f[2], stride_bytes >= 20 ? f[4] : (stride_bytes >= 16 ? f[3] : 0.f));
    }
    if (cudaMemcpyAsync(k.p, dst, n * sizeof(float4), cudaMemcpyHostToDevice, e.stream) != cudaSuccess ||
        cudaStreamSynchronize(e.stream) != cudaSuccess) {
      cudaFree(k.p);
      return e.fail("keyframe upload failed");
    }
  }
  e.kfs.push_back(k);
  if (kf_id) *kf_id = (int)e.kfs.size() - 1;
  return 0;
}
int ndtb_batch_clear_keyframes(ndtb_batch* b) {
  Engine& e = b->e;
  cudaSetDevice(e.device);
  cudaStreamSynchronize(e.stream);
  for (auto& k : e.kfs) cudaFree(k.p);
  e.kfs.clear();
  return 0;
}
int ndtb_batch_num_keyframes(const ndtb_batch* b) { return (int)b->e.kfs.size(); }

int ndtb_batch_add_submap(ndtb_batch* b, int as_target, int root, int half_window, const float* poses, int n_poses, float leaf,
                          int* id, size_t* n_points) {
  Engine& e = b->e;
  if (cudaSetDevice(e.device) != cudaSuccess) return e.fail("cudaSetDevice failed");
  if (n_points) *n_points = 0;
  if (half_window < 0) return e.fail("half_window must be >= 0");
  std::vector<SubmapSeg> segs;
  size_t total = 0, max_n = 0;
  for (long long i = -(long long)half_window; i <= (long long)half_window; ++i) {
    const long long k = (long long)root + i;
    if (k < 0 || k >= (long long)e.kfs.size() || k >= (long long)n_poses) continue;   // pgo_node.cpp:538-539
    const Engine::KeyFrame& kf = e.kfs[(size_t)k];
    if (kf.n == 0) continue;
    SubmapSeg sg;
    sg.src = kf.p;
    sg.offset = (long long)total;
    sg.n = kf.n;
    const float* T = poses + (size_t)k * 16;
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 4; ++c) sg.T[r * 4 + c] = T[c * 4 + r];
    segs.push_back(sg);
    total += kf.n;
    if (kf.n > max_n) max_n = kf.n;
  }
  // destination: a fresh batch target, or the tail of the source pool
  Target* t = nullptr;
  float4* dst = nullptr;
  if (as_target) {
    if (e.n_targets_live == (int)e.targets.size()) e.targets.emplace_back();
    t = &e.targets[e.n_targets_live];
    t->n = 0;
    t->use = nullptr;
    t->pending_upload = false;
    t->built = false;
    t->ts.have_bbox = false;
    if (e.reserve_target(*t, total)) return 1;
    dst = t->pts;
  } else {
    if (e.reserve_sources(total)) return 1;
    dst = e.src_pool + e.src_used;
  }
  size_t m = 0;
  if (total > 0) {
    const bool ds = leaf > 0.f;
    if (ds && e.ensure(e.kf_cat, e.kf_cat_cap, total)) return 1;
    if (e.ensure(e.d_segs, e.segs_cap, segs.size())) return 1;
    // `segs` is pageable: the copy is staged before cudaMemcpyAsync returns
    if (cudaMemcpyAsync(e.d_segs, segs.data(), segs.size() * sizeof(SubmapSeg), cudaMemcpyHostToDevice, e.stream) != cudaSuccess)
      return e.fail("upload failed");
    if (launch_submap_gather(e.d_segs, (int)segs.size(), max_n, ds ? e.kf_cat : dst, e.stream) != cudaSuccess) return e.fail("submap gather failed");
    ++e.stats.kernel_launches;
    m = total;
    if (ds) {
      cudaError_t ce = voxelgrid_downsample(e.ws, e.kf_cat, total, leaf, dst, &m, e.stream, &e.stats.kernel_launches);
      if (ce != cudaSuccess) return e.fail(std::string("voxelgrid_downsample: ") + cudaGetErrorString(ce));
    } else if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  }
  if (as_target) {
    t->n = m;
    if (id) *id = e.n_targets_live;
    e.n_targets_live++;
  } else {
    e.sources.push_back(Source{(long long)e.src_used, m, nullptr});
    if (id) *id = (int)e.sources.size() - 1;
    e.src_used += m;
  }
  if (n_points) *n_points = m;
  return 0;
}

// Many submaps in one call: destinations are reserved up front, then the build lanes (host threads with their own streams
// and sort workspaces) assemble different submaps concurrently -- a submap is ~15 short launches and two blocking reads.
int ndtb_batch_add_submaps(ndtb_batch* b, int n, const int* as_target, const int* root, const int* half_window, const float* poses,
                           int n_poses, float leaf, int* ids, size_t* n_points) {
  Engine& e = b->e;
  if (cudaSetDevice(e.device) != cudaSuccess) return e.fail("cudaSetDevice failed");
  if (n <= 0) return 0;
  struct Job { std::vector<SubmapSeg> segs; size_t total = 0, max_n = 0, m = 0; float4* dst = nullptr; int tslot = -1; long long soff = -1; };
  std::vector<Job> jobs((size_t)n);
  size_t src_total = 0;
  for (int j = 0; j < n; ++j) {
    if (half_window[j] < 0) return e.fail("half_window must be >= 0");
    Job& jb = jobs[j];
    for (long long i = -(long long)half_window[j]; i <= (long long)half_window[j]; ++i) {
      const long long k = (long long)root[j] + i;
      if (k < 0 || k >= (long long)e.kfs.size() || k >= (long long)n_poses) continue;   // pgo_node.cpp:538-539
      const Engine::KeyFrame& kf = e.kfs[(size_t)k];
      if (kf.n == 0) continue;
      SubmapSeg sg;
      sg.src = kf.p;
      sg.offset = (long long)jb.total;
      sg.n = kf.n;
      const float* T = poses + (size_t)k * 16;
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 4; ++c) sg.T[r * 4 + c] = T[c * 4 + r];
      jb.segs.push_back(sg);
      jb.total += kf.n;
      if (kf.n > jb.max_n) jb.max_n = kf.n;
    }
    if (!as_target[j]) src_total += jb.total;
  }
  // destinations: target slots first (the vector may grow), one reservation of the source pool for all source submaps
  for (int j = 0; j < n; ++j) {
    if (!as_target[j]) continue;
    if (e.n_targets_live == (int)e.targets.size()) e.targets.emplace_back();
    jobs[j].tslot = e.n_targets_live++;
    if (ids) ids[j] = jobs[j].tslot;
  }
  for (int j = 0; j < n; ++j) {
    if (jobs[j].tslot < 0) continue;
    Target& t = e.targets[jobs[j].tslot];
    t.n = 0; t.use = nullptr; t.pending_upload = false; t.built = false; t.ts.have_bbox = false;
    if (e.reserve_target(t, jobs[j].total)) return 1;
    jobs[j].dst = t.pts;
  }
  if (e.reserve_sources(src_total)) return 1;
  for (int j = 0; j < n; ++j) {
    if (as_target[j]) continue;
    jobs[j].soff = (long long)e.src_used;
    jobs[j].dst = e.src_pool + e.src_used;
    e.src_used += jobs[j].total;                 // a source keeps the room of its undownsampled stack
  }
  const int lanes = (int)std::min<size_t>((size_t)Engine::build_lanes(), (size_t)n);
  if (e.ensure_lanes(lanes)) return 1;
  if (cudaEventRecord(e.bstart, e.stream) != cudaSuccess) return e.fail("event record failed");
  std::string errs[Engine::kBuildLanes];
  long long launches[Engine::kBuildLanes] = {};
  std::atomic<int> next{0};
  const bool ds = leaf > 0.f;
  auto work = [&](int g) {
    if (cudaSetDevice(e.device) != cudaSuccess) { errs[g] = "cudaSetDevice failed"; return; }
    cudaStream_t bs = e.bstream[g];
    if (cudaStreamWaitEvent(bs, e.bstart, 0) != cudaSuccess) { errs[g] = "stream wait failed"; return; }
    for (int j = next.fetch_add(1); j < n; j = next.fetch_add(1)) {
      Job& jb = jobs[j];
      if (jb.total == 0) continue;
      if (ds && jb.total > e.lane_cat_cap[g]) {
        cudaFree(e.lane_cat[g]);
        e.lane_cat[g] = nullptr;
        e.lane_cat_cap[g] = jb.total + jb.total / 4 + 1024;
        if (cudaMalloc(&e.lane_cat[g], e.lane_cat_cap[g] * sizeof(float4)) != cudaSuccess) { e.lane_cat_cap[g] = 0; errs[g] = "submap buffer allocation failed"; return; }
      }
      if (jb.segs.size() > e.lane_segs_cap[g]) {
        cudaFree(e.lane_segs[g]);
        e.lane_segs[g] = nullptr;
        e.lane_segs_cap[g] = jb.segs.size() + 64;
        if (cudaMalloc(&e.lane_segs[g], e.lane_segs_cap[g] * sizeof(SubmapSeg)) != cudaSuccess) { e.lane_segs_cap[g] = 0; errs[g] = "segment table allocation failed"; return; }
      }
      cudaError_t ce = cudaMemcpyAsync(e.lane_segs[g], jb.segs.data(), jb.segs.size() * sizeof(SubmapSeg), cudaMemcpyHostToDevice, bs);
      if (ce == cudaSuccess) ce = launch_submap_gather(e.lane_segs[g], (int)jb.segs.size(), jb.max_n, ds ? e.lane_cat[g] : jb.dst, bs);
      ++launches[g];
      jb.m = jb.total;
      if (ce == cudaSuccess && ds) ce = voxelgrid_downsample(e.bws[g], e.lane_cat[g], jb.total, leaf, jb.dst, &jb.m, bs, &launches[g]);
      if (ce == cudaSuccess) ce = cudaStreamSynchronize(bs);     // jb.segs and the lane's buffers are reused by the next job
      if (ce != cudaSuccess) { errs[g] = std::string("submap assembly: ") + cudaGetErrorString(ce); return; }
    }
    if (cudaEventRecord(e.bdone[g], bs) != cudaSuccess) errs[g] = "event record failed";
  };
  std::vector<std::thread> th;
  for (int g = 1; g < lanes; ++g) th.emplace_back(work, g);
  work(0);
  for (auto& x : th) x.join();
  for (int g = 0; g < lanes; ++g) {
    if (!errs[g].empty()) return e.fail(errs[g]);
    e.stats.kernel_launches += launches[g];
    if (cudaStreamWaitEvent(e.stream, e.bdone[g], 0) != cudaSuccess) return e.fail("stream wait failed");
  }
  for (int j = 0; j < n; ++j) {
    if (jobs[j].tslot >= 0) {
      e.targets[jobs[j].tslot].n = jobs[j].m;
    } else {
      e.sources.push_back(Source{jobs[j].soff, jobs[j].m, nullptr});
      if (ids) ids[j] = (int)e.sources.size() - 1;
    }
    if (n_points) n_points[j] = jobs[j].m;
  }
  return 0;
}

int ndtb_batch_download(ndtb_batch* b, int is_target, int id, void* pts, size_t cap, size_t stride_bytes, size_t* n) {
  Engine& e = b->e;
  if (cudaSetDevice(e.device) != cudaSuccess) return e.fail("cudaSetDevice failed");
  if (stride_bytes < 12) return e.fail("stride_bytes must be >= 12");
  const float4* d = nullptr;
  size_t have = 0;
  if (is_target) {
    if (id < 0 || id >= e.n_targets_live) return e.fail("target id out of range");
    const Target& t = e.targets[id];
    if (t.pending_upload && cudaStreamSynchronize(e.copy_stream) != cudaSuccess) return e.fail("upload failed");
    d = t.use ? t.use : t.pts;
    have = t.n;
  } else {
    if (id < 0 || id >= (int)e.sources.size()) return e.fail("source id out of range");
    if (e.sources_pending && cudaStreamSynchronize(e.copy_stream) != cudaSuccess) return e.fail("upload failed");
    d = e.source_ptr(e.sources[id]);
    have = e.sources[id].n;
  }
  const size_t m = have < cap ? have : cap;
  if (n) *n = m;
  if (m == 0) return 0;
  if (copy_cloud_to_host(e, d, m, pts, stride_bytes)) return 1;
  if (cudaStreamSynchronize(e.stream) != cudaSuccess) return e.fail("sync failed");
  return 0;
}

int ndtb_batch_get_stats(ndtb_batch* b, ndtb_batch_stats* s) { *s = b->e.stats; return 0; }
int ndtb_batch_get_diag(ndtb_batch* b, int* replayed, int* borderline, int* n_rise, double* last_norm) {
  Engine& e = b->e;
  const size_t n = e.last_item_target.size();
  if (n == 0 || !e.h_states) return e.fail("ndtb_batch_get_diag needs a completed ndtb_batch_run");
  for (size_t i = 0; i < n; ++i) {
    const AlignState& st = e.h_states[i];
    if (replayed) replayed[i] = i < e.replay_flag.size() ? e.replay_flag[i] : 0;
    if (borderline) borderline[i] = st.borderline;
    if (n_rise) n_rise[i] = st.n_rise;
    if (last_norm) last_norm[i] = st.last_norm;
  }
  return 0;
}
int ndtb_batch_get_step_history(ndtb_batch* b, float* norms, int max_steps) {
  Engine& e = b->e;
  const size_t n = e.last_item_target.size();
  if (n == 0 || !e.h_states) return e.fail("ndtb_batch_get_step_history needs a completed ndtb_batch_run");
  if (!norms || max_steps < 1) return e.fail("ndtb_batch_get_step_history: norms must hold items x max_steps floats");
  for (size_t i = 0; i < n; ++i)
    for (int k = 0; k < max_steps; ++k) norms[i * (size_t)max_steps + k] = k < 32 ? e.h_states[i].norm_hist[k] : 0.f;
  return 0;
}
int ndtb_batch_set_stream(ndtb_batch* b, void* cuda_stream) {
  Engine& e = b->e;
  cudaSetDevice(e.device);
  if (e.stream && e.own_stream) { cudaStreamSynchronize(e.stream); cudaStreamDestroy(e.stream); }
  e.stream = (cudaStream_t)cuda_stream;
  e.own_stream = false;
  return 0;
}

}  // extern "C"

// ---- one job over several devices ---------------------------------------------------------------
#include <chrono>
struct ndtb_multi {
  std::vector<int> devices;
  std::vector<ndtb_batch*> batches;
  std::string err;
  std::vector<int> st_items, st_targets;
  std::vector<double> st_dev_ms, st_wall_ms;
};

extern "C" {

int ndtb_multi_create(int n_devices, const int* devices, ndtb_multi** out) {
  if (!out) return 1;
  *out = nullptr;
  if (n_devices <= 0 || !devices) { g_create_error = "ndtb_multi_create needs at least one device"; return 1; }
  ndtb_multi* m = new ndtb_multi();
  for (int k = 0; k < n_devices; ++k) {
    ndtb_batch* b = nullptr;
    if (ndtb_batch_create(devices[k], &b)) {
      for (auto* q : m->batches) ndtb_batch_destroy(q);
      delete m;
      return 1;                                         // g_create_error holds the batch's message
    }
    m->devices.push_back(devices[k]);
    m->batches.push_back(b);
  }
  m->st_items.assign(n_devices, 0); m->st_targets.assign(n_devices, 0);
  m->st_dev_ms.assign(n_devices, 0.0); m->st_wall_ms.assign(n_devices, 0.0);
  *out = m;
  return 0;
}
int ndtb_multi_destroy(ndtb_multi* m) {
  if (!m) return 0;
  for (auto* b : m->batches) ndtb_batch_destroy(b);
  delete m;
  return 0;
}
const char* ndtb_multi_last_error(const ndtb_multi* m) { return m ? m->err.c_str() : g_create_error.c_str(); }
int ndtb_multi_set_params(ndtb_multi* m, const ndtb_params* p) {
  for (auto* b : m->batches)
    if (ndtb_batch_set_params(b, p)) { m->err = ndtb_batch_last_error(b); return 1; }
  return 0;
}

int ndtb_multi_run(ndtb_multi* m, int n_targets, const void* const* target_pts, const size_t* target_n, int n_sources,
                   const void* const* source_pts, const size_t* source_n, size_t stride_bytes, int n_items, const int* item_target,
                   const int* item_source, const float* guesses, ndtb_result* results) {
  const int D = (int)m->batches.size();
  for (int d = 0; d < D; ++d) { m->st_items[d] = m->st_targets[d] = 0; m->st_dev_ms[d] = m->st_wall_ms[d] = 0.0; }
  if (n_items <= 0) return 0;
  for (int i = 0; i < n_items; ++i) {
    if (item_target[i] < 0 || item_target[i] >= n_targets) { m->err = "item target id out of range"; return 1; }
    if (item_source[i] < 0 || item_source[i] >= n_sources) { m->err = "item source id out of range"; return 1; }
  }
  // plan (shard.plan_shards in Python): whole submaps dealt round-robin in target-id order when that leaves the devices' item
  // counts within max(1, 2 %) of each other -- every table built once, neighbouring submaps (hard or easy together) on different
  // devices -- else items by target id (stable) in contiguous blocks of n/D (+1 for the first n%D devices)
  std::vector<int> order((size_t)n_items);
  for (int i = 0; i < n_items; ++i) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return item_target[a] < item_target[b]; });
  std::vector<int> dev_first((size_t)D, 0), dev_count((size_t)D, 0);
  {
    std::vector<int> rank_of_target((size_t)n_targets, -1), sizes((size_t)D, 0);
    int distinct = 0;
    for (int k = 0; k < n_items; ++k) {
      int& r = rank_of_target[item_target[order[k]]];
      if (r < 0) r = distinct++ % D;
      ++sizes[r];
    }
    const int mx = *std::max_element(sizes.begin(), sizes.end()), mn = *std::min_element(sizes.begin(), sizes.end());
    if (D > 1 && distinct >= D && mx - mn <= std::max(1, (int)(0.02 * n_items / D))) {
      std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return rank_of_target[item_target[a]] < rank_of_target[item_target[b]]; });
      int pos = 0;
      for (int d = 0; d < D; ++d) { dev_first[d] = pos; dev_count[d] = sizes[d]; pos += sizes[d]; }
    } else {
      const int base = n_items / D, extra = n_items % D;
      int pos = 0;
      for (int d = 0; d < D; ++d) { dev_first[d] = pos; dev_count[d] = base + (d < extra ? 1 : 0); pos += dev_count[d]; }
    }
  }
  std::vector<std::string> errs((size_t)D);
  auto work = [&](int d, int first, int count) {
    const auto t0 = std::chrono::steady_clock::now();
    ndtb_batch* b = m->batches[d];
    auto failb = [&]() { errs[d] = std::string("device ") + std::to_string(m->devices[d]) + ": " + ndtb_batch_last_error(b); };
    if (ndtb_batch_clear(b)) return failb();
    // local ids in order of first use
    std::vector<int> tmap((size_t)n_targets, -1), smap((size_t)n_sources, -1), lt((size_t)count), ls((size_t)count);
    std::vector<const void*> tp, sp;
    std::vector<size_t> tn, sn;
    std::vector<float> G((size_t)count * 16);
    for (int k = 0; k < count; ++k) {
      const int i = order[first + k];
      int& a = tmap[item_target[i]];
      if (a < 0) { a = (int)tp.size(); tp.push_back(target_pts[item_target[i]]); tn.push_back(target_n[item_target[i]]); }
      int& c = smap[item_source[i]];
      if (c < 0) { c = (int)sp.size(); sp.push_back(source_pts[item_source[i]]); sn.push_back(source_n[item_source[i]]); }
      lt[k] = a; ls[k] = c;
      std::memcpy(&G[(size_t)k * 16], guesses + (size_t)i * 16, 64);
    }
    if (ndtb_batch_add_targets(b, (int)tp.size(), tp.data(), tn.data(), stride_bytes, 0, nullptr)) return failb();
    if (ndtb_batch_add_sources(b, (int)sp.size(), sp.data(), sn.data(), stride_bytes, 0, nullptr)) return failb();
    std::vector<ndtb_result> res((size_t)count);
    if (ndtb_batch_run(b, count, lt.data(), ls.data(), G.data(), res.data())) return failb();
    for (int k = 0; k < count; ++k) results[order[first + k]] = res[k];      // the gather: disjoint entries per device
    ndtb_batch_stats st;
    ndtb_batch_get_stats(b, &st);
    m->st_items[d] = count; m->st_targets[d] = (int)tp.size(); m->st_dev_ms[d] = st.total_ms;
    m->st_wall_ms[d] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  };
  std::vector<std::thread> th;
  for (int d = 0; d < D; ++d)
    if (dev_count[d] > 0) th.emplace_back(work, d, dev_first[d], dev_count[d]);
  for (auto& t : th) t.join();
  for (int d = 0; d < D; ++d)
    if (!errs[d].empty()) { m->err = errs[d]; return 1; }
  return 0;
}

int ndtb_multi_get_stats(ndtb_multi* m, int* items, int* targets, double* device_ms, double* wall_ms) {
  for (size_t d = 0; d < m->batches.size(); ++d) {
    if (items) items[d] = m->st_items[d];
    if (targets) targets[d] = m->st_targets[d];
    if (device_ms) device_ms[d] = m->st_dev_ms[d];
    if (wall_ms) wall_ms[d] = m->st_wall_ms[d];
  }
  return 0;
}

}  // extern "C"
