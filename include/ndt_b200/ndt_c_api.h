/* ndt_c_api.h -- C ABI of the B200-native NDT scan-matching backend (libndt_b200.so).
 *
 * Drop-in boundary for the one hot path of lowmee/xchu_slam: the registration backend that
 * odom_node calls.  The reference has no FFI; its interface for this path is the C++ class
 * pclomp::NormalDistributionsTransform (reference xchu_mapping/include/pclomp/ndt_omp.h:70-204,
 * NDT_H below) over pcl::Registration.  Every entry point cites the member it replaces.
 * include/pclomp/ndt_omp.h in this repo re-creates that class on top of this ABI so that
 * xchu_mapping/src/odom_node.cpp:55-83, 263-284 compiles unchanged.
 *
 * Conventions: every function returns 0 on success, non-zero on failure (ndtb_last_error
 * gives the text); nothing throws across the ABI.  All pointers are HOST pointers unless a
 * function says otherwise; the library copies inside the call and retains no caller pointer.
 * Point clouds are arrays of records whose first three floats are x, y, z; `stride_bytes`
 * is the record size (16 for float4 / pcl::PointXYZ, 32 for pcl::PointXYZI).  4x4 matrices
 * are float, COLUMN-major (Eigen::Matrix4f memory order).  A handle owns one CUDA stream and
 * is not thread-safe; distinct handles may be used from distinct threads.
 * There is NO CPU fallback: without a CUDA device ndtb_create fails.
 */
#ifndef NDT_B200_C_API_H_
#define NDT_B200_C_API_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ndtb_handle ndtb_handle;
typedef struct ndtb_batch ndtb_batch;

/* pclomp::NeighborSearchMethod, NDT_H:52-57 (same numeric values). */
enum { NDTB_KDTREE = 0, NDTB_DIRECT26 = 1, NDTB_DIRECT7 = 2, NDTB_DIRECT1 = 3 };

/* Everything the reference configures through setters.  Defaults = the reference ctor
 * (ndt_omp_impl.hpp:46-69) and VoxelGridCovariance ctor (voxel_grid_covariance_omp.h:199-207). */
typedef struct ndtb_params {
  float resolution;      /* setResolution,            NDT_H:127-137   default 1.0  */
  double step_size;      /* setStepSize,              NDT_H:160-164   default 0.1  */
  double trans_eps;      /* setTransformationEpsilon, pcl::Registration default 0.1 (ctor NDT_I:64) */
  int max_iter;          /* setMaximumIterations,     default 35 (ctor NDT_I:65)   */
  double outlier_ratio;  /* setOulierRatio,           NDT_H:178-182   default 0.55 */
  int search_method;     /* setNeighborhoodSearchMethod, NDT_H:184-186 default DIRECT7 */
  int min_points;        /* min_points_per_voxel_,    VGC_H:204       default 6    */
  double eig_mult;       /* min_covar_eigvalue_mult_, VGC_H:205       default 0.01 */
  int strict_math;       /* which derivative kernel runs.  -1 (default) and 1: the term-by-term kernel that reproduces every
                            float operation of NDT_I:491-548 in the reference's order (bit-identical per-pair terms; only the
                            grouping of the FP64 sums differs, 1e-16): every registration -- a handle's or a batch's -- ends on the
                            reference arithmetic's iteration count and transform, and a whole odometry sequence replays the
                            reference's trajectory exactly.  0: the per-point factorised kernel only (4x faster per evaluation;
                            score / gradient / Hessian within 1e-7 relative of the reference arithmetic, but ~1 % of the
                            registrations of a C5-like job -- those whose Newton iteration amplifies any perturbation -- end
                            outside 1e-4 m / 1e-5 rad of the reference's result).  -2: HYBRID, for an ndtb_batch only (a handle
                            runs the strict kernel): every registration runs with the factorised kernel, and those that took
                            more than `replay_iters` iterations, did not converge, ended on a Newton step shorter than
                            `replay_min_norm`, solved an ill-conditioned system, or took a discrete decision of the loop (the
                            |step| < eps test NDT_I:152-157, the descent-direction flip NDT_I:774-784) within 1e-3 of its
                            threshold are replayed from their initial guess with the term-by-term kernel inside the same
                            ndtb_batch_run.  1.9x the throughput of the default; measured on six synthetic worlds of 4096
                            registrations each, 24571 of 24576 results are inside the gate (the criterion was tuned on the
                            first world, which has no miss) */
  long long dense_max_cells; /* cell -> voxel lookups go through a dense int32 index over the target's bounding box
                            when that box (padded by 2 cells) holds at most this many cells, else through the
                            hash table; 0 = hash only.  default 2^26 (256 MB of index per target at most) */
  int replay_iters;      /* hybrid batches (strict_math = -2): iteration count above which a registration is replayed with
                            the term-by-term kernel; <= 0: only non-converged / borderline registrations.  default 11 */
  double replay_min_norm;/* hybrid batches: a registration whose LAST Newton step |delta_p| (NDT_I:119-125) was shorter than this
                            is replayed too -- its final move is the clamped eps / 2 along delta_p / |delta_p|, which a tiny
                            |delta_p| leaves ill-defined.  < 0 (default): 0.05 * trans_eps */
} ndtb_params;

typedef struct ndtb_result {
  float T[16];              /* getFinalTransformation()                                   */
  int iterations;           /* getFinalNumIteration(), NDT_H:200-204                      */
  int converged;            /* hasConverged()                                             */
  double score;             /* last computeDerivatives score                              */
  double trans_probability; /* getTransformationProbability(), NDT_H:191-195              */
  int n_evals;              /* computeDerivatives/computeHessian evaluations performed    */
} ndtb_result;

/* ---- lifetime ------------------------------------------------------------------------ */
int ndtb_create(int device, ndtb_handle** out);            /* NormalDistributionsTransform(), NDT_I:46-69 */
int ndtb_destroy(ndtb_handle* h);
const char* ndtb_last_error(const ndtb_handle* h);          /* h may be NULL: last create error */
void ndtb_default_params(ndtb_params* p);

/* ---- configuration (NDT_H:110-186 setters; a change of resolution rebuilds the voxel grid
 *      when a target is present, as init() does, NDT_H:127-137, 271-278) ------------------ */
int ndtb_set_params(ndtb_handle* h, const ndtb_params* p);
int ndtb_get_params(const ndtb_handle* h, ndtb_params* p);

/* ---- target: VoxelGridCovariance build (NDT_H:117-122 setInputTarget -> init() ->
 *      voxel_grid_covariance_omp_impl.hpp:48-370 applyFilter) ----------------------------- 
 * Points with a NaN / Inf coordinate are skipped, as the reference's first pass skips them for non-dense clouds
 * (voxel_grid_covariance_omp_impl.hpp:209-215); the finite points keep their relative order. */
int ndtb_set_target(ndtb_handle* h, const void* pts, size_t n, size_t stride_bytes);
/* If the first points of `pts` are bit-identical to the current target (odom_node only ever
 * appends between window resets, src/odom_node.cpp:336), only the new tail is merged into the
 * voxel table; otherwise a full rebuild.  *was_incremental (may be NULL) reports which. */
int ndtb_set_target_auto(ndtb_handle* h, const void* pts, size_t n, size_t stride_bytes, int* was_incremental);
/* Append points to the current target: cpu::NormalDistributionsTransform::updateVoxelGrid
 * semantics (include/ndt_cpu/NormalDistributionsTransform.h:39); result == full rebuild over the
 * concatenated cloud. */
int ndtb_update_target(ndtb_handle* h, const void* pts, size_t n, size_t stride_bytes);

/* ---- source + align (pcl::Registration::setInputSource / align ->
 *      ndt_omp_impl.hpp:73-164 computeTransformation) ------------------------------------- */
int ndtb_set_source(ndtb_handle* h, const void* pts, size_t n, size_t stride_bytes);
int ndtb_align(ndtb_handle* h, const float guess[16], ndtb_result* out);
/* Transformed source of the last align (the `output` cloud of align()); x,y,z written at the
 * start of each `stride_bytes` record, other bytes untouched. */
int ndtb_get_output(ndtb_handle* h, void* pts, size_t stride_bytes);
/* pcl::Registration::getFitnessScore(max_range): mean squared 1-NN distance of the source moved
 * by the final transformation (or by T if not NULL) to the RAW target cloud. */
int ndtb_fitness(ndtb_handle* h, double max_range, const float* T, double* out);
/* calculateScore, ndt_omp_impl.hpp:919-952 (cloud = source transformed by T). */
int ndtb_calculate_score(ndtb_handle* h, const float T[16], double* out);

/* ---- inspection hooks used by the parity tests ----------------------------------------- */
int ndtb_num_voxels(ndtb_handle* h, size_t* n_all);  /* all occupied voxels (any point count)  */
/* Arrays of n_all entries (any may be NULL): absolute cell coordinates, nr_points (-1 = flagged
 * invalid, VGC_I:337-341/360-364), inserted point count, mean (f64), inverse covariance as the
 * derivative kernel reads it (f32, row-major 3x3), f32 centroid (VGC_I:242-243, 289). */
int ndtb_dump_voxels(ndtb_handle* h, int* ijk, int* nr_points, int* n_raw, double* mean, float* icov,
                     float* centroid);

/* Leaf::icov_ as the device holds it in f64 (VGC_H:213 getInverseCov; the kernels of computeHessian / calculateScore read this
 * copy, the derivative kernels its float rounding), [n_all][9] row-major, same voxel order as ndtb_dump_voxels. */
int ndtb_dump_icov64(ndtb_handle* h, double* icov);
/* NormalDistributionsTransform::convertTransform (NDT_H:210-229): [x y z roll pitch yaw] -> column-major float 4x4,
 * Translation * AngleAxis(X) * AngleAxis(Y) * AngleAxis(Z) in float, the chain the device uses.  Host only, no device call. */
void ndtb_pose_to_matrix(const double p[6], float T[16]);
/* Page-locked host memory (cudaHostAlloc, portable across devices) for clouds handed to the upload entry points: pageable
 * memory works too but is staged by the driver at a fraction of the link rate.  0 on success. */
int ndtb_host_alloc(size_t bytes, void** out);
int ndtb_host_free(void* p);
/* Neighbour voxels (VGC_I:373-442 / VGC_H:470-499) of n already transformed xyz points;
 * cnt[n], ijk[n][27][3] in visiting order. */
int ndtb_neighbors(ndtb_handle* h, const float* xyz, size_t n, int method, int* cnt, int* ijk);
/* getFitnessScore without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The value is
 * ndtb_fitness's (mean, in double, of the float squared distances <= max_range from every source point moved by T -- column-major
 * 4x4, host memory -- to its exact nearest target point; DBL_MAX when no point is in range), computed by the exhaustive query
 * kernel (n_source x n_target distance evaluations) instead of the cell grid, whose build allocates.
 *   source_dev / target_dev: DEVICE pointers on `device`, 16-byte records (x y z _), caller-owned; out_dev: DEVICE pointer to one
 *   double; cuda_stream: a cudaStream_t (NULL = the default stream).  Enqueues three kernels and returns WITHOUT synchronising;
 *   calls no cudaMalloc / cudaFree / allocating library routine.
 * Scratch: a file-scope __device__ array of NDTB_NOALLOC_MAX_SOURCE floats + 256 B (1 MiB; every GPU that loads the library holds
 * it until exit).  n_source beyond that limit is rejected (returns 1, text from ndtb_noalloc_last_error).  Every call shares the
 * array, which keeps the last call's contents -- each call re-arms what must start clean -- so two calls that overlap on the GPU
 * (different streams or threads) RACE: serialise them.  Returns 0 on success. */
#define NDTB_NOALLOC_MAX_SOURCE 262144
int ndtb_fitness_noalloc(int device, const void* source_dev, size_t n_source, const void* target_dev, size_t n_target, const float T[16],
                         double max_range, double* out_dev, void* cuda_stream);
const char* ndtb_noalloc_last_error(void);   /* per calling thread */
/* One computeDerivatives (ndt_omp_impl.hpp:175-283) at transform T and pose vector p. */
int ndtb_eval_derivatives(ndtb_handle* h, const float T[16], const double p[6], int compute_hessian,
                          double* score, double g[6], double H[36]);
/* computeHessian (ndt_omp_impl.hpp:551-641): FP64, radius-search neighbours. */
int ndtb_compute_hessian(ndtb_handle* h, const float T[16], const double p[6], double H[36]);
/* Per-evaluation trace of the last ndtb_align: records of 52 doubles
 * {p[6], score, g[6], H[36], step, compute_hessian, kind}; kind 0 = initial, 1 = first trial of a
 * line search, 2 = More-Thuente trial, 3 = computeHessian.  *n = records available. */
int ndtb_get_trace(ndtb_handle* h, double* out, int max_recs, int* n);
/* Device time (ms, CUDA events on the handle's stream) of the last set_target / update / align. */
int ndtb_last_timing(ndtb_handle* h, double* build_ms, double* align_ms);

/* ---- odom_node's localmap kept on the device ------------------
 * What LidarOdom::OdomEstimate does to its clouds either side of align() (reference
 * xchu_mapping/src/odom_node.cpp, ODOM below), with the clouds resident in HBM: the only per-scan
 * host<->device traffic left is the scan itself (ndtb_set_source) and the 4x4 result.
 *   localmap / tmp_map : pcl::PointCloud<PointXYZI> localmap, tmp_map        (ODOM members)
 *   snapshot           : pc_target_.reset(new PointCloud(localmap))           (ODOM:233)
 * Points carry intensity (ndtb_set_source keeps the float at byte offset 16 of 32-byte records). */
/* localmap.clear(); tmp_map.clear(); snapshot = empty. */
int ndtb_localmap_clear(ndtb_handle* h);
/* transformed = T * current source (pcl::transformPointCloud, ODOM:220 / :290); when leaf > 0 it is then
 * voxel-centroid downsampled like downSizeFilterLocalmap (pcl::VoxelGrid, leaf^3 cells, ODOM:98, 334-335);
 * localmap += transformed (ODOM:221 / :336) and, when also_tmp != 0, tmp_map += transformed (ODOM:337).
 * *n_added (may be NULL) = points appended. */
int ndtb_localmap_append_source(ndtb_handle* h, const float T[16], float leaf, int also_tmp, size_t* n_added);
/* pc_target_.reset(new PointCloud(localmap)) (ODOM:233): remembers the localmap as it is now. */
int ndtb_localmap_snapshot(ndtb_handle* h);
/* setInputTarget(pc_target_) (ODOM:228 / :349) with pc_target_ = the last snapshot.  Only the points the
 * snapshot gained since the previous call are merged into the voxel table (result identical to a full
 * rebuild); after ndtb_localmap_roll, or when `force_full` != 0, the table is rebuilt.  No host traffic. */
int ndtb_localmap_set_target(ndtb_handle* h, int force_full, int* was_incremental);
/* localmap = tmp_map; tmp_map.clear() (ODOM:352-356). */
int ndtb_localmap_roll(ndtb_handle* h);
/* sizes[0..2] = points in localmap, tmp_map, snapshot. */
int ndtb_localmap_sizes(ndtb_handle* h, size_t sizes[3]);
/* Copies a cloud to the host: which = 0 localmap, 1 tmp_map, 2 the cloud appended by the last
 * ndtb_localmap_append_source.  Records of stride_bytes: x, y, z at offset 0 and, when stride_bytes >= 20,
 * intensity at offset 16.  `cap` = records `pts` can hold; *n = records written. */
int ndtb_localmap_download(ndtb_handle* h, int which, void* pts, size_t cap, size_t stride_bytes, size_t* n);

/* ---- the keyframe-count (sliding window) localmap on the device -------------------------------
 * LidarOdom::ExtractSurroundKeyframes (reference xchu_mapping/src/odom_node.cpp:737-802, ODOM below): the target is the
 * newest surround_search_num_ (20) keyframes, each moved by its pose and downsampled with downSizeFilterLocalmap (1.0 m,
 * ODOM:749-753 / :767-771), concatenated newest first (ODOM:778-781), downsampled with downSizeFilterKeyFrames (0.5 m,
 * ODOM:783-784) and handed to setInputTarget (ODOM:786-791) -- a NON-prefix cloud at every refresh, i.e. a full voxel-table
 * build at keyframe rate.  The clouds stay in HBM; which keyframes are in the window (recent_keyframes_, a deque the caller
 * pushes and pops, including the reference's `static latest_frame_id` quirk) stays the caller's business: it holds ids. */
/* tf_cloud = downSizeFilterLocalmap(T * current source) kept on the device; *id names it (ids count up from 0; an id is not reused until
 * ndtb_window_clear, which drops every cloud and restarts the count at 0).
 * leaf <= 0: no downsample.  *n_points (may be NULL) = its size. */
int ndtb_window_add_source(ndtb_handle* h, const float T[16], float leaf, int* id, size_t* n_points);
int ndtb_window_drop(ndtb_handle* h, int id);      /* the cloud leaves recent_keyframes_ for good */
int ndtb_window_clear(ndtb_handle* h);             /* every cloud dropped, ids restart at 0 */
/* pc_target_ = the listed clouds concatenated in the order given, downsampled at `leaf` (<= 0: not), setInputTarget(pc_target_).
 * *n_target (may be NULL) = points of pc_target_. */
int ndtb_window_set_target(ndtb_handle* h, const int* ids, int n_ids, float leaf, size_t* n_target);
/* Copies the current target cloud to the host (whatever set it); records as in ndtb_localmap_download.  Test hook. */
int ndtb_get_target(ndtb_handle* h, void* pts, size_t cap, size_t stride_bytes, size_t* n);

/* ---- filter_node's front end on the device ---------------------
 * What CloudFilter::PcCB does to a raw scan before odom_node sees it (reference xchu_mapping/src/filter_node.cpp:236-262):
 * pcl::removeNaNFromPointCloud, the range gate r_min < sqrt(x^2 + y^2) < r_max (:242-247), downSizeFilterKeyFrames
 * (pcl::VoxelGrid, leaf 0.5 m, :33-34, 249-251) and pcl::StatisticalOutlierRemoval (mean_k 30, stddev_mul 1.0, :254-262).
 * The filtered scan becomes the handle's source (as if passed to ndtb_set_source) and stays in HBM. */
typedef struct ndtb_filter_params {
  double r_min, r_max;   /* 1.0, 60.0 */
  float leaf;            /* 0.5; <= 0 skips the downsample */
  int mean_k;            /* 30; <= 0 skips the outlier removal; at most 63 */
  double stddev_mul;     /* 1.0 */
} ndtb_filter_params;
void ndtb_default_filter_params(ndtb_filter_params* p);
/* pts: n raw records (x y z at offset 0; intensity at byte offset 16 when stride_bytes >= 20, else the 4th float of
 * 16-byte records).  counts[0..2] (may be NULL) = points after the gate, after the downsample, after the outlier removal. */
int ndtb_filter_scan(ndtb_handle* h, const void* pts, size_t n, size_t stride_bytes, const ndtb_filter_params* p, size_t counts[3]);
/* Copies the handle's current source to the host (records as in ndtb_localmap_download). */
int ndtb_get_source(ndtb_handle* h, void* pts, size_t cap, size_t stride_bytes, size_t* n);

/* ---- batched registration on one device (north_star item 4) ------------------------------
 * A batch holds targets (submaps) and sources (scans) resident in HBM and runs any number of
 * independent registrations {target id, source id, initial guess} concurrently; no collective
 * inside a registration.  `on_device`: 0 = `pts` is a HOST pointer (uploaded asynchronously on the
 * batch's copy stream, overlapped with the voxel-table builds of earlier targets: the buffer must stay
 * valid until ndtb_batch_run returns); 1 = DEVICE pointer on the batch's device, copied device-to-device;
 * 2 = DEVICE pointer BORROWED in place (16-byte records, no copy; valid until ndtb_batch_clear). */
int ndtb_batch_create(int device, ndtb_batch** out);
int ndtb_batch_destroy(ndtb_batch* b);
const char* ndtb_batch_last_error(const ndtb_batch* b);
int ndtb_batch_set_params(ndtb_batch* b, const ndtb_params* p);
int ndtb_batch_clear(ndtb_batch* b);
int ndtb_batch_add_target(ndtb_batch* b, const void* pts, size_t n, size_t stride_bytes, int on_device, int* id);
int ndtb_batch_add_source(ndtb_batch* b, const void* pts, size_t n, size_t stride_bytes, int on_device, int* id);
/* The same for n clouds in one call (pts[i]: ns[i] records of stride_bytes): ids first_id .. first_id + n - 1.  A batch of
 * thousands of scans is registered with two calls instead of thousands. */
int ndtb_batch_add_targets(ndtb_batch* b, int n, const void* const* pts, const size_t* ns, size_t stride_bytes, int on_device, int* first_id);
int ndtb_batch_add_sources(ndtb_batch* b, int n, const void* const* pts, const size_t* ns, size_t stride_bytes, int on_device, int* first_id);
/* Blocks until every host upload enqueued by ndtb_batch_add_* so far has landed in HBM (lets a caller with several
 * batches in flight stage one batch's uploads behind another batch's compute). */
int ndtb_batch_wait_uploads(ndtb_batch* b);
/* Builds every target's voxel table (if not built yet) and runs n_items registrations.
 * guesses: n_items x 16 floats.  results: n_items records (host). */
int ndtb_batch_run(ndtb_batch* b, int n_items, const int* item_target, const int* item_source,
                   const float* guesses, ndtb_result* results);
/* pcl::Registration::getFitnessScore(max_range) of every registration of the last ndtb_batch_run, at its final
 * transform: loop-closure candidates are accepted on it (reference xchu_mapping/src/pgo_node.cpp:460-476, fitness <= 0.3).
 * out: one double per item of that run (host).  Exact 1-NN over each target; one launch per distinct target. */
int ndtb_batch_fitness(ndtb_batch* b, double max_range, double* out);
/* ---- loop-closure submaps on the device --------------------
 * Replaces PGO::LoopFindNearKeyframesCloud (reference xchu_mapping/src/pgo_node.cpp:530-554) and feeds the batch the
 * clouds PGO::ICPRefine (pgo_node.cpp:404-476) registers: keyframe clouds (sensor frame) stay resident in HBM
 * (keyframeLaserClouds, pgo_node.cpp:216); a submap is the concatenation, in ascending keyframe order, of the keyframes
 * root - half_window .. root + half_window that exist, each moved into the map frame by its pose as
 * PGO::TransformCloud2Map does (pgo_node.cpp:568-590: x' = ((m00 x + m01 y) + m02 z) + m03 in float, intensity kept),
 * then pcl::VoxelGrid-downsampled at `leaf` (downSizeFilterICP, pgo_node.cpp:75, 551-552; leaf <= 0: no downsample).
 * Keyframes survive ndtb_batch_clear; ids count from 0 in the order added. */
int ndtb_batch_add_keyframe(ndtb_batch* b, const void* pts, size_t n, size_t stride_bytes, int* kf_id);
int ndtb_batch_clear_keyframes(ndtb_batch* b);
int ndtb_batch_num_keyframes(const ndtb_batch* b);
/* poses: one column-major 4x4 float matrix per keyframe id (what pcl::getTransformation makes of keyframePosesUpdated[id]),
 * n_poses of them; keyframes with id >= n_poses or id >= the number stored are skipped like the reference's range test
 * (pgo_node.cpp:538-539).  as_target != 0: the submap becomes a batch target (id = target id), else a batch source.
 * n_points (may be NULL) = points of the submap; an empty submap (no keyframe in range) still gets an id and holds 0 points. */
int ndtb_batch_add_submap(ndtb_batch* b, int as_target, int root, int half_window, const float* poses, int n_poses, float leaf,
                          int* id, size_t* n_points);
/* n submaps in one call (arrays of n; poses / n_poses / leaf shared): the batch's build lanes assemble them concurrently.
 * ids[j] / n_points[j] as ndtb_batch_add_submap returns them. */
int ndtb_batch_add_submaps(ndtb_batch* b, int n, const int* as_target, const int* root, const int* half_window, const float* poses,
                           int n_poses, float leaf, int* ids, size_t* n_points);
/* Copies target / source `id` of the batch to the host (x y z at offset 0, intensity at byte 12 for 16-byte records, at
 * byte 16 for stride_bytes >= 20): test hook, and what a caller publishing the submap needs. */
int ndtb_batch_download(ndtb_batch* b, int is_target, int id, void* pts, size_t cap, size_t stride_bytes, size_t* n);

/* Device times (ms) of the last run: voxel-table builds, derivative kernels, whole run; number of
 * kernels launched; derivative-kernel launches; (point, voxel) pairs and points evaluated. */
typedef struct ndtb_batch_stats {
  double build_ms, deriv_ms, total_ms;
  long long kernel_launches, deriv_launches;
  long long points_evaluated, pairs_evaluated, probes;
  /* hybrid batches: registrations replayed with the term-by-term kernel, device time of that pass (ms), of its derivative
   * launches, their number, and the points / pairs they evaluated.  deriv_ms / deriv_launches / points_evaluated /
   * pairs_evaluated above count the first (factorised-kernel) pass only. */
  long long replayed;
  double replay_ms, replay_deriv_ms;
  long long replay_deriv_launches, replay_points, replay_pairs;
} ndtb_batch_stats;
int ndtb_batch_get_stats(ndtb_batch* b, ndtb_batch_stats* s);
/* Per item of the last ndtb_batch_run (arrays of that many entries, any may be NULL): replayed[i] = 1 when the hybrid mode ran
 * registration i again with the term-by-term kernel; borderline[i] / n_rise[i] / last_norm[i] = the loop diagnostics of its FINAL
 * pass (a discrete decision taken within 1e-3 of its threshold; line-search steps that grew after having dropped below
 * step_size; |delta_p| of the last Newton solve, NDT_I:119-125).  Test / tuning hook. */
int ndtb_batch_get_diag(ndtb_batch* b, int* replayed, int* borderline, int* n_rise, double* last_norm);
/* norms[i * max_steps + k] = |delta_p| of the Newton solve of iteration k of item i's FINAL pass (NDT_I:119-125; the first 32
 * iterations, 0 beyond the last one).  Test / tuning hook: the data the replay criterion of the hybrid mode is studied on. */
int ndtb_batch_get_step_history(ndtb_batch* b, float* norms, int max_steps);
/* Run all of the batch's copies and kernels on the caller's CUDA stream (a cudaStream_t) instead of
 * the batch's own, so that the caller's CUDA events bracket the work. */
int ndtb_batch_set_stream(ndtb_batch* b, void* cuda_stream);

/* ---- one job over several GPUs of one box (north_star item 4) ---------
 * The C5 job -- n_items independent registrations over n_targets submaps -- cut over the listed devices: items are ordered by
 * target id (stable) and dealt in contiguous blocks whose sizes differ by at most one, so that every submap's voxel table is
 * built on as few devices as possible; a device uploads and builds only the targets and sources its block names.  One host
 * thread per device inside the library, one ndtb_batch each; there is NO collective: a single registration does not shard, the
 * only exchange is the gather of the result records into `results` (item order), which is a host memcpy here.  All pointers
 * are HOST pointers.  The same device may be listed more than once (two batches on one GPU; what the tests on a one-GPU box do).
 * Results do not depend on how the job was cut up to the FP64 summation order of a derivative evaluation (block sizes follow
 * the number of points a device has in flight), i.e. to ~1e-16 relative per evaluation. */
typedef struct ndtb_multi ndtb_multi;
int ndtb_multi_create(int n_devices, const int* devices, ndtb_multi** out);
int ndtb_multi_destroy(ndtb_multi* m);
const char* ndtb_multi_last_error(const ndtb_multi* m);
int ndtb_multi_set_params(ndtb_multi* m, const ndtb_params* p);
int ndtb_multi_run(ndtb_multi* m, int n_targets, const void* const* target_pts, const size_t* target_n, int n_sources,
                   const void* const* source_pts, const size_t* source_n, size_t stride_bytes, int n_items, const int* item_target,
                   const int* item_source, const float* guesses, ndtb_result* results);
/* Per device of the last ndtb_multi_run (arrays of n_devices, any may be NULL): items it ran, targets it built, its device
 * time (ms, CUDA events) and the host wall time of its thread (ms). */
int ndtb_multi_get_stats(ndtb_multi* m, int* items, int* targets, double* device_ms, double* wall_ms);

#ifdef __cplusplus
}
#endif
#endif /* NDT_B200_C_API_H_ */
