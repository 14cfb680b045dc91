// odom_sequence.cpp -- headless C++ mirror of xchu_slam's scan-to-localmap odometry loop
// (reference xchu_mapping/src/odom_node.cpp:208-356 LidarOdom::OdomEstimate, :42-83 ParamInitial)
// driving the drop-in pclomp::NormalDistributionsTransform of include/pclomp/ndt_omp.h.
//
// It is the caller side of BASELINE config C2 (a sequence of scans registered against a localmap
// that is refreshed by the distance-threshold strategy, every refresh going through
// setInputTarget -> incremental voxel rebuild) and doubles as the compile/run test of the shim:
// the NDT calls below are the reference's own lines (odom_node.cpp:70-78, 228, 278-283, 349).
// ROS plumbing, IMU/odometry guesses and publishers are out of scope.
//
//   odom_sequence <sequence.bin> <out.txt> [--res R] [--refresh distance|surround] [--no-incremental] [--device-localmap] [--strict|--fast]
//                 [--threads N] [--max-scans N] [--device D]
//
// --refresh distance (default): LidarOdom::Run as shipped -- the localmap is refreshed by OdomEstimate's distance-threshold
//   strategy alone (odom_node.cpp:329-356).
// --refresh surround: the keyframe-count sliding window, i.e. LidarOdom::Run with its commented call
//   `ExtractSurroundKeyframes();` (odom_node.cpp:186, 201) enabled in front of OdomEstimate and nothing else changed:
//   the newest surround_search_num_ = 20 keyframes, each moved by its pose and 1.0 m-downsampled, concatenated newest first,
//   0.5 m-downsampled, setInputTarget (odom_node.cpp:737-802) -- a full voxel-table build per scan, because every scan is
//   pushed to cloud_keyframes_ (:303-304).  OdomEstimate's own setInputTarget(pc_target_) (:349) still runs, as in the
//   reference; the window's target replaces it before the next align.
//
// sequence.bin: "NDTSEQ1\0", int32 n_scans, then per scan: int32 n, double gt[6] (sensor pose
// x y z roll pitch yaw), float xyzi[n][4] (filter_node output, sensor frame).
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <deque>
#include <map>
#include <string>
#include <vector>

#ifdef ODOM_USE_REFERENCE
// The SAME caller compiled against the reference's own pclomp::NormalDistributionsTransform
// (/root/reference/xchu_mapping/include/pclomp/*, over the stand-in Eigen/PCL headers of oracle/ref_stubs/):
// the CPU arm of config C2 and the trajectory oracle for the drop-in.  Test infrastructure; built as
// examples/odom_sequence_ref only where /root/reference exists.
#include <pcl/point_types.h>
#include <pcl/registration/registration.h>
#include <pclomp/ndt_omp.h>
#include <pclomp/ndt_omp_impl.hpp>
#include <pclomp/voxel_grid_covariance_omp_impl.hpp>
#else
#include "../include/pclomp/ndt_omp.h"
#endif

typedef pcl::PointXYZI PointT;
typedef pcl::PointCloud<PointT> Cloud;

struct Pose6D {   // include/xchu_mapping/common.h
  double x = 0, y = 0, z = 0, roll = 0, pitch = 0, yaw = 0;
  Pose6D operator+(const Pose6D& o) const { Pose6D r; r.x = x + o.x; r.y = y + o.y; r.z = z + o.z; r.roll = roll + o.roll; r.pitch = pitch + o.pitch; r.yaw = yaw + o.yaw; return r; }
  Pose6D operator-(const Pose6D& o) const {
    Pose6D r; r.x = x - o.x; r.y = y - o.y; r.z = z - o.z; r.roll = roll - o.roll; r.pitch = pitch - o.pitch;
    double d = yaw - o.yaw;   // calcDiffForRadian
    if (d >= M_PI) d -= 2 * M_PI; else if (d < -M_PI) d += 2 * M_PI;
    r.yaw = d;
    return r;
  }
};

// common.h:64-71: Translation * Rz * Ry * Rx (double), cast to float by the caller
static Eigen::Matrix4f Pose6D2Matrix(const Pose6D& p) {
  const double cr = cos(p.roll), sr = sin(p.roll), cp = cos(p.pitch), sp = sin(p.pitch), cy = cos(p.yaw), sy = sin(p.yaw);
  const double R[3][3] = {{cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr},
                          {sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr},
                          {-sp, cp * sr, cp * cr}};
  Eigen::Matrix4f m = Eigen::Matrix4f::Identity();
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) m(r, c) = (float)R[r][c];
  m(0, 3) = (float)p.x; m(1, 3) = (float)p.y; m(2, 3) = (float)p.z;
  return m;
}
// common.h:51-63: tf getRPY of the rotation block
static Pose6D Matrix2Pose6D(const Eigen::Matrix4f& m) {
  Pose6D p;
  p.x = m(0, 3); p.y = m(1, 3); p.z = m(2, 3);
  p.pitch = asin(fmax(-1.0, fmin(1.0, -(double)m(2, 0))));
  p.roll = atan2((double)m(2, 1), (double)m(2, 2));
  p.yaw = atan2((double)m(1, 0), (double)m(0, 0));
  return p;
}
// pcl::transformPointCloud (float, left to right)
static void transformPointCloud(const Cloud& in, Cloud& out, const Eigen::Matrix4f& T) {
  out.points.resize(in.points.size());
  for (size_t i = 0; i < in.points.size(); ++i) {
    const PointT& p = in.points[i];
    PointT q = p;
    q.x = ((T(0, 0) * p.x + T(0, 1) * p.y) + T(0, 2) * p.z) + T(0, 3);
    q.y = ((T(1, 0) * p.x + T(1, 1) * p.y) + T(1, 2) * p.z) + T(1, 3);
    q.z = ((T(2, 0) * p.x + T(2, 1) * p.y) + T(2, 2) * p.z) + T(2, 3);
    out.points[i] = q;
  }
  out.width = (decltype(out.width))out.points.size();
  out.height = 1;
}
// downSizeFilterLocalmap (odom_node.cpp:98, 334-335): pcl::VoxelGrid<PointXYZI>, downsample_all_data_ = true -- one point
// per occupied leaf in ascending leaf index (k, j, i), float centroids of x, y, z, intensity over the leaf's points in
// input order, divided by float(count) (the restatement of oracle/thirdparty_math.h; PCL is not in this image).
static void voxelDownsample(const Cloud& in, Cloud& out, float leaf) {
  const size_t n = in.points.size();
  Cloud r;
  if (n) {
    const float inv = 1.0f / leaf;
    float mn[3] = {3.402823466e38f, 3.402823466e38f, 3.402823466e38f}, mx[3] = {-3.402823466e38f, -3.402823466e38f, -3.402823466e38f};
    for (const PointT& p : in.points) {
      const float v[3] = {p.x, p.y, p.z};
      for (int a = 0; a < 3; ++a) { mn[a] = v[a] < mn[a] ? v[a] : mn[a]; mx[a] = v[a] > mx[a] ? v[a] : mx[a]; }
    }
    int min_b[3], max_b[3];
    for (int a = 0; a < 3; ++a) { min_b[a] = (int)floorf(mn[a] * inv); max_b[a] = (int)floorf(mx[a] * inv); }
    const int d0 = max_b[0] - min_b[0] + 1, d1 = max_b[1] - min_b[1] + 1;
    std::vector<std::pair<unsigned, unsigned>> iv(n);
    for (size_t i = 0; i < n; ++i) {
      const PointT& p = in.points[i];
      const int a = (int)(floorf(p.x * inv) - (float)min_b[0]), b = (int)(floorf(p.y * inv) - (float)min_b[1]), c = (int)(floorf(p.z * inv) - (float)min_b[2]);
      iv[i] = std::make_pair((unsigned)(a + b * d0 + c * d0 * d1), (unsigned)i);
    }
    std::stable_sort(iv.begin(), iv.end(), [](const std::pair<unsigned, unsigned>& x, const std::pair<unsigned, unsigned>& y) { return x.first < y.first; });
    for (size_t b = 0; b < n;) {
      size_t e = b + 1;
      while (e < n && iv[e].first == iv[b].first) ++e;
      const PointT& p0 = in.points[iv[b].second];
      float c[4] = {p0.x, p0.y, p0.z, p0.intensity};
      for (size_t q = b + 1; q < e; ++q) {
        const PointT& p = in.points[iv[q].second];
        c[0] += p.x; c[1] += p.y; c[2] += p.z; c[3] += p.intensity;
      }
      const float cnt = (float)(e - b);
      PointT o;
      o.x = c[0] / cnt; o.y = c[1] / cnt; o.z = c[2] / cnt; o.intensity = c[3] / cnt;
      r.points.push_back(o);
      b = e;
    }
  }
  r.width = (decltype(r.width))r.points.size();
  r.height = 1;
  out = r;
}

int main(int argc, char** argv) {
  if (argc < 3) { fprintf(stderr, "usage: odom_sequence <sequence.bin> <out.txt> [--res R] [--no-incremental] [--device-localmap] [--strict|--fast] [--threads N] [--max-scans N] [--device D]\n"); return 2; }
  float ndt_res = 2.0f;              // odom_node.cpp:43 default
  const double step_size = 0.1, trans_eps = 0.01, min_add_scan_shift = 0.5;   // odom_node.cpp:44-47
  const int max_iter = 30;
  double max_localmap_size = 5;      // odom_node.cpp:49
  bool incremental = true, device_localmap = false, surround = false;
  const int surround_search_num_ = 20;   // odom_node.cpp:48
  const float down_size = 0.5f;          // odom_node.cpp:97-98: downSizeFilterKeyFrames 0.5 m, downSizeFilterLocalmap 1.0 m
  int strict = -1;   // library default: a handle runs the term-by-term kernel
  int max_scans = 1 << 30, device = 0, threads = 1;   // odom_node.cpp:74 runs ndt_omp with one thread
  for (int a = 3; a < argc; ++a) {
    if (!strcmp(argv[a], "--res") && a + 1 < argc) ndt_res = (float)atof(argv[++a]);
    else if (!strcmp(argv[a], "--no-incremental")) incremental = false;
    else if (!strcmp(argv[a], "--device-localmap")) device_localmap = true;   // localmap kept in HBM (ndtb_localmap_*)
    else if (!strcmp(argv[a], "--refresh") && a + 1 < argc) surround = !strcmp(argv[++a], "surround");
    else if (!strcmp(argv[a], "--strict")) strict = 1;                         // term-by-term derivative kernel (the default for a handle)
    else if (!strcmp(argv[a], "--fast")) strict = 0;                           // factorised derivative kernel
    else if (!strcmp(argv[a], "--max-scans") && a + 1 < argc) max_scans = atoi(argv[++a]);
    else if (!strcmp(argv[a], "--device") && a + 1 < argc) device = atoi(argv[++a]);
    else if (!strcmp(argv[a], "--threads") && a + 1 < argc) threads = atoi(argv[++a]);
    else if (!strcmp(argv[a], "--submap") && a + 1 < argc) max_localmap_size = atof(argv[++a]);
  }
  FILE* f = fopen(argv[1], "rb");
  if (!f) { perror(argv[1]); return 2; }
  char magic[8];
  int n_scans = 0;
  if (fread(magic, 1, 8, f) != 8 || memcmp(magic, "NDTSEQ1", 7) != 0 || fread(&n_scans, 4, 1, f) != 1) { fprintf(stderr, "bad sequence file\n"); return 2; }
  if (n_scans > max_scans) n_scans = max_scans;

  // ---- ParamInitial, odom_node.cpp:67-80 ---------------------------------------------------
  pclomp::NormalDistributionsTransform<PointT, PointT>::Ptr omp_ndt;
  try {
#ifdef ODOM_USE_REFERENCE
    (void)device;
    pclomp::NormalDistributionsTransform<PointT, PointT>::Ptr ndt(new pclomp::NormalDistributionsTransform<PointT, PointT>());
#else
    pclomp::NormalDistributionsTransform<PointT, PointT>::Ptr ndt(new pclomp::NormalDistributionsTransform<PointT, PointT>(device));
#endif
#ifndef ODOM_USE_REFERENCE
    if (strict >= 0) ndt->setStrictMath(strict != 0);
#endif
    ndt->setNeighborhoodSearchMethod(pclomp::DIRECT7);
    ndt->setNumThreads(threads);
    ndt->setTransformationEpsilon(trans_eps);
    ndt->setStepSize(step_size);
    ndt->setResolution(ndt_res);
    ndt->setMaximumIterations(max_iter);
    omp_ndt = ndt;
  } catch (const std::exception& e) {
    fprintf(stderr, "%s\n", e.what());
    return 3;
  }
  FILE* out = fopen(argv[2], "w");
  if (!out) { perror(argv[2]); return 2; }

  Cloud localmap, tmp_map;
  Cloud::Ptr pc_target_(new Cloud()), transformed_scan_ptr(new Cloud());
  Pose6D previous_pose, current_pose, diff_pose, guess_pose;
  double localmap_size = 0, shift_dis = 0;
  int initial_scan_loaded = 0, n_incremental = 0, n_target_sets = 0;
  // ExtractSurroundKeyframes state (odom_node.h:128-135)
  std::vector<Cloud::Ptr> cloud_keyframes_;
  std::vector<Eigen::Matrix4f> cloud_keyposes_;
  std::deque<Cloud::Ptr> recent_keyframes_;
  std::deque<int> recent_ids_;          // device mode: the window clouds' ids, same deque discipline
  std::vector<int> keyframe_win_id_;    // device mode: window cloud of keyframe k
  int latest_frame_id = -1;             // the function-local `static int latest_frame_id` of odom_node.cpp:762
  bool latest_frame_id_set = false;
  int n_surround_sets = 0;
  double total_surround_ms = 0;
  double total_align_ms = 0, total_target_ms = 0, total_fitness_ms = 0, total_source_ms = 0;
  const auto wall0 = std::chrono::steady_clock::now();

  for (int s = 0; s < n_scans; ++s) {
    int n = 0;
    double gt[6];
    if (fread(&n, 4, 1, f) != 1 || fread(gt, 8, 6, f) != 6) { fprintf(stderr, "truncated sequence\n"); return 2; }
    Cloud::Ptr filtered_scan_ptr(new Cloud());
    filtered_scan_ptr->points.resize(n);
    std::vector<float> buf((size_t)n * 4);
    if (fread(buf.data(), 16, n, f) != (size_t)n) { fprintf(stderr, "truncated sequence\n"); return 2; }
    for (int i = 0; i < n; ++i) { PointT& p = filtered_scan_ptr->points[i]; p.x = buf[i * 4]; p.y = buf[i * 4 + 1]; p.z = buf[i * 4 + 2]; p.intensity = buf[i * 4 + 3]; }
    filtered_scan_ptr->width = n; filtered_scan_ptr->height = 1;
    if (s == 0) {   // start the odometry at the first scan's pose (the reference starts at init_x.. = 0)
      previous_pose.x = gt[0]; previous_pose.y = gt[1]; previous_pose.z = gt[2]; previous_pose.roll = gt[3]; previous_pose.pitch = gt[4]; previous_pose.yaw = gt[5];
      current_pose = previous_pose;
    }
    // ---- ExtractSurroundKeyframes, odom_node.cpp:737-802 (only with --refresh surround) ------
    if (surround && !cloud_keyposes_.empty()) {
      const auto tw0 = std::chrono::steady_clock::now();
      bool target_updated = false;
      const size_t have = device_localmap ? recent_ids_.size() : recent_keyframes_.size();
      if ((int)have < surround_search_num_) {                           // :743-760
        recent_keyframes_.clear();
        recent_ids_.clear();
        for (int i = (int)cloud_keyposes_.size() - 1; i >= 0; --i) {
          const int this_key_id = i;
          if (device_localmap) {
#ifndef ODOM_USE_REFERENCE
            recent_ids_.push_back(keyframe_win_id_[this_key_id]);       // the cloud :749-753 computes, made when the keyframe was stored
#endif
          } else {
            Cloud::Ptr tf_cloud(new Cloud());
            transformPointCloud(*cloud_keyframes_[this_key_id], *tf_cloud, cloud_keyposes_[this_key_id]);
            voxelDownsample(*tf_cloud, *tf_cloud, down_size * 2);
            recent_keyframes_.push_back(tf_cloud);
          }
          if ((int)(device_localmap ? recent_ids_.size() : recent_keyframes_.size()) >= surround_search_num_) break;
        }
        target_updated = true;
      } else {                                                          // :761-775
        if (!latest_frame_id_set) { latest_frame_id = (int)cloud_keyposes_.size() - 1; latest_frame_id_set = true; }
        if (latest_frame_id != (int)cloud_keyposes_.size() - 1) {
          latest_frame_id = (int)cloud_keyposes_.size() - 1;
          if (device_localmap) {
#ifndef ODOM_USE_REFERENCE
            recent_ids_.pop_back();
            recent_ids_.push_front(keyframe_win_id_[latest_frame_id]);
#endif
          } else {
            recent_keyframes_.pop_back();
            Cloud::Ptr tf_cloud(new Cloud());
            transformPointCloud(*cloud_keyframes_[latest_frame_id], *tf_cloud, cloud_keyposes_[latest_frame_id]);
            voxelDownsample(*tf_cloud, *tf_cloud, down_size * 2);
            recent_keyframes_.push_front(tf_cloud);
          }
          target_updated = true;
        }
      }
      if (target_updated) {                                             // :777-792
        if (device_localmap) {
#ifndef ODOM_USE_REFERENCE
          omp_ndt->windowSetTarget(recent_ids_, down_size);
#endif
        } else {
          pc_target_->clear();
          for (auto keyframe : recent_keyframes_) *pc_target_ += *keyframe;
          voxelDownsample(*pc_target_, *pc_target_, down_size);
          omp_ndt->setInputTarget(pc_target_);
        }
        ++n_surround_sets;
      }
      total_surround_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tw0).count();
    }
    transformed_scan_ptr->clear();
    // ---- OdomEstimate, odom_node.cpp:218-231 ----------------------------------------------
#ifndef ODOM_USE_REFERENCE
    if (device_localmap) {   // the scan is the only cloud that crosses the bus
      const auto ts0 = std::chrono::steady_clock::now();
      omp_ndt->setInputSource(filtered_scan_ptr);
      total_source_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - ts0).count();
    }
#endif
    if (initial_scan_loaded == 0) {
#ifndef ODOM_USE_REFERENCE
      if (device_localmap) {
        omp_ndt->localmapAppendSource(Pose6D2Matrix(previous_pose), 0.f, false);
        omp_ndt->localmapSnapshot();
        omp_ndt->localmapSetTarget();
      } else
#endif
      {
        transformPointCloud(*filtered_scan_ptr, *transformed_scan_ptr, Pose6D2Matrix(previous_pose));
        localmap += *transformed_scan_ptr;
        *pc_target_ += *transformed_scan_ptr;
        omp_ndt->setInputTarget(pc_target_);
      }
      initial_scan_loaded = 1;
    }
    size_t target_points = 0;
#ifndef ODOM_USE_REFERENCE
    if (device_localmap) { omp_ndt->localmapSnapshot(); target_points = omp_ndt->localmapSize(2); } else
#endif
    { pc_target_.reset(new Cloud(localmap)); target_points = pc_target_->size(); }   // :233
    guess_pose = previous_pose + diff_pose;                              // :237-239
    guess_pose.pitch = previous_pose.pitch;
    guess_pose.roll = previous_pose.roll;
    const Eigen::Matrix4f init_guess = Pose6D2Matrix(guess_pose);        // :256
    Cloud::Ptr output_cloud(new Cloud());
    const auto t0 = std::chrono::steady_clock::now();
#ifndef ODOM_USE_REFERENCE
    if (device_localmap) omp_ndt->alignNoOutput(init_guess); else
#endif
    {
      omp_ndt->setInputSource(filtered_scan_ptr);                        // :278-283
      omp_ndt->align(*output_cloud, init_guess);
    }
    const auto tf0 = std::chrono::steady_clock::now();
    const double fitness_score = omp_ndt->getFitnessScore();
    total_fitness_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tf0).count();
    const Eigen::Matrix4f t_localizer = omp_ndt->getFinalTransformation();
    const bool has_converged = omp_ndt->hasConverged();
    const int final_num_iteration = omp_ndt->getFinalNumIteration();
    const auto t1 = std::chrono::steady_clock::now();
    if (!device_localmap) transformPointCloud(*filtered_scan_ptr, *transformed_scan_ptr, t_localizer);   // :290
    current_pose = Matrix2Pose6D(t_localizer);                           // :291-296 (tf_l2b = identity)
    if (surround) {                                                      // :303-304: EVERY scan is a keyframe of the window
      cloud_keyposes_.push_back(t_localizer);
      if (device_localmap) {
#ifndef ODOM_USE_REFERENCE
        keyframe_win_id_.push_back(omp_ndt->windowAddSource(t_localizer, down_size * 2));
        // a keyframe the deque can no longer reach is dropped: the window only ever looks at the newest surround_search_num_ + 1
        const int dead = (int)keyframe_win_id_.size() - 1 - (surround_search_num_ + 2);
        if (dead >= 0 && std::find(recent_ids_.begin(), recent_ids_.end(), keyframe_win_id_[dead]) == recent_ids_.end() && keyframe_win_id_[dead] >= 0) {
          omp_ndt->windowDrop(keyframe_win_id_[dead]);
          keyframe_win_id_[dead] = -1;
        }
#endif
      } else {
        cloud_keyframes_.push_back(filtered_scan_ptr);
      }
    }
    diff_pose = current_pose - previous_pose;                            // :309
    shift_dis = sqrt(pow(current_pose.x - previous_pose.x, 2.0) + pow(current_pose.y - previous_pose.y, 2.0));   // :322
    previous_pose = current_pose;
    double target_ms = 0;
    int was_inc = -1;
    if (shift_dis >= min_add_scan_shift) {                               // :329-350
      localmap_size += shift_dis;
      auto t2 = std::chrono::steady_clock::now();
#ifndef ODOM_USE_REFERENCE
      if (device_localmap) {
        omp_ndt->localmapAppendSource(t_localizer, 1.0f, true);          // :290, 334-337 on the device
        omp_ndt->localmapSetTarget(!incremental);                        // :349
        was_inc = omp_ndt->lastTargetUpdateWasIncremental() ? 1 : 0;
      } else
#endif
      {
      voxelDownsample(*transformed_scan_ptr, *transformed_scan_ptr, 1.0f);
      localmap += *transformed_scan_ptr;
      tmp_map += *transformed_scan_ptr;
      t2 = std::chrono::steady_clock::now();
#ifdef ODOM_USE_REFERENCE
      omp_ndt->setInputTarget(pc_target_);                               // :349, full rebuild
      was_inc = 0;
#else
      if (incremental) omp_ndt->setInputTarget(pc_target_);             // prefix-append detection inside
      else { ndtb_set_target(omp_ndt->handle(), pc_target_->points.data(), pc_target_->points.size(), sizeof(PointT)); }
      was_inc = incremental && omp_ndt->lastTargetUpdateWasIncremental() ? 1 : 0;
#endif
      }
      target_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t2).count();
      n_incremental += was_inc;
      n_target_sets++;
    }
    if (localmap_size >= max_localmap_size) {                            // :352-356
#ifndef ODOM_USE_REFERENCE
      if (device_localmap) omp_ndt->localmapRoll(); else
#endif
      { localmap = tmp_map; tmp_map.clear(); }
      localmap_size = 0.0;
    }
    const double align_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    total_align_ms += align_ms;
    total_target_ms += target_ms;
    fprintf(out, "%d %d %d %.9g %.9f %.9f %.9f %.9f %.9f %.9f %zu %d %.4f %.4f %.6f %.6f %.6f\n", s, final_num_iteration, has_converged ? 1 : 0,
            fitness_score, current_pose.x, current_pose.y, current_pose.z, current_pose.roll, current_pose.pitch, current_pose.yaw,
            target_points, was_inc, align_ms, target_ms, gt[0], gt[1], gt[2]);
  }
  const double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - wall0).count();
  fprintf(out, "# scans %d wall_s %.4f scans_per_s %.2f align_ms_mean %.4f target_ms_mean %.4f target_sets %d incremental %d fitness_ms_mean %.4f source_ms_mean %.4f surround_sets %d surround_ms_mean %.4f\n",
          n_scans, wall, n_scans / wall, total_align_ms / n_scans, n_target_sets ? total_target_ms / n_target_sets : 0.0, n_target_sets, n_incremental,
          total_fitness_ms / n_scans, total_source_ms / n_scans, n_surround_sets, n_surround_sets ? total_surround_ms / n_surround_sets : 0.0);
  fclose(out);
  fclose(f);
  return 0;
}
