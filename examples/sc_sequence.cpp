// sc_sequence.cpp -- headless mirror of what pgo_node does with its SCManager (reference src/pgo_node.cpp:58, 236, 344):
// every keyframe of a sequence file is described (makeAndSaveScancontextAndKeys) and queried (detectLoopClosureID); the loop
// id and yaw difference of every keyframe go to the output file.  Compiled twice from this one source:
//   default              over include/scancontext/Scancontext.h  (device implementation behind the C ABI)
//   -DSC_USE_REFERENCE   over the reference's own scancontext/Scancontext.{h,cpp} (CPU; test infrastructure, built by
//                        xchu_slam_b200/build.py against the stand-in headers of oracle/ref_stubs/)
// usage: sc_sequence <sequence.bin> <out.txt> [--every K] [--revisit a,b,...] [--thres T]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <string>
#include <vector>

#ifdef SC_USE_REFERENCE
#include "scancontext/Scancontext.cpp"
#else
#include <scancontext/Scancontext.h>
#endif

static int run(int argc, char** argv) {
  if (argc < 3) { fprintf(stderr, "usage: sc_sequence <sequence.bin> <out.txt> [--every K] [--revisit a,b,...] [--thres T]\n"); return 2; }
  int every = 1;
  double thres = -1;
  std::vector<int> revisit;
  for (int i = 3; i < argc; ++i) {
    if (!strcmp(argv[i], "--every") && i + 1 < argc) every = atoi(argv[++i]);
    else if (!strcmp(argv[i], "--thres") && i + 1 < argc) thres = atof(argv[++i]);
    else if (!strcmp(argv[i], "--revisit") && i + 1 < argc) {
      std::string s = argv[++i];
      size_t pos = 0;
      while (pos < s.size()) { revisit.push_back(atoi(s.c_str() + pos)); pos = s.find(',', pos); if (pos == std::string::npos) break; ++pos; }
    }
  }
  FILE* f = fopen(argv[1], "rb");
  if (!f) { fprintf(stderr, "cannot open %s\n", argv[1]); return 2; }
  char magic[8];
  int n_scans = 0;
  if (fread(magic, 1, 8, f) != 8 || memcmp(magic, "NDTSEQ1", 7) != 0 || fread(&n_scans, 4, 1, f) != 1) { fprintf(stderr, "bad sequence file\n"); return 2; }
  std::vector<pcl::PointCloud<SCPointType>> keyframes;
  std::vector<float> buf;
  for (int s = 0; s < n_scans; ++s) {
    int n = 0;
    double gt[6];
    if (fread(&n, 4, 1, f) != 1 || fread(gt, 8, 6, f) != 6) { fprintf(stderr, "truncated sequence\n"); return 2; }
    buf.resize((size_t)n * 4);
    if (fread(buf.data(), 16, n, f) != (size_t)n) { fprintf(stderr, "truncated sequence\n"); return 2; }
    if (s % every) continue;
    pcl::PointCloud<SCPointType> c;
    c.points.resize(n);
    for (int i = 0; i < n; ++i) { c.points[i].x = buf[4 * i]; c.points[i].y = buf[4 * i + 1]; c.points[i].z = buf[4 * i + 2]; c.points[i].intensity = buf[4 * i + 3]; }
    keyframes.push_back(c);
  }
  fclose(f);
  const size_t n_drive = keyframes.size();
  for (int r : revisit)
    if (r >= 0 && (size_t)r < n_drive) keyframes.push_back(keyframes[r]);
  FILE* out = fopen(argv[2], "w");
  if (!out) { fprintf(stderr, "cannot open %s\n", argv[2]); return 2; }
  SCManager scManager;
  if (thres > 0) scManager.setSCdistThres(thres);                    // pgo_node.cpp:58
  int loops = 0;
  const auto t0 = std::chrono::steady_clock::now();
  for (size_t k = 0; k < keyframes.size(); ++k) {
    scManager.makeAndSaveScancontextAndKeys(keyframes[k]);            // pgo_node.cpp:236
    auto detectResult = scManager.detectLoopClosureID();              // pgo_node.cpp:344
    fprintf(out, "%zu %d %.9g\n", k, detectResult.first, (double)detectResult.second);
    if (detectResult.first >= 0) ++loops;
  }
  const double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  fprintf(out, "# keyframes %zu loops %d wall_s %.4f keyframes_per_s %.1f\n", keyframes.size(), loops, wall, keyframes.size() / wall);
  fclose(out);
  return 0;
}

int main(int argc, char** argv) {
  try {
    return run(argc, argv);
  } catch (const std::exception& e) {     // e.g. no CUDA device: the device library has no CPU fallback
    fprintf(stderr, "sc_sequence: %s\n", e.what());
    return 3;
  }
}
