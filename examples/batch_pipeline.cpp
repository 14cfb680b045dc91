// examples/batch_pipeline.cpp -- the loop-closure / relocalisation batch (config C5) driven from C++ over the C ABI alone:
// host clouds in page-locked memory -> ndtb_batch_add_targets / _sources -> ndtb_batch_run, as a two-stage pipeline (one batch
// uploads while another computes), timed on the host clock around whole steps.  This is the end-to-end number with no
// interpreter in the loop; bench.py runs the same pipeline from Python and reports both.
//
//   batch_pipeline <job file> [--steps K] [--warmup W] [--inflight N] [--strict-math M] [--device D] [--out results.bin]
//
// The job file is written by xchu_slam_b200/jobfile.py (format in its docstring).  Prints one JSON line.
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../include/ndt_b200/ndt_c_api.h"

namespace {

struct Job {
  int n_targets = 0, n_sources = 0, n_items = 0, stride = 16;
  std::vector<size_t> target_n, source_n;
  std::vector<int> item_target, item_source;
  std::vector<float> guesses;
  void* pinned = nullptr;                 // every cloud, one allocation
  std::vector<const void*> target_p, source_p;
  size_t cloud_bytes = 0;
};

bool read_exact(FILE* f, void* p, size_t n) { return n == 0 || std::fread(p, 1, n, f) == n; }

int load(const char* path, Job& j) {
  FILE* f = std::fopen(path, "rb");
  if (!f) { std::fprintf(stderr, "cannot open %s\n", path); return 1; }
  char magic[8];
  int hdr[4];
  if (!read_exact(f, magic, 8) || std::memcmp(magic, "NDTBJOB1", 8) != 0 || !read_exact(f, hdr, sizeof(hdr))) {
    std::fprintf(stderr, "not a job file\n");
    return 1;
  }
  j.n_targets = hdr[0]; j.n_sources = hdr[1]; j.n_items = hdr[2]; j.stride = hdr[3];
  std::vector<long long> tn(j.n_targets), sn(j.n_sources);
  j.item_target.resize(j.n_items); j.item_source.resize(j.n_items); j.guesses.resize((size_t)j.n_items * 16);
  if (!read_exact(f, tn.data(), tn.size() * 8) || !read_exact(f, sn.data(), sn.size() * 8) ||
      !read_exact(f, j.item_target.data(), (size_t)j.n_items * 4) || !read_exact(f, j.item_source.data(), (size_t)j.n_items * 4) ||
      !read_exact(f, j.guesses.data(), j.guesses.size() * 4)) {
    std::fprintf(stderr, "short job file\n");
    return 1;
  }
  size_t total = 0;
  for (long long n : tn) { j.target_n.push_back((size_t)n); total += (size_t)n; }
  for (long long n : sn) { j.source_n.push_back((size_t)n); total += (size_t)n; }
  j.cloud_bytes = total * (size_t)j.stride;
  if (ndtb_host_alloc(j.cloud_bytes, &j.pinned) != 0) {
    std::fprintf(stderr, "page-locked allocation of %zu bytes failed (no CUDA device?): this program has no CPU fallback\n", j.cloud_bytes);
    return 3;
  }
  if (!read_exact(f, j.pinned, j.cloud_bytes)) { std::fprintf(stderr, "short job file\n"); return 1; }
  std::fclose(f);
  const char* p = static_cast<const char*>(j.pinned);
  for (size_t n : j.target_n) { j.target_p.push_back(p); p += n * (size_t)j.stride; }
  for (size_t n : j.source_n) { j.source_p.push_back(p); p += n * (size_t)j.stride; }
  return 0;
}

double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

}  // namespace

int main(int argc, char** argv) {
  if (argc < 2) {
    std::fprintf(stderr, "usage: batch_pipeline <job> [--steps K] [--warmup W] [--inflight N] [--strict-math M] [--device D] [--out file]\n");
    return 2;
  }
  int steps = 5, warmup = 3, inflight = 2, strict = -1, device = 0;
  std::string out_path;
  for (int i = 2; i < argc; ++i) {
    const std::string a = argv[i];
    auto val = [&]() { return i + 1 < argc ? argv[++i] : "0"; };
    if (a == "--steps") steps = std::atoi(val());
    else if (a == "--warmup") warmup = std::atoi(val());
    else if (a == "--inflight") inflight = std::atoi(val());
    else if (a == "--strict-math") strict = std::atoi(val());
    else if (a == "--device") device = std::atoi(val());
    else if (a == "--out") out_path = val();
  }
  if (inflight < 1) inflight = 1;
  std::vector<ndtb_batch*> batches(inflight, nullptr);
  for (auto& b : batches)
    if (ndtb_batch_create(device, &b) != 0) {
      std::fprintf(stderr, "batch_pipeline: %s (no CPU fallback)\n", ndtb_batch_last_error(nullptr));
      return 3;
    }
  Job job;
  if (int rc = load(argv[1], job)) return rc;
  ndtb_params p;
  ndtb_default_params(&p);
  p.resolution = 1.0f; p.step_size = 0.1; p.trans_eps = 0.01; p.max_iter = 30; p.search_method = 2; p.strict_math = strict;   // config C5
  for (auto b : batches) ndtb_batch_set_params(b, &p);

  std::mutex upload_lock, compute_lock;
  std::vector<std::vector<ndtb_result>> results(inflight, std::vector<ndtb_result>((size_t)job.n_items));
  std::atomic<int> failed{0};
  std::vector<ndtb_batch_stats> last_stats(inflight);
  auto one_step = [&](int k) {
    ndtb_batch* b = batches[k];
    int rc = 0;
    {
      std::lock_guard<std::mutex> g(upload_lock);       // stage 1: this batch's clouds cross the link
      rc |= ndtb_batch_clear(b);
      rc |= ndtb_batch_add_targets(b, job.n_targets, job.target_p.data(), job.target_n.data(), (size_t)job.stride, 0, nullptr);
      rc |= ndtb_batch_add_sources(b, job.n_sources, job.source_p.data(), job.source_n.data(), (size_t)job.stride, 0, nullptr);
      rc |= ndtb_batch_wait_uploads(b);
    }
    {
      std::lock_guard<std::mutex> g(compute_lock);      // stage 2: table builds + registrations, results back on the host
      rc |= ndtb_batch_run(b, job.n_items, job.item_target.data(), job.item_source.data(), job.guesses.data(), results[k].data());
      ndtb_batch_get_stats(b, &last_stats[k]);
    }
    if (rc) { std::fprintf(stderr, "batch_pipeline: %s\n", ndtb_batch_last_error(b)); failed = 1; }
  };
  auto run_steps = [&](int n) {   // n steps shared by `inflight` workers: worker k takes steps k, k + inflight, ...
    std::vector<std::thread> th;
    for (int k = 0; k < inflight; ++k)
      th.emplace_back([&, k]() { for (int s = k; s < n && !failed; s += inflight) one_step(k); });
    for (auto& t : th) t.join();
  };
  run_steps(warmup > inflight ? warmup : inflight);
  if (failed) return 1;
  const double t0 = now_ms();
  run_steps(steps);
  const double ms = (now_ms() - t0) / steps;
  if (failed) return 1;
  int converged = 0;
  for (const auto& r : results[0]) converged += r.converged;
  if (!out_path.empty()) {
    FILE* f = std::fopen(out_path.c_str(), "wb");
    for (const auto& r : results[0]) {
      double row[21];
      for (int i = 0; i < 16; ++i) row[i] = r.T[i];
      row[16] = r.iterations; row[17] = r.converged; row[18] = r.score; row[19] = r.trans_probability; row[20] = r.n_evals;
      std::fwrite(row, sizeof(double), 21, f);
    }
    std::fclose(f);
  }
  const ndtb_batch_stats& st = last_stats[0];
  std::printf("{\"caller\": \"examples/batch_pipeline.cpp\", \"aligns_per_s\": %.1f, \"ms_per_step\": %.3f, \"steps\": %d, \"warmup\": %d, "
              "\"batches_in_flight\": %d, \"items\": %d, \"stride_bytes\": %d, \"h2d_bytes_per_step\": %zu, \"d2h_bytes_per_step\": %zu, "
              "\"converged\": %d, \"replayed\": %lld, \"device_total_ms\": %.3f, \"device_build_ms\": %.3f}\n",
              1000.0 * job.n_items / ms, ms, steps, warmup, inflight, job.n_items, job.stride, job.cloud_bytes + (size_t)job.n_items * 64,
              (size_t)job.n_items * sizeof(ndtb_result), converged, st.replayed, st.total_ms, st.build_ms);
  for (auto b : batches) ndtb_batch_destroy(b);
  ndtb_host_free(job.pinned);
  return 0;
}
