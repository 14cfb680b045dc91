// Compile-and-link check of include/pclomp/ndt_omp.h in its PCL build (NDT_B200_WITH_PCL): the class derives from
// pcl::Registration, so it can be held and driven through the base pointer the way any PCL matcher is, and the static
// convertTransform overloads of the reference header (xchu_mapping/include/pclomp/ndt_omp.h:210-229) exist.
// PCL and Eigen are not in this image: the test compiles against the stand-in headers under oracle/ref_stubs (test
// infrastructure), which declare pcl::Registration with the members the reference header uses.
#include <pclomp/ndt_omp.h>

#include <cstdio>
#include <type_traits>

#ifndef NDT_B200_WITH_PCL
#error "this translation unit must see <pcl/point_cloud.h> and <Eigen/Core>"
#endif

typedef pcl::PointXYZI P;
typedef pclomp::NormalDistributionsTransform<P, P> NDT;
static_assert(std::is_base_of<pcl::Registration<P, P>, NDT>::value, "pclomp NDT derives from pcl::Registration (ndt_omp.h:70-71)");

int main(int argc, char** argv) {
  // static members need no device
  Eigen::Matrix<double, 6, 1> x;
  x << 1.0, 2.0, 3.0, 0.1, -0.2, 0.3;
  Eigen::Matrix4f T;
  NDT::convertTransform(x, T);
  Eigen::Affine3f A;
  NDT::convertTransform(x, A);
  for (int i = 0; i < 16; ++i)
    if (T.data()[i] != A.matrix().data()[i]) return 4;
  std::printf("T %.9g %.9g %.9g %.9g\n", T(0, 0), T(1, 0), T(2, 1), T(0, 3));
  if (argc > 1) return 0;   // "--static-only"
  try {
    boost::shared_ptr<pcl::Registration<P, P>> reg(new NDT());
    reg->setTransformationEpsilon(0.02);      // the base's non-virtual setters: picked up at align time
    reg->setMaximumIterations(12);
    NDT* ndt = static_cast<NDT*>(reg.get());
    ndt->setResolution(2.0f);
    ndt->setNeighborhoodSearchMethod(pclomp::DIRECT7);
    pcl::PointCloud<P>::Ptr tgt(new pcl::PointCloud<P>()), src(new pcl::PointCloud<P>());
    for (int i = 0; i < 4000; ++i) {          // three walls of a corner, so that every direction is observed
      P p;
      const float u = (float)(i % 64) * 0.25f, v = (float)((i / 64) % 21) * 0.2f;
      const int wall = i % 3;
      p.x = wall == 0 ? 0.f : u; p.y = wall == 1 ? 0.f : (wall == 0 ? u : v); p.z = wall == 2 ? 0.f : (wall == 0 ? v : u * 0.3f);
      p.intensity = 1.f;
      tgt->push_back(p);
      if (i % 4 == 0) { P q = p; q.x += 0.15f; q.y -= 0.1f; src->push_back(q); }
    }
    reg->setInputTarget(tgt);                 // virtual: reaches the device upload
    reg->setInputSource(src);
    pcl::PointCloud<P> out;
    reg->align(out, Eigen::Matrix4f::Identity());     // the base's align() -> computeTransformation()
    const Eigen::Matrix4f F = reg->getFinalTransformation();
    std::printf("base-pointer align: converged %d iterations %d t %.4f %.4f %.4f out %zu\n", (int)reg->hasConverged(),
                ndt->getFinalNumIteration(), F(0, 3), F(1, 3), F(2, 3), out.points.size());
    if (ndt->getFinalNumIteration() < 1 || ndt->getFinalNumIteration() > 12) return 5;
    if (out.points.size() != src->points.size()) return 6;
    const double s1 = ndt->calculateScore(out);       // ndt_omp.h:233: an already transformed cloud
    const double s2 = ndt->calculateScore(F);         // the same cloud as T * source
    std::printf("calculateScore %.12g %.12g\n", s1, s2);
    if (!(s1 == s1) || std::fabs(s1 - s2) > 1e-5 * std::fabs(s2)) return 7;
  } catch (const std::exception& e) {
    std::fprintf(stderr, "%s\n", e.what());
    return 3;
  }
  return 0;
}
