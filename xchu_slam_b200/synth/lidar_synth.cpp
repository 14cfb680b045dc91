// lidar_synth.cpp -- deterministic synthetic LiDAR data for the NDT hot path.
//
// Test/bench INPUT tooling (not on the product path, not part of the oracle): a procedural
// urban block world ray-cast with an HDL-64E-shaped sensor, the filter_node-style range gate
// and voxel-centroid downsample, and a KITTI-00-shaped street-grid trajectory.  Spec:
// SURVEY.md section 8(d).  Plain arithmetic (ray/plane, ray/box, ray/cylinder, ray/sphere);
// no graphics API.  Built by xchu_slam_b200/build.py into _lidar_synth.so, bound via ctypes.

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

struct Rng {  // splitmix64
  uint64_t s;
  explicit Rng(uint64_t seed) : s(seed) {}
  uint64_t next() {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  double uni() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
  double uni(double a, double b) { return a + (b - a) * uni(); }
  double gauss() {
    const double u1 = std::max(uni(), 1e-300), u2 = uni();
    return std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586 * u2);
  }
};
static inline uint64_t mix(uint64_t a, uint64_t b) {
  Rng r(a * 0x9E3779B97F4A7C15ull ^ (b + 0x7F4A7C15ull));
  return r.next();
}

enum PrimType { BOX = 0, CYL = 1, SPH = 2 };
struct Prim {
  int type;
  double a[6];  // BOX: min xyz, max xyz.  CYL: cx, cy, r, z0, z1.  SPH: cx, cy, cz, r.
  double bb[4]; // xy AABB: minx, miny, maxx, maxy
};

struct World {
  uint64_t seed = 0;
  int half_blocks = 20;
  double block = 80.0, cell = 10.0;
  double ext = 0;  // world spans [-ext, ext]^2
  int ncell = 0;
  std::vector<Prim> prims;
  std::vector<std::vector<int>> grid;

  void add(const Prim& p) { prims.push_back(p); }
  void add_box(double x0, double y0, double z0, double x1, double y1, double z1) {
    Prim p; p.type = BOX;
    p.a[0] = x0; p.a[1] = y0; p.a[2] = z0; p.a[3] = x1; p.a[4] = y1; p.a[5] = z1;
    p.bb[0] = x0; p.bb[1] = y0; p.bb[2] = x1; p.bb[3] = y1;
    add(p);
  }
  void add_cyl(double cx, double cy, double r, double z0, double z1) {
    Prim p; p.type = CYL;
    p.a[0] = cx; p.a[1] = cy; p.a[2] = r; p.a[3] = z0; p.a[4] = z1; p.a[5] = 0;
    p.bb[0] = cx - r; p.bb[1] = cy - r; p.bb[2] = cx + r; p.bb[3] = cy + r;
    add(p);
  }
  void add_sph(double cx, double cy, double cz, double r) {
    Prim p; p.type = SPH;
    p.a[0] = cx; p.a[1] = cy; p.a[2] = cz; p.a[3] = r; p.a[4] = p.a[5] = 0;
    p.bb[0] = cx - r; p.bb[1] = cy - r; p.bb[2] = cx + r; p.bb[3] = cy + r;
    add(p);
  }

  void build() {
    ext = half_blocks * block;
    for (int bi = -half_blocks; bi < half_blocks; ++bi)
      for (int bj = -half_blocks; bj < half_blocks; ++bj) {
        Rng r(mix(seed, (uint64_t)((bi + 4096) * 8192 + (bj + 4096))));
        const double x0 = bi * block, y0 = bj * block;
        // buildings: 2x2 lots inside the block, set back 8-15 m from the street axes
        for (int li = 0; li < 2; ++li)
          for (int lj = 0; lj < 2; ++lj) {
            if (r.uni() < 0.12) continue;  // empty lot
            const double sb = r.uni(8.0, 15.0);
            const double lot_x0 = x0 + (li == 0 ? sb : block * 0.5 + 1.0), lot_x1 = x0 + (li == 0 ? block * 0.5 - 1.0 : block - sb);
            const double lot_y0 = y0 + (lj == 0 ? sb : block * 0.5 + 1.0), lot_y1 = y0 + (lj == 0 ? block * 0.5 - 1.0 : block - sb);
            const double fx = std::min(r.uni(10.0, 40.0), lot_x1 - lot_x0), fy = std::min(r.uni(10.0, 40.0), lot_y1 - lot_y0);
            const double h = r.uni(6.0, 30.0);
            // hug the street-side corner of the lot
            const double bx0 = (li == 0) ? lot_x0 : lot_x1 - fx, by0 = (lj == 0) ? lot_y0 : lot_y1 - fy;
            add_box(bx0, by0, 0.0, bx0 + fx, by0 + fy, h);
          }
        // street furniture along the two street axes owned by this block (x = x0 and y = y0)
        for (int axis = 0; axis < 2; ++axis) {
          for (double s = 12.5; s < block; s += 25.0) {  // poles every 25 m, both kerbs
            for (int side = -1; side <= 1; side += 2) {
              const double off = side * 6.5;
              const double px = axis == 0 ? x0 + off : x0 + s, py = axis == 0 ? y0 + s : y0 + off;
              add_cyl(px, py, 0.15, 0.0, 6.0);
            }
          }
          for (double s = 10.0; s + 6.0 < block - 8.0; s += 6.0) {  // parked cars, 30 % of kerb slots
            for (int side = -1; side <= 1; side += 2) {
              if (r.uni() >= 0.30) continue;
              const double off = side * 4.6;
              if (axis == 0) add_box(x0 + off - 0.9, y0 + s, 0.0, x0 + off + 0.9, y0 + s + 4.5, 1.5);
              else add_box(x0 + s, y0 + off - 0.9, 0.0, x0 + s + 4.5, y0 + off + 0.9, 1.5);
            }
          }
          for (double s = 20.0; s < block - 10.0; s += 20.0) {  // trees
            for (int side = -1; side <= 1; side += 2) {
              if (r.uni() < 0.4) continue;
              const double off = side * 7.6;
              const double tx = axis == 0 ? x0 + off : x0 + s + r.uni(-2, 2), ty = axis == 0 ? y0 + s + r.uni(-2, 2) : y0 + off;
              const double cr = r.uni(1.4, 2.4);
              add_cyl(tx, ty, 0.2, 0.0, 3.2);
              add_sph(tx, ty, 3.0 + cr, cr);
            }
          }
        }
      }
    ncell = (int)std::ceil(2 * ext / cell);
    grid.assign((size_t)ncell * ncell, {});
    for (size_t i = 0; i < prims.size(); ++i) {
      const Prim& p = prims[i];
      const int cx0 = std::max(0, (int)std::floor((p.bb[0] + ext) / cell)), cx1 = std::min(ncell - 1, (int)std::floor((p.bb[2] + ext) / cell));
      const int cy0 = std::max(0, (int)std::floor((p.bb[1] + ext) / cell)), cy1 = std::min(ncell - 1, (int)std::floor((p.bb[3] + ext) / cell));
      for (int cy = cy0; cy <= cy1; ++cy)
        for (int cx = cx0; cx <= cx1; ++cx) grid[(size_t)cy * ncell + cx].push_back((int)i);
    }
  }

  static double hit_prim(const Prim& p, const double o[3], const double d[3], double tmax) {
    if (p.type == BOX) {
      double t0 = 1e-6, t1 = tmax;
      for (int a = 0; a < 3; ++a) {
        if (std::fabs(d[a]) < 1e-12) {
          if (o[a] < p.a[a] || o[a] > p.a[a + 3]) return -1;
        } else {
          double ta = (p.a[a] - o[a]) / d[a], tb = (p.a[a + 3] - o[a]) / d[a];
          if (ta > tb) std::swap(ta, tb);
          t0 = std::max(t0, ta);
          t1 = std::min(t1, tb);
          if (t0 > t1) return -1;
        }
      }
      return t0;
    } else if (p.type == CYL) {
      const double ox = o[0] - p.a[0], oy = o[1] - p.a[1];
      const double A = d[0] * d[0] + d[1] * d[1];
      if (A < 1e-14) return -1;
      const double B = ox * d[0] + oy * d[1], C = ox * ox + oy * oy - p.a[2] * p.a[2];
      const double disc = B * B - A * C;
      if (disc < 0) return -1;
      const double t = (-B - std::sqrt(disc)) / A;
      if (t <= 1e-6 || t >= tmax) return -1;
      const double z = o[2] + t * d[2];
      if (z < p.a[3] || z > p.a[4]) return -1;
      return t;
    } else {
      const double ox = o[0] - p.a[0], oy = o[1] - p.a[1], oz = o[2] - p.a[2];
      const double B = ox * d[0] + oy * d[1] + oz * d[2], C = ox * ox + oy * oy + oz * oz - p.a[3] * p.a[3];
      const double disc = B * B - C;
      if (disc < 0) return -1;
      const double t = -B - std::sqrt(disc);
      if (t <= 1e-6 || t >= tmax) return -1;
      return t;
    }
  }

  // nearest hit along o + t d (|d| = 1), t in (0, tmax); returns -1 if none
  double cast(const double o[3], const double d[3], double tmax) const {
    double best = tmax;
    bool hit = false;
    if (d[2] < -1e-9) {  // ground plane z = 0
      const double t = -o[2] / d[2];
      if (t > 1e-6 && t < best) { best = t; hit = true; }
    }
    // 2D DDA over the primitive grid
    const double gx = (o[0] + ext) / cell, gy = (o[1] + ext) / cell;
    int cx = (int)std::floor(gx), cy = (int)std::floor(gy);
    const int sx = d[0] > 0 ? 1 : -1, sy = d[1] > 0 ? 1 : -1;
    const double inv_dx = std::fabs(d[0]) > 1e-12 ? 1.0 / std::fabs(d[0]) : 1e30;
    const double inv_dy = std::fabs(d[1]) > 1e-12 ? 1.0 / std::fabs(d[1]) : 1e30;
    double tx = ((d[0] > 0 ? (cx + 1 - gx) : (gx - cx)) * cell) * inv_dx;
    double ty = ((d[1] > 0 ? (cy + 1 - gy) : (gy - cy)) * cell) * inv_dy;
    const double dtx = cell * inv_dx, dty = cell * inv_dy;
    double t_enter = 0;
    while (t_enter < best) {
      if (cx < 0 || cy < 0 || cx >= ncell || cy >= ncell) break;
      for (int pi : grid[(size_t)cy * ncell + cx]) {
        const double t = hit_prim(prims[pi], o, d, best);
        if (t > 0 && t < best) { best = t; hit = true; }
      }
      if (tx < ty) { t_enter = tx; tx += dtx; cx += sx; }
      else { t_enter = ty; ty += dty; cy += sy; }
    }
    return hit ? best : -1;
  }
};

struct Key3 { int64_t k; uint32_t i; };

}  // namespace

extern "C" {

void* synth_world_create(uint64_t seed, int half_blocks) {
  World* w = new World();
  w->seed = seed;
  w->half_blocks = half_blocks;
  w->build();
  return w;
}
void synth_world_destroy(void* w) { delete (World*)w; }
int synth_world_num_prims(void* w) { return (int)((World*)w)->prims.size(); }

// KITTI-00-shaped drive on the street grid: straight segments along street axes with
// quarter-circle turns (radius 8 m) at seeded intersections, `step` metres per scan.
// out: n x 3 (x, y, yaw).
void synth_trajectory(void* wv, int n, double step, uint64_t seed, double* out) {
  World* w = (World*)wv;
  Rng r(mix(w->seed, seed ^ 0x7261ull));
  double x = 0, y = 0;
  int heading = 0;  // 0:+x 1:+y 2:-x 3:-y
  const double R = 8.0;
  int i = 0;
  const double lim = w->ext - 3 * w->block;
  auto emit = [&](double px, double py, double yaw) {
    if (i < n) { out[i * 3] = px; out[i * 3 + 1] = py; out[i * 3 + 2] = yaw; ++i; }
  };
  double carry = 0;  // distance already travelled into the next sample
  while (i < n) {
    // drive to R before the next intersection
    const double hx = heading == 0 ? 1 : heading == 2 ? -1 : 0, hy = heading == 1 ? 1 : heading == 3 ? -1 : 0;
    const double pos = hx != 0 ? x * hx : y * hy;  // coordinate along travel direction (increasing)
    double next_axis = (std::floor(pos / w->block + 1e-9) + 1) * w->block;
    double seg = next_axis - R - pos;
    if (seg < 0) { next_axis += w->block; seg += w->block; }
    double s = carry;
    const double yaw0 = heading * 1.5707963267948966;
    for (; s <= seg && i < n; s += step) emit(x + hx * s, y + hy * s, yaw0);
    carry = s - seg;
    x += hx * seg; y += hy * seg;
    // choose: straight / left / right (stay inside the world)
    int turn = 0;
    const double u = r.uni();
    if (u < 0.30) turn = 1; else if (u < 0.55) turn = -1;
    const int nh = (heading + turn + 4) % 4;
    const double nhx = nh == 0 ? 1 : nh == 2 ? -1 : 0, nhy = nh == 1 ? 1 : nh == 3 ? -1 : 0;
    if (std::fabs(x + nhx * 3 * w->block) > lim || std::fabs(y + nhy * 3 * w->block) > lim) {
      // head back toward the origin instead
      turn = ((-y * hx + x * hy) > 0) ? -1 : 1;
      if (std::fabs(x + hx * 3 * w->block) <= lim && std::fabs(y + hy * 3 * w->block) <= lim && r.uni() < 0.5) turn = 0;
    }
    if (turn == 0) {
      // go straight through the intersection (2R)
      double s2 = carry;
      for (; s2 <= 2 * R && i < n; s2 += step) emit(x + hx * s2, y + hy * s2, yaw0);
      carry = s2 - 2 * R;
      x += hx * 2 * R; y += hy * 2 * R;
    } else {
      // quarter circle of radius R, centre to the left (turn=+1) or right (turn=-1)
      const double cxx = x - hy * R * turn, cyy = y + hx * R * turn;
      const double arc = R * 1.5707963267948966;
      double s2 = carry;
      for (; s2 <= arc && i < n; s2 += step) {
        const double a = s2 / R * turn;
        const double ang0 = std::atan2(y - cyy, x - cxx);
        emit(cxx + R * std::cos(ang0 + a), cyy + R * std::sin(ang0 + a), yaw0 + a);
      }
      carry = s2 - arc;
      const double ang0 = std::atan2(y - cyy, x - cxx);
      x = cxx + R * std::cos(ang0 + turn * 1.5707963267948966);
      y = cyy + R * std::sin(ang0 + turn * 1.5707963267948966);
      heading = (heading + turn + 4) % 4;
      // snap to the street axis to avoid drift
      if (heading == 0 || heading == 2) y = std::round(y / w->block) * w->block; else x = std::round(x / w->block) * w->block;
    }
  }
}

// One HDL-64E-shaped revolution from a vehicle pose (x, y, z, roll, pitch, yaw; sensor 1.73 m above
// the vehicle origin).  Output points are in the SENSOR frame, float xyz + intensity(=0), stride 4.
// Returns the number of returns written (<= max_pts).
int synth_scan(void* wv, const double* pose6, uint64_t scan_seed, float* out, int max_pts) {
  const World* w = (const World*)wv;
  const int n_beams = 64, n_az = 2083;
  const double el_top = 2.0 * M_PI / 180.0, el_bot = -24.8 * M_PI / 180.0;
  const double max_range = 120.0, sigma = 0.02, sensor_h = 1.73;
  const double cr = std::cos(pose6[3]), sr = std::sin(pose6[3]), cp = std::cos(pose6[4]), sp = std::sin(pose6[4]);
  const double cy = std::cos(pose6[5]), sy = std::sin(pose6[5]);
  // R = Rz(yaw) Ry(pitch) Rx(roll)
  const double R[3][3] = {{cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr},
                          {sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr},
                          {-sp, cp * sr, cp * cr}};
  const double o[3] = {pose6[0] + R[0][2] * sensor_h, pose6[1] + R[1][2] * sensor_h, pose6[2] + R[2][2] * sensor_h};
  std::vector<float> buf((size_t)n_beams * n_az * 4);
  std::vector<uint8_t> ok((size_t)n_beams * n_az, 0);
#pragma omp parallel for schedule(dynamic, 16)
  for (int az = 0; az < n_az; ++az) {
    const double phi = 2 * M_PI * az / n_az;
    for (int b = 0; b < n_beams; ++b) {
      const double el = el_top + (el_bot - el_top) * b / (n_beams - 1);
      const double ds[3] = {std::cos(el) * std::cos(phi), std::cos(el) * std::sin(phi), std::sin(el)};
      double d[3];
      for (int a = 0; a < 3; ++a) d[a] = R[a][0] * ds[0] + R[a][1] * ds[1] + R[a][2] * ds[2];
      const double t = w->cast(o, d, max_range);
      if (t <= 0) continue;
      // ground roughness (+-0.03 m) and range noise N(0, 0.02)
      Rng r(mix(scan_seed, (uint64_t)az * 64 + b));
      double tt = t + sigma * r.gauss();
      const double zhit = o[2] + t * d[2];
      if (zhit < 0.02 && d[2] < -1e-3) tt += r.uni(-0.03, 0.03) / -d[2] * 0.5;
      const size_t k = (size_t)az * n_beams + b;
      buf[k * 4] = (float)(ds[0] * tt); buf[k * 4 + 1] = (float)(ds[1] * tt); buf[k * 4 + 2] = (float)(ds[2] * tt); buf[k * 4 + 3] = 0.f;
      ok[k] = 1;
    }
  }
  int cnt = 0;
  for (size_t k = 0; k < ok.size() && cnt < max_pts; ++k)
    if (ok[k]) { std::memcpy(out + (size_t)cnt * 4, &buf[k * 4], 16); ++cnt; }
  return cnt;
}

// Range gate lo < r_xy < hi (filter_node) ; in place compaction, returns new count.
int synth_range_gate(float* pts, int n, float lo, float hi) {
  int c = 0;
  for (int i = 0; i < n; ++i) {
    const float r = std::sqrt(pts[i * 4] * pts[i * 4] + pts[i * 4 + 1] * pts[i * 4 + 1]);
    if (r > lo && r < hi) { if (c != i) std::memcpy(pts + (size_t)c * 4, pts + (size_t)i * 4, 16); ++c; }
  }
  return c;
}

// pcl::VoxelGrid-style centroid downsample (one float centroid per occupied voxel, voxel order).
int synth_voxel_downsample(const float* pts, int n, float leaf, float* out, int max_out) {
  if (n == 0) return 0;
  const float inv = 1.0f / leaf;
  std::vector<Key3> keys(n);
  for (int i = 0; i < n; ++i) {
    const int64_t a = (int64_t)std::floor(pts[i * 4] * inv), b = (int64_t)std::floor(pts[i * 4 + 1] * inv), c = (int64_t)std::floor(pts[i * 4 + 2] * inv);
    keys[i].k = ((c + (1 << 20)) << 42) | ((b + (1 << 20)) << 21) | (a + (1 << 20));
    keys[i].i = (uint32_t)i;
  }
  std::stable_sort(keys.begin(), keys.end(), [](const Key3& x, const Key3& y) { return x.k < y.k; });
  int cnt = 0;
  for (int s = 0; s < n && cnt < max_out;) {
    int e = s;
    float sx = 0, sy = 0, sz = 0;
    while (e < n && keys[e].k == keys[s].k) { const float* p = pts + (size_t)keys[e].i * 4; sx += p[0]; sy += p[1]; sz += p[2]; ++e; }
    const float m = (float)(e - s);
    out[cnt * 4] = sx / m; out[cnt * 4 + 1] = sy / m; out[cnt * 4 + 2] = sz / m; out[cnt * 4 + 3] = 0.f;
    ++cnt;
    s = e;
  }
  return cnt;
}

// Rigid transform (x, y, z, roll, pitch, yaw; R = Rz Ry Rx) of stride-4 points, float math.
void synth_transform(const float* in, int n, const double* pose6, float* out) {
  const double cr = std::cos(pose6[3]), sr = std::sin(pose6[3]), cp = std::cos(pose6[4]), sp = std::sin(pose6[4]);
  const double cy = std::cos(pose6[5]), sy = std::sin(pose6[5]);
  const float R[3][3] = {{(float)(cy * cp), (float)(cy * sp * sr - sy * cr), (float)(cy * sp * cr + sy * sr)},
                         {(float)(sy * cp), (float)(sy * sp * sr + cy * cr), (float)(sy * sp * cr - cy * sr)},
                         {(float)(-sp), (float)(cp * sr), (float)(cp * cr)}};
  const float t[3] = {(float)pose6[0], (float)pose6[1], (float)pose6[2]};
  for (int i = 0; i < n; ++i) {
    const float x = in[i * 4], y = in[i * 4 + 1], z = in[i * 4 + 2];
    out[i * 4] = R[0][0] * x + R[0][1] * y + R[0][2] * z + t[0];
    out[i * 4 + 1] = R[1][0] * x + R[1][1] * y + R[1][2] * z + t[1];
    out[i * 4 + 2] = R[2][0] * x + R[2][1] * y + R[2][2] * z + t[2];
    out[i * 4 + 3] = in[i * 4 + 3];
  }
}

}  // extern "C"
