// ndt_core.cu -- the per-iteration hot path: fused derivative kernel (K3), FP64 Hessian / score
// variants (K4), device-resident Newton + More-Thuente state machine (K7), neighbour-set and
// transform helpers.  Compiled with -fmad=false: every float expression below is evaluated in the
// written order with separate multiplies and adds, which is what makes the voxel assignment of
// transformed points and each (point, voxel) term bit-identical to the reference arithmetic
// (reference built for baseline x86-64: no FMA contraction).
//
// Reference: xchu_mapping/include/pclomp/ndt_omp_impl.hpp (NDT_I), ndt_omp.h (NDT_H),
// voxel_grid_covariance_omp_impl.hpp (VGC_I), voxel_grid_covariance_omp.h (VGC_H).
#include <atomic>
#include <cfloat>
#include <cmath>

#include "ndt_internal.h"
#include "ndt_device.cuh"

namespace ndtb {

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return e_; } while (0)

// Block shape of the term-by-term kernel: three 128-thread blocks per SM leave 168 registers per thread, which hold 25 of the 43
// FP64 accumulators beside the pair arithmetic without spills (two 256-thread blocks at 128 registers held 7): 58 % less
// shared-memory accumulator traffic for 12 instead of 16 resident warps, -3.4 % kernel time (profiles/r02/r02_c12_strict_variants.txt)
#ifndef NDTB_THREADS
#define NDTB_THREADS 128
#endif
#ifndef NDTB_MINBLOCKS
#define NDTB_MINBLOCKS 3
#endif
#ifndef NDTB_PD_SMEM      // 1: per-point derivative vectors parked in shared memory, 0: in registers
#define NDTB_PD_SMEM 0
#endif
#ifndef NDTB_PREFETCH     // 1: fetch the next voxel record behind the current pair's math
#define NDTB_PREFETCH 1
#endif
#ifndef NDTB_NOUNROLL          // 1: the neighbour loop of the strict kernel is not unrolled (code size 4552 -> 1816 instructions)
#define NDTB_NOUNROLL 1         //    measured -6.7 % kernel time: 17 % of the stall samples were instruction-cache misses (profiles/r02)
#endif
#ifndef NDTB_STRICT_ACC_SMEM   // 1: the 43 FP64 accumulators of a thread live in shared memory ([k][thread], conflict-free), which
#define NDTB_STRICT_ACC_SMEM 1  //    takes 86 registers off the kernel -> two resident blocks per SM instead of one (with the
                                //    loop not unrolled: -10.7 % kernel time; with it unrolled +23 %: profiles/r02/r02_c5_strict_variants.txt)
#endif
#ifndef NDTB_STRICT_REG_ACCS   // with NDTB_STRICT_ACC_SMEM: the first N of the 43 accumulators (score, gradient, then Hessian rows) stay
#define NDTB_STRICT_REG_ACCS 25 //    in registers, the rest live in shared memory: N / 43 less shared-memory traffic per pair
#endif
constexpr int kThreads = NDTB_THREADS;
#if NDTB_STRICT_ACC_SMEM
constexpr int kAccStride = kThreads + 1;
constexpr int kRegAccs = NDTB_STRICT_REG_ACCS;
#define SACC(k) acc[(k) * kAccStride]
#define ACC(k) (*((k) < kRegAccs ? &gacc[(k) < kRegAccs ? (k) : 0] : &acc[(k) * kAccStride]))
#else
#define SACC(k) acc[(k)]
#define ACC(k) acc[(k)]
#endif

struct EvalShared {
  float T[12];
  float jang[24];
  float hang[45];
  float d2f;
  double d1, d2, d3;
  int mode, method;
  VoxelTable tab;
};

// ---------------------------------------------------------------------------------------
// K3: computeDerivatives (NDT_I:175-283) = transform + neighbour lookup + computePointDerivatives
// (NDT_I:401-445) + updateDerivatives (NDT_I:491-548), float per-term math, double accumulation,
// warp-shuffle + shared-memory block reduction to one partial per block (fixed order, no atomics).
// ---------------------------------------------------------------------------------------
// Record fetch: the 64 B record as four 128-bit loads (issued together, consumed later)
struct RecRegs { double2 a, b; float4 c, d; };
__device__ __forceinline__ RecRegs load_rec(const VoxelRec* __restrict__ recp) {
  RecRegs r;
  r.a = __ldg(reinterpret_cast<const double2*>(recp));
  r.b = __ldg(reinterpret_cast<const double2*>(recp) + 1);
  r.c = __ldg(reinterpret_cast<const float4*>(recp) + 2);
  r.d = __ldg(reinterpret_cast<const float4*>(recp) + 3);
  return r;
}

template <bool CH>
__device__ __forceinline__ void pair_update_r(double* acc, double* gacc, const RecRegs& rr, float xtx, float xty, float xtz,
                                              const float* __restrict__ pd /* smem column, stride kThreads */, float d2f, double d1) {
#if NDTB_PD_SMEM
#define PGV(r) pd[(r) * kThreads]
#define PHV(r) pd[(8 + (r)) * kThreads]
#else
#define PGV(r) pd[(r)]
#define PHV(r) pd[8 + (r)]
#endif
  const double2 m01 = rr.a, m2c = rr.b;
  const float4 c1 = rr.c, c2 = rr.d;
  const float C00 = __int_as_float(__double2loint(m2c.y)), C01 = __int_as_float(__double2hiint(m2c.y));
  const float C02 = c1.x, C10 = c1.y, C11 = c1.z, C12 = c1.w, C20 = c2.x, C21 = c2.y, C22 = c2.z;
  // x_trans = f64(x_trans_pt) - mean, narrowed to float (NDT_I:257-260, 500)
  const float x0 = (float)((double)xtx - m01.x), x1 = (float)((double)xty - m01.y), x2 = (float)((double)xtz - m2c.x);
  const float xC0 = (x0 * C00 + x1 * C10) + x2 * C20;
  const float xC1 = (x0 * C01 + x1 * C11) + x2 * C21;
  const float xC2 = (x0 * C02 + x1 * C12) + x2 * C22;
  const float q = (x0 * xC0 + x1 * xC1) + x2 * xC2;
  float e = (float)exp((double)((-d2f * q) * 0.5f));            // NDT_I:507
  const float score_inc = (float)(-d1 * (double)e);             // NDT_I:509
  e = d2f * e;                                                  // NDT_I:511
  if (e > 1.f || e < 0.f || e != e) return;                     // NDT_I:514-515 (score not counted either)
  e = (float)((double)e * d1);                                  // NDT_I:518
  ACC(0) += (double)score_inc;
  const float g13 = PGV(0), g23 = PGV(1), g04 = PGV(2), g14 = PGV(3), g24 = PGV(4), g05 = PGV(5), g15 = PGV(6), g25 = PGV(7);
  // CJ = c_inv4 * point_gradient4 (NDT_I:520); columns 0..2 are the columns of C
  float CJ[3][6];
  CJ[0][0] = C00; CJ[0][1] = C01; CJ[0][2] = C02;
  CJ[1][0] = C10; CJ[1][1] = C11; CJ[1][2] = C12;
  CJ[2][0] = C20; CJ[2][1] = C21; CJ[2][2] = C22;
  CJ[0][3] = C01 * g13 + C02 * g23; CJ[1][3] = C11 * g13 + C12 * g23; CJ[2][3] = C21 * g13 + C22 * g23;
  CJ[0][4] = (C00 * g04 + C01 * g14) + C02 * g24; CJ[1][4] = (C10 * g04 + C11 * g14) + C12 * g24; CJ[2][4] = (C20 * g04 + C21 * g14) + C22 * g24;
  CJ[0][5] = (C00 * g05 + C01 * g15) + C02 * g25; CJ[1][5] = (C10 * g05 + C11 * g15) + C12 * g25; CJ[2][5] = (C20 * g05 + C21 * g15) + C22 * g25;
  float xCJ[6];
  xCJ[0] = xC0; xCJ[1] = xC1; xCJ[2] = xC2;
#pragma unroll
  for (int c = 3; c < 6; ++c) xCJ[c] = (x0 * CJ[0][c] + x1 * CJ[1][c]) + x2 * CJ[2][c];   // NDT_I:521
#pragma unroll
  for (int c = 0; c < 6; ++c) ACC(1 + c) += (double)(e * xCJ[c]);                          // NDT_I:523
  if (!CH) return;
  // xCH(i,j) = (x_trans4 * c_inv4) * point_hessian_.block<4,1>(4i, j)  (NDT_I:534-535); non-zero for i,j >= 3
  const float ha = xC1 * PHV(0) + xC2 * PHV(1);
  const float hb = xC1 * PHV(2) + xC2 * PHV(3);
  const float hc = xC1 * PHV(4) + xC2 * PHV(5);
  const float hd = (xC0 * PHV(6) + xC1 * PHV(7)) + xC2 * PHV(8);
  const float he = (xC0 * PHV(9) + xC1 * PHV(10)) + xC2 * PHV(11);
  const float hf = (xC0 * PHV(12) + xC1 * PHV(13)) + xC2 * PHV(14);
  const float xCH[3][3] = {{ha, hb, hc}, {hb, hd, he}, {hc, he, hf}};
  const float nd2 = -d2f;
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    const float nda = nd2 * xCJ[i];
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      // JtCJ(j,i) = point_gradient4.col(j) . CJ.col(i)   (NDT_I:527-528, 542)
      float jt;
      if (j < 3) jt = CJ[j][i];
      else if (j == 3) jt = g13 * CJ[1][i] + g23 * CJ[2][i];
      else if (j == 4) jt = (g04 * CJ[0][i] + g14 * CJ[1][i]) + g24 * CJ[2][i];
      else jt = (g05 * CJ[0][i] + g15 * CJ[1][i]) + g25 * CJ[2][i];
      float v = nda * xCJ[j];
      if (i >= 3 && j >= 3) v = v + xCH[i - 3][j - 3];
      v = v + jt;
      ACC(7 + i * 6 + j) += (double)(e * v);                                               // NDT_I:539-542
    }
  }
}
#undef PGV
#undef PHV

// Block reduction of the 44 per-thread accumulators through shared memory: every warp parks its
// accumulators column-wise, 44 lanes each add one column in lane order, then 44 threads add the
// warps in warp order.  Fixed order => bit-reproducible regardless of batch layout.
constexpr int kRedPad = 33;
__device__ __forceinline__ void block_reduce_store(const double* acc, int npairs, double* __restrict__ out) {
  extern __shared__ double red_sm[];                 // [warps][44][33] + [warps][44]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
  double* mine = red_sm + (size_t)warp * kPartialStride * kRedPad;
#pragma unroll
  for (int k = 0; k < kAccN; ++k) mine[k * kRedPad + lane] = acc[k];
  mine[kAccN * kRedPad + lane] = (double)npairs;
  __syncwarp();
  double* wsum = red_sm + (size_t)nwarps * kPartialStride * kRedPad;
  for (int k = lane; k < kPartialStride; k += 32) {
    double v = 0;
#pragma unroll 8
    for (int l = 0; l < 32; ++l) v += mine[k * kRedPad + l];
    wsum[warp * kPartialStride + k] = v;
  }
  __syncthreads();
  if (tid < kPartialStride) {
    double v = 0;
    for (int w = 0; w < nwarps; ++w) v += wsum[w * kPartialStride + tid];
    out[tid] = v;
  }
}
#if NDTB_STRICT_ACC_SMEM
// The same reduction when the accumulators already sit in shared memory as [k][thread]: lane k of every warp adds the warp's 32
// columns of row k in lane order, then 44 threads add the warps in warp order -- the additions of block_reduce_store, one for one.
__device__ __forceinline__ void block_reduce_store_resident(int npairs, double* __restrict__ out) {
  extern __shared__ double red_sm[];                 // [44][kAccStride] + [warps][44]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
  red_sm[kAccN * kAccStride + tid] = (double)npairs;
  __syncthreads();
  double* wsum = red_sm + (size_t)kPartialStride * kAccStride;
  for (int k = lane; k < kPartialStride; k += 32) {
    const double* row = red_sm + (size_t)k * kAccStride + warp * 32;
    double v = 0;
#pragma unroll 8
    for (int l = 0; l < 32; ++l) v += row[l];
    wsum[warp * kPartialStride + k] = v;
  }
  __syncthreads();
  if (tid < kPartialStride) {
    double v = 0;
    for (int w = 0; w < nwarps; ++w) v += wsum[w * kPartialStride + tid];
    out[tid] = v;
  }
}
static size_t reduce_smem_bytes(int threads) {
  const int nw = threads / 32;
  return ((size_t)kPartialStride * (threads + 1) + (size_t)nw * kPartialStride) * sizeof(double);
}
#else
static size_t reduce_smem_bytes(int threads) {
  const int nw = threads / 32;
  return ((size_t)nw * kPartialStride * kRedPad + (size_t)nw * kPartialStride) * sizeof(double);
}
#endif

template <int METHOD, bool CH>
__global__ void __launch_bounds__(kThreads, NDTB_MINBLOCKS) k_derivatives(const AlignState* __restrict__ states, const VoxelTable* __restrict__ tables,
                                                          const float4* __restrict__ src_pool, const int2* __restrict__ blockmap,
                                                          int chunk, double* __restrict__ partials) {
  const int2 bm = blockmap[blockIdx.x];
  const AlignState& st = states[bm.x];
  const int mode = st.req.mode;
  if (mode != (CH ? 1 : 2) || st.done || st.req.search_method != METHOD) return;
  __shared__ EvalShared sh;
#if NDTB_PD_SMEM
  __shared__ float pd_sm[23 * kThreads];
#endif
  const int tid = threadIdx.x;
  for (int k = tid; k < 12; k += kThreads) sh.T[k] = st.req.T[k];
  for (int k = tid; k < 24; k += kThreads) sh.jang[k] = (&st.req.jang[0][0])[k];
  for (int k = tid; k < 45; k += kThreads) sh.hang[k] = (&st.req.hang[0][0])[k];
  if (tid == kThreads - 1) {
    sh.d1 = st.req.d1; sh.d2 = st.req.d2; sh.d2f = (float)st.req.d2;   // float gauss_d2 = gauss_d2_ (NDT_I:504)
    sh.mode = mode; sh.method = METHOD;
    sh.tab = tables[st.target];
  }
  __syncthreads();
#if NDTB_STRICT_ACC_SMEM
  extern __shared__ double red_sm[];
  double* acc = red_sm + tid;
#else
  double acc[kAccN];
#endif
#pragma unroll
  for (int k = 0; k < kAccN; ++k) SACC(k) = 0.0;
#if NDTB_STRICT_ACC_SMEM
  double gacc[kRegAccs > 0 ? kRegAccs : 1];
#pragma unroll
  for (int k = 0; k < (kRegAccs > 0 ? kRegAccs : 1); ++k) gacc[k] = 0.0;
#else
  double* gacc = nullptr;
  (void)gacc;
#endif
  int npairs = 0;
  const float4* __restrict__ src = st.src;
  const int end = min(bm.y + chunk, st.n_src);
  const VoxelRec* __restrict__ recs = sh.tab.rec;
  const float d2f = sh.d2f;
  const double d1 = sh.d1;
  for (int i = bm.y + tid; i < end; i += kThreads) {
    const float4 p = __ldg(src + i);
    float xt, yt, zt;
    xform(sh.T, p.x, p.y, p.z, xt, yt, zt);
    int nb[METHOD == 2 ? 7 : 27];
    int cnt = 0;
    cnt = find_neighbors(sh.tab, METHOD, xt, yt, zt, nb);
    if (cnt == 0) continue;
    // computePointDerivatives (NDT_I:401-445): x_j_ang = j_ang * x4, x_h_ang = h_ang * x4.  The 23
    // per-point values are parked in shared memory (one column per thread) instead of registers.
#if NDTB_PD_SMEM
    float* pdv = pd_sm + tid;
#define PDI(r) pdv[(r) * kThreads]
#else
    float pdv[23];
#define PDI(r) pdv[(r)]
#endif
#pragma unroll
    for (int r = 0; r < 8; ++r) PDI(r) = (sh.jang[r * 3] * p.x + sh.jang[r * 3 + 1] * p.y) + sh.jang[r * 3 + 2] * p.z;
    if (CH) {
#pragma unroll
      for (int r = 0; r < 15; ++r) PDI(8 + r) = (sh.hang[r * 3] * p.x + sh.hang[r * 3 + 1] * p.y) + sh.hang[r * 3 + 2] * p.z;
    }
#undef PDI
#if NDTB_PREFETCH
    RecRegs cur = load_rec(recs + nb[0]);
#if NDTB_NOUNROLL
#pragma unroll 1
#endif
    for (int k = 0; k < cnt; ++k) {
      const RecRegs use = cur;
      if (k + 1 < cnt) cur = load_rec(recs + nb[k + 1]);
      pair_update_r<CH>(acc, gacc, use, xt, yt, zt, pdv, d2f, d1);
    }
#else
    for (int k = 0; k < cnt; ++k) {
      const RecRegs use = load_rec(recs + nb[k]);
      pair_update_r<CH>(acc, gacc, use, xt, yt, zt, pdv, d2f, d1);
    }
#endif
    npairs += cnt;
  }
#if NDTB_STRICT_ACC_SMEM
#pragma unroll
  for (int k = 0; k < kRegAccs; ++k) SACC(k) = gacc[k];
#endif
#if NDTB_STRICT_ACC_SMEM
  block_reduce_store_resident(npairs, partials + (size_t)blockIdx.x * kPartialStride);
#else
  block_reduce_store(acc, npairs, partials + (size_t)blockIdx.x * kPartialStride);
#endif
}

// ---------------------------------------------------------------------------------------
// K4: computeHessian / updateHessian (NDT_I:551-641) and calculateScore (NDT_I:919-952): all FP64,
// neighbours always by radius search.  Only launched when the More-Thuente loop can run
// (step_size <= trans_eps / 2) or on explicit request.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_hessian_f64(const AlignState* __restrict__ states, const VoxelTable* __restrict__ tables,
                                                     const float4* __restrict__ src_pool, const int2* __restrict__ blockmap, int chunk,
                                                     double* __restrict__ partials) {
  const int2 bm = blockmap[blockIdx.x];
  const AlignState& st = states[bm.x];
  const int mode = st.req.mode;
  if ((mode != 3 && mode != 4) || st.done) return;
  __shared__ double red[4][kPartialStride];
  const VoxelTable tab = tables[st.target];
  const int tid = threadIdx.x;
  double acc[kAccN];
  for (int k = 0; k < kAccN; ++k) acc[k] = 0.0;
  int npairs = 0;
  const float4* __restrict__ src = st.src;
  const int end = min(bm.y + chunk, st.n_src);
  const double d1 = st.req.d1, d2 = st.req.d2, d3 = st.req.d3;
  for (int i = bm.y + tid; i < end; i += 128) {
    const float4 p = src[i];
    float xt, yt, zt;
    xform(st.req.T, p.x, p.y, p.z, xt, yt, zt);
    int nb[27];
    const int cnt = find_neighbors(tab, 0, xt, yt, zt, nb);
    npairs += cnt;
    const double x[3] = {(double)p.x, (double)p.y, (double)p.z};
    double pg[3][6];
    double hv[6][3];  // a b c d e f
    if (mode == 3 && cnt > 0) {
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 6; ++c) pg[r][c] = (r == c) ? 1.0 : 0.0;
      auto dj = [&](int r) { return (x[0] * st.req.jang_d[r][0] + x[1] * st.req.jang_d[r][1]) + x[2] * st.req.jang_d[r][2]; };
      auto dh = [&](int r) { return (x[0] * st.req.hang_d[r][0] + x[1] * st.req.hang_d[r][1]) + x[2] * st.req.hang_d[r][2]; };
      pg[1][3] = dj(0); pg[2][3] = dj(1); pg[0][4] = dj(2); pg[1][4] = dj(3); pg[2][4] = dj(4); pg[0][5] = dj(5); pg[1][5] = dj(6); pg[2][5] = dj(7);
      hv[0][0] = 0; hv[0][1] = dh(0); hv[0][2] = dh(1);
      hv[1][0] = 0; hv[1][1] = dh(2); hv[1][2] = dh(3);
      hv[2][0] = 0; hv[2][1] = dh(4); hv[2][2] = dh(5);
      hv[3][0] = dh(6); hv[3][1] = dh(7); hv[3][2] = dh(8);
      hv[4][0] = dh(9); hv[4][1] = dh(10); hv[4][2] = dh(11);
      hv[5][0] = dh(12); hv[5][1] = dh(13); hv[5][2] = dh(14);
    }
    for (int k = 0; k < cnt; ++k) {
      const int v = nb[k];
      const VoxelRec* r = tab.rec + v;
      const double* C = tab.icov64 + (size_t)v * 9;
      const double xr[3] = {(double)xt - r->mean[0], (double)yt - r->mean[1], (double)zt - r->mean[2]};
      double Cx[3];
      for (int a = 0; a < 3; ++a) Cx[a] = (C[a * 3] * xr[0] + C[a * 3 + 1] * xr[1]) + C[a * 3 + 2] * xr[2];
      const double q = (xr[0] * Cx[0] + xr[1] * Cx[1]) + xr[2] * Cx[2];
      if (mode == 4) {
        const double e = exp(-d2 * q / 2);
        const double score_inc = -d1 * e - d3;
        acc[0] += score_inc / cnt;
        continue;
      }
      double e = d2 * exp(-d2 * q / 2);   // NDT_I:619
      if (e > 1 || e < 0 || e != e) continue;
      e *= d1;
      // block (i, j) of point_hessian_: rows 3i..3i+2, column j (NDT_I:478-486)
      const int hidx[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
      for (int a = 0; a < 6; ++a) {
        double cd[3];
        for (int rr = 0; rr < 3; ++rr) cd[rr] = (C[rr * 3] * pg[0][a] + C[rr * 3 + 1] * pg[1][a]) + C[rr * 3 + 2] * pg[2][a];
        const double xcd = (xr[0] * cd[0] + xr[1] * cd[1]) + xr[2] * cd[2];
        for (int b = 0; b < 6; ++b) {
          double cb[3];
          for (int rr = 0; rr < 3; ++rr) cb[rr] = (C[rr * 3] * pg[0][b] + C[rr * 3 + 1] * pg[1][b]) + C[rr * 3 + 2] * pg[2][b];
          const double xcb = (xr[0] * cb[0] + xr[1] * cb[1]) + xr[2] * cb[2];
          double xch = 0.0;
          if (a >= 3 && b >= 3) {
            const double* h = hv[hidx[a - 3][b - 3]];
            double chh[3];
            for (int rr = 0; rr < 3; ++rr) chh[rr] = (C[rr * 3] * h[0] + C[rr * 3 + 1] * h[1]) + C[rr * 3 + 2] * h[2];
            xch = (xr[0] * chh[0] + xr[1] * chh[1]) + xr[2] * chh[2];
          }
          const double pc = (pg[0][b] * cd[0] + pg[1][b] * cd[1]) + pg[2][b] * cd[2];
          acc[7 + a * 6 + b] += e * (-d2 * xcd * xcb + xch + pc);   // NDT_I:634-637
        }
      }
    }
  }
  const int lane = tid & 31, warp = tid >> 5;
  for (int k = 0; k < kAccN; ++k) {
    double v = acc[k];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp][k] = v;
  }
  {
    int v = npairs;
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp][kAccN] = (double)v;
  }
  __syncthreads();
  if (tid < kPartialStride) {
    double v = 0;
    for (int w = 0; w < 4; ++w) v += red[w][tid];
    partials[(size_t)blockIdx.x * kPartialStride + tid] = v;
  }
}

// ---------------------------------------------------------------------------------------
// pose helpers (float; Eigen::Translation * AngleAxis(x) * AngleAxis(y) * AngleAxis(z))
// ---------------------------------------------------------------------------------------
__host__ __device__ void angle_axis_matrix(float angle, int axis_id, float R[3][3]) {
  // Eigen AngleAxis<float>::toRotationMatrix() for a unit coordinate axis; sinf/cosf evaluated in
  // double and rounded (the libm-independent definition shared with the oracle)
  float ax[3] = {0.f, 0.f, 0.f};
  ax[axis_id] = 1.f;
#ifdef __CUDA_ARCH__
  double sd, cd;
  sincos((double)angle, &sd, &cd);
#else
  const double sd = std::sin((double)angle), cd = std::cos((double)angle);
#endif
  const float s = (float)sd, c = (float)cd;
  const float sa[3] = {s * ax[0], s * ax[1], s * ax[2]};
  const float one_c = 1.f - c;
  const float ca[3] = {one_c * ax[0], one_c * ax[1], one_c * ax[2]};
  float tmp;
  tmp = ca[0] * ax[1]; R[0][1] = tmp - sa[2]; R[1][0] = tmp + sa[2];
  tmp = ca[0] * ax[2]; R[0][2] = tmp + sa[1]; R[2][0] = tmp - sa[1];
  tmp = ca[1] * ax[2]; R[1][2] = tmp - sa[0]; R[2][1] = tmp + sa[0];
  for (int i = 0; i < 3; ++i) R[i][i] = ca[i] * ax[i] + c;
}
__host__ __device__ void mul3f(const float A[3][3], const float B[3][3], float C[3][3]) {
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C[i][j] = (A[i][0] * B[0][j] + A[i][1] * B[1][j]) + A[i][2] * B[2][j];
}
// NDT_H:210-217, NDT_I:815-819: column-major float 4x4
__host__ __device__ void pose_to_matrix(const double p[6], float T[16]) {
  float Rx[3][3], Ry[3][3], Rz[3][3], Rxy[3][3], R[3][3];
  angle_axis_matrix((float)p[3], 0, Rx);
  angle_axis_matrix((float)p[4], 1, Ry);
  angle_axis_matrix((float)p[5], 2, Rz);
  mul3f(Rx, Ry, Rxy);
  mul3f(Rxy, Rz, R);
  for (int c = 0; c < 3; ++c) {
    for (int r = 0; r < 3; ++r) T[c * 4 + r] = R[r][c];
    T[c * 4 + 3] = 0.f;
  }
  T[12] = (float)p[0]; T[13] = (float)p[1]; T[14] = (float)p[2]; T[15] = 1.f;
}

void pose_to_matrix_host(const double p[6], float T[16]) { pose_to_matrix(p, T); }

// computeAngleDerivatives (NDT_I:286-398)
__device__ void angle_derivatives(const double p[6], EvalReq& rq) {
  double cx, cy, cz, sx, sy, sz;
  if (fabs(p[3]) < 10e-5) { cx = 1.0; sx = 0.0; } else sincos(p[3], &sx, &cx);
  if (fabs(p[4]) < 10e-5) { cy = 1.0; sy = 0.0; } else sincos(p[4], &sy, &cy);
  if (fabs(p[5]) < 10e-5) { cz = 1.0; sz = 0.0; } else sincos(p[5], &sz, &cz);
  const double J[8][3] = {
      {(-sx * sz + cx * sy * cz), (-sx * cz - cx * sy * sz), (-cx * cy)},
      {(cx * sz + sx * sy * cz), (cx * cz - sx * sy * sz), (-sx * cy)},
      {(-sy * cz), sy * sz, cy},
      {sx * cy * cz, (-sx * cy * sz), sx * sy},
      {(-cx * cy * cz), cx * cy * sz, (-cx * sy)},
      {(-cy * sz), (-cy * cz), 0},
      {(cx * cz - sx * sy * sz), (-cx * sz - sx * sy * cz), 0},
      {(sx * cz + cx * sy * sz), (cx * sy * cz - sx * sz), 0}};
  for (int r = 0; r < 8; ++r)
    for (int c = 0; c < 3; ++c) { rq.jang_d[r][c] = J[r][c]; rq.jang[r][c] = (float)J[r][c]; }
  const double Hh[15][3] = {
      {(-cx * sz - sx * sy * cz), (-cx * cz + sx * sy * sz), sx * cy},
      {(-sx * sz + cx * sy * cz), (-cx * sy * sz - sx * cz), (-cx * cy)},
      {(cx * cy * cz), (-cx * cy * sz), (cx * sy)},
      {(sx * cy * cz), (-sx * cy * sz), (sx * sy)},
      {(-sx * cz - cx * sy * sz), (sx * sz - cx * sy * cz), 0},
      {(cx * cz - sx * sy * sz), (-sx * sy * cz - cx * sz), 0},
      {(-cy * cz), (cy * sz), (sy)},
      {(-sx * sy * cz), (sx * sy * sz), (sx * cy)},
      {(cx * sy * cz), (-cx * sy * sz), (-cx * cy)},
      {(sy * sz), (sy * cz), 0},
      {(-sx * cy * sz), (-sx * cy * cz), 0},
      {(cx * cy * sz), (cx * cy * cz), 0},
      {(-cy * cz), (cy * sz), 0},
      {(-cx * sz - sx * sy * cz), (-cx * cz + sx * sy * sz), 0},
      {(-sx * sz + cx * sy * cz), (-cx * sy * sz - sx * cz), 0}};
  for (int r = 0; r < 15; ++r)
    for (int c = 0; c < 3; ++c) { rq.hang_d[r][c] = Hh[r][c]; rq.hang[r][c] = (float)Hh[r][c]; }
}

__device__ void set_req_transform(AlignState& st) {
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 4; ++c) st.req.T[r * 4 + c] = st.final_T[c * 4 + r];
}

// x = pinv(H) b by a one-sided Jacobi SVD -- Eigen::JacobiSVD(H, FullU|FullV).solve(b), NDT_I:119-121.
__device__ void svd_solve6(const double* H, const double* b, double* x) {
  double U[6][6], V[6][6];
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) { U[i][j] = H[i * 6 + j]; V[i][j] = (i == j) ? 1.0 : 0.0; }
  for (int sweep = 0; sweep < 60; ++sweep) {
    bool rotated = false;
    for (int p = 0; p < 5; ++p)
      for (int q = p + 1; q < 6; ++q) {
        double alpha = 0, beta = 0, gamma = 0;
        for (int k = 0; k < 6; ++k) { alpha += U[k][p] * U[k][p]; beta += U[k][q] * U[k][q]; gamma += U[k][p] * U[k][q]; }
        if (gamma == 0.0 || fabs(gamma) <= 1e-16 * sqrt(alpha * beta)) continue;
        rotated = true;
        const double zeta = (beta - alpha) / (2.0 * gamma);
        const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
        for (int k = 0; k < 6; ++k) {
          const double up = U[k][p], uq = U[k][q];
          U[k][p] = c * up - s * uq; U[k][q] = s * up + c * uq;
          const double vp = V[k][p], vq = V[k][q];
          V[k][p] = c * vp - s * vq; V[k][q] = s * vp + c * vq;
        }
      }
    if (!rotated) break;
  }
  double sig[6], smax = 0;
  for (int j = 0; j < 6; ++j) {
    double n2 = 0;
    for (int k = 0; k < 6; ++k) n2 += U[k][j] * U[k][j];
    sig[j] = sqrt(n2);
    smax = fmax(smax, sig[j]);
  }
  const double thresh = 6.0 * DBL_EPSILON * smax;   // Eigen's default rank threshold: diagSize * epsilon
  for (int i = 0; i < 6; ++i) x[i] = 0;
  for (int j = 0; j < 6; ++j) {
    if (!(sig[j] > thresh)) continue;
    double ub = 0;
    for (int k = 0; k < 6; ++k) ub += U[k][j] * b[k];
    const double coef = ub / (sig[j] * sig[j]);
    for (int i = 0; i < 6; ++i) x[i] += V[i][j] * coef;
  }
}

// Newton step H x = b.  Eigen::JacobiSVD::solve is a pseudo-inverse solve; for the full-rank,
// moderately conditioned Hessians of a registration that is the ordinary solution, which Gaussian
// elimination with row pivoting plus one step of iterative refinement delivers to the same ~1e-15 in a
// few hundred FP64 operations instead of the ~10^5 of the Jacobi sweeps.  Everything is unrolled with
// compile-time indices (row swaps are predicated selects), so the factors live in registers: the version
// with run-time indices went through local memory and made k_newton 15-25 % of a batched step.
// Rank-deficient or non-finite systems (empty grid: H = 0) return false and take the SVD path, which keeps
// the minimum-norm / zero-step behaviour of NDT_I:119-130.
struct LU6 {
  double a[6][6];   // L (unit diagonal, below) and U (on and above the diagonal) of the row-permuted matrix
  int piv[6];       // row chosen at step k
};
__device__ __forceinline__ bool lu_factor6(const double* H, LU6& f, double* pivot_ratio) {
  *pivot_ratio = 0.0;
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j < 6; ++j) f.a[i][j] = H[i * 6 + j];
  double pmax = 0.0, pmin = DBL_MAX;
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    int pi = k;
    double best = fabs(f.a[k][k]);
#pragma unroll
    for (int i = k + 1; i < 6; ++i) {
      const double v = fabs(f.a[i][k]);
      if (v > best) { best = v; pi = i; }
    }
    if (!(best > 0.0) || !(best <= DBL_MAX)) return false;   // singular column, NaN or inf
    f.piv[k] = pi;
#pragma unroll
    for (int i = k + 1; i < 6; ++i) {       // swap rows k and pi (pi is a run-time value: predicated selects)
      const bool sw = (i == pi);
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        const double t = f.a[k][j], u = f.a[i][j];
        f.a[k][j] = sw ? u : t;
        f.a[i][j] = sw ? t : u;
      }
    }
    pmax = fmax(pmax, best);
    pmin = fmin(pmin, best);
    const double inv = 1.0 / f.a[k][k];
#pragma unroll
    for (int i = k + 1; i < 6; ++i) {
      const double m = f.a[i][k] * inv;
      f.a[i][k] = m;
#pragma unroll
      for (int j = k + 1; j < 6; ++j) f.a[i][j] -= m * f.a[k][j];
    }
  }
  *pivot_ratio = pmin / pmax;
  return pmin > 1e-10 * pmax;
}
__device__ __forceinline__ void lu_apply6(const LU6& f, const double* b, double* x) {
  double y[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) y[i] = b[i];
#pragma unroll
  for (int k = 0; k < 6; ++k) {             // all row swaps first: the stored L is in fully permuted row order
#pragma unroll
    for (int i = k + 1; i < 6; ++i) {
      const bool sw = (i == f.piv[k]);
      const double t = y[k], u = y[i];
      y[k] = sw ? u : t;
      y[i] = sw ? t : u;
    }
  }
#pragma unroll
  for (int k = 0; k < 6; ++k) {             // forward substitution with unit-diagonal L
#pragma unroll
    for (int i = k + 1; i < 6; ++i) y[i] -= f.a[i][k] * y[k];
  }
#pragma unroll
  for (int i = 5; i >= 0; --i) {
    double v = y[i];
#pragma unroll
    for (int j = i + 1; j < 6; ++j) v -= f.a[i][j] * x[j];
    x[i] = v / f.a[i][i];
  }
}
// returns true when the pseudo-inverse (SVD) branch produced the step, or when the system is so ill-conditioned (smallest /
// largest pivot below 1e-6; a registration's Hessian sits around 1e-3) that a 1e-7 perturbation of H decides the step: the
// factorised kernel's arithmetic noise turns an exactly rank-deficient Hessian (one source point: rank 3) into such a system
__device__ bool solve6(const double* H, const double* b, double* x) {
  LU6 f;
  double ratio;
  if (!lu_factor6(H, f, &ratio)) { svd_solve6(H, b, x); return true; }
  const bool ill = ratio < 1e-6;
  lu_apply6(f, b, x);
  double r[6], d[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) {   // one refinement step: r = b - H x
    double v = b[i];
#pragma unroll
    for (int j = 0; j < 6; ++j) v -= H[i * 6 + j] * x[j];
    r[i] = v;
  }
  lu_apply6(f, r, d);
  bool ok = true;
#pragma unroll
  for (int i = 0; i < 6; ++i) { x[i] += d[i]; ok = ok && (fabs(x[i]) <= DBL_MAX); }
  if (!ok) svd_solve6(H, b, x);
  return !ok || ill;
}

// ---------------------------------------------------------------------------------------
// More-Thuente pieces (NDT_I:644-757), verbatim semantics
// ---------------------------------------------------------------------------------------
__device__ bool update_interval_mt(double& a_l, double& f_l, double& g_l, double& a_u, double& f_u, double& g_u, double a_t,
                                   double f_t, double g_t) {
  if (f_t > f_l) { a_u = a_t; f_u = f_t; g_u = g_t; return false; }
  else if (g_t * (a_l - a_t) > 0) { a_l = a_t; f_l = f_t; g_l = g_t; return false; }
  else if (g_t * (a_l - a_t) < 0) { a_u = a_l; f_u = f_l; g_u = g_l; a_l = a_t; f_l = f_t; g_l = g_t; return false; }
  return true;
}
__device__ double trial_value_mt(double a_l, double f_l, double g_l, double a_u, double f_u, double g_u, double a_t, double f_t,
                                 double g_t) {
  if (f_t > f_l) {
    const double z = 3 * (f_t - f_l) / (a_t - a_l) - g_t - g_l;
    const double w = sqrt(z * z - g_t * g_l);
    const double a_c = a_l + (a_t - a_l) * (w - g_l - z) / (g_t - g_l + 2 * w);
    const double a_q = a_l - 0.5 * (a_l - a_t) * g_l / (g_l - (f_l - f_t) / (a_l - a_t));
    if (fabs(a_c - a_l) < fabs(a_q - a_l)) return a_c;
    return 0.5 * (a_q + a_c);
  } else if (g_t * g_l < 0) {
    const double z = 3 * (f_t - f_l) / (a_t - a_l) - g_t - g_l;
    const double w = sqrt(z * z - g_t * g_l);
    const double a_c = a_l + (a_t - a_l) * (w - g_l - z) / (g_t - g_l + 2 * w);
    const double a_s = a_l - (a_l - a_t) / (g_l - g_t) * g_l;
    if (fabs(a_c - a_t) >= fabs(a_s - a_t)) return a_c;
    return a_s;
  } else if (fabs(g_t) <= fabs(g_l)) {
    const double z = 3 * (f_t - f_l) / (a_t - a_l) - g_t - g_l;
    const double w = sqrt(z * z - g_t * g_l);
    const double a_c = a_l + (a_t - a_l) * (w - g_l - z) / (g_t - g_l + 2 * w);
    const double a_s = a_l - (a_l - a_t) / (g_l - g_t) * g_l;
    const double a_t_next = (fabs(a_c - a_t) < fabs(a_s - a_t)) ? a_c : a_s;
    if (a_t > a_l) return fmin(a_t + 0.66 * (a_u - a_t), a_t_next);
    return fmax(a_t + 0.66 * (a_u - a_t), a_t_next);
  } else {
    const double z = 3 * (f_t - f_u) / (a_t - a_u) - g_t - g_u;
    const double w = sqrt(z * z - g_t * g_u);
    return a_u + (a_t - a_u) * (w - g_u - z) / (g_t - g_u + 2 * w);
  }
}

// ---------------------------------------------------------------------------------------
// K7: the device-resident Newton / line-search state machine.  One block per registration; after
// every derivative launch it (1) sums that registration's per-block partials in a fixed order and
// (2) advances computeTransformation / computeStepLengthMT (NDT_I:73-164, 760-916) until the next
// evaluation request is ready, with no host round trip.
// ---------------------------------------------------------------------------------------
constexpr int PH_START = kPhaseStart, PH_EVAL_ONLY = kPhaseEvalOnly;

struct NewtonCtx {
  AlignState* st;
  double* trace;
  int trace_cap;
  int* n_active;
};

__device__ void record(NewtonCtx& c, const double* p, double step, int chflag, int kind) {
  AlignState& st = *c.st;
  if (c.trace && st.n_evals < c.trace_cap) {
    double* d = c.trace + (size_t)st.n_evals * kTraceRec;
    for (int k = 0; k < 6; ++k) d[k] = p[k];
    d[6] = st.score;
    for (int k = 0; k < 6; ++k) d[7 + k] = st.g[k];
    for (int k = 0; k < 36; ++k) d[13 + k] = st.H[k];
    d[49] = step; d[50] = chflag; d[51] = kind;
  }
  st.n_evals++;
}
__device__ void finish(NewtonCtx& c, int converged) {
  AlignState& st = *c.st;
  st.trans_probability = st.score / (double)st.n_src;   // NDT_I:127, 163
  st.converged = converged;
  st.done = 1;
  st.phase = PH_DONE;
  st.req.mode = 0;
  atomicSub(c.n_active, 1);
}
__device__ void request_eval(AlignState& st, int mode) {
  pose_to_matrix(st.x_t, st.final_T);     // final_transformation_ = T(x_t), NDT_I:815-819 / 857-862
  set_req_transform(st);
  angle_derivatives(st.x_t, st.req);      // computeDerivatives -> computeAngleDerivatives(p), NDT_I:199
  st.req.mode = mode;
}

// top of the while loop of computeTransformation (NDT_I:114-134) + prologue of computeStepLengthMT
__device__ void newton_top(NewtonCtx& c) {
  AlignState& st = *c.st;
  while (true) {
    double neg_g[6], delta[6];
    for (int k = 0; k < 6; ++k) neg_g[k] = -st.g[k];
    // a rank-deficient Hessian sends the step through the pseudo-inverse, where a 1e-7 difference in H is amplified
    // without bound: such a registration is one the hybrid mode replays with the term-by-term kernel
    if (solve6(st.H, neg_g, delta)) st.borderline = 1;
    double n2 = 0;
    for (int k = 0; k < 6; ++k) n2 += delta[k] * delta[k];
    const double norm = sqrt(n2);
    if (norm == 0 || norm != norm) { finish(c, norm == norm ? 1 : 0); return; }   // NDT_I:126-130
    for (int k = 0; k < 6; ++k) st.dir[k] = delta[k] / norm;                       // delta_p.normalize()
    st.dir_norm = norm;
    // computeStepLengthMT(p, dir, norm, step_size_, eps/2, ...)
    const double step_max = st.step_size, step_min = st.trans_eps / 2;
    st.phi_0 = -st.score;
    double dp = 0;
    for (int k = 0; k < 6; ++k) dp += st.g[k] * st.dir[k];
    st.d_phi_0 = -dp;
    {   // the sign test below is a discrete decision: note when it is taken within 1e-3 (relative to |g|, dir is a unit vector) of zero
      double g2 = 0;
      for (int k = 0; k < 6; ++k) g2 += st.g[k] * st.g[k];
      if (fabs(st.d_phi_0) <= 1e-3 * sqrt(g2)) st.borderline = 1;
    }
    if (st.d_phi_0 >= 0) {
      if (st.d_phi_0 == 0) {
        // return 0 (NDT_I:776-777): delta_p = 0, p unchanged, no evaluation
        const double a = 0.0;
        if (st.nr_iterations > st.max_iter || (st.nr_iterations && (fabs(a) < st.trans_eps))) { st.nr_iterations++; finish(c, 1); return; }
        st.nr_iterations++;
        continue;
      }
      st.d_phi_0 *= -1;
      for (int k = 0; k < 6; ++k) st.dir[k] *= -1;
    }
    const double mu = 1.e-4;
    st.step_iterations = 0;
    st.a_l = 0; st.a_u = 0;
    st.f_l = st.phi_0 - st.phi_0 - mu * st.d_phi_0 * st.a_l;   // auxilaryFunction_PsiMT, NDT_H:425-429
    st.g_l = st.d_phi_0 - mu * st.d_phi_0;                     // auxilaryFunction_dPsiMT, NDT_H:437-441
    st.f_u = st.phi_0 - st.phi_0 - mu * st.d_phi_0 * st.a_u;
    st.g_u = st.d_phi_0 - mu * st.d_phi_0;
    st.interval_converged = (step_max - step_min) > 0 ? 1 : 0;  // NDT_I:807 (the reference's inverted test, kept)
    st.open_interval = 1;
    double a_t = norm;
    a_t = fmin(a_t, step_max);
    a_t = fmax(a_t, step_min);
    st.a_t = a_t;
    st.last_norm = norm;
    if (st.nr_iterations >= 0 && st.nr_iterations < 32) st.norm_hist[st.nr_iterations] = (float)norm;
    if (st.nr_iterations > 0 && a_t > st.prev_at && st.prev_at < step_max) st.n_rise++;
    st.prev_at = a_t;
    st.min_at = st.nr_iterations > 0 ? fmin(st.min_at, a_t) : a_t;
    for (int k = 0; k < 6; ++k) st.x_t[k] = st.p[k] + st.dir[k] * a_t;
    request_eval(st, 1);
    st.phase = PH_MT_FIRST;
    return;
  }
}

// after the line search returned a_t (NDT_I:135-157)
__device__ void post_step(NewtonCtx& c) {
  AlignState& st = *c.st;
  const double a = st.a_t;
  for (int k = 0; k < 6; ++k) st.p[k] = st.p[k] + st.dir[k] * a;
  int conv = 0;
  if (st.nr_iterations > st.max_iter || (st.nr_iterations && (fabs(a) < st.trans_eps))) conv = 1;
  if (fabs(fabs(a) - st.trans_eps) <= 1e-3 * st.trans_eps) st.borderline = 1;   // the |step| < eps test taken at its threshold
  st.nr_iterations++;
  if (conv) { finish(c, 1); return; }
  // first pass of a hybrid batch: a registration past the replay threshold is going to be run again with the term-by-term
  // kernel whatever it does from here on -- it leaves now (unconverged, which flags it) instead of riding out the tail
  if (st.abort_iters > 0 && st.nr_iterations > st.abort_iters) { finish(c, 0); return; }
  newton_top(c);
}

__device__ void mt_continue(NewtonCtx& c, double phi_t, double d_phi_t, double psi_t, double d_psi_t) {
  AlignState& st = *c.st;
  const double nu = 0.9;
  const double step_max = st.step_size, step_min = st.trans_eps / 2;
  if (!st.interval_converged && st.step_iterations < 10 && !(psi_t <= 0 && d_phi_t <= -nu * st.d_phi_0)) {   // NDT_I:839-840
    double a_t;
    if (st.open_interval) a_t = trial_value_mt(st.a_l, st.f_l, st.g_l, st.a_u, st.f_u, st.g_u, st.a_t, psi_t, d_psi_t);
    else a_t = trial_value_mt(st.a_l, st.f_l, st.g_l, st.a_u, st.f_u, st.g_u, st.a_t, phi_t, d_phi_t);
    a_t = fmin(a_t, step_max);
    a_t = fmax(a_t, step_min);
    st.a_t = a_t;
    for (int k = 0; k < 6; ++k) st.x_t[k] = st.p[k] + st.dir[k] * a_t;
    request_eval(st, 2);
    st.phase = PH_MT_LOOP;
    return;
  }
  if (st.step_iterations) {   // NDT_I:912-913
    st.req.mode = 3;          // same transform, same angle tables ("unnecessary", NDT_I:574)
    st.phase = PH_MT_HESS;
    return;
  }
  post_step(c);
}

// The state machine proper: runs on one thread, on a SHARED-MEMORY copy of the registration's state
// (global-memory latency on every st.* access made this kernel 15 % of a batched step).
__device__ void newton_advance(AlignState& st, const double* sums, double* trace, int trace_cap, int* n_active) {
  NewtonCtx c{&st, trace, trace_cap, n_active};
  const double mu = 1.e-4;
  if (st.phase == PH_START) {
    // first computeDerivatives of computeTransformation (NDT_I:112): cloud already moved by the guess
    set_req_transform(st);
    angle_derivatives(st.p, st.req);
    st.req.mode = 1;
    st.phase = PH_INIT;
    return;
  }
  st.pairs = (long long)sums[kAccN];
  st.pairs_total += st.pairs;
  switch (st.phase) {
    case PH_EVAL_ONLY: {
      st.score = sums[0];
      for (int k = 0; k < 6; ++k) st.g[k] = sums[1 + k];
      for (int k = 0; k < 36; ++k) st.H[k] = sums[7 + k];
      st.req.mode = 0;
      st.done = 1;
      atomicSub(n_active, 1);
      return;
    }
    case PH_INIT: {
      st.score = sums[0];
      for (int k = 0; k < 6; ++k) st.g[k] = sums[1 + k];
      for (int k = 0; k < 36; ++k) st.H[k] = sums[7 + k];
      record(c, st.p, 0.0, 1, 0);
      newton_top(c);
      return;
    }
    case PH_MT_FIRST:
    case PH_MT_LOOP: {
      st.score = sums[0];
      for (int k = 0; k < 6; ++k) st.g[k] = sums[1 + k];
      for (int k = 0; k < 36; ++k) st.H[k] = sums[7 + k];   // zeros when compute_hessian == false (NDT_I:184)
      record(c, st.x_t, st.a_t, st.phase == PH_MT_FIRST ? 1 : 0, st.phase == PH_MT_FIRST ? 1 : 2);
      const double phi_t = -st.score;
      double dp = 0;
      for (int k = 0; k < 6; ++k) dp += st.g[k] * st.dir[k];
      const double d_phi_t = -dp;
      const double psi_t = phi_t - st.phi_0 - mu * st.d_phi_0 * st.a_t;
      const double d_psi_t = d_phi_t - mu * st.d_phi_0;
      if (st.phase == PH_MT_LOOP) {   // tail of the loop body, NDT_I:881-906
        if (st.open_interval && (psi_t <= 0 && d_psi_t >= 0)) {
          st.open_interval = 0;
          st.f_l = st.f_l + st.phi_0 - mu * st.d_phi_0 * st.a_l;
          st.g_l = st.g_l + mu * st.d_phi_0;
          st.f_u = st.f_u + st.phi_0 - mu * st.d_phi_0 * st.a_u;
          st.g_u = st.g_u + mu * st.d_phi_0;
        }
        if (st.open_interval) st.interval_converged = update_interval_mt(st.a_l, st.f_l, st.g_l, st.a_u, st.f_u, st.g_u, st.a_t, psi_t, d_psi_t) ? 1 : 0;
        else st.interval_converged = update_interval_mt(st.a_l, st.f_l, st.g_l, st.a_u, st.f_u, st.g_u, st.a_t, phi_t, d_phi_t) ? 1 : 0;
        st.step_iterations++;
      }
      mt_continue(c, phi_t, d_phi_t, psi_t, d_psi_t);
      return;
    }
    case PH_MT_HESS: {
      for (int k = 0; k < 36; ++k) st.H[k] = sums[7 + k];
      record(c, st.x_t, st.a_t, 1, 3);
      post_step(c);
      return;
    }
    default: return;
  }
}

// cond != 0 (a single registration run as ONE CUDA graph, ndt_api.cu AlignGraph): the kernel is the last node of the body of a
// WHILE conditional node and decides whether the body runs again -- until the registration is done, or eval_cap evaluations
// have run (a loop that cannot end would hang the device).
__global__ void __launch_bounds__(64) k_newton(AlignState* __restrict__ states, int n_items, const double* __restrict__ partials,
                                               double* __restrict__ trace, int trace_cap, int* __restrict__ n_active,
                                               unsigned long long cond, int eval_cap) {
  const int item = blockIdx.x;
  if (item >= n_items) return;
  if (states[item].done) {
    if (cond && threadIdx.x == 0) cudaGraphSetConditional((cudaGraphConditionalHandle)cond, 0u);
    return;
  }
  __shared__ double sums[kPartialStride];
  __shared__ AlignState st;
  static_assert(sizeof(AlignState) % 8 == 0, "AlignState is copied as 64-bit words");
  const int tid = threadIdx.x;
  {
    const unsigned long long* src = reinterpret_cast<const unsigned long long*>(states + item);
    unsigned long long* dst = reinterpret_cast<unsigned long long*>(&st);
    for (int k = tid; k < (int)(sizeof(AlignState) / 8); k += 64) dst[k] = src[k];
  }
  __syncthreads();
  if (st.req.mode != 0 && tid < kPartialStride) {
    double v = 0;
    const double* p = partials + (size_t)st.first_block * kPartialStride + tid;
    // fixed order (deterministic); the loads of eight partials are issued together so that the chain is eight
    // additions per memory round trip, not one
    const int nb = st.n_blocks;
    int b = 0;
    for (; b + 8 <= nb; b += 8) {
      double q[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) q[u] = __ldg(p + (size_t)(b + u) * kPartialStride);
#pragma unroll
      for (int u = 0; u < 8; ++u) v += q[u];
    }
    for (; b < nb; ++b) v += __ldg(p + (size_t)b * kPartialStride);
    sums[tid] = v;
  }
  __syncthreads();
  if (tid == 0) {
    newton_advance(st, sums, trace ? trace + (size_t)item * trace_cap * kTraceRec : nullptr, trace_cap, n_active);
    if (cond) cudaGraphSetConditional((cudaGraphConditionalHandle)cond, (!st.done && st.n_evals < eval_cap) ? 1u : 0u);
  }
  __syncthreads();
  {
    const unsigned long long* src = reinterpret_cast<const unsigned long long*>(&st);
    unsigned long long* dst = reinterpret_cast<unsigned long long*>(states + item);
    for (int k = tid; k < (int)(sizeof(AlignState) / 8); k += 64) dst[k] = src[k];
  }
}

// ---------------------------------------------------------------------------------------
// small utility kernels
// ---------------------------------------------------------------------------------------
__global__ void k_neighbors(const VoxelTable* __restrict__ table, const float* __restrict__ xyz, size_t n, int method,
                            int* __restrict__ cnt, int* __restrict__ ijk) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const VoxelTable t = *table;
  int nb[27];
  const int c = find_neighbors(t, method, xyz[i * 3], xyz[i * 3 + 1], xyz[i * 3 + 2], nb);
  cnt[i] = c;
  for (int k = 0; k < c; ++k) {
    int a, b, d;
    unpack_key(t.vkey[nb[k]], a, b, d);
    ijk[(i * 27 + k) * 3] = a; ijk[(i * 27 + k) * 3 + 1] = b; ijk[(i * 27 + k) * 3 + 2] = d;
  }
}
__global__ void k_transform(const float4* __restrict__ src, size_t n, const float* __restrict__ T16, float4* __restrict__ out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float T[12];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 4; ++c) T[r * 4 + c] = T16[c * 4 + r];
  const float4 p = src[i];
  float4 o;
  xform(T, p.x, p.y, p.z, o.x, o.y, o.z);
  o.w = p.w;
  out[i] = o;
}
__global__ void k_prepare_eval(AlignState* st, const float* T16, const double* p6, int mode) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  for (int k = 0; k < 16; ++k) st->final_T[k] = T16[k];
  set_req_transform(*st);
  double p[6];
  for (int k = 0; k < 6; ++k) p[k] = p6[k];
  angle_derivatives(p, st->req);
  st->req.mode = mode;
  st->phase = PH_EVAL_ONLY;
  st->done = 0;
}
__global__ void k_prefix_equal(const uint4* __restrict__ a, const uint4* __restrict__ b, size_t n, int* __restrict__ flag) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const uint4 x = a[i], y = b[i];
    if (x.x != y.x || x.y != y.y || x.z != y.z) { *flag = 0; }
  }
}

template <int METHOD>
static cudaError_t launch_deriv_m(const AlignState* states, const VoxelTable* tables, const float4* src_pool, const int2* blockmap,
                                  int n_blocks, int chunk, double* partials, bool both, cudaStream_t s) {
  const size_t sm = reduce_smem_bytes(kThreads);
  // the attribute is per device: one flag per device of the process (handles / batches on distinct GPUs, distinct host threads)
  static std::atomic<unsigned long long> configured{0};
  int dev = 0;
  cudaGetDevice(&dev);
  const unsigned long long bit = 1ull << (dev & 63);
  if (!(configured.load(std::memory_order_acquire) & bit)) {
    cudaFuncSetAttribute(k_derivatives<METHOD, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    cudaFuncSetAttribute(k_derivatives<METHOD, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
#if NDTB_STRICT_ACC_SMEM
    // two 93 KB blocks per SM: ask for the large shared-memory carve-out
    cudaFuncSetAttribute(k_derivatives<METHOD, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(k_derivatives<METHOD, false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
#endif
    configured.fetch_or(bit, std::memory_order_release);
  }
  k_derivatives<METHOD, true><<<n_blocks, kThreads, sm, s>>>(states, tables, src_pool, blockmap, chunk, partials);
  if (both) k_derivatives<METHOD, false><<<n_blocks, kThreads, sm, s>>>(states, tables, src_pool, blockmap, chunk, partials);
  return cudaGetLastError();
}
// `method` = the batch's neighbour search method; `with_gradient_only` additionally launches the
// compute_hessian == false variant (More-Thuente trial evaluations, NDT_I:869).
cudaError_t launch_derivatives(const AlignState* states, const VoxelTable* tables, const float4* src_pool, const int2* blockmap,
                               int n_blocks, int chunk, double* partials, int method, bool with_gradient_only, cudaStream_t s) {
  if (n_blocks <= 0) return cudaSuccess;
  switch (method) {
    case 0: return launch_deriv_m<0>(states, tables, src_pool, blockmap, n_blocks, chunk, partials, with_gradient_only, s);
    case 1: return launch_deriv_m<1>(states, tables, src_pool, blockmap, n_blocks, chunk, partials, with_gradient_only, s);
    case 3: return launch_deriv_m<3>(states, tables, src_pool, blockmap, n_blocks, chunk, partials, with_gradient_only, s);
    default: return launch_deriv_m<2>(states, tables, src_pool, blockmap, n_blocks, chunk, partials, with_gradient_only, s);
  }
}
cudaError_t launch_hessian_f64(const AlignState* states, const VoxelTable* tables, const float4* src_pool, const int2* blockmap,
                               int n_blocks, int chunk, double* partials, cudaStream_t s) {
  if (n_blocks <= 0) return cudaSuccess;
  k_hessian_f64<<<n_blocks, 128, 0, s>>>(states, tables, src_pool, blockmap, chunk, partials);
  return cudaGetLastError();
}
cudaError_t launch_newton(AlignState* states, int n_items, const double* partials, double* trace, int trace_cap, int* n_active,
                          cudaStream_t s, unsigned long long cond, int eval_cap) {
  if (n_items <= 0) return cudaSuccess;
  k_newton<<<n_items, 64, 0, s>>>(states, n_items, partials, trace, trace_cap, n_active, cond, eval_cap);
  return cudaGetLastError();
}
cudaError_t launch_neighbors(const VoxelTable* table, const float* xyz, size_t n, int method, int* cnt, int* ijk, cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  k_neighbors<<<(unsigned)((n + 127) / 128), 128, 0, s>>>(table, xyz, n, method, cnt, ijk);
  return cudaGetLastError();
}
cudaError_t launch_transform(const float4* src, size_t n, const float* T16, float4* out, cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  k_transform<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(src, n, T16, out);
  return cudaGetLastError();
}
// PGO::LoopFindNearKeyframesCloud (reference src/pgo_node.cpp:530-546): every keyframe of the window moved into the map
// frame by its own pose (TransformCloud2Map, pgo_node.cpp:568-590) and stacked; one launch, blockIdx.y = keyframe.
__global__ void k_submap_gather(const SubmapSeg* __restrict__ segs, float4* __restrict__ out) {
  const SubmapSeg& sg = segs[blockIdx.y];
  const float4* __restrict__ src = sg.src;
  float4* __restrict__ dst = out + sg.offset;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < sg.n; i += (size_t)gridDim.x * blockDim.x) {
    const float4 p = src[i];
    float4 o;
    xform(sg.T, p.x, p.y, p.z, o.x, o.y, o.z);
    o.w = p.w;
    dst[i] = o;
  }
}
cudaError_t launch_submap_gather(const SubmapSeg* segs_dev, int n_seg, size_t max_n, float4* out, cudaStream_t s) {
  if (n_seg <= 0 || max_n == 0) return cudaSuccess;
  const unsigned bx = (unsigned)((max_n + 255) / 256 < 64 ? (max_n + 255) / 256 : 64);
  k_submap_gather<<<dim3(bx, (unsigned)n_seg), 256, 0, s>>>(segs_dev, out);
  return cudaGetLastError();
}
cudaError_t launch_prepare_eval(AlignState* st, const float* T16, const double* p6, int mode, cudaStream_t s) {
  k_prepare_eval<<<1, 32, 0, s>>>(st, T16, p6, mode);
  return cudaGetLastError();
}
// Small host -> device transfers of the run set-up (registration states, block map, table descriptors) are PULLED by
// a kernel out of mapped pinned host memory instead of being pushed with cudaMemcpyAsync: an H2D memcpy queues on the
// copy engine behind any other batch's multi-megabyte cloud uploads and would stall this batch for their duration.
__global__ void k_pull(uint4* __restrict__ dst, const uint4* __restrict__ src, size_t n16) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}
cudaError_t launch_pull(void* dst_dev, const void* src_mapped_dev, size_t bytes, cudaStream_t s) {
  if (bytes == 0) return cudaSuccess;
  const size_t n16 = (bytes + 15) / 16;   // both buffers are allocated in multiples of 16 bytes
  const unsigned blocks = (unsigned)((n16 + 255) / 256 < 1184 ? (n16 + 255) / 256 : 1184);
  k_pull<<<blocks, 256, 0, s>>>(static_cast<uint4*>(dst_dev), static_cast<const uint4*>(src_mapped_dev), n16);
  return cudaGetLastError();
}
// Work decomposition of the NEXT evaluations, redone on the device whenever the host polls: only registrations still
// running get blocks, and the block size follows the live point count (the same rule the host applies to the whole batch
// at the start), so the tail iterations -- a few stragglers of a batch of thousands -- neither walk a grid of finished
// blocks nor serialise 1024 points behind one block.  One block; plan = {blocks, chunk} in mapped host memory.
__global__ void __launch_bounds__(1024) k_replan(AlignState* __restrict__ states, int n_items, int2* __restrict__ blockmap,
                                                 int* __restrict__ plan_mapped, int target_blocks) {
  __shared__ unsigned long long s_pts[32];
  __shared__ int s_scan[32];
  __shared__ int s_base, s_chunk;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned long long pts = 0;
  for (int i = tid; i < n_items; i += 1024)
    if (!states[i].done) pts += (unsigned long long)states[i].n_src;
  for (int o = 16; o; o >>= 1) pts += __shfl_down_sync(0xffffffffu, pts, o);
  if (lane == 0) s_pts[warp] = pts;
  __syncthreads();
  if (tid == 0) {
    unsigned long long tot = 0;
    for (int w = 0; w < 32; ++w) tot += s_pts[w];
    long long c = (long long)((tot / (unsigned long long)target_blocks + 255) / 256 * 256);
    s_chunk = (int)(c < 256 ? 256 : (c > 1024 ? 1024 : c));
    s_base = 0;
  }
  __syncthreads();
  const int chunk = s_chunk;
  for (int i0 = 0; i0 < n_items; i0 += 1024) {
    const int i = i0 + tid;
    int nb = 0, n_src = 0;
    if (i < n_items && !states[i].done) { n_src = states[i].n_src; nb = (n_src + chunk - 1) / chunk; }
    int incl = nb;
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) s_scan[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int w = s_scan[lane];
      for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += t; }
      s_scan[lane] = w;   // inclusive over warps
    }
    __syncthreads();
    const int first = s_base + (warp ? s_scan[warp - 1] : 0) + incl - nb;
    if (i < n_items && nb > 0) {
      states[i].first_block = first;
      states[i].n_blocks = nb;
      for (int b = 0; b < nb; ++b) blockmap[first + b] = make_int2(i, b * chunk);
    }
    __syncthreads();
    if (tid == 0) s_base += s_scan[31];
    __syncthreads();
  }
  if (tid == 0) {
    plan_mapped[0] = s_base;
    plan_mapped[1] = chunk;
    __threadfence_system();
  }
}
cudaError_t launch_replan(AlignState* states, int n_items, int2* blockmap, int* plan_mapped, int target_blocks, cudaStream_t s) {
  k_replan<<<1, 1024, 0, s>>>(states, n_items, blockmap, plan_mapped, target_blocks < 1 ? 1 : target_blocks);
  return cudaGetLastError();
}
// One block per selected registration: its state record copied out of mapped pinned host memory (the replay pass of a
// hybrid batch re-arms only the registrations it runs again; the others keep their finished state on the device).
__global__ void k_pull_items(AlignState* __restrict__ states, const AlignState* __restrict__ src, const int* __restrict__ idx, int n) {
  if ((int)blockIdx.x >= n) return;
  const int i = idx[blockIdx.x];
  const unsigned long long* s = reinterpret_cast<const unsigned long long*>(src + i);
  unsigned long long* d = reinterpret_cast<unsigned long long*>(states + i);
  for (int k = threadIdx.x; k < (int)(sizeof(AlignState) / 8); k += blockDim.x) d[k] = s[k];
}
cudaError_t launch_pull_items(AlignState* states, const AlignState* src_mapped_dev, const int* idx_mapped_dev, int n, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  k_pull_items<<<n, 64, 0, s>>>(states, src_mapped_dev, idx_mapped_dev, n);
  return cudaGetLastError();
}
// packed x y z records (12 bytes) -> the 16-byte records the kernels read (w = 0)
__global__ void k_widen_xyz(const float* __restrict__ packed, float4* __restrict__ dst, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  dst[i] = make_float4(packed[i * 3], packed[i * 3 + 1], packed[i * 3 + 2], 0.f);
}
cudaError_t launch_widen_xyz(const float* packed_dev, float4* dst, size_t n, cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  k_widen_xyz<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(packed_dev, dst, n);
  return cudaGetLastError();
}
__global__ void k_set_int(int* dst, int v) { *dst = v; }
cudaError_t launch_set_int(int* dst, int v, cudaStream_t s) {
  k_set_int<<<1, 1, 0, s>>>(dst, v);
  return cudaGetLastError();
}
__global__ void k_publish(const int* __restrict__ src, int* __restrict__ dst) { *dst = *src; __threadfence_system(); }
cudaError_t launch_publish(const int* src_dev, int* dst_mapped_host, cudaStream_t s) {
  k_publish<<<1, 1, 0, s>>>(src_dev, dst_mapped_host);
  return cudaGetLastError();
}
cudaError_t launch_prefix_equal(const float4* a, const float4* b, size_t n, int* flag, cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  k_prefix_equal<<<296, 256, 0, s>>>(reinterpret_cast<const uint4*>(a), reinterpret_cast<const uint4*>(b), n, flag);
  return cudaGetLastError();
}

}  // namespace ndtb
