// ndt_fast.cu -- K3, production variant: computeDerivatives (reference
// xchu_mapping/include/pclomp/ndt_omp_impl.hpp:175-283 = NDT_I) with the per-pair work of
// updateDerivatives (NDT_I:491-548) factorised per POINT.
//
// For one source point x with Jacobian J (3x6, NDT_I:401-445) and second-derivative vectors
// H_ij, and its neighbour voxels k with w_k = C_k (T x - mu_k), f_k = d1 d2 exp(-d2/2 q_k):
//     score    += -d1 sum_k exp(..)
//     gradient += J^T a                                   a = sum_k f_k w_k          (3)
//     Hessian  += J^T (M - d2 B) J + [a . H_ij]           M = sum_k f_k C_k          (6, symmetric)
//                                                         B = sum_k f_k w_k w_k^T    (6, symmetric)
// because J and H_ij depend on the point only.  A pair costs ~50 float instructions (three
// 128-bit loads of a 48 B record, a 3x3 symmetric mat-vec, one expf, 16 FMAs) instead of the
// ~600 of the term-by-term form, and the 6x6 algebra runs once per point on the 21 unique
// entries.  Per-point results are accumulated in FP64 per thread and reduced in a fixed order
// (warp columns through shared memory, then warps), so results do not depend on batch layout.
//
// Numerics contract: WHICH voxels a point sees is bit-identical to the reference (ndt_device.cuh);
// the per-point float arithmetic differs from the reference's operation order (FMAs, symmetric C,
// factorisation, expf instead of exp(double)), which moves score / gradient / Hessian by ~1e-7
// relative -- inside the 1e-6 parity gate, tested against the oracle and against the strict
// term-by-term kernel in ndt_core.cu.  This file is compiled WITHOUT -fmad=false.
#include <atomic>
#include <cfloat>
#include <cmath>

#include "ndt_device.cuh"
#include "ndt_internal.h"

namespace ndtb {

#ifndef NDTB_FAST_THREADS
#define NDTB_FAST_THREADS 128
#endif
#ifndef NDTB_FAST_MINBLOCKS
#define NDTB_FAST_MINBLOCKS 4
#endif
constexpr int kFT = NDTB_FAST_THREADS;
constexpr int kFAcc = 28;          // score, g[6], H upper triangle [21]
constexpr int kFCols = kFAcc + 1;  // + pair count
constexpr int kFPad = 33;

struct FastShared {
  float T[12];
  float4 jh[23];   // rows of j_ang (8) then h_ang (15), NDT_I:329-336, 362-396
  float d2f, d1d2f, nhalf_d2f;
  double neg_d1;
  VoxelTable tab;
};

__device__ __forceinline__ int hidx(int i, int j) { return i * 6 - (i * (i - 1)) / 2 + (j - i); }   // i <= j

struct FRec { float4 a, b, c; };
__device__ __forceinline__ FRec load_frec(const FastRec* __restrict__ r) {
  FRec f;
  const float4* p = reinterpret_cast<const float4*>(r);
  f.a = __ldg(p); f.b = __ldg(p + 1); f.c = __ldg(p + 2);
  return f;
}

// Fixed-order block reduction of kFCols per-thread doubles; thread k < 44 then writes entry k of the
// 44-double partial in the layout k_newton sums (score, g[6], H[36] row-major, pair count).
__device__ __forceinline__ void fast_reduce_store(const double* acc, int npairs, double* __restrict__ out) {
  extern __shared__ double red_sm[];                 // [warps][kFCols][kFPad] + [warps][kFCols]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = kFT >> 5;
  double* mine = red_sm + (size_t)warp * kFCols * kFPad;
#pragma unroll
  for (int k = 0; k < kFAcc; ++k) mine[k * kFPad + lane] = acc[k];
  mine[kFAcc * kFPad + lane] = (double)npairs;
  __syncwarp();
  double* wsum = red_sm + (size_t)nwarps * kFCols * kFPad;
  if (lane < kFCols) {
    double v = 0;
#pragma unroll 8
    for (int l = 0; l < 32; ++l) v += mine[lane * kFPad + l];
    wsum[warp * kFCols + lane] = v;
  }
  __syncthreads();
  if (tid < kPartialStride) {
    int col;
    if (tid < 7) col = tid;
    else if (tid == 43) col = kFAcc;
    else {
      const int i = (tid - 7) / 6, j = (tid - 7) % 6;
      col = 7 + (i <= j ? hidx(i, j) : hidx(j, i));
    }
    double v = 0;
    for (int w = 0; w < nwarps; ++w) v += wsum[w * kFCols + col];
    out[tid] = v;
  }
}
static size_t fast_smem_bytes() {
  const int nw = kFT / 32;
  return ((size_t)nw * kFCols * kFPad + (size_t)nw * kFCols) * sizeof(double);
}

// DIRECT7 over the dense cell index, block-uniform set-up held in registers.  x / leaf == x * (1 / leaf) bit for bit when leaf
// is a power of two (1 / leaf is exact, both are the correctly rounded value of the same real number).
struct Dense7 {
  const int* dense;
  int dlo0, dlo1, dlo2, dn0, dn1, dn2, sxy;
  float leaf, inv_leaf;
  bool dense32, pow2;
};
__device__ __forceinline__ Dense7 make_dense7(const VoxelTable& t, bool direct7) {
  Dense7 d;
  d.dense = t.dense;
  d.dlo0 = t.dlo[0]; d.dlo1 = t.dlo[1]; d.dlo2 = t.dlo[2];
  d.dn0 = t.dn[0]; d.dn1 = t.dn[1]; d.dn2 = t.dn[2];
  d.sxy = d.dn0 * d.dn1;
  d.dense32 = direct7 && t.dense != nullptr && t.slots != nullptr && (long long)d.dn0 * d.dn1 < (1ll << 31) &&
              (long long)d.sxy * d.dn2 < (1ll << 31);
  d.leaf = t.leaf;
  const unsigned leaf_bits = __float_as_uint(d.leaf);
  d.pow2 = (leaf_bits & 0x807fffffu) == 0u && (leaf_bits >> 23) >= 64u && (leaf_bits >> 23) <= 190u;
  d.inv_leaf = __fdiv_rn(1.0f, d.leaf);
  return d;
}
// neighbors7() (ndt_device.cuh) with the dense branch on the hoisted quantities and 32-bit index arithmetic
__device__ __forceinline__ void lookup7(const Dense7& d, const VoxelTable& t, float xt, float yt, float zt, int* v) {
  if (d.dense32) {
    const int ci = (int)floorf(d.pow2 ? xt * d.inv_leaf : __fdiv_rn(xt, d.leaf)), cj = (int)floorf(d.pow2 ? yt * d.inv_leaf : __fdiv_rn(yt, d.leaf)),
              ck = (int)floorf(d.pow2 ? zt * d.inv_leaf : __fdiv_rn(zt, d.leaf));
    const unsigned ui = (unsigned)(ci - d.dlo0) - 1u, uj = (unsigned)(cj - d.dlo1) - 1u, uk = (unsigned)(ck - d.dlo2) - 1u;
    if (ui >= (unsigned)(d.dn0 - 2) || uj >= (unsigned)(d.dn1 - 2) || uk >= (unsigned)(d.dn2 - 2)) {
#pragma unroll
      for (int o = 0; o < 7; ++o) v[o] = -1;
      return;
    }
    const int* __restrict__ c = d.dense + (((ck - d.dlo2) * d.dn1 + (cj - d.dlo1)) * d.dn0 + (ci - d.dlo0));
    v[0] = __ldg(c); v[1] = __ldg(c + 1); v[2] = __ldg(c - 1); v[3] = __ldg(c + d.dn0); v[4] = __ldg(c - d.dn0);
    v[5] = __ldg(c + d.sxy); v[6] = __ldg(c - d.sxy);
  } else {
    neighbors7(t, xt, yt, zt, v);
  }
}

template <int METHOD, bool CH>
__global__ void __launch_bounds__(kFT, NDTB_FAST_MINBLOCKS)
k_derivatives_fast(const AlignState* __restrict__ states, const VoxelTable* __restrict__ tables, const float4* __restrict__ src_pool,
                   const int2* __restrict__ blockmap, int chunk, double* __restrict__ partials) {
  const int2 bm = blockmap[blockIdx.x];
  const AlignState& st = states[bm.x];
  if (st.req.mode != (CH ? 1 : 2) || st.done || st.req.search_method != METHOD) return;
  __shared__ FastShared sh;
  const int tid = threadIdx.x;
  if (tid < 12) sh.T[tid] = st.req.T[tid];
  if (tid >= 32 && tid < 32 + 23) {
    const int r = tid - 32;
    const float* row = r < 8 ? st.req.jang[r] : st.req.hang[r - 8];
    sh.jh[r] = make_float4(row[0], row[1], row[2], 0.f);
  }
  if (tid == kFT - 1) {
    const double d1 = st.req.d1, d2 = st.req.d2;
    const float d2f = (float)d2;                       // float gauss_d2 = gauss_d2_ (NDT_I:504)
    sh.d2f = d2f;
    sh.nhalf_d2f = -0.5f * d2f;
    sh.d1d2f = (float)(d1 * (double)d2f);
    sh.neg_d1 = -d1;
    sh.tab = tables[st.target];
  }
  __syncthreads();
  double acc[kFAcc];
#define ACC(k) acc[k]
#pragma unroll
  for (int k = 0; k < kFAcc; ++k) ACC(k) = 0.0;
  int npairs = 0;
  const float4* __restrict__ src = st.src;
  const int end = min(bm.y + chunk, st.n_src);
  const FastRec* __restrict__ recs = sh.tab.frec;
  const float d2f = sh.d2f, nhalf_d2f = sh.nhalf_d2f, d1d2f = sh.d1d2f;
  const Dense7 d7 = make_dense7(sh.tab, METHOD == 2);
  // the source record (PFP >= 1) and the seven cell codes (PFP == 2) of the thread's NEXT point are fetched before this point's
  // pairs are evaluated: one / two of the three dependent global loads of a point leave the critical path
  float4 p_nx = make_float4(0.f, 0.f, 0.f, 0.f);
  if (bm.y + tid < end) {
    p_nx = __ldg(src + bm.y + tid);
  }
  for (int i = bm.y + tid; i < end; i += kFT) {
    const float4 p = p_nx;
    float xt, yt, zt;
    xform(sh.T, p.x, p.y, p.z, xt, yt, zt);
    if (i + kFT < end) {
      p_nx = __ldg(src + i + kFT);
    }
    float S = 0.f, a0 = 0.f, a1 = 0.f, a2 = 0.f;
    float M00 = 0.f, M01 = 0.f, M02 = 0.f, M11 = 0.f, M12 = 0.f, M22 = 0.f;
    float B00 = 0.f, B01 = 0.f, B02 = 0.f, B11 = 0.f, B12 = 0.f, B22 = 0.f;
    // one (point, voxel) pair of updateDerivatives (NDT_I:491-548), accumulated per point
#define NDTB_PAIR(r)                                                                                         \
    {                                                                                                        \
      const float x0 = (xt - (r).a.x) - (r).a.w, x1 = (yt - (r).a.y) - (r).b.x, x2 = (zt - (r).a.z) - (r).b.y; \
      const float C00 = (r).b.z, C01 = (r).b.w, C02 = (r).c.x, C11 = (r).c.y, C12 = (r).c.z, C22 = (r).c.w;  \
      const float w0 = fmaf(x2, C02, fmaf(x1, C01, x0 * C00));                                               \
      const float w1 = fmaf(x2, C12, fmaf(x1, C11, x0 * C01));                                               \
      const float w2 = fmaf(x2, C22, fmaf(x1, C12, x0 * C02));                                               \
      const float q = fmaf(x2, w2, fmaf(x1, w1, x0 * w0));                                                   \
      const float es = expf(nhalf_d2f * q);                      /* NDT_I:507 */                             \
      const float eg = d2f * es;                                 /* NDT_I:511 */                             \
      if (!(eg > 1.f || eg < 0.f || eg != eg)) {                 /* NDT_I:514-515: else the pair adds nothing */ \
        const float f = d1d2f * es;                              /* NDT_I:518 */                             \
        S += es;                                                                                             \
        const float t0 = f * w0, t1 = f * w1, t2 = f * w2;                                                   \
        a0 += t0; a1 += t1; a2 += t2;                                                                        \
        if (CH) {                                                                                            \
          M00 = fmaf(f, C00, M00); M01 = fmaf(f, C01, M01); M02 = fmaf(f, C02, M02);                         \
          M11 = fmaf(f, C11, M11); M12 = fmaf(f, C12, M12); M22 = fmaf(f, C22, M22);                         \
          B00 = fmaf(t0, w0, B00); B01 = fmaf(t0, w1, B01); B02 = fmaf(t0, w2, B02);                         \
          B11 = fmaf(t1, w1, B11); B12 = fmaf(t1, w2, B12); B22 = fmaf(t2, w2, B22);                         \
        }                                                                                                    \
      }                                                                                                      \
    }
    if (METHOD == 2) {
      // DIRECT7: the seven cells in the reference's order, fully unrolled and predicated (no compaction,
      // no local memory); the record of cell o+1 is in flight while cell o is being evaluated.
      int v[7];
      lookup7(d7, sh.tab, xt, yt, zt, v);
      int cnt = 0;
#pragma unroll
      for (int o = 0; o < 7; ++o) cnt += v[o] >= 0 ? 1 : 0;
      if (cnt == 0) continue;
      npairs += cnt;
      FRec nxt = {};
      if (v[0] >= 0) nxt = load_frec(recs + v[0]);
#pragma unroll
      for (int o = 0; o < 7; ++o) {
        const FRec r = nxt;
        if (o + 1 < 7) { if (v[o + 1] >= 0) nxt = load_frec(recs + v[o + 1]); }
        if (v[o] >= 0) NDTB_PAIR(r)
      }
    } else {
      int nb[27];
      const int cnt = find_neighbors(sh.tab, METHOD, xt, yt, zt, nb);
      if (cnt == 0) continue;
      npairs += cnt;
      FRec cur = load_frec(recs + nb[0]);
      for (int k = 0; k < cnt; ++k) {
        const FRec r = cur;
        if (k + 1 < cnt) cur = load_frec(recs + nb[k + 1]);
        NDTB_PAIR(r)
      }
    }
#undef NDTB_PAIR
    // ---- once per point: computePointDerivatives (NDT_I:401-445) and the 6x6 algebra ----------
    ACC(0) += sh.neg_d1 * (double)S;                             // NDT_I:509
#define JH(r) fmaf(sh.jh[r].z, p.z, fmaf(sh.jh[r].y, p.y, sh.jh[r].x * p.x))
    const float g13 = JH(0), g23 = JH(1), g04 = JH(2), g14 = JH(3), g24 = JH(4), g05 = JH(5), g15 = JH(6), g25 = JH(7);
    ACC(1) += (double)a0; ACC(2) += (double)a1; ACC(3) += (double)a2;                         // NDT_I:521-523
    ACC(4) += (double)fmaf(g23, a2, g13 * a1);
    ACC(5) += (double)fmaf(g24, a2, fmaf(g14, a1, g04 * a0));
    ACC(6) += (double)fmaf(g25, a2, fmaf(g15, a1, g05 * a0));
    if (CH) {
      const float nd2 = -d2f;
      const float N00 = fmaf(nd2, B00, M00), N01 = fmaf(nd2, B01, M01), N02 = fmaf(nd2, B02, M02);
      const float N11 = fmaf(nd2, B11, M11), N12 = fmaf(nd2, B12, M12), N22 = fmaf(nd2, B22, M22);
      // P = N * [j3 j4 j5], j3 = (0, g13, g23), j4 = (g04, g14, g24), j5 = (g05, g15, g25)
      const float P00 = fmaf(N02, g23, N01 * g13), P10 = fmaf(N12, g23, N11 * g13), P20 = fmaf(N22, g23, N12 * g13);
      const float P01 = fmaf(N02, g24, fmaf(N01, g14, N00 * g04)), P11 = fmaf(N12, g24, fmaf(N11, g14, N01 * g04)),
                  P21 = fmaf(N22, g24, fmaf(N12, g14, N02 * g04));
      const float P02 = fmaf(N02, g25, fmaf(N01, g15, N00 * g05)), P12 = fmaf(N12, g25, fmaf(N11, g15, N01 * g05)),
                  P22 = fmaf(N22, g25, fmaf(N12, g15, N02 * g05));
      // a . H_ij (NDT_I:534-535): rows 8.. of jh are a2 a3 b2 b3 c2 c3 d1 d2 d3 e1 e2 e3 f1 f2 f3
      const float h33 = fmaf(a2, JH(9), a1 * JH(8));
      const float h34 = fmaf(a2, JH(11), a1 * JH(10));
      const float h35 = fmaf(a2, JH(13), a1 * JH(12));
      const float h44 = fmaf(a2, JH(16), fmaf(a1, JH(15), a0 * JH(14)));
      const float h45 = fmaf(a2, JH(19), fmaf(a1, JH(18), a0 * JH(17)));
      const float h55 = fmaf(a2, JH(22), fmaf(a1, JH(21), a0 * JH(20)));
      ACC(7) += (double)N00; ACC(8) += (double)N01; ACC(9) += (double)N02; ACC(10) += (double)P00; ACC(11) += (double)P01; ACC(12) += (double)P02;
      ACC(13) += (double)N11; ACC(14) += (double)N12; ACC(15) += (double)P10; ACC(16) += (double)P11; ACC(17) += (double)P12;
      ACC(18) += (double)N22; ACC(19) += (double)P20; ACC(20) += (double)P21; ACC(21) += (double)P22;
      ACC(22) += (double)(fmaf(g23, P20, g13 * P10) + h33);
      ACC(23) += (double)(fmaf(g23, P21, g13 * P11) + h34);
      ACC(24) += (double)(fmaf(g23, P22, g13 * P12) + h35);
      ACC(25) += (double)(fmaf(g24, P21, fmaf(g14, P11, g04 * P01)) + h44);
      ACC(26) += (double)(fmaf(g24, P22, fmaf(g14, P12, g04 * P02)) + h45);
      ACC(27) += (double)(fmaf(g25, P22, fmaf(g15, P12, g05 * P02)) + h55);
    }
#undef JH
  }
#undef ACC
  fast_reduce_store(acc, npairs, partials + (size_t)blockIdx.x * kPartialStride);
}

template <int METHOD>
static cudaError_t launch_fast_m(const AlignState* states, const VoxelTable* tables, const float4* src_pool, const int2* blockmap,
                                 int n_blocks, int chunk, double* partials, bool both, cudaStream_t s) {
  const size_t sm = fast_smem_bytes();
  // the attribute is per device: one flag per device of the process (handles / batches on distinct GPUs, distinct host threads)
  static std::atomic<unsigned long long> configured{0};
  int dev = 0;
  cudaGetDevice(&dev);
  const unsigned long long bit = 1ull << (dev & 63);
  if (!(configured.load(std::memory_order_acquire) & bit)) {
    cudaFuncSetAttribute(k_derivatives_fast<METHOD, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    cudaFuncSetAttribute(k_derivatives_fast<METHOD, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    configured.fetch_or(bit, std::memory_order_release);
  }
  k_derivatives_fast<METHOD, true><<<n_blocks, kFT, sm, s>>>(states, tables, src_pool, blockmap, chunk, partials);
  if (both) k_derivatives_fast<METHOD, false><<<n_blocks, kFT, sm, s>>>(states, tables, src_pool, blockmap, chunk, partials);
  return cudaGetLastError();
}

cudaError_t launch_derivatives_fast(const AlignState* states, const VoxelTable* tables, const float4* src_pool, const int2* blockmap,
                                    int n_blocks, int chunk, double* partials, int method, bool with_gradient_only, cudaStream_t s) {
  if (n_blocks <= 0) return cudaSuccess;
  switch (method) {
    case 0: return launch_fast_m<0>(states, tables, src_pool, blockmap, n_blocks, chunk, partials, with_gradient_only, s);
    case 1: return launch_fast_m<1>(states, tables, src_pool, blockmap, n_blocks, chunk, partials, with_gradient_only, s);
    case 3: return launch_fast_m<3>(states, tables, src_pool, blockmap, n_blocks, chunk, partials, with_gradient_only, s);
    default: return launch_fast_m<2>(states, tables, src_pool, blockmap, n_blocks, chunk, partials, with_gradient_only, s);
  }
}

}  // namespace ndtb
