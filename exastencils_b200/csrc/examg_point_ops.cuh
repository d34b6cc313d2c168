// examg_point_ops.cuh -- the per-point arithmetic of the hot-path loops, shared by the grid
// kernels (examg_kernels.cu, examg_transfer.cu), the coarse CG (examg_cg.cu) and the
// single-CTA coarse V-cycle (examg_coarse.cu), so that every path has ONE floating-point
// association (the contract stated at the top of examg_kernels.cu / examg_transfer.cu).
#pragma once

#include "examg_device.cuh"

namespace examg {

// A*u at linear index p: acc = c0*u0; acc = fma(ck, uk, acc) in declaration order
template <int NST, bool VAR>
__device__ __forceinline__ double conv(const StencilArgs &st, const double *__restrict__ sf, long long sfvol,
                                       const double *u, long long p) {
  const int nst = NST > 0 ? NST : st.nst;
  double acc;
  if (VAR) {
    acc = sf[p] * u[p + st.lin[0]];
#pragma unroll
    for (int k = 1; k < nst; ++k) acc = fma(sf[k * sfvol + p], u[p + st.lin[k]], acc);
  } else {
    acc = st.coef[0] * u[p + st.lin[0]];
#pragma unroll
    for (int k = 1; k < nst; ++k) acc = fma(st.coef[k], u[p + st.lin[k]], acc);
  }
  return acc;
}

// full-weighting restriction at the coarse point whose fine counterpart has linear index q
template <int DIM>
__device__ __forceinline__ double restrict_point(const double *__restrict__ rf, long long q, int fpitch,
                                                 long long fplane, double w0, double w1, double w2, double w3) {
  double S[4];
  bool first[4] = {true, true, true, true};
#pragma unroll
  for (int dz = (DIM == 3 ? -1 : 0); dz <= (DIM == 3 ? 1 : 0); ++dz)
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
      for (int dx = -1; dx <= 1; ++dx) {
        const int m = (dx != 0) + (dy != 0) + (dz != 0);
        const double v = rf[q + dx + (long long)fpitch * dy + fplane * dz];
        if (first[m]) { S[m] = v; first[m] = false; } else S[m] += v;
      }
  double acc = w0 * S[0];
  acc = fma(w1, S[1], acc);
  acc = fma(w2, S[2], acc);
  if (DIM == 3) acc = fma(w3, S[3], acc);
  return acc;
}

// (bi/tri)linear interpolation of the coarse field at fine node (i,j,k), added to uf
template <int DIM>
__device__ __forceinline__ double prolong_point(const double *__restrict__ uc, int cpitch, long long cplane, int i,
                                                int j, int k, double uf) {
  const int ox = i & 1, ci = i >> 1, oy = j & 1, cj = j >> 1;
  const int oz = DIM == 3 ? (k & 1) : 0, ck = DIM == 3 ? (k >> 1) : 0;
  const long long q = ci + (long long)cpitch * cj + cplane * ck;
  double s = 0.0;
  bool first = true;
  for (int dz = 0; dz <= oz; ++dz)
    for (int dy = 0; dy <= oy; ++dy)
      for (int dx = 0; dx <= ox; ++dx) {
        const double v = uc[q + dx + (long long)cpitch * dy + cplane * dz];
        if (first) { s = v; first = false; } else s += v;
      }
  const int m = ox + oy + oz;
  const double wgt = m == 0 ? 1.0 : (m == 1 ? 0.5 : (m == 2 ? 0.25 : 0.125));
  return fma(wgt, s, uf);
}

}  // namespace examg
