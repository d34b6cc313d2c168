// examg_transfer.cu -- inter-grid transfer: full-weighting restriction and
// (bi/tri)linear prolongation with correction.
//
// Reference: Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4:49-107,221-236;
// default stencils C/s/e/operator/l4/L4_DefaultRestriction.scala:29-90 and
// L4_DefaultProlongation.scala:30-45; parity split of the prolongation loop
// C/s/e/stencil/ir/IR_FindStencilConvolutions.scala:122-170.
//
// Arithmetic contract (same as oracle/examg_oracle.c og_restrict/og_prolong_add):
//   restriction : entries grouped by number m of non-zero offsets (weight
//                 scale*2^-(dim+m)); each group summed in lexicographic (dz,dy,dx)
//                 order; acc = w0*S0; acc = fma(w1,S1,acc); ... fma(w_dim,S_dim,acc)
//   prolongation: s = sum of the 2^m contributing coarse values (m = number of odd
//                 fine indices) in lexicographic (dz,dy,dx) order; uf = fma(2^-m, s, uf)
#include "examg_device.cuh"

namespace examg {

template <int DIM>
__global__ void __launch_bounds__(256)
k_restrict(const GridArgs gc, int fpitch, long long fplane, double w0, double w1, double w2, double w3,
           const double *__restrict__ rf, double *__restrict__ rhsc, int kz) {
  const int i = gc.lo[0] + blockIdx.x * blockDim.x + threadIdx.x;
  const int z0 = gc.lo[2] + blockIdx.z * kz;
  const int z1 = min(z0 + kz, gc.hi[2]);
  if (i >= gc.hi[0]) return;
  for (int j = gc.lo[1] + blockIdx.y * blockDim.y + threadIdx.y; j < gc.hi[1]; j += gridDim.y * blockDim.y)
    for (int k = z0; k < z1; ++k) {
      const long long q = 2 * i + (long long)fpitch * (2 * j) + (DIM == 3 ? fplane * (2 * k) : 0);
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
      rhsc[i + (long long)gc.pitch * j + gc.plane * k] = acc;
    }
}

template <int DIM>
__global__ void __launch_bounds__(256)
k_prolong(const GridArgs gf, int cpitch, long long cplane, const double *__restrict__ uc, double *uf, int kz) {
  const int i = gf.lo[0] + blockIdx.x * blockDim.x + threadIdx.x;
  const int z0 = gf.lo[2] + blockIdx.z * kz;
  const int z1 = min(z0 + kz, gf.hi[2]);
  if (i >= gf.hi[0]) return;
  const int ox = i & 1, ci = i >> 1;
  for (int j = gf.lo[1] + blockIdx.y * blockDim.y + threadIdx.y; j < gf.hi[1]; j += gridDim.y * blockDim.y) {
    const int oy = j & 1, cj = j >> 1;
    for (int k = z0; k < z1; ++k) {
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
      const long long p = i + (long long)gf.pitch * j + gf.plane * k;
      uf[p] = fma(wgt, s, uf[p]);
    }
  }
}

int launch_restrict_generic(examg_ctx *c, int lvl_fine, double scale, const double *rf, double *rhsc) {
  const LevelGeom &gf = *geom(c, lvl_fine);
  const LevelGeom &gc = *geom(c, lvl_fine - 1);
  const GridArgs ga = grid_args(gc);
  const int ex = gc.hi[0] - gc.lo[0], ey = gc.hi[1] - gc.lo[1], ez = gc.hi[2] - gc.lo[2];
  if (ex <= 0 || ey <= 0 || ez <= 0) return EXAMG_OK;
  const Shape sh = make_shape(ex, ey, ez);
  const int dim = c->cfg.dim;
  double w[4];
  for (int m = 0; m <= 3; ++m) w[m] = scale * ldexp(1.0, -(dim + m));
  if (dim == 3)
    k_restrict<3><<<sh.grid, sh.block, 0, c->stream>>>(ga, gf.pitch, gf.plane, w[0], w[1], w[2], w[3], rf, rhsc, sh.kz);
  else
    k_restrict<2><<<sh.grid, sh.block, 0, c->stream>>>(ga, gf.pitch, gf.plane, w[0], w[1], w[2], w[3], rf, rhsc, sh.kz);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

int launch_prolong_generic(examg_ctx *c, int lvl_fine, const double *uc, double *uf) {
  const LevelGeom &gf = *geom(c, lvl_fine);
  const LevelGeom &gc = *geom(c, lvl_fine - 1);
  const GridArgs ga = grid_args(gf);
  const int ex = gf.hi[0] - gf.lo[0], ey = gf.hi[1] - gf.lo[1], ez = gf.hi[2] - gf.lo[2];
  if (ex <= 0 || ey <= 0 || ez <= 0) return EXAMG_OK;
  const Shape sh = make_shape(ex, ey, ez);
  if (c->cfg.dim == 3)
    k_prolong<3><<<sh.grid, sh.block, 0, c->stream>>>(ga, gc.pitch, gc.plane, uc, uf, sh.kz);
  else
    k_prolong<2><<<sh.grid, sh.block, 0, c->stream>>>(ga, gc.pitch, gc.plane, uc, uf, sh.kz);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

}  // namespace examg
