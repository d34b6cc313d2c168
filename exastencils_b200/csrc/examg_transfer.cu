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
#include <cstdlib>

#include "examg_point_ops.cuh"

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
      const double acc = restrict_point<DIM>(rf, q, fpitch, fplane, w0, w1, w2, w3);
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
  for (int j = gf.lo[1] + blockIdx.y * blockDim.y + threadIdx.y; j < gf.hi[1]; j += gridDim.y * blockDim.y)
    for (int k = z0; k < z1; ++k) {
      const long long p = i + (long long)gf.pitch * j + gf.plane * k;
      uf[p] = prolong_point<DIM>(uc, cpitch, cplane, i, j, k, uf[p]);
    }
}

// 3D fast path of the prolongation: one thread per coarse cell (ci, cj) marching over coarse
// planes; it owns the 2x2 fine points (2ci..2ci+1, 2cj..2cj+1) of the fine planes 2ck and 2ck+1,
// keeps the coarse values of planes ck / ck+1 in registers (the x+1 neighbours come from the
// next lane) and moves the fine field with 128-bit loads and stores.  Same sums, same order
// as k_prolong (lexicographic (dz,dy,dx)), so the results are bit-identical.
struct ProlongArgs {
  int lo[3], hi[3];   // fine update range
  int ncx, ncy, ncz;  // coarse nodes
  int fpitch, cpitch;
  long long fplane, cplane;
  int kz;             // coarse planes per block
};

__global__ void __launch_bounds__(256)
k_prolong3_fast(const ProlongArgs a, const double *__restrict__ uc, double *__restrict__ uf) {
  const int lane = threadIdx.x;
  const int ci = blockIdx.x * 32 + lane;
  const int cj = blockIdx.y * blockDim.y + threadIdx.y;
  const int ck0 = blockIdx.z * a.kz, ck1 = min(ck0 + a.kz, a.ncz);
  const bool active = ci < a.ncx && cj < a.ncy;
  // clamp so that the whole warp takes part in the shuffles; inactive threads never store
  const int cic = min(ci, a.ncx - 1), cjc = min(cj, a.ncy - 1);
  const double *pc = uc + cic + (long long)a.cpitch * cjc;
  const int fx = 2 * ci, fy = 2 * cj;
  const bool x0ok = active && fx >= a.lo[0] && fx < a.hi[0], x1ok = active && fx + 1 >= a.lo[0] && fx + 1 < a.hi[0];
  const bool y0ok = fy >= a.lo[1] && fy < a.hi[1], y1ok = fy + 1 >= a.lo[1] && fy + 1 < a.hi[1];
  const bool row0 = (x0ok || x1ok) && y0ok, row1 = (x0ok || x1ok) && y1ok;
  double *pf = uf + fx + (long long)a.fpitch * fy;

  // coarse values of the current plane: c[dy][dx]
  double c00, c01, c10, c11;  // c<dy><dx>
  {
    const double *q = pc + a.cplane * ck0;
    c00 = q[0];
    c10 = q[a.cpitch];
    c01 = __shfl_down_sync(0xffffffffu, c00, 1);
    c11 = __shfl_down_sync(0xffffffffu, c10, 1);
    if (lane == 31) { c01 = q[1]; c11 = q[a.cpitch + 1]; }
  }
  for (int ck = ck0; ck < ck1; ++ck) {
    // next coarse plane (ghost plane ncz exists in the allocation; its values only reach masked points)
    double d00, d01, d10, d11;
    {
      const double *q = pc + a.cplane * (ck + 1);
      d00 = q[0];
      d10 = q[a.cpitch];
      d01 = __shfl_down_sync(0xffffffffu, d00, 1);
      d11 = __shfl_down_sync(0xffffffffu, d10, 1);
      if (lane == 31) { d01 = q[1]; d11 = q[a.cpitch + 1]; }
    }
    const int fz = 2 * ck;
    // ---- fine plane 2ck: coarse plane ck only ----
    if (fz >= a.lo[2] && fz < a.hi[2]) {
      double *p = pf + a.fplane * fz;
      if (row0) {
        double2 v = *reinterpret_cast<double2 *>(p);
        if (x0ok) v.x = fma(1.0, c00, v.x);
        if (x1ok) v.y = fma(0.5, c00 + c01, v.y);
        *reinterpret_cast<double2 *>(p) = v;
      }
      if (row1) {
        double2 v = *reinterpret_cast<double2 *>(p + a.fpitch);
        if (x0ok) v.x = fma(0.5, c00 + c10, v.x);
        if (x1ok) v.y = fma(0.25, ((c00 + c01) + c10) + c11, v.y);
        *reinterpret_cast<double2 *>(p + a.fpitch) = v;
      }
    }
    // ---- fine plane 2ck+1: coarse planes ck and ck+1 ----
    if (fz + 1 >= a.lo[2] && fz + 1 < a.hi[2]) {
      double *p = pf + a.fplane * (fz + 1);
      if (row0) {
        double2 v = *reinterpret_cast<double2 *>(p);
        if (x0ok) v.x = fma(0.5, c00 + d00, v.x);
        if (x1ok) v.y = fma(0.25, ((c00 + c01) + d00) + d01, v.y);
        *reinterpret_cast<double2 *>(p) = v;
      }
      if (row1) {
        double2 v = *reinterpret_cast<double2 *>(p + a.fpitch);
        if (x0ok) v.x = fma(0.25, ((c00 + c10) + d00) + d10, v.x);
        if (x1ok) v.y = fma(0.125, ((((((c00 + c01) + c10) + c11) + d00) + d01) + d10) + d11, v.y);
        *reinterpret_cast<double2 *>(p + a.fpitch) = v;
      }
    }
    c00 = d00; c01 = d01; c10 = d10; c11 = d11;
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
  static const bool no_fast = getenv("EXAMG_NO_FAST_PROLONG") && getenv("EXAMG_NO_FAST_PROLONG")[0] == '1';
  if (c->cfg.dim == 3 && gc.n[0] >= 32 && !no_fast) {
    ProlongArgs a;
    for (int d = 0; d < 3; ++d) { a.lo[d] = gf.lo[d]; a.hi[d] = gf.hi[d]; }
    a.ncx = gc.n[0]; a.ncy = gc.n[1]; a.ncz = gc.n[2];
    a.fpitch = gf.pitch; a.cpitch = gc.pitch;
    a.fplane = gf.plane; a.cplane = gc.plane;
    const int by = 8;
    const int gx = (a.ncx + 31) / 32, gy = (a.ncy + by - 1) / by;
    int kz = 1;  // enough blocks for ~8 waves of 148 SMs x 8 resident blocks, at least 4 planes per block
    while (kz < 16 && (long long)gx * gy * ((a.ncz + 2 * kz - 1) / (2 * kz)) >= 148 * 8 * 4) kz *= 2;
    if (kz < 4) kz = a.ncz < 4 ? a.ncz : 4;
    a.kz = kz;
    k_prolong3_fast<<<dim3(gx, gy, (a.ncz + kz - 1) / kz), dim3(32, by, 1), 0, c->stream>>>(a, uc, uf);
    EXAMG_LAUNCH_CHECK(c);
    return EXAMG_OK;
  }
  const Shape sh = make_shape(ex, ey, ez);
  if (c->cfg.dim == 3)
    k_prolong<3><<<sh.grid, sh.block, 0, c->stream>>>(ga, gc.pitch, gc.plane, uc, uf, sh.kz);
  else
    k_prolong<2><<<sh.grid, sh.block, 0, c->stream>>>(ga, gc.pitch, gc.plane, uc, uf, sh.kz);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

}  // namespace examg
