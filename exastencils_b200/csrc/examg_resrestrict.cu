// examg_resrestrict.cu -- residual and full-weighting restriction in ONE pass over the fine grid
// (3D, 7-point constant coefficients).
//
// Reference loops being replaced (Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4:216-224):
//   communicate Solution; loop over Residual { Residual = RHS - A * Solution }; apply bc to Residual
//   communicate Residual; loop over RHS@coarser { RHS@coarser = Restriction * Residual }
//   restriction stencil: C/s/e/operator/l4/L4_DefaultRestriction.scala:29-37,63-90
// Inside mgCycle the Residual of a level only feeds the restriction, so it is not materialised here:
// 16 B read + 1 B written per fine point instead of 24 + 9.  (UpResidual@finest of the Solve loop and
// examg_residual still write the field.)
//
// Design: a thread owns the coarse node (ci,cj) and the fine 2x2 block (2ci..2ci+1, 2cj..2cj+1) and marches
// in z.  Per fine plane it computes the four residuals of its block from a 3-plane register window of
// Solution (in-plane neighbours: warp shuffles in x, L1-cached loads in y), hands the upper row to the
// thread above through shared memory (double buffered, one barrier per plane) and the right column to
// the next lane by shuffle, and accumulates the 27-point sums of two coarse planes at a time.
// Arithmetic: residual as in examg_kernels.cu (declaration order, fma chain), restriction sums as
// restrict_point (examg_point_ops.cuh: four sums by number of non-zero offsets, lexicographic order,
// then w0*S0, fma(w1,S1), fma(w2,S2), fma(w3,S3)) -- bit-identical to the two separate kernels.
#include <cmath>
#include <cstdlib>

#include "examg_device.cuh"

namespace examg {

constexpr int RRX = 32, RRY = 8;  // threads = coarse nodes per CTA tile; column 0 and row 0 are halo only

struct RrArgs {
  int clo[3], chi[3];  // coarse update range
  int fpitch, cpitch;
  long long fplane, cplane;
  double coef[7];
  double w[4];
  int kz;              // coarse planes per CTA
};

template <int ORD>
__device__ __forceinline__ double rr_resid(double c, double xm, double xp, double ym, double yp, double zm, double zp,
                                           double f, const double (&coef)[7]) {
  double acc = coef[0] * c;
  if (ORD == 0) {
    acc = fma(coef[1], xm, acc);
    acc = fma(coef[2], xp, acc);
    acc = fma(coef[3], ym, acc);
    acc = fma(coef[4], yp, acc);
    acc = fma(coef[5], zm, acc);
    acc = fma(coef[6], zp, acc);
  } else {
    acc = fma(coef[1], xp, acc);
    acc = fma(coef[2], xm, acc);
    acc = fma(coef[3], yp, acc);
    acc = fma(coef[4], ym, acc);
    acc = fma(coef[5], zp, acc);
    acc = fma(coef[6], zm, acc);
  }
  return f - acc;
}

template <int ORD>
__global__ void __launch_bounds__(RRX *RRY)
k_resrestrict3(const RrArgs a, const double *__restrict__ u, const double *__restrict__ f, double *__restrict__ rhsc) {
  __shared__ double sm[2][RRY][2 * RRX + 2];  // upper fine row of every thread's block, column 2ci at [2*tx + 2]
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int ci = a.clo[0] - 1 + blockIdx.x * (RRX - 1) + tx;
  const int cj = a.clo[1] - 1 + blockIdx.y * (RRY - 1) + ty;
  const int ck0 = a.clo[2] + blockIdx.z * a.kz, ck1 = min(ck0 + a.kz, a.chi[2]);
  const bool act = ci < a.chi[0] && cj < a.chi[1];  // coarse nodes below hi, incl. the halo column / row at lo-1
  const bool out = act && tx >= 1 && ty >= 1;
  double coef[7];
#pragma unroll
  for (int k = 0; k < 7; ++k) coef[k] = a.coef[k];
  const long long fp = a.fplane;
  const int pitch = a.fpitch;
  const long long base = 2 * ci + (long long)pitch * (2 * cj);

  // Software pipeline: the own block of plane z+2 and the in-plane neighbours / RHS of plane z+1 are
  // loaded while plane z is computed, so every load has a whole plane of arithmetic to complete.
  double um[4], uc[4], up[4], un[4];  // Solution on planes z-1, z, z+1 (resident) and z+2 (in flight)
  auto load_block = [&](int z, double (&dst)[4]) {
    if (act) {
      const double2 lo = *reinterpret_cast<const double2 *>(u + base + fp * z);
      const double2 hi = *reinterpret_cast<const double2 *>(u + base + pitch + fp * z);
      dst[0] = lo.x; dst[1] = lo.y; dst[2] = hi.x; dst[3] = hi.y;
    } else {
      dst[0] = dst[1] = dst[2] = dst[3] = 0.0;
    }
  };
  struct Ring {  // what a plane needs besides the blocks: rows y0-1 and y0+2, RHS, and the columns beyond the warp
    double2 s, n, f0, f1;
    double w0, w1, e0, e1;
  };
  const bool wedge = tx == 0, eedge = tx == RRX - 1 || ci == a.chi[0] - 1;
  auto load_ring = [&](int z, Ring &r) {
    if (act) {
      const double *pz = u + base + fp * z;
      r.s = *reinterpret_cast<const double2 *>(pz - pitch);
      r.n = *reinterpret_cast<const double2 *>(pz + 2 * pitch);
      r.f0 = *reinterpret_cast<const double2 *>(f + base + fp * z);
      r.f1 = *reinterpret_cast<const double2 *>(f + base + pitch + fp * z);
      if (wedge) { r.w0 = pz[-1]; r.w1 = pz[pitch - 1]; }
      if (eedge) { r.e0 = pz[2]; r.e1 = pz[pitch + 2]; }
    }
  };
  const int zs = 2 * ck0 - 1, ze = 2 * ck1 - 1;  // fine planes zs..ze, both odd
  Ring rc, rn;
  rc.s = rc.n = rc.f0 = rc.f1 = rn.s = rn.n = rn.f0 = rn.f1 = make_double2(0.0, 0.0);
  rc.w0 = rc.w1 = rc.e0 = rc.e1 = rn.w0 = rn.w1 = rn.e0 = rn.e1 = 0.0;
  load_block(zs - 1, um);
  load_block(zs, uc);
  load_block(zs + 1, up);
  load_ring(zs, rc);
  double S0 = 0.0, S1 = 0.0, S2 = 0.0, S3 = 0.0;  // sums by number of non-zero offsets

  for (int z = zs; z <= ze; ++z) {
    if (z < ze) {
      load_block(z + 2, un);
      load_ring(z + 1, rn);
    }
    double r[4] = {0.0, 0.0, 0.0, 0.0};
    // x neighbours: the adjacent lanes' blocks (own loads at the warp edges and at the last column)
    double w0 = __shfl_up_sync(0xffffffffu, uc[1], 1), w1 = __shfl_up_sync(0xffffffffu, uc[3], 1);
    double e0 = __shfl_down_sync(0xffffffffu, uc[0], 1), e1 = __shfl_down_sync(0xffffffffu, uc[2], 1);
    if (act) {
      if (wedge) { w0 = rc.w0; w1 = rc.w1; }
      if (eedge) { e0 = rc.e0; e1 = rc.e1; }
      r[0] = rr_resid<ORD>(uc[0], w0, uc[1], rc.s.x, uc[2], um[0], up[0], rc.f0.x, coef);
      r[1] = rr_resid<ORD>(uc[1], uc[0], e0, rc.s.y, uc[3], um[1], up[1], rc.f0.y, coef);
      r[2] = rr_resid<ORD>(uc[2], w1, uc[3], uc[0], rc.n.x, um[2], up[2], rc.f1.x, coef);
      r[3] = rr_resid<ORD>(uc[3], uc[2], e1, uc[1], rc.n.y, um[3], up[3], rc.f1.y, coef);
    }
    // the restriction at (ci,cj) reads fine x = 2ci-1..2ci+1, y = 2cj-1..2cj+1 of this plane:
    // row 2cj-1 is the upper row of the block below (shared memory), column 2ci-1 the right column of lane-1
    const int b = z & 1;
    *reinterpret_cast<double2 *>(&sm[b][ty][2 * tx + 2]) = make_double2(r[2], r[3]);
    const double v3 = __shfl_up_sync(0xffffffffu, r[1], 1), v6 = __shfl_up_sync(0xffffffffu, r[3], 1);
    __syncthreads();
    if (out) {
      const double *row = &sm[b][ty - 1][2 * tx];
      const double v0 = row[1], v1 = row[2], v2 = row[3], v4 = r[0], v5 = r[1], v7 = r[2], v8 = r[3];
      if (b) {  // z = 2ck+1: last plane of coarse plane ck = (z-1)/2, first plane of ck+1
        if (z > zs) {
          S3 += v0; S2 += v1; S3 += v2; S2 += v3; S1 += v4; S2 += v5; S3 += v6; S2 += v7; S3 += v8;
          double acc = a.w[0] * S0;
          acc = fma(a.w[1], S1, acc);
          acc = fma(a.w[2], S2, acc);
          acc = fma(a.w[3], S3, acc);
          rhsc[ci + (long long)a.cpitch * cj + a.cplane * ((z - 1) >> 1)] = acc;
        }
        S3 = v0; S2 = v1; S3 += v2; S2 += v3; S1 = v4; S2 += v5; S3 += v6; S2 += v7; S3 += v8;
      } else {  // z = 2ck: the middle plane
        S2 += v0; S1 += v1; S2 += v2; S1 += v3; S0 = v4; S1 += v5; S2 += v6; S1 += v7; S2 += v8;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) { um[j] = uc[j]; uc[j] = up[j]; up[j] = un[j]; }
    rc = rn;
  }
}

int stencil_order(const Stencil &s);  // examg_fast7.cu

static bool rr_enabled() {
  static int v = -1;
  if (v < 0) {
    const char *e = getenv("EXAMG_NO_FUSED_RESTRICT");
    v = (e && e[0] == '1') ? 0 : 1;
  }
  return v == 1;
}

bool residual_restrict_usable(examg_ctx *c, const Stencil &s) {
  if (!rr_enabled() || c->cfg.dim != 3 || !s.is7 || s.coeff_field >= 0 || stencil_order(s) < 0) return false;
  if (s.level <= c->cfg.min_level || c->periodic_any) return false;
  const LevelGeom &g = *geom(c, s.level);
  return g.n[0] >= 65 && g.n[1] >= 65 && g.n[2] >= 65;
}

// RHS@coarser = scale * Restriction * (RHS - A * Solution); Solution and RHS must carry 2 resp. 1 valid ghost layers
// on the sides with neighbour blocks (the caller exchanges)
int launch_residual_restrict(examg_ctx *c, const Stencil &s, const double *u, const double *rhs, double *rhsc,
                             double scale) {
  const LevelGeom &gf = *geom(c, s.level), &gc = *geom(c, s.level - 1);
  RrArgs a;
  for (int d = 0; d < 3; ++d) { a.clo[d] = gc.lo[d]; a.chi[d] = gc.hi[d]; }
  a.fpitch = gf.pitch; a.cpitch = gc.pitch;
  a.fplane = gf.plane; a.cplane = gc.plane;
  for (int k = 0; k < 7; ++k) a.coef[k] = s.coef[k];
  for (int m = 0; m <= 3; ++m) a.w[m] = scale * ldexp(1.0, -(3 + m));
  const int ex = gc.hi[0] - gc.lo[0], ey = gc.hi[1] - gc.lo[1], ez = gc.hi[2] - gc.lo[2];
  if (ex <= 0 || ey <= 0 || ez <= 0) return EXAMG_OK;
  const int gx = (ex + RRX - 2) / (RRX - 1), gy = (ey + RRY - 2) / (RRY - 1);
  int kz = 32;  // coarse planes per CTA: one extra fine plane per segment; enough CTAs for several waves
  while (kz > 4 && (long long)gx * gy * ((ez + kz - 1) / kz) < 148 * 8 * 3) kz /= 2;
  a.kz = kz;
  const dim3 grid(gx, gy, (ez + kz - 1) / kz), block(RRX, RRY);
  if (stencil_order(s) == 0)
    k_resrestrict3<0><<<grid, block, 0, c->stream>>>(a, u, rhs, rhsc);
  else
    k_resrestrict3<1><<<grid, block, 0, c->stream>>>(a, u, rhs, rhsc);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

}  // namespace examg
