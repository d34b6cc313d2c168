// examg_cg.cu -- mgCycle@coarsest: un-preconditioned conjugate gradients on the
// coarsest level, as ONE single-CTA kernel (the grid is a few hundred nodes; the
// reference spends ~8 loops + 3 reductions per iteration on it, each a separate
// OpenMP region / kernel launch + host synchronisation).
//
// Reference: Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4:152-201; generator
// C/s/e/solver/l3/L3_ConjugateGradientForEquation.scala:37-151.  Same recurrences
// and stopping rule as oracle/examg_oracle_solver.c og_cg().  The iteration
// count is data dependent and decided on the device, so the whole V-cycle stays
// free of host round-trips (CUDA-graph capturable).
#include "examg_cg.cuh"

namespace examg {

template <bool VAR>
__global__ void __launch_bounds__(1024)
k_cg(const __grid_constant__ CgArgs a, const double *__restrict__ sf, double *u, const double *__restrict__ rhs,
     double *res, double *p, double *ap, int *iters_out) {
  const int iters = cg_device<VAR>(a, sf, u, rhs, res, p, ap);
  if (threadIdx.x == 0) *iters_out = iters;
}

template <bool VAR>
__global__ void __launch_bounds__(1024)
k_bicgstab(const __grid_constant__ CgArgs a, const double *__restrict__ sf, double *u, const double *__restrict__ rhs,
           double *res, double *v, double *p, double *sv, double *tv, double *rhat, int *iters_out) {
  const int iters = bicgstab_device<VAR>(a, sf, u, rhs, res, v, p, sv, tv, rhat);
  if (threadIdx.x == 0) *iters_out = iters;
}

template <bool VAR>
__global__ void __launch_bounds__(1024)
k_cr(const __grid_constant__ CgArgs a, const double *__restrict__ sf, double *u, const double *__restrict__ rhs,
     double *res, double *p, double *ap, double *ar, int *iters_out) {
  const int iters = cr_device<VAR>(a, sf, u, rhs, res, p, ap, ar);
  if (threadIdx.x == 0) *iters_out = iters;
}

// solver_cgs = "Smoother" (C/s/e/solver/l3/L3_VankaAsSolverForEquation.scala:31-94): residual + norm, then up to max_it
// times { ten smoother steps; residual; norm; stop at rel_tol * initRes }, whole solve in one CTA.  kind 0: damped Jacobi
// between u and `other` (ten steps leave the result in u), kind 1: red-black Gauss-Seidel in place.  w = omega (stencil
// fields) or omega / diag.
template <bool VAR>
__global__ void __launch_bounds__(1024)
k_smoother_solve(const __grid_constant__ CgArgs a, const double *__restrict__ sf, double *u, const double *__restrict__ rhs,
                 double *res, double *other, int kind, double w, int *iters_out) {
  const GridArgs &g = a.g;
  const int ex = g.hi[0] - g.lo[0], ey = g.hi[1] - g.lo[1], ez = g.hi[2] - g.lo[2];
  const int total = ex * ey * ez;
  const int tid = threadIdx.x, nt = blockDim.x;
  auto resnorm = [&]() {
    dd part = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      const double r = rhs[q] - cg_conv<VAR>(a, sf, u, q);
      res[q] = r;
      dd_add(part, r * r);
    }
    return sqrt(cg_sum(part));
  };
  const double initRes = resnorm();
  int iters = 0;
  if (initRes != 0.0) {
    for (int step = 0; step < a.max_it; ++step) {
      iters = step + 1;
      for (int in = 0; in < 10; ++in) {
        if (kind == 0) {
          const double *src = (in & 1) ? other : u;
          double *dst = (in & 1) ? u : other;
          for (int t = tid; t < total; t += nt) {
            const long long q = cg_point(g, t, ex, ey);
            const double ww = VAR ? w / sf[a.st.centre * a.sfvol + q] : w;
            dst[q] = fma(ww, rhs[q] - cg_conv<VAR>(a, sf, src, q), src[q]);
          }
          __syncthreads();
        } else {
          for (int colour = 0; colour < 2; ++colour) {
            for (int t = tid; t < total; t += nt) {
              const int i = g.lo[0] + t % ex, j = g.lo[1] + (t / ex) % ey, k = g.lo[2] + t / (ex * ey);
              if (((i + j + k + g.par) & 1) != colour) continue;
              const long long q = i + (long long)g.pitch * j + g.plane * k;
              const double ww = VAR ? w / sf[a.st.centre * a.sfvol + q] : w;
              u[q] = fma(ww, rhs[q] - cg_conv<VAR>(a, sf, u, q), u[q]);
            }
            __syncthreads();
          }
        }
      }
      const double nextRes = resnorm();
      if (nextRes <= a.rel_tol * initRes) break;
    }
  }
  if (tid == 0) *iters_out = iters;
}

int launch_smoother_solve(examg_ctx *c, const Stencil &s, double *u, const double *rhs, double *res, double *other, int kind,
                          double omega, int max_it, double rel_tol) {
  const LevelGeom &g = *geom(c, s.level);
  if (c->world > 1)
    return set_error(EXAMG_EUNSUP, "examg_smoother_coarse: the coarse solve runs on the agglomerated (single-block) grid");
  if (c->periodic_any) return set_error(EXAMG_EUNSUP, "examg_smoother_coarse: periodic dimensions are not covered");
  const long long total = (long long)(g.hi[0] - g.lo[0]) * (g.hi[1] - g.lo[1]) * (g.hi[2] - g.lo[2]);
  if (total > (1LL << 21)) return set_error(EXAMG_EUNSUP, "examg_smoother_coarse: coarsest level has %lld unknowns; lower minLevel", total);
  CgArgs a;
  a.g = grid_args(g);
  a.st = stencil_args(g, s);
  a.sfvol = g.vol;
  a.max_it = max_it;
  a.rel_tol = rel_tol;
  for (int d = 0; d < 3; ++d) a.per[d] = 0;
  if (total <= 0) {
    EXAMG_CUDA(cudaMemsetAsync(c->int_out, 0, sizeof(int), c->stream));
    return EXAMG_OK;
  }
  int nt = 32;
  while (nt < 1024 && nt < total) nt <<= 1;
  if (s.coeff_field >= 0) {
    const double *sf = field_ptr(c, &c->fields[s.coeff_field], 0, 0);
    k_smoother_solve<true><<<1, nt, 0, c->stream>>>(a, sf, u, rhs, res, other, kind, omega, c->int_out);
  } else {
    k_smoother_solve<false><<<1, nt, 0, c->stream>>>(a, nullptr, u, rhs, res, other, kind, omega / s.coef[s.centre], c->int_out);
  }
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

int launch_cr(examg_ctx *c, const Stencil &s, double *u, const double *rhs, double *res, double *const tmp[3], int max_it,
              double rel_tol) {
  const LevelGeom &g = *geom(c, s.level);
  if (c->world > 1)
    return set_error(EXAMG_EUNSUP, "examg_cr_coarse: the coarse solve runs on the agglomerated (single-block) grid");
  const long long total = (long long)(g.hi[0] - g.lo[0]) * (g.hi[1] - g.lo[1]) * (g.hi[2] - g.lo[2]);
  if (total > (1LL << 21)) return set_error(EXAMG_EUNSUP, "examg_cr_coarse: coarsest level has %lld unknowns; lower minLevel", total);
  CgArgs a;
  a.g = grid_args(g);
  a.st = stencil_args(g, s);
  a.sfvol = g.vol;
  a.max_it = max_it;
  a.rel_tol = rel_tol;
  for (int d = 0; d < 3; ++d) a.per[d] = c->per[d] ? 1 : 0;
  if (total <= 0) {
    EXAMG_CUDA(cudaMemsetAsync(c->int_out, 0, sizeof(int), c->stream));
    return EXAMG_OK;
  }
  int nt = 32;
  while (nt < 1024 && nt < total) nt <<= 1;
  if (s.coeff_field >= 0) {
    const double *sf = field_ptr(c, &c->fields[s.coeff_field], 0, 0);
    k_cr<true><<<1, nt, 0, c->stream>>>(a, sf, u, rhs, res, tmp[0], tmp[1], tmp[2], c->int_out);
  } else {
    k_cr<false><<<1, nt, 0, c->stream>>>(a, nullptr, u, rhs, res, tmp[0], tmp[1], tmp[2], c->int_out);
  }
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

template <bool VAR>
__global__ void __launch_bounds__(1024)
k_minres(const __grid_constant__ CgArgs a, const double *__restrict__ sf, double *u, const double *__restrict__ rhs,
         double *res, double *w0, double *w1, double *w2, double *w3, double *w4, double *w5, double *av, int *iters_out) {
  const int iters = minres_device<VAR>(a, sf, u, rhs, res, w0, w1, w2, w3, w4, w5, av);
  if (threadIdx.x == 0) *iters_out = iters;
}

int launch_minres(examg_ctx *c, const Stencil &s, double *u, const double *rhs, double *res, double *const tmp[7],
                  int max_it, double rel_tol) {
  const LevelGeom &g = *geom(c, s.level);
  if (c->world > 1)
    return set_error(EXAMG_EUNSUP, "examg_minres_coarse: the coarse solve runs on the agglomerated (single-block) grid");
  const long long total = (long long)(g.hi[0] - g.lo[0]) * (g.hi[1] - g.lo[1]) * (g.hi[2] - g.lo[2]);
  if (total > (1LL << 21)) return set_error(EXAMG_EUNSUP, "examg_minres_coarse: coarsest level has %lld unknowns; lower minLevel", total);
  CgArgs a;
  a.g = grid_args(g);
  a.st = stencil_args(g, s);
  a.sfvol = g.vol;
  a.max_it = max_it;
  a.rel_tol = rel_tol;
  for (int d = 0; d < 3; ++d) a.per[d] = c->per[d] ? 1 : 0;
  if (total <= 0) {
    EXAMG_CUDA(cudaMemsetAsync(c->int_out, 0, sizeof(int), c->stream));
    return EXAMG_OK;
  }
  int nt = 32;
  while (nt < 1024 && nt < total) nt <<= 1;
  if (s.coeff_field >= 0) {
    const double *sf = field_ptr(c, &c->fields[s.coeff_field], 0, 0);
    k_minres<true><<<1, nt, 0, c->stream>>>(a, sf, u, rhs, res, tmp[0], tmp[1], tmp[2], tmp[3], tmp[4], tmp[5], tmp[6], c->int_out);
  } else {
    k_minres<false><<<1, nt, 0, c->stream>>>(a, nullptr, u, rhs, res, tmp[0], tmp[1], tmp[2], tmp[3], tmp[4], tmp[5], tmp[6], c->int_out);
  }
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

int launch_bicgstab(examg_ctx *c, const Stencil &s, double *u, const double *rhs, double *res, double *const tmp[5],
                    int max_it, double rel_tol) {
  const LevelGeom &g = *geom(c, s.level);
  if (c->world > 1)
    return set_error(EXAMG_EUNSUP, "examg_bicgstab_coarse: the coarse solve runs on the agglomerated (single-block) grid");
  const long long total = (long long)(g.hi[0] - g.lo[0]) * (g.hi[1] - g.lo[1]) * (g.hi[2] - g.lo[2]);
  if (total > (1LL << 21))
    return set_error(EXAMG_EUNSUP, "examg_bicgstab_coarse: coarsest level has %lld unknowns; lower minLevel", total);
  CgArgs a;
  a.g = grid_args(g);
  a.st = stencil_args(g, s);
  a.sfvol = g.vol;
  a.max_it = max_it;
  a.rel_tol = rel_tol;
  for (int d = 0; d < 3; ++d) a.per[d] = c->per[d] ? 1 : 0;
  if (total <= 0) {
    EXAMG_CUDA(cudaMemsetAsync(c->int_out, 0, sizeof(int), c->stream));
    return EXAMG_OK;
  }
  int nt = 32;
  while (nt < 1024 && nt < total) nt <<= 1;
  if (s.coeff_field >= 0) {
    const double *sf = field_ptr(c, &c->fields[s.coeff_field], 0, 0);
    k_bicgstab<true><<<1, nt, 0, c->stream>>>(a, sf, u, rhs, res, tmp[0], tmp[1], tmp[2], tmp[3], tmp[4], c->int_out);
  } else {
    k_bicgstab<false><<<1, nt, 0, c->stream>>>(a, nullptr, u, rhs, res, tmp[0], tmp[1], tmp[2], tmp[3], tmp[4], c->int_out);
  }
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

int launch_cg(examg_ctx *c, const Stencil &s, double *u, const double *rhs, double *res, double *p, double *ap,
              int max_it, double rel_tol) {
  const LevelGeom &g = *geom(c, s.level);
  if (c->world > 1)
    return set_error(EXAMG_EUNSUP, "examg_cg_coarse: the coarse solve runs on the agglomerated (single-block) grid");
  const long long total = (long long)(g.hi[0] - g.lo[0]) * (g.hi[1] - g.lo[1]) * (g.hi[2] - g.lo[2]);
  if (total > (1LL << 21))
    return set_error(EXAMG_EUNSUP, "examg_cg_coarse: coarsest level has %lld unknowns; lower minLevel", total);
  CgArgs a;
  a.g = grid_args(g);
  a.st = stencil_args(g, s);
  a.sfvol = g.vol;
  a.max_it = max_it;
  a.rel_tol = rel_tol;
  for (int d = 0; d < 3; ++d) a.per[d] = c->per[d] ? 1 : 0;
  if (total <= 0) {
    EXAMG_CUDA(cudaMemsetAsync(c->int_out, 0, sizeof(int), c->stream));
    return EXAMG_OK;
  }
  int nt = 32;
  while (nt < 1024 && nt < total) nt <<= 1;
  if (s.coeff_field >= 0) {
    const double *sf = field_ptr(c, &c->fields[s.coeff_field], 0, 0);
    k_cg<true><<<1, nt, 0, c->stream>>>(a, sf, u, rhs, res, p, ap, c->int_out);
  } else {
    k_cg<false><<<1, nt, 0, c->stream>>>(a, nullptr, u, rhs, res, p, ap, c->int_out);
  }
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

}  // namespace examg
