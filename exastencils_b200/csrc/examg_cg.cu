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
#include "examg_device.cuh"

namespace examg {

struct CgArgs {
  GridArgs g;
  StencilArgs st;
  long long sfvol;
  int max_it;
  double rel_tol;
};

template <bool VAR>
__device__ __forceinline__ double cg_conv(const CgArgs &a, const double *__restrict__ sf, const double *u,
                                          long long p) {
  double acc;
  if (VAR) {
    acc = sf[p] * u[p + a.st.lin[0]];
    for (int k = 1; k < a.st.nst; ++k) acc = fma(sf[k * a.sfvol + p], u[p + a.st.lin[k]], acc);
  } else {
    acc = a.st.coef[0] * u[p + a.st.lin[0]];
    for (int k = 1; k < a.st.nst; ++k) acc = fma(a.st.coef[k], u[p + a.st.lin[k]], acc);
  }
  return acc;
}

__device__ __forceinline__ long long cg_point(const GridArgs &g, int t, int ex, int ey) {
  const int i = g.lo[0] + t % ex;
  const int j = g.lo[1] + (t / ex) % ey;
  const int k = g.lo[2] + t / (ex * ey);
  return i + (long long)g.pitch * j + g.plane * k;
}

// block-wide sum, result broadcast to all threads
__device__ __forceinline__ double cg_sum(dd part) {
  __shared__ double bc;
  part = dd_block_reduce(part);
  if (threadIdx.x == 0) bc = part.hi + part.lo;
  __syncthreads();
  const double r = bc;
  __syncthreads();
  return r;
}

template <bool VAR>
__global__ void __launch_bounds__(1024)
k_cg(const __grid_constant__ CgArgs a, const double *__restrict__ sf, double *u, const double *__restrict__ rhs,
     double *res, double *p, double *ap, int *iters_out) {
  const GridArgs &g = a.g;
  const int ex = g.hi[0] - g.lo[0], ey = g.hi[1] - g.lo[1], ez = g.hi[2] - g.lo[2];
  const int total = ex * ey * ez;
  const int tid = threadIdx.x, nt = blockDim.x;

  // Residual = RHS - A*Solution; curRes = ||Residual||; cgTmp0 = Residual
  dd part = {0.0, 0.0};
  for (int t = tid; t < total; t += nt) {
    const long long q = cg_point(g, t, ex, ey);
    const double r = rhs[q] - cg_conv<VAR>(a, sf, u, q);
    res[q] = r;
    p[q] = r;
    dd_add(part, r * r);
  }
  double curRes = sqrt(cg_sum(part));
  const double initRes = curRes;
  int iters = 0;
  for (int step = 0; step < a.max_it; ++step) {
    iters = step + 1;
    // cgTmp1 = A * cgTmp0 ; alphaNom = <r,r> ; alphaDenom = <p, Ap>
    dd pn = {0.0, 0.0}, pd = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      const double v = cg_conv<VAR>(a, sf, p, q);
      ap[q] = v;
      const double r = res[q];
      dd_add(pn, r * r);
      dd_add(pd, p[q] * v);
    }
    const double alphaNom = cg_sum(pn);
    const double alphaDenom = cg_sum(pd);
    const double alpha = alphaNom / alphaDenom;
    // Solution += alpha*p ; Residual -= alpha*Ap ; nextRes = ||Residual||
    dd pr = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      u[q] = fma(alpha, p[q], u[q]);
      const double r = fma(-alpha, ap[q], res[q]);
      res[q] = r;
      dd_add(pr, r * r);
    }
    const double nextRes = sqrt(cg_sum(pr));
    if (nextRes <= a.rel_tol * initRes) break;
    const double beta = (nextRes * nextRes) / (curRes * curRes);
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      p[q] = fma(beta, p[q], res[q]);
    }
    curRes = nextRes;
    __syncthreads();
  }
  if (tid == 0) *iters_out = iters;
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
