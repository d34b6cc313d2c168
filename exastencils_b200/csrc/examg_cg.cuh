// examg_cg.cuh -- un-preconditioned CG as a block-wide device function, shared by the
// stand-alone coarse solve (examg_cg.cu) and the single-CTA coarse V-cycle (examg_coarse.cu).
#pragma once

#include "examg_device.cuh"

namespace examg {

struct CgArgs {
  GridArgs g;
  StencilArgs st;
  long long sfvol;
  int max_it;
  double rel_tol;
  int per[3];  // periodic dimensions: the block is its own neighbour (ghost layers refreshed inside the kernel)
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

// loop point t counts in reductions (a point shared by the two ends of a periodic dimension counts once)
__device__ __forceinline__ bool cg_counts(const GridArgs &g, int t, int ex, int ey) {
  const int i = g.lo[0] + t % ex;
  const int j = g.lo[1] + (t / ex) % ey;
  const int k = g.lo[2] + t / (ex * ey);
  return i >= g.rlo[0] && j >= g.rlo[1] && k >= g.rlo[2];
}

// `communicate x` of a single periodic block: duplicate + ghost layer on the low side from the high end, ghost layer
// on the high side from the low end, one dimension after the other over the orthogonal extents incl. ghost layers
__device__ __forceinline__ void cg_refresh(const CgArgs &a, double *x) {
  const GridArgs &g = a.g;
  __syncthreads();
  for (int d = 0; d < 3; ++d) {
    if (!a.per[d]) continue;
    const int da = d == 0 ? 1 : 0, db = d == 2 ? 1 : 2;
    const int ga = g.n[da] > 1 ? 1 : 0, gb = g.n[db] > 1 ? 1 : 0;
    const int ea = g.n[da] + 2 * ga, eb = g.n[db] + 2 * gb;
    const int total = 3 * ea * eb;
    for (int t = threadIdx.x; t < total; t += blockDim.x) {
      const int l = t % 3, ia = (t / 3) % ea - ga, ib = t / (3 * ea) - gb;
      int idx[3], src[3];
      idx[da] = src[da] = ia;
      idx[db] = src[db] = ib;
      idx[d] = l == 0 ? -1 : (l == 1 ? 0 : g.n[d]);
      src[d] = l == 0 ? g.n[d] - 2 : (l == 1 ? g.n[d] - 1 : 1);
      x[idx[0] + (long long)g.pitch * idx[1] + g.plane * idx[2]] = x[src[0] + (long long)g.pitch * src[1] + g.plane * src[2]];
    }
    __syncthreads();
  }
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

// All threads of the block must call it; returns the number of iterations (same value in every thread).
template <bool VAR>
__device__ __forceinline__ int cg_device(const CgArgs &a, const double *__restrict__ sf, double *u, const double *rhs,
                                         double *res, double *p, double *ap) {
  const GridArgs &g = a.g;
  const int ex = g.hi[0] - g.lo[0], ey = g.hi[1] - g.lo[1], ez = g.hi[2] - g.lo[2];
  const int total = ex * ey * ez;
  const int tid = threadIdx.x, nt = blockDim.x;
  const bool per = a.per[0] || a.per[1] || a.per[2];

  // Residual = RHS - A*Solution; curRes = ||Residual||; cgTmp0 = Residual
  if (per) cg_refresh(a, u);
  dd part = {0.0, 0.0};
  for (int t = tid; t < total; t += nt) {
    const long long q = cg_point(g, t, ex, ey);
    const double r = rhs[q] - cg_conv<VAR>(a, sf, u, q);
    res[q] = r;
    p[q] = r;
    dd_add(part, !per || cg_counts(g, t, ex, ey) ? r * r : 0.0);
  }
  if (per) cg_refresh(a, p);
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
      const bool cnt = !per || cg_counts(g, t, ex, ey);
      dd_add(pn, cnt ? r * r : 0.0);
      dd_add(pd, cnt ? p[q] * v : 0.0);
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
      dd_add(pr, !per || cg_counts(g, t, ex, ey) ? r * r : 0.0);
    }
    const double nextRes = sqrt(cg_sum(pr));
    if (nextRes <= a.rel_tol * initRes) break;
    const double beta = (nextRes * nextRes) / (curRes * curRes);
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      p[q] = fma(beta, p[q], res[q]);
    }
    curRes = nextRes;
    if (per) cg_refresh(a, p);
    __syncthreads();
  }
  return iters;
}

// BiCGStab of the layer-3 generator (solver_cgs = "BiCGStab", C/s/e/solver/l3/L3_BiCGStabForEquation.scala:36-198):
// resHat = r0; rho = rhoOld = alpha = beta = omega = 1; v = p = 0; per step
//   rho = <resHat, r>; beta = (rho/rhoOld)*(alpha/omega); p = r + beta*(p - omega*v); v = A p; alpha = rho/<resHat, v>;
//   s = r - alpha*v; t = A s; omega = <t,s>/<t,t>; u += alpha*p + omega*s; r = s - omega*t;
//   stop at ||r|| <= rel_tol * ||r0||.
// Association: p = fma(beta, fma(-omega, v, p), r); s = fma(-alpha, v, r); u = fma(omega, s, fma(alpha, p, u));
// r = fma(-omega, t, s); sums in double-double.  All threads of the block must call it; returns the step count.
template <bool VAR>
__device__ __forceinline__ int bicgstab_device(const CgArgs &a, const double *__restrict__ sf, double *u, const double *rhs,
                                               double *res, double *v, double *p, double *sv, double *tv, double *rhat) {
  const GridArgs &g = a.g;
  const int ex = g.hi[0] - g.lo[0], ey = g.hi[1] - g.lo[1], ez = g.hi[2] - g.lo[2];
  const int total = ex * ey * ez;
  const int tid = threadIdx.x, nt = blockDim.x;
  const bool per = a.per[0] || a.per[1] || a.per[2];
  if (per) cg_refresh(a, u);
  dd part = {0.0, 0.0};
  for (int t = tid; t < total; t += nt) {
    const long long q = cg_point(g, t, ex, ey);
    const double r = rhs[q] - cg_conv<VAR>(a, sf, u, q);
    res[q] = r;
    rhat[q] = r;
    v[q] = 0.0;
    p[q] = 0.0;
    dd_add(part, !per || cg_counts(g, t, ex, ey) ? r * r : 0.0);
  }
  double curRes = sqrt(cg_sum(part));
  const double initRes = curRes;
  if (curRes <= 0.0) return 0;  // solver_cgs_absResThreshold = 0
  double rho = 1.0, rhoOld = 1.0, alpha = 1.0, omega = 1.0;
  int iters = 0;
  for (int step = 0; step < a.max_it; ++step) {
    iters = step + 1;
    rhoOld = rho;
    dd pr = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      dd_add(pr, !per || cg_counts(g, t, ex, ey) ? rhat[q] * res[q] : 0.0);
    }
    rho = cg_sum(pr);
    const double beta = (rho / rhoOld) * (alpha / omega);
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      p[q] = fma(beta, fma(-omega, v[q], p[q]), res[q]);
    }
    if (per) cg_refresh(a, p);
    __syncthreads();
    dd pa = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      const double w = cg_conv<VAR>(a, sf, p, q);
      v[q] = w;
      dd_add(pa, !per || cg_counts(g, t, ex, ey) ? rhat[q] * w : 0.0);
    }
    alpha = rho / cg_sum(pa);
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      sv[q] = fma(-alpha, v[q], res[q]);
    }
    if (per) cg_refresh(a, sv);
    __syncthreads();
    dd pn = {0.0, 0.0}, pd = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      const double w = cg_conv<VAR>(a, sf, sv, q);
      tv[q] = w;
      const bool cnt = !per || cg_counts(g, t, ex, ey);
      dd_add(pn, cnt ? w * sv[q] : 0.0);
      dd_add(pd, cnt ? w * w : 0.0);
    }
    const double omegaNom = cg_sum(pn);
    const double omegaDenom = cg_sum(pd);
    omega = omegaNom / omegaDenom;
    dd pq = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      u[q] = fma(omega, sv[q], fma(alpha, p[q], u[q]));
      const double r = fma(-omega, tv[q], sv[q]);
      res[q] = r;
      dd_add(pq, !per || cg_counts(g, t, ex, ey) ? r * r : 0.0);
    }
    curRes = sqrt(cg_sum(pq));
    if (curRes <= a.rel_tol * initRes) break;
    __syncthreads();
  }
  return iters;
}

// Conjugate residuals of the layer-3 generator (solver_cgs = "ConjugateResidual",
// C/s/e/solver/l3/L3_ConjugateResidualForEquation.scala:36-170): p = r; ap = A p; ar = A r; per step
//   alpha = <r, ar> / <ap, ap>; u += alpha p; r -= alpha ap; stop at ||r|| <= rel_tol ||r0||;
//   ar = A r; beta = <r, ar> / <r, ar>_old; p = r + beta p; ap = ar + beta ap.
// Association: u = fma(alpha, p, u); r = fma(-alpha, ap, r); p = fma(beta, p, r); ap = fma(beta, ap, ar).
template <bool VAR>
__device__ __forceinline__ int cr_device(const CgArgs &a, const double *__restrict__ sf, double *u, const double *rhs,
                                         double *res, double *p, double *ap, double *ar) {
  const GridArgs &g = a.g;
  const int ex = g.hi[0] - g.lo[0], ey = g.hi[1] - g.lo[1], ez = g.hi[2] - g.lo[2];
  const int total = ex * ey * ez;
  const int tid = threadIdx.x, nt = blockDim.x;
  const bool per = a.per[0] || a.per[1] || a.per[2];
  if (per) cg_refresh(a, u);
  dd part = {0.0, 0.0};
  for (int t = tid; t < total; t += nt) {
    const long long q = cg_point(g, t, ex, ey);
    const double r = rhs[q] - cg_conv<VAR>(a, sf, u, q);
    res[q] = r;
    p[q] = r;
    dd_add(part, !per || cg_counts(g, t, ex, ey) ? r * r : 0.0);
  }
  double curRes = sqrt(cg_sum(part));
  const double initRes = curRes;
  if (curRes <= 0.0) return 0;
  if (per) cg_refresh(a, p);
  if (per) cg_refresh(a, res);
  __syncthreads();
  for (int t = tid; t < total; t += nt) {
    const long long q = cg_point(g, t, ex, ey);
    const double w = cg_conv<VAR>(a, sf, p, q);
    ap[q] = w;
    ar[q] = w;  // p == r
  }
  __syncthreads();
  int iters = 0;
  for (int step = 0; step < a.max_it; ++step) {
    iters = step + 1;
    dd pn = {0.0, 0.0}, pd = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      const bool cnt = !per || cg_counts(g, t, ex, ey);
      dd_add(pn, cnt ? res[q] * ar[q] : 0.0);
      dd_add(pd, cnt ? ap[q] * ap[q] : 0.0);
    }
    const double alphaNom = cg_sum(pn);
    const double alphaDenom = cg_sum(pd);
    const double alpha = alphaNom / alphaDenom;
    dd pr = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      u[q] = fma(alpha, p[q], u[q]);
      const double r = fma(-alpha, ap[q], res[q]);
      res[q] = r;
      dd_add(pr, !per || cg_counts(g, t, ex, ey) ? r * r : 0.0);
    }
    const double nextRes = sqrt(cg_sum(pr));
    if (nextRes <= a.rel_tol * initRes) break;
    if (per) cg_refresh(a, res);
    __syncthreads();
    dd pb = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      const double w = cg_conv<VAR>(a, sf, res, q);
      ar[q] = w;
      dd_add(pb, !per || cg_counts(g, t, ex, ey) ? res[q] * w : 0.0);
    }
    const double beta = cg_sum(pb) / alphaNom;
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      p[q] = fma(beta, p[q], res[q]);
      ap[q] = fma(beta, ap[q], ar[q]);
    }
    curRes = nextRes;
    __syncthreads();
  }
  return iters;
}

// MINRES of the layer-3 generator (solver_cgs = "MinRes", C/s/e/solver/l3/L3_MinResForEquation.scala:36-230): Lanczos
// vectors vOld, v, vNew, search directions pOld, p, pNew (buffers rotated instead of copied), Givens rotations in scalars,
// the residual norm from the recurrence curRes = -sNew * curRes.  Association: vNew = fma(-beta, vOld, A v);
// vNew = fma(-alpha, v, vNew); vNew /= betaNew; pNew = fma(-rho2, p, fma(-rho1, pOld, v)) / rho3;
// u = fma(cNew * curRes, pNew, u); scalars exactly as written below.  w = { v0, v1, v2, p0, p1, p2, av }.
template <bool VAR>
__device__ __forceinline__ int minres_device(const CgArgs &a, const double *__restrict__ sf, double *u, const double *rhs,
                                             double *res, double *w0, double *w1, double *w2, double *w3, double *w4,
                                             double *w5, double *av) {
  const GridArgs &g = a.g;
  const int ex = g.hi[0] - g.lo[0], ey = g.hi[1] - g.lo[1], ez = g.hi[2] - g.lo[2];
  const int total = ex * ey * ez;
  const int tid = threadIdx.x, nt = blockDim.x;
  const bool per = a.per[0] || a.per[1] || a.per[2];
  double *vOld = w0, *v = w1, *vNew = w2, *pOld = w3, *pp = w4, *pNew = w5;
  if (per) cg_refresh(a, u);
  dd part = {0.0, 0.0};
  for (int t = tid; t < total; t += nt) {
    const long long q = cg_point(g, t, ex, ey);
    const double r = rhs[q] - cg_conv<VAR>(a, sf, u, q);
    res[q] = r;
    dd_add(part, !per || cg_counts(g, t, ex, ey) ? r * r : 0.0);
  }
  double curRes = sqrt(cg_sum(part));
  const double initRes = curRes;
  if (curRes <= 0.0) return 0;
  for (int t = tid; t < total; t += nt) {
    const long long q = cg_point(g, t, ex, ey);
    pp[q] = 0.0;
    pNew[q] = 0.0;
    v[q] = 0.0;
    vNew[q] = res[q] / initRes;
  }
  double beta = 0.0, betaNew = 0.0, cOld = 1.0, c = 1.0, cNew = 1.0, sOld = 0.0, sn = 0.0, sNew = 0.0;
  int iters = 0;
  for (int step = 0; step < a.max_it; ++step) {
    iters = step + 1;
    beta = betaNew;
    { double *tp = vOld; vOld = v; v = vNew; vNew = tp; }  // vOld = v; v = vNew
    if (per) cg_refresh(a, v);
    __syncthreads();
    dd pa = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      const double avq = cg_conv<VAR>(a, sf, v, q);
      av[q] = avq;
      const double vn = fma(-beta, vOld[q], avq);
      vNew[q] = vn;
      dd_add(pa, !per || cg_counts(g, t, ex, ey) ? vn * v[q] : 0.0);
    }
    const double alpha = cg_sum(pa);
    dd pb = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      const double vn = fma(-alpha, v[q], vNew[q]);
      vNew[q] = vn;
      dd_add(pb, !per || cg_counts(g, t, ex, ey) ? vn * vn : 0.0);
    }
    betaNew = sqrt(cg_sum(pb));
    cOld = c;
    c = cNew;
    sOld = sn;
    sn = sNew;
    const double rho1 = sOld * beta;
    const double rho2 = c * cOld * beta + sn * alpha;
    const double rho3Tilde = c * alpha - sn * cOld * beta;
    const double tau = fabs(rho3Tilde) + fabs(betaNew);
    const double nu = tau * sqrt((rho3Tilde / tau) * (rho3Tilde / tau) + (betaNew / tau) * (betaNew / tau));
    cNew = rho3Tilde / nu;
    sNew = betaNew / nu;
    const double rho3 = nu;
    { double *tp = pOld; pOld = pp; pp = pNew; pNew = tp; }  // pOld = p; p = pNew
    const double scale = cNew * curRes;
    for (int t = tid; t < total; t += nt) {
      const long long q = cg_point(g, t, ex, ey);
      vNew[q] = vNew[q] / betaNew;
      const double pn = fma(-rho2, pp[q], fma(-rho1, pOld[q], v[q])) / rho3;
      pNew[q] = pn;
      u[q] = fma(scale, pn, u[q]);
    }
    curRes = -1.0 * sNew * curRes;
    if (fabs(curRes) <= a.rel_tol * initRes) break;
    __syncthreads();
  }
  return iters;
}

}  // namespace examg
