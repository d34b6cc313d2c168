/*
 * examg_oracle_optflow.c -- CPU restatement of the ExaStencils optical-flow solver
 * (2D, cell-centred, two coupled unknowns u,v, red-black 2x2 block smoother).
 * TEST INFRASTRUCTURE ONLY (see examg_oracle.h).
 *
 * Follows (file:line under /root/reference):
 *   program text        Benchmark/OptFlow2D/2D_FD_OptFlow.exa4 (whole file; `solve locally`
 *                       form), Testing/Application/OpticalFlow2D.exa4 (Vec2 form, the one the
 *                       golden OpticalFlow2D.results was produced from; in exact arithmetic the
 *                       two smoothers are the same 2x2 block solve)
 *   cell layout         ghostLayers 1, duplicateLayers 0: Layout cellWComm / defLayoutForCell
 *   loop range          all cells, C/s/e/baseExt/ir/IR_LoopOverPointsInOneFragment.scala:86-101
 *   Neumann (order 1)   ghost cell = adjacent inner cell, faces in the order -x,+x,-y,+y over the
 *                       full orthogonal extent incl. ghosts:
 *                       C/s/e/boundary/ir/IR_NeumannBC.scala:70-80,
 *                       C/s/e/boundary/ir/IR_ApplyBCFunction.scala:53-85
 *   local solve         C/s/e/solver/ir/IR_LocalSolve.scala:88-236 (unknowns at the cell -> matrix,
 *                       everything else -> right-hand side), IR_LocalDirectInvert.scala:44-140
 *                       (write back u*(1-relax) + relax*u_new), 2x2 system by the explicit inverse
 *                       (C/s/e/solver/ir/IR_SolveMatrixSystem.scala compile-time branch ->
 *                       IR_CompiletimeMatOps.scala:796-803)
 *   cell restriction    [1/2,1/2]^d, prolongation = transpose * 2^d = piecewise constant:
 *                       C/s/e/operator/l4/L4_DefaultRestriction.scala:40-47,
 *                       L4_DefaultProlongation.scala:30-45
 *   grid                cell centre (i + 1/2) h, h = 1 / (F 2^level)
 *                       C/s/e/domain/ir/IR_DomainFromAABB.scala:31-40
 *
 * Dense layout of every array: (ny+2) x (nx+2) doubles, x fastest, one ghost cell per side;
 * cell (i,j) is element (i+1) + (nx+2)*(j+1).
 *
 * Arithmetic contract (shared with the CUDA kernels, csrc/examg_optflow.cu):
 *   L(w)   = Lc*w_c, then fma(Lw,w_w,.), fma(Le,w_e,.), fma(Ls,w_s,.), fma(Ln,w_n,.)   [declaration order]
 *   A_u    = fma(a2, L(u), fma(cuv, v, cuu*u))        residual_u = rhs_u - A_u   (v alike with cuv,cvv)
 *   smoother at one cell:
 *     o(w)  = Lw*w_w, then fma(Le,..), fma(Ls,..), fma(Ln,..)       (off-centre part)
 *     f0 = rhs_u - a2*o(u); f1 = rhs_v - a2*o(v); A00 = fma(a2,Lc,cuu); A11 = fma(a2,Lc,cvv); A01 = A10 = cuv
 *     the 2x2 inverse as the generator's compile-time branch writes it
 *     (C/s/e/baseExt/ir/IR_MatOperations/IR_CompiletimeMatOps.scala:796-803):
 *     rdet = 1/(A00*A11 - A01*A10); I00 = rdet*A11; I01 = I10 = -(rdet*A01); I11 = rdet*A00
 *     nu = I00*f0 + I01*f1;  nv = I10*f0 + I11*f1          (products rounded, then added)
 *     u = u*(1-relax) + relax*nu;  v alike
 *   sums: double-double, as in examg_oracle.c
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "examg_oracle.h"

typedef struct { double hi, lo; } ogf_dd;
static inline void ogf_dd_add(ogf_dd *a, double x) {
  double s = a->hi + x;
  double bb = s - a->hi;
  double err = (a->hi - (s - bb)) + (x - bb);
  a->hi = s;
  a->lo += err;
}
static inline void ogf_dd_merge(ogf_dd *a, const ogf_dd *b) {
  ogf_dd_add(a, b->hi);
  a->lo += b->lo;
}

#define IDX(i, j) ((size_t)((i) + 1) + (size_t)(nx + 2) * (size_t)((j) + 1))

void ogf_neumann(int nx, int ny, double *a) {
  for (int j = -1; j <= ny; ++j) a[IDX(-1, j)] = a[IDX(0, j)];
  for (int j = -1; j <= ny; ++j) a[IDX(nx, j)] = a[IDX(nx - 1, j)];
  for (int i = -1; i <= nx; ++i) a[IDX(i, -1)] = a[IDX(i, 0)];
  for (int i = -1; i <= nx; ++i) a[IDX(i, ny)] = a[IDX(i, ny - 1)];
}

static inline double ogf_lap(const double L[5], const double *w, size_t p, size_t sy) {
  double acc = L[0] * w[p];
  acc = fma(L[1], w[p - 1], acc);
  acc = fma(L[2], w[p + 1], acc);
  acc = fma(L[3], w[p - sy], acc);
  acc = fma(L[4], w[p + sy], acc);
  return acc;
}

static inline double ogf_off(const double L[5], const double *w, size_t p, size_t sy) {
  double acc = L[1] * w[p - 1];
  acc = fma(L[2], w[p + 1], acc);
  acc = fma(L[3], w[p - sy], acc);
  acc = fma(L[4], w[p + sy], acc);
  return acc;
}

/* out_u = A_u(u,v), out_v = A_v(u,v) on all cells */
void ogf_apply(int nx, int ny, const double L[5], double a2, const double *cuu, const double *cuv,
               const double *cvv, const double *u, const double *v, double *ou, double *ov) {
  const size_t sy = (size_t)nx + 2;
#pragma omp parallel for schedule(static)
  for (int j = 0; j < ny; ++j)
    for (int i = 0; i < nx; ++i) {
      const size_t p = IDX(i, j);
      ou[p] = fma(a2, ogf_lap(L, u, p, sy), fma(cuv[p], v[p], cuu[p] * u[p]));
      ov[p] = fma(a2, ogf_lap(L, v, p, sy), fma(cvv[p], v[p], cuv[p] * u[p]));
    }
}

void ogf_residual(int nx, int ny, const double L[5], double a2, const double *cuu, const double *cuv,
                  const double *cvv, const double *u, const double *v, const double *fu, const double *fv,
                  double *ru, double *rv) {
  const size_t sy = (size_t)nx + 2;
#pragma omp parallel for schedule(static)
  for (int j = 0; j < ny; ++j)
    for (int i = 0; i < nx; ++i) {
      const size_t p = IDX(i, j);
      ru[p] = fu[p] - fma(a2, ogf_lap(L, u, p, sy), fma(cuv[p], v[p], cuu[p] * u[p]));
      rv[p] = fv[p] - fma(a2, ogf_lap(L, v, p, sy), fma(cvv[p], v[p], cuv[p] * u[p]));
    }
}

/* one colour of the red-black block smoother: cells with (i+j)%2 == color */
void ogf_smooth_color(int nx, int ny, const double L[5], double a2, const double *cuu, const double *cuv,
                      const double *cvv, double *u, double *v, const double *fu, const double *fv,
                      double relax, int color) {
  const size_t sy = (size_t)nx + 2;
  const double keep = 1.0 - relax;
#pragma omp parallel for schedule(static)
  for (int j = 0; j < ny; ++j)
    for (int i = (j + color) & 1; i < nx; i += 2) {
      const size_t p = IDX(i, j);
      const double f0 = fu[p] - a2 * ogf_off(L, u, p, sy);
      const double f1 = fv[p] - a2 * ogf_off(L, v, p, sy);
      const double a00 = fma(a2, L[0], cuu[p]), a11 = fma(a2, L[0], cvv[p]), a01 = cuv[p];
      const double rdet = 1.0 / (a00 * a11 - a01 * a01);
      const double i00 = rdet * a11, i01 = -(rdet * a01), i11 = rdet * a00;
      const double nu = i00 * f0 + i01 * f1;
      const double nv = i01 * f0 + i11 * f1;
      u[p] = u[p] * keep + relax * nu;
      v[p] = v[p] * keep + relax * nv;
    }
}

/* coarse = RestrictionCell * fine: mean of the four children (lexicographic sum) */
void ogf_restrict(int nxc, int nyc, const double *fine, double *coarse) {
  const int nxf = 2 * nxc;
#pragma omp parallel for schedule(static)
  for (int j = 0; j < nyc; ++j)
    for (int i = 0; i < nxc; ++i) {
      const size_t q = (size_t)(2 * i + 1) + (size_t)(nxf + 2) * (size_t)(2 * j + 1);
      const double s = ((fine[q] + fine[q + 1]) + fine[q + nxf + 2]) + fine[q + nxf + 3];
      coarse[(size_t)(i + 1) + (size_t)(nxc + 2) * (size_t)(j + 1)] = 0.25 * s;
    }
}

/* fine += ProlongationCell * coarse: every child gets its parent's value */
void ogf_prolong_add(int nxc, int nyc, const double *coarse, double *fine) {
  const int nxf = 2 * nxc;
#pragma omp parallel for schedule(static)
  for (int j = 0; j < 2 * nyc; ++j)
    for (int i = 0; i < nxf; ++i)
      fine[(size_t)(i + 1) + (size_t)(nxf + 2) * (size_t)(j + 1)] +=
          coarse[(size_t)(i / 2 + 1) + (size_t)(nxc + 2) * (size_t)(j / 2 + 1)];
}

/* sum over all cells of a*b (+ c*d if c != NULL, accumulated a*b first then c*d per cell) */
double ogf_dot2(int nx, int ny, const double *a, const double *b, const double *c, const double *d) {
  ogf_dd total = {0.0, 0.0};
#pragma omp parallel
  {
    ogf_dd part = {0.0, 0.0};
#pragma omp for schedule(static) nowait
    for (int j = 0; j < ny; ++j)
      for (int i = 0; i < nx; ++i) {
        const size_t p = IDX(i, j);
        ogf_dd_add(&part, a[p] * b[p]);
        if (c) ogf_dd_add(&part, c[p] * d[p]);
      }
#pragma omp critical
    ogf_dd_merge(&total, &part);
  }
  return total.hi + total.lo;
}

/* ---- whole solver -------------------------------------------------------- */

typedef struct {
  int nx, ny;
  double h[2], L[5];
  double *u, *v, *fu, *fv, *ru, *rv, *cuu, *cuv, *cvv;
} ogf_level;

struct ogf_solver {
  ogf_config cfg;
  int nlev;
  ogf_level *lev;
  double *img0, *img1, *ix, *iy, *it; /* finest */
  double *pu, *pv, *apu, *apv;        /* coarsest: cgTmp0, cgTmp1 */
  double a2;
  int last_cg_iters;
};

static double *ogf_alloc(int nx, int ny) { return (double *)calloc((size_t)(nx + 2) * (size_t)(ny + 2), sizeof(double)); }

static ogf_level *ogf_lv(ogf_solver *s, int lvl) { return &s->lev[lvl - s->cfg.min_level]; }

ogf_solver *ogf_solver_create(const ogf_config *cfg) {
  ogf_solver *s = (ogf_solver *)calloc(1, sizeof(*s));
  s->cfg = *cfg;
  s->nlev = cfg->max_level - cfg->min_level + 1;
  s->lev = (ogf_level *)calloc((size_t)s->nlev, sizeof(ogf_level));
  s->a2 = cfg->alpha * cfg->alpha; /* alpha ** 2 */
  for (int lvl = cfg->min_level; lvl <= cfg->max_level; ++lvl) {
    ogf_level *l = ogf_lv(s, lvl);
    l->nx = cfg->F[0] << lvl;
    l->ny = cfg->F[1] << lvl;
    l->h[0] = 1.0 / (double)l->nx;
    l->h[1] = 1.0 / (double)l->ny;
    const double hx2 = l->h[0] * l->h[0], hy2 = l->h[1] * l->h[1];
    l->L[0] = 2.0 / hx2 + 2.0 / hy2; /* Stencil Laplace, declaration order */
    l->L[1] = -1.0 / hx2;
    l->L[2] = -1.0 / hx2;
    l->L[3] = -1.0 / hy2;
    l->L[4] = -1.0 / hy2;
    double **all[] = {&l->u, &l->v, &l->fu, &l->fv, &l->ru, &l->rv, &l->cuu, &l->cuv, &l->cvv};
    for (size_t k = 0; k < sizeof(all) / sizeof(all[0]); ++k) *all[k] = ogf_alloc(l->nx, l->ny);
  }
  ogf_level *f = ogf_lv(s, cfg->max_level), *c = ogf_lv(s, cfg->min_level);
  s->img0 = ogf_alloc(f->nx, f->ny);
  s->img1 = ogf_alloc(f->nx, f->ny);
  s->ix = ogf_alloc(f->nx, f->ny);
  s->iy = ogf_alloc(f->nx, f->ny);
  s->it = ogf_alloc(f->nx, f->ny);
  s->pu = ogf_alloc(c->nx, c->ny);
  s->pv = ogf_alloc(c->nx, c->ny);
  s->apu = ogf_alloc(c->nx, c->ny);
  s->apv = ogf_alloc(c->nx, c->ny);
  return s;
}

void ogf_solver_destroy(ogf_solver *s) {
  if (!s) return;
  for (int k = 0; k < s->nlev; ++k) {
    ogf_level *l = &s->lev[k];
    free(l->u); free(l->v); free(l->fu); free(l->fv); free(l->ru); free(l->rv);
    free(l->cuu); free(l->cuv); free(l->cvv);
  }
  free(s->img0); free(s->img1); free(s->ix); free(s->iy); free(s->it);
  free(s->pu); free(s->pv); free(s->apu); free(s->apv);
  free(s->lev);
  free(s);
}

void ogf_solver_dims(const ogf_solver *s, int lvl, int n[2]) {
  n[0] = s->lev[lvl - s->cfg.min_level].nx;
  n[1] = s->lev[lvl - s->cfg.min_level].ny;
}

double *ogf_solver_field(ogf_solver *s, int lvl, int which) {
  ogf_level *l = ogf_lv(s, lvl);
  switch (which) {
    case OGF_U: return l->u;
    case OGF_V: return l->v;
    case OGF_RHS_U: return l->fu;
    case OGF_RHS_V: return l->fv;
    case OGF_RES_U: return l->ru;
    case OGF_RES_V: return l->rv;
    case OGF_CUU: return l->cuu;
    case OGF_CUV: return l->cuv;
    case OGF_CVV: return l->cvv;
    case OGF_IMG0: return s->img0;
    case OGF_IMG1: return s->img1;
    default: return NULL;
  }
}

/* initImgs@finest with the program's analytic images (Benchmark/OptFlow2D/2D_FD_OptFlow.exa4 `initImgs`) */
void ogf_solver_init_images(ogf_solver *s) {
  ogf_level *l = ogf_lv(s, s->cfg.max_level);
  const int nx = l->nx, ny = l->ny;
  for (int j = 0; j < ny; ++j)
    for (int i = 0; i < nx; ++i) {
      const double x = ((double)i + 0.5) * l->h[0], y = ((double)j + 0.5) * l->h[1];
      s->img0[IDX(i, j)] = sin(x) + cos(y);
      s->img1[IDX(i, j)] = sin(x - 0.01) + cos(y + 0.01);
    }
}

/* apply bc to img0/img1; initGradients@all; rhs@finest  (`InitFields` after the images are set) */
void ogf_solver_init_fields(ogf_solver *s) {
  ogf_level *l = ogf_lv(s, s->cfg.max_level);
  const int nx = l->nx, ny = l->ny;
  const size_t sy = (size_t)nx + 2;
  ogf_neumann(nx, ny, s->img0);
  ogf_neumann(nx, ny, s->img1);
  const double gx1 = 1.0 / l->h[0], gx0 = -1.0 / l->h[0], gy1 = 1.0 / l->h[1], gy0 = -1.0 / l->h[1];
  for (int j = 0; j < ny; ++j)
    for (int i = 0; i < nx; ++i) {
      const size_t p = IDX(i, j);
      s->ix[p] = fma(gx0, s->img0[p], gx1 * s->img0[p + 1]);   /* GradX: [1,0] then [0,0] */
      s->iy[p] = fma(gy0, s->img0[p], gy1 * s->img0[p + sy]);
      s->it[p] = s->img1[p] - s->img0[p];
      l->cuu[p] = s->ix[p] * s->ix[p];
      l->cuv[p] = s->ix[p] * s->iy[p];
      l->cvv[p] = s->iy[p] * s->iy[p];
      l->fu[p] = -s->ix[p] * s->it[p];
      l->fv[p] = -s->iy[p] * s->it[p];
    }
  for (int lvl = s->cfg.max_level - 1; lvl >= s->cfg.min_level; --lvl) {
    ogf_level *c = ogf_lv(s, lvl), *f = ogf_lv(s, lvl + 1);
    ogf_restrict(c->nx, c->ny, f->cuu, c->cuu);
    ogf_restrict(c->nx, c->ny, f->cuv, c->cuv);
    ogf_restrict(c->nx, c->ny, f->cvv, c->cvv);
  }
}

static void ogf_up_residual(ogf_solver *s, int lvl) {
  ogf_level *l = ogf_lv(s, lvl);
  ogf_residual(l->nx, l->ny, l->L, s->a2, l->cuu, l->cuv, l->cvv, l->u, l->v, l->fu, l->fv, l->ru, l->rv);
  ogf_neumann(l->nx, l->ny, l->ru);
  ogf_neumann(l->nx, l->ny, l->rv);
}

static double ogf_res_norm(ogf_solver *s, int lvl) {
  ogf_level *l = ogf_lv(s, lvl);
  return sqrt(ogf_dot2(l->nx, l->ny, l->ru, l->ru, l->rv, l->rv));
}

double ogf_solver_update_residual_norm(ogf_solver *s, int lvl) {
  ogf_up_residual(s, lvl);
  return ogf_res_norm(s, lvl);
}

static void ogf_smooth(ogf_solver *s, int lvl, int n) {
  ogf_level *l = ogf_lv(s, lvl);
  for (int it = 0; it < n; ++it)
    for (int color = 0; color < 2; ++color) {
      ogf_smooth_color(l->nx, l->ny, l->L, s->a2, l->cuu, l->cuv, l->cvv, l->u, l->v, l->fu, l->fv, s->cfg.relax, color);
      ogf_neumann(l->nx, l->ny, l->u);
      ogf_neumann(l->nx, l->ny, l->v);
    }
}

static void ogf_axpy(int nx, int ny, double alpha, const double *x, double *y) {
  for (int j = 0; j < ny; ++j)
    for (int i = 0; i < nx; ++i) y[IDX(i, j)] = fma(alpha, x[IDX(i, j)], y[IDX(i, j)]);
}

static void ogf_cg(ogf_solver *s) {
  ogf_level *l = ogf_lv(s, s->cfg.min_level);
  const int nx = l->nx, ny = l->ny;
  ogf_up_residual(s, s->cfg.min_level);
  double curRes = ogf_res_norm(s, s->cfg.min_level);
  const double initRes = curRes;
  memcpy(s->pu, l->ru, sizeof(double) * (size_t)(nx + 2) * (size_t)(ny + 2));
  memcpy(s->pv, l->rv, sizeof(double) * (size_t)(nx + 2) * (size_t)(ny + 2));
  ogf_neumann(nx, ny, s->pu);
  ogf_neumann(nx, ny, s->pv);
  s->last_cg_iters = 0;
  for (int step = 0; step < s->cfg.cg_max_it; ++step) {
    s->last_cg_iters = step + 1;
    ogf_apply(nx, ny, l->L, s->a2, l->cuu, l->cuv, l->cvv, s->pu, s->pv, s->apu, s->apv);
    const double nom = ogf_dot2(nx, ny, l->ru, l->ru, NULL, NULL) + ogf_dot2(nx, ny, l->rv, l->rv, NULL, NULL);
    const double den = ogf_dot2(nx, ny, s->pu, s->apu, NULL, NULL) + ogf_dot2(nx, ny, s->pv, s->apv, NULL, NULL);
    const double alpha = nom / den;
    ogf_axpy(nx, ny, alpha, s->pu, l->u);
    ogf_neumann(nx, ny, l->u);
    ogf_axpy(nx, ny, -alpha, s->apu, l->ru);
    ogf_neumann(nx, ny, l->ru);
    ogf_axpy(nx, ny, alpha, s->pv, l->v);
    ogf_neumann(nx, ny, l->v);
    ogf_axpy(nx, ny, -alpha, s->apv, l->rv);
    ogf_neumann(nx, ny, l->rv);
    const double nextRes = ogf_res_norm(s, s->cfg.min_level);
    if (nextRes <= s->cfg.cg_rel_tol * initRes) return;
    const double beta = (nextRes * nextRes) / (curRes * curRes);
    for (int j = 0; j < ny; ++j)
      for (int i = 0; i < nx; ++i) {
        const size_t p = IDX(i, j);
        s->pu[p] = fma(beta, s->pu[p], l->ru[p]);
        s->pv[p] = fma(beta, s->pv[p], l->rv[p]);
      }
    ogf_neumann(nx, ny, s->pu);
    ogf_neumann(nx, ny, s->pv);
    curRes = nextRes;
  }
}

void ogf_solver_vcycle(ogf_solver *s, int lvl) {
  if (lvl == s->cfg.min_level) {
    ogf_cg(s);
    return;
  }
  ogf_level *l = ogf_lv(s, lvl), *c = ogf_lv(s, lvl - 1);
  ogf_smooth(s, lvl, s->cfg.nu_pre);
  ogf_up_residual(s, lvl);
  ogf_restrict(c->nx, c->ny, l->ru, c->fu);
  ogf_restrict(c->nx, c->ny, l->rv, c->fv);
  memset(c->u, 0, sizeof(double) * (size_t)(c->nx + 2) * (size_t)(c->ny + 2)); /* u@coarser = 0; apply bc */
  memset(c->v, 0, sizeof(double) * (size_t)(c->nx + 2) * (size_t)(c->ny + 2));
  ogf_solver_vcycle(s, lvl - 1);
  ogf_prolong_add(c->nx, c->ny, c->u, l->u);
  ogf_neumann(l->nx, l->ny, l->u);
  ogf_prolong_add(c->nx, c->ny, c->v, l->v);
  ogf_neumann(l->nx, l->ny, l->v);
  ogf_smooth(s, lvl, s->cfg.nu_post);
}

int ogf_solver_last_cg_iters(const ogf_solver *s) { return s->last_cg_iters; }
