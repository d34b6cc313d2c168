/*
 * examg_oracle_solver.c -- level loop / coarse-grid solve of the oracle.
 * TEST INFRASTRUCTURE ONLY (see examg_oracle.h).
 *
 * Restates mgCycle@(all but coarsest), mgCycle@coarsest and the residual/norm
 * part of Solve@finest of Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4:121-249
 * (identical structure: Examples/Poisson/2D_FD_Poisson_fromL4.exa4,
 * Testing/Smoothers/{RBGS,Jac}.exa4, Testing/SISC/{2D,3D}_{Const,Var}Coeff.exa4, Testing/Opts/base.exa4;
 * generator: C/s/e/solver/l3/L3_SolverForEquation.scala:341-395,
 * C/s/e/solver/l3/L3_ConjugateGradientForEquation.scala:37-151).
 */
#include <omp.h>

#include "examg_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  int n[3];
  size_t size;
  double h[3];
  og_stencil st;
  double *sf;   /* stencil-field coefficients [nst][size] or NULL */
  double *u[2]; /* Solution slots (slot 1 only for Jacobi) */
  int cur;      /* currentSlot (C/s/e/field/ir/IR_Slot.scala:33-62) */
  double *rhs, *res;
  double *p, *ap; /* cgTmp0 / cgTmp1 (coarsest only) */
  double *approx; /* gen_approx (solver_useFAS, all levels but the finest) */
} og_level;

struct og_solver {
  og_config cfg;
  og_level *lv; /* indexed by level - min_level */
  int last_cg_iters;
};

static og_level *L(og_solver *s, int lvl) { return &s->lv[lvl - s->cfg.min_level]; }

static double *og_calloc(size_t n) {
  double *p = (double *)calloc(n, sizeof(double)); /* initFieldsWithZero */
  if (!p) { fprintf(stderr, "og: out of memory\n"); exit(1); }
  return p;
}

/* getCoefficient of Testing/SISC/3D_VarCoeff.exa4 */
static double og_coeff3(double kappa, double x, double y, double z) {
  return exp(kappa * (((x - (x * x)) * (y - (y * y))) * (z - (z * z))));
}
/* getCoefficient of Testing/SISC/2D_VarCoeff.exa4:66-68 */
static double og_coeff2(double kappa, double x, double y) {
  return exp(kappa * ((x - (x * x)) * (y - (y * y))));
}

static void og_build_stencil(const og_config *c, og_level *l) {
  const int dim = c->dim;
  og_stencil *st = &l->st;
  memset(st, 0, sizeof(*st));
  if (c->op == OG_OP_VARCOEFF27) { /* 27 entries, lexicographic (dx fastest, dz slowest): centre is entry 13 */
    st->nst = 27;
    for (int e = 0; e < 27; ++e) { st->off[e][0] = e % 3 - 1; st->off[e][1] = (e / 3) % 3 - 1; st->off[e][2] = e / 9 - 1; }
    return;
  }
  st->nst = 2 * dim + 1;
  /* entry order as declared in the .exa4 files */
  int e = 1;
  for (int d = 0; d < dim; ++d) {
    int first = (c->order == OG_ORDER_BENCH) ? -1 : 1;
    st->off[e][d] = first; ++e;
    st->off[e][d] = -first; ++e;
  }
  if (c->op == OG_OP_LAPLACE_UNSCALED) {
    st->coef[0] = 2.0 * dim;
    for (int k = 1; k < st->nst; ++k) st->coef[k] = -1.0;
  } else {
    /* 2.0/(hx**2) + 2.0/(hy**2) [+ 2.0/(hz**2)], -1.0/(h**2) */
    double diag = 0.0;
    for (int d = 0; d < dim; ++d) diag = (d == 0) ? 2.0 / (l->h[d] * l->h[d])
                                                  : diag + 2.0 / (l->h[d] * l->h[d]);
    st->coef[0] = diag;
    for (int k = 1; k < st->nst; ++k) {
      int d = (k - 1) / 2;
      st->coef[k] = -1.0 / (l->h[d] * l->h[d]);
    }
  }
}

/* InitLaplace@lvl of Testing/SISC/3D_VarCoeff.exa4 (3D, 7 entries) */
static void og_init_varcoeff(const og_config *c, og_level *l) {
  const double kap = c->kappa;
  const double hx = l->h[0], hy = l->h[1], hz = l->h[2];
  const og_stencil *st = &l->st;
  l->sf = og_calloc((size_t)st->nst * l->size);
#pragma omp parallel for schedule(static) collapse(2)
  for (int k = 0; k < l->n[2]; ++k)
    for (int j = 0; j < l->n[1]; ++j)
      for (int i = 0; i < l->n[0]; ++i) {
        size_t p = (size_t)i + (size_t)l->n[0] * ((size_t)j + (size_t)l->n[1] * (size_t)k);
        double x = i * hx, y = j * hy, z = k * hz;
        double cxp, cxm, cyp, cym, czp = 0.0, czm = 0.0;
        if (c->dim == 3) {
          cxp = og_coeff3(kap, x + 0.5 * hx, y, z); cxm = og_coeff3(kap, x - 0.5 * hx, y, z);
          cyp = og_coeff3(kap, x, y + 0.5 * hy, z); cym = og_coeff3(kap, x, y - 0.5 * hy, z);
          czp = og_coeff3(kap, x, y, z + 0.5 * hz); czm = og_coeff3(kap, x, y, z - 0.5 * hz);
        } else {
          cxp = og_coeff2(kap, x + 0.5 * hx, y); cxm = og_coeff2(kap, x - 0.5 * hx, y);
          cyp = og_coeff2(kap, x, y + 0.5 * hy); cym = og_coeff2(kap, x, y - 0.5 * hy);
        }
        for (int e = 0; e < st->nst; ++e) {
          double v;
          const int *o = st->off[e];
          if (!o[0] && !o[1] && !o[2])
            v = (c->dim == 3)
                    ? ((cxp + cxm) / (hx * hx) + (cyp + cym) / (hy * hy)) + (czp + czm) / (hz * hz)
                    : (cxp + cxm) / (hx * hx) + (cyp + cym) / (hy * hy);
          else if (o[0] == 1) v = (-1.0 * cxp) / (hx * hx);
          else if (o[0] == -1) v = (-1.0 * cxm) / (hx * hx);
          else if (o[1] == 1) v = (-1.0 * cyp) / (hy * hy);
          else if (o[1] == -1) v = (-1.0 * cym) / (hy * hy);
          else if (o[2] == 1) v = (-1.0 * czp) / (hz * hz);
          else v = (-1.0 * czm) / (hz * hz);
          l->sf[(size_t)e * l->size + p] = v;
        }
      }
}

/* BASELINE config 3: "3D 27-point variable-coefficient Poisson".  The reference has no such program (its
 * variable-coefficient tests are 7-point, Testing/SISC/3D_VarCoeff.exa4); the operator is defined here in the style of
 * that file's InitLaplace: the compact 27-point Laplacian (faces 14/30, edges 3/30, corners 1/30 over h^2, centre
 * 128/30) with every off-centre entry weighted by getCoefficient at the midpoint towards that neighbour, and the
 * centre = -(sum of the others, accumulated in entry order), so rows sum to zero.  Meant for cubic cells (all measured
 * configurations); on other grids the midpoints use each dimension's width and the scaling stays 1/hx^2. */
static void og_init_varcoeff27(const og_config *c, og_level *l) {
  const double kap = c->kappa, h = l->h[0], hy = l->h[1], hz = l->h[2];
  static const double w[4] = {0.0, 14.0 / 30.0, 3.0 / 30.0, 1.0 / 30.0};
  l->sf = og_calloc((size_t)27 * l->size);
#pragma omp parallel for schedule(static) collapse(2)
  for (int k = 0; k < l->n[2]; ++k)
    for (int j = 0; j < l->n[1]; ++j)
      for (int i = 0; i < l->n[0]; ++i) {
        size_t p = (size_t)i + (size_t)l->n[0] * ((size_t)j + (size_t)l->n[1] * (size_t)k);
        double x = i * h, y = j * hy, z = k * hz;
        double sum = 0.0;
        for (int e = 0; e < 27; ++e) {
          if (e == 13) continue;
          const int dx = e % 3 - 1, dy = (e / 3) % 3 - 1, dz = e / 9 - 1;
          const int m = (dx != 0) + (dy != 0) + (dz != 0);
          const double a = og_coeff3(kap, x + (0.5 * dx) * h, y + (0.5 * dy) * hy, z + (0.5 * dz) * hz);
          const double v = ((-1.0 * w[m]) * a) / (h * h);
          l->sf[(size_t)e * l->size + p] = v;
          sum = sum + v;
        }
        l->sf[(size_t)13 * l->size + p] = -sum;
      }
}

void og_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }
int og_get_max_threads(void) { return omp_get_max_threads(); }
int og_count_threads(void) {
  int n = 1;
#pragma omp parallel
  {
#pragma omp master
    n = omp_get_num_threads();
  }
  return n;
}

og_solver *og_solver_create(const og_config *cfg) {
  og_solver *s = (og_solver *)calloc(1, sizeof(*s));
  s->cfg = *cfg;
  int nl = cfg->max_level - cfg->min_level + 1;
  s->lv = (og_level *)calloc((size_t)nl, sizeof(og_level));
  for (int lvl = cfg->min_level; lvl <= cfg->max_level; ++lvl) {
    og_level *l = L(s, lvl);
    for (int d = 0; d < 3; ++d) {
      if (d < cfg->dim) {
        int cells = cfg->F[d] << lvl;
        l->n[d] = cells + 1;
        const double hi = cfg->domain_hi[d] > 0.0 ? cfg->domain_hi[d] : 1.0;
        l->h[d] = hi / (double)cells; /* IR_DomainFromAABB.scala:31-40, domain [0,hi]^d */
      } else { l->n[d] = 1; l->h[d] = 1.0; }
    }
    l->size = (size_t)l->n[0] * l->n[1] * l->n[2];
    og_build_stencil(cfg, l);
    if (cfg->op == OG_OP_VARCOEFF) og_init_varcoeff(cfg, l);
    if (cfg->op == OG_OP_VARCOEFF27) og_init_varcoeff27(cfg, l);
    l->u[0] = og_calloc(l->size);
    if (cfg->smoother == OG_SMOOTH_JACOBI) l->u[1] = og_calloc(l->size);
    l->rhs = og_calloc(l->size);
    l->res = og_calloc(l->size);
    if (lvl == cfg->min_level) { l->p = og_calloc(l->size); l->ap = og_calloc(l->size); }
    if (cfg->use_fas && lvl != cfg->max_level) l->approx = og_calloc(l->size);
  }
  return s;
}

void og_solver_destroy(og_solver *s) {
  if (!s) return;
  int nl = s->cfg.max_level - s->cfg.min_level + 1;
  for (int i = 0; i < nl; ++i) {
    og_level *l = &s->lv[i];
    free(l->sf); free(l->u[0]); free(l->u[1]); free(l->rhs); free(l->res); free(l->p); free(l->ap); free(l->approx);
  }
  free(s->lv);
  free(s);
}

double *og_solver_solution(og_solver *s, int lvl) { return L(s, lvl)->u[L(s, lvl)->cur]; }
double *og_solver_rhs(og_solver *s, int lvl) { return L(s, lvl)->rhs; }
double *og_solver_residual(og_solver *s, int lvl) { return L(s, lvl)->res; }
double *og_solver_coeff(og_solver *s, int lvl) { return L(s, lvl)->sf; }
void og_solver_dims(const og_solver *s, int lvl, int n[3]) {
  const og_level *l = &s->lv[lvl - s->cfg.min_level];
  n[0] = l->n[0]; n[1] = l->n[1]; n[2] = l->n[2];
}
void og_solver_stencil(const og_solver *s, int lvl, og_stencil *out) {
  *out = s->lv[lvl - s->cfg.min_level].st;
}
int og_solver_last_cg_iters(const og_solver *s) { return s->last_cg_iters; }

/* `apply bc to Solution<0>; apply bc to Solution<1>` after the caller wrote the
 * Dirichlet values into the active slot (Testing/Smoothers/Jac.exa4 Application) */
void og_solver_sync_slots(og_solver *s) {
  for (int lvl = s->cfg.min_level; lvl <= s->cfg.max_level; ++lvl) {
    og_level *l = L(s, lvl);
    if (l->u[1]) memcpy(l->u[1 - l->cur], l->u[l->cur], l->size * sizeof(double));
  }
}

static void og_smooth(og_solver *s, og_level *l, int nsweeps) {
  for (int it = 0; it < nsweeps; ++it) {
    if (s->cfg.smoother == OG_SMOOTH_RBGS) {
      og_smooth_rb(l->n, &l->st, l->sf, s->cfg.omega, l->u[l->cur], l->rhs, 0);
      og_smooth_rb(l->n, &l->st, l->sf, s->cfg.omega, l->u[l->cur], l->rhs, 1);
    } else {
      og_smooth_jacobi(l->n, &l->st, l->sf, s->cfg.omega, l->u[l->cur], l->rhs, l->u[1 - l->cur]);
      l->cur = 1 - l->cur; /* advance Solution */
    }
  }
}

double og_solver_update_residual_norm(og_solver *s, int lvl) {
  og_level *l = L(s, lvl);
  og_residual(l->n, &l->st, l->sf, l->u[l->cur], l->rhs, l->res);
  return sqrt(og_dot(l->n, l->res, l->res));
}

/* mgCycle@coarsest: un-preconditioned CG (…fromL4.exa4:152-201) */
static void og_cg(og_solver *s, og_level *l) {
  double *u = l->u[l->cur];
  og_residual(l->n, &l->st, l->sf, u, l->rhs, l->res);
  double curRes = sqrt(og_dot(l->n, l->res, l->res));
  const double initRes = curRes;
  og_copy_interior(l->n, l->res, l->p);
  s->last_cg_iters = 0;
  for (int step = 0; step < s->cfg.cg_max_it; ++step) {
    s->last_cg_iters = step + 1;
    og_apply(l->n, &l->st, l->sf, l->p, l->ap);
    double alphaNom = og_dot(l->n, l->res, l->res);
    double alphaDenom = og_dot(l->n, l->p, l->ap);
    double alpha = alphaNom / alphaDenom;
    og_axpy(l->n, alpha, l->p, u);
    og_axpy(l->n, -alpha, l->ap, l->res);
    double nextRes = sqrt(og_dot(l->n, l->res, l->res));
    if (nextRes <= s->cfg.cg_rel_tol * initRes) return;
    double beta = (nextRes * nextRes) / (curRes * curRes);
    og_xpay(l->n, l->res, beta, l->p);
    curRes = nextRes;
  }
}

void og_solver_vcycle(og_solver *s, int lvl) {
  og_level *l = L(s, lvl);
  if (lvl == s->cfg.min_level) { og_cg(s, l); return; }
  og_level *c = L(s, lvl - 1);
  og_smooth(s, l, s->cfg.nu_pre);
  og_residual(l->n, &l->st, l->sf, l->u[l->cur], l->rhs, l->res);
  og_restrict(s->cfg.dim, l->n, c->n, s->cfg.restrict_scale, l->res, c->rhs);
  if (s->cfg.use_fas) {
    /* L3_SolverForEquation.scala:401-452: Approx@coarser = Restriction * Solution; Solution@coarser = Approx@coarser;
     * RHS@coarser = Restriction * Residual + A@coarser * Solution@coarser; ...; Solution@coarser -= Approx@coarser */
    og_restrict(s->cfg.dim, l->n, c->n, 1.0, l->u[l->cur], c->approx);
    og_copy_interior(c->n, c->approx, c->u[c->cur]);
    og_apply(c->n, &c->st, c->sf, c->u[c->cur], c->res);
    og_axpy(c->n, 1.0, c->res, c->rhs);
    og_solver_vcycle(s, lvl - 1);
    og_axpy(c->n, -1.0, c->approx, c->u[c->cur]);
  } else {
    og_set_interior(c->n, c->u[c->cur], 0.0);
    og_solver_vcycle(s, lvl - 1);
  }
  og_prolong_add(s->cfg.dim, l->n, c->n, c->u[c->cur], l->u[l->cur]);
  og_smooth(s, l, s->cfg.nu_post);
}

void og_print_reduced_prec(double v, char *buf, size_t buflen) {
  if (v <= 1e-12) { snprintf(buf, buflen, "EFFECTIVELY ZERO"); return; }
  int prec = 4;
  if (v <= 1e-11) prec = 1;
  else if (v <= 1e-10) prec = 2;
  else if (v <= 1e-9) prec = 3;
  snprintf(buf, buflen, "%.*g", prec, v);
}
