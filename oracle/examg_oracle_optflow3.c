/*
 * examg_oracle_optflow3.c -- CPU restatement of the ExaStencils 3D optical-flow solver
 * (cell-centred, three coupled unknowns in a Vec3 field, red-black block-Jacobi-type smoother with the
 * 3x3 diagonal block inverted per cell).  TEST INFRASTRUCTURE ONLY (see examg_oracle.h).
 *
 * Follows (file:line under /root/reference):
 *   program text        Testing/Application/OpticalFlow3D.exa4 (whole file; golden OpticalFlow3D.results),
 *                       Testing/Application/OpticalFlow3D.knowledge (levels 2..6: 64^3 cells)
 *   smoother            `flow = flow + inverse ( diag ( combinedOp ) ) * ( rhs - combinedOp * flow )` under
 *                       `repeat with { 0 == (i0+i1+i2) % 2, 1 == ... }` (OpticalFlow3D.exa4:219-231): diag() of a
 *                       matrix-valued stencil is its centre entry, a 3x3 block; inverse() of a 3x3 expression is
 *                       resolved at compile time by the cofactor formulas of
 *                       C/s/e/baseExt/ir/IR_MatOperations/IR_CompiletimeMatOps.scala:804-826
 *   Neumann (order 1)   ghost cell = adjacent inner cell, faces -x,+x,-y,+y,-z,+z over the full orthogonal extent
 *                       incl. ghosts: C/s/e/boundary/ir/IR_NeumannBC.scala:70-80, IR_ApplyBCFunction.scala:53-85
 *   cell restriction    [1/2,1/2]^3 = mean of the 8 children, prolongation piecewise constant:
 *                       C/s/e/operator/l4/L4_DefaultRestriction.scala:40-47, L4_DefaultProlongation.scala:30-45
 *   grid                cell centre (i + 1/2) h, h = 1 / (F 2^level)   C/s/e/domain/ir/IR_DomainFromAABB.scala:31-40
 *
 * Dense layout of every array: (nz+2) x (ny+2) x (nx+2) doubles, x fastest, one ghost cell per side;
 * cell (i,j,k) is element (i+1) + (nx+2)*((j+1) + (ny+2)*(k+1)).
 *
 * Arithmetic contract (shared with the CUDA kernels, csrc/examg_optflow3.cu); G = the symmetric 3x3 block of
 * gradContribution (xx,xy,xz,yy,yz,zz), f = flow component c, a2 = alpha**2:
 *   L(w)   = Lc*w_c, then fma over -x,+x,-y,+y,-z,+z in declaration order
 *   A_c    = fma(a2, L(f_c), t),  t = G_c0*f_0, fma(G_c1, f_1, t), fma(G_c2, f_2, t)
 *   residual_c = rhs_c - A_c
 *   smoother at one cell: r = residual there; D = G + a2*Lc on the diagonal (D_cc = fma(a2, Lc, G_cc));
 *     cofactors A = e*i - f*h, B = -(d*i - f*g), C = d*h - e*g, D' = -(b*i - c*h), E = a*i - c*g,
 *     F = -(a*h - b*g), G' = b*f - c*e, H = -(a*f - c*d), I = a*e - b*d; det = a*A + b*B + c*C;
 *     inverse rows (A/det, D'/det, G'/det), (B/det, E/det, H/det), (C/det, F/det, I/det);
 *     f_c += inv_c0*r_0 + inv_c1*r_1 + inv_c2*r_2           (products rounded, added left to right)
 *   sums: per cell d = x0*y0 + x1*y1 + x2*y2 (left to right), accumulated in double-double
 * Parity: pinned by Testing/Application/OpticalFlow3D.results (5 printed values, tests/test_oracle_golden.py).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "examg_oracle.h"

typedef struct { double hi, lo; } og3_dd;
static inline void og3_dd_add(og3_dd *a, double x) {
  double s = a->hi + x;
  double bb = s - a->hi;
  double err = (a->hi - (s - bb)) + (x - bb);
  a->hi = s;
  a->lo += err;
}
static inline void og3_dd_merge(og3_dd *a, const og3_dd *b) {
  og3_dd_add(a, b->hi);
  a->lo += b->lo;
}

typedef struct {
  int n[3];
  size_t sy, sz, vol;
  double h[3], L[7];
  double *f[15]; /* flow 0..2, rhs 3..5, residual 6..8, G xx,xy,xz,yy,yz,zz 9..14 */
} og3_level;

struct og3_solver {
  og3_config cfg;
  int nlev;
  og3_level *lev;
  double *img0, *img1;  /* finest */
  double *p[3], *ap[3]; /* coarsest: cgTmp0, cgTmp1 */
  double a2;
  int last_cg_iters;
};

#define P3(l, i, j, k) ((size_t)((i) + 1) + (l)->sy * (size_t)((j) + 1) + (l)->sz * (size_t)((k) + 1))

static og3_level *og3_lv(og3_solver *s, int lvl) { return &s->lev[lvl - s->cfg.min_level]; }

/* apply bc (Neumann): every ghost cell = the inner cell with clamped indices */
static void og3_neumann(const og3_level *l, double *a) {
  const int nx = l->n[0], ny = l->n[1], nz = l->n[2];
  for (int k = -1; k <= nz; ++k)
    for (int j = -1; j <= ny; ++j)
      for (int i = -1; i <= nx; ++i) {
        if (i >= 0 && i < nx && j >= 0 && j < ny && k >= 0 && k < nz) continue;
        const int ci = i < 0 ? 0 : (i >= nx ? nx - 1 : i), cj = j < 0 ? 0 : (j >= ny ? ny - 1 : j);
        const int ck = k < 0 ? 0 : (k >= nz ? nz - 1 : k);
        a[P3(l, i, j, k)] = a[P3(l, ci, cj, ck)];
      }
}

static inline double og3_lap(const og3_level *l, const double *w, size_t p) {
  double acc = l->L[0] * w[p];
  acc = fma(l->L[1], w[p - 1], acc);
  acc = fma(l->L[2], w[p + 1], acc);
  acc = fma(l->L[3], w[p - l->sy], acc);
  acc = fma(l->L[4], w[p + l->sy], acc);
  acc = fma(l->L[5], w[p - l->sz], acc);
  acc = fma(l->L[6], w[p + l->sz], acc);
  return acc;
}

/* index of G_cd in f[9..14] */
static const int og3_g[3][3] = {{9, 10, 11}, {10, 12, 13}, {11, 13, 14}};

/* (combinedOp * w)_c at p */
static inline double og3_apply_c(const og3_level *l, double a2, double *const w[3], int c, size_t p) {
  double t = l->f[og3_g[c][0]][p] * w[0][p];
  t = fma(l->f[og3_g[c][1]][p], w[1][p], t);
  t = fma(l->f[og3_g[c][2]][p], w[2][p], t);
  return fma(a2, og3_lap(l, w[c], p), t);
}

/* residual = rhs - combinedOp * flow; apply bc to residual */
static void og3_up_residual(og3_solver *s, int lvl) {
  og3_level *l = og3_lv(s, lvl);
  double *w[3] = {l->f[0], l->f[1], l->f[2]};
#pragma omp parallel for schedule(static)
  for (int k = 0; k < l->n[2]; ++k)
    for (int j = 0; j < l->n[1]; ++j)
      for (int i = 0; i < l->n[0]; ++i) {
        const size_t p = P3(l, i, j, k);
        for (int c = 0; c < 3; ++c) l->f[6 + c][p] = l->f[3 + c][p] - og3_apply_c(l, s->a2, w, c, p);
      }
  for (int c = 0; c < 3; ++c) og3_neumann(l, l->f[6 + c]);
}

/* sum over the cells of x . y (Vec3 dot product per cell) */
static double og3_dot(const og3_level *l, double *const x[3], double *const y[3]) {
  og3_dd total = {0.0, 0.0};
  for (int k = 0; k < l->n[2]; ++k)
    for (int j = 0; j < l->n[1]; ++j)
      for (int i = 0; i < l->n[0]; ++i) {
        const size_t p = P3(l, i, j, k);
        og3_dd_add(&total, x[0][p] * y[0][p] + x[1][p] * y[1][p] + x[2][p] * y[2][p]);
      }
  og3_dd z = {0.0, 0.0};
  og3_dd_merge(&z, &total);
  return z.hi + z.lo;
}

static double og3_res_norm(og3_solver *s, int lvl) {
  og3_level *l = og3_lv(s, lvl);
  double *r[3] = {l->f[6], l->f[7], l->f[8]};
  return sqrt(og3_dot(l, r, r));
}

double og3_solver_update_residual_norm(og3_solver *s, int lvl) {
  og3_up_residual(s, lvl);
  return og3_res_norm(s, lvl);
}

/* one colour of the smoother (in place), then apply bc to flow */
void og3_solver_smooth_color(og3_solver *s, int lvl, int color) {
  og3_level *l = og3_lv(s, lvl);
  double *w[3] = {l->f[0], l->f[1], l->f[2]};
  const double a2 = s->a2;
#pragma omp parallel for schedule(static)
  for (int k = 0; k < l->n[2]; ++k)
    for (int j = 0; j < l->n[1]; ++j)
      for (int i = (j + k + color) & 1; i < l->n[0]; i += 2) {
        const size_t p = P3(l, i, j, k);
        double r[3];
        for (int c = 0; c < 3; ++c) r[c] = l->f[3 + c][p] - og3_apply_c(l, a2, w, c, p);
        const double a = fma(a2, l->L[0], l->f[9][p]), b = l->f[10][p], cc = l->f[11][p];
        const double d = b, e = fma(a2, l->L[0], l->f[12][p]), f = l->f[13][p];
        const double g = cc, h = f, ii = fma(a2, l->L[0], l->f[14][p]);
        const double A = e * ii - f * h, B = -(d * ii - f * g), C = d * h - e * g;
        const double D = -(b * ii - cc * h), E = a * ii - cc * g, F = -(a * h - b * g);
        const double G = b * f - cc * e, H = -(a * f - cc * d), I = a * e - b * d;
        const double det = a * A + b * B + cc * C;
        const double inv[3][3] = {{A / det, D / det, G / det}, {B / det, E / det, H / det}, {C / det, F / det, I / det}};
        for (int c = 0; c < 3; ++c) w[c][p] = w[c][p] + (inv[c][0] * r[0] + inv[c][1] * r[1] + inv[c][2] * r[2]);
      }
  for (int c = 0; c < 3; ++c) og3_neumann(l, w[c]);
}

void og3_solver_smooth(og3_solver *s, int lvl, int n) {
  for (int it = 0; it < n; ++it)
    for (int color = 0; color < 2; ++color) og3_solver_smooth_color(s, lvl, color);
}

/* coarse = RestrictionCell * fine: 0.125 * (sum of the 8 children in lexicographic z, y, x order) */
static void og3_restrict(const og3_level *c, const og3_level *f, const double *fine, double *coarse) {
#pragma omp parallel for schedule(static)
  for (int k = 0; k < c->n[2]; ++k)
    for (int j = 0; j < c->n[1]; ++j)
      for (int i = 0; i < c->n[0]; ++i) {
        double sum = 0.0;
        int first = 1;
        for (int dz = 0; dz < 2; ++dz)
          for (int dy = 0; dy < 2; ++dy)
            for (int dx = 0; dx < 2; ++dx) {
              const double v = fine[P3(f, 2 * i + dx, 2 * j + dy, 2 * k + dz)];
              sum = first ? v : sum + v;
              first = 0;
            }
        coarse[P3(c, i, j, k)] = 0.125 * sum;
      }
}

/* fine += ProlongationCell * coarse (piecewise constant) */
static void og3_prolong_add(const og3_level *c, const og3_level *f, const double *coarse, double *fine) {
#pragma omp parallel for schedule(static)
  for (int k = 0; k < f->n[2]; ++k)
    for (int j = 0; j < f->n[1]; ++j)
      for (int i = 0; i < f->n[0]; ++i) fine[P3(f, i, j, k)] += coarse[P3(c, i >> 1, j >> 1, k >> 1)];
}

/* ---- whole solver -------------------------------------------------------- */

og3_solver *og3_solver_create(const og3_config *cfg) {
  og3_solver *s = (og3_solver *)calloc(1, sizeof(*s));
  s->cfg = *cfg;
  s->nlev = cfg->max_level - cfg->min_level + 1;
  s->lev = (og3_level *)calloc((size_t)s->nlev, sizeof(og3_level));
  s->a2 = cfg->alpha * cfg->alpha; /* alpha ** 2 */
  for (int lvl = cfg->min_level; lvl <= cfg->max_level; ++lvl) {
    og3_level *l = og3_lv(s, lvl);
    for (int d = 0; d < 3; ++d) {
      l->n[d] = cfg->F[d] << lvl;
      l->h[d] = 1.0 / (double)l->n[d];
    }
    l->sy = (size_t)l->n[0] + 2;
    l->sz = l->sy * ((size_t)l->n[1] + 2);
    l->vol = l->sz * ((size_t)l->n[2] + 2);
    const double hx2 = l->h[0] * l->h[0], hy2 = l->h[1] * l->h[1], hz2 = l->h[2] * l->h[2];
    l->L[0] = 2.0 / hx2 + 2.0 / hy2 + 2.0 / hz2; /* Stencil Laplace, declaration order */
    l->L[1] = l->L[2] = -1.0 / hx2;
    l->L[3] = l->L[4] = -1.0 / hy2;
    l->L[5] = l->L[6] = -1.0 / hz2;
    for (int k = 0; k < 15; ++k) l->f[k] = (double *)calloc(l->vol, sizeof(double));
  }
  og3_level *f = og3_lv(s, cfg->max_level), *c = og3_lv(s, cfg->min_level);
  s->img0 = (double *)calloc(f->vol, sizeof(double));
  s->img1 = (double *)calloc(f->vol, sizeof(double));
  for (int k = 0; k < 3; ++k) {
    s->p[k] = (double *)calloc(c->vol, sizeof(double));
    s->ap[k] = (double *)calloc(c->vol, sizeof(double));
  }
  return s;
}

void og3_solver_destroy(og3_solver *s) {
  if (!s) return;
  for (int q = 0; q < s->nlev; ++q)
    for (int k = 0; k < 15; ++k) free(s->lev[q].f[k]);
  free(s->img0);
  free(s->img1);
  for (int k = 0; k < 3; ++k) {
    free(s->p[k]);
    free(s->ap[k]);
  }
  free(s->lev);
  free(s);
}

void og3_solver_dims(const og3_solver *s, int lvl, int n[3]) {
  for (int d = 0; d < 3; ++d) n[d] = s->lev[lvl - s->cfg.min_level].n[d];
}

/* which: 0..14 the level's fields, OG3_IMG0 / OG3_IMG1 the images (finest) */
double *og3_solver_field(og3_solver *s, int lvl, int which) {
  if (which == OG3_IMG0) return s->img0;
  if (which == OG3_IMG1) return s->img1;
  if (which < 0 || which >= 15) return NULL;
  return og3_lv(s, lvl)->f[which];
}

/* initImgs@finest (OpticalFlow3D.exa4:334-346) */
void og3_solver_init_images(og3_solver *s) {
  og3_level *l = og3_lv(s, s->cfg.max_level);
  for (int k = 0; k < l->n[2]; ++k)
    for (int j = 0; j < l->n[1]; ++j)
      for (int i = 0; i < l->n[0]; ++i) {
        const double x = ((double)i + 0.5) * l->h[0], y = ((double)j + 0.5) * l->h[1], z = ((double)k + 0.5) * l->h[2];
        s->img0[P3(l, i, j, k)] = sin(x) + cos(y) + sin(z);
        s->img1[P3(l, i, j, k)] = sin(x - 0.01) + cos(y + 0.01) + sin(z - 0.005);
      }
}

/* apply bc to the images; initGradients@all; rhs@finest (`InitFields` after the images are set) */
void og3_solver_init_fields(og3_solver *s) {
  og3_level *l = og3_lv(s, s->cfg.max_level);
  og3_neumann(l, s->img0);
  og3_neumann(l, s->img1);
  const double g1[3] = {1.0 / l->h[0], 1.0 / l->h[1], 1.0 / l->h[2]};
  const double g0[3] = {-1.0 / l->h[0], -1.0 / l->h[1], -1.0 / l->h[2]};
  const size_t st[3] = {1, l->sy, l->sz};
  for (int k = 0; k < l->n[2]; ++k)
    for (int j = 0; j < l->n[1]; ++j)
      for (int i = 0; i < l->n[0]; ++i) {
        const size_t p = P3(l, i, j, k);
        double I[3];
        for (int d = 0; d < 3; ++d) I[d] = fma(g0[d], s->img0[p], g1[d] * s->img0[p + st[d]]); /* Grad: [+1] then [0] */
        const double It = s->img1[p] - s->img0[p];
        l->f[9][p] = I[0] * I[0];
        l->f[10][p] = I[0] * I[1];
        l->f[11][p] = I[0] * I[2];
        l->f[12][p] = I[1] * I[1];
        l->f[13][p] = I[1] * I[2];
        l->f[14][p] = I[2] * I[2];
        for (int d = 0; d < 3; ++d) l->f[3 + d][p] = -I[d] * It;
      }
  for (int lvl = s->cfg.max_level - 1; lvl >= s->cfg.min_level; --lvl) {
    og3_level *c = og3_lv(s, lvl), *f = og3_lv(s, lvl + 1);
    for (int k = 9; k < 15; ++k) og3_restrict(c, f, f->f[k], c->f[k]);
  }
}

/* mgCycle@coarsest (OpticalFlow3D.exa4:266-318) */
static void og3_cg(og3_solver *s) {
  og3_level *l = og3_lv(s, s->cfg.min_level);
  double *r[3] = {l->f[6], l->f[7], l->f[8]};
  og3_up_residual(s, s->cfg.min_level);
  double curRes = og3_res_norm(s, s->cfg.min_level);
  const double initRes = curRes;
  for (int c = 0; c < 3; ++c) {
    memcpy(s->p[c], r[c], sizeof(double) * l->vol);
    og3_neumann(l, s->p[c]);
  }
  s->last_cg_iters = 0;
  for (int step = 0; step < s->cfg.cg_max_it; ++step) {
    s->last_cg_iters = step + 1;
    for (int k = 0; k < l->n[2]; ++k)
      for (int j = 0; j < l->n[1]; ++j)
        for (int i = 0; i < l->n[0]; ++i) {
          const size_t p = P3(l, i, j, k);
          for (int c = 0; c < 3; ++c) s->ap[c][p] = og3_apply_c(l, s->a2, s->p, c, p);
        }
    const double nom = og3_dot(l, r, r), den = og3_dot(l, s->p, s->ap);
    const double alpha = nom / den;
    for (int c = 0; c < 3; ++c) {
      for (int k = 0; k < l->n[2]; ++k)
        for (int j = 0; j < l->n[1]; ++j)
          for (int i = 0; i < l->n[0]; ++i) {
            const size_t p = P3(l, i, j, k);
            l->f[c][p] = fma(alpha, s->p[c][p], l->f[c][p]);
            r[c][p] = fma(-alpha, s->ap[c][p], r[c][p]);
          }
      og3_neumann(l, l->f[c]);
      og3_neumann(l, r[c]);
    }
    const double nextRes = og3_res_norm(s, s->cfg.min_level);
    if (nextRes <= s->cfg.cg_rel_tol * initRes) return;
    const double beta = (nextRes * nextRes) / (curRes * curRes);
    for (int c = 0; c < 3; ++c) {
      for (int k = 0; k < l->n[2]; ++k)
        for (int j = 0; j < l->n[1]; ++j)
          for (int i = 0; i < l->n[0]; ++i) {
            const size_t p = P3(l, i, j, k);
            s->p[c][p] = fma(beta, s->p[c][p], r[c][p]);
          }
      og3_neumann(l, s->p[c]);
    }
    curRes = nextRes;
  }
}

/* residual@lvl -> rhs@(lvl-1); flow@(lvl-1) = 0 */
void og3_solver_restrict(og3_solver *s, int lvl) {
  og3_level *l = og3_lv(s, lvl), *c = og3_lv(s, lvl - 1);
  for (int k = 0; k < 3; ++k) {
    og3_restrict(c, l, l->f[6 + k], c->f[3 + k]);
    memset(c->f[k], 0, sizeof(double) * c->vol);
  }
}

/* flow@lvl += ProlongationCell * flow@(lvl-1); apply bc */
void og3_solver_correct(og3_solver *s, int lvl) {
  og3_level *l = og3_lv(s, lvl), *c = og3_lv(s, lvl - 1);
  for (int k = 0; k < 3; ++k) {
    og3_prolong_add(c, l, c->f[k], l->f[k]);
    og3_neumann(l, l->f[k]);
  }
}

void og3_solver_vcycle(og3_solver *s, int lvl) {
  if (lvl == s->cfg.min_level) {
    og3_cg(s);
    return;
  }
  og3_solver_smooth(s, lvl, s->cfg.nu_pre);
  og3_up_residual(s, lvl);
  og3_solver_restrict(s, lvl);
  og3_solver_vcycle(s, lvl - 1);
  og3_solver_correct(s, lvl);
  og3_solver_smooth(s, lvl, s->cfg.nu_post);
}

int og3_solver_last_cg_iters(const og3_solver *s) { return s->last_cg_iters; }
