/* examg_oracle_stokes.c -- CPU restatement of the staggered-grid Stokes programs of the reference: TEST INFRASTRUCTURE ONLY
 * (only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline legs may use anything under oracle/).
 *
 * Follows Testing/Application/ExaStokes_2D.exa4 and ExaStokes_3D.exa4 (lid-driven cavity on the unit square / cube):
 *   fields      velocity component d at Face_d (nodes in dimension d, cells in the others), p at cell centres  2D:3-45
 *   operators   Laplace (star, declaration order centre,-x,+x,-y,+y[,-z,+z]), d?Left / d?Right          2D:49-77, 3D:54-92
 *   smoother    color with { i0 % 3, i1 % 3 [, i2 % 3] } loop over p { solve locally 2 dim + 1 unknowns: the two faces
 *               of the cell in every dimension and p }, then apply bc                          2D:228-303, 3D:272-381
 *               local system: C/s/e/solver/ir/IR_LocalSolve.scala:88-236 (terms with unknowns -> matrix, the rest ->
 *               right-hand side), rows of unknowns that are no valid computation points (wall faces) are identities and
 *               are not written back: IR_LocalDirectInvert.scala:76-103,127-147, IR_IsOnBoundary.scala:53-97
 *   loops       node-like dimensions skip the boundary node, cell-like ones run over all cells
 *               (C/s/e/baseExt/ir/IR_LoopOverPointsInOneFragment.scala:76-112)
 *   apply bc    Dirichlet on a face field: value on the wall faces of its own dimension, ghost = 2 g - inner in the others;
 *               tangential range ghost to ghost, faces in the order -x,+x,-y,+y,-z,+z
 *               (C/s/e/boundary/ir/IR_BoundaryCondition.scala:38-45, IR_ApplyBCFunction.scala:55-88, IR_DirichletBC.scala:37-50);
 *               u on the finest level: the program's ApplyBC_u (ghost rows in y, then the wall faces in x; its `only dup
 *               [0,0,+-1]` loops are empty for a Face_x field, so the ghost layers in z keep their initial zeros)
 *   transfer    default 'linear' restriction = kron of [1/4,1/2,1/4] (node-like) and [1/2,1/2] (cell-like), prolongation =
 *               2^dim * transpose  (C/s/e/operator/l4/L4_DefaultRestriction.scala:29-47,79-100, L4_DefaultProlongation.scala:30-45)
 *   cycle       V(3,3), 10 smoother steps on the coarsest level, SetSolution@coarser(0) incl. one layer around
 * Pinned by tests/test_oracle_golden.py::test_golden_stokes against Testing/Application/ExaStokes_{2D,3D}.results.
 *
 * Storage: every field is an (n+3)^dim array, entry (i, j, k) at (i+1) + (n+3) (j+1) + (n+3)^2 (k+1).
 * Arithmetic contract (what libexamg's kernels reproduce bit for bit; built with -ffp-contract=off): stated at each loop.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  int dim;
  int min_level, max_level;
  int nu_pre, nu_post, nu_coarse;
  double top_vel;
} ogs_config;

/* field ids: velocity d, rhs of velocity d, residual of velocity d (d = 0..2), then p, rhs_p, residual_p */
enum { OGS_VEL = 0, OGS_RHS = 3, OGS_RES = 6, OGS_P = 9, OGS_RHS_P = 10, OGS_RES_P = 11, OGS_NFIELDS = 12 };

typedef struct {
  int n;        /* cells per dimension */
  size_t s[3];  /* strides; s[2] = 0 in 2D */
  size_t size;
  double h;
  double *f[OGS_NFIELDS];
} ogs_level;

typedef struct ogs_solver {
  ogs_config cfg;
  ogs_level *lev;
} ogs_solver;

static ogs_level *ogs_lv(ogs_solver *s, int lvl) { return &s->lev[lvl - s->cfg.min_level]; }

static inline size_t ogs_at(const ogs_level *L, const int idx[3]) {
  return (size_t)(idx[0] + 1) * L->s[0] + (size_t)(idx[1] + 1) * L->s[1] + (L->s[2] ? (size_t)(idx[2] + 1) * L->s[2] : 0);
}

typedef struct { double hi, lo; } ogs_dd;
static inline void ogs_dd_add(ogs_dd *a, double x) {
  double s = a->hi + x;
  double bb = s - a->hi;
  double err = (a->hi - (s - bb)) + (x - bb);
  a->hi = s;
  a->lo += err;
}

ogs_solver *ogs_solver_create(const ogs_config *cfg) {
  ogs_solver *s = (ogs_solver *)calloc(1, sizeof(*s));
  s->cfg = *cfg;
  const int nl = cfg->max_level - cfg->min_level + 1;
  s->lev = (ogs_level *)calloc((size_t)nl, sizeof(ogs_level));
  for (int lvl = cfg->min_level; lvl <= cfg->max_level; ++lvl) {
    ogs_level *L = ogs_lv(s, lvl);
    L->n = 1 << lvl;
    const size_t m = (size_t)L->n + 3;
    L->s[0] = 1;
    L->s[1] = m;
    L->s[2] = cfg->dim == 3 ? m * m : 0;
    L->size = cfg->dim == 3 ? m * m * m : m * m;
    L->h = 1.0 / (double)L->n;
    for (int k = 0; k < OGS_NFIELDS; ++k) L->f[k] = (double *)calloc(L->size, sizeof(double));
  }
  return s;
}

void ogs_solver_destroy(ogs_solver *s) {
  if (!s) return;
  const int nl = s->cfg.max_level - s->cfg.min_level + 1;
  for (int l = 0; l < nl; ++l)
    for (int k = 0; k < OGS_NFIELDS; ++k) free(s->lev[l].f[k]);
  free(s->lev);
  free(s);
}

double *ogs_solver_field(ogs_solver *s, int lvl, int which) { return ogs_lv(s, lvl)->f[which]; }
int ogs_solver_cells(ogs_solver *s, int lvl) { return ogs_lv(s, lvl)->n; }

/* loop over the box lo..hi (inclusive) of the first `dim` dimensions */
#define OGS_BOX(dim, lo, hi, idx)                                                       \
  for (idx[2] = (dim) == 3 ? lo[2] : 0; idx[2] <= ((dim) == 3 ? hi[2] : 0); ++idx[2])   \
    for (idx[1] = lo[1]; idx[1] <= hi[1]; ++idx[1])                                     \
      for (idx[0] = lo[0]; idx[0] <= hi[0]; ++idx[0])

/* the faces of dimension e of velocity component d.  e == d: wall faces idx[e] = 0, n take `value` (kind 0);
 * e != d: ghost = g2 - inner (kind 1) or -inner (kind 2), g2 given per side.  tlo/thi: tangential ranges. */
static void ogs_bc_faces(const ogs_level *L, int dim, double *w, int e, int kind, double g2_lo, double g2_hi,
                         const int tlo[3], const int thi[3]) {
  const int n = L->n;
  for (int side = 0; side < 2; ++side) {
    int lo[3] = {tlo[0], tlo[1], tlo[2]}, hi[3] = {thi[0], thi[1], thi[2]}, idx[3];
    if (kind == 0) lo[e] = hi[e] = side ? n : 0;
    else lo[e] = hi[e] = side ? n : -1;
    OGS_BOX(dim, lo, hi, idx) {
      const size_t q = ogs_at(L, idx);
      if (kind == 0) {
        w[q] = 0.0;
      } else {
        const double inner = side ? w[q - L->s[e]] : w[q + L->s[e]];
        w[q] = kind == 2 ? -inner : (side ? g2_hi : g2_lo) - inner;
      }
    }
  }
}

void ogs_solver_apply_bc(ogs_solver *s, int lvl) {
  ogs_level *L = ogs_lv(s, lvl);
  const int n = L->n, dim = s->cfg.dim;
  for (int d = 0; d < dim; ++d) {
    double *w = L->f[OGS_VEL + d];
    if (d == 0 && lvl == s->cfg.max_level) {
      /* ApplyBC_u: ghost [0,1] = 2 topVel - inner, ghost [0,-1] = - inner (x ghost to ghost, z ghost to ghost), then
       * dup [-1,0] and dup [1,0] = 0 on the inner rows */
      int tlo[3] = {-1, 0, -1}, thi[3] = {n + 1, 0, n};
      {
        int lo[3] = {tlo[0], n, tlo[2]}, hi[3] = {thi[0], n, thi[2]}, idx[3];
        OGS_BOX(dim, lo, hi, idx) { const size_t q = ogs_at(L, idx); w[q] = 2.0 * s->cfg.top_vel - w[q - L->s[1]]; }
        lo[1] = hi[1] = -1;
        OGS_BOX(dim, lo, hi, idx) { const size_t q = ogs_at(L, idx); w[q] = -w[q + L->s[1]]; }
      }
      int wlo[3] = {0, 0, 0}, whi[3] = {0, n - 1, n - 1};
      ogs_bc_faces(L, dim, w, 0, 0, 0.0, 0.0, wlo, whi);
      continue;
    }
    for (int e = 0; e < dim; ++e) { /* Dirichlet 0.0, faces -x,+x,-y,+y,-z,+z */
      int tlo[3], thi[3];
      for (int t = 0; t < 3; ++t) { tlo[t] = -1; thi[t] = t == d ? n + 1 : n; }
      if (e == d) ogs_bc_faces(L, dim, w, e, 0, 0.0, 0.0, tlo, thi);
      else ogs_bc_faces(L, dim, w, e, 1, 2.0 * 0.0, 2.0 * 0.0, tlo, thi);
    }
  }
  /* p: None */
}

/* Laplace * w at q: acc = c0*w; fma in declaration order -x, +x, -y, +y[, -z, +z] */
static inline double ogs_lap(const ogs_level *L, int dim, const double *w, size_t q, double c0, double cx) {
  double acc = c0 * w[q];
  for (int e = 0; e < dim; ++e) {
    acc = fma(cx, w[q - L->s[e]], acc);
    acc = fma(cx, w[q + L->s[e]], acc);
  }
  return acc;
}

static inline double ogs_c0(int dim, double h) {
  double c0 = 2.0 / (h * h) + 2.0 / (h * h);
  if (dim == 3) c0 = c0 + 2.0 / (h * h);
  return c0;
}

/* one colour of the smoother: cells with idx[d] % 3 == col[d].
 * Unknowns 2d, 2d+1 = the faces of the cell in dimension d (at the cell, one further), last = p.
 * Row 2d:   c0 on itself, cx on the face one further, -1/h on p;  rhs = rhs_d - cx * (the other neighbours, in the
 *           stencil's declaration order) - (1/h) p(one back)        [ Laplace * vel - d?Left * p == rhs ]
 * Row 2d+1: cx on the face at the cell, c0 on itself, +1/h on p;     rhs = rhs_d - cx * (others) + (1/h) p(one further)
 * Last row: -1/h, +1/h per dimension; rhs = rhs_p.
 * Gaussian elimination without pivoting in row order, back substitution. */
void ogs_smooth_color(ogs_solver *s, int lvl, int col0, int col1, int col2) {
  ogs_level *L = ogs_lv(s, lvl);
  const int n = L->n, dim = s->cfg.dim, NU = 2 * dim + 1;
  const double h = L->h;
  const double c0 = ogs_c0(dim, h), cx = -1.0 / (h * h), ih = 1.0 / h;
  double *p = L->f[OGS_P];
  const int col[3] = {col0, col1, col2};
  const int k0 = dim == 3 ? col[2] : 0, k1 = dim == 3 ? n : 1;
#pragma omp parallel for schedule(static) collapse(2)
  for (int kk = k0; kk < k1; kk += 3)
    for (int jj = col[1]; jj < n; jj += 3)
      for (int ii = col[0]; ii < n; ii += 3) {
        const int idx[3] = {ii, jj, kk};
        const size_t q = ogs_at(L, idx);
        double A[7][7], f[7], cur[7];
        int ok[7];
        for (int r = 0; r < NU; ++r)
          for (int c = 0; c < NU; ++c) A[r][c] = 0.0;
        for (int d = 0; d < dim; ++d) {
          double *w = L->f[OGS_VEL + d];
          const double *fw = L->f[OGS_RHS + d];
          const size_t sd = L->s[d];
          const int a = 2 * d, b = 2 * d + 1;
          A[a][a] = c0; A[a][b] = cx; A[a][NU - 1] = -ih;
          A[b][a] = cx; A[b][b] = c0; A[b][NU - 1] = ih;
          A[NU - 1][a] = -ih; A[NU - 1][b] = ih;
          ok[a] = idx[d] >= 1;
          ok[b] = idx[d] + 1 <= n - 1;
          cur[a] = w[q];
          cur[b] = w[q + sd];
          double fa = fw[q], fb = fw[q + sd];
          for (int e = 0; e < dim; ++e) {
            const size_t se = L->s[e];
            fa = fma(-cx, w[q - se], fa);
            if (e != d) fa = fma(-cx, w[q + se], fa);
            if (e != d) fb = fma(-cx, w[q + sd - se], fb);
            fb = fma(-cx, w[q + sd + se], fb);
          }
          f[a] = fma(-ih, p[q - sd], fa);
          f[b] = fma(ih, p[q + sd], fb);
        }
        f[NU - 1] = L->f[OGS_RHS_P][q];
        ok[NU - 1] = 1;
        cur[NU - 1] = p[q];
        for (int k = 0; k < NU; ++k)
          if (!ok[k]) {
            for (int c = 0; c < NU; ++c) A[k][c] = c == k ? 1.0 : 0.0;
            f[k] = cur[k];
          }
        for (int k = 0; k < NU; ++k)
          for (int r = k + 1; r < NU; ++r) {
            const double m = A[r][k] / A[k][k];
            for (int c = k + 1; c < NU; ++c) A[r][c] = fma(-m, A[k][c], A[r][c]);
            f[r] = fma(-m, f[k], f[r]);
          }
        double x[7];
        for (int r = NU - 1; r >= 0; --r) {
          double t = f[r];
          for (int c = r + 1; c < NU; ++c) t = fma(-A[r][c], x[c], t);
          x[r] = t / A[r][r];
        }
        for (int d = 0; d < dim; ++d) {
          double *w = L->f[OGS_VEL + d];
          if (ok[2 * d]) w[q] = x[2 * d];
          if (ok[2 * d + 1]) w[q + L->s[d]] = x[2 * d + 1];
        }
        p[q] = x[NU - 1];
      }
}

void ogs_solver_smooth(ogs_solver *s, int lvl, int nsweeps) {
  const int nz = s->cfg.dim == 3 ? 3 : 1;
  for (int it = 0; it < nsweeps; ++it) {
    for (int c2 = 0; c2 < nz; ++c2)
      for (int c1 = 0; c1 < 3; ++c1)
        for (int c0 = 0; c0 < 3; ++c0) ogs_smooth_color(s, lvl, c0, c1, c2);
    ogs_solver_apply_bc(s, lvl);
  }
}

/* UpdateResidual_*: residual_d = rhs_d - ( Laplace * vel_d - d?Left * p ) on the inner faces;
 * residual_p = rhs_p - ( dxRight * u + dyRight * v [+ dzRight * w] ) */
void ogs_solver_residual(ogs_solver *s, int lvl) {
  ogs_level *L = ogs_lv(s, lvl);
  const int n = L->n, dim = s->cfg.dim;
  const double h = L->h;
  const double c0 = ogs_c0(dim, h), cx = -1.0 / (h * h), ih = 1.0 / h;
  const double *p = L->f[OGS_P];
  const int lo[3] = {0, 0, 0}, hi[3] = {n - 1, n - 1, n - 1};
  int idx[3];
  OGS_BOX(dim, lo, hi, idx) {
    const size_t q = ogs_at(L, idx);
    double div = 0.0;
    for (int d = 0; d < dim; ++d) {
      const double *w = L->f[OGS_VEL + d];
      const size_t sd = L->s[d];
      if (idx[d] >= 1) {
        const double dl = fma(ih, p[q], (-ih) * p[q - sd]);
        L->f[OGS_RES + d][q] = L->f[OGS_RHS + d][q] - (ogs_lap(L, dim, w, q, c0, cx) - dl);
      }
      const double dr = fma(ih, w[q + sd], (-ih) * w[q]);
      div = d == 0 ? dr : div + dr;
    }
    L->f[OGS_RES_P][q] = L->f[OGS_RHS_P][q] - div;
  }
}

/* sqrt of the sum of squares over the loop points of each residual field (u, v[, w], p); double-double sums */
void ogs_solver_norms(ogs_solver *s, int lvl, double out[4]) {
  ogs_level *L = ogs_lv(s, lvl);
  const int n = L->n, dim = s->cfg.dim;
  ogs_dd acc[4] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}};
  const int lo[3] = {0, 0, 0}, hi[3] = {n - 1, n - 1, n - 1};
  int idx[3];
  OGS_BOX(dim, lo, hi, idx) {
    const size_t q = ogs_at(L, idx);
    for (int d = 0; d < dim; ++d)
      if (idx[d] >= 1) { const double r = L->f[OGS_RES + d][q]; ogs_dd_add(&acc[d], r * r); }
    const double r = L->f[OGS_RES_P][q];
    ogs_dd_add(&acc[dim], r * r);
  }
  for (int k = 0; k <= dim; ++k) out[k] = sqrt(acc[k].hi + acc[k].lo);
}

/* value of Restriction_<loc> * src at coarse index I: kron of [1/4 1/2 1/4] (fine 2I-1..2I+1) in dimension `own`
 * (own < 0: none) and [1/2 1/2] (fine 2I, 2I+1) in the others; entries in lexicographic order (x fastest), fma chain */
static double ogs_restrict_point(const ogs_level *F, int dim, const double *src, const int I[3], int own) {
  int lo[3], cnt[3];
  for (int t = 0; t < 3; ++t) { lo[t] = t == own ? -1 : 0; cnt[t] = t < dim ? (t == own ? 3 : 2) : 1; }
  double acc = 0.0;
  int first = 1;
  for (int c = 0; c < cnt[2]; ++c)
    for (int b = 0; b < cnt[1]; ++b)
      for (int a = 0; a < cnt[0]; ++a) {
        const int off[3] = {lo[0] + a, lo[1] + b, lo[2] + c};
        double wgt = 1.0;
        int fi[3] = {0, 0, 0};
        for (int t = 0; t < dim; ++t) {
          wgt *= t == own ? (off[t] == 0 ? 0.5 : 0.25) : 0.5;
          fi[t] = 2 * I[t] + off[t];
        }
        const double v = src[ogs_at(F, fi)];
        acc = first ? wgt * v : fma(wgt, v, acc);
        first = 0;
      }
  return acc;
}

/* Restriction_*: rhs@coarser = Restriction * residual */
void ogs_solver_restrict(ogs_solver *s, int lvl) {
  ogs_level *F = ogs_lv(s, lvl), *C = ogs_lv(s, lvl - 1);
  const int nc = C->n, dim = s->cfg.dim;
  const int lo[3] = {0, 0, 0}, hi[3] = {nc - 1, nc - 1, nc - 1};
  int I[3];
  OGS_BOX(dim, lo, hi, I) {
    const size_t qc = ogs_at(C, I);
    for (int d = 0; d < dim; ++d)
      if (I[d] >= 1) C->f[OGS_RHS + d][qc] = ogs_restrict_point(F, dim, F->f[OGS_RES + d], I, d);
    C->f[OGS_RHS_P][qc] = ogs_restrict_point(F, dim, F->f[OGS_RES_P], I, -1);
  }
}

/* SetSolution@lvl ( value ): loop over target starting [-1,..] ending [-1,..] -- the loop points and one layer around */
void ogs_solver_set_solution(ogs_solver *s, int lvl, double value) {
  ogs_level *L = ogs_lv(s, lvl);
  const int n = L->n, dim = s->cfg.dim;
  const int lo[3] = {-1, -1, -1}, hi[3] = {n, n, n};
  int idx[3];
  OGS_BOX(dim, lo, hi, idx) {
    const size_t q = ogs_at(L, idx);
    for (int d = 0; d < dim; ++d)
      if (idx[d] >= 0) L->f[OGS_VEL + d][q] = value;
    L->f[OGS_P][q] = value;
  }
}

/* Correction_*: dest += Correction * source@coarser (linear in the component's own dimension, constant in the others;
 * p: the parent's value); apply bc */
void ogs_solver_correct(ogs_solver *s, int lvl) {
  ogs_level *F = ogs_lv(s, lvl), *C = ogs_lv(s, lvl - 1);
  const int n = F->n, dim = s->cfg.dim;
  const int lo[3] = {0, 0, 0}, hi[3] = {n - 1, n - 1, n - 1};
  int idx[3];
  OGS_BOX(dim, lo, hi, idx) {
    const size_t q = ogs_at(F, idx);
    const int ci[3] = {idx[0] >> 1, idx[1] >> 1, idx[2] >> 1};
    const size_t qc = ogs_at(C, ci);
    for (int d = 0; d < dim; ++d)
      if (idx[d] >= 1) {
        const double *wc = C->f[OGS_VEL + d];
        double *w = &F->f[OGS_VEL + d][q];
        *w = (idx[d] & 1) ? fma(0.5, wc[qc] + wc[qc + C->s[d]], *w) : fma(1.0, wc[qc], *w);
      }
    F->f[OGS_P][q] = fma(1.0, C->f[OGS_P][qc], F->f[OGS_P][q]);
  }
  ogs_solver_apply_bc(s, lvl);
}

void ogs_solver_vcycle(ogs_solver *s, int lvl) {
  if (lvl == s->cfg.min_level) {
    ogs_solver_smooth(s, lvl, s->cfg.nu_coarse);
    return;
  }
  ogs_solver_smooth(s, lvl, s->cfg.nu_pre);
  ogs_solver_residual(s, lvl);
  ogs_solver_restrict(s, lvl);
  ogs_solver_set_solution(s, lvl - 1, 0.0);
  ogs_solver_vcycle(s, lvl - 1);
  ogs_solver_correct(s, lvl);
  ogs_solver_smooth(s, lvl, s->cfg.nu_post);
}
