/* examg_oracle_stokes.c -- CPU restatement of the staggered-grid Stokes program of the reference: TEST INFRASTRUCTURE ONLY
 * (only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline legs may use anything under oracle/).
 *
 * Follows Testing/Application/ExaStokes_2D.exa4 (lid-driven cavity on the unit square, levels 2..8):
 *   fields      u at Face_x (nodes in x, cells in y), v at Face_y, p at cell centres         ...exa4:3-45
 *   operators   Laplace (5-point, declaration order centre,-x,+x,-y,+y), dxLeft/dxRight/dyLeft/dyRight  ...exa4:49-77
 *   smoother    color with { i0 % 3, i1 % 3 } loop over p { solve locally 5 unknowns: u@[0,0], u@[1,0], v@[0,0],
 *               v@[0,1], p@[0,0] }, then apply bc to u, v, p                                 ...exa4:228-303
 *               local system: C/s/e/solver/ir/IR_LocalSolve.scala:88-236 (terms with unknowns -> matrix, the rest ->
 *               right-hand side), rows of unknowns that are no valid computation points (wall faces) are identities and
 *               are not written back: IR_LocalDirectInvert.scala:76-103,127-147, IR_IsOnBoundary.scala:53-97
 *   loops       node-like dimensions skip the boundary node, cell-like ones run over all cells
 *               (C/s/e/baseExt/ir/IR_LoopOverPointsInOneFragment.scala:76-112)
 *   apply bc    Dirichlet on a face field: value on the wall faces of its own dimension, ghost = 2 g - inner in the others;
 *               tangential range ghost to ghost, faces in the order -x,+x,-y,+y
 *               (C/s/e/boundary/ir/IR_BoundaryCondition.scala:38-45, IR_ApplyBCFunction.scala:55-88, IR_DirichletBC.scala:37-50)
 *   transfer    default 'linear' restriction = kron of [1/4,1/2,1/4] (node-like) and [1/2,1/2] (cell-like), prolongation =
 *               2^dim * transpose  (C/s/e/operator/l4/L4_DefaultRestriction.scala:29-47,79-100, L4_DefaultProlongation.scala:30-45)
 *   cycle       V(3,3), 10 smoother steps on the coarsest level, SetSolution@coarser(0) incl. one layer around  ...exa4:206-226
 * Pinned by tests/test_oracle_golden.py::test_golden_stokes_2d against Testing/Application/ExaStokes_2D.results.
 *
 * Storage: every field is an (n+3) x (n+3) array, entry (i, j) at (i+1) + (n+3) (j+1); u uses i = -1..n+1, j = -1..n;
 * v uses i = -1..n, j = -1..n+1; p uses -1..n in both.
 * Arithmetic contract (what libexamg's kernels reproduce bit for bit; built with -ffp-contract=off): stated at each loop.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  int min_level, max_level;
  int nu_pre, nu_post, nu_coarse;
  double top_vel;
} ogs_config;

enum { OGS_U = 0, OGS_V, OGS_P, OGS_RHS_U, OGS_RHS_V, OGS_RHS_P, OGS_RES_U, OGS_RES_V, OGS_RES_P, OGS_NFIELDS };

typedef struct {
  int n;      /* cells per dimension */
  int pitch;  /* n + 3 */
  double h;
  double *f[OGS_NFIELDS];
} ogs_level;

typedef struct ogs_solver {
  ogs_config cfg;
  ogs_level *lev;
} ogs_solver;

#define OGS_AT(L, i, j) ((size_t)((i) + 1) + (size_t)(L)->pitch * (size_t)((j) + 1))

static ogs_level *ogs_lv(ogs_solver *s, int lvl) { return &s->lev[lvl - s->cfg.min_level]; }

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
    L->pitch = L->n + 3;
    L->h = 1.0 / (double)L->n;
    for (int k = 0; k < OGS_NFIELDS; ++k) L->f[k] = (double *)calloc((size_t)L->pitch * (size_t)L->pitch, sizeof(double));
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

/* apply bc to u: finest = ApplyBC_u (lid), coarser = Dirichlet 0.0 */
static void ogs_bc_u(ogs_solver *s, int lvl) {
  ogs_level *L = ogs_lv(s, lvl);
  const int n = L->n;
  double *u = L->f[OGS_U];
  if (lvl == s->cfg.max_level) {
    const double top = s->cfg.top_vel;
    for (int i = -1; i <= n + 1; ++i) u[OGS_AT(L, i, n)] = 2.0 * top - u[OGS_AT(L, i, n - 1)];
    for (int i = -1; i <= n + 1; ++i) u[OGS_AT(L, i, -1)] = -u[OGS_AT(L, i, 0)];
    for (int j = 0; j < n; ++j) u[OGS_AT(L, 0, j)] = 0.0;
    for (int j = 0; j < n; ++j) u[OGS_AT(L, n, j)] = 0.0;
  } else {
    for (int j = -1; j <= n; ++j) u[OGS_AT(L, 0, j)] = 0.0;
    for (int j = -1; j <= n; ++j) u[OGS_AT(L, n, j)] = 0.0;
    for (int i = -1; i <= n + 1; ++i) u[OGS_AT(L, i, -1)] = (2.0 * 0.0) - u[OGS_AT(L, i, 0)];
    for (int i = -1; i <= n + 1; ++i) u[OGS_AT(L, i, n)] = (2.0 * 0.0) - u[OGS_AT(L, i, n - 1)];
  }
}

/* apply bc to v: Dirichlet 0.0 on every level */
static void ogs_bc_v(ogs_solver *s, int lvl) {
  ogs_level *L = ogs_lv(s, lvl);
  const int n = L->n;
  double *v = L->f[OGS_V];
  for (int j = -1; j <= n + 1; ++j) v[OGS_AT(L, -1, j)] = (2.0 * 0.0) - v[OGS_AT(L, 0, j)];
  for (int j = -1; j <= n + 1; ++j) v[OGS_AT(L, n, j)] = (2.0 * 0.0) - v[OGS_AT(L, n - 1, j)];
  for (int i = -1; i <= n; ++i) v[OGS_AT(L, i, 0)] = 0.0;
  for (int i = -1; i <= n; ++i) v[OGS_AT(L, i, n)] = 0.0;
}

void ogs_solver_apply_bc(ogs_solver *s, int lvl) {
  ogs_bc_u(s, lvl);
  ogs_bc_v(s, lvl);
  /* p: None */
}

/* Laplace * w at q: acc = c0*w; fma in declaration order -x, +x, -y, +y */
static inline double ogs_lap(const double *w, size_t q, size_t sy, double c0, double cx) {
  double acc = c0 * w[q];
  acc = fma(cx, w[q - 1], acc);
  acc = fma(cx, w[q + 1], acc);
  acc = fma(cx, w[q - sy], acc);
  acc = fma(cx, w[q + sy], acc);
  return acc;
}

/* one colour of the smoother: cells with i % 3 == c0 and j % 3 == c1 */
void ogs_smooth_color(ogs_solver *s, int lvl, int col0, int col1) {
  ogs_level *L = ogs_lv(s, lvl);
  const int n = L->n;
  const size_t sy = (size_t)L->pitch;
  const double h = L->h;
  const double c0 = 2.0 / (h * h) + 2.0 / (h * h), cx = -1.0 / (h * h), ih = 1.0 / h;
  double *u = L->f[OGS_U], *v = L->f[OGS_V], *p = L->f[OGS_P];
  const double *fu = L->f[OGS_RHS_U], *fv = L->f[OGS_RHS_V], *fp = L->f[OGS_RHS_P];
#pragma omp parallel for schedule(static)
  for (int j = col1; j < n; j += 3)
    for (int i = col0; i < n; i += 3) {
      const size_t q = OGS_AT(L, i, j);
      /* unknowns: 0 u(i,j), 1 u(i+1,j), 2 v(i,j), 3 v(i,j+1), 4 p(i,j); valid = not a wall face */
      const int ok[5] = {i >= 1, i + 1 <= n - 1, j >= 1, j + 1 <= n - 1, 1};
      double A[5][5] = {{c0, cx, 0, 0, -ih}, {cx, c0, 0, 0, ih}, {0, 0, c0, cx, -ih}, {0, 0, cx, c0, ih}, {-ih, ih, -ih, ih, 0}};
      double f[5];
      f[0] = fu[q];
      f[0] = fma(-cx, u[q - 1], f[0]);
      f[0] = fma(-cx, u[q - sy], f[0]);
      f[0] = fma(-cx, u[q + sy], f[0]);
      f[0] = fma(-ih, p[q - 1], f[0]);
      f[1] = fu[q + 1];
      f[1] = fma(-cx, u[q + 2], f[1]);
      f[1] = fma(-cx, u[q + 1 - sy], f[1]);
      f[1] = fma(-cx, u[q + 1 + sy], f[1]);
      f[1] = fma(ih, p[q + 1], f[1]);
      f[2] = fv[q];
      f[2] = fma(-cx, v[q - 1], f[2]);
      f[2] = fma(-cx, v[q + 1], f[2]);
      f[2] = fma(-cx, v[q - sy], f[2]);
      f[2] = fma(-ih, p[q - sy], f[2]);
      f[3] = fv[q + sy];
      f[3] = fma(-cx, v[q + sy - 1], f[3]);
      f[3] = fma(-cx, v[q + sy + 1], f[3]);
      f[3] = fma(-cx, v[q + 2 * sy], f[3]);
      f[3] = fma(ih, p[q + sy], f[3]);
      f[4] = fp[q];
      const double cur[5] = {u[q], u[q + 1], v[q], v[q + sy], p[q]};
      for (int k = 0; k < 5; ++k)
        if (!ok[k]) {
          for (int c = 0; c < 5; ++c) A[k][c] = c == k ? 1.0 : 0.0;
          f[k] = cur[k];
        }
      /* Gaussian elimination without pivoting, rows in order; x by back substitution */
      for (int k = 0; k < 5; ++k)
        for (int r = k + 1; r < 5; ++r) {
          const double m = A[r][k] / A[k][k];
          for (int c = k + 1; c < 5; ++c) A[r][c] = fma(-m, A[k][c], A[r][c]);
          f[r] = fma(-m, f[k], f[r]);
        }
      double x[5];
      for (int r = 4; r >= 0; --r) {
        double t = f[r];
        for (int c = r + 1; c < 5; ++c) t = fma(-A[r][c], x[c], t);
        x[r] = t / A[r][r];
      }
      if (ok[0]) u[q] = x[0];
      if (ok[1]) u[q + 1] = x[1];
      if (ok[2]) v[q] = x[2];
      if (ok[3]) v[q + sy] = x[3];
      p[q] = x[4];
    }
}

void ogs_solver_smooth(ogs_solver *s, int lvl, int nsweeps) {
  for (int it = 0; it < nsweeps; ++it) {
    for (int c1 = 0; c1 < 3; ++c1)
      for (int c0 = 0; c0 < 3; ++c0) ogs_smooth_color(s, lvl, c0, c1);
    ogs_solver_apply_bc(s, lvl);
  }
}

/* UpdateResidual_{u,v,p} */
void ogs_solver_residual(ogs_solver *s, int lvl) {
  ogs_level *L = ogs_lv(s, lvl);
  const int n = L->n;
  const size_t sy = (size_t)L->pitch;
  const double h = L->h;
  const double c0 = 2.0 / (h * h) + 2.0 / (h * h), cx = -1.0 / (h * h), ih = 1.0 / h;
  const double *u = L->f[OGS_U], *v = L->f[OGS_V], *p = L->f[OGS_P];
#pragma omp parallel for schedule(static)
  for (int j = 0; j < n; ++j) {
    for (int i = 1; i < n; ++i) {  /* residual_u = rhs_u - ( Laplace * u - dxLeft * p ) */
      const size_t q = OGS_AT(L, i, j);
      const double dx = fma(ih, p[q], (-ih) * p[q - 1]);
      L->f[OGS_RES_U][q] = L->f[OGS_RHS_U][q] - (ogs_lap(u, q, sy, c0, cx) - dx);
    }
    if (j >= 1)
      for (int i = 0; i < n; ++i) {  /* residual_v = rhs_v - ( Laplace * v - dyLeft * p ) */
        const size_t q = OGS_AT(L, i, j);
        const double dy = fma(ih, p[q], (-ih) * p[q - sy]);
        L->f[OGS_RES_V][q] = L->f[OGS_RHS_V][q] - (ogs_lap(v, q, sy, c0, cx) - dy);
      }
    for (int i = 0; i < n; ++i) {  /* residual_p = rhs_p - ( dxRight * u + dyRight * v ) */
      const size_t q = OGS_AT(L, i, j);
      const double dx = fma(ih, u[q + 1], (-ih) * u[q]);
      const double dy = fma(ih, v[q + sy], (-ih) * v[q]);
      L->f[OGS_RES_P][q] = L->f[OGS_RHS_P][q] - (dx + dy);
    }
  }
}

/* sqrt of the sum of squares over the loop points of each residual field; double-double sums, rows in order */
void ogs_solver_norms(ogs_solver *s, int lvl, double out[3]) {
  ogs_level *L = ogs_lv(s, lvl);
  const int n = L->n;
  ogs_dd a = {0, 0}, b = {0, 0}, c = {0, 0};
  for (int j = 0; j < n; ++j) {
    for (int i = 1; i < n; ++i) { const double r = L->f[OGS_RES_U][OGS_AT(L, i, j)]; ogs_dd_add(&a, r * r); }
    if (j >= 1)
      for (int i = 0; i < n; ++i) { const double r = L->f[OGS_RES_V][OGS_AT(L, i, j)]; ogs_dd_add(&b, r * r); }
    for (int i = 0; i < n; ++i) { const double r = L->f[OGS_RES_P][OGS_AT(L, i, j)]; ogs_dd_add(&c, r * r); }
  }
  out[0] = sqrt(a.hi + a.lo);
  out[1] = sqrt(b.hi + b.lo);
  out[2] = sqrt(c.hi + c.lo);
}

/* Restriction_{u,v,p}: rhs@coarser = Restriction * residual.  Entries in lexicographic order (x fastest), fma chain. */
void ogs_solver_restrict(ogs_solver *s, int lvl) {
  ogs_level *F = ogs_lv(s, lvl), *C = ogs_lv(s, lvl - 1);
  const int nc = C->n;
  const size_t fy = (size_t)F->pitch;
  const double *ru = F->f[OGS_RES_U], *rv = F->f[OGS_RES_V], *rp = F->f[OGS_RES_P];
#pragma omp parallel for schedule(static)
  for (int J = 0; J < nc; ++J) {
    for (int I = 1; I < nc; ++I) {  /* Face_x: [1/4 1/2 1/4] in x (fine 2I-1..2I+1), [1/2 1/2] in y (fine 2J, 2J+1) */
      const size_t q = OGS_AT(F, 2 * I, 2 * J);
      double acc = 0.125 * ru[q - 1];
      acc = fma(0.25, ru[q], acc);
      acc = fma(0.125, ru[q + 1], acc);
      acc = fma(0.125, ru[q + fy - 1], acc);
      acc = fma(0.25, ru[q + fy], acc);
      acc = fma(0.125, ru[q + fy + 1], acc);
      C->f[OGS_RHS_U][OGS_AT(C, I, J)] = acc;
    }
    if (J >= 1)
      for (int I = 0; I < nc; ++I) {  /* Face_y: [1/2 1/2] in x (fine 2I, 2I+1), [1/4 1/2 1/4] in y (fine 2J-1..2J+1) */
        const size_t q = OGS_AT(F, 2 * I, 2 * J);
        double acc = 0.125 * rv[q - fy];
        acc = fma(0.125, rv[q - fy + 1], acc);
        acc = fma(0.25, rv[q], acc);
        acc = fma(0.25, rv[q + 1], acc);
        acc = fma(0.125, rv[q + fy], acc);
        acc = fma(0.125, rv[q + fy + 1], acc);
        C->f[OGS_RHS_V][OGS_AT(C, I, J)] = acc;
      }
    for (int I = 0; I < nc; ++I) {  /* Cell: the mean of the four children */
      const size_t q = OGS_AT(F, 2 * I, 2 * J);
      double acc = 0.25 * rp[q];
      acc = fma(0.25, rp[q + 1], acc);
      acc = fma(0.25, rp[q + fy], acc);
      acc = fma(0.25, rp[q + fy + 1], acc);
      C->f[OGS_RHS_P][OGS_AT(C, I, J)] = acc;
    }
  }
}

/* SetSolution@lvl ( value ): loop over target starting [-1,-1] ending [-1,-1] -- the loop points and one layer around */
void ogs_solver_set_solution(ogs_solver *s, int lvl, double value) {
  ogs_level *L = ogs_lv(s, lvl);
  const int n = L->n;
  for (int j = -1; j <= n; ++j)
    for (int i = 0; i <= n; ++i) L->f[OGS_U][OGS_AT(L, i, j)] = value;
  for (int j = 0; j <= n; ++j)
    for (int i = -1; i <= n; ++i) L->f[OGS_V][OGS_AT(L, i, j)] = value;
  for (int j = -1; j <= n; ++j)
    for (int i = -1; i <= n; ++i) L->f[OGS_P][OGS_AT(L, i, j)] = value;
}

/* Correction_{u,v,p}: dest += Correction * source@coarser; apply bc */
void ogs_solver_correct(ogs_solver *s, int lvl) {
  ogs_level *F = ogs_lv(s, lvl), *C = ogs_lv(s, lvl - 1);
  const int n = F->n;
  const size_t cy = (size_t)C->pitch;
  const double *uc = C->f[OGS_U], *vc = C->f[OGS_V], *pc = C->f[OGS_P];
#pragma omp parallel for schedule(static)
  for (int j = 0; j < n; ++j) {
    for (int i = 1; i < n; ++i) {  /* Face_x: linear in x, constant in y */
      const size_t q = OGS_AT(F, i, j), qc = OGS_AT(C, i >> 1, j >> 1);
      double *d = &F->f[OGS_U][q];
      *d = (i & 1) ? fma(0.5, uc[qc] + uc[qc + 1], *d) : fma(1.0, uc[qc], *d);
    }
    if (j >= 1)
      for (int i = 0; i < n; ++i) {  /* Face_y: constant in x, linear in y */
        const size_t q = OGS_AT(F, i, j), qc = OGS_AT(C, i >> 1, j >> 1);
        double *d = &F->f[OGS_V][q];
        *d = (j & 1) ? fma(0.5, vc[qc] + vc[qc + cy], *d) : fma(1.0, vc[qc], *d);
      }
    for (int i = 0; i < n; ++i) {  /* Cell: the parent's value */
      const size_t q = OGS_AT(F, i, j), qc = OGS_AT(C, i >> 1, j >> 1);
      F->f[OGS_P][q] = fma(1.0, pc[qc], F->f[OGS_P][q]);
    }
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
