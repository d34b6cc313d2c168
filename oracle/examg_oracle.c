/*
 * examg_oracle.c -- CPU restatement of the ExaStencils-generated V-cycle.
 * TEST INFRASTRUCTURE ONLY (see examg_oracle.h).  Build: oracle/Makefile
 * (gcc -O3 -fopenmp -ffp-contract=off: only the explicit fma() calls fuse).
 *
 * Parallelisation mirrors the reference OpenMP build of the benchmark
 * (omp_parallelizeLoopOverDimensions = true -> `#pragma omp parallel for
 * schedule(static)` on the outermost dimension,
 * C/s/e/parallelization/api/omp/OMP_Loop.scala:101-110).
 */
#include "examg_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---------- helpers ------------------------------------------------------ */

static inline size_t og_size(const int n[3]) {
  return (size_t)n[0] * (size_t)n[1] * (size_t)n[2];
}

/* domain_rect_periodic_{x,y,z} (Knowledge; Testing/BC/{2D,3D}_Periodic.knowledge).  In a periodic dimension the
 * first and the last node are the same point: the boundary fragments are each other's neighbours, so loops cover
 * both copies (iteration offsets 0, C/s/e/domain/ir/IR_ConnectFragments.scala:60-93), reductions count the point
 * once (i >= numDupLeft, C/s/e/baseExt/ir/IR_LoopOverPointsInOneFragment.scala:116-125), and a stencil leaving the
 * array on one side continues one node inside the other (`communicate ghost`).  Process-wide switch of the oracle. */
static int og_per[3] = {0, 0, 0};
void og_set_periodic(const int per[3]) { for (int d = 0; d < 3; ++d) og_per[d] = per[d]; }
static inline int og_per_any(void) { return og_per[0] | og_per[1] | og_per[2]; }

/* loop range of dimension d: 1..n-2 (0..n-1 if periodic), or the single plane 0 if n == 1 (2D) */
static inline void og_range(const int n[3], int d, int *lo, int *hi) {
  if (n[d] == 1) { *lo = 0; *hi = 1; }
  else if (og_per[d]) { *lo = 0; *hi = n[d]; }
  else { *lo = 1; *hi = n[d] - 1; }
}
/* reductions: a periodic dimension counts its duplicated point once */
static inline void og_range_red(const int n[3], int d, int *lo, int *hi) {
  og_range(n, d, lo, hi);
  if (n[d] > 1 && og_per[d]) *lo = 1;
}
/* neighbour index in dimension d with the periodic continuation */
static inline int og_wrap(const int n[3], int d, int i) {
  if (n[d] > 1 && og_per[d]) { if (i < 0) return n[d] - 2 - (-1 - i); if (i > n[d] - 1) return i - (n[d] - 1); }
  return i;
}

/* linear offsets of the stencil entries */
static void og_lin_offsets(const int n[3], const og_stencil *st, ptrdiff_t *lo) {
  for (int k = 0; k < st->nst; ++k)
    lo[k] = (ptrdiff_t)st->off[k][0] +
            (ptrdiff_t)n[0] * ((ptrdiff_t)st->off[k][1] +
                               (ptrdiff_t)n[1] * (ptrdiff_t)st->off[k][2]);
}

/* index of the centre entry (diag(), C/s/e/stencil/ir/IR_StencilFunction.scala:46-60) */
static int og_centre(const og_stencil *st) {
  for (int k = 0; k < st->nst; ++k)
    if (!st->off[k][0] && !st->off[k][1] && !st->off[k][2]) return k;
  fprintf(stderr, "og: stencil without centre entry\n");
  exit(1);
}

/* canonical A*u at node p */
static inline double og_conv(int nst, const ptrdiff_t *lo, const double *c,
                             const double *sf, size_t size, const double *u,
                             size_t p) {
  double acc;
  if (sf) {
    acc = sf[p] * u[(ptrdiff_t)p + lo[0]];
    for (int k = 1; k < nst; ++k)
      acc = fma(sf[(size_t)k * size + p], u[(ptrdiff_t)p + lo[k]], acc);
  } else {
    acc = c[0] * u[(ptrdiff_t)p + lo[0]];
    for (int k = 1; k < nst; ++k) acc = fma(c[k], u[(ptrdiff_t)p + lo[k]], acc);
  }
  return acc;
}

/* the same fold at a node on a periodic boundary: neighbours by wrapped coordinates */
static double og_conv_wrap(const int n[3], const og_stencil *st, const double *sf, size_t size, const double *u,
                           int i, int j, int k) {
  const size_t p = (size_t)i + (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
  double acc = 0.0;
  for (int e = 0; e < st->nst; ++e) {
    const int ii = og_wrap(n, 0, i + st->off[e][0]), jj = og_wrap(n, 1, j + st->off[e][1]),
              kk = og_wrap(n, 2, k + st->off[e][2]);
    const double v = u[(size_t)ii + (size_t)n[0] * ((size_t)jj + (size_t)n[1] * (size_t)kk)];
    const double c = sf ? sf[(size_t)e * size + p] : st->coef[e];
    acc = e == 0 ? c * v : fma(c, v, acc);
  }
  return acc;
}
static inline int og_on_per_boundary(const int n[3], int i, int j, int k) {
  return (og_per[0] && n[0] > 1 && (i == 0 || i == n[0] - 1)) || (og_per[1] && n[1] > 1 && (j == 0 || j == n[1] - 1)) ||
         (og_per[2] && n[2] > 1 && (k == 0 || k == n[2] - 1));
}
#define OG_CONV(i, j, k, p) \
  (og_on_per_boundary(n, i, j, k) ? og_conv_wrap(n, st, sf, size, u, i, j, k) : og_conv(nst, lo, st->coef, sf, size, u, p))

#define OG_LOOP_BEGIN(n)                                         \
  int xl, xh, yl, yh, zl, zh;                                    \
  og_range(n, 0, &xl, &xh);                                      \
  og_range(n, 1, &yl, &yh);                                      \
  og_range(n, 2, &zl, &zh);                                      \
  const int outer_is_z = (n[2] > 1);                             \
  (void)outer_is_z;

/* ---------- stencil loops ------------------------------------------------- */

void og_apply(const int n[3], const og_stencil *st, const double *sf,
              const double *u, double *out) {
  ptrdiff_t lo[OG_MAX_ST];
  og_lin_offsets(n, st, lo);
  const size_t size = og_size(n);
  const int nst = st->nst;
  OG_LOOP_BEGIN(n)
#pragma omp parallel for schedule(static) collapse(2)
  for (int k = zl; k < zh; ++k)
    for (int j = yl; j < yh; ++j) {
      size_t row = (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
      for (int i = xl; i < xh; ++i)
        out[row + i] = OG_CONV(i, j, k, row + i);
    }
}

void og_residual(const int n[3], const og_stencil *st, const double *sf,
                 const double *u, const double *rhs, double *res) {
  ptrdiff_t lo[OG_MAX_ST];
  og_lin_offsets(n, st, lo);
  const size_t size = og_size(n);
  const int nst = st->nst;
  OG_LOOP_BEGIN(n)
#pragma omp parallel for schedule(static) collapse(2)
  for (int k = zl; k < zh; ++k)
    for (int j = yl; j < yh; ++j) {
      size_t row = (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
      for (int i = xl; i < xh; ++i)
        res[row + i] = rhs[row + i] - OG_CONV(i, j, k, row + i);
    }
}

void og_smooth_jacobi(const int n[3], const og_stencil *st, const double *sf,
                      double omega, const double *u, const double *rhs,
                      double *unext) {
  ptrdiff_t lo[OG_MAX_ST];
  og_lin_offsets(n, st, lo);
  const size_t size = og_size(n);
  const int nst = st->nst;
  const int c = og_centre(st);
  const double wconst = omega / st->coef[c];
  OG_LOOP_BEGIN(n)
#pragma omp parallel for schedule(static) collapse(2)
  for (int k = zl; k < zh; ++k)
    for (int j = yl; j < yh; ++j) {
      size_t row = (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
      for (int i = xl; i < xh; ++i) {
        size_t p = row + i;
        double w = sf ? omega / sf[(size_t)c * size + p] : wconst;
        double r = rhs[p] - OG_CONV(i, j, k, p);
        unext[p] = fma(w, r, u[p]);
      }
    }
}

void og_smooth_rb(const int n[3], const og_stencil *st, const double *sf,
                  double omega, double *u, const double *rhs, int color) {
  ptrdiff_t lo[OG_MAX_ST];
  og_lin_offsets(n, st, lo);
  const size_t size = og_size(n);
  const int nst = st->nst;
  const int c = og_centre(st);
  const double wconst = omega / st->coef[c];
  OG_LOOP_BEGIN(n)
#pragma omp parallel for schedule(static) collapse(2)
  for (int k = zl; k < zh; ++k)
    for (int j = yl; j < yh; ++j) {
      size_t row = (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
      /* first i >= xl with (i + j + k) % 2 == color */
      int i0 = xl + (((xl + j + k) & 1) != color);
      for (int i = i0; i < xh; i += 2) {
        size_t p = row + i;
        double w = sf ? omega / sf[(size_t)c * size + p] : wconst;
        double r = rhs[p] - OG_CONV(i, j, k, p);
        u[p] = fma(w, r, u[p]);
      }
    }
}

/* One colour (cx, cy, cz) of an ncol^dim-way coloured Gauss-Seidel sweep: in place on the loop points with
 * i_d % ncol == c_d (solver_smoother_coloring "4-way" ... "27-way", C/s/e/solver/l3/L3_VankaForEquation.scala:154-183) */
void og_smooth_color(const int n[3], const og_stencil *st, const double *sf, double omega, double *u,
                     const double *rhs, int ncol, int cx, int cy, int cz) {
  ptrdiff_t lo[OG_MAX_ST];
  og_lin_offsets(n, st, lo);
  const size_t size = og_size(n);
  const int nst = st->nst;
  const int c = og_centre(st);
  const double wconst = omega / st->coef[c];
  OG_LOOP_BEGIN(n)
  for (int k = zl; k < zh; ++k) {
    if (n[2] > 1 && k % ncol != cz) continue;
    for (int j = yl; j < yh; ++j) {
      if (j % ncol != cy) continue;
      size_t row = (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
      for (int i = xl; i < xh; ++i) {
        if (i % ncol != cx) continue;
        size_t p = row + i;
        double w = sf ? omega / sf[(size_t)c * size + p] : wconst;
        double r = rhs[p] - OG_CONV(i, j, k, p);
        u[p] = fma(w, r, u[p]);
      }
    }
  }
}

/* ---------- reductions ---------------------------------------------------- */

/* double-double accumulator: (hi, lo) += x via Knuth TwoSum */
typedef struct { double hi, lo; } og_dd;

static inline void og_dd_add(og_dd *a, double x) {
  double s = a->hi + x;
  double bb = s - a->hi;
  double err = (a->hi - (s - bb)) + (x - bb);
  a->hi = s;
  a->lo += err;
}

static inline void og_dd_merge(og_dd *a, const og_dd *b) {
  og_dd_add(a, b->hi);
  a->lo += b->lo;
}

double og_dot(const int n[3], const double *a, const double *b) {
  int xl, xh, yl, yh, zl, zh;
  og_range_red(n, 0, &xl, &xh);
  og_range_red(n, 1, &yl, &yh);
  og_range_red(n, 2, &zl, &zh);
  og_dd total = {0.0, 0.0};
#pragma omp parallel
  {
    og_dd part = {0.0, 0.0};
#pragma omp for schedule(static) collapse(2) nowait
    for (int k = zl; k < zh; ++k)
      for (int j = yl; j < yh; ++j) {
        size_t row = (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
        for (int i = xl; i < xh; ++i) og_dd_add(&part, a[row + i] * b[row + i]);
      }
#pragma omp critical
    og_dd_merge(&total, &part);
  }
  return total.hi + total.lo;
}

void og_glibc_rand_fill(double *out, long long n) {
  srand(1); /* the state a fresh process starts with */
  for (long long i = 0; i < n; ++i) out[i] = ((double)rand() / RAND_MAX);
}

double og_max_abs_diff(const int n[3], const double *a, const double *b) {
  OG_LOOP_BEGIN(n)
  double m = 0.0;
#pragma omp parallel for schedule(static) collapse(2) reduction(max : m)
  for (int k = zl; k < zh; ++k)
    for (int j = yl; j < yh; ++j) {
      size_t row = (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
      for (int i = xl; i < xh; ++i) {
        double d = fabs(a[row + i] - b[row + i]);
        if (d > m) m = d;
      }
    }
  return m;
}

/* ---------- inter-grid transfer ------------------------------------------- */

/*
 * Full weighting: 1-D [1/4,1/2,1/4] Kronecker product
 * (L4_DefaultRestriction.scala:29-37,63-90).  The generator normalises the sum
 * into "coefficient x (sum of terms with that coefficient)"
 * (C/s/e/optimization/ir/IR_SimplifyExpression.scala:579-881); restated here as
 * groups by number of non-zero offsets m (weight 2^-(dim+m)), each group summed
 * in lexicographic (dz,dy,dx) order, then
 *   acc = w0*S0; acc = fma(w1,S1,acc); acc = fma(w2,S2,acc); acc = fma(w3,S3,acc).
 */
void og_restrict(int dim, const int nf[3], const int nc[3], double scale,
                 const double *rf, double *rhsc) {
  OG_LOOP_BEGIN(nc)
  const int dzr = (dim == 3) ? 1 : 0;
  double w[4];
  for (int m = 0; m <= 3; ++m) w[m] = scale * ldexp(1.0, -(dim + m));
#pragma omp parallel for schedule(static) collapse(2)
  for (int k = zl; k < zh; ++k)
    for (int j = yl; j < yh; ++j)
      for (int i = xl; i < xh; ++i) {
        double S[4] = {0.0, 0.0, 0.0, 0.0};
        int first[4] = {1, 1, 1, 1};
        for (int dz = -dzr; dz <= dzr; ++dz)
          for (int dy = -1; dy <= 1; ++dy)
            for (int dx = -1; dx <= 1; ++dx) {
              int m = (dx != 0) + (dy != 0) + (dz != 0);
              size_t q = (size_t)og_wrap(nf, 0, 2 * i + dx) +
                         (size_t)nf[0] * ((size_t)og_wrap(nf, 1, 2 * j + dy) +
                                          (size_t)nf[1] * (size_t)og_wrap(nf, 2, (dim == 3 ? 2 * k : 0) + dz));
              if (first[m]) { S[m] = rf[q]; first[m] = 0; } else S[m] += rf[q];
            }
        double acc = w[0] * S[0];
        acc = fma(w[1], S[1], acc);
        acc = fma(w[2], S[2], acc);
        if (dim == 3) acc = fma(w[3], S[3], acc);
        rhsc[(size_t)i + (size_t)nc[0] * ((size_t)j + (size_t)nc[1] * (size_t)k)] = acc;
      }
}

/*
 * Trilinear interpolation + correction.  The generator splits the loop into
 * 2^dim parity cases (C/s/e/stencil/ir/IR_FindStencilConvolutions.scala:122-170).
 * Restated: s = sum of the 2^m contributing coarse values (m = number of odd
 * fine indices) in lexicographic (z,y,x) order; uf = fma(2^-m, s, uf).
 */
void og_prolong_add(int dim, const int nf[3], const int nc[3], const double *uc,
                    double *uf) {
  OG_LOOP_BEGIN(nf)
#pragma omp parallel for schedule(static) collapse(2)
  for (int k = zl; k < zh; ++k)
    for (int j = yl; j < yh; ++j)
      for (int i = xl; i < xh; ++i) {
        int ox = i & 1, oy = j & 1, oz = (dim == 3) ? (k & 1) : 0;
        int ci = i >> 1, cj = j >> 1, ck = (dim == 3) ? (k >> 1) : 0;
        double s = 0.0;
        int first = 1;
        for (int dz = 0; dz <= oz; ++dz)
          for (int dy = 0; dy <= oy; ++dy)
            for (int dx = 0; dx <= ox; ++dx) {
              double v = uc[(size_t)(ci + dx) +
                            (size_t)nc[0] * ((size_t)(cj + dy) + (size_t)nc[1] * (size_t)(ck + dz))];
              if (first) { s = v; first = 0; } else s += v;
            }
        size_t p = (size_t)i + (size_t)nf[0] * ((size_t)j + (size_t)nf[1] * (size_t)k);
        uf[p] = fma(ldexp(1.0, -(ox + oy + oz)), s, uf[p]);
      }
}

/* ---------- streaming loops ----------------------------------------------- */

void og_set_interior(const int n[3], double *a, double v) {
  OG_LOOP_BEGIN(n)
#pragma omp parallel for schedule(static) collapse(2)
  for (int k = zl; k < zh; ++k)
    for (int j = yl; j < yh; ++j) {
      size_t row = (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
      for (int i = xl; i < xh; ++i) a[row + i] = v;
    }
}

void og_loop_eval(const int n[3], const double h[3], const double x0[3], int nsrc, const double *const *src,
                  double *dst, const int *ops, const double *vals, int nops) {
  (void)nsrc;
  OG_LOOP_BEGIN(n)
#pragma omp parallel for schedule(static) collapse(2)
  for (int k = zl; k < zh; ++k)
    for (int j = yl; j < yh; ++j) {
      size_t row = (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
      const double y = (double)j * h[1] + x0[1], z = n[2] > 1 ? (double)k * h[2] + x0[2] : 0.0;
      for (int i = xl; i < xh; ++i) {
        const double x = (double)i * h[0] + x0[0];
        double stk[16];
        int sp = 0;
        for (int q = 0; q < nops; ++q)
          switch (ops[q]) {
            case 0: stk[sp++] = vals[q]; break;
            case 1: stk[sp++] = x; break;
            case 2: stk[sp++] = y; break;
            case 3: stk[sp++] = z; break;
            case 4: --sp; stk[sp - 1] = stk[sp - 1] + stk[sp]; break;
            case 5: --sp; stk[sp - 1] = stk[sp - 1] - stk[sp]; break;
            case 6: --sp; stk[sp - 1] = stk[sp - 1] * stk[sp]; break;
            case 7: --sp; stk[sp - 1] = stk[sp - 1] / stk[sp]; break;
            case 8: stk[sp - 1] = -stk[sp - 1]; break;
            case 9: stk[sp - 1] = exp(stk[sp - 1]); break;
            case 10: stk[sp - 1] = sin(stk[sp - 1]); break;
            case 11: stk[sp - 1] = cos(stk[sp - 1]); break;
            case 12: stk[sp - 1] = sinh(stk[sp - 1]); break;
            case 13: stk[sp - 1] = sqrt(stk[sp - 1]); break;
            case 15: stk[sp++] = src[(int)vals[q]][row + i]; break;
            default: break;
          }
        dst[row + i] = stk[0];
      }
    }
}

void og_copy_interior(const int n[3], const double *src, double *dst) {
  OG_LOOP_BEGIN(n)
#pragma omp parallel for schedule(static) collapse(2)
  for (int k = zl; k < zh; ++k)
    for (int j = yl; j < yh; ++j) {
      size_t row = (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
      for (int i = xl; i < xh; ++i) dst[row + i] = src[row + i];
    }
}

void og_axpy(const int n[3], double alpha, const double *x, double *y) {
  OG_LOOP_BEGIN(n)
#pragma omp parallel for schedule(static) collapse(2)
  for (int k = zl; k < zh; ++k)
    for (int j = yl; j < yh; ++j) {
      size_t row = (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
      for (int i = xl; i < xh; ++i) y[row + i] = fma(alpha, x[row + i], y[row + i]);
    }
}

void og_xpay(const int n[3], const double *x, double beta, double *y) {
  OG_LOOP_BEGIN(n)
#pragma omp parallel for schedule(static) collapse(2)
  for (int k = zl; k < zh; ++k)
    for (int j = yl; j < yh; ++j) {
      size_t row = (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
      for (int i = xl; i < xh; ++i) y[row + i] = fma(beta, y[row + i], x[row + i]);
    }
}

/* ---------- cell-centred fields (the Testing/CellBased programs) ------------------------------------------------
 * A cell field with one ghost layer is stored as an array of (nc+2)^dim entries -- which is a node array whose
 * "boundary nodes" are the ghost cells: every loop above (og_apply, og_residual, og_smooth_jacobi, og_dot, og_axpy,
 * ...) runs over exactly the cells and reads the ghosts through the stencil.  Only the boundary update and the
 * inter-grid transfer are specific to the localisation.
 */

/* `apply bc` on a cell field.  kind 1: Dirichlet, ghost = 2 * value - inner cell (order 1,
 * C/s/e/boundary/ir/IR_DirichletBC.scala:42-50); face values g[f] per ghost cell of face f = 2*d + side in the
 * face's own lexicographic order, NULL = 0.  kind 2: Neumann(1), ghost = inner cell
 * (C/s/e/boundary/ir/IR_NeumannBC.scala:70-80).  n = padded extents (cells + 2; 1 for the unused dimension). */
void og_cell_apply_bc(int dim, const int n[3], double *u, int kind, const double *const g[6]) {
  for (int d = 0; d < dim; ++d)
    for (int side = 0; side < 2; ++side) {
      const int da = d == 0 ? 1 : 0, db = d == 2 ? 1 : 2;
      const int na = n[da] > 1 ? n[da] - 2 : 1, nb = (db < dim && n[db] > 1) ? n[db] - 2 : 1;
      const double *gv = g ? g[2 * d + side] : NULL;
      for (int b = 0; b < nb; ++b)
        for (int a = 0; a < na; ++a) {
          int idx[3] = {0, 0, 0}, inner[3];
          idx[da] = a + 1;
          if (db < dim) idx[db] = b + 1;
          idx[d] = side ? n[d] - 1 : 0;
          inner[0] = idx[0]; inner[1] = idx[1]; inner[2] = idx[2];
          inner[d] = side ? n[d] - 2 : 1;
          const size_t pg = (size_t)idx[0] + (size_t)n[0] * ((size_t)idx[1] + (size_t)n[1] * (size_t)idx[2]);
          const size_t pi = (size_t)inner[0] + (size_t)n[0] * ((size_t)inner[1] + (size_t)n[1] * (size_t)inner[2]);
          if (kind == 2) u[pg] = u[pi];
          else u[pg] = (2.0 * (gv ? gv[(size_t)a + (size_t)na * (size_t)b] : 0.0)) - u[pi];
        }
    }
}

/* RHS@coarser = RestrictionCell * Residual: mean of the 2^dim children, summed in lexicographic (z,y,x) order
 * (C/s/e/operator/l4/L4_DefaultRestriction.scala:40-47); padded extents */
void og_cell_restrict(int dim, const int nf[3], const int nc[3], const double *rf, double *rc) {
  const int ncz = dim == 3 ? nc[2] - 2 : 1;
#pragma omp parallel for schedule(static) collapse(2)
  for (int k = 0; k < ncz; ++k)
    for (int j = 0; j < nc[1] - 2; ++j)
      for (int i = 0; i < nc[0] - 2; ++i) {
        double s = 0.0;
        int first = 1;
        for (int dz = 0; dz <= (dim == 3); ++dz)
          for (int dy = 0; dy <= 1; ++dy)
            for (int dx = 0; dx <= 1; ++dx) {
              const double v = rf[(size_t)(2 * i + dx + 1) +
                                  (size_t)nf[0] * ((size_t)(2 * j + dy + 1) + (size_t)nf[1] * (size_t)(dim == 3 ? 2 * k + dz + 1 : 0))];
              if (first) { s = v; first = 0; } else s += v;
            }
        rc[(size_t)(i + 1) + (size_t)nc[0] * ((size_t)(j + 1) + (size_t)nc[1] * (size_t)(dim == 3 ? k + 1 : 0))] =
            ldexp(1.0, -dim) * s;
      }
}

/* Solution += CorrectionCell * Solution@coarser: every child takes its parent's value
 * (transpose of the restriction times 2^dim, C/s/e/operator/l4/L4_DefaultProlongation.scala:30-45) */
void og_cell_prolong_add(int dim, const int nf[3], const int nc[3], const double *uc, double *uf) {
  const int nfz = dim == 3 ? nf[2] - 2 : 1;
#pragma omp parallel for schedule(static) collapse(2)
  for (int k = 0; k < nfz; ++k)
    for (int j = 0; j < nf[1] - 2; ++j)
      for (int i = 0; i < nf[0] - 2; ++i) {
        const double v = uc[(size_t)(i / 2 + 1) + (size_t)nc[0] * ((size_t)(j / 2 + 1) + (size_t)nc[1] * (size_t)(dim == 3 ? k / 2 + 1 : 0))];
        uf[(size_t)(i + 1) + (size_t)nf[0] * ((size_t)(j + 1) + (size_t)nf[1] * (size_t)(dim == 3 ? k + 1 : 0))] += v;
      }
}

/* ---------- BiCGStab of the layer-3 generator ----------------------------------------------------------------
 * solver_cgs = "BiCGStab", C/s/e/solver/l3/L3_BiCGStabForEquation.scala:36-198 (statement order and start values as
 * generated; solver_cgs_absResThreshold = 0).  Elementwise updates in the association the library uses:
 * p = fma(beta, fma(-omega, v, p), r); s = fma(-alpha, v, r); u = fma(omega, s, fma(alpha, p, u)); r = fma(-omega, t, s).
 * Returns the number of steps. */
int og_bicgstab(const int n[3], const og_stencil *st, const double *sf, double *u, const double *rhs, double *res,
                int max_it, double rel_tol) {
  const size_t size = og_size(n);
  double *v = (double *)calloc(size, sizeof(double)), *p = (double *)calloc(size, sizeof(double));
  double *s = (double *)calloc(size, sizeof(double)), *t = (double *)calloc(size, sizeof(double));
  double *rhat = (double *)calloc(size, sizeof(double));
  OG_LOOP_BEGIN(n)
  og_residual(n, st, sf, u, rhs, res);
  double curRes = sqrt(og_dot(n, res, res));
  const double initRes = curRes;
  int iters = 0;
  if (curRes > 0.0) {
    og_copy_interior(n, res, rhat);
    double rho = 1.0, rhoOld = 1.0, alpha = 1.0, omega = 1.0;
    for (int step = 0; step < max_it; ++step) {
      iters = step + 1;
      rhoOld = rho;
      rho = og_dot(n, rhat, res);
      const double beta = (rho / rhoOld) * (alpha / omega);
      for (int k = zl; k < zh; ++k)
        for (int j = yl; j < yh; ++j)
          for (int i = xl; i < xh; ++i) {
            size_t q = (size_t)i + (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
            p[q] = fma(beta, fma(-omega, v[q], p[q]), res[q]);
          }
      og_apply(n, st, sf, p, v);
      alpha = rho / og_dot(n, rhat, v);
      for (int k = zl; k < zh; ++k)
        for (int j = yl; j < yh; ++j)
          for (int i = xl; i < xh; ++i) {
            size_t q = (size_t)i + (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
            s[q] = fma(-alpha, v[q], res[q]);
          }
      og_apply(n, st, sf, s, t);
      const double omegaNom = og_dot(n, t, s), omegaDenom = og_dot(n, t, t);
      omega = omegaNom / omegaDenom;
      for (int k = zl; k < zh; ++k)
        for (int j = yl; j < yh; ++j)
          for (int i = xl; i < xh; ++i) {
            size_t q = (size_t)i + (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
            u[q] = fma(omega, s[q], fma(alpha, p[q], u[q]));
            res[q] = fma(-omega, t[q], s[q]);
          }
      curRes = sqrt(og_dot(n, res, res));
      if (curRes <= rel_tol * initRes) break;
    }
  }
  free(v); free(p); free(s); free(t); free(rhat);
  return iters;
}

/* ---------- conjugate residuals of the layer-3 generator -----------------------------------------------------
 * solver_cgs = "ConjugateResidual", C/s/e/solver/l3/L3_ConjugateResidualForEquation.scala:36-170.  Association as in the
 * library: u = fma(alpha, p, u); r = fma(-alpha, ap, r); p = fma(beta, p, r); ap = fma(beta, ap, ar). */
int og_cr(const int n[3], const og_stencil *st, const double *sf, double *u, const double *rhs, double *res, int max_it,
          double rel_tol) {
  const size_t size = og_size(n);
  double *p = (double *)calloc(size, sizeof(double)), *ap = (double *)calloc(size, sizeof(double));
  double *ar = (double *)calloc(size, sizeof(double));
  og_residual(n, st, sf, u, rhs, res);
  double curRes = sqrt(og_dot(n, res, res));
  const double initRes = curRes;
  int iters = 0;
  if (curRes > 0.0) {
    og_copy_interior(n, res, p);
    og_apply(n, st, sf, p, ap);
    og_apply(n, st, sf, res, ar);
    for (int step = 0; step < max_it; ++step) {
      iters = step + 1;
      const double alphaNom = og_dot(n, res, ar), alphaDenom = og_dot(n, ap, ap);
      const double alpha = alphaNom / alphaDenom;
      og_axpy(n, alpha, p, u);
      og_axpy(n, -alpha, ap, res);
      const double nextRes = sqrt(og_dot(n, res, res));
      if (nextRes <= rel_tol * initRes) break;
      og_apply(n, st, sf, res, ar);
      const double beta = og_dot(n, res, ar) / alphaNom;
      og_xpay(n, res, beta, p);   /* p = r + beta p */
      og_xpay(n, ar, beta, ap);   /* ap = ar + beta ap */
      curRes = nextRes;
    }
  }
  free(p); free(ap); free(ar);
  return iters;
}

/* ---------- the smoother as coarse-grid solver -----------------------------------------------------------------
 * solver_cgs = "Smoother", C/s/e/solver/l3/L3_VankaAsSolverForEquation.scala:31-94: residual and norm; then up to max_it
 * times { 10 smoother steps (its numInnerSteps); residual; nextRes = norm; return if nextRes <= rel_tol * initRes }.
 * kind 0: damped Jacobi (two buffers, `advance` after every step; ten steps leave the active buffer where it was),
 * kind 1: red-black Gauss-Seidel (colour 0, colour 1).  Returns the number of outer steps done (0: zero residual). */
int og_smoother_solve(const int n[3], const og_stencil *st, const double *sf, double *u, const double *rhs, double *res,
                      int kind, double omega, int max_it, double rel_tol) {
  const size_t size = og_size(n);
  double *other = NULL;
  if (kind == 0) { other = (double *)malloc(size * sizeof(double)); memcpy(other, u, size * sizeof(double)); }
  og_residual(n, st, sf, u, rhs, res);
  const double initRes = sqrt(og_dot(n, res, res));
  int iters = 0;
  if (initRes != 0.0) {
    for (int step = 0; step < max_it; ++step) {
      iters = step + 1;
      for (int in = 0; in < 10; ++in) {
        if (kind == 0) {
          if (in % 2 == 0) og_smooth_jacobi(n, st, sf, omega, u, rhs, other);
          else og_smooth_jacobi(n, st, sf, omega, other, rhs, u);
        } else {
          og_smooth_rb(n, st, sf, omega, u, rhs, 0);
          og_smooth_rb(n, st, sf, omega, u, rhs, 1);
        }
      }
      og_residual(n, st, sf, u, rhs, res);
      const double nextRes = sqrt(og_dot(n, res, res));
      if (nextRes <= rel_tol * initRes) break;
    }
  }
  free(other);
  return iters;
}

/* ---------- MINRES of the layer-3 generator -------------------------------------------------------------------
 * solver_cgs = "MinRes", C/s/e/solver/l3/L3_MinResForEquation.scala:36-230: statements and start values as generated
 * (vOld = v; v = vNew and pOld = p; p = pNew done by rotating buffers).  Association as in the library. */
int og_minres(const int n[3], const og_stencil *st, const double *sf, double *u, const double *rhs, double *res,
              int max_it, double rel_tol) {
  const size_t size = og_size(n);
  double *w[7];
  for (int k = 0; k < 7; ++k) w[k] = (double *)calloc(size, sizeof(double));
  double *vOld = w[0], *v = w[1], *vNew = w[2], *pOld = w[3], *p = w[4], *pNew = w[5], *av = w[6];
  OG_LOOP_BEGIN(n)
  og_residual(n, st, sf, u, rhs, res);
  double curRes = sqrt(og_dot(n, res, res));
  const double initRes = curRes;
  int iters = 0;
  if (curRes > 0.0) {
    for (int k = zl; k < zh; ++k)
      for (int j = yl; j < yh; ++j)
        for (int i = xl; i < xh; ++i) {
          size_t q = (size_t)i + (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
          vNew[q] = res[q] / initRes;
        }
    double beta = 0.0, betaNew = 0.0, cOld = 1.0, c = 1.0, cNew = 1.0, sOld = 0.0, s = 0.0, sNew = 0.0;
    for (int step = 0; step < max_it; ++step) {
      iters = step + 1;
      beta = betaNew;
      { double *tp = vOld; vOld = v; v = vNew; vNew = tp; }
      og_apply(n, st, sf, v, av);
      for (int k = zl; k < zh; ++k)
        for (int j = yl; j < yh; ++j)
          for (int i = xl; i < xh; ++i) {
            size_t q = (size_t)i + (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
            vNew[q] = fma(-beta, vOld[q], av[q]);
          }
      const double alpha = og_dot(n, vNew, v);
      og_axpy(n, -alpha, v, vNew);
      betaNew = sqrt(og_dot(n, vNew, vNew));
      cOld = c; c = cNew; sOld = s; s = sNew;
      const double rho1 = sOld * beta;
      const double rho2 = c * cOld * beta + s * alpha;
      const double rho3Tilde = c * alpha - s * cOld * beta;
      const double tau = fabs(rho3Tilde) + fabs(betaNew);
      const double nu = tau * sqrt((rho3Tilde / tau) * (rho3Tilde / tau) + (betaNew / tau) * (betaNew / tau));
      cNew = rho3Tilde / nu;
      sNew = betaNew / nu;
      const double rho3 = nu;
      { double *tp = pOld; pOld = p; p = pNew; pNew = tp; }
      const double scale = cNew * curRes;
      for (int k = zl; k < zh; ++k)
        for (int j = yl; j < yh; ++j)
          for (int i = xl; i < xh; ++i) {
            size_t q = (size_t)i + (size_t)n[0] * ((size_t)j + (size_t)n[1] * (size_t)k);
            vNew[q] = vNew[q] / betaNew;
            const double pn = fma(-rho2, p[q], fma(-rho1, pOld[q], v[q])) / rho3;
            pNew[q] = pn;
            u[q] = fma(scale, pn, u[q]);
          }
      curRes = -1.0 * sNew * curRes;
      if (fabs(curRes) <= rel_tol * initRes) break;
    }
  }
  for (int k = 0; k < 7; ++k) free(w[k]);
  return iters;
}
