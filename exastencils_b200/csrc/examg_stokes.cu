// examg_stokes.cu -- staggered-grid Stokes: face-centred velocities, cell-centred pressure, and the mixed-localisation
// 5 x 5 "Vanka" smoother of the reference's Stokes programs.
//
// Reference program: Testing/Application/ExaStokes_2D.exa4 (lid-driven cavity; Examples/Stokes/2D_FD_Stokes_fromL4.exa4 has
// the same loops):
//   Field u < Face_x >, v < Face_y >, p < Cell >, rhs_*, residual_*                          ...exa4:3-45
//   Smoother    color with { i0 % 3, i1 % 3 } loop over p { solve locally { u@[0,0], u@[1,0], v@[0,0], v@[0,1], p@[0,0] } }
//               communicate; apply bc                                                         ...exa4:228-303
//               -> C/s/e/solver/ir/IR_LocalSolve.scala:88-236, IR_LocalDirectInvert.scala:44-150 (identity rows for
//               unknowns that are no valid computation points: IR_IsOnBoundary.scala:53-97)
//   UpdateResidual_{u,v,p}, Restriction / Correction with the default 'linear' operators of each localisation
//               (C/s/e/operator/l4/L4_DefaultRestriction.scala:29-47,79-100, L4_DefaultProlongation.scala:30-45),
//               SetSolution, apply bc on face fields (C/s/e/boundary/ir/IR_BoundaryCondition.scala:38-45,
//               IR_ApplyBCFunction.scala:55-88), V(3,3) cycle with 10 smoother steps at the bottom   ...exa4:206-226
// Loop ranges per localisation: C/s/e/baseExt/ir/IR_LoopOverPointsInOneFragment.scala:76-112 (node-like dimensions skip the
// physical boundary node, cell-like ones cover all cells).
//
// Layout: every field of a level is an (n+3) x (n+3) array of doubles, entry (i, j) at (i+1) + (n+3)(j+1): u uses
// i = -1..n+1 (faces 0..n, walls at 0 and n), j = -1..n; v the transpose; p -1..n in both -- one ghost layer each.
// Arithmetic contract = oracle/examg_oracle_stokes.c, bit for bit: fma chains in declaration order, the local system by
// Gaussian elimination without pivoting in row order and back substitution (built with --fmad=false).
// One block (one GPU); the grids of the reference test (256^2 cells on the finest level) live in L2.
#include <cmath>
#include <cstdlib>
#include <vector>

#include "examg_device.cuh"

using namespace examg;

namespace {

struct StLevel {
  int n, pitch;
  double h;
  double *f[EXAMG_ST_NFIELDS];
};

}  // namespace

struct examg_stokes {
  examg_ctx *c = nullptr;
  examg_stokes_desc desc;
  std::vector<StLevel> lev;
  dd *partials = nullptr;  // [3][ST_RED_BLOCKS]
  dd *sums = nullptr;      // [3]
  // VCycle@finest as one CUDA graph (542 kernels of a few microseconds each at 256^2 cells: launch-bound otherwise);
  // EXAMG_NO_GRAPH=1 launches them one by one
  cudaGraphExec_t graph = nullptr;
  long long graph_launches = 0;
};

namespace {

constexpr int ST_RED_BLOCKS = 256;

struct StArgs {
  int n, pitch;
  double c0, cx, ih;
  double *u, *v, *p;
  const double *fu, *fv, *fp;
  double *ru, *rv, *rp;
};

__device__ __forceinline__ double st_lap(const double *w, long long q, int sy, double c0, double cx) {
  double acc = c0 * w[q];
  acc = fma(cx, w[q - 1], acc);
  acc = fma(cx, w[q + 1], acc);
  acc = fma(cx, w[q - sy], acc);
  acc = fma(cx, w[q + sy], acc);
  return acc;
}

// one colour: cells i = col0 + 3 a, j = col1 + 3 b
__global__ void __launch_bounds__(128) k_st_smooth_color(const StArgs a, int col0, int col1, int na, int nb) {
  const int ta = blockIdx.x * blockDim.x + threadIdx.x, tb = blockIdx.y * blockDim.y + threadIdx.y;
  if (ta >= na || tb >= nb) return;
  const int i = col0 + 3 * ta, j = col1 + 3 * tb, n = a.n, sy = a.pitch;
  const long long q = (i + 1) + (long long)sy * (j + 1);
  const double c0 = a.c0, cx = a.cx, ih = a.ih;
  double *u = a.u, *v = a.v, *p = a.p;
  const bool ok[5] = {i >= 1, i + 1 <= n - 1, j >= 1, j + 1 <= n - 1, true};
  double A[5][5] = {{c0, cx, 0, 0, -ih}, {cx, c0, 0, 0, ih}, {0, 0, c0, cx, -ih}, {0, 0, cx, c0, ih}, {-ih, ih, -ih, ih, 0}};
  double f[5];
  f[0] = a.fu[q];
  f[0] = fma(-cx, u[q - 1], f[0]);
  f[0] = fma(-cx, u[q - sy], f[0]);
  f[0] = fma(-cx, u[q + sy], f[0]);
  f[0] = fma(-ih, p[q - 1], f[0]);
  f[1] = a.fu[q + 1];
  f[1] = fma(-cx, u[q + 2], f[1]);
  f[1] = fma(-cx, u[q + 1 - sy], f[1]);
  f[1] = fma(-cx, u[q + 1 + sy], f[1]);
  f[1] = fma(ih, p[q + 1], f[1]);
  f[2] = a.fv[q];
  f[2] = fma(-cx, v[q - 1], f[2]);
  f[2] = fma(-cx, v[q + 1], f[2]);
  f[2] = fma(-cx, v[q - sy], f[2]);
  f[2] = fma(-ih, p[q - sy], f[2]);
  f[3] = a.fv[q + sy];
  f[3] = fma(-cx, v[q + sy - 1], f[3]);
  f[3] = fma(-cx, v[q + sy + 1], f[3]);
  f[3] = fma(-cx, v[q + 2 * sy], f[3]);
  f[3] = fma(ih, p[q + sy], f[3]);
  f[4] = a.fp[q];
  const double cur[5] = {u[q], u[q + 1], v[q], v[q + sy], p[q]};
#pragma unroll
  for (int k = 0; k < 5; ++k)
    if (!ok[k]) {
#pragma unroll
      for (int cc = 0; cc < 5; ++cc) A[k][cc] = cc == k ? 1.0 : 0.0;
      f[k] = cur[k];
    }
#pragma unroll
  for (int k = 0; k < 5; ++k)
#pragma unroll
    for (int r = k + 1; r < 5; ++r) {
      const double m = A[r][k] / A[k][k];
#pragma unroll
      for (int cc = k + 1; cc < 5; ++cc) A[r][cc] = fma(-m, A[k][cc], A[r][cc]);
      f[r] = fma(-m, f[k], f[r]);
    }
  double x[5];
#pragma unroll
  for (int r = 4; r >= 0; --r) {
    double t = f[r];
#pragma unroll
    for (int cc = r + 1; cc < 5; ++cc) t = fma(-A[r][cc], x[cc], t);
    x[r] = t / A[r][r];
  }
  if (ok[0]) u[q] = x[0];
  if (ok[1]) u[q + 1] = x[1];
  if (ok[2]) v[q] = x[2];
  if (ok[3]) v[q + sy] = x[3];
  p[q] = x[4];
}

// apply bc in the order of the program text, two phases (the corners of the ghost layer are written by both):
//   u below the finest level (Dirichlet 0.0): phase 0 the wall faces i = 0, n (rows -1..n), phase 1 the ghost rows;
//   u on the finest level (ApplyBC_u): phase 0 the ghost rows (lid: 2 * top - inner; bottom: - inner), phase 1 the wall
//   faces (rows 0..n-1);   v (Dirichlet 0.0): phase 0 the ghost columns (2 * 0 - inner), phase 1 the wall faces j = 0, n
__global__ void __launch_bounds__(256) k_st_bc(const StArgs a, int phase, int lid, double top) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x - 1;  // -1 .. n+1
  const int n = a.n, sy = a.pitch;
  if (t > n + 1) return;
#define ST_AT(i, j) ((i) + 1 + (long long)sy * ((j) + 1))
  if ((phase == 0) != (lid != 0)) {  // wall faces of u
    const bool urow = lid ? (t >= 0 && t <= n - 1) : (t <= n);
    if (urow) {
      a.u[ST_AT(0, t)] = 0.0;
      a.u[ST_AT(n, t)] = 0.0;
    }
  } else {  // ghost rows of u
    if (lid) {
      a.u[ST_AT(t, n)] = 2.0 * top - a.u[ST_AT(t, n - 1)];
      a.u[ST_AT(t, -1)] = -a.u[ST_AT(t, 0)];
    } else {
      a.u[ST_AT(t, -1)] = (2.0 * 0.0) - a.u[ST_AT(t, 0)];
      a.u[ST_AT(t, n)] = (2.0 * 0.0) - a.u[ST_AT(t, n - 1)];
    }
  }
  if (phase == 0) {
    a.v[ST_AT(-1, t)] = (2.0 * 0.0) - a.v[ST_AT(0, t)];
    a.v[ST_AT(n, t)] = (2.0 * 0.0) - a.v[ST_AT(n - 1, t)];
  } else if (t <= n) {
    a.v[ST_AT(t, 0)] = 0.0;
    a.v[ST_AT(t, n)] = 0.0;
  }
}

__global__ void __launch_bounds__(256) k_st_residual(const StArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y * blockDim.y + threadIdx.y;
  const int n = a.n, sy = a.pitch;
  if (i >= n || j >= n) return;
  const long long q = ST_AT(i, j);
  const double c0 = a.c0, cx = a.cx, ih = a.ih;
  if (i >= 1) {
    const double dx = fma(ih, a.p[q], (-ih) * a.p[q - 1]);
    a.ru[q] = a.fu[q] - (st_lap(a.u, q, sy, c0, cx) - dx);
  }
  if (j >= 1) {
    const double dy = fma(ih, a.p[q], (-ih) * a.p[q - sy]);
    a.rv[q] = a.fv[q] - (st_lap(a.v, q, sy, c0, cx) - dy);
  }
  {
    const double dx = fma(ih, a.u[q + 1], (-ih) * a.u[q]);
    const double dy = fma(ih, a.v[q + sy], (-ih) * a.v[q]);
    a.rp[q] = a.fp[q] - (dx + dy);
  }
}

// sums of squares of the three residual fields over their loop points
__global__ void __launch_bounds__(256) k_st_norms(const StArgs a, dd *__restrict__ partials) {
  const int n = a.n, sy = a.pitch;
  dd su = {0.0, 0.0}, sv = {0.0, 0.0}, sp = {0.0, 0.0};
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < (long long)n * n; t += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(t % n), j = (int)(t / n);
    const long long q = ST_AT(i, j);
    if (i >= 1) { const double r = a.ru[q]; dd_add(su, r * r); }
    if (j >= 1) { const double r = a.rv[q]; dd_add(sv, r * r); }
    { const double r = a.rp[q]; dd_add(sp, r * r); }
  }
  su = dd_block_reduce(su);
  sv = dd_block_reduce(sv);
  sp = dd_block_reduce(sp);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = su;
    partials[ST_RED_BLOCKS + blockIdx.x] = sv;
    partials[2 * ST_RED_BLOCKS + blockIdx.x] = sp;
  }
}

__global__ void __launch_bounds__(32) k_st_norms_final(const dd *__restrict__ partials, int nblocks, dd *__restrict__ out) {
  const int k = threadIdx.x;
  if (k >= 3) return;
  dd s = {0.0, 0.0};
  for (int b = 0; b < nblocks; ++b) dd_merge(s, partials[k * ST_RED_BLOCKS + b]);
  out[k] = s;
}

// rhs@coarser = Restriction * residual, one thread per coarse cell (I, J)
__global__ void __launch_bounds__(256) k_st_restrict(const StArgs f, const StArgs c) {
  const int I = blockIdx.x * blockDim.x + threadIdx.x, J = blockIdx.y * blockDim.y + threadIdx.y;
  const int nc = c.n, fy = f.pitch;
  if (I >= nc || J >= nc) return;
  const long long q = (2 * I + 1) + (long long)fy * (2 * J + 1);
  const long long qc = (I + 1) + (long long)c.pitch * (J + 1);
  if (I >= 1) {
    double acc = 0.125 * f.ru[q - 1];
    acc = fma(0.25, f.ru[q], acc);
    acc = fma(0.125, f.ru[q + 1], acc);
    acc = fma(0.125, f.ru[q + fy - 1], acc);
    acc = fma(0.25, f.ru[q + fy], acc);
    acc = fma(0.125, f.ru[q + fy + 1], acc);
    const_cast<double *>(c.fu)[qc] = acc;
  }
  if (J >= 1) {
    double acc = 0.125 * f.rv[q - fy];
    acc = fma(0.125, f.rv[q - fy + 1], acc);
    acc = fma(0.25, f.rv[q], acc);
    acc = fma(0.25, f.rv[q + 1], acc);
    acc = fma(0.125, f.rv[q + fy], acc);
    acc = fma(0.125, f.rv[q + fy + 1], acc);
    const_cast<double *>(c.fv)[qc] = acc;
  }
  {
    double acc = 0.25 * f.rp[q];
    acc = fma(0.25, f.rp[q + 1], acc);
    acc = fma(0.25, f.rp[q + fy], acc);
    acc = fma(0.25, f.rp[q + fy + 1], acc);
    const_cast<double *>(c.fp)[qc] = acc;
  }
}

// SetSolution ( value ): the loop points of u, v, p and one layer around
__global__ void __launch_bounds__(256) k_st_set(const StArgs a, double value) {
  const int i = (int)(blockIdx.x * blockDim.x + threadIdx.x) - 1, j = (int)(blockIdx.y * blockDim.y + threadIdx.y) - 1;
  const int n = a.n, sy = a.pitch;
  if (i > n || j > n) return;
  const long long q = ST_AT(i, j);
  if (i >= 0) a.u[q] = value;
  if (j >= 0) a.v[q] = value;
  a.p[q] = value;
}

// dest += Correction * source@coarser, one thread per fine cell (i, j)
__global__ void __launch_bounds__(256) k_st_correct(const StArgs f, const StArgs c) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y * blockDim.y + threadIdx.y;
  const int n = f.n, sy = f.pitch, cy = c.pitch;
  if (i >= n || j >= n) return;
  const long long q = ST_AT(i, j);
  const long long qc = ((i >> 1) + 1) + (long long)cy * ((j >> 1) + 1);
  if (i >= 1) f.u[q] = (i & 1) ? fma(0.5, c.u[qc] + c.u[qc + 1], f.u[q]) : fma(1.0, c.u[qc], f.u[q]);
  if (j >= 1) f.v[q] = (j & 1) ? fma(0.5, c.v[qc] + c.v[qc + cy], f.v[q]) : fma(1.0, c.v[qc], f.v[q]);
  f.p[q] = fma(1.0, c.p[qc], f.p[q]);
}
#undef ST_AT

StArgs st_args(const StLevel &L) {
  StArgs a;
  a.n = L.n;
  a.pitch = L.pitch;
  const double h = L.h;
  a.c0 = 2.0 / (h * h) + 2.0 / (h * h);
  a.cx = -1.0 / (h * h);
  a.ih = 1.0 / h;
  a.u = L.f[EXAMG_ST_U]; a.v = L.f[EXAMG_ST_V]; a.p = L.f[EXAMG_ST_P];
  a.fu = L.f[EXAMG_ST_RHS_U]; a.fv = L.f[EXAMG_ST_RHS_V]; a.fp = L.f[EXAMG_ST_RHS_P];
  a.ru = L.f[EXAMG_ST_RES_U]; a.rv = L.f[EXAMG_ST_RES_V]; a.rp = L.f[EXAMG_ST_RES_P];
  return a;
}

StLevel *st_level(examg_stokes *s, int level) {
  if (!s) { set_error(EXAMG_EINVAL, "null Stokes solver"); return nullptr; }
  const int lo = s->c->cfg.min_level, hi = s->c->cfg.max_level;
  if (level < 0) level = hi;
  if (level < lo || level > hi) { set_error(EXAMG_EINVAL, "Stokes solver: level %d outside %d..%d", level, lo, hi); return nullptr; }
  return &s->lev[level - lo];
}

dim3 grid2(int nx, int ny, dim3 block) { return dim3((nx + block.x - 1) / block.x, (ny + block.y - 1) / block.y, 1); }

}  // namespace

extern "C" {

examg_stokes *examg_stokes_create(examg_ctx *c, const examg_stokes_desc *d) {
  if (!c || !d) { set_error(EXAMG_EINVAL, "examg_stokes_create: null argument"); return nullptr; }
  if (c->cfg.dim != 2 || c->world != 1 || c->periodic_any) {
    set_error(EXAMG_EUNSUP, "examg_stokes_create: 2D, one block, no periodic dimension");
    return nullptr;
  }
  if (c->cfg.block_cells[0] != c->cfg.block_cells[1] || c->cfg.min_level + (c->cfg.block_cells[0] > 1 ? 1 : 0) < 1) {
    set_error(EXAMG_EUNSUP, "examg_stokes_create: square grids with at least 2 cells per dimension on the coarsest level");
    return nullptr;
  }
  if (d->nu_pre < 0 || d->nu_post < 0 || d->nu_coarse < 0) { set_error(EXAMG_EINVAL, "examg_stokes_create: negative sweep count"); return nullptr; }
  examg_stokes *s = new examg_stokes;
  s->c = c;
  s->desc = *d;
  bool ok = true;
  for (int level = c->cfg.min_level; level <= c->cfg.max_level && ok; ++level) {
    StLevel L;
    L.n = c->cfg.block_cells[0] << level;
    L.pitch = L.n + 3;
    L.h = (c->cfg.domain_hi[0] - c->cfg.domain_lo[0]) / (double)L.n;
    const size_t bytes = sizeof(double) * (size_t)L.pitch * (size_t)L.pitch;
    for (int k = 0; k < EXAMG_ST_NFIELDS; ++k) {
      L.f[k] = nullptr;
      ok = ok && cudaMalloc(&L.f[k], bytes) == cudaSuccess && cudaMemsetAsync(L.f[k], 0, bytes, c->stream) == cudaSuccess;
    }
    s->lev.push_back(L);
  }
  ok = ok && cudaMalloc(&s->partials, sizeof(dd) * 3 * ST_RED_BLOCKS) == cudaSuccess && cudaMalloc(&s->sums, sizeof(dd) * 3) == cudaSuccess;
  if (!ok) {
    set_error(EXAMG_ECUDA, "examg_stokes_create: device allocation failed");
    examg_stokes_destroy(s);
    return nullptr;
  }
  return s;
}

int examg_stokes_destroy(examg_stokes *s) {
  if (!s) return EXAMG_OK;
  for (StLevel &L : s->lev)
    for (int k = 0; k < EXAMG_ST_NFIELDS; ++k)
      if (L.f[k]) cudaFree(L.f[k]);
  if (s->partials) cudaFree(s->partials);
  if (s->sums) cudaFree(s->sums);
  if (s->graph) cudaGraphExecDestroy(s->graph);
  delete s;
  return EXAMG_OK;
}

int examg_stokes_cells(examg_stokes *s, int level) {
  StLevel *L = st_level(s, level);
  return L ? L->n : -1;
}

int examg_stokes_upload(examg_stokes *s, int level, int which, const double *host) {
  StLevel *L = st_level(s, level);
  if (!L) return EXAMG_EINVAL;
  if (which < 0 || which >= EXAMG_ST_NFIELDS || !host) return set_error(EXAMG_EINVAL, "examg_stokes_upload: field %d", which);
  EXAMG_CUDA(cudaMemcpyAsync(L->f[which], host, sizeof(double) * (size_t)L->pitch * L->pitch, cudaMemcpyHostToDevice, s->c->stream));
  EXAMG_CUDA(cudaStreamSynchronize(s->c->stream));
  return EXAMG_OK;
}

int examg_stokes_download(examg_stokes *s, int level, int which, double *host) {
  StLevel *L = st_level(s, level);
  if (!L) return EXAMG_EINVAL;
  if (which < 0 || which >= EXAMG_ST_NFIELDS || !host) return set_error(EXAMG_EINVAL, "examg_stokes_download: field %d", which);
  EXAMG_CUDA(cudaMemcpyAsync(host, L->f[which], sizeof(double) * (size_t)L->pitch * L->pitch, cudaMemcpyDeviceToHost, s->c->stream));
  EXAMG_CUDA(cudaStreamSynchronize(s->c->stream));
  return EXAMG_OK;
}

int examg_stokes_apply_bc(examg_stokes *s, int level) {
  StLevel *L = st_level(s, level);
  if (!L) return EXAMG_EINVAL;
  examg_ctx *c = s->c;
  const StArgs a = st_args(*L);
  const int lid = (level < 0 || level == c->cfg.max_level) ? 1 : 0;
  const int blocks = (L->n + 3 + 255) / 256;
  for (int phase = 0; phase < 2; ++phase) {
    k_st_bc<<<blocks, 256, 0, c->stream>>>(a, phase, lid, s->desc.top_vel);
    EXAMG_LAUNCH_CHECK(c);
  }
  return EXAMG_OK;
}

int examg_stokes_smooth_color(examg_stokes *s, int level, int c0, int c1) {
  StLevel *L = st_level(s, level);
  if (!L) return EXAMG_EINVAL;
  if (c0 < 0 || c0 > 2 || c1 < 0 || c1 > 2) return set_error(EXAMG_EINVAL, "examg_stokes_smooth_color: colour (%d, %d)", c0, c1);
  examg_ctx *c = s->c;
  const int na = L->n > c0 ? (L->n - c0 + 2) / 3 : 0, nb = L->n > c1 ? (L->n - c1 + 2) / 3 : 0;
  if (na <= 0 || nb <= 0) return EXAMG_OK;
  const dim3 block(32, 4, 1);
  k_st_smooth_color<<<grid2(na, nb, block), block, 0, c->stream>>>(st_args(*L), c0, c1, na, nb);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

int examg_stokes_smooth(examg_stokes *s, int level, int nsweeps) {
  for (int it = 0; it < nsweeps; ++it) {
    for (int c1 = 0; c1 < 3; ++c1)
      for (int c0 = 0; c0 < 3; ++c0) {
        const int rc = examg_stokes_smooth_color(s, level, c0, c1);
        if (rc) return rc;
      }
    const int rc = examg_stokes_apply_bc(s, level);
    if (rc) return rc;
  }
  return EXAMG_OK;
}

int examg_stokes_residual(examg_stokes *s, int level, double *norms_out) {
  StLevel *L = st_level(s, level);
  if (!L) return EXAMG_EINVAL;
  examg_ctx *c = s->c;
  const StArgs a = st_args(*L);
  const dim3 block(32, 8, 1);
  k_st_residual<<<grid2(L->n, L->n, block), block, 0, c->stream>>>(a);
  EXAMG_LAUNCH_CHECK(c);
  if (!norms_out) return EXAMG_OK;
  long long nb = ((long long)L->n * L->n + 255) / 256;
  if (nb > ST_RED_BLOCKS) nb = ST_RED_BLOCKS;
  k_st_norms<<<(int)nb, 256, 0, c->stream>>>(a, s->partials);
  EXAMG_LAUNCH_CHECK(c);
  k_st_norms_final<<<1, 32, 0, c->stream>>>(s->partials, (int)nb, s->sums);
  EXAMG_LAUNCH_CHECK(c);
  dd host[3];
  EXAMG_CUDA(cudaMemcpyAsync(host, s->sums, sizeof(host), cudaMemcpyDeviceToHost, c->stream));
  EXAMG_CUDA(cudaStreamSynchronize(c->stream));
  for (int k = 0; k < 3; ++k) norms_out[k] = sqrt(host[k].hi + host[k].lo);
  return EXAMG_OK;
}

int examg_stokes_restrict(examg_stokes *s, int level) {
  StLevel *F = st_level(s, level);
  if (!F) return EXAMG_EINVAL;
  if (F == &s->lev[0]) return set_error(EXAMG_EINVAL, "examg_stokes_restrict: no level below the coarsest");
  StLevel *C = F - 1;
  const dim3 block(32, 8, 1);
  k_st_restrict<<<grid2(C->n, C->n, block), block, 0, s->c->stream>>>(st_args(*F), st_args(*C));
  EXAMG_LAUNCH_CHECK(s->c);
  return EXAMG_OK;
}

int examg_stokes_set_solution(examg_stokes *s, int level, double value) {
  StLevel *L = st_level(s, level);
  if (!L) return EXAMG_EINVAL;
  const dim3 block(32, 8, 1);
  k_st_set<<<grid2(L->n + 2, L->n + 2, block), block, 0, s->c->stream>>>(st_args(*L), value);
  EXAMG_LAUNCH_CHECK(s->c);
  return EXAMG_OK;
}

int examg_stokes_correct(examg_stokes *s, int level) {
  StLevel *F = st_level(s, level);
  if (!F) return EXAMG_EINVAL;
  if (F == &s->lev[0]) return set_error(EXAMG_EINVAL, "examg_stokes_correct: no level below the coarsest");
  StLevel *C = F - 1;
  const dim3 block(32, 8, 1);
  k_st_correct<<<grid2(F->n, F->n, block), block, 0, s->c->stream>>>(st_args(*F), st_args(*C));
  EXAMG_LAUNCH_CHECK(s->c);
  return examg_stokes_apply_bc(s, level);
}

static int st_vcycle_rec(examg_stokes *s, int level) {
  int rc;
  if (level == s->c->cfg.min_level) return examg_stokes_smooth(s, level, s->desc.nu_coarse);
  if ((rc = examg_stokes_smooth(s, level, s->desc.nu_pre))) return rc;
  if ((rc = examg_stokes_residual(s, level, nullptr))) return rc;
  if ((rc = examg_stokes_restrict(s, level))) return rc;
  if ((rc = examg_stokes_set_solution(s, level - 1, 0.0))) return rc;
  if ((rc = st_vcycle_rec(s, level - 1))) return rc;
  if ((rc = examg_stokes_correct(s, level))) return rc;
  return examg_stokes_smooth(s, level, s->desc.nu_post);
}

int examg_stokes_vcycle(examg_stokes *s, int level) {
  if (!s) return set_error(EXAMG_EINVAL, "null Stokes solver");
  examg_ctx *c = s->c;
  if (level < 0) level = c->cfg.max_level;
  if (level < c->cfg.min_level || level > c->cfg.max_level) return set_error(EXAMG_EINVAL, "examg_stokes_vcycle: level %d", level);
  static const bool use_graph = !(getenv("EXAMG_NO_GRAPH") && getenv("EXAMG_NO_GRAPH")[0] == '1');
  if (!use_graph || level != c->cfg.max_level || level == c->cfg.min_level) return st_vcycle_rec(s, level);
  if (!s->graph) {  // the buffers never move: one capture serves every later cycle
    const long long l0 = c->launches;
    cudaGraph_t graph = nullptr;
    EXAMG_CUDA(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    const int rc = st_vcycle_rec(s, level);
    const cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
    s->graph_launches = c->launches - l0;
    c->launches = l0;  // nothing ran yet
    if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
    if (e != cudaSuccess) return set_error(EXAMG_ECUDA, "graph capture of VCycle failed: %s", cudaGetErrorString(e));
    const cudaError_t e2 = cudaGraphInstantiate(&s->graph, graph, 0);
    cudaGraphDestroy(graph);
    if (e2 != cudaSuccess) { s->graph = nullptr; return set_error(EXAMG_ECUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(e2)); }
  }
  EXAMG_CUDA(cudaGraphLaunch(s->graph, c->stream));
  c->launches += s->graph_launches;
  return EXAMG_OK;
}

}  // extern "C"
