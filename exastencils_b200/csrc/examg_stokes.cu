// examg_stokes.cu -- staggered-grid Stokes in 2D and 3D: face-centred velocities, cell-centred pressure, and the
// mixed-localisation (2 dim + 1) x (2 dim + 1) "Vanka" smoother of the reference's Stokes programs.
//
// Reference programs: Testing/Application/ExaStokes_2D.exa4 and ExaStokes_3D.exa4 (lid-driven cavity;
// Examples/Stokes/{2D,3D}_FD_Stokes_fromL4.exa4 have the same loops):
//   Field u < Face_x >, v < Face_y > [, w < Face_z >], p < Cell >, rhs_*, residual_*            2D:3-45, 3D:3-52
//   Smoother    color with { i0 % 3, i1 % 3 [, i2 % 3] } loop over p { solve locally { the two faces of the cell in
//               every dimension, p@[0,0] } }; communicate; apply bc                               2D:228-303, 3D:272-381
//               -> C/s/e/solver/ir/IR_LocalSolve.scala:88-236, IR_LocalDirectInvert.scala:44-150 (identity rows for
//               unknowns that are no valid computation points: IR_IsOnBoundary.scala:53-97)
//   UpdateResidual_*, Restriction / Correction with the default 'linear' operators of each localisation
//               (C/s/e/operator/l4/L4_DefaultRestriction.scala:29-47,79-100, L4_DefaultProlongation.scala:30-45),
//               SetSolution, apply bc on face fields (C/s/e/boundary/ir/IR_BoundaryCondition.scala:38-45,
//               IR_ApplyBCFunction.scala:55-88), V(3,3) cycle with 10 smoother steps at the bottom
// Loop ranges per localisation: C/s/e/baseExt/ir/IR_LoopOverPointsInOneFragment.scala:76-112 (node-like dimensions skip the
// physical boundary node, cell-like ones cover all cells).
//
// Layout: every field of a level is an (n+3)^dim array of doubles, entry (i, j, k) at (i+1) + (n+3)(j+1) + (n+3)^2 (k+1):
// velocity component d uses the faces 0..n in dimension d (walls at 0 and n) and the cells 0..n-1 in the others, p the
// cells; one ghost layer around each.
// Arithmetic contract = oracle/examg_oracle_stokes.c, bit for bit: fma chains in declaration order, the local system by
// Gaussian elimination without pivoting in row order and back substitution (built with --fmad=false).
// One block (one GPU).
#include <cmath>
#include <cstdlib>
#include <vector>

#include "examg_device.cuh"

using namespace examg;

namespace {

struct StLevel {
  int n;
  long long s[3];  // strides (s[2] = 0 in 2D)
  size_t size;
  double h;
  double *f[EXAMG_ST_NFIELDS];
};

}  // namespace

struct examg_stokes {
  examg_ctx *c = nullptr;
  examg_stokes_desc desc;
  int dim = 2;
  std::vector<StLevel> lev;
  dd *partials = nullptr;  // [4][ST_RED_BLOCKS]
  dd *sums = nullptr;      // [4]
  // VCycle@finest as one CUDA graph (542 kernels of a few microseconds each at 256^2 cells: launch-bound otherwise);
  // EXAMG_NO_GRAPH=1 launches them one by one
  cudaGraphExec_t graph = nullptr;
  long long graph_launches = 0;
};

namespace {

constexpr int ST_RED_BLOCKS = 256;

struct StArgs {
  int n;
  long long s[3];
  double c0, cx, ih;
  double *vel[3], *p;
  double *fvel[3], *fp;  // right-hand sides
  double *rvel[3], *rp;  // residuals
};

template <int D>
__device__ __forceinline__ long long st_at(const StArgs &a, int i, int j, int k) {
  return (i + 1) + a.s[1] * (j + 1) + (D == 3 ? a.s[2] * (k + 1) : 0);
}

// Laplace * w at q: acc = c0*w; fma in declaration order -x, +x, -y, +y[, -z, +z]
template <int D>
__device__ __forceinline__ double st_lap(const StArgs &a, const double *w, long long q) {
  double acc = a.c0 * w[q];
#pragma unroll
  for (int e = 0; e < D; ++e) {
    acc = fma(a.cx, w[q - a.s[e]], acc);
    acc = fma(a.cx, w[q + a.s[e]], acc);
  }
  return acc;
}

// one colour: cells col[d] + 3 t_d.  Unknowns 2d, 2d+1 = the faces of the cell in dimension d, last = p.
// The oracle eliminates the dense (2D+1)^2 matrix; here only the entries that can be non-zero are kept -- per dimension
// the 2 x 2 block of the two faces, their column of p and their entries of the last row (aa ab ap / ba bb bp / la lb),
// plus the diagonal of p.  Eliminating with a multiplier 0 leaves a row bit for bit as it was (fma(-0, x, y) = y), and no
// fill leaves this pattern, so the result equals the dense elimination in row order exactly; the identity rows of wall
// faces only add zeros.  (The dense form put the 7 x 7 matrix of the 3D program into local memory: 464 B of stack.)
template <int D>
__global__ void __launch_bounds__(128)
k_st_smooth_color(const StArgs a, int col0, int col1, int col2, int na, int nb, int nc) {
  const int ta = blockIdx.x * blockDim.x + threadIdx.x, tb = blockIdx.y * blockDim.y + threadIdx.y, tc = blockIdx.z;
  if (ta >= na || tb >= nb || tc >= nc) return;
  const int idx[3] = {col0 + 3 * ta, col1 + 3 * tb, D == 3 ? col2 + 3 * tc : 0};
  const int n = a.n;
  const long long q = st_at<D>(a, idx[0], idx[1], idx[2]);
  const double c0 = a.c0, cx = a.cx, ih = a.ih;
  double aa[D], ab[D], ap[D], ba[D], bb[D], bp[D], la[D], lb[D], fa[D], fb[D];
  bool oka[D], okb[D];
#pragma unroll
  for (int d = 0; d < D; ++d) {
    const double *w = a.vel[d];
    const double *fw = a.fvel[d];
    const long long sd = a.s[d];
    oka[d] = idx[d] >= 1;
    okb[d] = idx[d] + 1 <= n - 1;
    double ta_ = fw[q], tb_ = fw[q + sd];
#pragma unroll
    for (int e = 0; e < D; ++e) {
      const long long se = a.s[e];
      ta_ = fma(-cx, w[q - se], ta_);
      if (e != d) ta_ = fma(-cx, w[q + se], ta_);
      if (e != d) tb_ = fma(-cx, w[q + sd - se], tb_);
      tb_ = fma(-cx, w[q + sd + se], tb_);
    }
    ta_ = fma(-ih, a.p[q - sd], ta_);
    tb_ = fma(ih, a.p[q + sd], tb_);
    // rows of wall faces: identity, right-hand side = the current value
    aa[d] = oka[d] ? c0 : 1.0; ab[d] = oka[d] ? cx : 0.0; ap[d] = oka[d] ? -ih : 0.0; fa[d] = oka[d] ? ta_ : w[q];
    ba[d] = okb[d] ? cx : 0.0; bb[d] = okb[d] ? c0 : 1.0; bp[d] = okb[d] ? ih : 0.0; fb[d] = okb[d] ? tb_ : w[q + sd];
    la[d] = -ih;
    lb[d] = ih;
  }
  double pp = 0.0, fp = a.fp[q];
#pragma unroll
  for (int d = 0; d < D; ++d) {
    {  // pivot row 2d: rows 2d+1 and last
      const double m = ba[d] / aa[d];
      bb[d] = fma(-m, ab[d], bb[d]);
      bp[d] = fma(-m, ap[d], bp[d]);
      fb[d] = fma(-m, fa[d], fb[d]);
      const double ml = la[d] / aa[d];
      lb[d] = fma(-ml, ab[d], lb[d]);
      pp = fma(-ml, ap[d], pp);
      fp = fma(-ml, fa[d], fp);
    }
    {  // pivot row 2d+1: last row
      const double ml = lb[d] / bb[d];
      pp = fma(-ml, bp[d], pp);
      fp = fma(-ml, fb[d], fp);
    }
  }
  const double xp = fp / pp;
#pragma unroll
  for (int d = D - 1; d >= 0; --d) {
    const double xb = fma(-bp[d], xp, fb[d]) / bb[d];
    double t = fma(-ab[d], xb, fa[d]);
    t = fma(-ap[d], xp, t);
    const double xa = t / aa[d];
    if (oka[d]) a.vel[d][q] = xa;
    if (okb[d]) a.vel[d][q + a.s[d]] = xb;
  }
  a.p[q] = xp;
}

// One phase of `apply bc`: for every velocity component one pair of opposite faces (dimension e[d]; -1: nothing).
// kind 0: wall faces idx[e] = 0, n take 0; kind 1: ghost = g2 - inner; kind 2: ghost = -inner (per side).
// blockIdx.y = component.  lo / hi: the tangential box (the entry of dimension e is ignored).
struct StBcPhase {
  int e[3], kind_lo[3], kind_hi[3];
  int lo[3][3], hi[3][3];
  double g2_lo[3], g2_hi[3];
};

template <int D>
__global__ void __launch_bounds__(256) k_st_bc(const StArgs a, const __grid_constant__ StBcPhase ph) {
  const int d = blockIdx.y;
  const int e = ph.e[d];
  if (e < 0) return;
  int ext[3], total = 1;
#pragma unroll
  for (int t = 0; t < 3; ++t) {
    ext[t] = (t < D && t != e) ? ph.hi[d][t] - ph.lo[d][t] + 1 : 1;
    total *= ext[t];
  }
  const int n = a.n;
  double *w = a.vel[d];
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < 2 * total; t += gridDim.x * blockDim.x) {
    const int side = t >= total ? 1 : 0;
    int r = t - side * total, idx[3];
    idx[0] = ph.lo[d][0] + r % ext[0]; r /= ext[0];
    idx[1] = ph.lo[d][1] + r % ext[1]; r /= ext[1];
    idx[2] = D == 3 ? ph.lo[d][2] + r : 0;
    const int kind = side ? ph.kind_hi[d] : ph.kind_lo[d];
    idx[e] = kind == 0 ? (side ? n : 0) : (side ? n : -1);
    const long long q = st_at<D>(a, idx[0], idx[1], idx[2]);
    if (kind == 0) {
      w[q] = 0.0;
    } else {
      const double inner = side ? w[q - a.s[e]] : w[q + a.s[e]];
      w[q] = kind == 2 ? -inner : (side ? ph.g2_hi[d] : ph.g2_lo[d]) - inner;
    }
  }
}

// cell (i, j, k) of thread t in an n^D box
template <int D>
__device__ __forceinline__ bool st_cell(int n, int idx[3]) {
  idx[0] = blockIdx.x * blockDim.x + threadIdx.x;
  idx[1] = blockIdx.y * blockDim.y + threadIdx.y;
  idx[2] = D == 3 ? (int)blockIdx.z : 0;
  return idx[0] < n && idx[1] < n && idx[2] < n;
}

// UpdateResidual_*: residual_d = rhs_d - ( Laplace * vel_d - d?Left * p ) on the inner faces;
// residual_p = rhs_p - ( dxRight * u + dyRight * v [+ dzRight * w] )
template <int D>
__global__ void __launch_bounds__(256) k_st_residual(const StArgs a) {
  int idx[3];
  if (!st_cell<D>(a.n, idx)) return;
  const long long q = st_at<D>(a, idx[0], idx[1], idx[2]);
  const double ih = a.ih;
  double div = 0.0;
#pragma unroll
  for (int d = 0; d < D; ++d) {
    const double *w = a.vel[d];
    const long long sd = a.s[d];
    if (idx[d] >= 1) {
      const double dl = fma(ih, a.p[q], (-ih) * a.p[q - sd]);
      a.rvel[d][q] = a.fvel[d][q] - (st_lap<D>(a, w, q) - dl);
    }
    const double dr = fma(ih, w[q + sd], (-ih) * w[q]);
    div = d == 0 ? dr : div + dr;
  }
  a.rp[q] = a.fp[q] - div;
}

// sums of squares of the residual fields over their loop points: slots 0..D-1 velocities, D pressure
template <int D>
__global__ void __launch_bounds__(256) k_st_norms(const StArgs a, dd *__restrict__ partials) {
  const int n = a.n;
  const long long cells = D == 3 ? (long long)n * n * n : (long long)n * n;
  dd acc[D + 1];
#pragma unroll
  for (int k = 0; k <= D; ++k) acc[k] = dd{0.0, 0.0};
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < cells; t += (long long)gridDim.x * blockDim.x) {
    const int idx[3] = {(int)(t % n), (int)((t / n) % n), (int)(t / ((long long)n * n))};
    const long long q = st_at<D>(a, idx[0], idx[1], idx[2]);
#pragma unroll
    for (int d = 0; d < D; ++d)
      if (idx[d] >= 1) { const double r = a.rvel[d][q]; dd_add(acc[d], r * r); }
    const double r = a.rp[q];
    dd_add(acc[D], r * r);
  }
#pragma unroll
  for (int k = 0; k <= D; ++k) {
    const dd v = dd_block_reduce(acc[k]);
    if (threadIdx.x == 0) partials[k * ST_RED_BLOCKS + blockIdx.x] = v;
  }
}

__global__ void __launch_bounds__(32) k_st_norms_final(const dd *__restrict__ partials, int nblocks, int nsums, dd *__restrict__ out) {
  const int k = threadIdx.x;
  if (k >= nsums) return;
  dd s = {0.0, 0.0};
  for (int b = 0; b < nblocks; ++b) dd_merge(s, partials[k * ST_RED_BLOCKS + b]);
  out[k] = s;
}

// Restriction_<loc> * src at the coarse index I: kron of [1/4 1/2 1/4] (fine 2I-1..2I+1) in dimension OWN (OWN < 0:
// none) and [1/2 1/2] (fine 2I, 2I+1) in the others; entries in lexicographic order (x fastest), fma chain
template <int D, int OWN>
__device__ __forceinline__ double st_restrict_point(const StArgs &f, const double *src, const int I[3]) {
  constexpr int c0 = OWN == 0 ? 3 : 2, c1 = OWN == 1 ? 3 : 2, c2 = D == 3 ? (OWN == 2 ? 3 : 2) : 1;
  double acc = 0.0;
#pragma unroll
  for (int c = 0; c < c2; ++c)
#pragma unroll
    for (int b = 0; b < c1; ++b)
#pragma unroll
      for (int a = 0; a < c0; ++a) {
        const int ox = (OWN == 0 ? -1 : 0) + a, oy = (OWN == 1 ? -1 : 0) + b, oz = (OWN == 2 ? -1 : 0) + c;
        double wgt = OWN == 0 ? (ox == 0 ? 0.5 : 0.25) : 0.5;
        wgt *= OWN == 1 ? (oy == 0 ? 0.5 : 0.25) : 0.5;
        if (D == 3) wgt *= OWN == 2 ? (oz == 0 ? 0.5 : 0.25) : 0.5;
        const double v = src[st_at<D>(f, 2 * I[0] + ox, 2 * I[1] + oy, D == 3 ? 2 * I[2] + oz : 0)];
        acc = (a == 0 && b == 0 && c == 0) ? wgt * v : fma(wgt, v, acc);
      }
  return acc;
}

// rhs@coarser = Restriction * residual, one thread per coarse cell
template <int D>
__global__ void __launch_bounds__(256) k_st_restrict(const StArgs f, const StArgs c) {
  int I[3];
  if (!st_cell<D>(c.n, I)) return;
  const long long qc = st_at<D>(c, I[0], I[1], I[2]);
  if (I[0] >= 1) c.fvel[0][qc] = st_restrict_point<D, 0>(f, f.rvel[0], I);
  if (I[1] >= 1) c.fvel[1][qc] = st_restrict_point<D, 1>(f, f.rvel[1], I);
  if (D == 3 && I[2] >= 1) c.fvel[2][qc] = st_restrict_point<D, 2>(f, f.rvel[2], I);
  c.fp[qc] = st_restrict_point<D, -1>(f, f.rp, I);
}

// SetSolution ( value ): the loop points of the unknowns and one layer around
template <int D>
__global__ void __launch_bounds__(256) k_st_set(const StArgs a, double value) {
  int idx[3];
  if (!st_cell<D>(a.n + 2, idx)) return;
#pragma unroll
  for (int t = 0; t < D; ++t) idx[t] -= 1;
  const long long q = st_at<D>(a, idx[0], idx[1], idx[2]);
#pragma unroll
  for (int d = 0; d < D; ++d)
    if (idx[d] >= 0) a.vel[d][q] = value;
  a.p[q] = value;
}

// dest += Correction * source@coarser, one thread per fine cell: linear in the component's own dimension, constant in
// the others; p takes the parent's value
template <int D>
__global__ void __launch_bounds__(256) k_st_correct(const StArgs f, const StArgs c) {
  int idx[3];
  if (!st_cell<D>(f.n, idx)) return;
  const long long q = st_at<D>(f, idx[0], idx[1], idx[2]);
  const long long qc = st_at<D>(c, idx[0] >> 1, idx[1] >> 1, idx[2] >> 1);
#pragma unroll
  for (int d = 0; d < D; ++d)
    if (idx[d] >= 1) {
      const double *wc = c.vel[d];
      f.vel[d][q] = (idx[d] & 1) ? fma(0.5, wc[qc] + wc[qc + c.s[d]], f.vel[d][q]) : fma(1.0, wc[qc], f.vel[d][q]);
    }
  f.p[q] = fma(1.0, c.p[qc], f.p[q]);
}

StArgs st_args(const examg_stokes *s, const StLevel &L) {
  StArgs a;
  a.n = L.n;
  const double h = L.h;
  a.c0 = 2.0 / (h * h) + 2.0 / (h * h);
  if (s->dim == 3) a.c0 = a.c0 + 2.0 / (h * h);
  a.cx = -1.0 / (h * h);
  a.ih = 1.0 / h;
  for (int d = 0; d < 3; ++d) {
    a.s[d] = L.s[d];
    a.vel[d] = L.f[EXAMG_ST_VEL + d];
    a.fvel[d] = L.f[EXAMG_ST_RHS + d];
    a.rvel[d] = L.f[EXAMG_ST_RES + d];
  }
  a.p = L.f[EXAMG_ST_P];
  a.fp = L.f[EXAMG_ST_RHS_P];
  a.rp = L.f[EXAMG_ST_RES_P];
  return a;
}

StLevel *st_level(examg_stokes *s, int level) {
  if (!s) { set_error(EXAMG_EINVAL, "null Stokes solver"); return nullptr; }
  const int lo = s->c->cfg.min_level, hi = s->c->cfg.max_level;
  if (level < 0) level = hi;
  if (level < lo || level > hi) { set_error(EXAMG_EINVAL, "Stokes solver: level %d outside %d..%d", level, lo, hi); return nullptr; }
  return &s->lev[level - lo];
}

bool st_field_ok(const examg_stokes *s, int which) {
  if (which < 0 || which >= EXAMG_ST_NFIELDS) return false;
  return which >= EXAMG_ST_P || which % 3 < s->dim;
}

// one thread per cell of an m^dim box: blocks of 32 x 8, z in the grid
dim3 st_grid(int dim, int m) { return dim3((m + 31) / 32, (m + 7) / 8, dim == 3 ? m : 1); }
const dim3 ST_BLOCK(32, 8, 1);

#define ST_DISPATCH(s, KERNEL, GRID, BLOCK, ...)                                            \
  do {                                                                                      \
    if ((s)->dim == 3) KERNEL<3><<<GRID, BLOCK, 0, (s)->c->stream>>>(__VA_ARGS__);           \
    else KERNEL<2><<<GRID, BLOCK, 0, (s)->c->stream>>>(__VA_ARGS__);                         \
    EXAMG_LAUNCH_CHECK((s)->c);                                                             \
  } while (0)

}  // namespace

extern "C" {

examg_stokes *examg_stokes_create(examg_ctx *c, const examg_stokes_desc *d) {
  if (!c || !d) { set_error(EXAMG_EINVAL, "examg_stokes_create: null argument"); return nullptr; }
  if (c->world != 1 || c->periodic_any) {
    set_error(EXAMG_EUNSUP, "examg_stokes_create: one block, no periodic dimension");
    return nullptr;
  }
  const int dim = c->cfg.dim;
  for (int t = 1; t < dim; ++t)
    if (c->cfg.block_cells[t] != c->cfg.block_cells[0]) {
      set_error(EXAMG_EUNSUP, "examg_stokes_create: the same number of cells in every dimension");
      return nullptr;
    }
  if ((c->cfg.block_cells[0] << c->cfg.min_level) < 2) {
    set_error(EXAMG_EUNSUP, "examg_stokes_create: at least 2 cells per dimension on the coarsest level");
    return nullptr;
  }
  if (d->nu_pre < 0 || d->nu_post < 0 || d->nu_coarse < 0) { set_error(EXAMG_EINVAL, "examg_stokes_create: negative sweep count"); return nullptr; }
  examg_stokes *s = new examg_stokes;
  s->c = c;
  s->desc = *d;
  s->dim = dim;
  bool ok = true;
  for (int level = c->cfg.min_level; level <= c->cfg.max_level && ok; ++level) {
    StLevel L;
    L.n = c->cfg.block_cells[0] << level;
    const long long m = L.n + 3;
    L.s[0] = 1;
    L.s[1] = m;
    L.s[2] = dim == 3 ? m * m : 0;
    L.size = (size_t)(dim == 3 ? m * m * m : m * m);
    L.h = (c->cfg.domain_hi[0] - c->cfg.domain_lo[0]) / (double)L.n;
    for (int k = 0; k < EXAMG_ST_NFIELDS; ++k) {
      L.f[k] = nullptr;
      if (!st_field_ok(s, k)) continue;
      ok = ok && cudaMalloc(&L.f[k], sizeof(double) * L.size) == cudaSuccess &&
           cudaMemsetAsync(L.f[k], 0, sizeof(double) * L.size, c->stream) == cudaSuccess;
    }
    s->lev.push_back(L);
  }
  ok = ok && cudaMalloc(&s->partials, sizeof(dd) * 4 * ST_RED_BLOCKS) == cudaSuccess && cudaMalloc(&s->sums, sizeof(dd) * 4) == cudaSuccess;
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
  if (!st_field_ok(s, which) || !host) return set_error(EXAMG_EINVAL, "examg_stokes_upload: field %d", which);
  EXAMG_CUDA(cudaMemcpyAsync(L->f[which], host, sizeof(double) * L->size, cudaMemcpyHostToDevice, s->c->stream));
  EXAMG_CUDA(cudaStreamSynchronize(s->c->stream));
  return EXAMG_OK;
}

int examg_stokes_download(examg_stokes *s, int level, int which, double *host) {
  StLevel *L = st_level(s, level);
  if (!L) return EXAMG_EINVAL;
  if (!st_field_ok(s, which) || !host) return set_error(EXAMG_EINVAL, "examg_stokes_download: field %d", which);
  EXAMG_CUDA(cudaMemcpyAsync(host, L->f[which], sizeof(double) * L->size, cudaMemcpyDeviceToHost, s->c->stream));
  EXAMG_CUDA(cudaStreamSynchronize(s->c->stream));
  return EXAMG_OK;
}

int examg_stokes_apply_bc(examg_stokes *s, int level) {
  StLevel *L = st_level(s, level);
  if (!L) return EXAMG_EINVAL;
  examg_ctx *c = s->c;
  const StArgs a = st_args(s, *L);
  const int n = L->n, dim = s->dim;
  const bool lid = level < 0 || level == c->cfg.max_level;
  // phase ph: component d handles the faces of dimension ph (Dirichlet 0.0, faces in the order -x,+x,-y,+y,-z,+z); u on
  // the finest level follows ApplyBC_u instead: phase 0 the ghost rows in y (lid on top), phase 1 the wall faces in x
  for (int ph = 0; ph < dim; ++ph) {
    StBcPhase p;
    for (int d = 0; d < 3; ++d) {
      p.e[d] = d < dim ? ph : -1;
      p.kind_lo[d] = p.kind_hi[d] = ph == d ? 0 : 1;
      p.g2_lo[d] = p.g2_hi[d] = 2.0 * 0.0;
      for (int t = 0; t < 3; ++t) { p.lo[d][t] = -1; p.hi[d][t] = t == d ? n + 1 : n; }
    }
    if (lid) {
      if (ph == 0) {
        p.e[0] = 1; p.kind_lo[0] = 2; p.kind_hi[0] = 1; p.g2_hi[0] = 2.0 * s->desc.top_vel;
        p.lo[0][0] = -1; p.hi[0][0] = n + 1; p.lo[0][2] = -1; p.hi[0][2] = n;
      } else if (ph == 1) {
        p.e[0] = 0; p.kind_lo[0] = p.kind_hi[0] = 0;
        for (int t = 0; t < 3; ++t) { p.lo[0][t] = 0; p.hi[0][t] = n - 1; }
      } else {
        p.e[0] = -1;
      }
    }
    long long pts = 2;
    for (int t = 0; t < dim - 1; ++t) pts *= n + 3;
    int bx = (int)((pts + 255) / 256);
    if (bx > 1024) bx = 1024;
    ST_DISPATCH(s, k_st_bc, dim3(bx, dim, 1), 256, a, p);
  }
  return EXAMG_OK;
}

int examg_stokes_smooth_color(examg_stokes *s, int level, int c0, int c1, int c2) {
  StLevel *L = st_level(s, level);
  if (!L) return EXAMG_EINVAL;
  if (c0 < 0 || c0 > 2 || c1 < 0 || c1 > 2 || c2 < 0 || c2 > (s->dim == 3 ? 2 : 0))
    return set_error(EXAMG_EINVAL, "examg_stokes_smooth_color: colour (%d, %d, %d)", c0, c1, c2);
  const int n = L->n;
  const int na = n > c0 ? (n - c0 + 2) / 3 : 0, nb = n > c1 ? (n - c1 + 2) / 3 : 0;
  const int nc = s->dim == 3 ? (n > c2 ? (n - c2 + 2) / 3 : 0) : 1;
  if (na <= 0 || nb <= 0 || nc <= 0) return EXAMG_OK;
  const dim3 block(32, 4, 1), grid((na + 31) / 32, (nb + 3) / 4, nc);
  ST_DISPATCH(s, k_st_smooth_color, grid, block, st_args(s, *L), c0, c1, c2, na, nb, nc);
  return EXAMG_OK;
}

int examg_stokes_smooth(examg_stokes *s, int level, int nsweeps) {
  if (!s) return set_error(EXAMG_EINVAL, "null Stokes solver");
  const int nz = s->dim == 3 ? 3 : 1;
  for (int it = 0; it < nsweeps; ++it) {
    for (int c2 = 0; c2 < nz; ++c2)
      for (int c1 = 0; c1 < 3; ++c1)
        for (int c0 = 0; c0 < 3; ++c0) {
          const int rc = examg_stokes_smooth_color(s, level, c0, c1, c2);
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
  const StArgs a = st_args(s, *L);
  ST_DISPATCH(s, k_st_residual, st_grid(s->dim, L->n), ST_BLOCK, a);
  if (!norms_out) return EXAMG_OK;
  long long cells = (long long)L->n * L->n * (s->dim == 3 ? L->n : 1);
  long long nb = (cells + 255) / 256;
  if (nb > ST_RED_BLOCKS) nb = ST_RED_BLOCKS;
  ST_DISPATCH(s, k_st_norms, (int)nb, 256, a, s->partials);
  k_st_norms_final<<<1, 32, 0, c->stream>>>(s->partials, (int)nb, s->dim + 1, s->sums);
  EXAMG_LAUNCH_CHECK(c);
  dd host[4];
  EXAMG_CUDA(cudaMemcpyAsync(host, s->sums, sizeof(dd) * (s->dim + 1), cudaMemcpyDeviceToHost, c->stream));
  EXAMG_CUDA(cudaStreamSynchronize(c->stream));
  for (int k = 0; k <= s->dim; ++k) norms_out[k] = sqrt(host[k].hi + host[k].lo);
  return EXAMG_OK;
}

int examg_stokes_restrict(examg_stokes *s, int level) {
  StLevel *F = st_level(s, level);
  if (!F) return EXAMG_EINVAL;
  if (F == &s->lev[0]) return set_error(EXAMG_EINVAL, "examg_stokes_restrict: no level below the coarsest");
  StLevel *C = F - 1;
  ST_DISPATCH(s, k_st_restrict, st_grid(s->dim, C->n), ST_BLOCK, st_args(s, *F), st_args(s, *C));
  return EXAMG_OK;
}

int examg_stokes_set_solution(examg_stokes *s, int level, double value) {
  StLevel *L = st_level(s, level);
  if (!L) return EXAMG_EINVAL;
  ST_DISPATCH(s, k_st_set, st_grid(s->dim, L->n + 2), ST_BLOCK, st_args(s, *L), value);
  return EXAMG_OK;
}

int examg_stokes_correct(examg_stokes *s, int level) {
  StLevel *F = st_level(s, level);
  if (!F) return EXAMG_EINVAL;
  if (F == &s->lev[0]) return set_error(EXAMG_EINVAL, "examg_stokes_correct: no level below the coarsest");
  StLevel *C = F - 1;
  ST_DISPATCH(s, k_st_correct, st_grid(s->dim, F->n), ST_BLOCK, st_args(s, *F), st_args(s, *C));
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
