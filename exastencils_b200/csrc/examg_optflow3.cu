// examg_optflow3.cu -- the 3D optical-flow program: a Vec3 unknown on cell centres, the operator
//   combinedOp = alpha**2 * Laplace (x) I_3 + gradContribution (a symmetric 3 x 3 block of cell fields at the centre),
// red-black sweeps `flow += inverse ( diag ( combinedOp ) ) * ( rhs - combinedOp * flow )`, cell restriction and
// piecewise-constant correction, CG on the coarsest level.
//
// Reference program: Testing/Application/OpticalFlow3D.exa4 (whole file; .knowledge: levels 2..6, 64^3 cells; golden
// OpticalFlow3D.results).  diag() of the matrix-valued stencil is its centre block; inverse() of a 3 x 3 expression is
// resolved at compile time by the cofactor formulas of C/s/e/baseExt/ir/IR_MatOperations/IR_CompiletimeMatOps.scala:804-826;
// Neumann ghost cells C/s/e/boundary/ir/IR_NeumannBC.scala:70-80; cell transfer
// C/s/e/operator/l4/L4_DefaultRestriction.scala:40-47, L4_DefaultProlongation.scala:30-45.
//
// Layout: every array of a level is dense (nz+2) x (ny+2) x (nx+2), x fastest, one ghost cell per side; cell (i,j,k) at
// (i+1) + (nx+2) ((j+1) + (ny+2) (k+1)).  The three components of flow / rhs / residual are separate arrays (SoA:
// the reference's Vec3 layout interleaves them; a warp reading one component of 32 cells then touches 3 x the lines).
// `apply bc` costs no launch: every kernel that writes a Neumann field writes the ghost cells next to the cells it
// updates (ghost = the inner cell with clamped indices, which is what the six face loops of the reference produce).
// Arithmetic contract = oracle/examg_oracle_optflow3.c, bit for bit for the pointwise loops (built with --fmad=false);
// sums in double-double.  One block (one GPU).
#include <cmath>
#include <cstdlib>
#include <map>
#include <vector>

#include "examg_cg.cuh"
#include "examg_device.cuh"

using namespace examg;

namespace {

constexpr int O3_NF = 15;  // flow 0..2, rhs 3..5, residual 6..8, G xx,xy,xz,yy,yz,zz 9..14
constexpr int O3_RED_BLOCKS = 1024;

struct O3Level {
  int n[3];
  long long sy, sz;
  size_t vol;
  double h[3], L[7];
  double *f[O3_NF];
  double *alt[3];  // second buffers of flow for the one-pass sweeps (levels of >= 32 cells in x; else nullptr)
  bool flipped;    // f[0..2] and alt[] have been exchanged an odd number of times
  // copies of the six gradContribution arrays with every row split into its even and its odd elements: a colour pass
  // reads consecutive doubles (levels of >= 2^15 cells with an even row length; else nullptr).  Valid after InitFields.
  double *gs[6];
  bool gs_valid;
};

// by value to the kernels
struct O3Args {
  int nx, ny, nz;
  long long sy, sz;
  double L[7];
  double a2;
  double *f[O3_NF];
  const double *gs[6];  // split-row copies of f[9..14] (nullptr: none)
};

// planes [o3_z0, o3_z1) of thread block z among gridDim.z
__device__ __forceinline__ int o3_z0(int nz) { return (int)((long long)nz * blockIdx.z / gridDim.z); }
__device__ __forceinline__ int o3_z1(int nz) { return (int)((long long)nz * (blockIdx.z + 1) / gridDim.z); }

__device__ __forceinline__ long long o3_idx(const O3Args &a, int i, int j, int k) {
  return (i + 1) + a.sy * (j + 1) + a.sz * (k + 1);
}

// w(i,j,k) = val and the Neumann ghost cells that mirror this cell (faces, edges, corner)
__device__ __forceinline__ void o3_store(const O3Args &a, double *w, int i, int j, int k, double val) {
  const int gi = i == 0 ? -1 : (i == a.nx - 1 ? a.nx : i), gj = j == 0 ? -1 : (j == a.ny - 1 ? a.ny : j);
  const int gk = k == 0 ? -1 : (k == a.nz - 1 ? a.nz : k);
  const bool ex = gi != i, ey = gj != j, ez = gk != k;
  w[o3_idx(a, i, j, k)] = val;
  if (ex) w[o3_idx(a, gi, j, k)] = val;
  if (ey) w[o3_idx(a, i, gj, k)] = val;
  if (ez) w[o3_idx(a, i, j, gk)] = val;
  if (ex && ey) w[o3_idx(a, gi, gj, k)] = val;
  if (ex && ez) w[o3_idx(a, gi, j, gk)] = val;
  if (ey && ez) w[o3_idx(a, i, gj, gk)] = val;
  if (ex && ey && ez) w[o3_idx(a, gi, gj, gk)] = val;
}

__device__ __forceinline__ double o3_lap(const O3Args &a, const double *w, long long p) {
  double acc = a.L[0] * w[p];
  acc = fma(a.L[1], w[p - 1], acc);
  acc = fma(a.L[2], w[p + 1], acc);
  acc = fma(a.L[3], w[p - a.sy], acc);
  acc = fma(a.L[4], w[p + a.sy], acc);
  acc = fma(a.L[5], w[p - a.sz], acc);
  acc = fma(a.L[6], w[p + a.sz], acc);
  return acc;
}

// rhs - combinedOp * w at p, all three components; g6 = gradContribution xx, xy, xz, yy, yz, zz at the cell
__device__ __forceinline__ void o3_residual_g(const O3Args &a, const double (&g6)[6], const double *w0, const double *w1,
                                              const double *w2, long long p, double (&r)[3]) {
  const double gxx = g6[0], gxy = g6[1], gxz = g6[2], gyy = g6[3], gyz = g6[4], gzz = g6[5];
  const double u = w0[p], v = w1[p], w = w2[p];
  double t = gxx * u;
  t = fma(gxy, v, t);
  t = fma(gxz, w, t);
  r[0] = a.f[3][p] - fma(a.a2, o3_lap(a, w0, p), t);
  t = gxy * u;
  t = fma(gyy, v, t);
  t = fma(gyz, w, t);
  r[1] = a.f[4][p] - fma(a.a2, o3_lap(a, w1, p), t);
  t = gxz * u;
  t = fma(gyz, v, t);
  t = fma(gzz, w, t);
  r[2] = a.f[5][p] - fma(a.a2, o3_lap(a, w2, p), t);
}

__device__ __forceinline__ void o3_residual_at(const O3Args &a, const double *w0, const double *w1, const double *w2,
                                               long long p, double (&r)[3]) {
  const double g6[6] = {a.f[9][p], a.f[10][p], a.f[11][p], a.f[12][p], a.f[13][p], a.f[14][p]};
  o3_residual_g(a, g6, w0, w1, w2, p, r);
}

// dst = src with every row of `row` doubles split into its even and its odd elements
__global__ void __launch_bounds__(256) k3_split_rows(const double *__restrict__ src, double *__restrict__ dst, int row,
                                                     long long total) {
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int ip = (int)(t % row);
    dst[t - ip + (ip & 1) * (row >> 1) + (ip >> 1)] = src[t];
  }
}

// combinedOp * w at p
__device__ __forceinline__ void o3_apply_at(const O3Args &a, const double *w0, const double *w1, const double *w2,
                                            long long p, double (&o)[3]) {
  const double gxx = a.f[9][p], gxy = a.f[10][p], gxz = a.f[11][p], gyy = a.f[12][p], gyz = a.f[13][p], gzz = a.f[14][p];
  const double u = w0[p], v = w1[p], w = w2[p];
  double t = gxx * u;
  t = fma(gxy, v, t);
  t = fma(gxz, w, t);
  o[0] = fma(a.a2, o3_lap(a, w0, p), t);
  t = gxy * u;
  t = fma(gyy, v, t);
  t = fma(gyz, w, t);
  o[1] = fma(a.a2, o3_lap(a, w1, p), t);
  t = gxz * u;
  t = fma(gyz, v, t);
  t = fma(gzz, w, t);
  o[2] = fma(a.a2, o3_lap(a, w2, p), t);
}

// every ghost cell = the inner cell with clamped indices (the images, once)
__global__ void __launch_bounds__(256) k3_ghost_fill(const O3Args a, double *w) {
  const long long total = (long long)(a.nx + 2) * (a.ny + 2) * (a.nz + 2);
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(t % (a.nx + 2)) - 1, j = (int)((t / (a.nx + 2)) % (a.ny + 2)) - 1;
    const int k = (int)(t / ((long long)(a.nx + 2) * (a.ny + 2))) - 1;
    if (i >= 0 && i < a.nx && j >= 0 && j < a.ny && k >= 0 && k < a.nz) continue;
    w[t] = w[o3_idx(a, min(max(i, 0), a.nx - 1), min(max(j, 0), a.ny - 1), min(max(k, 0), a.nz - 1))];
  }
}

// initGradients@finest + the right-hand side (OpticalFlow3D.exa4:348-412)
__global__ void __launch_bounds__(256) k3_init(const O3Args a, const double *__restrict__ img0, const double *__restrict__ img1,
                                               double gx1, double gx0, double gy1, double gy0, double gz1, double gz0) {
  const long long total = (long long)a.nx * a.ny * a.nz;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(t % a.nx), j = (int)((t / a.nx) % a.ny), k = (int)(t / ((long long)a.nx * a.ny));
    const long long p = o3_idx(a, i, j, k);
    const double c = img0[p];
    const double ix = fma(gx0, c, gx1 * img0[p + 1]);  // Grad: [+1] then [0]
    const double iy = fma(gy0, c, gy1 * img0[p + a.sy]);
    const double iz = fma(gz0, c, gz1 * img0[p + a.sz]);
    const double it = img1[p] - c;
    a.f[9][p] = ix * ix;
    a.f[10][p] = ix * iy;
    a.f[11][p] = ix * iz;
    a.f[12][p] = iy * iy;
    a.f[13][p] = iy * iz;
    a.f[14][p] = iz * iz;
    a.f[3][p] = -ix * it;
    a.f[4][p] = -iy * it;
    a.f[5][p] = -iz * it;
  }
}

struct O3Pairs {
  const double *src[6];
  double *dst[6];
};

// coarse = 0.125 * (sum of the 8 children in lexicographic z, y, x order); blockIdx.y = field of the list
__global__ void __launch_bounds__(256) k3_restrict(const O3Pairs pr, int nxc, int nyc, int nzc, long long syf, long long szf,
                                                   long long syc, long long szc) {
  const double *__restrict__ f = pr.src[blockIdx.y];
  double *__restrict__ cd = pr.dst[blockIdx.y];
  const long long total = (long long)nxc * nyc * nzc;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(t % nxc), j = (int)((t / nxc) % nyc), k = (int)(t / ((long long)nxc * nyc));
    const double *q = f + (2 * i + 1) + syf * (2 * j + 1) + szf * (2 * k + 1);
    double s = q[0] + q[1];
    s += q[syf];
    s += q[syf + 1];
    s += q[szf];
    s += q[szf + 1];
    s += q[szf + syf];
    s += q[szf + syf + 1];
    cd[(i + 1) + syc * (j + 1) + szc * (k + 1)] = 0.125 * s;
  }
}

// flow += ProlongationCell * flow@coarser (piecewise constant); apply bc
__global__ void __launch_bounds__(256) k3_correct(const O3Args a, const double *__restrict__ c0, const double *__restrict__ c1,
                                                  const double *__restrict__ c2, long long syc, long long szc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y * blockDim.y + threadIdx.y;
  if (i >= a.nx || j >= a.ny) return;
  for (int k = o3_z0(a.nz), k1 = o3_z1(a.nz); k < k1; ++k) {  // a contiguous chunk of planes: z neighbours stay in L1
    const long long p = o3_idx(a, i, j, k), q = ((i >> 1) + 1) + syc * ((j >> 1) + 1) + szc * ((k >> 1) + 1);
    o3_store(a, a.f[0], i, j, k, a.f[0][p] + c0[q]);
    o3_store(a, a.f[1], i, j, k, a.f[1][p] + c1[q]);
    o3_store(a, a.f[2], i, j, k, a.f[2][p] + c2[q]);
  }
}

// residual = rhs - combinedOp * flow; apply bc; NORM: partial sums of dot ( residual, residual )
// thread blocks of 64 x 4 cells in x, y; planes by grid stride (no integer division per cell)
template <bool NORM>
__global__ void __launch_bounds__(256) k3_residual(const O3Args a, dd *__restrict__ partials) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y * blockDim.y + threadIdx.y;
  dd part = {0.0, 0.0};
  if (i < a.nx && j < a.ny)
    for (int k = o3_z0(a.nz), k1 = o3_z1(a.nz); k < k1; ++k) {  // a contiguous chunk of planes: z neighbours stay in L1
      const long long p = o3_idx(a, i, j, k);
      double r[3];
      o3_residual_at(a, a.f[0], a.f[1], a.f[2], p, r);
      o3_store(a, a.f[6], i, j, k, r[0]);
      o3_store(a, a.f[7], i, j, k, r[1]);
      o3_store(a, a.f[8], i, j, k, r[2]);
      if (NORM) dd_add(part, r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
    }
  if (NORM) {
    part = dd_block_reduce(part);
    if (threadIdx.x == 0 && threadIdx.y == 0) partials[blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z)] = part;
  }
}

__global__ void __launch_bounds__(1024) k3_final_sum(const dd *__restrict__ partials, int n, dd *out) {
  dd v = {0.0, 0.0};
  for (int i = threadIdx.x; i < n; i += blockDim.x) dd_merge(v, partials[i]);
  v = dd_block_reduce(v);
  if (threadIdx.x == 0) *out = v;
}

// the inverse of the 3 x 3 centre block by cofactors, applied to r
// G = xx, xy, xz, yy, yz, zz at the cell
__device__ __forceinline__ void o3_block_solve(const O3Args &a, const double (&G6)[6], const double (&r)[3], double (&x)[3]) {
  // rows (a b c), (d e f), (g h i) with d = b, g = c, h = f
  const double a_ = fma(a.a2, a.L[0], G6[0]), b = G6[1], c = G6[2];
  const double d = b, e = fma(a.a2, a.L[0], G6[3]), f = G6[4];
  const double g = c, h = f, i_ = fma(a.a2, a.L[0], G6[5]);
  const double A = e * i_ - f * h, B = -(d * i_ - f * g), C = d * h - e * g;
  const double D = -(b * i_ - c * h), E = a_ * i_ - c * g, F = -(a_ * h - b * g);
  const double G = b * f - c * e, H = -(a_ * f - c * d), I = a_ * e - b * d;
  const double det = a_ * A + b * B + c * C;
  // d = b, g = c, h = f: D = B, G = C, H = F bit for bit (the products commute) -- six divisions
  (void)D; (void)G; (void)H;
  const double i00 = A / det, i01 = B / det, i02 = C / det, i11 = E / det, i12 = F / det, i22 = I / det;
  x[0] = i00 * r[0] + i01 * r[1] + i02 * r[2];
  x[1] = i01 * r[0] + i11 * r[1] + i12 * r[2];
  x[2] = i02 * r[0] + i12 * r[1] + i22 * r[2];
}

// one colour: flow += inverse ( diag ( combinedOp ) ) * ( rhs - combinedOp * flow ) on the cells with
// (i + j + k) % 2 == color; apply bc
// SPLIT: gradContribution from the split-row copies (the lanes of a warp read consecutive doubles)
template <bool SPLIT>
__global__ void __launch_bounds__(256) k3_smooth_color(const O3Args a, int color) {
  const int tx = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y * blockDim.y + threadIdx.y;
  if (j >= a.ny) return;
  for (int k = o3_z0(a.nz), k1 = o3_z1(a.nz); k < k1; ++k) {  // a contiguous chunk of planes: z neighbours stay in L1
    const int i = 2 * tx + ((j + k + color) & 1);
    if (i >= a.nx) continue;
    const long long p = o3_idx(a, i, j, k);
    double r[3], d[3], g6[6];
    if (SPLIT) {
      const int ip = i + 1;
      const long long sp = p - ip + (ip & 1) * ((a.nx + 2) >> 1) + (ip >> 1);
#pragma unroll
      for (int q = 0; q < 6; ++q) g6[q] = a.gs[q][sp];
    } else {
#pragma unroll
      for (int q = 0; q < 6; ++q) g6[q] = a.f[9 + q][p];
    }
    o3_residual_g(a, g6, a.f[0], a.f[1], a.f[2], p, r);
    o3_block_solve(a, g6, r, d);
    o3_store(a, a.f[0], i, j, k, a.f[0][p] + d[0]);
    o3_store(a, a.f[1], i, j, k, a.f[1][p] + d[1]);
    o3_store(a, a.f[2], i, j, k, a.f[2][p] + d[2]);
  }
}

// ---- one whole sweep (colour 0, then colour 1) in one pass over HBM ---------------------------------------------------
// The per-colour kernel fetches every sector of the 15 arrays twice per sweep and uses half of it each time.  Here a CTA of
// 32 x 16 threads owns a tile (thread = cell column) and marches in z with a ring of five planes of the three flow
// components in shared memory (34 x 18 with the halo).  Colour 0 ("red") of plane t is updated at step t from the old
// values of planes t-1, t, t+1 -- in the ring, in place; colour 1 ("black") of plane t-2 at the same step from the
// updated red cells of planes t-3, t-2, t-1 (see the schedule below); plane t-2 is final then and goes to the second
// buffers of flow (out of place: neighbouring tiles recompute the red cells of their one-cell ring from old values).
// A cell's coefficients and right-hand side are read once, one plane ahead, into registers.  Neumann ghost cells in the
// ring never need a refresh: a ghost cell mirrors the very cell that reads it.  Same per-cell arithmetic as
// k3_smooth_color => bit-identical.
constexpr int F3W = 32, F3H = 16, F3NT = F3W * F3H;  // threads = red region of a tile
constexpr int F3IW = F3W - 2, F3IH = F3H - 2;        // output region
constexpr int F3SW = F3W + 2, F3SH = F3H + 2, F3PL = F3SW * F3SH;  // staged plane with halo
constexpr int F3SL = 5;  // ring slots: planes t-3 .. t+1
constexpr size_t F3_SMEM = sizeof(double) * 3 * F3SL * F3PL;

struct O3Fused {
  double *dst[3];
  int tiles_x, tiles_y, nseg, seg_planes;
};

__global__ void __launch_bounds__(F3NT, F3NT <= 256 ? 2 : 1) k3_sweep(const O3Args a, const O3Fused fz) {
  extern __shared__ double ring[];  // [comp][slot][F3PL]
  const int tid = threadIdx.x, tx = tid % F3W, ty = tid / F3W;
  const int tile = blockIdx.x % (fz.tiles_x * fz.tiles_y), seg = blockIdx.x / (fz.tiles_x * fz.tiles_y);
  const int ox = F3IW * (tile % fz.tiles_x), oy = F3IH * (tile / fz.tiles_x);
  const int za = seg * fz.seg_planes, zb = min(za + fz.seg_planes, a.nz);
  const int i = ox - 1 + tx, j = oy - 1 + ty;
  const bool indom = i >= 0 && i < a.nx && j >= 0 && j < a.ny;
  const bool interior = indom && tx >= 1 && tx <= F3IW && ty >= 1 && ty <= F3IH;
  const int lo = (ty + 1) * F3SW + tx + 1;  // own cell inside a staged plane
  const long long col = (i + 1) + a.sy * (j + 1);  // own column, plane k at col + sz (k + 1)
  // halo elements staged by this thread: e = tid and tid + F3NT
  const int e0 = tid, e1 = tid + F3NT;
  const int gx0 = ox - 2 + e0 % F3SW, gy0 = oy - 2 + e0 / F3SW, gx1 = ox - 2 + e1 % F3SW, gy1 = oy - 2 + e1 / F3SW;
  const bool v0 = gx0 >= -1 && gx0 <= a.nx && gy0 >= -1 && gy0 <= a.ny;
  const bool v1 = e1 < F3PL && gx1 >= -1 && gx1 <= a.nx && gy1 >= -1 && gy1 <= a.ny;
  const long long c0 = (gx0 + 1) + a.sy * (gy0 + 1), c1 = (gx1 + 1) + a.sy * (gy1 + 1);

  auto slot = [](int plane) { return (plane + 2 * F3SL) % F3SL; };
  auto stage = [&](int plane) {  // flow planes -> ring
    const bool pv = plane >= -1 && plane <= a.nz;
    const long long zo = a.sz * (plane + 1);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      double *dstp = ring + (c * F3SL + slot(plane)) * F3PL;
      dstp[e0] = (pv && v0) ? a.f[c][c0 + zo] : 0.0;
      if (e1 < F3PL) dstp[e1] = (pv && v1) ? a.f[c][c1 + zo] : 0.0;
    }
  };
  // coefficients and right-hand side of the own cell: G xx,xy,xz,yy,yz,zz, rhs 0..2
  auto fetch = [&](int plane, double (&cf)[9]) {
    if (indom && plane >= 0 && plane < a.nz) {
      const long long p = col + a.sz * (plane + 1);
#pragma unroll
      for (int q = 0; q < 6; ++q) cf[q] = a.f[9 + q][p];
#pragma unroll
      for (int q = 0; q < 3; ++q) cf[6 + q] = a.f[3 + q][p];
    }
  };
  // the update of the own cell on `plane`, in the ring
  auto update = [&](int plane, const double (&cf)[9]) {
    const double *m[3], *c[3], *pp[3];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      m[q] = ring + (q * F3SL + slot(plane - 1)) * F3PL + lo;
      c[q] = ring + (q * F3SL + slot(plane)) * F3PL + lo;
      pp[q] = ring + (q * F3SL + slot(plane + 1)) * F3PL + lo;
    }
    auto lap = [&](int q) {
      double acc = a.L[0] * c[q][0];
      acc = fma(a.L[1], c[q][-1], acc);
      acc = fma(a.L[2], c[q][1], acc);
      acc = fma(a.L[3], c[q][-F3SW], acc);
      acc = fma(a.L[4], c[q][F3SW], acc);
      acc = fma(a.L[5], m[q][0], acc);
      acc = fma(a.L[6], pp[q][0], acc);
      return acc;
    };
    const double u = c[0][0], v = c[1][0], w = c[2][0];
    double r[3], x[3];
    double t = cf[0] * u;
    t = fma(cf[1], v, t);
    t = fma(cf[2], w, t);
    r[0] = cf[6] - fma(a.a2, lap(0), t);
    t = cf[1] * u;
    t = fma(cf[3], v, t);
    t = fma(cf[4], w, t);
    r[1] = cf[7] - fma(a.a2, lap(1), t);
    t = cf[2] * u;
    t = fma(cf[4], v, t);
    t = fma(cf[5], w, t);
    r[2] = cf[8] - fma(a.a2, lap(2), t);
    const double g6[6] = {cf[0], cf[1], cf[2], cf[3], cf[4], cf[5]};
    o3_block_solve(a, g6, r, x);
    const_cast<double *>(c[0])[0] = u + x[0];
    const_cast<double *>(c[1])[0] = v + x[1];
    const_cast<double *>(c[2])[0] = w + x[2];
  };

  // Schedule without divergence: at step t the lanes with (i + j + t) even update colour 0 on plane t, the others colour 1
  // on plane t - 2 (whose red neighbours on planes t-3, t-2, t-1 are done) -- every lane runs the same update on its own
  // (plane, coefficients).  A lane holds the coefficients it uses now, the set it uses next and the set in flight:
  //   even at t: now = plane t,   next = plane t-1 (it is odd at t+1: colour 1 on plane t-1)
  //   odd  at t: now = plane t-2, next = plane t+1 (in flight; it is even at t+1)
  const int t0 = max(za - 1, 0), t1 = zb + 1;  // colour 0 on planes t0 .. min(zb, nz - 1), colour 1 on za .. zb - 1
  double use[9], saved[9], nxt[9];
#pragma unroll
  for (int q = 0; q < 9; ++q) use[q] = saved[q] = nxt[q] = 0.0;
  for (int pl = t0 - 3; pl <= t0 + 1; ++pl) stage(pl);
  const bool even0 = ((i + j + t0) & 1) == 0;
  // even at t0: now = plane t0, saved = plane t0 - 1; odd at t0: now = plane t0 - 2, saved = plane t0
  fetch(even0 ? t0 : t0 - 2, use);
  fetch(even0 ? t0 - 1 : t0, saved);
  __syncthreads();
  for (int t = t0; t <= t1; ++t) {
    const bool even = ((i + j + t) & 1) == 0;
    // requests of the next step first: plane t+2 of flow, plane t+1 of the coefficients
    double h0[3], h1[3];
    {
      const int plane = t + 2;
      const bool pv = plane <= a.nz && t < t1;
      const long long zo = a.sz * (plane + 1);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        h0[c] = (pv && v0) ? a.f[c][c0 + zo] : 0.0;
        h1[c] = (pv && v1) ? a.f[c][c1 + zo] : 0.0;
      }
    }
    if (t < t1) fetch(t + 1, nxt);
    const int plane = even ? t : t - 2;
    const bool act = even ? (indom && t < a.nz && t <= zb) : (interior && plane >= za && plane < zb);
    if (act) update(plane, use);
    // plane t - 2 is final
    const int pz = t - 2;
    if (interior && pz >= za && pz < zb) {
#pragma unroll
      for (int c = 0; c < 3; ++c) o3_store(a, fz.dst[c], i, j, pz, ring[(c * F3SL + slot(pz)) * F3PL + lo]);
    }
    __syncthreads();
    if (t < t1) {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        double *dstp = ring + (c * F3SL + slot(t + 2)) * F3PL;
        dstp[e0] = h0[c];
        if (e1 < F3PL) dstp[e1] = h1[c];
      }
#pragma unroll
      for (int q = 0; q < 9; ++q) {
        const double sv = saved[q], nx = nxt[q];
        use[q] = even ? sv : nx;
        saved[q] = even ? nx : sv;
      }
    }
    __syncthreads();
  }
}

// mgCycle@coarsest (OpticalFlow3D.exa4:266-318): the whole CG in one CTA (4^3 cells in the reference's configuration)
// p, ap: cgTmp0, cgTmp1 (three arrays each)
struct O3Cg {
  double *p[3], *ap[3];
};

__global__ void __launch_bounds__(1024) k3_cg(const O3Args a, const O3Cg w, int max_it, double rel_tol, int *iters_out) {
  const int total = a.nx * a.ny * a.nz, tid = threadIdx.x, nt = blockDim.x;
  double *r0 = a.f[6], *r1 = a.f[7], *r2 = a.f[8];
  dd part = {0.0, 0.0};
  for (int t = tid; t < total; t += nt) {
    const int i = t % a.nx, j = (t / a.nx) % a.ny, k = t / (a.nx * a.ny);
    const long long p = o3_idx(a, i, j, k);
    double r[3];
    o3_residual_at(a, a.f[0], a.f[1], a.f[2], p, r);
    o3_store(a, r0, i, j, k, r[0]);
    o3_store(a, r1, i, j, k, r[1]);
    o3_store(a, r2, i, j, k, r[2]);
    o3_store(a, w.p[0], i, j, k, r[0]);
    o3_store(a, w.p[1], i, j, k, r[1]);
    o3_store(a, w.p[2], i, j, k, r[2]);
    dd_add(part, r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  }
  double curRes = sqrt(cg_sum(part));
  const double initRes = curRes;
  int iters = 0;
  for (int step = 0; step < max_it; ++step) {
    iters = step + 1;
    dd pn = {0.0, 0.0}, pd = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const int i = t % a.nx, j = (t / a.nx) % a.ny, k = t / (a.nx * a.ny);
      const long long p = o3_idx(a, i, j, k);
      double o[3];
      o3_apply_at(a, w.p[0], w.p[1], w.p[2], p, o);
      w.ap[0][p] = o[0];
      w.ap[1][p] = o[1];
      w.ap[2][p] = o[2];
      dd_add(pn, r0[p] * r0[p] + r1[p] * r1[p] + r2[p] * r2[p]);
      dd_add(pd, w.p[0][p] * o[0] + w.p[1][p] * o[1] + w.p[2][p] * o[2]);
    }
    const double nom = cg_sum(pn), den = cg_sum(pd);
    const double alpha = nom / den;
    dd pr = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const int i = t % a.nx, j = (t / a.nx) % a.ny, k = t / (a.nx * a.ny);
      const long long p = o3_idx(a, i, j, k);
      o3_store(a, a.f[0], i, j, k, fma(alpha, w.p[0][p], a.f[0][p]));
      o3_store(a, a.f[1], i, j, k, fma(alpha, w.p[1][p], a.f[1][p]));
      o3_store(a, a.f[2], i, j, k, fma(alpha, w.p[2][p], a.f[2][p]));
      const double n0 = fma(-alpha, w.ap[0][p], r0[p]), n1 = fma(-alpha, w.ap[1][p], r1[p]), n2 = fma(-alpha, w.ap[2][p], r2[p]);
      o3_store(a, r0, i, j, k, n0);
      o3_store(a, r1, i, j, k, n1);
      o3_store(a, r2, i, j, k, n2);
      dd_add(pr, n0 * n0 + n1 * n1 + n2 * n2);
    }
    const double nextRes = sqrt(cg_sum(pr));
    if (nextRes <= rel_tol * initRes) break;
    const double beta = (nextRes * nextRes) / (curRes * curRes);
    for (int t = tid; t < total; t += nt) {
      const int i = t % a.nx, j = (t / a.nx) % a.ny, k = t / (a.nx * a.ny);
      const long long p = o3_idx(a, i, j, k);
      o3_store(a, w.p[0], i, j, k, fma(beta, w.p[0][p], r0[p]));
      o3_store(a, w.p[1], i, j, k, fma(beta, w.p[1][p], r1[p]));
      o3_store(a, w.p[2], i, j, k, fma(beta, w.p[2][p], r2[p]));
    }
    curRes = nextRes;
    __syncthreads();
  }
  if (tid == 0 && iters_out) *iters_out = iters;
}

}  // namespace

struct examg_optflow3 {
  examg_ctx *c = nullptr;
  examg_optflow3_desc desc;
  std::vector<O3Level> lev;
  double *img[2] = {nullptr, nullptr};  // finest
  double *cg[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // coarsest: cgTmp0, cgTmp1
  dd *partials = nullptr, *sum = nullptr;
  dd *sum_host = nullptr;
  int *iters = nullptr, *iters_host = nullptr;
  // mgCycle@finest as a CUDA graph per buffer state (the one-pass sweeps exchange flow with its second buffers)
  struct Graph { cudaGraphExec_t exec = nullptr; long long launches = 0; std::vector<char> exit_state; };
  std::map<std::vector<char>, Graph> graphs;
};

namespace {

O3Level *o3_level(examg_optflow3 *o, int level) {
  if (!o) { set_error(EXAMG_EINVAL, "null optflow3 object"); return nullptr; }
  if (level < 0) level = o->c->cfg.max_level;
  if (level < o->c->cfg.min_level || level > o->c->cfg.max_level) {
    set_error(EXAMG_EINVAL, "optflow3: level %d outside %d..%d", level, o->c->cfg.min_level, o->c->cfg.max_level);
    return nullptr;
  }
  return &o->lev[level - o->c->cfg.min_level];
}

O3Args o3_args(const examg_optflow3 *o, const O3Level &L) {
  O3Args a;
  a.nx = L.n[0];
  a.ny = L.n[1];
  a.nz = L.n[2];
  a.sy = L.sy;
  a.sz = L.sz;
  for (int k = 0; k < 7; ++k) a.L[k] = L.L[k];
  a.a2 = o->desc.alpha * o->desc.alpha;  // alpha ** 2
  for (int k = 0; k < O3_NF; ++k) a.f[k] = L.f[k];
  for (int k = 0; k < 6; ++k) a.gs[k] = L.gs_valid ? L.gs[k] : nullptr;
  return a;
}

// 64 x 4 cells (x, y) per thread block, a chunk of planes each
dim3 o3_shape(int ex, int ny, int nz) {
  const int gx = (ex + 63) / 64, gy = (ny + 3) / 4;
  // measured at 256^3 / 512^3: 2048 blocks 0.458 / 3.65 ms for the residual, 8192 blocks 0.406 / 3.20 ms (less of a tail)
  static const int target = getenv("EXAMG_OF3_BLOCKS") ? atoi(getenv("EXAMG_OF3_BLOCKS")) : 8192;
  int gz = target / (gx * gy);
  gz = gz < 1 ? 1 : (gz > nz ? nz : gz);
  // chunks of at least 8 planes (L1 reuse of the z neighbours) where that still leaves enough thread blocks
  if (gz > nz / 8 && (long long)gx * gy * (nz / 8) >= 1024) gz = nz / 8;
  return dim3(gx, gy, gz);
}

int o3_grid(long long total) {
  long long g = (total + 255) / 256;
  return (int)(g < 1 ? 1 : (g > O3_RED_BLOCKS ? O3_RED_BLOCKS : g));
}

// which -> device array and its element count (0: bad id)
double *o3_array(examg_optflow3 *o, int level, int which, size_t *count) {
  if (which == EXAMG_OF3_IMG0 || which == EXAMG_OF3_IMG1) {
    *count = o->lev.back().vol;
    return o->img[which - EXAMG_OF3_IMG0];
  }
  O3Level *L = o3_level(o, level);
  if (!L || which < 0 || which >= O3_NF) { *count = 0; return nullptr; }
  *count = L->vol;
  return L->f[which];
}

int o3_smooth_color(examg_optflow3 *o, O3Level &L, int color) {
  examg_ctx *c = o->c;
  const O3Args a = o3_args(o, L);
  if (a.gs[0])
    k3_smooth_color<true><<<o3_shape((L.n[0] + 1) / 2, L.n[1], L.n[2]), dim3(64, 4), 0, c->stream>>>(a, color);
  else
    k3_smooth_color<false><<<o3_shape((L.n[0] + 1) / 2, L.n[1], L.n[2]), dim3(64, 4), 0, c->stream>>>(a, color);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

// Opt-in (EXAMG_OF3_FUSED=1): measured on B200 at 256^3 the one-pass sweep takes 0.98 ms against 0.84 ms for the two
// colour launches -- 27 doubles of per-thread state force one 512-thread CTA per SM (128 registers), and with 25 % of
// the warp slots the dependent FP64 chains of the block inverse, the two barriers per plane and the global loads are
// not hidden (ncu: stalls barrier 2.3, wait 2.3, long scoreboard 2.0 warps per issue, FP64 pipe 16 %, DRAM 5 % on a
// 128^3 level); profiles/r02_notes.md.
bool o3_fused_on() {
  static const bool on = getenv("EXAMG_OF3_FUSED") && getenv("EXAMG_OF3_FUSED")[0] == '1';
  return on;
}

// one sweep: colour 0, colour 1 -- in one pass where the level has second buffers
int o3_sweep(examg_optflow3 *o, O3Level &L) {
  examg_ctx *c = o->c;
  if (!L.alt[0] || !o3_fused_on()) {
    int rc;
    if ((rc = o3_smooth_color(o, L, 0)) || (rc = o3_smooth_color(o, L, 1))) return rc;
    return EXAMG_OK;
  }
  EXAMG_ONCE_PER_DEVICE(c, {
    EXAMG_CUDA(cudaFuncSetAttribute(k3_sweep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)F3_SMEM));
  });
  O3Fused fz;
  for (int k = 0; k < 3; ++k) fz.dst[k] = L.alt[k];
  fz.tiles_x = (L.n[0] + F3IW - 1) / F3IW;
  fz.tiles_y = (L.n[1] + F3IH - 1) / F3IH;
  const int tiles = fz.tiles_x * fz.tiles_y;
  int nseg = (4 * 148 + tiles - 1) / tiles;  // about four CTAs per SM over the launch, segments of >= 16 planes
  if (nseg > L.n[2] / 16) nseg = L.n[2] / 16;
  if (nseg < 1) nseg = 1;
  fz.seg_planes = (L.n[2] + nseg - 1) / nseg;
  fz.nseg = (L.n[2] + fz.seg_planes - 1) / fz.seg_planes;
  k3_sweep<<<tiles * fz.nseg, F3NT, F3_SMEM, c->stream>>>(o3_args(o, L), fz);
  EXAMG_LAUNCH_CHECK(c);
  for (int k = 0; k < 3; ++k) {
    double *t = L.f[k];
    L.f[k] = L.alt[k];
    L.alt[k] = t;
  }
  L.flipped = !L.flipped;
  return EXAMG_OK;
}

std::vector<char> o3_state(const examg_optflow3 *o) {
  std::vector<char> key;
  for (const O3Level &L : o->lev) key.push_back(L.flipped ? 1 : 0);
  return key;
}

void o3_restore(examg_optflow3 *o, const std::vector<char> &key) {
  for (size_t l = 0; l < o->lev.size(); ++l) {
    O3Level &L = o->lev[l];
    if ((L.flipped ? 1 : 0) == key[l]) continue;
    for (int k = 0; k < 3; ++k) {
      double *t = L.f[k];
      L.f[k] = L.alt[k];
      L.alt[k] = t;
    }
    L.flipped = !L.flipped;
  }
}

int o3_residual(examg_optflow3 *o, O3Level &L, bool norm) {
  examg_ctx *c = o->c;
  const dim3 grid = o3_shape(L.n[0], L.n[1], L.n[2]);
  if (norm) {
    k3_residual<true><<<grid, dim3(64, 4), 0, c->stream>>>(o3_args(o, L), o->partials);
    EXAMG_LAUNCH_CHECK(c);
    k3_final_sum<<<1, 1024, 0, c->stream>>>(o->partials, (int)(grid.x * grid.y * grid.z), o->sum);
    EXAMG_LAUNCH_CHECK(c);
  } else {
    k3_residual<false><<<grid, dim3(64, 4), 0, c->stream>>>(o3_args(o, L), nullptr);
    EXAMG_LAUNCH_CHECK(c);
  }
  return EXAMG_OK;
}

// n fields of `fine` -> `coarse`
int o3_restrict(examg_optflow3 *o, const O3Level &f, const O3Level &cl, const int *src, const int *dst, int n) {
  examg_ctx *c = o->c;
  O3Pairs pr;
  for (int k = 0; k < n; ++k) {
    pr.src[k] = f.f[src[k]];
    pr.dst[k] = cl.f[dst[k]];
  }
  const dim3 grid(o3_grid((long long)cl.n[0] * cl.n[1] * cl.n[2]), n);
  k3_restrict<<<grid, 256, 0, c->stream>>>(pr, cl.n[0], cl.n[1], cl.n[2], f.sy, f.sz, cl.sy, cl.sz);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

int o3_cg(examg_optflow3 *o) {
  examg_ctx *c = o->c;
  O3Level &L = o->lev.front();
  O3Cg w;
  for (int k = 0; k < 3; ++k) {
    w.p[k] = o->cg[k];
    w.ap[k] = o->cg[3 + k];
  }
  int nt = 32;
  while (nt < 1024 && nt < L.n[0] * L.n[1] * L.n[2]) nt <<= 1;
  k3_cg<<<1, nt, 0, c->stream>>>(o3_args(o, L), w, o->desc.cg_max_it, o->desc.cg_rel_tol, o->iters);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

int o3_restrict_residual(examg_optflow3 *o, int level) {
  examg_ctx *c = o->c;
  O3Level &L = o->lev[level - c->cfg.min_level], &C = o->lev[level - 1 - c->cfg.min_level];
  static const int src[3] = {6, 7, 8}, dst[3] = {3, 4, 5};
  int rc = o3_restrict(o, L, C, src, dst, 3);
  if (rc) return rc;
  for (int k = 0; k < 3; ++k) {  // flow@coarser = 0; apply bc: the whole arrays
    EXAMG_CUDA(cudaMemsetAsync(C.f[k], 0, sizeof(double) * C.vol, c->stream));
    c->launches++;
  }
  return EXAMG_OK;
}

int o3_correct(examg_optflow3 *o, int level) {
  examg_ctx *c = o->c;
  O3Level &L = o->lev[level - c->cfg.min_level], &C = o->lev[level - 1 - c->cfg.min_level];
  k3_correct<<<o3_shape(L.n[0], L.n[1], L.n[2]), dim3(64, 4), 0, c->stream>>>(o3_args(o, L), C.f[0], C.f[1], C.f[2], C.sy, C.sz);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

int o3_vcycle_rec(examg_optflow3 *o, int level) {
  examg_ctx *c = o->c;
  if (level == c->cfg.min_level) return o3_cg(o);
  O3Level &L = o->lev[level - c->cfg.min_level];
  int rc;
  for (int it = 0; it < o->desc.nu_pre; ++it)
    if ((rc = o3_sweep(o, L))) return rc;
  if ((rc = o3_residual(o, L, false)) || (rc = o3_restrict_residual(o, level))) return rc;
  if ((rc = o3_vcycle_rec(o, level - 1)) || (rc = o3_correct(o, level))) return rc;
  for (int it = 0; it < o->desc.nu_post; ++it)
    if ((rc = o3_sweep(o, L))) return rc;
  return EXAMG_OK;
}

}  // namespace

extern "C" {

examg_optflow3 *examg_optflow3_create(examg_ctx *c, const examg_optflow3_desc *d) {
  if (!c || !d) { set_error(EXAMG_EINVAL, "examg_optflow3_create: null argument"); return nullptr; }
  if (c->cfg.dim != 3 || c->world != 1 || c->periodic_any) {
    set_error(EXAMG_EUNSUP, "examg_optflow3_create: 3D, one block, no periodic dimension");
    return nullptr;
  }
  if (d->nu_pre < 0 || d->nu_post < 0 || d->cg_max_it < 1) { set_error(EXAMG_EINVAL, "examg_optflow3_create: sweep / iteration counts"); return nullptr; }
  for (int t = 0; t < 3; ++t)
    if ((c->cfg.block_cells[t] << c->cfg.min_level) < 2) {
      set_error(EXAMG_EUNSUP, "examg_optflow3_create: at least 2 cells per dimension on the coarsest level");
      return nullptr;
    }
  examg_optflow3 *o = new examg_optflow3;
  o->c = c;
  o->desc = *d;
  bool ok = true;
  for (int level = c->cfg.min_level; level <= c->cfg.max_level && ok; ++level) {
    O3Level L;
    for (int t = 0; t < 3; ++t) {
      L.n[t] = c->cfg.block_cells[t] << level;
      L.h[t] = (c->cfg.domain_hi[t] - c->cfg.domain_lo[t]) / (double)L.n[t];
    }
    L.sy = L.n[0] + 2;
    L.sz = L.sy * (L.n[1] + 2);
    L.vol = (size_t)L.sz * (size_t)(L.n[2] + 2);
    const double hx2 = L.h[0] * L.h[0], hy2 = L.h[1] * L.h[1], hz2 = L.h[2] * L.h[2];
    L.L[0] = 2.0 / hx2 + 2.0 / hy2 + 2.0 / hz2;  // Stencil Laplace@all, declaration order
    L.L[1] = L.L[2] = -1.0 / hx2;
    L.L[3] = L.L[4] = -1.0 / hy2;
    L.L[5] = L.L[6] = -1.0 / hz2;
    for (int k = 0; k < O3_NF; ++k) {
      L.f[k] = nullptr;
      ok = ok && cudaMalloc(&L.f[k], sizeof(double) * L.vol) == cudaSuccess &&
           cudaMemsetAsync(L.f[k], 0, sizeof(double) * L.vol, c->stream) == cudaSuccess;
    }
    L.gs_valid = false;
    static const bool split_on = !(getenv("EXAMG_OF3_SPLIT") && getenv("EXAMG_OF3_SPLIT")[0] == '0');
    for (int k = 0; k < 6; ++k) {
      L.gs[k] = nullptr;
      if (split_on && L.n[0] % 2 == 0 && (long long)L.n[0] * L.n[1] * L.n[2] >= (1 << 15))
        ok = ok && cudaMalloc(&L.gs[k], sizeof(double) * L.vol) == cudaSuccess;
    }
    L.flipped = false;
    for (int k = 0; k < 3; ++k) {
      L.alt[k] = nullptr;
      if (L.n[0] >= F3W && L.n[1] >= F3H && L.n[2] >= 4 && o3_fused_on())
        ok = ok && cudaMalloc(&L.alt[k], sizeof(double) * L.vol) == cudaSuccess &&
             cudaMemsetAsync(L.alt[k], 0, sizeof(double) * L.vol, c->stream) == cudaSuccess;
    }
    o->lev.push_back(L);
  }
  if (ok) {
    const size_t vf = o->lev.back().vol, vc = o->lev.front().vol;
    for (int k = 0; k < 2 && ok; ++k)
      ok = cudaMalloc(&o->img[k], sizeof(double) * vf) == cudaSuccess && cudaMemsetAsync(o->img[k], 0, sizeof(double) * vf, c->stream) == cudaSuccess;
    for (int k = 0; k < 6 && ok; ++k)
      ok = cudaMalloc(&o->cg[k], sizeof(double) * vc) == cudaSuccess && cudaMemsetAsync(o->cg[k], 0, sizeof(double) * vc, c->stream) == cudaSuccess;
    const dim3 gs = o3_shape(o->lev.back().n[0], o->lev.back().n[1], o->lev.back().n[2]);
    size_t nblocks = (size_t)gs.x * gs.y * gs.z;
    if (nblocks < O3_RED_BLOCKS) nblocks = O3_RED_BLOCKS;
    ok = ok && cudaMalloc(&o->partials, sizeof(dd) * nblocks) == cudaSuccess && cudaMalloc(&o->sum, sizeof(dd)) == cudaSuccess &&
         cudaMalloc(&o->iters, sizeof(int)) == cudaSuccess && cudaMallocHost(&o->sum_host, sizeof(dd)) == cudaSuccess &&
         cudaMallocHost(&o->iters_host, sizeof(int)) == cudaSuccess;
    if (ok) ok = cudaMemsetAsync(o->iters, 0, sizeof(int), c->stream) == cudaSuccess;
  }
  if (!ok) {
    cudaGetLastError();
    set_error(EXAMG_ENOMEM, "examg_optflow3_create: device allocation failed");
    examg_optflow3_destroy(o);
    return nullptr;
  }
  return o;
}

int examg_optflow3_destroy(examg_optflow3 *o) {
  if (!o) return EXAMG_OK;
  if (o->c) cudaStreamSynchronize(o->c->stream);
  for (O3Level &L : o->lev) {
    for (int k = 0; k < O3_NF; ++k)
      if (L.f[k]) cudaFree(L.f[k]);
    for (int k = 0; k < 3; ++k)
      if (L.alt[k]) cudaFree(L.alt[k]);
    for (int k = 0; k < 6; ++k)
      if (L.gs[k]) cudaFree(L.gs[k]);
  }
  for (double *p : o->img)
    if (p) cudaFree(p);
  for (double *p : o->cg)
    if (p) cudaFree(p);
  if (o->partials) cudaFree(o->partials);
  if (o->sum) cudaFree(o->sum);
  if (o->iters) cudaFree(o->iters);
  if (o->sum_host) cudaFreeHost(o->sum_host);
  if (o->iters_host) cudaFreeHost(o->iters_host);
  for (auto &kv : o->graphs)
    if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
  delete o;
  return EXAMG_OK;
}

int examg_optflow3_cells(examg_optflow3 *o, int level, int n[3]) {
  O3Level *L = o3_level(o, level);
  if (!L || !n) return L ? set_error(EXAMG_EINVAL, "examg_optflow3_cells: null argument") : EXAMG_EINVAL;
  for (int t = 0; t < 3; ++t) n[t] = L->n[t];
  return EXAMG_OK;
}

int examg_optflow3_upload(examg_optflow3 *o, int level, int which, const double *host) {
  if (!o || !host) return set_error(EXAMG_EINVAL, "examg_optflow3_upload: null argument");
  size_t count = 0;
  double *dev = o3_array(o, level, which, &count);
  if (!dev) return set_error(EXAMG_EINVAL, "examg_optflow3_upload: level %d field %d", level, which);
  if (which >= EXAMG_OF3_G && which < EXAMG_OF3_G + 6) {  // the split-row copies follow at once (cycles may be graphs)
    O3Level *L = o3_level(o, level);
    EXAMG_CUDA(cudaMemcpyAsync(dev, host, sizeof(double) * count, cudaMemcpyHostToDevice, o->c->stream));
    if (L->gs[which - EXAMG_OF3_G]) {
      k3_split_rows<<<o3_grid((long long)L->vol), 256, 0, o->c->stream>>>(dev, L->gs[which - EXAMG_OF3_G], L->n[0] + 2, (long long)L->vol);
      EXAMG_LAUNCH_CHECK(o->c);
    }
    EXAMG_CUDA(cudaStreamSynchronize(o->c->stream));
    return EXAMG_OK;
  }
  EXAMG_CUDA(cudaMemcpyAsync(dev, host, sizeof(double) * count, cudaMemcpyHostToDevice, o->c->stream));
  EXAMG_CUDA(cudaStreamSynchronize(o->c->stream));
  return EXAMG_OK;
}

int examg_optflow3_download(examg_optflow3 *o, int level, int which, double *host) {
  if (!o || !host) return set_error(EXAMG_EINVAL, "examg_optflow3_download: null argument");
  size_t count = 0;
  double *dev = o3_array(o, level, which, &count);
  if (!dev) return set_error(EXAMG_EINVAL, "examg_optflow3_download: level %d field %d", level, which);
  EXAMG_CUDA(cudaMemcpyAsync(host, dev, sizeof(double) * count, cudaMemcpyDeviceToHost, o->c->stream));
  EXAMG_CUDA(cudaStreamSynchronize(o->c->stream));
  return EXAMG_OK;
}

int examg_optflow3_init_fields(examg_optflow3 *o) {
  if (!o) return set_error(EXAMG_EINVAL, "null optflow3 object");
  examg_ctx *c = o->c;
  O3Level &L = o->lev.back();
  const O3Args a = o3_args(o, L);
  const int gall = o3_grid((long long)L.vol);
  for (int k = 0; k < 2; ++k) {  // apply bc to img0, img1
    k3_ghost_fill<<<gall, 256, 0, c->stream>>>(a, o->img[k]);
    EXAMG_LAUNCH_CHECK(c);
  }
  k3_init<<<o3_grid((long long)L.n[0] * L.n[1] * L.n[2]), 256, 0, c->stream>>>(a, o->img[0], o->img[1], 1.0 / L.h[0], -1.0 / L.h[0],
                                                                          1.0 / L.h[1], -1.0 / L.h[1], 1.0 / L.h[2], -1.0 / L.h[2]);
  EXAMG_LAUNCH_CHECK(c);
  static const int g[6] = {9, 10, 11, 12, 13, 14};
  for (int l = (int)o->lev.size() - 2; l >= 0; --l) {  // initGradients@(all but finest)
    const int rc = o3_restrict(o, o->lev[l + 1], o->lev[l], g, g, 6);
    if (rc) return rc;
  }
  for (O3Level &Lv : o->lev) {  // split-row copies for the colour passes
    if (!Lv.gs[0]) continue;
    for (int k = 0; k < 6; ++k) {
      k3_split_rows<<<o3_grid((long long)Lv.vol), 256, 0, c->stream>>>(Lv.f[9 + k], Lv.gs[k], Lv.n[0] + 2, (long long)Lv.vol);
      EXAMG_LAUNCH_CHECK(c);
    }
    Lv.gs_valid = true;
  }
  return EXAMG_OK;
}

int examg_optflow3_smooth_color(examg_optflow3 *o, int level, int color) {
  O3Level *L = o3_level(o, level);
  if (!L) return EXAMG_EINVAL;
  if (color != 0 && color != 1) return set_error(EXAMG_EINVAL, "examg_optflow3_smooth_color: colour %d", color);
  return o3_smooth_color(o, *L, color);
}

int examg_optflow3_smooth(examg_optflow3 *o, int level, int nsweeps) {
  O3Level *L = o3_level(o, level);
  if (!L) return EXAMG_EINVAL;
  if (nsweeps < 0) return set_error(EXAMG_EINVAL, "examg_optflow3_smooth: %d sweeps", nsweeps);
  for (int it = 0; it < nsweeps; ++it) {
    const int rc = o3_sweep(o, *L);
    if (rc) return rc;
  }
  return EXAMG_OK;
}

int examg_optflow3_residual(examg_optflow3 *o, int level, double *norm_out) {
  O3Level *L = o3_level(o, level);
  if (!L) return EXAMG_EINVAL;
  examg_ctx *c = o->c;
  int rc = o3_residual(o, *L, norm_out != nullptr);
  if (rc || !norm_out) return rc;
  EXAMG_CUDA(cudaMemcpyAsync(o->sum_host, o->sum, sizeof(dd), cudaMemcpyDeviceToHost, c->stream));
  EXAMG_SYNC(c);
  *norm_out = std::sqrt(o->sum_host->hi + o->sum_host->lo);
  return EXAMG_OK;
}

int examg_optflow3_restrict(examg_optflow3 *o, int level) {
  O3Level *L = o3_level(o, level);
  if (!L) return EXAMG_EINVAL;
  if (level < 0) level = o->c->cfg.max_level;
  if (level == o->c->cfg.min_level) return set_error(EXAMG_EINVAL, "examg_optflow3_restrict: no level below %d", level);
  return o3_restrict_residual(o, level);
}

int examg_optflow3_correct(examg_optflow3 *o, int level) {
  O3Level *L = o3_level(o, level);
  if (!L) return EXAMG_EINVAL;
  if (level < 0) level = o->c->cfg.max_level;
  if (level == o->c->cfg.min_level) return set_error(EXAMG_EINVAL, "examg_optflow3_correct: no level below %d", level);
  return o3_correct(o, level);
}

int examg_optflow3_cg(examg_optflow3 *o, int *iters_out) {
  if (!o) return set_error(EXAMG_EINVAL, "null optflow3 object");
  const int rc = o3_cg(o);
  if (rc || !iters_out) return rc;
  return examg_optflow3_last_cg_iters(o, iters_out);
}

int examg_optflow3_last_cg_iters(examg_optflow3 *o, int *iters_out) {
  if (!o || !iters_out) return set_error(EXAMG_EINVAL, "examg_optflow3_last_cg_iters: null argument");
  examg_ctx *c = o->c;
  EXAMG_CUDA(cudaMemcpyAsync(o->iters_host, o->iters, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  EXAMG_SYNC(c);
  *iters_out = *o->iters_host;
  return EXAMG_OK;
}

int examg_optflow3_vcycle(examg_optflow3 *o, int level) {
  O3Level *L = o3_level(o, level);
  if (!L) return EXAMG_EINVAL;
  examg_ctx *c = o->c;
  if (level < 0) level = c->cfg.max_level;
  static const bool use_graph = !(getenv("EXAMG_NO_GRAPH") && getenv("EXAMG_NO_GRAPH")[0] == '1');
  if (!use_graph || level != c->cfg.max_level || level == c->cfg.min_level) return o3_vcycle_rec(o, level);
  const std::vector<char> key = o3_state(o);
  auto it = o->graphs.find(key);
  if (it == o->graphs.end()) {
    const long long l0 = c->launches;
    cudaGraph_t graph = nullptr;
    EXAMG_CUDA(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    const int rc = o3_vcycle_rec(o, level);
    const cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
    examg_optflow3::Graph g;
    g.launches = c->launches - l0;
    c->launches = l0;  // nothing ran yet
    g.exit_state = o3_state(o);
    o3_restore(o, key);
    if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
    if (e != cudaSuccess) return set_error(EXAMG_ECUDA, "graph capture of mgCycle failed: %s", cudaGetErrorString(e));
    const cudaError_t e2 = cudaGraphInstantiate(&g.exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e2 != cudaSuccess) return set_error(EXAMG_ECUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(e2));
    it = o->graphs.emplace(key, g).first;
  }
  EXAMG_CUDA(cudaGraphLaunch(it->second.exec, c->stream));
  c->launches += it->second.launches;
  o3_restore(o, it->second.exit_state);
  return EXAMG_OK;
}

int examg_optflow3_cycles(examg_optflow3 *o, int n, double *norm_out) {
  for (int i = 0; i < n; ++i) {
    const int rc = examg_optflow3_vcycle(o, -1);
    if (rc) return rc;
  }
  return norm_out ? examg_optflow3_residual(o, -1, norm_out) : EXAMG_OK;
}

}  // extern "C"
