// examg_optflow.cu -- the coupled two-unknown system of the optical-flow programs on 2D cell-centred fields:
// red-black 2x2 block smoother (`solve locally`), residual (+ fused norm), cell restriction / prolongation,
// Neumann ghost cells, coupled CG on the coarsest level, and the whole cycle as one object.
//
// Reference loops being replaced (Benchmark/OptFlow2D/2D_FD_OptFlow.exa4; Vec2 form:
// Testing/Application/OpticalFlow2D.exa4):
//   smoother     :144-161 / :216-231   repeat with { colour 0, colour 1; communicate u, v; loop over u { solve
//                locally relax r { u => uuSten*u + uvSten*v + alpha*alpha*Laplace*u == rhs_u; v => ... } }; apply bc }
//                C/s/e/solver/ir/IR_LocalSolve.scala:88-236, IR_LocalDirectInvert.scala:44-140
//   residual     :167-177   residual_u = rhs_u - (uuSten*u + uvSten*v + alpha**2 * (Laplace*u)); apply bc
//   restriction  :180-188   rhs@coarser = RestrictionCell * residual   (L4_DefaultRestriction.scala:40-47: [1/2,1/2]^d)
//   correction   :204-214   u += ProlongationCell@coarser * u@coarser  (L4_DefaultProlongation.scala:30-45: piecewise
//                constant); apply bc
//   coarse CG    :234-335
//   Neumann      C/s/e/boundary/ir/IR_NeumannBC.scala:70-80 (ghost cell = adjacent inner cell), faces in the order
//                -x,+x,-y,+y over the full orthogonal extent (IR_ApplyBCFunction.scala:53-85) => ghost (i,j) = cell
//                (clamp i, clamp j), corners included
//
// Design (B200): every kernel that writes a Neumann field also writes the ghost cells next to the cells it
// updates, so `apply bc` costs no launch and no pass; the two unknowns are processed together (the reference
// makes one loop per unknown for residual/restriction/correction); the cycle is one CUDA graph.
// Arithmetic: oracle/examg_oracle_optflow.c header ("Arithmetic contract"); fields are bit-identical.
// Layout: cell (i,j) of a cell field sits where node (i,j) of a node field of the level does, i.e.
// n-1 cells use the n-node allocation; ghost cells are elements -1 and n_cells.
#include <cmath>
#include <map>
#include <type_traits>
#include <utility>
#include <vector>

#include "examg_cg.cuh"

using namespace examg;

namespace examg {

// fn(integral_constant<K>) for K = 0, 1, ... until it returns false
template <int... Ks, class F>
__device__ __forceinline__ bool of_steps(std::integer_sequence<int, Ks...>, F &&fn) {
  return (fn(std::integral_constant<int, Ks>{}) && ...);
}

struct OfArgs {
  int nx, ny, pitch;
  int par;  // parity of the block's first cell in the global grid: the colour of cell (i, j) is (i + j + par) & 1
  // k_of_rbw on several blocks: the cells [xlo, xhi) x [ylo, yhi) it works on -- the block's own cells plus S layers of
  // the neighbour blocks' cells where there is a neighbour (exchanged beforehand; updated redundantly, one layer less per
  // half-sweep) -- and which sides are physical (Neumann: clamped neighbour access, ghost cells mirror the result)
  int xlo, xhi, ylo, yhi;
  int phys;  // bit 0 west, 1 east, 2 south, 3 north
  double L[5];
  double a2;
  const double *cuu, *cuv, *cvv;
};

// store a cell of a Neumann field together with the ghost cells that mirror it
__device__ __forceinline__ void store_neumann(double *a, int i, int j, int nx, int ny, int pitch, double val) {
  const long long p = i + (long long)pitch * j;
  a[p] = val;
  const int gi = i == 0 ? -1 : (i == nx - 1 ? nx : i), gj = j == 0 ? -1 : (j == ny - 1 ? ny : j);
  const bool ex = i == 0 || i == nx - 1, ey = j == 0 || j == ny - 1;
  if (ex) a[gi + (long long)pitch * j] = val;
  if (ey) a[i + (long long)pitch * gj] = val;
  if (ex && ey) a[gi + (long long)pitch * gj] = val;
}

__device__ __forceinline__ double of_lap(const double (&L)[5], const double *w, long long p, int pitch) {
  double acc = L[0] * w[p];
  acc = fma(L[1], w[p - 1], acc);
  acc = fma(L[2], w[p + 1], acc);
  acc = fma(L[3], w[p - pitch], acc);
  acc = fma(L[4], w[p + pitch], acc);
  return acc;
}

__device__ __forceinline__ double of_off(const double (&L)[5], const double *w, long long p, int pitch) {
  double acc = L[1] * w[p - 1];
  acc = fma(L[2], w[p + 1], acc);
  acc = fma(L[3], w[p - pitch], acc);
  acc = fma(L[4], w[p + pitch], acc);
  return acc;
}

// the 2x2 local solve at one cell
__device__ __forceinline__ void of_local_solve(const OfArgs &a, const double *u, const double *v, const double *fu,
                                               const double *fv, long long p, double relax, double keep, double &nu,
                                               double &nv) {
  const double f0 = fu[p] - a.a2 * of_off(a.L, u, p, a.pitch);
  const double f1 = fv[p] - a.a2 * of_off(a.L, v, p, a.pitch);
  const double a01 = a.cuv[p];
  const double a00 = fma(a.a2, a.L[0], a.cuu[p]), a11 = fma(a.a2, a.L[0], a.cvv[p]);
  const double rdet = 1.0 / (a00 * a11 - a01 * a01);
  const double i00 = rdet * a11, i01 = -(rdet * a01), i11 = rdet * a00;
  const double su = i00 * f0 + i01 * f1;
  const double sv = i01 * f0 + i11 * f1;
  nu = u[p] * keep + relax * su;
  nv = v[p] * keep + relax * sv;
}

// ghost (i,j) = cell (clamp i, clamp j): the result of the four face loops in the reference's order
__global__ void __launch_bounds__(256) k_of_neumann(double *a, int nx, int ny, int pitch) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int per = 2 * (nx + 2) + 2 * ny;
  if (t >= per) return;
  int i, j;
  if (t < nx + 2) { i = t - 1; j = -1; }
  else if (t < 2 * (nx + 2)) { i = t - (nx + 2) - 1; j = ny; }
  else if (t < 2 * (nx + 2) + ny) { i = -1; j = t - 2 * (nx + 2); }
  else { i = nx; j = t - 2 * (nx + 2) - ny; }
  const int ci = min(max(i, 0), nx - 1), cj = min(max(j, 0), ny - 1);
  a[i + (long long)pitch * j] = a[ci + (long long)pitch * cj];
}

// one colour of the block smoother; thread per updated cell
__global__ void __launch_bounds__(256)
k_of_smooth_color(const OfArgs a, double *__restrict__ u, double *__restrict__ v, const double *__restrict__ fu,
                  const double *__restrict__ fv, double relax, int color) {
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  if (j >= a.ny) return;
  const int i = 2 * (blockIdx.x * blockDim.x + threadIdx.x) + ((j + color + a.par) & 1);
  if (i >= a.nx) return;
  const long long p = i + (long long)a.pitch * j;
  double nu, nv;
  of_local_solve(a, u, v, fu, fv, p, relax, 1.0 - relax, nu, nv);
  store_neumann(u, i, j, a.nx, a.ny, a.pitch, nu);
  store_neumann(v, i, j, a.nx, a.ny, a.pitch, nv);
}

// the same on the cells [xlo, xhi) x [ylo, yhi) of a block among several: its own cells plus layers of the neighbours'
// cells (exchanged beforehand, updated redundantly); Neumann ghost cells on physical sides only
__global__ void __launch_bounds__(256)
k_of_smooth_color_mb(const OfArgs a, double *__restrict__ u, double *__restrict__ v, const double *__restrict__ fu,
                     const double *__restrict__ fv, double relax, int color) {
  const int j = a.ylo + blockIdx.y * blockDim.y + threadIdx.y;
  if (j >= a.yhi) return;
  const int i = a.xlo + ((a.xlo + j + color + a.par) & 1) + 2 * (blockIdx.x * blockDim.x + threadIdx.x);
  if (i >= a.xhi) return;
  const long long p = i + (long long)a.pitch * j;
  // neighbours across a physical side: the cell itself (what its Neumann ghost cell holds; the ghost cells next to the
  // neighbours' layers are not part of the exchange)
  const bool cw = i == 0 && (a.phys & 1), ce = i == a.nx - 1 && (a.phys & 2);
  const bool cs = j == 0 && (a.phys & 4), cn = j == a.ny - 1 && (a.phys & 8);
  const long long pw = cw ? p : p - 1, pe = ce ? p : p + 1, ps = cs ? p : p - a.pitch, pn = cn ? p : p + a.pitch;
  double ou = a.L[1] * u[pw], ov = a.L[1] * v[pw];
  ou = fma(a.L[2], u[pe], ou); ov = fma(a.L[2], v[pe], ov);
  ou = fma(a.L[3], u[ps], ou); ov = fma(a.L[3], v[ps], ov);
  ou = fma(a.L[4], u[pn], ou); ov = fma(a.L[4], v[pn], ov);
  const double f0 = fu[p] - a.a2 * ou, f1 = fv[p] - a.a2 * ov;
  const double a01 = a.cuv[p];
  const double a00 = fma(a.a2, a.L[0], a.cuu[p]), a11 = fma(a.a2, a.L[0], a.cvv[p]);
  const double rdet = 1.0 / (a00 * a11 - a01 * a01);
  const double i00 = rdet * a11, i01 = -(rdet * a01), i11 = rdet * a00;
  const double su = i00 * f0 + i01 * f1, sv = i01 * f0 + i11 * f1;
  const double keep = 1.0 - relax;
  const double nu = u[p] * keep + relax * su, nv = v[p] * keep + relax * sv;
  u[p] = nu;
  v[p] = nv;
  const int gi = (i == 0 && (a.phys & 1)) ? -1 : ((i == a.nx - 1 && (a.phys & 2)) ? a.nx : i);
  const int gj = (j == 0 && (a.phys & 4)) ? -1 : ((j == a.ny - 1 && (a.phys & 8)) ? a.ny : j);
  if (gi != i) { u[gi + (long long)a.pitch * j] = nu; v[gi + (long long)a.pitch * j] = nv; }
  if (gj != j) { u[i + (long long)a.pitch * gj] = nu; v[i + (long long)a.pitch * gj] = nv; }
  if (gi != i && gj != j) { u[gi + (long long)a.pitch * gj] = nu; v[gi + (long long)a.pitch * gj] = nv; }
}

// residual of both unknowns (+ sum of squares); thread per cell, rows by grid stride
template <bool NORM>
__global__ void __launch_bounds__(256)
k_of_residual(const OfArgs a, const double *__restrict__ u, const double *__restrict__ v, const double *__restrict__ fu,
              const double *__restrict__ fv, double *__restrict__ ru, double *__restrict__ rv, dd *partials) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  dd part = {0.0, 0.0};
  for (int j = blockIdx.y * blockDim.y + threadIdx.y; j < a.ny; j += gridDim.y * blockDim.y) {
    if (i >= a.nx) break;
    const long long p = i + (long long)a.pitch * j;
    const double r0 = fu[p] - fma(a.a2, of_lap(a.L, u, p, a.pitch), fma(a.cuv[p], v[p], a.cuu[p] * u[p]));
    const double r1 = fv[p] - fma(a.a2, of_lap(a.L, v, p, a.pitch), fma(a.cvv[p], v[p], a.cuv[p] * u[p]));
    store_neumann(ru, i, j, a.nx, a.ny, a.pitch, r0);
    store_neumann(rv, i, j, a.nx, a.ny, a.pitch, r1);
    if (NORM) {
      dd_add(part, r0 * r0);
      dd_add(part, r1 * r1);
    }
  }
  if (NORM) {
    part = dd_block_reduce(part);
    if (threadIdx.x == 0 && threadIdx.y == 0) partials[blockIdx.x + gridDim.x * blockIdx.y] = part;
  }
}

__global__ void __launch_bounds__(1024) k_of_final_sum(const dd *__restrict__ partials, int n, dd *out) {
  dd v = {0.0, 0.0};
  for (int i = threadIdx.x; i < n; i += blockDim.x) dd_merge(v, partials[i]);
  v = dd_block_reduce(v);
  if (threadIdx.x == 0) *out = v;
}

struct Pair {
  const double *src[3];
  double *dst[3];
};

// coarse = 0.25 * (((f00 + f10) + f01) + f11) for up to three fields (blockIdx.z)
__global__ void __launch_bounds__(256) k_of_restrict(const Pair pr, int nxc, int nyc, int pitch_f, int pitch_c) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const double *__restrict__ f = pr.src[blockIdx.z];
  double *__restrict__ cdst = pr.dst[blockIdx.z];
  for (int j = blockIdx.y * blockDim.y + threadIdx.y; j < nyc; j += gridDim.y * blockDim.y) {
    if (i >= nxc) break;
    const long long q = 2 * i + (long long)pitch_f * (2 * j);
    const double2 lo = *reinterpret_cast<const double2 *>(f + q);
    const double2 hi = *reinterpret_cast<const double2 *>(f + q + pitch_f);
    cdst[i + (long long)pitch_c * j] = 0.25 * (((lo.x + lo.y) + hi.x) + hi.y);
  }
}

// fine += coarse(parent) for up to three fields; thread per coarse cell; Neumann ghosts refreshed
__global__ void __launch_bounds__(256) k_of_prolong_add(const Pair pr, int nxc, int nyc, int pitch_f, int pitch_c) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const double *__restrict__ cs = pr.src[blockIdx.z];
  double *__restrict__ f = pr.dst[blockIdx.z];
  const int nxf = 2 * nxc, nyf = 2 * nyc;
  for (int j = blockIdx.y * blockDim.y + threadIdx.y; j < nyc; j += gridDim.y * blockDim.y) {
    if (i >= nxc) break;
    const double cv = cs[i + (long long)pitch_c * j];
    const long long q = 2 * i + (long long)pitch_f * (2 * j);
    double2 lo = *reinterpret_cast<double2 *>(f + q);
    double2 hi = *reinterpret_cast<double2 *>(f + q + pitch_f);
    lo.x += cv; lo.y += cv; hi.x += cv; hi.y += cv;
    const bool edge = i == 0 || i == nxc - 1 || j == 0 || j == nyc - 1;
    if (!edge) {
      *reinterpret_cast<double2 *>(f + q) = lo;
      *reinterpret_cast<double2 *>(f + q + pitch_f) = hi;
    } else {
      store_neumann(f, 2 * i, 2 * j, nxf, nyf, pitch_f, lo.x);
      store_neumann(f, 2 * i + 1, 2 * j, nxf, nyf, pitch_f, lo.y);
      store_neumann(f, 2 * i, 2 * j + 1, nxf, nyf, pitch_f, hi.x);
      store_neumann(f, 2 * i + 1, 2 * j + 1, nxf, nyf, pitch_f, hi.y);
    }
  }
}

// 3D cell transfer (Testing/CellBased/3D_*.exa4): coarse = 0.125 * (sum of the 8 children in lexicographic z,y,x order);
// fine += parent.  Thread per coarse cell, x fastest; ghost cells are the business of `apply bc`.
__global__ void __launch_bounds__(256)
k_cell_restrict3(const Pair pr, int nxc, int nyc, int nzc, int pitch_f, long long plane_f, int pitch_c, long long plane_c) {
  const double *__restrict__ f = pr.src[blockIdx.z];
  double *__restrict__ cdst = pr.dst[blockIdx.z];
  const long long total = (long long)nxc * nyc * nzc;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(t % nxc), j = (int)((t / nxc) % nyc), k = (int)(t / ((long long)nxc * nyc));
    const double *q = f + 2 * i + (long long)pitch_f * (2 * j) + plane_f * (2 * k);
    const double2 a = *reinterpret_cast<const double2 *>(q), b = *reinterpret_cast<const double2 *>(q + pitch_f);
    const double2 cc = *reinterpret_cast<const double2 *>(q + plane_f), dd2 = *reinterpret_cast<const double2 *>(q + plane_f + pitch_f);
    cdst[i + (long long)pitch_c * j + plane_c * k] =
        0.125 * (((((((a.x + a.y) + b.x) + b.y) + cc.x) + cc.y) + dd2.x) + dd2.y);
  }
}

__global__ void __launch_bounds__(256)
k_cell_prolong3(const Pair pr, int nxc, int nyc, int nzc, int pitch_f, long long plane_f, int pitch_c, long long plane_c) {
  const double *__restrict__ cs = pr.src[blockIdx.z];
  double *__restrict__ f = pr.dst[blockIdx.z];
  const long long total = (long long)nxc * nyc * nzc;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(t % nxc), j = (int)((t / nxc) % nyc), k = (int)(t / ((long long)nxc * nyc));
    const double cv = cs[i + (long long)pitch_c * j + plane_c * k];
    double *q = f + 2 * i + (long long)pitch_f * (2 * j) + plane_f * (2 * k);
#pragma unroll
    for (int dz = 0; dz < 2; ++dz)
#pragma unroll
      for (int dy = 0; dy < 2; ++dy) {
        double2 *p2 = reinterpret_cast<double2 *>(q + plane_f * dz + (long long)pitch_f * dy);
        double2 v = *p2;
        v.x += cv;
        v.y += cv;
        *p2 = v;
      }
  }
}

// InitFields: Ix, Iy, It -> IxIx, IxIy, IyIy and rhs (images carry valid Neumann ghosts)
__global__ void __launch_bounds__(256)
k_of_init(const double *__restrict__ img0, const double *__restrict__ img1, int nx, int ny, int pitch, double gx1,
          double gx0, double gy1, double gy0, double *__restrict__ cuu, double *__restrict__ cuv,
          double *__restrict__ cvv, double *__restrict__ fu, double *__restrict__ fv) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  for (int j = blockIdx.y * blockDim.y + threadIdx.y; j < ny; j += gridDim.y * blockDim.y) {
    if (i >= nx) break;
    const long long p = i + (long long)pitch * j;
    const double ix = fma(gx0, img0[p], gx1 * img0[p + 1]);  // GradX: [1,0] then [0,0]
    const double iy = fma(gy0, img0[p], gy1 * img0[p + pitch]);
    const double it = img1[p] - img0[p];
    cuu[p] = ix * ix;
    cuv[p] = ix * iy;
    cvv[p] = iy * iy;
    fu[p] = -ix * it;
    fv[p] = -iy * it;
  }
}


// ---- fused red-black block smoother: S half-sweeps (S/2 sweeps) per pass over HBM ---------------------
// "Warp-private wavefront": a warp owns a strip of 64 cells in x (S-deep halo each side included) and
// marches in y; a lane owns two adjacent cells (one of each colour) and keeps S+2 rows of u, v and S+1
// rows of the five read-only arrays (rhs_u, rhs_v, IxIx, IxIy, IyIy) in registers.  Half-sweep s works on
// row t-s, one row behind half-sweep s-1: the y neighbours are the lane's own registers, the x neighbours
// its other cell and one warp shuffle.  No shared memory, no block-wide barrier; 128-bit coalesced loads
// (512 B per row, array and warp) issued one row ahead.  Out of place (u, v -> second buffers, exchanged
// afterwards), because neighbouring strips read each other's halo.  Neumann boundaries = clamped
// neighbour access; the ghost cells of the result are written with the cells they mirror.
// HBM traffic per cell: 7 reads + 2 writes = 72 B per pass for S/2 sweeps of 80 B algorithmic each
// (halo columns come from L2: adjacent strips run in the same CTA).
template <int S, bool MB>
__global__ void __launch_bounds__(128)
k_of_rbw(const OfArgs a, const double *__restrict__ u, const double *__restrict__ v, const double *__restrict__ fu,
         const double *__restrict__ fv, double *__restrict__ uo, double *__restrict__ vo, double relax, int seg_rows,
         int nstrips) {
  constexpr int NW = S + 2, IW = 64 - 2 * S;
  const int lane = threadIdx.x & 31;
  const int strip = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (strip >= nstrips) return;  // whole warp
  const int nx = a.nx, ny = a.ny, pitch = a.pitch;
  const int ox = IW * strip, x0 = ox - S + 2 * lane;  // x0 even: cell A = (x0, .), cell B = (x0+1, .)
  const int za = blockIdx.y * seg_rows, zb = min(ny, za + seg_rows);
  int t0 = za - S;
  t0 -= t0 & 1;  // even: at odd steps every half-sweep updates cell A of its row, at even steps cell B
  const int tl = zb - 1 + S;
  // MB = false (one block): the block's own cells, all four sides physical -- compile-time constants
  const int xlo = MB ? a.xlo : 0, xhi = MB ? a.xhi : nx, ylo = MB ? a.ylo : 0, yhi = MB ? a.yhi : ny;
  const int phys = MB ? a.phys : 15;
  const bool inx = x0 >= xlo && x0 < xhi;
  const bool west_edge = x0 == 0 && (phys & 1), east_edge = x0 + 1 == nx - 1 && (phys & 2);
  const bool phys_s = (phys & 4) != 0, phys_n = (phys & 8) != 0;
  const bool st = inx && x0 >= ox && x0 < ox + IW && x0 < nx;
  const double keep = 1.0 - relax, a2 = a.a2;
  const double L0 = a.L[0], L1 = a.L[1], L2 = a.L[2], L3 = a.L[3], L4 = a.L[4];

  double uw[NW][2], vw[NW][2], cf[NW][5][2];  // cf: rhs_u, rhs_v and the inverse of the cell's 2x2 block: I00, I01, I11
#pragma unroll
  for (int q = 0; q < NW; ++q) {
    uw[q][0] = uw[q][1] = vw[q][0] = vw[q][1] = 0.0;
#pragma unroll
    for (int k = 0; k < 5; ++k) cf[q][k][0] = cf[q][k][1] = 0.0;
  }
  double2 nx_u, nx_v, nx_c[5];  // row t+1, in flight
  auto fetch = [&](int r) {
    if (inx && r >= ylo && r < yhi && r <= tl) {
      const long long p = x0 + (long long)pitch * r;
      nx_u = *reinterpret_cast<const double2 *>(u + p);
      nx_v = *reinterpret_cast<const double2 *>(v + p);
      nx_c[0] = __ldg(reinterpret_cast<const double2 *>(fu + p));
      nx_c[1] = __ldg(reinterpret_cast<const double2 *>(fv + p));
      nx_c[2] = __ldg(reinterpret_cast<const double2 *>(a.cuu + p));
      nx_c[3] = __ldg(reinterpret_cast<const double2 *>(a.cuv + p));
      nx_c[4] = __ldg(reinterpret_cast<const double2 *>(a.cvv + p));
    } else {
      nx_u = nx_v = make_double2(0.0, 0.0);
#pragma unroll
      for (int k = 0; k < 5; ++k) nx_c[k] = make_double2(0.0, 0.0);
    }
  };
  fetch(t0);

  auto step = [&](auto kc, const int t) {
    constexpr int K = decltype(kc)::value;
    constexpr int X = (K & 1) ? 0 : 1;  // the cell every half-sweep updates at this step
    uw[K][0] = nx_u.x; uw[K][1] = nx_u.y; vw[K][0] = nx_v.x; vw[K][1] = nx_v.y;
    cf[K][0][0] = nx_c[0].x; cf[K][0][1] = nx_c[0].y; cf[K][1][0] = nx_c[1].x; cf[K][1][1] = nx_c[1].y;
    {  // the block inverse depends on the coefficients only: once per cell and pass, off the critical path
      const double a00x = fma(a2, L0, nx_c[2].x), a11x = fma(a2, L0, nx_c[4].x), a01x = nx_c[3].x;
      const double a00y = fma(a2, L0, nx_c[2].y), a11y = fma(a2, L0, nx_c[4].y), a01y = nx_c[3].y;
      const double rx = 1.0 / (a00x * a11x - a01x * a01x), ry = 1.0 / (a00y * a11y - a01y * a01y);
      cf[K][2][0] = rx * a11x; cf[K][3][0] = -(rx * a01x); cf[K][4][0] = rx * a00x;
      cf[K][2][1] = ry * a11y; cf[K][3][1] = -(ry * a01y); cf[K][4][1] = ry * a00y;
    }
    fetch(t + 1);
#pragma unroll
    for (int s = 1; s <= S; ++s) {
      const int c = (K + NW - s) % NW, cm = (K + NW - s - 1) % NW, cp = (K + NW - s + 1) % NW;
      const int r = t - s;
      // x neighbour owned by the adjacent lane: A's west = B of lane-1, B's east = A of lane+1
      const double su = X == 0 ? __shfl_up_sync(0xffffffffu, uw[c][1], 1) : __shfl_down_sync(0xffffffffu, uw[c][0], 1);
      const double sv = X == 0 ? __shfl_up_sync(0xffffffffu, vw[c][1], 1) : __shfl_down_sync(0xffffffffu, vw[c][0], 1);
      if (inx && r >= ylo && r < yhi) {
        const double cu = uw[c][X], cv = vw[c][X];
        double uwst, uest, vwst, vest;
        if (X == 0) {
          uwst = west_edge ? cu : su; vwst = west_edge ? cv : sv;
          uest = uw[c][1]; vest = vw[c][1];
        } else {
          uwst = uw[c][0]; vwst = vw[c][0];
          uest = east_edge ? cu : su; vest = east_edge ? cv : sv;
        }
        const bool cs = r == 0 && phys_s, cn = r == ny - 1 && phys_n;
        const double usth = cs ? cu : uw[cm][X], vsth = cs ? cv : vw[cm][X];
        const double unth = cn ? cu : uw[cp][X], vnth = cn ? cv : vw[cp][X];
        double o0 = L1 * uwst;
        o0 = fma(L2, uest, o0);
        o0 = fma(L3, usth, o0);
        o0 = fma(L4, unth, o0);
        double o1 = L1 * vwst;
        o1 = fma(L2, vest, o1);
        o1 = fma(L3, vsth, o1);
        o1 = fma(L4, vnth, o1);
        const double f0 = cf[c][0][X] - a2 * o0, f1 = cf[c][1][X] - a2 * o1;
        const double nu = cf[c][2][X] * f0 + cf[c][3][X] * f1;
        const double nv = cf[c][3][X] * f0 + cf[c][4][X] * f1;
        uw[c][X] = cu * keep + relax * nu;
        vw[c][X] = cv * keep + relax * nv;
      }
    }
    const int zf = t - S;
    if (st && zf >= za && zf < zb) {
      const int c = (K + NW - S) % NW;
      const double2 ru = make_double2(uw[c][0], uw[c][1]), rv = make_double2(vw[c][0], vw[c][1]);
      const long long p = x0 + (long long)pitch * zf;
      *reinterpret_cast<double2 *>(uo + p) = ru;
      *reinterpret_cast<double2 *>(vo + p) = rv;
      // Neumann ghost cells mirror the cells just written
      if (west_edge) { uo[p - 1] = ru.x; vo[p - 1] = rv.x; }
      if (east_edge) { uo[p + 2] = ru.y; vo[p + 2] = rv.y; }
      if ((zf == 0 && phys_s) || (zf == ny - 1 && phys_n)) {
        const long long g = p + (zf == 0 ? -(long long)pitch : (long long)pitch);
        *reinterpret_cast<double2 *>(uo + g) = ru;
        *reinterpret_cast<double2 *>(vo + g) = rv;
        if (west_edge) { uo[g - 1] = ru.x; vo[g - 1] = rv.x; }
        if (east_edge) { uo[g + 2] = ru.y; vo[g + 2] = rv.y; }
      }
    }
  };

  for (int t = t0; t <= tl; t += NW) {
    const bool more = of_steps(std::make_integer_sequence<int, NW>{}, [&](auto kc) {
      constexpr int K = decltype(kc)::value;
      if (t + K > tl) return false;
      step(kc, t + K);
      return true;
    });
    if (!more) break;
  }
}

// mgCycle@coarsest: the coupled CG, whole solve in one CTA (the grid has a few dozen cells)
__global__ void __launch_bounds__(1024)
k_of_cg(const OfArgs a, double *u, double *v, const double *fu, const double *fv, double *ru, double *rv, double *pu,
        double *pv, double *apu, double *apv, int max_it, double rel_tol, int *iters_out) {
  const int total = a.nx * a.ny, tid = threadIdx.x, nt = blockDim.x;
  dd part = {0.0, 0.0};
  for (int t = tid; t < total; t += nt) {
    const int i = t % a.nx, j = t / a.nx;
    const long long p = i + (long long)a.pitch * j;
    const double r0 = fu[p] - fma(a.a2, of_lap(a.L, u, p, a.pitch), fma(a.cuv[p], v[p], a.cuu[p] * u[p]));
    const double r1 = fv[p] - fma(a.a2, of_lap(a.L, v, p, a.pitch), fma(a.cvv[p], v[p], a.cuv[p] * u[p]));
    store_neumann(ru, i, j, a.nx, a.ny, a.pitch, r0);
    store_neumann(rv, i, j, a.nx, a.ny, a.pitch, r1);
    store_neumann(pu, i, j, a.nx, a.ny, a.pitch, r0);
    store_neumann(pv, i, j, a.nx, a.ny, a.pitch, r1);
    dd_add(part, r0 * r0);
    dd_add(part, r1 * r1);
  }
  double curRes = sqrt(cg_sum(part));
  const double initRes = curRes;
  int iters = 0;
  for (int step = 0; step < max_it; ++step) {
    iters = step + 1;
    dd n0 = {0.0, 0.0}, n1 = {0.0, 0.0}, d0 = {0.0, 0.0}, d1 = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const int i = t % a.nx, j = t / a.nx;
      const long long p = i + (long long)a.pitch * j;
      const double au = fma(a.a2, of_lap(a.L, pu, p, a.pitch), fma(a.cuv[p], pv[p], a.cuu[p] * pu[p]));
      const double av = fma(a.a2, of_lap(a.L, pv, p, a.pitch), fma(a.cvv[p], pv[p], a.cuv[p] * pu[p]));
      apu[p] = au;
      apv[p] = av;
      dd_add(n0, ru[p] * ru[p]);
      dd_add(n1, rv[p] * rv[p]);
      dd_add(d0, pu[p] * au);
      dd_add(d1, pv[p] * av);
    }
    const double nom = cg_sum(n0) + cg_sum(n1);
    const double den = cg_sum(d0) + cg_sum(d1);
    const double alpha = nom / den;
    dd pr = {0.0, 0.0};
    for (int t = tid; t < total; t += nt) {
      const int i = t % a.nx, j = t / a.nx;
      const long long p = i + (long long)a.pitch * j;
      store_neumann(u, i, j, a.nx, a.ny, a.pitch, fma(alpha, pu[p], u[p]));
      store_neumann(v, i, j, a.nx, a.ny, a.pitch, fma(alpha, pv[p], v[p]));
      const double r0 = fma(-alpha, apu[p], ru[p]), r1 = fma(-alpha, apv[p], rv[p]);
      store_neumann(ru, i, j, a.nx, a.ny, a.pitch, r0);
      store_neumann(rv, i, j, a.nx, a.ny, a.pitch, r1);
      dd_add(pr, r0 * r0);
      dd_add(pr, r1 * r1);
    }
    const double nextRes = sqrt(cg_sum(pr));
    if (nextRes <= rel_tol * initRes) break;
    const double beta = (nextRes * nextRes) / (curRes * curRes);
    for (int t = tid; t < total; t += nt) {
      const int i = t % a.nx, j = t / a.nx;
      const long long p = i + (long long)a.pitch * j;
      store_neumann(pu, i, j, a.nx, a.ny, a.pitch, fma(beta, pu[p], ru[p]));
      store_neumann(pv, i, j, a.nx, a.ny, a.pitch, fma(beta, pv[p], rv[p]));
    }
    curRes = nextRes;
    __syncthreads();
  }
  if (tid == 0 && iters_out) *iters_out = iters;
}

// ---- the same CG on several blocks: one kernel per loop of k_of_cg, global sums in between ---------------------------
// (the scalars are formed from the same sums -- sum r_u^2 and sum r_v^2 rounded separately, then added, etc. -- in
// double-double, so they do not depend on how the cells are spread over blocks)

// cgTmp1 = A * cgTmp0; partial sums of r_u^2, r_v^2, p_u (A p)_u, p_v (A p)_v -> partials[k * nblocks + block]
__global__ void __launch_bounds__(256)
k_of_cg_apply(const OfArgs a, const double *__restrict__ ru, const double *__restrict__ rv, const double *__restrict__ pu,
              const double *__restrict__ pv, double *__restrict__ apu, double *__restrict__ apv, dd *partials) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  dd n0 = {0.0, 0.0}, n1 = {0.0, 0.0}, d0 = {0.0, 0.0}, d1 = {0.0, 0.0};
  for (int j = blockIdx.y * blockDim.y + threadIdx.y; j < a.ny; j += gridDim.y * blockDim.y) {
    if (i >= a.nx) break;
    const long long p = i + (long long)a.pitch * j;
    const double au = fma(a.a2, of_lap(a.L, pu, p, a.pitch), fma(a.cuv[p], pv[p], a.cuu[p] * pu[p]));
    const double av = fma(a.a2, of_lap(a.L, pv, p, a.pitch), fma(a.cvv[p], pv[p], a.cuv[p] * pu[p]));
    apu[p] = au;
    apv[p] = av;
    dd_add(n0, ru[p] * ru[p]);
    dd_add(n1, rv[p] * rv[p]);
    dd_add(d0, pu[p] * au);
    dd_add(d1, pv[p] * av);
  }
  const int nb = gridDim.x * gridDim.y, b = blockIdx.x + gridDim.x * blockIdx.y;
  dd v = dd_block_reduce(n0);
  if (threadIdx.x == 0 && threadIdx.y == 0) partials[b] = v;
  v = dd_block_reduce(n1);
  if (threadIdx.x == 0 && threadIdx.y == 0) partials[nb + b] = v;
  v = dd_block_reduce(d0);
  if (threadIdx.x == 0 && threadIdx.y == 0) partials[2 * nb + b] = v;
  v = dd_block_reduce(d1);
  if (threadIdx.x == 0 && threadIdx.y == 0) partials[3 * nb + b] = v;
}

// Solution += alpha * cgTmp0; Residual -= alpha * cgTmp1 (ghost cells with them); partial sums of r_u^2 + r_v^2
__global__ void __launch_bounds__(256)
k_of_cg_update(const OfArgs a, double alpha, double *__restrict__ u, double *__restrict__ v, double *__restrict__ ru,
               double *__restrict__ rv, const double *__restrict__ pu, const double *__restrict__ pv,
               const double *__restrict__ apu, const double *__restrict__ apv, dd *partials) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  dd pr = {0.0, 0.0};
  for (int j = blockIdx.y * blockDim.y + threadIdx.y; j < a.ny; j += gridDim.y * blockDim.y) {
    if (i >= a.nx) break;
    const long long p = i + (long long)a.pitch * j;
    store_neumann(u, i, j, a.nx, a.ny, a.pitch, fma(alpha, pu[p], u[p]));
    store_neumann(v, i, j, a.nx, a.ny, a.pitch, fma(alpha, pv[p], v[p]));
    const double r0 = fma(-alpha, apu[p], ru[p]), r1 = fma(-alpha, apv[p], rv[p]);
    store_neumann(ru, i, j, a.nx, a.ny, a.pitch, r0);
    store_neumann(rv, i, j, a.nx, a.ny, a.pitch, r1);
    dd_add(pr, r0 * r0);
    dd_add(pr, r1 * r1);
  }
  pr = dd_block_reduce(pr);
  if (threadIdx.x == 0 && threadIdx.y == 0) partials[blockIdx.x + gridDim.x * blockIdx.y] = pr;
}

// cgTmp0 = Residual + beta * cgTmp0 (ghost cells with it); beta = 0 with copy: cgTmp0 = Residual
__global__ void __launch_bounds__(256)
k_of_cg_dir(const OfArgs a, double beta, int copy, const double *__restrict__ ru, const double *__restrict__ rv,
            double *__restrict__ pu, double *__restrict__ pv) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  for (int j = blockIdx.y * blockDim.y + threadIdx.y; j < a.ny; j += gridDim.y * blockDim.y) {
    if (i >= a.nx) break;
    const long long p = i + (long long)a.pitch * j;
    store_neumann(pu, i, j, a.nx, a.ny, a.pitch, copy ? ru[p] : fma(beta, pu[p], ru[p]));
    store_neumann(pv, i, j, a.nx, a.ny, a.pitch, copy ? rv[p] : fma(beta, pv[p], rv[p]));
  }
}

// ---- the same CG on several blocks, coarsest level agglomerated: rank 0 runs k_of_cg on the global grid ---------------
// dense global arrays [gny][gnx] <-> rank 0's padded arrays (pitch gnx + 2, one ghost cell all round)

// in: 7 dense arrays u, v, rhs_u, rhs_v, cuu, cuv, cvv -> pad[0..3], pad[10..12] (u, v with their Neumann ghost cells)
__global__ void __launch_bounds__(256) k_of_agg_in(const double *__restrict__ dense, long long tot, int gnx, int gny,
                                                   double *pad, long long padsz) {
  static constexpr int which[7] = {0, 1, 2, 3, 10, 11, 12};
  const int pitch = gnx + 2;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < 7 * tot; t += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(t / tot);
    const long long q = t - k * tot;
    const int i = (int)(q % gnx), j = (int)(q / gnx);
    double *dst = pad + which[k] * padsz + pitch + 1;
    if (k < 2) store_neumann(dst, i, j, gnx, gny, pitch, dense[t]);
    else dst[i + (long long)pitch * j] = dense[t];
  }
}

// out: u, v -> 2 dense arrays, the iteration count behind them
__global__ void __launch_bounds__(256) k_of_agg_out(const double *__restrict__ pad, long long padsz, int gnx, double *dense,
                                                    long long tot, const int *iters) {
  const int pitch = gnx + 2;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < 2 * tot; t += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(t / tot);
    const long long q = t - k * tot;
    dense[t] = pad[k * padsz + pitch + 1 + (q % gnx) + (long long)pitch * (q / gnx)];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) dense[2 * tot] = (double)*iters;
}

__global__ void k_of_agg_iters(const double *src, int *iters) { *iters = (int)*src; }

// ---- host side ---------------------------------------------------------------------

static System2 *get_system(examg_ctx *c, int h) {
  if (!c || h < 0 || h >= (int)c->systems.size()) return nullptr;
  return &c->systems[h];
}

static Field *cell_field(examg_ctx *c, int h, int level) {
  Field *f = get_field(c, h);
  if (!f || f->loc != 1 || f->level != level || f->ncomp != 1) return nullptr;
  return f;
}

static OfArgs of_args(examg_ctx *c, const System2 &s) {
  const LevelGeom &g = *geom(c, s.level);
  OfArgs a;
  a.nx = g.n[0] - 1;
  a.ny = g.n[1] - 1;
  a.pitch = g.pitch;
  a.par = (g.goff[0] + g.goff[1]) & 1;
  a.xlo = a.ylo = 0;
  a.xhi = a.nx;
  a.yhi = a.ny;
  a.phys = (c->nb[0][0] ? 0 : 1) | (c->nb[0][1] ? 0 : 2) | (c->nb[1][0] ? 0 : 4) | (c->nb[1][1] ? 0 : 8);
  for (int k = 0; k < 5; ++k) a.L[k] = s.L[k];
  a.a2 = s.a2;
  a.cuu = field_ptr(c, &c->fields[s.cuu], -1);
  a.cuv = field_ptr(c, &c->fields[s.cuv], -1);
  a.cvv = field_ptr(c, &c->fields[s.cvv], -1);
  return a;
}

static dim3 rows_grid(int nx, int ny, dim3 block, int nz = 1) {
  int gx = (nx + block.x - 1) / block.x, gy = (ny + block.y - 1) / block.y;
  const int cap = 148 * 16;  // rows beyond that are covered by the grid stride
  if (gx * gy * nz > cap) gy = cap / (gx * nz) > 0 ? cap / (gx * nz) : 1;
  return dim3(gx, gy, nz);
}

int launch_bc_cell(examg_ctx *c, const Field &f, int slot) {
  if (c->cfg.dim != 2) return set_error(EXAMG_EUNSUP, "apply bc: cell-localised fields are 2D so far");
  const LevelGeom &g = *geom(c, f.level);
  const int nx = g.n[0] - 1, ny = g.n[1] - 1, per = 2 * (nx + 2) + 2 * ny;
  for (int comp = 0; comp < f.ncomp; ++comp) {
    k_of_neumann<<<(per + 255) / 256, 256, 0, c->stream>>>(field_ptr(c, &f, slot, comp), nx, ny, g.pitch);
    EXAMG_LAUNCH_CHECK(c);
  }
  return EXAMG_OK;
}

int ensure_shadow(examg_ctx *c, Field &fu);  // examg_fast7.cu

// sweeps fused into one pass (EXAMG_OF_FUSE_SWEEPS = 0..3; 0: one kernel per colour)
static int of_max_sweeps() {
  static int v = -1;
  if (v < 0) {
    const char *e = getenv("EXAMG_OF_FUSE_SWEEPS");
    v = e ? atoi(e) : 3;
    v = v < 0 ? 0 : v > 3 ? 3 : v;
  }
  return v;
}

static bool of_fused_usable(examg_ctx *c, const System2 &s) {
  const LevelGeom &g = *geom(c, s.level);
  const int nx = g.n[0] - 1, ny = g.n[1] - 1;
  // several blocks: blocks start on even cells (the colour pattern of the lanes), up to two sweeps per pass (S = 4 layers)
  if (c->world > 1 && (((g.goff[0] | g.goff[1]) & 1) || getenv("EXAMG_OF_NO_FUSED_BLOCKS"))) return false;
  return of_max_sweeps() > 0 && nx % 2 == 0 && nx >= 128 && ny >= 16;
}

// several blocks: `communicate` of both unknowns in one handshake; the ghost cells at block faces take the neighbour's
// cells (the Neumann copies the kernels left there are overwritten), the ones at the physical boundary stay
static int of_exchange2(examg_ctx *c, Field *fa, Field *fb) {
  if (c->world == 1) return EXAMG_OK;
  Field *fs[2] = {fa, fb};
  const int slots[2] = {fa->cur, fb->cur}, depths[2] = {1, 1};
  return do_exchange_many(c, fs, slots, depths, fa == fb ? 1 : 2);
}

template <int S>
static int launch_of_rbw_s(examg_ctx *c, const OfArgs &a, const double *u, const double *v, const double *bu,
                           const double *bv, double *uo, double *vo, double relax) {
  constexpr int IW = 64 - 2 * S;
  const int nstrips = (a.nx + IW - 1) / IW;
  const int gx = (nstrips + 3) / 4;
  // rows per segment: enough CTAs for a few waves, but segments long against their 2S warm-up rows
  int nseg = (148 * 12 + gx - 1) / gx;
  const int max_seg = a.ny / (16 * S) > 0 ? a.ny / (16 * S) : 1;
  if (nseg > max_seg) nseg = max_seg;
  const int seg_rows = (a.ny + nseg - 1) / nseg;
  nseg = (a.ny + seg_rows - 1) / seg_rows;
  if (c->world > 1) k_of_rbw<S, true><<<dim3(gx, nseg), 128, 0, c->stream>>>(a, u, v, bu, bv, uo, vo, relax, seg_rows, nstrips);
  else k_of_rbw<S, false><<<dim3(gx, nseg), 128, 0, c->stream>>>(a, u, v, bu, bv, uo, vo, relax, seg_rows, nstrips);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

// `sweeps` red-black sweeps in one pass; exchanges the buffers of u and v with their second buffers
static int launch_of_rbw(examg_ctx *c, const System2 &s, Field &fu, Field &fv, const double *bu, const double *bv,
                         double relax, int sweeps) {
  const LevelGeom &g = *geom(c, s.level);
  OfArgs a = of_args(c, s);
  const int S = 2 * sweeps;
  if (!(a.phys & 1)) a.xlo = -S;
  if (!(a.phys & 2)) a.xhi = a.nx + S;
  if (!(a.phys & 4)) a.ylo = -S;
  if (!(a.phys & 8)) a.yhi = a.ny + S;
  const double *u = field_ptr(c, &fu, -1), *v = field_ptr(c, &fv, -1);
  double *uo = fu.shadow + g.origin, *vo = fv.shadow + g.origin;
  int rc;
  switch (sweeps) {
    case 1: rc = launch_of_rbw_s<2>(c, a, u, v, bu, bv, uo, vo, relax); break;
    case 2: rc = launch_of_rbw_s<4>(c, a, u, v, bu, bv, uo, vo, relax); break;
    default: rc = launch_of_rbw_s<6>(c, a, u, v, bu, bv, uo, vo, relax); break;
  }
  if (rc) return rc;
  for (Field *f : {&fu, &fv}) {
    double *t = f->slot[f->cur];
    f->slot[f->cur] = f->shadow;
    f->shadow = t;
    f->swapped = !f->swapped;
  }
  return EXAMG_OK;
}

#define OF_SYS(var, h)                  \
  System2 *var = get_system(c, h);      \
  if (!var) return set_error(EXAMG_EINVAL, "%s: bad system handle %d", __func__, h)
#define OF_FIELD(var, h, lvl)                \
  Field *var = cell_field(c, h, lvl);        \
  if (!var) return set_error(EXAMG_EINVAL, "%s: field %d is not a scalar cell field of level %d", __func__, h, lvl)

}  // namespace examg

extern "C" {

int examg_system2_create(examg_ctx *c, const examg_system2_desc *d) {
  if (!c || !d) return set_error(EXAMG_EINVAL, "examg_system2_create: null argument");
  if (c->cfg.dim != 2) return set_error(EXAMG_EUNSUP, "examg_system2_create: 2D programs only");
  Stencil *st = get_stencil(c, d->laplace);
  if (!st || st->coeff_field >= 0 || st->nst != 5)
    return set_error(EXAMG_EINVAL, "examg_system2_create: laplace must be a constant 5-point stencil");
  static const int want[5][2] = {{0, 0}, {-1, 0}, {1, 0}, {0, -1}, {0, 1}};
  for (int k = 0; k < 5; ++k)
    if (st->off[k][0] != want[k][0] || st->off[k][1] != want[k][1] || st->off[k][2] != 0)
      return set_error(EXAMG_EUNSUP, "examg_system2_create: stencil entries must be declared centre,-x,+x,-y,+y");
  const LevelGeom &g = *geom(c, st->level);
  if (g.n[0] < 3 || g.n[1] < 3) return set_error(EXAMG_EUNSUP, "examg_system2_create: fewer than 2 cells per dimension");
  System2 s;
  s.level = st->level;
  s.a2 = d->scale;
  for (int k = 0; k < 5; ++k) s.L[k] = st->coef[k];
  const int hs[3] = {d->c_uu, d->c_uv, d->c_vv};
  for (int k = 0; k < 3; ++k)
    if (!cell_field(c, hs[k], s.level))
      return set_error(EXAMG_EINVAL, "examg_system2_create: coefficient %d is not a scalar cell field of level %d", k, s.level);
  s.cuu = d->c_uu;
  s.cuv = d->c_uv;
  s.cvv = d->c_vv;
  c->systems.push_back(s);
  return (int)c->systems.size() - 1;
}

int examg_system2_residual(examg_ctx *c, int sys, int u, int v, int rhs_u, int rhs_v, int res_u, int res_v,
                           double *norm2_out) {
  OF_SYS(s, sys);
  OF_FIELD(fu, u, s->level); OF_FIELD(fv, v, s->level); OF_FIELD(bu, rhs_u, s->level); OF_FIELD(bv, rhs_v, s->level);
  OF_FIELD(ru, res_u, s->level); OF_FIELD(rv, res_v, s->level);
  {
    const int rc = of_exchange2(c, fu, fv);
    if (rc) return rc;
  }
  const OfArgs a = of_args(c, *s);
  const dim3 block(64, 4), grid = rows_grid(a.nx, a.ny, block);
  if (norm2_out) {
    k_of_residual<true><<<grid, block, 0, c->stream>>>(a, field_ptr(c, fu, -1), field_ptr(c, fv, -1), field_ptr(c, bu, -1),
                                                       field_ptr(c, bv, -1), field_ptr(c, ru, -1), field_ptr(c, rv, -1),
                                                       c->partials);
    EXAMG_LAUNCH_CHECK(c);
    k_of_final_sum<<<1, 1024, 0, c->stream>>>(c->partials, (int)(grid.x * grid.y), c->red_out);
    EXAMG_LAUNCH_CHECK(c);
    return finish_sum(c, 0, norm2_out);
  }
  k_of_residual<false><<<grid, block, 0, c->stream>>>(a, field_ptr(c, fu, -1), field_ptr(c, fv, -1), field_ptr(c, bu, -1),
                                                      field_ptr(c, bv, -1), field_ptr(c, ru, -1), field_ptr(c, rv, -1),
                                                      nullptr);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

int examg_system2_smooth_rb_color(examg_ctx *c, int sys, int u, int v, int rhs_u, int rhs_v, double relax, int color) {
  OF_SYS(s, sys);
  OF_FIELD(fu, u, s->level); OF_FIELD(fv, v, s->level); OF_FIELD(bu, rhs_u, s->level); OF_FIELD(bv, rhs_v, s->level);
  if (color != 0 && color != 1) return set_error(EXAMG_EINVAL, "examg_system2_smooth_rb_color: colour %d", color);
  {
    const int rc = of_exchange2(c, fu, fv);
    if (rc) return rc;
  }
  const OfArgs a = of_args(c, *s);
  const dim3 block(64, 4), grid((a.nx / 2 + 1 + block.x - 1) / block.x, (a.ny + block.y - 1) / block.y);
  k_of_smooth_color<<<grid, block, 0, c->stream>>>(a, field_ptr(c, fu, -1), field_ptr(c, fv, -1), field_ptr(c, bu, -1),
                                                   field_ptr(c, bv, -1), relax, color);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

// several blocks, S half-sweeps in a row: S layers of u, v and one less of the right-hand sides and the coefficients
// (the redundant updates in the neighbours' layers read them), in one handshake; the static fields only when their
// layers are stale
static int of_exchange_deep(examg_ctx *c, const System2 &s, Field *fu, Field *fv, Field *bu, Field *bv, int S) {
  Field *fs[7] = {fu, fv, bu, bv, &c->fields[s.cuu], &c->fields[s.cuv], &c->fields[s.cvv]};
  Field *todo[7];
  int slots[7], depths[7], n = 0;
  for (int k = 0; k < 7; ++k) {
    const int d = k < 2 ? S : S - 1;
    if (k >= 2 && fs[k]->halo_depth >= d) continue;
    todo[n] = fs[k]; slots[n] = fs[k]->cur; depths[n++] = d;
  }
  return do_exchange_many(c, todo, slots, depths, n);
}

int examg_system2_smooth_rb(examg_ctx *c, int sys, int u, int v, int rhs_u, int rhs_v, double relax, int nsweeps) {
  {
    OF_SYS(s, sys);
    OF_FIELD(fu, u, s->level); OF_FIELD(fv, v, s->level); OF_FIELD(bu, rhs_u, s->level); OF_FIELD(bv, rhs_v, s->level);
    if (of_fused_usable(c, *s) && fu != fv) {  // up to three sweeps per pass over HBM
      int rc;
      if ((rc = ensure_shadow(c, *fu)) || (rc = ensure_shadow(c, *fv))) return rc;
      for (int it = 0; it < nsweeps;) {
        int plan = of_max_sweeps();
        if (c->world > 1 && plan > 2) plan = 2;  // four ghost layers per side
        if (plan > nsweeps - it) plan = nsweeps - it;
        if (c->world > 1 && (rc = of_exchange_deep(c, *s, fu, fv, bu, bv, 2 * plan))) return rc;
        if ((rc = launch_of_rbw(c, *s, *fu, *fv, field_ptr(c, bu, -1), field_ptr(c, bv, -1), relax, plan))) return rc;
        fu->halo_depth = fv->halo_depth = 0;
        it += plan;
      }
      return EXAMG_OK;
    }
  }
  {
    // several blocks, levels too small for the wavefront kernel: the same redundancy with the colour kernel -- one
    // handshake per two sweeps instead of one per colour; half-sweep h of S works on S - 1 - h layers of the
    // neighbours' cells
    OF_SYS(s, sys);
    OF_FIELD(fu, u, s->level); OF_FIELD(fv, v, s->level); OF_FIELD(bu, rhs_u, s->level); OF_FIELD(bv, rhs_v, s->level);
    const OfArgs a0 = of_args(c, *s);
    if (c->world > 1 && fu != fv && a0.nx >= GHOST && a0.ny >= GHOST && !getenv("EXAMG_OF_NO_FUSED_BLOCKS")) {
      for (int it = 0; it < nsweeps;) {
        const int plan = nsweeps - it >= 2 ? 2 : 1, S = 2 * plan;
        int rc = of_exchange_deep(c, *s, fu, fv, bu, bv, S);
        if (rc) return rc;
        for (int h = 0; h < S; ++h) {
          OfArgs a = a0;
          const int ext = S - 1 - h;
          if (!(a.phys & 1)) a.xlo = -ext;
          if (!(a.phys & 2)) a.xhi = a.nx + ext;
          if (!(a.phys & 4)) a.ylo = -ext;
          if (!(a.phys & 8)) a.yhi = a.ny + ext;
          const dim3 block(64, 4), grid(((a.xhi - a.xlo) / 2 + 1 + block.x - 1) / block.x, (a.yhi - a.ylo + block.y - 1) / block.y);
          k_of_smooth_color_mb<<<grid, block, 0, c->stream>>>(a, field_ptr(c, fu, -1), field_ptr(c, fv, -1),
                                                              field_ptr(c, bu, -1), field_ptr(c, bv, -1), relax, h & 1);
          EXAMG_LAUNCH_CHECK(c);
        }
        fu->halo_depth = fv->halo_depth = 0;
        it += plan;
      }
      return EXAMG_OK;
    }
  }
  for (int it = 0; it < nsweeps; ++it)
    for (int color = 0; color < 2; ++color) {
      int rc = examg_system2_smooth_rb_color(c, sys, u, v, rhs_u, rhs_v, relax, color);
      if (rc) return rc;
    }
  return EXAMG_OK;
}

static int cell_transfer(examg_ctx *c, const int *src, const int *dst, int n, bool restrict_op, const char *who) {
  if (n < 1 || n > 3) return set_error(EXAMG_EINVAL, "%s: 1..3 fields", who);
  Pair pr;
  int lf = -1, lc = -1;
  for (int k = 0; k < n; ++k) {
    Field *fs = get_field(c, src[k]), *fd = get_field(c, dst[k]);
    if (!fs || !fd || fs->loc != 1 || fd->loc != 1) return set_error(EXAMG_EINVAL, "%s: not a cell field", who);
    const int fine = restrict_op ? fs->level : fd->level, coarse = restrict_op ? fd->level : fs->level;
    if (coarse != fine - 1 || (lf >= 0 && fine != lf)) return set_error(EXAMG_EINVAL, "%s: levels %d -> %d", who, fs->level, fd->level);
    lf = fine;
    lc = coarse;
    pr.src[k] = field_ptr(c, fs, -1);
    pr.dst[k] = field_ptr(c, fd, -1);
    fd->halo_depth = 0;
  }
  const LevelGeom &gf = *geom(c, lf), &gc = *geom(c, lc);
  if (c->cfg.dim == 3) {
    const int nxc = gc.n[0] - 1, nyc = gc.n[1] - 1, nzc = gc.n[2] - 1;
    const long long total = (long long)nxc * nyc * nzc;
    long long gx = (total + 255) / 256;
    if (gx > 148 * 8) gx = 148 * 8;
    const dim3 grid((unsigned)gx, 1, n);
    if (restrict_op)
      k_cell_restrict3<<<grid, 256, 0, c->stream>>>(pr, nxc, nyc, nzc, gf.pitch, gf.plane, gc.pitch, gc.plane);
    else
      k_cell_prolong3<<<grid, 256, 0, c->stream>>>(pr, nxc, nyc, nzc, gf.pitch, gf.plane, gc.pitch, gc.plane);
    EXAMG_LAUNCH_CHECK(c);
    return EXAMG_OK;
  }
  const int nxc = gc.n[0] - 1, nyc = gc.n[1] - 1;
  const dim3 block(64, 4), grid = rows_grid(nxc, nyc, block, n);
  if (restrict_op)
    k_of_restrict<<<grid, block, 0, c->stream>>>(pr, nxc, nyc, gf.pitch, gc.pitch);
  else
    k_of_prolong_add<<<grid, block, 0, c->stream>>>(pr, nxc, nyc, gf.pitch, gc.pitch);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

int examg_restrict_cell(examg_ctx *c, const int *fine, const int *coarse, int nfields) {
  return cell_transfer(c, fine, coarse, nfields, true, "examg_restrict_cell");
}

int examg_prolong_add_cell(examg_ctx *c, const int *coarse, const int *fine, int nfields) {
  return cell_transfer(c, coarse, fine, nfields, false, "examg_prolong_add_cell");
}

}  // extern "C"

// mgCycle@coarsest on several blocks: the loops of k_of_cg as kernels over this block's cells, `communicate` of cgTmp0
// before A * cgTmp0, global sums (all-reduce + one host synchronisation) where the program has its reductions.
// fields = { u, v, rhs_u, rhs_v, residual_u, residual_v, cgTmp0_u, cgTmp0_v, cgTmp1_u, cgTmp1_v }
static int of_cg_blocks(examg_ctx *c, const System2 &s, const int *fields, int max_it, double rel_tol, int *iters_out) {
  Field *f[10];
  double *p[10];
  for (int k = 0; k < 10; ++k) {
    f[k] = cell_field(c, fields[k], s.level);
    p[k] = field_ptr(c, f[k], -1);
  }
  int rc;
  double n2 = 0.0;
  // Residual = RHS - A * Solution (+ its sum of squares); cgTmp0 = Residual
  const int sys = (int)(&s - c->systems.data());
  if ((rc = examg_system2_residual(c, sys, fields[0], fields[1], fields[2], fields[3], fields[4], fields[5], &n2))) return rc;
  const OfArgs a = of_args(c, s);
  const dim3 block(64, 4), grid = rows_grid(a.nx, a.ny, block);
  const int nb = (int)(grid.x * grid.y);
  k_of_cg_dir<<<grid, block, 0, c->stream>>>(a, 0.0, 1, p[4], p[5], p[6], p[7]);
  EXAMG_LAUNCH_CHECK(c);
  double curRes = std::sqrt(n2);
  const double initRes = curRes;
  int iters = 0;
  for (int step = 0; step < max_it; ++step) {
    iters = step + 1;
    if ((rc = of_exchange2(c, f[6], f[7]))) return rc;
    k_of_cg_apply<<<grid, block, 0, c->stream>>>(a, p[4], p[5], p[6], p[7], p[8], p[9], c->partials);
    EXAMG_LAUNCH_CHECK(c);
    for (int k = 0; k < 4; ++k) {
      k_of_final_sum<<<1, 1024, 0, c->stream>>>(c->partials + k * nb, nb, c->red_out + 4 + k);
      EXAMG_LAUNCH_CHECK(c);
    }
    double sums[4];
    if ((rc = finish_sums(c, 4, 4, sums))) return rc;
    const double alpha = (sums[0] + sums[1]) / (sums[2] + sums[3]);
    k_of_cg_update<<<grid, block, 0, c->stream>>>(a, alpha, p[0], p[1], p[4], p[5], p[6], p[7], p[8], p[9], c->partials);
    EXAMG_LAUNCH_CHECK(c);
    k_of_final_sum<<<1, 1024, 0, c->stream>>>(c->partials, nb, c->red_out + 4);
    EXAMG_LAUNCH_CHECK(c);
    double pr = 0.0;
    if ((rc = finish_sums(c, 4, 1, &pr))) return rc;
    const double nextRes = std::sqrt(pr);
    if (nextRes <= rel_tol * initRes) break;
    const double beta = (nextRes * nextRes) / (curRes * curRes);
    k_of_cg_dir<<<grid, block, 0, c->stream>>>(a, beta, 0, p[4], p[5], p[6], p[7]);
    EXAMG_LAUNCH_CHECK(c);
    curRes = nextRes;
  }
  for (int k : {0, 1, 4, 5, 6, 7}) f[k]->halo_depth = 0;
  c->int_host[0] = iters;
  EXAMG_CUDA(cudaMemcpyAsync(c->int_out, c->int_host, sizeof(int), cudaMemcpyHostToDevice, c->stream));
  EXAMG_SYNC(c);
  if (iters_out) *iters_out = iters;
  return EXAMG_OK;
}

// The level is agglomerated (comm_agg_setup_cells, examg_optflow_create): every block stores its cells of u, v, the
// right-hand sides and the coefficients into dense global arrays on rank 0, rank 0 runs the one-block kernel on the
// global grid, every block takes its cells of the result.  No host synchronisation: the cycle stays one CUDA graph.
static bool of_cg_agg_usable(examg_ctx *c, const System2 &s) {
  int gn[3];
  long long tot;
  return comm_agg_cells(c, gn, &tot) == s.level && !getenv("EXAMG_OF_NO_AGG");
}

static int of_cg_agg(examg_ctx *c, System2 &s, const int *fields, int max_it, double rel_tol, int *iters_out) {
  int gn[3], rc;
  long long tot;
  comm_agg_cells(c, gn, &tot);
  Field *in[7] = {cell_field(c, fields[0], s.level), cell_field(c, fields[1], s.level), cell_field(c, fields[2], s.level),
                  cell_field(c, fields[3], s.level), &c->fields[s.cuu], &c->fields[s.cuv], &c->fields[s.cvv]};
  for (int k = 0; k < 7; ++k)
    if ((rc = comm_agg_gather(c, *in[k], k))) return rc;
  if (c->cfg.rank == 0) {
    if (!s.agg_pad) return set_error(EXAMG_ESTATE, "examg_system2_cg_coarse: no global arrays for the agglomerated level");
    const int pitch = gn[0] + 2;
    const long long padsz = (long long)pitch * (gn[1] + 2);
    const int grid = (int)((7 * tot + 255) / 256 > 148 ? 148 : (7 * tot + 255) / 256);
    k_of_agg_in<<<grid, 256, 0, c->stream>>>(comm_agg_in(c), tot, gn[0], gn[1], s.agg_pad, padsz);
    EXAMG_LAUNCH_CHECK(c);
    double *p[13];
    for (int k = 0; k < 13; ++k) p[k] = s.agg_pad + k * padsz + pitch + 1;
    OfArgs a = of_args(c, s);
    a.nx = gn[0];
    a.ny = gn[1];
    a.pitch = pitch;
    a.par = 0;
    a.xlo = a.ylo = 0;
    a.xhi = a.nx;
    a.yhi = a.ny;
    a.phys = 15;
    a.cuu = p[10];
    a.cuv = p[11];
    a.cvv = p[12];
    int nt = 32;
    while (nt < 1024 && nt < a.nx * a.ny) nt <<= 1;
    k_of_cg<<<1, nt, 0, c->stream>>>(a, p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], p[8], p[9], max_it, rel_tol, c->int_out);
    EXAMG_LAUNCH_CHECK(c);
    k_of_agg_out<<<grid, 256, 0, c->stream>>>(s.agg_pad, padsz, gn[0], comm_agg_out(c), tot, c->int_out);
    EXAMG_LAUNCH_CHECK(c);
  }
  if ((rc = comm_agg_scatter(c, *in[0], 0)) || (rc = comm_agg_scatter(c, *in[1], 1))) return rc;
  if (c->cfg.rank != 0) {
    k_of_agg_iters<<<1, 1, 0, c->stream>>>(comm_agg_out_peer(c) + 2 * tot, c->int_out);
    EXAMG_LAUNCH_CHECK(c);
  }
  if ((rc = examg_apply_bc(c, fields[0], -1)) || (rc = examg_apply_bc(c, fields[1], -1))) return rc;
  if (iters_out) {
    EXAMG_CUDA(cudaMemcpyAsync(c->int_host, c->int_out, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    EXAMG_SYNC(c);
    *iters_out = c->int_host[0];
  }
  return EXAMG_OK;
}

// global arrays of the agglomerated level on rank 0 (allocation: not inside a captured cycle)
static int of_agg_prepare(examg_ctx *c, System2 &s) {
  int gn[3];
  long long tot;
  if (comm_agg_cells(c, gn, &tot) != s.level || c->cfg.rank != 0 || s.agg_pad) return EXAMG_OK;
  const size_t padsz = (size_t)(gn[0] + 2) * (gn[1] + 2);
  s.agg_pad = dev_alloc(c, 13 * padsz * sizeof(double));
  if (!s.agg_pad) return set_error(EXAMG_ENOMEM, "global arrays of the agglomerated level %d", s.level);
  EXAMG_CUDA(cudaMemsetAsync(s.agg_pad, 0, 13 * padsz * sizeof(double), c->stream));
  return EXAMG_OK;
}

extern "C" {

int examg_system2_cg_coarse(examg_ctx *c, int sys, const int *fields, int max_it, double rel_tol, int *iters_out) {
  OF_SYS(s, sys);
  if (!fields) return set_error(EXAMG_EINVAL, "examg_system2_cg_coarse: null field list");
  double *p[10];
  for (int k = 0; k < 10; ++k) {
    Field *f = cell_field(c, fields[k], s->level);
    if (!f) return set_error(EXAMG_EINVAL, "examg_system2_cg_coarse: field %d (#%d) is not a cell field of level %d", fields[k], k, s->level);
    p[k] = field_ptr(c, f, -1);
  }
  if (c->world > 1)
    return of_cg_agg_usable(c, *s) ? of_cg_agg(c, *s, fields, max_it, rel_tol, iters_out)
                                   : of_cg_blocks(c, *s, fields, max_it, rel_tol, iters_out);
  const OfArgs a = of_args(c, *s);
  int nt = 32;
  while (nt < 1024 && nt < a.nx * a.ny) nt <<= 1;
  k_of_cg<<<1, nt, 0, c->stream>>>(a, p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], p[8], p[9], max_it, rel_tol, c->int_out);
  EXAMG_LAUNCH_CHECK(c);
  if (iters_out) {
    EXAMG_CUDA(cudaMemcpyAsync(c->int_host, c->int_out, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    EXAMG_SYNC(c);
    *iters_out = c->int_host[0];
  }
  return EXAMG_OK;
}

}  // extern "C"

// ---- the whole program: InitFields / doSolve@finest / mgCycle@all -----------------------------

struct examg_optflow {
  examg_ctx *c = nullptr;
  examg_optflow_desc desc;
  int nlev = 0;
  std::vector<int> f[9];   // per level: u, v, rhs_u, rhs_v, residual_u, residual_v, IxIx, IxIy, IyIy
  std::vector<int> sys;    // per level
  int img0 = -1, img1 = -1;
  int cg[4] = {-1, -1, -1, -1};  // cgTmp0_u, cgTmp0_v, cgTmp1_u, cgTmp1_v @coarsest
  struct Graph { cudaGraphExec_t exec = nullptr; long long launches = 0; std::vector<int> exit_state; };
  std::map<std::vector<int>, Graph> graphs;  // keyed by which buffer of u, v is active on every level
  // several blocks: the legs before and after the coarse solve (which synchronises with the host for its global sums)
  std::map<std::vector<int>, Graph> graphs_down, graphs_up;
};

namespace {

int of_lv(examg_optflow *o, int level) { return level - o->c->cfg.min_level; }

int of_zero(examg_ctx *c, int h) {  // u@coarser = 0; apply bc: the whole allocation, ghosts included
  Field *f = get_field(c, h);
  const LevelGeom &g = *geom(c, f->level);
  EXAMG_CUDA(cudaMemsetAsync(f->slot[f->cur], 0, sizeof(double) * (size_t)g.vol, c->stream));
  c->launches++;
  return EXAMG_OK;
}

// which buffer of u, v is active, and how many ghost layers of every field hold the neighbour blocks' values (the
// host-side decisions of a captured leg depend on both, and a replay has to leave both as the capture did)
std::vector<int> of_state(examg_optflow *o) {
  std::vector<int> key;
  for (int k = 0; k < 2; ++k)
    for (int h : o->f[k]) key.push_back(o->c->fields[h].swapped ? 1 : 0);
  if (o->c->world > 1)
    for (int k = 0; k < 9; ++k)
      for (int h : o->f[k]) key.push_back(o->c->fields[h].halo_depth);
  return key;
}

void of_restore(examg_optflow *o, const std::vector<int> &key) {
  size_t pos = 0;
  for (int k = 0; k < 2; ++k)
    for (int h : o->f[k]) {
      Field &f = o->c->fields[h];
      if ((f.swapped ? 1 : 0) != key[pos++]) {
        double *t = f.slot[f.cur];
        f.slot[f.cur] = f.shadow;
        f.shadow = t;
        f.swapped = !f.swapped;
      }
    }
  if (o->c->world > 1)
    for (int k = 0; k < 9; ++k)
      for (int h : o->f[k]) o->c->fields[h].halo_depth = key[pos++];
}

int of_vcycle_rec(examg_optflow *o, int level) {
  examg_ctx *c = o->c;
  const int l = of_lv(o, level);
  if (level == c->cfg.min_level) {
    const int fl[10] = {o->f[0][l], o->f[1][l], o->f[2][l], o->f[3][l], o->f[4][l], o->f[5][l],
                        o->cg[0], o->cg[1], o->cg[2], o->cg[3]};
    return examg_system2_cg_coarse(c, o->sys[l], fl, o->desc.cg_max_it, o->desc.cg_rel_tol, nullptr);
  }
  int rc;
  if ((rc = examg_system2_smooth_rb(c, o->sys[l], o->f[0][l], o->f[1][l], o->f[2][l], o->f[3][l], o->desc.relax,
                                    o->desc.nu_pre))) return rc;
  if ((rc = examg_system2_residual(c, o->sys[l], o->f[0][l], o->f[1][l], o->f[2][l], o->f[3][l], o->f[4][l],
                                   o->f[5][l], nullptr))) return rc;
  const int res[2] = {o->f[4][l], o->f[5][l]}, rhsc[2] = {o->f[2][l - 1], o->f[3][l - 1]};
  if ((rc = examg_restrict_cell(c, res, rhsc, 2))) return rc;
  if ((rc = of_zero(c, o->f[0][l - 1])) || (rc = of_zero(c, o->f[1][l - 1]))) return rc;
  if ((rc = of_vcycle_rec(o, level - 1))) return rc;
  const int uc[2] = {o->f[0][l - 1], o->f[1][l - 1]}, uf[2] = {o->f[0][l], o->f[1][l]};
  if ((rc = examg_prolong_add_cell(c, uc, uf, 2))) return rc;
  return examg_system2_smooth_rb(c, o->sys[l], o->f[0][l], o->f[1][l], o->f[2][l], o->f[3][l], o->desc.relax,
                                 o->desc.nu_post);
}

// the same cycle as two loops around the coarse solve
int of_down(examg_optflow *o, int level) {
  examg_ctx *c = o->c;
  int rc;
  for (; level > c->cfg.min_level; --level) {
    const int l = of_lv(o, level);
    if ((rc = examg_system2_smooth_rb(c, o->sys[l], o->f[0][l], o->f[1][l], o->f[2][l], o->f[3][l], o->desc.relax,
                                      o->desc.nu_pre))) return rc;
    if ((rc = examg_system2_residual(c, o->sys[l], o->f[0][l], o->f[1][l], o->f[2][l], o->f[3][l], o->f[4][l],
                                     o->f[5][l], nullptr))) return rc;
    const int res[2] = {o->f[4][l], o->f[5][l]}, rhsc[2] = {o->f[2][l - 1], o->f[3][l - 1]};
    if ((rc = examg_restrict_cell(c, res, rhsc, 2))) return rc;
    if ((rc = of_zero(c, o->f[0][l - 1])) || (rc = of_zero(c, o->f[1][l - 1]))) return rc;
  }
  return EXAMG_OK;
}

int of_up(examg_optflow *o, int level) {
  examg_ctx *c = o->c;
  int rc;
  for (int lv = c->cfg.min_level + 1; lv <= level; ++lv) {
    const int l = of_lv(o, lv);
    const int uc[2] = {o->f[0][l - 1], o->f[1][l - 1]}, uf[2] = {o->f[0][l], o->f[1][l]};
    if ((rc = examg_prolong_add_cell(c, uc, uf, 2))) return rc;
    if ((rc = examg_system2_smooth_rb(c, o->sys[l], o->f[0][l], o->f[1][l], o->f[2][l], o->f[3][l], o->desc.relax,
                                      o->desc.nu_post))) return rc;
  }
  return EXAMG_OK;
}

// capture `leg` once per state, replay it afterwards
int of_replay(examg_optflow *o, std::map<std::vector<int>, examg_optflow::Graph> &graphs, int (*leg)(examg_optflow *, int),
              int level) {
  examg_ctx *c = o->c;
  const std::vector<int> key = of_state(o);
  auto it = graphs.find(key);
  if (it == graphs.end()) {
    const long long l0 = c->launches;
    cudaGraph_t graph = nullptr;
    EXAMG_CUDA(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    const int rc = leg(o, level);
    const cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
    if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
    if (e != cudaSuccess) return set_error(EXAMG_ECUDA, "graph capture of mgCycle failed: %s", cudaGetErrorString(e));
    examg_optflow::Graph g;
    const cudaError_t e2 = cudaGraphInstantiate(&g.exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e2 != cudaSuccess) return set_error(EXAMG_ECUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(e2));
    g.launches = c->launches - l0;
    c->launches = l0;  // nothing ran yet
    g.exit_state = of_state(o);
    of_restore(o, key);
    it = graphs.emplace(key, g).first;
  }
  EXAMG_CUDA(cudaGraphLaunch(it->second.exec, c->stream));
  c->launches += it->second.launches;
  of_restore(o, it->second.exit_state);
  return EXAMG_OK;
}

}  // namespace

extern "C" {

examg_optflow *examg_optflow_create(examg_ctx *c, const examg_optflow_desc *d) {
  if (!c || !d) { set_error(EXAMG_EINVAL, "examg_optflow_create: null argument"); return nullptr; }
  if (c->cfg.dim != 2) {
    set_error(EXAMG_EUNSUP, "examg_optflow_create: 2D programs only");
    return nullptr;
  }
  examg_optflow *o = new examg_optflow;
  o->c = c;
  o->desc = *d;
  o->nlev = c->cfg.max_level - c->cfg.min_level + 1;
  static const char *names[9] = {"u", "v", "rhs_u", "rhs_v", "residual_u", "residual_v", "IxIx", "IxIy", "IyIy"};
  static const int offs[5][3] = {{0, 0, 0}, {-1, 0, 0}, {1, 0, 0}, {0, -1, 0}, {0, 1, 0}};
  bool ok = true;
  for (int level = c->cfg.min_level; level <= c->cfg.max_level && ok; ++level) {
    for (int k = 0; k < 9; ++k) {
      const int h = examg_field_create_cell(c, names[k], level, 1, 1);
      ok = ok && h >= 0;
      o->f[k].push_back(h);
      // u, v, residual_u, residual_v: Neumann; rhs and the coefficients: None
      if (ok && (k < 2 || k == 4 || k == 5)) examg_field_set_bc(c, h, EXAMG_BC_NEUMANN, 0.0);
    }
    if (!ok) break;
    // Stencil Laplace@all (Benchmark/OptFlow2D/2D_FD_OptFlow.exa4:75-81)
    const LevelGeom &g = *geom(c, level);
    const double hx2 = g.h[0] * g.h[0], hy2 = g.h[1] * g.h[1];
    const double coef[5] = {2.0 / hx2 + 2.0 / hy2, -1.0 / hx2, -1.0 / hx2, -1.0 / hy2, -1.0 / hy2};
    const int st = examg_stencil_const(c, level, 5, offs, coef);
    examg_system2_desc sd;
    sd.laplace = st;
    sd.scale = d->alpha * d->alpha;  // alpha ** 2
    sd.c_uu = o->f[6].back();
    sd.c_uv = o->f[7].back();
    sd.c_vv = o->f[8].back();
    const int sh = st >= 0 ? examg_system2_create(c, &sd) : -1;
    ok = ok && sh >= 0;
    o->sys.push_back(sh);
    // second buffers of the fused sweeps: now, graph capture cannot allocate
    if (ok && of_fused_usable(c, c->systems[sh]))
      ok = ensure_shadow(c, c->fields[o->f[0].back()]) == EXAMG_OK && ensure_shadow(c, c->fields[o->f[1].back()]) == EXAMG_OK;
  }
  // several blocks: the coarsest level as one grid on rank 0 for the coarse solve (7 arrays in, 2 out + the iteration
  // count; a few dozen cells)
  if (ok && c->world > 1 && !getenv("EXAMG_OF_NO_AGG"))
    ok = comm_agg_setup_cells(c, c->cfg.min_level, 7, 2, 1) == EXAMG_OK && of_agg_prepare(c, c->systems[o->sys[0]]) == EXAMG_OK;
  if (ok) {
    o->img0 = examg_field_create_cell(c, "img0", c->cfg.max_level, 1, 1);
    o->img1 = examg_field_create_cell(c, "img1", c->cfg.max_level, 1, 1);
    ok = o->img0 >= 0 && o->img1 >= 0;
    static const char *cgn[4] = {"cgTmp0_u", "cgTmp0_v", "cgTmp1_u", "cgTmp1_v"};
    for (int k = 0; k < 4 && ok; ++k) {
      o->cg[k] = examg_field_create_cell(c, cgn[k], c->cfg.min_level, 1, 1);
      ok = o->cg[k] >= 0;
    }
    if (ok) {
      examg_field_set_bc(c, o->img0, EXAMG_BC_NEUMANN, 0.0);
      examg_field_set_bc(c, o->img1, EXAMG_BC_NEUMANN, 0.0);
      examg_field_set_bc(c, o->cg[0], EXAMG_BC_NEUMANN, 0.0);
      examg_field_set_bc(c, o->cg[1], EXAMG_BC_NEUMANN, 0.0);
    }
  }
  if (!ok) {
    delete o;  // fields stay with the context and are released by examg_finalize
    return nullptr;
  }
  return o;
}

int examg_optflow_destroy(examg_optflow *o) {
  if (!o) return EXAMG_OK;
  for (auto *m : {&o->graphs, &o->graphs_down, &o->graphs_up})
    for (auto &kv : *m)
      if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
  delete o;
  return EXAMG_OK;
}

int examg_optflow_field(examg_optflow *o, int level, int which) {
  if (!o) return set_error(EXAMG_EINVAL, "null optflow object");
  examg_ctx *c = o->c;
  if (which == EXAMG_OF_IMG0) return o->img0;
  if (which == EXAMG_OF_IMG1) return o->img1;
  if (level < c->cfg.min_level || level > c->cfg.max_level || which < 0 || which > EXAMG_OF_IMG1)
    return set_error(EXAMG_EINVAL, "examg_optflow_field: level %d which %d", level, which);
  return o->f[which][of_lv(o, level)];
}

int examg_optflow_init_fields(examg_optflow *o) {
  if (!o) return set_error(EXAMG_EINVAL, "null optflow object");
  examg_ctx *c = o->c;
  int rc;
  if ((rc = examg_apply_bc(c, o->img0, -1)) || (rc = examg_apply_bc(c, o->img1, -1))) return rc;
  if ((rc = of_exchange2(c, &c->fields[o->img0], &c->fields[o->img1]))) return rc;  // the gradients read [1,0], [0,1]
  const int L = c->cfg.max_level, l = of_lv(o, L);
  const LevelGeom &g = *geom(c, L);
  const int nx = g.n[0] - 1, ny = g.n[1] - 1;
  const dim3 block(64, 4), grid = rows_grid(nx, ny, block);
  auto ptr = [&](int h) { return field_ptr(c, &c->fields[h], -1); };
  k_of_init<<<grid, block, 0, c->stream>>>(ptr(o->img0), ptr(o->img1), nx, ny, g.pitch, 1.0 / g.h[0], -1.0 / g.h[0],
                                           1.0 / g.h[1], -1.0 / g.h[1], ptr(o->f[6][l]), ptr(o->f[7][l]),
                                           ptr(o->f[8][l]), ptr(o->f[2][l]), ptr(o->f[3][l]));
  EXAMG_LAUNCH_CHECK(c);
  for (int k : {2, 3, 6, 7, 8}) c->fields[o->f[k][l]].halo_depth = 0;
  for (int level = L - 1; level >= c->cfg.min_level; --level) {  // initGradients@(all but finest)
    const int k = of_lv(o, level);
    const int fine[3] = {o->f[6][k + 1], o->f[7][k + 1], o->f[8][k + 1]}, coarse[3] = {o->f[6][k], o->f[7][k], o->f[8][k]};
    if ((rc = examg_restrict_cell(c, fine, coarse, 3))) return rc;
  }
  return EXAMG_OK;
}

int examg_optflow_vcycle(examg_optflow *o, int level) {
  if (!o) return set_error(EXAMG_EINVAL, "null optflow object");
  examg_ctx *c = o->c;
  if (level < 0) level = c->cfg.max_level;
  if (level < c->cfg.min_level || level > c->cfg.max_level) return set_error(EXAMG_EINVAL, "examg_optflow_vcycle: level %d", level);
  if (level != c->cfg.max_level || getenv("EXAMG_NO_GRAPH")) return of_vcycle_rec(o, level);
  if (c->world > 1 && !of_cg_agg_usable(c, c->systems[o->sys[0]])) {
    // several blocks, coarse solve on the blocks: it synchronises with the host for its global sums; one graph for
    // the legs on either side of it
    int rc;
    if ((rc = of_replay(o, o->graphs_down, of_down, level))) return rc;
    const int l = 0;
    const int fl[10] = {o->f[0][l], o->f[1][l], o->f[2][l], o->f[3][l], o->f[4][l], o->f[5][l],
                        o->cg[0], o->cg[1], o->cg[2], o->cg[3]};
    if ((rc = examg_system2_cg_coarse(c, o->sys[l], fl, o->desc.cg_max_it, o->desc.cg_rel_tol, nullptr))) return rc;
    return of_replay(o, o->graphs_up, of_up, level);
  }
  // one graph per buffer state (the fused sweeps exchange u, v with their second buffers)
  return of_replay(o, o->graphs, of_vcycle_rec, level);
}

int examg_optflow_residual_norm(examg_optflow *o, double *norm_out) {
  if (!o || !norm_out) return set_error(EXAMG_EINVAL, "examg_optflow_residual_norm: null argument");
  const int l = of_lv(o, o->c->cfg.max_level);
  double n2 = 0.0;
  const int rc = examg_system2_residual(o->c, o->sys[l], o->f[0][l], o->f[1][l], o->f[2][l], o->f[3][l], o->f[4][l],
                                        o->f[5][l], &n2);
  if (rc) return rc;
  *norm_out = std::sqrt(n2);
  return EXAMG_OK;
}

int examg_optflow_cycles(examg_optflow *o, int n, double *norm_out) {
  for (int i = 0; i < n; ++i) {
    const int rc = examg_optflow_vcycle(o, -1);
    if (rc) return rc;
  }
  return norm_out ? examg_optflow_residual_norm(o, norm_out) : EXAMG_OK;
}

int examg_optflow_last_cg_iters(examg_optflow *o, int *iters_out) {
  if (!o || !iters_out) return set_error(EXAMG_EINVAL, "examg_optflow_last_cg_iters: null argument");
  examg_ctx *c = o->c;
  EXAMG_CUDA(cudaMemcpyAsync(c->int_host, c->int_out, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  EXAMG_SYNC(c);
  *iters_out = c->int_host[0];
  return EXAMG_OK;
}

}  // extern "C"
