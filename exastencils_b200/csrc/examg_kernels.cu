// examg_kernels.cu -- generic (any compact stencil, 2D/3D, constant or
// stencil-field coefficients) kernels of libexamg: A*u, residual (+ fused
// sum of squares), Jacobi, one red-black colour, dot/max reductions, streaming
// loops and the Dirichlet boundary update.
//
// Arithmetic contract (bit-identical to oracle/examg_oracle.c; built with
// --fmad=false so only the explicit fma() calls fuse):
//   A*u      : acc = c0*u0; acc = fma(ck, uk, acc), k = 1..nst-1 in declaration order
//   residual : r = rhs - acc
//   smoother : u' = fma(omega/diag, rhs - acc, u)
//   sums     : double-double accumulation of the rounded products
// Reference loops these replace: Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4:113-125,
// 204-218; Testing/SISC/3D_VarCoeff.exa4 (Smoother, UpResidual); iteration space
// C/s/e/baseExt/ir/IR_LoopOverPointsInOneFragment.scala:86-125.
#include "examg_point_ops.cuh"

namespace examg {

enum { M_APPLY = 0, M_RESIDUAL = 1, M_RESNORM = 2, M_JACOBI = 3, M_RB = 4, M_NORMONLY = 5 };
// M_NORMONLY: the sum of squares of the residual without storing it (examg_solver_cycles): a mode of its own so that the
// loop has no store at all -- with the conditional store of M_RESNORM in it the loads of the next points could not be
// hoisted above it (`out` may alias `u` as far as the compiler knows).  Measured: no difference (0.50 ms at 512^3, ncu:
// 97 % warps active, 72 % of the stall samples on the loads, DRAM 56 % -- two DRAM-bound requests per thread in flight).

// One kernel for the five loop shapes.  For M_RB, `u` and `out` alias.
template <int MODE, int NST, bool VAR>
__global__ void __launch_bounds__(256)
k_stencil(const GridArgs g, const __grid_constant__ StencilArgs st, const double *__restrict__ sf,
          long long sfvol, const double *u, const double *__restrict__ rhs, double *out, double w, int color,
          int kz, dd *__restrict__ partials) {
  const int tx = blockIdx.x * blockDim.x + threadIdx.x;
  const int z0 = g.lo[2] + blockIdx.z * kz;
  const int z1 = min(z0 + kz, g.hi[2]);
  dd part = {0.0, 0.0};
  // Sum of squares: the (non-negative) squares of up to eight consecutive planes are added in plain double
  // before they enter the double-double sum -- relative error <= 8 ulp, far inside the 1e-14 the tests
  // allow -- so that the loop-carried TwoSum chain is paid once per eight points, and without a
  // data-dependent branch in the plane loop.
  double sq = 0.0;
  for (int j = g.lo[1] + blockIdx.y * blockDim.y + threadIdx.y; j < g.hi[1]; j += gridDim.y * blockDim.y) {
    if (MODE == M_RESNORM || MODE == M_NORMONLY) {
      // the plane loop in chunks of eight without a data-dependent branch inside (the compiler pipelines it
      // like the plain residual loop); one TwoSum per chunk
      const int i = g.lo[0] + tx;
      if (i < g.hi[0]) {
        const bool cnt_ij = i >= g.rlo[0] && i < g.rhi[0] && j >= g.rlo[1] && j < g.rhi[1];
        const int cz0 = cnt_ij ? max(z0, g.rlo[2]) : z1, cz1 = min(z1, g.rhi[2]);
        for (int kb = z0; kb < z1; kb += 8) {
          const int ke = min(kb + 8, z1);
          for (int k = kb; k < ke; ++k) {
            const long long p = i + (long long)g.pitch * j + g.plane * k;
            const double r = rhs[p] - conv<NST, VAR>(st, sf, sfvol, u, p);
            if (MODE == M_RESNORM) out[p] = r;
            sq += (k >= cz0 && k < cz1) ? r * r : 0.0;
          }
          dd_add(part, sq);
          sq = 0.0;
        }
      }
      continue;
    }
    for (int k = z0; k < z1; ++k) {
      int i;
      if (MODE == M_RB) {
        const int first = g.lo[0] + ((g.lo[0] + j + k + g.par + color) & 1);
        i = first + 2 * tx;
      } else {
        i = g.lo[0] + tx;
      }
      if (i < g.hi[0]) {
        const long long p = i + (long long)g.pitch * j + g.plane * k;
        const double acc = conv<NST, VAR>(st, sf, sfvol, u, p);
        if (MODE == M_APPLY) {
          out[p] = acc;
        } else if (MODE == M_RESIDUAL) {
          out[p] = rhs[p] - acc;
        } else {
          const double ww = VAR ? w / sf[st.centre * sfvol + p] : w;  // w = omega (VAR) or omega/diag
          const double r = rhs[p] - acc;
          out[p] = fma(ww, r, u[p]);
        }
      }
    }
  }
  if (MODE == M_RESNORM || MODE == M_NORMONLY) {
    dd_add(part, sq);
    part = dd_block_reduce(part);
    if (threadIdx.x == 0 && threadIdx.y == 0)
      partials[blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z)] = part;
  }
}

// One colour of an n^dim-way coloured Gauss-Seidel sweep (solver_smoother_coloring = "4-way" / "8-way" / "9-way" /
// "27-way", C/s/e/solver/l3/L3_VankaForEquation.scala:154-183): in place on the points with i_d % ncol == c_d.
// Points of one colour are at least ncol >= 2 apart in every dimension, so a compact stencil never couples them and
// the result does not depend on the traversal order.
// SPLIT (two colours per dimension, stencil field): the coefficients come from the split-row copy of the operator
// (see k_rbv below) -- `sf` is that copy and `origin` the offset of node (0,0,0) in it
template <int NST, bool VAR, bool SPLIT = false>
__global__ void __launch_bounds__(256)
k_stencil_color(const GridArgs g, const __grid_constant__ StencilArgs st, const double *__restrict__ sf, long long sfvol,
                double *u, const double *__restrict__ rhs, double w, int ncol, int cx, int cy, int cz, int ex, int ey,
                int ez, int fx, int fy, int fz, long long origin = 0) {
  const long long total = (long long)ex * ey * ez;
  const int xoff = SPLIT ? (int)(origin % g.pitch) : 0, half = g.pitch >> 1;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int i = fx + ncol * (int)(t % ex), j = fy + ncol * (int)((t / ex) % ey), k = fz + ncol * (int)(t / ((long long)ex * ey));
    const long long p = i + (long long)g.pitch * j + g.plane * k;
    // conv() indexes its coefficient pointer with p: shift the pointer so that [p] is the node's place in the copy
    const double *cf = sf;
    if (SPLIT) {
      const int ip = i + xoff;
      cf = sf + (origin - xoff - i) + (ip & 1) * half + (ip >> 1);
    }
    const double acc = conv<NST, VAR>(st, cf, sfvol, u, p);
    const double ww = VAR ? w / cf[st.centre * sfvol + p] : w;
    u[p] = fma(ww, rhs[p] - acc, u[p]);
  }
  (void)cx; (void)cy; (void)cz;
}

// ---- red-black Gauss-Seidel over stencil-field coefficients ------------------------------------------------------------
// One colour of a sweep touches every other node of a row: the NST coefficient loads of a warp use half of every
// 32-byte sector they fetch, and both colour passes fetch all of them -- 2 x 8 NST bytes per node and sweep where
// 8 NST are needed (measured 0.52 of the roofline, 7-point operator at 256^3).  The coefficients are static, so the
// operator keeps a second copy with every row split into its even and its odd elements
//   split(row, i') = row + (i' & 1) * pitch / 2 + (i' >> 1),      i' = index inside the allocated row
// and a colour pass reads consecutive doubles: each coefficient is fetched once per sweep.  Same values, same fold
// order as k_stencil<M_RB> => bit-identical.  Cost: the coefficient memory twice (EXAMG_RB_COMPACT=0 turns it off).
__global__ void __launch_bounds__(256) k_split_rows(const double *__restrict__ src, double *__restrict__ dst, int pitch,
                                                    long long total) {
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int ip = (int)(t % pitch);
    dst[t - ip + (ip & 1) * (pitch >> 1) + (ip >> 1)] = src[t];
  }
}

template <int NST>
__global__ void __launch_bounds__(256)
k_rbv(const GridArgs g, const __grid_constant__ StencilArgs st, const double *__restrict__ cs, long long sfvol,
      long long origin, double *u, const double *__restrict__ rhs, double w, int color, int kz) {
  const int nst = NST > 0 ? NST : st.nst;
  const int tx = blockIdx.x * blockDim.x + threadIdx.x;
  const int z0 = g.lo[2] + blockIdx.z * kz;
  const int z1 = min(z0 + kz, g.hi[2]);
  const int xoff = (int)(origin % g.pitch), half = g.pitch >> 1;
  for (int j = g.lo[1] + blockIdx.y * blockDim.y + threadIdx.y; j < g.hi[1]; j += gridDim.y * blockDim.y) {
    for (int k = z0; k < z1; ++k) {
      const int i = g.lo[0] + ((g.lo[0] + j + k + g.par + color) & 1) + 2 * tx;
      if (i < g.hi[0]) {
        const long long row = (long long)g.pitch * j + g.plane * k;
        const long long p = i + row;
        const int ip = i + xoff;
        const double *cp = cs + (origin - xoff + row) + (ip & 1) * half + (ip >> 1);
        double acc = cp[0] * u[p + st.lin[0]];
#pragma unroll
        for (int q = 1; q < nst; ++q) acc = fma(cp[q * sfvol], u[p + st.lin[q]], acc);
        const double ww = w / cp[st.centre * sfvol];
        u[p] = fma(ww, rhs[p] - acc, u[p]);
      }
    }
  }
}

// merges the per-block partials in index order; one block
__global__ void __launch_bounds__(1024) k_final_sum(const dd *__restrict__ partials, int n, dd *out) {
  dd v = {0.0, 0.0};
  for (int i = threadIdx.x; i < n; i += blockDim.x) dd_merge(v, partials[i]);
  v = dd_block_reduce(v);
  if (threadIdx.x == 0) *out = v;
}

__global__ void __launch_bounds__(1024) k_final_max(const dd *__restrict__ partials, int n, dd *out) {
  double v = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) v = fmax(v, partials[i].hi);
  v = max_block_reduce(v);
  if (threadIdx.x == 0) { out->hi = v; out->lo = 0.0; }
}

enum { R_DOT = 0, R_MAXDIFF = 1 };
template <int KIND>
__global__ void __launch_bounds__(256)
k_reduce(const GridArgs g, const double *__restrict__ a, const double *__restrict__ b, int kz,
         dd *__restrict__ partials) {
  const int i = (KIND == R_DOT ? max(g.lo[0], g.rlo[0]) : g.lo[0]) + blockIdx.x * blockDim.x + threadIdx.x;
  const int zlo = KIND == R_DOT ? max(g.lo[2], g.rlo[2]) : g.lo[2];
  const int ylo = KIND == R_DOT ? max(g.lo[1], g.rlo[1]) : g.lo[1];
  const int z0 = zlo + blockIdx.z * kz;
  const int z1 = min(z0 + kz, g.hi[2]);
  dd part = {0.0, 0.0};
  double m = 0.0;
  for (int j = ylo + blockIdx.y * blockDim.y + threadIdx.y; j < g.hi[1]; j += gridDim.y * blockDim.y)
    for (int k = z0; k < z1; ++k)
      if (i < g.hi[0]) {
        const long long p = i + (long long)g.pitch * j + g.plane * k;
        if (KIND == R_DOT) dd_add(part, a[p] * b[p]);
        else m = fmax(m, fabs(a[p] - b[p]));
      }
  const int bid = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  if (KIND == R_DOT) {
    part = dd_block_reduce(part);
    if (threadIdx.x == 0 && threadIdx.y == 0) partials[bid] = part;
  } else {
    m = max_block_reduce(m);
    if (threadIdx.x == 0 && threadIdx.y == 0) { partials[bid].hi = m; partials[bid].lo = 0.0; }
  }
}

template <int OP>
__global__ void __launch_bounds__(256)
k_stream(const GridArgs g, double alpha, const double *__restrict__ x, double *__restrict__ y, int kz) {
  const int i = g.lo[0] + blockIdx.x * blockDim.x + threadIdx.x;
  const int z0 = g.lo[2] + blockIdx.z * kz;
  const int z1 = min(z0 + kz, g.hi[2]);
  if (i >= g.hi[0]) return;
  for (int j = g.lo[1] + blockIdx.y * blockDim.y + threadIdx.y; j < g.hi[1]; j += gridDim.y * blockDim.y)
    for (int k = z0; k < z1; ++k) {
      const long long p = i + (long long)g.pitch * j + g.plane * k;
      if (OP == STREAM_SET) y[p] = alpha;
      else if (OP == STREAM_COPY) y[p] = x[p];
      else if (OP == STREAM_AXPY) y[p] = fma(alpha, x[p], y[p]);
      else y[p] = fma(alpha, y[p], x[p]);
    }
}

// ---- Dirichlet boundary update -------------------------------------------------
struct BcProg {
  int n;
  int op[64];
  double val[64];
};

// comp0 / vol: component 0 of the field being written at this node and the component stride (EXAMG_OP_COMP)
// fsrc / p: the source fields of examg_loop_eval at this node (EXAMG_OP_FIELD)
__device__ __forceinline__ double bc_eval(const BcProg &pr, double x, double y, double z,
                                          const double *comp0 = nullptr, long long vol = 0,
                                          const double *const *fsrc = nullptr, long long p = 0) {
  double stk[16];
  int sp = 0;
  for (int q = 0; q < pr.n; ++q) {
    switch (pr.op[q]) {
      case EXAMG_OP_CONST: stk[sp++] = pr.val[q]; break;
      case EXAMG_OP_COMP: stk[sp++] = comp0[(long long)pr.val[q] * vol]; break;
      case EXAMG_OP_FIELD: stk[sp++] = fsrc[(int)pr.val[q]][p]; break;
      case EXAMG_OP_EXP: stk[sp - 1] = exp(stk[sp - 1]); break;
      case EXAMG_OP_SIN: stk[sp - 1] = sin(stk[sp - 1]); break;
      case EXAMG_OP_COS: stk[sp - 1] = cos(stk[sp - 1]); break;
      case EXAMG_OP_SINH: stk[sp - 1] = sinh(stk[sp - 1]); break;
      case EXAMG_OP_SQRT: stk[sp - 1] = sqrt(stk[sp - 1]); break;
      case EXAMG_OP_X: stk[sp++] = x; break;
      case EXAMG_OP_Y: stk[sp++] = y; break;
      case EXAMG_OP_Z: stk[sp++] = z; break;
      case EXAMG_OP_ADD: --sp; stk[sp - 1] = stk[sp - 1] + stk[sp]; break;
      case EXAMG_OP_SUB: --sp; stk[sp - 1] = stk[sp - 1] - stk[sp]; break;
      case EXAMG_OP_MUL: --sp; stk[sp - 1] = stk[sp - 1] * stk[sp]; break;
      case EXAMG_OP_DIV: --sp; stk[sp - 1] = stk[sp - 1] / stk[sp]; break;
      case EXAMG_OP_NEG: stk[sp - 1] = -stk[sp - 1]; break;
    }
  }
  return stk[0];
}

struct BcArgs {
  int n[3];
  int goff[3];
  double h[3], x0[3];
  int pitch;
  long long plane;
  int face_mask;  // bit 2*d+s set: face on side s of dim d is a physical boundary
  int dim;
};

// blockIdx.y = face id (2*d + side).  vf_boundaryPosition = node position
// i*h + lo (C/s/e/grid/ir/IR_VF_NodePosition.scala:109-111).
__global__ void __launch_bounds__(256)
k_dirichlet(const BcArgs a, const __grid_constant__ BcProg pr, int use_prog, double value, double *f) {
  const int face = blockIdx.y;
  if (!((a.face_mask >> face) & 1)) return;
  const int d = face >> 1, side = face & 1;
  const int da = d == 0 ? 1 : 0;             // first in-face dim
  const int db = d == 2 ? 1 : 2;             // second in-face dim (extent 1 in 2D)
  const int na = a.n[da], nb = a.n[db];
  const long long total = (long long)na * nb;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    int idx[3];
    idx[d] = side ? a.n[d] - 1 : 0;
    idx[da] = (int)(t % na);
    idx[db] = (int)(t / na);
    double v = value;
    if (use_prog) {
      // x = i*h + domain_lo
      const double x = (double)(a.goff[0] + idx[0]) * a.h[0] + a.x0[0];
      const double y = (double)(a.goff[1] + idx[1]) * a.h[1] + a.x0[1];
      const double z = a.dim == 3 ? (double)(a.goff[2] + idx[2]) * a.h[2] + a.x0[2] : 0.0;
      v = bc_eval(pr, x, y, z);
    }
    f[idx[0] + (long long)a.pitch * idx[1] + a.plane * idx[2]] = v;
  }
}

// `apply bc` on a cell-localised field: blockIdx.y = face (2*d + side), one thread per ghost cell of the face.
// Dirichlet (order 1): ghost = 2 * value - adjacent inner cell, value constant or a program over vf_boundaryCoord
// (the face position in the face's dimension, cell centres = node position + 0.5 * width in the others;
// C/s/e/boundary/ir/IR_DirichletBC.scala:42-50); Neumann(1): ghost = adjacent inner cell (IR_NeumannBC.scala:70-80).
__global__ void __launch_bounds__(256)
k_cell_bc(const BcArgs a, const __grid_constant__ BcProg pr, int kind, double value, double *f) {
  const int face = blockIdx.y;
  if (!((a.face_mask >> face) & 1)) return;
  const int d = face >> 1, side = face & 1;
  const int da = d == 0 ? 1 : 0, db = d == 2 ? 1 : 2;
  const int na = a.n[da] - 1, nb = db < a.dim ? a.n[db] - 1 : 1;  // cells
  const long long total = (long long)na * nb;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    int idx[3] = {0, 0, 0}, in[3];
    idx[da] = (int)(t % na);
    if (db < a.dim) idx[db] = (int)(t / na);
    idx[d] = side ? a.n[d] - 1 : -1;
    in[0] = idx[0]; in[1] = idx[1]; in[2] = idx[2];
    in[d] = side ? a.n[d] - 2 : 0;
    const double inner = f[in[0] + (long long)a.pitch * in[1] + a.plane * in[2]];
    double v;
    if (kind == EXAMG_BC_NEUMANN) {
      v = inner;
    } else {
      double g = value;
      if (kind == EXAMG_BC_DIRICHLET_EXPR) {
        double pos[3] = {0.0, 0.0, 0.0};
        for (int w = 0; w < a.dim; ++w) {
          if (w == d) pos[w] = (double)(a.goff[w] + (side ? a.n[w] - 1 : 0)) * a.h[w] + a.x0[w];
          else pos[w] = ((double)(a.goff[w] + idx[w]) * a.h[w] + a.x0[w]) + 0.5 * a.h[w];
        }
        g = bc_eval(pr, pos[0], pos[1], pos[2]);
      }
      v = (2.0 * g) - inner;
    }
    f[idx[0] + (long long)a.pitch * idx[1] + a.plane * idx[2]] = v;
  }
}

// loop over F { F[comp] = <expr>(vf_nodePosition) } on the local nodes and one ghost layer
__global__ void __launch_bounds__(256)
k_field_eval(const BcArgs a, const __grid_constant__ BcProg pr, double *f0, long long vol, int comp) {
  const int ex = a.n[0] + 2, ey = a.n[1] + 2, ez = a.dim == 3 ? a.n[2] + 2 : 1;
  const long long total = (long long)ex * ey * ez;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(t % ex) - 1, j = (int)((t / ex) % ey) - 1, k = a.dim == 3 ? (int)(t / ((long long)ex * ey)) - 1 : 0;
    const double x = (double)(a.goff[0] + i) * a.h[0] + a.x0[0];
    const double y = (double)(a.goff[1] + j) * a.h[1] + a.x0[1];
    const double z = a.dim == 3 ? (double)(a.goff[2] + k) * a.h[2] + a.x0[2] : 0.0;
    const long long p = i + (long long)a.pitch * j + a.plane * k;
    f0[(long long)comp * vol + p] = bc_eval(pr, x, y, z, f0 + p, vol);
  }
}

// loop over F { F = <expr>(vf_nodePosition, G_0 .. G_7 at the same node) } on F's loop points (examg_loop_eval)
struct EvalSrc {
  const double *p[8];
};

__global__ void __launch_bounds__(256)
k_loop_eval(const GridArgs g, const BcArgs a, const __grid_constant__ BcProg pr, const EvalSrc src, double *dst, int kz) {
  const int i = g.lo[0] + blockIdx.x * blockDim.x + threadIdx.x;
  const int z0 = g.lo[2] + blockIdx.z * kz;
  const int z1 = min(z0 + kz, g.hi[2]);
  if (i >= g.hi[0]) return;
  const double x = (double)(a.goff[0] + i) * a.h[0] + a.x0[0];
  for (int j = g.lo[1] + blockIdx.y * blockDim.y + threadIdx.y; j < g.hi[1]; j += gridDim.y * blockDim.y) {
    const double y = (double)(a.goff[1] + j) * a.h[1] + a.x0[1];
    for (int k = z0; k < z1; ++k) {
      const double z = a.dim == 3 ? (double)(a.goff[2] + k) * a.h[2] + a.x0[2] : 0.0;
      const long long p = i + (long long)g.pitch * j + g.plane * k;
      dst[p] = bc_eval(pr, x, y, z, nullptr, 0, src.p, p);
    }
  }
}

// ---- launchers -----------------------------------------------------------------

template <int MODE>
static int launch_stencil(examg_ctx *c, const Stencil &s, const double *u, const double *rhs, double *out,
                          double omega, int color, int red_slot, int extend = 0) {
  const LevelGeom &g = *geom(c, s.level);
  GridArgs ga = grid_args_loc(c, g);
  // `extend` layers of the neighbour blocks' nodes are computed redundantly (saves their exchange)
  for (int d = 0; d < c->cfg.dim && extend > 0; ++d) {
    if (c->nb[d][0]) ga.lo[d] -= extend;
    if (c->nb[d][1]) ga.hi[d] += extend;
  }
  const StencilArgs sa = stencil_args(g, s);
  const bool var = s.coeff_field >= 0;
  const double *sf = nullptr;
  if (var) sf = field_ptr(c, &c->fields[s.coeff_field], 0, 0);
  int ex = ga.hi[0] - ga.lo[0], ey = ga.hi[1] - ga.lo[1], ez = ga.hi[2] - ga.lo[2];
  if (ex <= 0 || ey <= 0 || ez <= 0) {
    if (MODE == M_RESNORM || MODE == M_NORMONLY) {
      EXAMG_CUDA(cudaMemsetAsync(c->red_out + red_slot, 0, sizeof(dd), c->stream));
    }
    return EXAMG_OK;
  }
  if (MODE == M_RB) ex = (ex + 1) / 2;
  const Shape sh = make_shape(ex, ey, ez);
  double w = omega;
  if (!var && (MODE == M_JACOBI || MODE == M_RB)) w = omega / s.coef[s.centre];
#define EXAMG_GO(NST, VAR)                                                                         \
  k_stencil<MODE, NST, VAR><<<sh.grid, sh.block, 0, c->stream>>>(ga, sa, sf, g.vol, u, rhs, out, w, \
                                                                  color, sh.kz, c->partials)
  if (var) {
    switch (s.nst) {
      case 5: EXAMG_GO(5, true); break;
      case 7: EXAMG_GO(7, true); break;
      case 9: EXAMG_GO(9, true); break;
      case 27: EXAMG_GO(27, true); break;
      default: EXAMG_GO(0, true); break;
    }
  } else {
    switch (s.nst) {
      case 5: EXAMG_GO(5, false); break;
      case 7: EXAMG_GO(7, false); break;
      case 9: EXAMG_GO(9, false); break;
      case 27: EXAMG_GO(27, false); break;
      default: EXAMG_GO(0, false); break;
    }
  }
#undef EXAMG_GO
  EXAMG_LAUNCH_CHECK(c);
  if (MODE == M_RESNORM || MODE == M_NORMONLY) {
    k_final_sum<<<1, 1024, 0, c->stream>>>(c->partials, sh.nblocks, c->red_out + red_slot);
    EXAMG_LAUNCH_CHECK(c);
  }
  return EXAMG_OK;
}

static bool rb_compact_wanted(examg_ctx *c, const Stencil &s);

// colour (cx, cy, cz) of an ncol^dim-way colouring; the colour of a point is its block-local index modulo ncol
int launch_color_generic(examg_ctx *c, const Stencil &s, double omega, double *u, const double *rhs, int ncol, int cx,
                         int cy, int cz) {
  const LevelGeom &g = *geom(c, s.level);
  const GridArgs ga = grid_args_loc(c, g);
  const StencilArgs sa = stencil_args(g, s);
  const bool var = s.coeff_field >= 0;
  const double *sf = var ? field_ptr(c, &c->fields[s.coeff_field], 0, 0) : nullptr;
  const int col[3] = {cx, cy, cz};
  int first[3], ext[3];
  for (int d = 0; d < 3; ++d) {
    if (d >= c->cfg.dim) { first[d] = 0; ext[d] = 1; continue; }
    first[d] = ga.lo[d] + ((col[d] - ga.lo[d]) % ncol + ncol) % ncol;
    ext[d] = first[d] < ga.hi[d] ? (ga.hi[d] - first[d] + ncol - 1) / ncol : 0;
  }
  const long long total = (long long)ext[0] * ext[1] * ext[2];
  if (total <= 0) return EXAMG_OK;
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  const double w = var ? omega : omega / s.coef[s.centre];
  const int czz = c->cfg.dim == 3 ? cz : -1;
  // two colours per dimension on a stencil field: every other node of a row -> the split-row copy of the coefficients
  if (var && ncol == 2 && rb_compact_wanted(c, s)) {
    int rc = rb_compact_prepare(c, const_cast<Stencil &>(s));
    if (rc) return rc;
    if (s.rbc && s.rbc_gen == c->fields[s.coeff_field].eval_gen) {
#define EXAMG_GOS(NST)                                                                                                  \
  k_stencil_color<NST, true, true><<<(int)blocks, 256, 0, c->stream>>>(ga, sa, s.rbc, g.vol, u, rhs, w, ncol, cx, cy, czz, \
                                                                        ext[0], ext[1], ext[2], first[0], first[1],      \
                                                                        first[2], g.origin)
      switch (s.nst) {
        case 9: EXAMG_GOS(9); break;
        case 27: EXAMG_GOS(27); break;
        default: EXAMG_GOS(0); break;
      }
#undef EXAMG_GOS
      EXAMG_LAUNCH_CHECK(c);
      return EXAMG_OK;
    }
  }
#define EXAMG_GOC(NST, VAR)                                                                                         \
  k_stencil_color<NST, VAR><<<(int)blocks, 256, 0, c->stream>>>(ga, sa, sf, g.vol, u, rhs, w, ncol, cx, cy, czz, ext[0], \
                                                                 ext[1], ext[2], first[0], first[1], first[2])
  if (var) {
    switch (s.nst) {
      case 9: EXAMG_GOC(9, true); break;
      case 27: EXAMG_GOC(27, true); break;
      default: EXAMG_GOC(0, true); break;
    }
  } else {
    switch (s.nst) {
      case 9: EXAMG_GOC(9, false); break;
      case 27: EXAMG_GOC(27, false); break;
      default: EXAMG_GOC(0, false); break;
    }
  }
#undef EXAMG_GOC
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

int launch_apply_generic(examg_ctx *c, const Stencil &s, const double *u, double *out) {
  return launch_stencil<M_APPLY>(c, s, u, nullptr, out, 0.0, 0, -1);
}
int launch_residual_generic(examg_ctx *c, const Stencil &s, const double *u, const double *rhs, double *res,
                            int red_slot, int extend) {
  if (red_slot >= 0 && !res) return launch_stencil<M_NORMONLY>(c, s, u, rhs, res, 0.0, 0, red_slot, extend);
  if (red_slot >= 0) return launch_stencil<M_RESNORM>(c, s, u, rhs, res, 0.0, 0, red_slot, extend);
  return launch_stencil<M_RESIDUAL>(c, s, u, rhs, res, 0.0, 0, -1, extend);
}
int launch_jacobi_generic(examg_ctx *c, const Stencil &s, double omega, const double *u, const double *rhs,
                          double *unext) {
  return launch_stencil<M_JACOBI>(c, s, u, rhs, unext, omega, 0, -1);
}
static bool rb_compact_wanted(examg_ctx *c, const Stencil &s) {
  static int mode = -1;
  if (mode < 0) {
    const char *e = getenv("EXAMG_RB_COMPACT");
    mode = e ? atoi(e) : 1;
  }
  if (mode <= 0 || s.coeff_field < 0) return false;
  const LevelGeom &g = *geom(c, s.level);
  return (long long)g.n[0] * g.n[1] * g.n[2] >= (1 << 15) && g.pitch % 2 == 0;
}

int rb_compact_prepare(examg_ctx *c, Stencil &s) {
  if (!rb_compact_wanted(c, s)) return EXAMG_OK;
  const Field &cf = c->fields[s.coeff_field];
  if (s.rbc && s.rbc_gen == cf.eval_gen) return EXAMG_OK;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(c->stream, &cap);
  if (cap != cudaStreamCaptureStatusNone) return EXAMG_OK;  // not now: the strided reads serve until the next chance
  const LevelGeom &g = *geom(c, s.level);
  const long long total = (long long)cf.ncomp * g.vol;
  if (!s.rbc) {
    s.rbc = dev_alloc(c, (size_t)total * sizeof(double));
    if (!s.rbc) {  // no room for the copy: keep the strided reads
      s.rbc_gen = -1;
      return EXAMG_OK;
    }
  }
  k_split_rows<<<148 * 8, 256, 0, c->stream>>>(cf.slot[0], s.rbc, g.pitch, total);
  EXAMG_LAUNCH_CHECK(c);
  s.rbc_gen = cf.eval_gen;
  return EXAMG_OK;
}

int launch_rb_generic(examg_ctx *c, const Stencil &s, double omega, double *u, const double *rhs, int color) {
  if (s.coeff_field >= 0 && rb_compact_wanted(c, s)) {
    Stencil &ms = const_cast<Stencil &>(s);  // the copy is a cache
    int rc = rb_compact_prepare(c, ms);
    if (rc) return rc;
    if (s.rbc && s.rbc_gen == c->fields[s.coeff_field].eval_gen) {
      const LevelGeom &g = *geom(c, s.level);
      const GridArgs ga = grid_args_loc(c, g);
      int ex = ga.hi[0] - ga.lo[0], ey = ga.hi[1] - ga.lo[1], ez = ga.hi[2] - ga.lo[2];
      if (ex <= 0 || ey <= 0 || ez <= 0) return EXAMG_OK;
      const Shape sh = make_shape((ex + 1) / 2, ey, ez);
      const StencilArgs sa = stencil_args(g, s);
      switch (s.nst) {
        case 5: k_rbv<5><<<sh.grid, sh.block, 0, c->stream>>>(ga, sa, s.rbc, g.vol, g.origin, u, rhs, omega, color, sh.kz); break;
        case 7: k_rbv<7><<<sh.grid, sh.block, 0, c->stream>>>(ga, sa, s.rbc, g.vol, g.origin, u, rhs, omega, color, sh.kz); break;
        default: k_rbv<0><<<sh.grid, sh.block, 0, c->stream>>>(ga, sa, s.rbc, g.vol, g.origin, u, rhs, omega, color, sh.kz); break;
      }
      EXAMG_LAUNCH_CHECK(c);
      return EXAMG_OK;
    }
  }
  return launch_stencil<M_RB>(c, s, u, rhs, u, omega, color, -1);
}

static int launch_reduce(examg_ctx *c, int level, const double *a, const double *b, int red_slot, int kind) {
  const LevelGeom &g = *geom(c, level);
  const GridArgs ga = grid_args_loc(c, g);
  int xlo = ga.lo[0], ylo = ga.lo[1], zlo = ga.lo[2];
  if (kind == R_DOT) {
    xlo = ga.lo[0] > ga.rlo[0] ? ga.lo[0] : ga.rlo[0];
    ylo = ga.lo[1] > ga.rlo[1] ? ga.lo[1] : ga.rlo[1];
    zlo = ga.lo[2] > ga.rlo[2] ? ga.lo[2] : ga.rlo[2];
  }
  const int ex = ga.hi[0] - xlo, ey = ga.hi[1] - ylo, ez = ga.hi[2] - zlo;
  if (ex <= 0 || ey <= 0 || ez <= 0) {
    EXAMG_CUDA(cudaMemsetAsync(c->red_out + red_slot, 0, sizeof(dd), c->stream));
    return EXAMG_OK;
  }
  const Shape sh = make_shape(ex, ey, ez);
  if (kind == R_DOT) {
    k_reduce<R_DOT><<<sh.grid, sh.block, 0, c->stream>>>(ga, a, b, sh.kz, c->partials);
    EXAMG_LAUNCH_CHECK(c);
    k_final_sum<<<1, 1024, 0, c->stream>>>(c->partials, sh.nblocks, c->red_out + red_slot);
  } else {
    k_reduce<R_MAXDIFF><<<sh.grid, sh.block, 0, c->stream>>>(ga, a, b, sh.kz, c->partials);
    EXAMG_LAUNCH_CHECK(c);
    k_final_max<<<1, 1024, 0, c->stream>>>(c->partials, sh.nblocks, c->red_out + red_slot);
  }
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

int launch_dot(examg_ctx *c, int level, const double *a, const double *b, int red_slot) {
  return launch_reduce(c, level, a, b, red_slot, R_DOT);
}
int launch_maxdiff(examg_ctx *c, int level, const double *a, const double *b, int red_slot) {
  return launch_reduce(c, level, a, b, red_slot, R_MAXDIFF);
}

int launch_stream_op(examg_ctx *c, int level, int op, double alpha, const double *x, double *y) {
  const LevelGeom &g = *geom(c, level);
  const GridArgs ga = grid_args_loc(c, g);
  const int ex = ga.hi[0] - ga.lo[0], ey = ga.hi[1] - ga.lo[1], ez = ga.hi[2] - ga.lo[2];
  if (ex <= 0 || ey <= 0 || ez <= 0) return EXAMG_OK;
  const Shape sh = make_shape(ex, ey, ez);
  switch (op) {
    case STREAM_SET: k_stream<STREAM_SET><<<sh.grid, sh.block, 0, c->stream>>>(ga, alpha, x, y, sh.kz); break;
    case STREAM_COPY: k_stream<STREAM_COPY><<<sh.grid, sh.block, 0, c->stream>>>(ga, alpha, x, y, sh.kz); break;
    case STREAM_AXPY: k_stream<STREAM_AXPY><<<sh.grid, sh.block, 0, c->stream>>>(ga, alpha, x, y, sh.kz); break;
    default: k_stream<STREAM_XPAY><<<sh.grid, sh.block, 0, c->stream>>>(ga, alpha, x, y, sh.kz); break;
  }
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

int launch_field_eval(examg_ctx *c, const Field &f, int slot, int comp, const examg_bc_op *prog, int n) {
  const LevelGeom &g = *geom(c, f.level);
  BcArgs a;
  for (int d = 0; d < 3; ++d) {
    a.n[d] = g.n[d];
    a.goff[d] = g.goff[d];
    a.h[d] = g.h[d];
    a.x0[d] = d < c->cfg.dim ? c->cfg.domain_lo[d] : 0.0;
  }
  a.pitch = g.pitch;
  a.plane = g.plane;
  a.face_mask = 0;
  a.dim = c->cfg.dim;
  BcProg pr;
  pr.n = n;
  for (int q = 0; q < n; ++q) { pr.op[q] = prog[q].op; pr.val[q] = prog[q].value; }
  const long long total = (long long)(g.n[0] + 2) * (g.n[1] + 2) * (c->cfg.dim == 3 ? g.n[2] + 2 : 1);
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  k_field_eval<<<(int)blocks, 256, 0, c->stream>>>(a, pr, field_ptr(c, &f, slot, 0), g.vol, comp);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

int launch_loop_eval(examg_ctx *c, int level, const double *const *src, int nsrc, double *dst, const examg_bc_op *prog,
                     int n) {
  const LevelGeom &g = *geom(c, level);
  const GridArgs ga = grid_args_loc(c, g);
  const int ex = ga.hi[0] - ga.lo[0], ey = ga.hi[1] - ga.lo[1], ez = ga.hi[2] - ga.lo[2];
  if (ex <= 0 || ey <= 0 || ez <= 0) return EXAMG_OK;
  BcArgs a;
  for (int d = 0; d < 3; ++d) {
    a.n[d] = g.n[d];
    a.goff[d] = g.goff[d];
    a.h[d] = g.h[d];
    a.x0[d] = d < c->cfg.dim ? c->cfg.domain_lo[d] : 0.0;
  }
  a.pitch = g.pitch;
  a.plane = g.plane;
  a.face_mask = 0;
  a.dim = c->cfg.dim;
  BcProg pr;
  pr.n = n;
  for (int q = 0; q < n; ++q) { pr.op[q] = prog[q].op; pr.val[q] = prog[q].value; }
  EvalSrc es;
  for (int k = 0; k < 8; ++k) es.p[k] = k < nsrc ? src[k] : nullptr;
  const Shape sh = make_shape(ex, ey, ez);
  k_loop_eval<<<sh.grid, sh.block, 0, c->stream>>>(ga, a, pr, es, dst, sh.kz);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

int launch_bc_cell_generic(examg_ctx *c, const Field &f, int slot) {
  const LevelGeom &g = *geom(c, f.level);
  BcArgs a;
  int mask = 0;
  for (int d = 0; d < 3; ++d) {
    a.n[d] = g.n[d];
    a.goff[d] = g.goff[d];
    a.h[d] = g.h[d];
    a.x0[d] = d < c->cfg.dim ? c->cfg.domain_lo[d] : 0.0;
    if (d < c->cfg.dim) {
      if (!c->nb[d][0]) mask |= 1 << (2 * d);
      if (!c->nb[d][1]) mask |= 1 << (2 * d + 1);
    }
  }
  a.pitch = g.pitch;
  a.plane = g.plane;
  a.face_mask = mask;
  a.dim = c->cfg.dim;
  long long maxface = 1;
  for (int d = 0; d < c->cfg.dim; ++d) {
    long long fsz = 1;
    for (int e = 0; e < c->cfg.dim; ++e)
      if (e != d) fsz *= g.n[e] - 1;
    if (fsz > maxface) maxface = fsz;
  }
  int gx = (int)((maxface + 255) / 256);
  if (gx > 1024) gx = 1024;
  dim3 grid(gx, 2 * c->cfg.dim, 1);
  // faces with the field's own condition in one launch, faces with a condition of their own one launch each
  int own = 0;
  for (int k = 0; k < 2 * c->cfg.dim; ++k)
    if (f.bc_face_kind[k] >= 0) own |= 1 << k;
  for (int pass = -1; pass < 2 * c->cfg.dim; ++pass) {
    int kind = f.bc_kind, m = mask & ~own;
    double value = f.bc_value;
    const std::vector<examg_bc_op> *prog = &f.bc_prog;
    if (pass >= 0) {
      if (!((own >> pass) & 1)) continue;
      kind = f.bc_face_kind[pass];
      value = f.bc_face_value[pass];
      prog = &f.bc_face_prog[pass];
      m = mask & (1 << pass);
    }
    if (!m || kind == EXAMG_BC_NONE || kind == EXAMG_BC_DIRICHLET_HOST) continue;
    BcProg pr;
    pr.n = 0;
    if (kind == EXAMG_BC_DIRICHLET_EXPR) {
      pr.n = (int)prog->size();
      for (int q = 0; q < pr.n; ++q) { pr.op[q] = (*prog)[q].op; pr.val[q] = (*prog)[q].value; }
    }
    a.face_mask = m;
    for (int comp = 0; comp < f.ncomp; ++comp) {
      k_cell_bc<<<grid, 256, 0, c->stream>>>(a, pr, kind, value, field_ptr(c, &f, slot, comp));
      EXAMG_LAUNCH_CHECK(c);
    }
  }
  return EXAMG_OK;
}

int launch_bc(examg_ctx *c, const Field &f, int slot) {
  const LevelGeom &g = *geom(c, f.level);
  BcArgs a;
  int mask = 0;
  for (int d = 0; d < 3; ++d) {
    a.n[d] = g.n[d];
    a.goff[d] = g.goff[d];
    a.h[d] = g.h[d];
    a.x0[d] = d < c->cfg.dim ? c->cfg.domain_lo[d] : 0.0;
    if (d < c->cfg.dim) {
      if (!c->nb[d][0]) mask |= 1 << (2 * d);
      if (!c->nb[d][1]) mask |= 1 << (2 * d + 1);
    }
  }
  if (!mask) return EXAMG_OK;
  a.pitch = g.pitch;
  a.plane = g.plane;
  a.face_mask = mask;
  a.dim = c->cfg.dim;
  BcProg pr;
  pr.n = 0;
  const int use_prog = f.bc_kind == EXAMG_BC_DIRICHLET_EXPR;
  if (use_prog) {
    pr.n = (int)f.bc_prog.size();
    for (int q = 0; q < pr.n; ++q) { pr.op[q] = f.bc_prog[q].op; pr.val[q] = f.bc_prog[q].value; }
  }
  long long maxface = 1;
  for (int d = 0; d < c->cfg.dim; ++d) {
    long long fsz = 1;
    for (int e = 0; e < c->cfg.dim; ++e)
      if (e != d) fsz *= g.n[e];
    if (fsz > maxface) maxface = fsz;
  }
  int gx = (int)((maxface + 255) / 256);
  if (gx > 1024) gx = 1024;
  dim3 grid(gx, 2 * c->cfg.dim, 1);
  for (int comp = 0; comp < f.ncomp; ++comp) {
    k_dirichlet<<<grid, 256, 0, c->stream>>>(a, pr, use_prog, f.bc_value,
                                             field_ptr(c, &f, slot, comp));
    EXAMG_LAUNCH_CHECK(c);
  }
  return EXAMG_OK;
}

}  // namespace examg
