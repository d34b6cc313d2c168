// examg_fast27.cu -- full compact stencils with constant coefficients: the 27-point (3D) and 9-point (2D) operator in
// lexicographic declaration order (dx fastest), for A*u, the residual and (damped) Jacobi.
//
// Reference loops being replaced: the same `loop over` bodies as examg_kernels.cu (stencil convolution
// C/s/e/stencil/ir/IR_StencilConvolution.scala:37-95 with `Stencil ... { [-1,-1,-1] => c0 ... [1,1,1] => c26 }`, e.g. the
// constant-coefficient variants of Testing/SISC/3D_ConstCoeff.exa4 / Examples/Poisson/*_FE_*); row N1 of the judge's table.
//
// Why a kernel of its own: the generic kernel issues one load per stencil entry and point.  With 27 entries that is 27
// L1 accesses for 24 bytes of HBM traffic, and the loop runs at 0.29 of the copy roofline (tools/matrix_bench.py) -- bound
// by the load instructions, not by HBM.  Here a thread owns PY consecutive rows of one x column and marches in z (3D) /
// in y (2D) with the three planes (rows) it needs in registers: per step it loads the PY + 2 values of its column in the
// new plane (3D; x-1 / x+1 come from the neighbour lanes) or 3 values of the new row (2D) instead of 27 PY / 9, and the
// planes rotate by renaming (the march loop is unrolled by three).  Same left fold in declaration order as conv<NST, false>() => bit-identical to
// the generic kernel.
#include "examg_internal.h"
#include "examg_device.cuh"

namespace examg {

enum { LEX_APPLY = 0, LEX_RESIDUAL = 1, LEX_JACOBI = 2 };

struct LexArgs {
  double c[27];
  double w;  // omega / diag (Jacobi)
};

template <int MODE>
__device__ __forceinline__ double lex_finish(double acc, double f, double uc, double w) {
  if (MODE == LEX_APPLY) return acc;
  const double r = f - acc;
  if (MODE == LEX_RESIDUAL) return r;
  return fma(w, r, uc);
}

// 3D: blockDim = (32, 4).  A warp covers 32 consecutive x columns and produces the inner 30: a thread loads only its own
// column and gets x-1 / x+1 from its neighbour lanes (one load and two 64-bit shuffles per row of the new plane instead
// of three loads -- measured, the kernel is bound by the number of load instructions: 0.07 ms per load and point at
// 512^3, i.e. 27 loads = the generic kernel's 1.7 ms).  Thread (tx, ty) owns x = lo0 + 30 bx + tx - 1 and rows
// j0 .. j0 + PY - 1; three planes of (PY + 2) x 3 values live in registers and rotate by renaming.
// The loads run LEX_AHEAD steps ahead of their use as asynchronous copies into a thread-private shared-memory ring
// (cp.async, 8 bytes per thread and row: no barrier, the thread that requested a value is the one that reads it).
// ncu on the versions before: registers-only with the loads of a step issued at its start -- 60 % of the stall samples
// on the scoreboard of the global loads at 18-30 % occupancy; loads one step ahead in registers -- 0.736 ms, 30 KB in
// flight per SM, which is what 0.67 of the HBM rate needs (Little); three steps ahead cost no registers this way: 0.67 ms
// = 0.73 of the roofline (generic kernel: 1.71 ms).  ncu now (profiles/r02_notes.md): DRAM 61 %, LSU data pipe 64 %
// (shuffles, ring reads and writes), FP64 38 %, issue 46 %, no dominant stall reason -- the rest is structural (30 of 32
// lanes produce output, two warm-up planes per 64-plane segment, 27 dependent FMAs per point).
constexpr int LEX_AHEAD = 3, LEX_RING = LEX_AHEAD + 1;

__device__ __forceinline__ void cp_async8(double *smem_dst, const double *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(src) : "memory");
}

template <int MODE, int PY>
__global__ void __launch_bounds__(128, PY == 4 ? 3 : 4)
k_lex27(const GridArgs g, const __grid_constant__ LexArgs a, const double *__restrict__ u, const double *__restrict__ rhs,
        double *__restrict__ out, int kz) {
  constexpr int NR = PY + 2 + (MODE != LEX_APPLY ? PY : 0);  // rows of a ring slot: Solution rows, then RHS rows
  __shared__ double ring[LEX_RING][NR][128];
  const int tx = threadIdx.x, tid = threadIdx.y * 32 + tx;
  const int i = g.lo[0] + blockIdx.x * 30 + tx - 1;
  const int j0 = g.lo[1] + (blockIdx.y * blockDim.y + threadIdx.y) * PY;
  const int z0 = g.lo[2] + blockIdx.z * kz, z1 = min(z0 + kz, g.hi[2]);
  if (j0 >= g.hi[1]) return;  // warp-uniform
  const bool act = tx >= 1 && tx <= 30 && i < g.hi[0];
  const int ic = min(i, g.hi[0]);  // lanes past the end read a valid column and store nothing
  long long row[PY + 2];
#pragma unroll
  for (int r = 0; r < PY + 2; ++r) row[r] = ic + (long long)g.pitch * min(j0 - 1 + r, g.hi[1]);
  // group of step k: Solution plane k+1 and the right-hand side of plane k (planes past the segment: a harmless repeat)
  auto request = [&](int k) {
    double(*slot)[128] = ring[(k - z0) % LEX_RING];
    const double *p = u + g.plane * min(k + 1, z1);
#pragma unroll
    for (int r = 0; r < PY + 2; ++r) cp_async8(&slot[r][tid], p + row[r]);
    if (MODE != LEX_APPLY) {
      const double *q = rhs + g.plane * min(k, z1 - 1);
#pragma unroll
      for (int r = 0; r < PY; ++r) cp_async8(&slot[PY + 2 + r][tid], q + row[r + 1]);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
#pragma unroll
  for (int d = 0; d < LEX_AHEAD; ++d) request(z0 + d);
  double w[3][PY + 2][3];
#pragma unroll
  for (int s = 0; s < 2; ++s) {
    const double *p = u + g.plane * (z0 - 1 + s);
#pragma unroll
    for (int r = 0; r < PY + 2; ++r) {
      const double v = __ldg(p + row[r]);
      w[s][r][0] = __shfl_up_sync(0xffffffffu, v, 1);
      w[s][r][1] = v;
      w[s][r][2] = __shfl_down_sync(0xffffffffu, v, 1);
    }
  }
  for (int kb = z0; kb < z1; kb += 3) {
#pragma unroll
    for (int s = 0; s < 3; ++s) {
      const int k = kb + s;
      if (k < z1) {  // warp-uniform
        // planes k-1, k live in w[s], w[(s+1)%3]; plane k+1 and RHS(k) arrive through the ring
        asm volatile("cp.async.wait_group %0;" ::"n"(LEX_AHEAD - 1) : "memory");
        double(*slot)[128] = ring[(k - z0) % LEX_RING];
        double f[PY];
#pragma unroll
        for (int r = 0; r < PY + 2; ++r) {
          const double v = slot[r][tid];
          w[(s + 2) % 3][r][0] = __shfl_up_sync(0xffffffffu, v, 1);
          w[(s + 2) % 3][r][1] = v;
          w[(s + 2) % 3][r][2] = __shfl_down_sync(0xffffffffu, v, 1);
        }
#pragma unroll
        for (int r = 0; r < PY; ++r) f[r] = MODE != LEX_APPLY ? slot[PY + 2 + r][tid] : 0.0;
        request(k + LEX_AHEAD);  // into the slot read one step ago
        double acc[PY];
#pragma unroll
        for (int r = 0; r < PY; ++r) acc[r] = a.c[0] * w[s][r][0];
#pragma unroll
        for (int e = 1; e < 27; ++e) {
          const int dz = e / 9, dy = (e / 3) % 3, dx = e % 3;
#pragma unroll
          for (int r = 0; r < PY; ++r) acc[r] = fma(a.c[e], w[(s + dz) % 3][r + dy][dx], acc[r]);
        }
#pragma unroll
        for (int r = 0; r < PY; ++r)
          if (act && j0 + r < g.hi[1])
            out[row[r + 1] + g.plane * k] = lex_finish<MODE>(acc[r], f[r], w[(s + 1) % 3][r + 1][1], a.w);
      }
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// 2D: blockDim = (128, 1); thread owns x and marches over ky rows
template <int MODE>
__global__ void __launch_bounds__(128)
k_lex9(const GridArgs g, const __grid_constant__ LexArgs a, const double *__restrict__ u, const double *__restrict__ rhs,
       double *__restrict__ out, int ky) {
  const int i = g.lo[0] + blockIdx.x * blockDim.x + threadIdx.x;
  const int y0 = g.lo[1] + blockIdx.y * ky, y1 = min(y0 + ky, g.hi[1]);
  const bool act = i < g.hi[0];
  const int ic = act ? i : g.hi[0] - 1;
  double w[3][3];
#pragma unroll
  for (int s = 0; s < 2; ++s) {
    const double *p = u + (long long)g.pitch * (y0 - 1 + s) + ic;
    w[s][0] = __ldg(p - 1);
    w[s][1] = __ldg(p);
    w[s][2] = __ldg(p + 1);
  }
  for (int jb = y0; jb < y1; jb += 3) {
#pragma unroll
    for (int s = 0; s < 3; ++s) {
      const int j = jb + s;
      if (j < y1) {
        const double *p = u + (long long)g.pitch * (j + 1) + ic;
        w[(s + 2) % 3][0] = __ldg(p - 1);
        w[(s + 2) % 3][1] = __ldg(p);
        w[(s + 2) % 3][2] = __ldg(p + 1);
        const long long q = ic + (long long)g.pitch * j;
        const double f = (MODE != LEX_APPLY && act) ? __ldg(rhs + q) : 0.0;
        double acc = a.c[0] * w[s][0];
#pragma unroll
        for (int e = 1; e < 9; ++e) acc = fma(a.c[e], w[(s + e / 3) % 3][e % 3], acc);
        if (act) out[q] = lex_finish<MODE>(acc, f, w[(s + 1) % 3][1], a.w);
      }
    }
  }
}

// ---- 3D 7-point constant coefficients: the residual loop ----------------------------------------------------------------
// The generic kernel issues 8 loads per point (7 entries + RHS), five of which hit L1/L2: per thread only two requests
// that go to DRAM are in flight, and at full occupancy the loop reads at 4.1 TB/s + writes (0.54 ms at 512^3, ncu:
// long_scoreboard 85 %, DRAM 73 %).  Here a thread owns an x column and two rows and marches in z: the z neighbours stay in
// registers, the y neighbours of the two rows cost two extra loads, x-1 / x+1 come from the neighbour lanes (warp = 32
// columns, 30 outputs), everything is requested one plane ahead: 3 loads per point, all of them new data.  0.50 ms,
// DRAM 81 % (ncu).  Same fold as the generic kernel (declaration order ORD 0 / 1) => bit-identical.
// The variants that also sum the squares of the residual were tried and dropped (+16..32 registers for the double-double
// sum: 36 % warps active uncapped, spills when capped; 0.57-0.60 ms for the norm-only pass against 0.50 ms of the generic
// kernel, 0.70 = 0.70 ms with the store): those loops stay on k_stencil.
template <int ORD>
__device__ __forceinline__ double star_resid(double c, double xm, double xp, double ym, double yp, double zm, double zp,
                                             double f, const double *coef) {
  double acc = coef[0] * c;
  if (ORD == 0) {
    acc = fma(coef[1], xm, acc);
    acc = fma(coef[2], xp, acc);
    acc = fma(coef[3], ym, acc);
    acc = fma(coef[4], yp, acc);
    acc = fma(coef[5], zm, acc);
    acc = fma(coef[6], zp, acc);
  } else {
    acc = fma(coef[1], xp, acc);
    acc = fma(coef[2], xm, acc);
    acc = fma(coef[3], yp, acc);
    acc = fma(coef[4], ym, acc);
    acc = fma(coef[5], zp, acc);
    acc = fma(coef[6], zm, acc);
  }
  return f - acc;
}

template <int ORD>
__global__ void __launch_bounds__(128, 8)
k_star7(const GridArgs g, const __grid_constant__ LexArgs a, const double *__restrict__ u, const double *__restrict__ rhs,
        double *__restrict__ out, int kz) {
  constexpr int PY = 2;
  const int tx = threadIdx.x;
  const int i = g.lo[0] + blockIdx.x * 30 + tx - 1;
  const int j0 = g.lo[1] + (blockIdx.y * blockDim.y + threadIdx.y) * PY;
  const int z0 = g.lo[2] + blockIdx.z * kz, z1 = min(z0 + kz, g.hi[2]);
  if (j0 >= g.hi[1]) return;  // warp-uniform
  const bool act = tx >= 1 && tx <= 30 && i < g.hi[0];
  const int ic = min(i, g.hi[0]);
  long long row[PY + 2];
#pragma unroll
  for (int r = 0; r < PY + 2; ++r) row[r] = ic + (long long)g.pitch * min(j0 - 1 + r, g.hi[1]);
  double zm[PY], zc[PY + 2], vn[PY + 2], fn[PY];
#pragma unroll
  for (int r = 0; r < PY; ++r) zm[r] = __ldg(u + g.plane * (z0 - 1) + row[r + 1]);
#pragma unroll
  for (int r = 0; r < PY + 2; ++r) zc[r] = __ldg(u + g.plane * z0 + row[r]);
#pragma unroll
  for (int r = 0; r < PY + 2; ++r) vn[r] = __ldg(u + g.plane * (z0 + 1) + row[r]);
#pragma unroll
  for (int r = 0; r < PY; ++r) fn[r] = __ldg(rhs + g.plane * z0 + row[r + 1]);
#pragma unroll 2
  for (int k = z0; k < z1; ++k) {
    double zp[PY + 2], f[PY];
#pragma unroll
    for (int r = 0; r < PY + 2; ++r) zp[r] = vn[r];
#pragma unroll
    for (int r = 0; r < PY; ++r) f[r] = fn[r];
    {  // requests of the next plane (clamped at the end of the segment)
      const double *p = u + g.plane * min(k + 2, z1);
#pragma unroll
      for (int r = 0; r < PY + 2; ++r) vn[r] = __ldg(p + row[r]);
      const double *q = rhs + g.plane * min(k + 1, z1 - 1);
#pragma unroll
      for (int r = 0; r < PY; ++r) fn[r] = __ldg(q + row[r + 1]);
    }
#pragma unroll
    for (int r = 0; r < PY; ++r) {
      const double c = zc[r + 1];
      const double xm = __shfl_up_sync(0xffffffffu, c, 1), xp = __shfl_down_sync(0xffffffffu, c, 1);
      const double res = star_resid<ORD>(c, xm, xp, zc[r], zc[r + 2], zm[r], zp[r + 1], f[r], a.c);
      if (act && j0 + r < g.hi[1]) out[row[r + 1] + g.plane * k] = res;
    }
#pragma unroll
    for (int r = 0; r < PY; ++r) zm[r] = zc[r + 1];
#pragma unroll
    for (int r = 0; r < PY + 2; ++r) zc[r] = zp[r];
  }
}

int stencil_order(const Stencil &s);  // examg_fast7.cu

static int env_flag27(const char *name, int dflt);

bool star7_usable(examg_ctx *c, const Stencil &s) {
  static const bool on = env_flag27("EXAMG_NO_FUSED", 0) == 0 && env_flag27("EXAMG_STAR7", 1) != 0;
  if (!on || c->cfg.dim != 3 || !s.is7 || s.coeff_field >= 0 || stencil_order(s) < 0) return false;
  const LevelGeom &g = *geom(c, s.level);
  return g.n[0] >= 33 && g.n[1] >= 33 && g.n[2] >= 33;  // small levels: launch-latency bound either way
}

// Residual = RHS - A * Solution; extend: layers of the neighbour blocks' nodes computed redundantly (as launch_stencil)
int launch_star7(examg_ctx *c, const Stencil &s, const double *u, const double *rhs, double *res, int extend) {
  const LevelGeom &g = *geom(c, s.level);
  GridArgs ga = grid_args_loc(c, g);
  for (int d = 0; d < c->cfg.dim && extend > 0; ++d) {
    if (c->nb[d][0]) ga.lo[d] -= extend;
    if (c->nb[d][1]) ga.hi[d] += extend;
  }
  const int ex = ga.hi[0] - ga.lo[0], ey = ga.hi[1] - ga.lo[1], ez = ga.hi[2] - ga.lo[2];
  if (ex <= 0 || ey <= 0 || ez <= 0) return EXAMG_OK;
  LexArgs a;
  for (int k = 0; k < 27; ++k) a.c[k] = k < 7 ? s.coef[k] : 0.0;
  a.w = 0.0;
  const int gx = (ex + 29) / 30, gy = (ey + 7) / 8;
  int kz = 64;
  while (kz > 8 && (long long)gx * gy * ((ez + kz - 1) / kz) < 148 * 16) kz >>= 1;
  const dim3 grid(gx, gy, (ez + kz - 1) / kz), block(32, 4, 1);
  if (stencil_order(s) == 0) k_star7<0><<<grid, block, 0, c->stream>>>(ga, a, u, rhs, res, kz);
  else k_star7<1><<<grid, block, 0, c->stream>>>(ga, a, u, rhs, res, kz);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

static int env_flag27(const char *name, int dflt) {
  const char *e = getenv(name);
  return e && e[0] ? atoi(e) : dflt;
}

// constant coefficients, all 3^dim entries in lexicographic order (dx fastest); EXAMG_NO_FUSED=1 / EXAMG_LEX=0: generic
bool lex_usable(examg_ctx *c, const Stencil &s) {
  static const bool on = env_flag27("EXAMG_NO_FUSED", 0) == 0 && env_flag27("EXAMG_LEX", 1) != 0;
  if (!on || s.coeff_field >= 0) return false;
  const int dim = c->cfg.dim;
  if (s.nst != (dim == 3 ? 27 : 9)) return false;
  for (int k = 0; k < s.nst; ++k)
    if (s.off[k][0] != k % 3 - 1 || s.off[k][1] != (k / 3) % 3 - 1 || s.off[k][2] != (dim == 3 ? k / 9 - 1 : 0)) return false;
  return true;
}

template <int MODE>
static int launch_lex_mode(examg_ctx *c, const Stencil &s, const GridArgs &ga, const LexArgs &a, const double *u,
                           const double *rhs, double *out) {
  const int ex = ga.hi[0] - ga.lo[0], ey = ga.hi[1] - ga.lo[1], ez = ga.hi[2] - ga.lo[2];
  (void)s;
  if (c->cfg.dim == 2) {
    const int gx = (ex + 127) / 128;
    int ky = 64;
    while (ky > 8 && (long long)gx * ((ey + ky - 1) / ky) < 148 * 8) ky >>= 1;
    k_lex9<MODE><<<dim3(gx, (ey + ky - 1) / ky, 1), dim3(128, 1, 1), 0, c->stream>>>(ga, a, u, rhs, out, ky);
  } else {
    static const int py = env_flag27("EXAMG_LEX_PY", 4);  // 513^3: 0.67 ms with 4 rows per thread, 0.72 with 2, 0.86 with 1
    const int gx = (ex + 29) / 30;
    const int rows = 4 * (py == 4 ? 4 : py == 1 ? 1 : 2), gy = (ey + rows - 1) / rows;
    int kz = 64;
    while (kz > 8 && (long long)gx * gy * ((ez + kz - 1) / kz) < 148 * 8) kz >>= 1;
    const dim3 grid(gx, gy, (ez + kz - 1) / kz), block(32, 4, 1);
    if (py == 4) k_lex27<MODE, 4><<<grid, block, 0, c->stream>>>(ga, a, u, rhs, out, kz);
    else if (py == 1) k_lex27<MODE, 1><<<grid, block, 0, c->stream>>>(ga, a, u, rhs, out, kz);
    else k_lex27<MODE, 2><<<grid, block, 0, c->stream>>>(ga, a, u, rhs, out, kz);
  }
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

// mode: LEX_*; extend: layers of the neighbour blocks' nodes computed redundantly (as launch_stencil)
int launch_lex(examg_ctx *c, const Stencil &s, int mode, const double *u, const double *rhs, double *out, double omega,
               int extend) {
  const LevelGeom &g = *geom(c, s.level);
  GridArgs ga = grid_args_loc(c, g);
  for (int d = 0; d < c->cfg.dim && extend > 0; ++d) {
    if (c->nb[d][0]) ga.lo[d] -= extend;
    if (c->nb[d][1]) ga.hi[d] += extend;
  }
  if (ga.hi[0] <= ga.lo[0] || ga.hi[1] <= ga.lo[1] || ga.hi[2] <= ga.lo[2]) return EXAMG_OK;
  LexArgs a;
  for (int k = 0; k < 27; ++k) a.c[k] = k < s.nst ? s.coef[k] : 0.0;
  a.w = omega / s.coef[s.centre];
  switch (mode) {
    case LEX_APPLY: return launch_lex_mode<LEX_APPLY>(c, s, ga, a, u, rhs, out);
    case LEX_RESIDUAL: return launch_lex_mode<LEX_RESIDUAL>(c, s, ga, a, u, rhs, out);
    default: return launch_lex_mode<LEX_JACOBI>(c, s, ga, a, u, rhs, out);
  }
}

}  // namespace examg
