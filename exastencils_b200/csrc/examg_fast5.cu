// examg_fast5.cu -- 2D 5-point constant-coefficient fast path: S red-black half-sweeps (S = 2, 4, 6: one to three full
// sweeps) per pass.  BASELINE config 1 (Examples/Poisson 2D 5-point FD, V(3,3) RBGS, max level 10).
//
// Reference loops being replaced (per sweep two masked full-grid passes, each with its communicate and apply bc):
//   repeat 3 times { color with { ( i0 + i1 ) % 2, communicate Solution
//     loop over Solution { Solution += 0.8 / diag ( Laplace ) * ( RHS - Laplace * Solution ) } apply bc to Solution } }
//                                                  Examples/Poisson/2D_FD_Poisson_fromL4.exa4:189-224
//   colour order / duplication                     C/s/e/baseExt/l4/L4_ColorLoops.scala:44-70
//
// A 1025^2 level is 8.4 MB per field: the whole hierarchy lives in L2 and the cycle is launch-latency bound (round 1:
// 131 graph nodes, 0.28 ms per V(3,3) cycle, 12 smoother launches per level).  So the 2D kernel is built for few
// launches and short critical paths, not for HBM: "warp-private wavefront", the scheme of k_of_rbw (examg_optflow.cu):
//   * a warp owns a strip of 64 nodes in x (S-deep halo included) and a segment of rows, and marches in y; a lane owns
//     two adjacent nodes (one of each colour) and keeps S+2 rows of Solution and RHS in registers
//   * half-sweep s works two rows behind half-sweep s-1 (see k_rb2d): y neighbours are the lane's own registers, the x
//     neighbour inside the pair too, the other one comes from the adjacent lane by ONE shuffle; no shared memory, no
//     block barrier; 128-bit loads two rows ahead
//   * out of place (Solution -> its shadow buffer, swapped afterwards); row t-S is final and stored with 128-bit stores
// Three sweeps = one launch where the generic path needs six; the arithmetic is rb_update's (bit-identical to the
// unfused order, DESIGN.md "Arithmetic contract").
#include <type_traits>
#include <utility>

#include "examg_device.cuh"

namespace examg {

struct Rb2Args {
  int nx, ny, pitch;
  int ulo[2], uhi[2];  // update range (physical-boundary nodes excluded)
  int par;
  double coef[5];
  double w;            // omega / diag
};

template <int... Ks, class F>
__device__ __forceinline__ bool rb2_steps(std::integer_sequence<int, Ks...>, F &&fn) {
  return (fn(std::integral_constant<int, Ks>{}) && ...);
}

// ORD = declaration order of the stencil entries: 0 = centre,-x,+x,-y,+y (Examples/Poisson), 1 = centre,+x,-x,+y,-y
// (Testing/*)
template <int ORD>
__device__ __forceinline__ double rb2_update(double c, double xm, double xp, double ym, double yp, double f,
                                             const double (&coef)[5], double w) {
  double acc = coef[0] * c;
  if (ORD == 0) {
    acc = fma(coef[1], xm, acc);
    acc = fma(coef[2], xp, acc);
    acc = fma(coef[3], ym, acc);
    acc = fma(coef[4], yp, acc);
  } else {
    acc = fma(coef[1], xp, acc);
    acc = fma(coef[2], xm, acc);
    acc = fma(coef[3], yp, acc);
    acc = fma(coef[4], ym, acc);
  }
  return fma(w, f - acc, c);
}

// Half-sweep s works on row t - 2s + 1, TWO rows behind half-sweep s-1: the row above it (t - 2s + 2) then got its
// half-sweep s-1 one step EARLIER, so the S half-sweeps of a step do not depend on one another and their FP64 chains
// overlap (with a lag of one row they form one chain of S dependent updates per step; measured 27 us for three sweeps at
// 1025^2 -- slower than six generic launches).  Column updated by half-sweep s at step K (steps numbered so that
// (t0 + par) is even): X = (s + K) & 1.
template <int S, int ORD>
__global__ void __launch_bounds__(128)
k_rb2d(const Rb2Args a, const double *__restrict__ u, const double *__restrict__ f, double *__restrict__ uo,
       int seg_rows, int nstrips) {
  // PF rows are in flight: row t + PF is loaded straight into its window slot at step t (no register moves, so the
  // march does not wait for a load before its data is due; an L2 round trip is ~4 steps long)
  constexpr int PF = 4, NW = 2 * S + 2 + PF, IW = 64 - 2 * S;
  const int lane = threadIdx.x & 31;
  const int strip = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (strip >= nstrips) return;  // whole warp
  const int nx = a.nx, ny = a.ny, pitch = a.pitch;
  const int ox = IW * strip, x0 = ox - S + 2 * lane;  // even: node A = (x0, .), node B = (x0 + 1, .)
  const int za = blockIdx.y * seg_rows, zb = min(ny, za + seg_rows);
  // colour of node (x, y) = (x + y + par) & 1
  int t0 = za - S - 1;           // rows za-S .. are needed in their original state; one more keeps the window simple
  t0 -= (t0 + a.par) & 1;
  const int tl = zb - 1 + 2 * S - 1;  // the step at which half-sweep S reaches row zb-1
  const bool inx = x0 >= 0 && x0 < nx;
  const bool st = inx && x0 >= ox && x0 < ox + IW;
  const bool upA = x0 >= a.ulo[0] && x0 < a.uhi[0], upB = x0 + 1 >= a.ulo[0] && x0 + 1 < a.uhi[0];
  double coef[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) coef[k] = a.coef[k];
  const double w = a.w;

  double uw[NW][2], fw[NW][2];
#pragma unroll
  for (int q = 0; q < NW; ++q) uw[q][0] = uw[q][1] = fw[q][0] = fw[q][1] = 0.0;
  auto fetch = [&](int r, double (&du)[2], double (&df)[2]) {
    double2 vu = make_double2(0.0, 0.0), vf = vu;
    if (inx && r >= 0 && r < ny && r <= tl) {
      const long long p = x0 + (long long)pitch * r;
      vu = *reinterpret_cast<const double2 *>(u + p);
      vf = __ldg(reinterpret_cast<const double2 *>(f + p));
    }
    du[0] = vu.x; du[1] = vu.y;
    df[0] = vf.x; df[1] = vf.y;
  };
#pragma unroll
  for (int i = 0; i < PF; ++i) fetch(t0 + i, uw[i], fw[i]);

  auto step = [&](auto kc, const int t) {
    constexpr int K = decltype(kc)::value;
    fetch(t + PF, uw[(K + PF) % NW], fw[(K + PF) % NW]);
#pragma unroll
    for (int s = 1; s <= S; ++s) {
      const int X = (s + K) & 1;  // compile-time after unrolling
      const int c = (K + 2 * NW - 2 * s + 1) % NW, cm = (c + NW - 1) % NW, cp = (c + 1) % NW;
      const int r = t - 2 * s + 1;
      // x neighbour owned by the adjacent lane: A's west = B of lane-1, B's east = A of lane+1
      const double sx = X == 0 ? __shfl_up_sync(0xffffffffu, uw[c][1], 1) : __shfl_down_sync(0xffffffffu, uw[c][0], 1);
      // no branch around the update: the S chains of a step then interleave (with a branch per half-sweep they ran one
      // after the other, 850 cycles per step); the select keeps nodes outside the update range
      const bool upd = (X == 0 ? upA : upB) && r >= a.ulo[1] && r < a.uhi[1];
      const double cu = uw[c][X];
      const double xm = X == 0 ? sx : uw[c][0], xp = X == 0 ? uw[c][1] : sx;
      const double nv = rb2_update<ORD>(cu, xm, xp, uw[cm][X], uw[cp][X], fw[c][X], coef, w);
      uw[c][X] = upd ? nv : cu;
    }
    const int zf = t - 2 * S + 1;  // final after this step's half-sweep S
    if (st && zf >= za && zf < zb) {
      const int c = (K + 2 * NW - 2 * S + 1) % NW;
      *reinterpret_cast<double2 *>(uo + x0 + (long long)pitch * zf) = make_double2(uw[c][0], uw[c][1]);
    }
  };

  // whole rounds of NW steps without an exit test in between: one basic block per round, so that the scheduler can
  // overlap the chains of successive steps (steps beyond tl load nothing and store nothing; the launcher sizes the
  // segments so that few are wasted)
  for (int t = t0; t <= tl; t += NW)
    rb2_steps(std::make_integer_sequence<int, NW>{}, [&](auto kc) {
      step(kc, t + decltype(kc)::value);
      return true;
    });
}

// ---- host side -------------------------------------------------------------------

int ensure_shadow(examg_ctx *c, Field &fu);

static int env_flag5(const char *name, int dflt) {
  const char *e = getenv(name);
  return e && e[0] ? atoi(e) : dflt;
}

// 0 / 1: one of the two known declaration orders of the 5-point stencil; -1: something else
static int stencil_order5(const Stencil &s) {
  static const int dx[2][5] = {{0, -1, 1, 0, 0}, {0, 1, -1, 0, 0}};
  static const int dy[2][5] = {{0, 0, 0, -1, 1}, {0, 0, 0, 1, -1}};
  if (s.nst != 5 || s.coeff_field >= 0) return -1;
  for (int o = 0; o < 2; ++o) {
    bool same = true;
    for (int k = 0; k < 5; ++k) same = same && s.off[k][0] == dx[o][k] && s.off[k][1] == dy[o][k] && s.off[k][2] == 0;
    if (same) return o;
  }
  return -1;
}

// smallest level extent the strip kernel is used for (below: one warp would do everything; the generic kernels do)
constexpr int FUSED2D_MIN_N = 9;

bool fused_rb2d_usable(examg_ctx *c, const Stencil &s, const Field &fu) {
  // Used where a level exceeds L2 (>= 2^22 nodes: 32 MB per field), i.e. where the sweep is HBM-bound: there three sweeps
  // per pass are worth 2-3x (tools/matrix_bench.py at 8193^2: 0.49 of the one-pass-per-sweep roofline with one launch per
  // colour).  On L2-resident levels a V(3,3) cycle is launch-latency bound and the fused kernel -- one warp's march
  // through ~18-40 fully unrolled steps per launch, 11-32 us -- is not faster than the generic colour kernels (0.294
  // against 0.285 ms per cycle at 1025^2, profiles/r02_notes.md): there it is opt-in (EXAMG_FUSED_2D=1; =0 switches
  // it off everywhere).  Bit-identical either way.
  static const int mode = env_flag5("EXAMG_NO_FUSED", 0) != 0 ? 0 : env_flag5("EXAMG_FUSED_2D", -1);
  const bool off = mode == 0;
  if (off || c->cfg.dim != 2 || fu.ncomp != 1 || fu.loc != 0 || stencil_order5(s) < 0) return false;
  if (c->world > 1 || c->periodic_any) return false;  // several blocks / periodic dimensions: generic kernels
  const LevelGeom &g = *geom(c, s.level);
  if (mode < 0 && (long long)g.n[0] * g.n[1] < (1LL << 22)) return false;
  return g.n[0] >= FUSED2D_MIN_N && g.n[1] >= FUSED2D_MIN_N;
}

// sweeps the next launch will fuse (1..3).  On HBM-sized levels two per launch are best (8193^2: 0.283 / 0.268 / 0.307 ms
// per sweep for one / two / three per launch -- the kernel's time grows with the sweeps it fuses, it is bound by its
// dependent FP64 chains, not by HBM any more); on L2-resident levels (opt-in) launches count: three.
int rb2d_fused_plan(int nsweeps) {
  static const int m = [] {
    const int dflt = env_flag5("EXAMG_FUSED_2D", -1) > 0 ? 3 : 2;
    int v = env_flag5("EXAMG_FUSE_SWEEPS_2D", dflt);
    return v < 1 ? 1 : v > 3 ? 3 : v;
  }();
  return nsweeps < m ? nsweeps : m;
}

template <int S>
static int launch_rb2d_s(examg_ctx *c, const Stencil &s, double omega, Field &fu, const double *rhs) {
  constexpr int IW = 64 - 2 * S;
  const LevelGeom &g = *geom(c, s.level);
  Rb2Args a;
  a.nx = g.n[0];
  a.ny = g.n[1];
  a.pitch = g.pitch;
  for (int d = 0; d < 2; ++d) {
    a.ulo[d] = g.lo[d];
    a.uhi[d] = g.hi[d];
  }
  a.par = g.par;
  for (int k = 0; k < 5; ++k) a.coef[k] = s.coef[k];
  a.w = omega / s.coef[s.centre];
  const int nstrips = (a.nx + IW - 1) / IW;
  const int gx = (nstrips + 3) / 4;
  // The level lives in L2 and a launch is as long as one warp's march: short segments (2S + 1 warm-up rows each) on the
  // small levels, where only latency counts; on the large ones the warm-up is FP64 work, so segments of >= 4S rows
  // a segment of R rows takes R + 2S + 1 (+ parity) steps, executed in rounds of NW = 2S + 6: R = NW - 2S - 2 = 4 rows for
  // one round (small levels: latency), 2 NW - 2S - 2 for two (large levels: less redundant FP64 work)
  constexpr int NWH = 2 * S + 6;
  int seg_rows = (long long)a.nx * a.ny >= (1 << 18) ? 2 * NWH - 2 * S - 2 : NWH - 2 * S - 2;
  (void)gx;
  const int nseg = (a.ny + seg_rows - 1) / seg_rows;
  const double *u = fu.slot[fu.cur] + g.origin;
  double *uo = fu.shadow + g.origin;
  if (stencil_order5(s) == 0)
    k_rb2d<S, 0><<<dim3(gx, nseg), 128, 0, c->stream>>>(a, u, rhs, uo, seg_rows, nstrips);
  else
    k_rb2d<S, 1><<<dim3(gx, nseg), 128, 0, c->stream>>>(a, u, rhs, uo, seg_rows, nstrips);
  EXAMG_LAUNCH_CHECK(c);
  double *t = fu.slot[fu.cur];
  fu.slot[fu.cur] = fu.shadow;
  fu.shadow = t;
  fu.swapped = !fu.swapped;
  return EXAMG_OK;
}

// up to three full sweeps by one launch; *done = sweeps done
int launch_rb2d_fused(examg_ctx *c, const Stencil &s, double omega, Field &fu, const double *rhs, int nsweeps, int *done) {
  int rc = ensure_shadow(c, fu);
  if (rc) return rc;
  const int plan = rb2d_fused_plan(nsweeps);
  *done = plan;
  if (plan == 3) return launch_rb2d_s<6>(c, s, omega, fu, rhs);
  if (plan == 2) return launch_rb2d_s<4>(c, s, omega, fu, rhs);
  return launch_rb2d_s<2>(c, s, omega, fu, rhs);
}

}  // namespace examg
