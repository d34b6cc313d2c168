// examg_coarse.cu -- mgCycle@lvl for the small levels as ONE single-CTA kernel with every
// field of those levels resident in shared memory.
//
// On levels of a few thousand nodes the cycle of the reference is pure latency: per level
// 2*(nu_pre+nu_post) colour loops + residual + restriction + set + prolongation, each its own
// OpenMP region / kernel launch, plus ~8 loops and 3 reductions per CG iteration
// (Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4:152-249; generator
// C/s/e/solver/l3/L3_SolverForEquation.scala:341-395).  Here the sub-cycle below a size
// threshold is one launch: load Solution/RHS/Residual of those levels into shared memory,
// run the statements of mgCycle@lvl ... mgCycle@coarsest in program order with a
// __syncthreads between dependent loops, write everything back.
//
// The per-point arithmetic is the shared one (examg_point_ops.cuh, examg_cg.cuh), so the
// fields are bit-identical to the kernel-per-loop path (tests/test_gpu_programs.py).
#include "examg_cg.cuh"
#include "examg_point_ops.cuh"

namespace examg {

constexpr int CMAXL = 6;

struct CLevel {
  GridArgs g;        // compact layout in shared memory: pitch = n[0], plane = n[0]*n[1]
  StencilArgs st;    // lin[] for the compact layout
  double w;          // omega / diag
  double *gu[2];     // global node-(0,0,0) pointers: Solution slots, RHS, Residual
  double *gf, *gr;
  int gpitch;
  long long gplane;
  int nslots, cur;
  int su[2], sf, sr;  // offsets (doubles) into shared memory
  int total;          // n[0]*n[1]*n[2]
};

struct CArgs {
  int nlev, dim, smoother, nu_pre, nu_post;
  double rw[4];      // restriction weights scale * 2^-(dim+m)
  int cg_max_it;
  double cg_rel_tol;
  double *gp, *gap;  // cgTmp0 / cgTmp1 on level 0
  int sp, sap;
  int *iters_out;
  CLevel lv[CMAXL];  // lv[0] = coarsest
};

__device__ __forceinline__ void c_copy(double *dst, const double *src, const CLevel &L, bool to_shared) {
  const int n0 = L.g.n[0], n01 = L.g.n[0] * L.g.n[1];
  for (int t = threadIdx.x; t < L.total; t += blockDim.x) {
    const int k = t / n01, r = t - k * n01, j = r / n0, i = r - j * n0;
    const long long gi = i + (long long)L.gpitch * j + L.gplane * k;
    if (to_shared) dst[t] = src[gi]; else dst[gi] = src[t];
  }
}

// loop points of a level: t -> (i,j,k,p)
struct CPoint { int i, j, k, p; };
__device__ __forceinline__ CPoint c_point(const GridArgs &g, int t, int ex, int ey) {
  CPoint q;
  q.i = g.lo[0] + t % ex;
  q.j = g.lo[1] + (t / ex) % ey;
  q.k = g.lo[2] + t / (ex * ey);
  q.p = q.i + g.pitch * q.j + (int)g.plane * q.k;
  return q;
}

__global__ void __launch_bounds__(1024, 1) k_coarse_cycle(const __grid_constant__ CArgs a) {
  extern __shared__ __align__(16) double sm[];
  const int tid = threadIdx.x, nt = blockDim.x;
  int cur[CMAXL];
  // ---- load ----
  for (int l = 0; l < a.nlev; ++l) {
    const CLevel &L = a.lv[l];
    cur[l] = L.cur;
    for (int s = 0; s < L.nslots; ++s) c_copy(sm + L.su[s], L.gu[s], L, true);
    c_copy(sm + L.sf, L.gf, L, true);
    c_copy(sm + L.sr, L.gr, L, true);
  }
  c_copy(sm + a.sp, a.gp, a.lv[0], true);
  c_copy(sm + a.sap, a.gap, a.lv[0], true);
  __syncthreads();

  auto smooth = [&](int l, int nsweeps) {
    const CLevel &L = a.lv[l];
    const GridArgs &g = L.g;
    const int ex = g.hi[0] - g.lo[0], ey = g.hi[1] - g.lo[1], ez = g.hi[2] - g.lo[2];
    const int total = ex > 0 && ey > 0 && ez > 0 ? ex * ey * ez : 0;
    const double *f = sm + L.sf;
    for (int it = 0; it < nsweeps; ++it) {
      if (a.smoother == EXAMG_SMOOTH_RBGS) {
        double *u = sm + L.su[cur[l]];
        for (int color = 0; color < 2; ++color) {
          for (int t = tid; t < total; t += nt) {
            const CPoint q = c_point(g, t, ex, ey);
            if (((q.i + q.j + q.k + g.par) & 1) != color) continue;
            const double acc = conv<0, false>(L.st, nullptr, 0, u, q.p);
            const double r = f[q.p] - acc;
            u[q.p] = fma(L.w, r, u[q.p]);
          }
          __syncthreads();
        }
      } else {
        const double *u = sm + L.su[cur[l]];
        const int nxt = (cur[l] + 1) % L.nslots;
        double *un = sm + L.su[nxt];
        for (int t = tid; t < total; t += nt) {
          const CPoint q = c_point(g, t, ex, ey);
          const double acc = conv<0, false>(L.st, nullptr, 0, u, q.p);
          const double r = f[q.p] - acc;
          un[q.p] = fma(L.w, r, u[q.p]);
        }
        cur[l] = nxt;
        __syncthreads();
      }
    }
  };

  // ---- down: mgCycle@l for l = nlev-1 .. 1 up to the recursive call ----
  for (int l = a.nlev - 1; l >= 1; --l) {
    const CLevel &L = a.lv[l], &C = a.lv[l - 1];
    smooth(l, a.nu_pre);
    {  // Residual = RHS - A * Solution
      const GridArgs &g = L.g;
      const int ex = g.hi[0] - g.lo[0], ey = g.hi[1] - g.lo[1], ez = g.hi[2] - g.lo[2];
      const int total = ex > 0 && ey > 0 && ez > 0 ? ex * ey * ez : 0;
      const double *u = sm + L.su[cur[l]], *f = sm + L.sf;
      double *r = sm + L.sr;
      for (int t = tid; t < total; t += nt) {
        const CPoint q = c_point(g, t, ex, ey);
        r[q.p] = f[q.p] - conv<0, false>(L.st, nullptr, 0, u, q.p);
      }
      __syncthreads();
    }
    {  // RHS@coarser = Restriction * Residual ; Solution@coarser = 0
      const GridArgs &g = C.g;
      const int ex = g.hi[0] - g.lo[0], ey = g.hi[1] - g.lo[1], ez = g.hi[2] - g.lo[2];
      const int total = ex > 0 && ey > 0 && ez > 0 ? ex * ey * ez : 0;
      const double *rf = sm + L.sr;
      double *fc = sm + C.sf, *uc = sm + C.su[cur[l - 1]];
      const int fp = L.g.pitch;
      const long long fpl = L.g.plane;
      for (int t = tid; t < total; t += nt) {
        const CPoint q = c_point(g, t, ex, ey);
        const long long qf = 2 * q.i + (long long)fp * (2 * q.j) + (a.dim == 3 ? fpl * (2 * q.k) : 0);
        fc[q.p] = a.dim == 3 ? restrict_point<3>(rf, qf, fp, fpl, a.rw[0], a.rw[1], a.rw[2], a.rw[3])
                             : restrict_point<2>(rf, qf, fp, fpl, a.rw[0], a.rw[1], a.rw[2], a.rw[3]);
        uc[q.p] = 0.0;
      }
      __syncthreads();
    }
  }
  // ---- mgCycle@coarsest ----
  {
    const CLevel &L = a.lv[0];
    __shared__ CgArgs ca;
    if (tid == 0) {
      ca.g = L.g;
      ca.st = L.st;
      ca.sfvol = 0;
      ca.max_it = a.cg_max_it;
      ca.rel_tol = a.cg_rel_tol;
      ca.per[0] = ca.per[1] = ca.per[2] = 0;  // periodic contexts do not use this kernel (coarse_cycle_top)
    }
    __syncthreads();
    const int iters = cg_device<false>(ca, nullptr, sm + L.su[cur[0]], sm + L.sf, sm + L.sr, sm + a.sp, sm + a.sap);
    if (tid == 0) *a.iters_out = iters;
    __syncthreads();
  }
  // ---- up ----
  for (int l = 1; l < a.nlev; ++l) {
    const CLevel &L = a.lv[l], &C = a.lv[l - 1];
    {  // Solution += Prolongation * Solution@coarser
      const GridArgs &g = L.g;
      const int ex = g.hi[0] - g.lo[0], ey = g.hi[1] - g.lo[1], ez = g.hi[2] - g.lo[2];
      const int total = ex > 0 && ey > 0 && ez > 0 ? ex * ey * ez : 0;
      const double *uc = sm + C.su[cur[l - 1]];
      double *uf = sm + L.su[cur[l]];
      for (int t = tid; t < total; t += nt) {
        const CPoint q = c_point(g, t, ex, ey);
        uf[q.p] = a.dim == 3 ? prolong_point<3>(uc, C.g.pitch, C.g.plane, q.i, q.j, q.k, uf[q.p])
                             : prolong_point<2>(uc, C.g.pitch, C.g.plane, q.i, q.j, q.k, uf[q.p]);
      }
      __syncthreads();
    }
    smooth(l, a.nu_post);
  }
  // ---- store ----
  for (int l = 0; l < a.nlev; ++l) {
    const CLevel &L = a.lv[l];
    for (int s = 0; s < L.nslots; ++s) c_copy(L.gu[s], sm + L.su[s], L, false);
    if (l < a.nlev - 1) c_copy(L.gf, sm + L.sf, L, false);  // RHS of the top level is an input
    c_copy(L.gr, sm + L.sr, L, false);
  }
  c_copy(a.gp, sm + a.sp, a.lv[0], false);
  c_copy(a.gap, sm + a.sap, a.lv[0], false);
}

// ---- host ---------------------------------------------------------------------------

constexpr size_t COARSE_SMEM_MAX = 200 * 1024;

// shared-memory bytes the levels min_level..top need (0: not eligible)
static size_t coarse_bytes(examg_ctx *c, int top, const int *sol) {
  size_t doubles = 0;
  for (int lvl = c->cfg.min_level; lvl <= top; ++lvl) {
    const LevelGeom &g = *geom(c, lvl);
    const size_t tot = (size_t)g.n[0] * g.n[1] * g.n[2];
    const Field &fu = c->fields[sol[lvl - c->cfg.min_level]];
    doubles += tot * (fu.nslots + 2);
    if (lvl == c->cfg.min_level) doubles += 2 * tot;
  }
  return doubles * sizeof(double);
}

// Largest level for which mgCycle@level can run in the single-CTA kernel; min_level - 1 if none.
int coarse_cycle_top(examg_ctx *c, const examg_cycle_desc &d, const int *stencils, const int *sol, int max_top) {
  // Opt-in (EXAMG_COARSE_FUSE=1): measured on B200 the single CTA needs 87 us for levels 2..4 of the
  // benchmark against ~55 us for the same levels as nodes of the cycle's CUDA graph (profiles/r01_notes.md).
  static const bool off = !(getenv("EXAMG_COARSE_FUSE") && getenv("EXAMG_COARSE_FUSE")[0] == '1');
  const int lo = c->cfg.min_level;
  if (off || c->world > 1 || c->periodic_any) return lo - 1;
  if (d.coarse_solver != EXAMG_CGS_CG || d.smoother > EXAMG_SMOOTH_RBGS || d.use_fas) return lo - 1;  // k_coarse_cycle ends in CG; BiCGStab / CR / MinRes run as graph nodes
  int top = lo - 1;
  for (int lvl = lo; lvl <= max_top && lvl - lo < CMAXL; ++lvl) {
    const Stencil &s = c->stencils[stencils[lvl - lo]];
    const Field &fu = c->fields[sol[lvl - lo]];
    const LevelGeom &g = *geom(c, lvl);
    if (s.coeff_field >= 0 || fu.ncomp != 1 || fu.nslots > 2) break;
    if (d.smoother == EXAMG_SMOOTH_JACOBI && fu.nslots != 2) break;
    if ((long long)g.n[0] * g.n[1] * g.n[2] > 6000) break;  // beyond that several CTAs win
    if (coarse_bytes(c, lvl, sol) > COARSE_SMEM_MAX) break;
    top = lvl;
  }
  return top > lo ? top : lo - 1;  // the coarsest level alone is the plain CG kernel
}

int launch_coarse_cycle(examg_ctx *c, const examg_cycle_desc &d, int top, const int *stencils, const int *sol,
                        const int *rhs, const int *res, int cg_p, int cg_ap) {
  EXAMG_ONCE_PER_DEVICE(c, {
    EXAMG_CUDA(cudaFuncSetAttribute(k_coarse_cycle, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)COARSE_SMEM_MAX));
  });
  const int lo = c->cfg.min_level;
  CArgs a;
  a.nlev = top - lo + 1;
  a.dim = c->cfg.dim;
  a.smoother = d.smoother;
  a.nu_pre = d.nu_pre;
  a.nu_post = d.nu_post;
  for (int m = 0; m <= 3; ++m) a.rw[m] = d.restrict_scale * ldexp(1.0, -(a.dim + m));
  a.cg_max_it = d.cg_max_it;
  a.cg_rel_tol = d.cg_rel_tol;
  a.iters_out = c->int_out;
  int off = 0;
  for (int l = 0; l < a.nlev; ++l) {
    const LevelGeom &g = *geom(c, lo + l);
    const Stencil &s = c->stencils[stencils[l]];
    Field &fu = c->fields[sol[l]];
    CLevel &L = a.lv[l];
    L.g = grid_args(g);
    L.g.pitch = g.n[0];
    L.g.plane = (long long)g.n[0] * g.n[1];
    L.st.nst = s.nst;
    L.st.centre = s.centre;
    for (int k = 0; k < EXAMG_MAX_STENCIL; ++k) { L.st.lin[k] = 0; L.st.coef[k] = 0.0; }
    for (int k = 0; k < s.nst; ++k) {
      L.st.lin[k] = s.off[k][0] + (long long)L.g.pitch * s.off[k][1] + L.g.plane * s.off[k][2];
      L.st.coef[k] = s.coef[k];
    }
    L.w = d.omega / s.coef[s.centre];
    L.nslots = fu.nslots;
    L.cur = fu.cur;
    L.total = g.n[0] * g.n[1] * g.n[2];
    for (int sl = 0; sl < 2; ++sl) {
      L.gu[sl] = sl < fu.nslots ? field_ptr(c, &fu, sl) : nullptr;
      L.su[sl] = off;
      if (sl < fu.nslots) off += L.total;
    }
    L.gf = field_ptr(c, &c->fields[rhs[l]], -1);
    L.gr = field_ptr(c, &c->fields[res[l]], -1);
    L.sf = off; off += L.total;
    L.sr = off; off += L.total;
    L.gpitch = g.pitch;
    L.gplane = g.plane;
  }
  a.gp = field_ptr(c, &c->fields[cg_p], -1);
  a.gap = field_ptr(c, &c->fields[cg_ap], -1);
  a.sp = off; off += a.lv[0].total;
  a.sap = off; off += a.lv[0].total;
  const size_t bytes = sizeof(double) * (size_t)off;
  if (bytes > COARSE_SMEM_MAX) return set_error(EXAMG_EUNSUP, "coarse cycle: %zu bytes of shared memory", bytes);
  k_coarse_cycle<<<1, 1024, bytes, c->stream>>>(a);
  EXAMG_LAUNCH_CHECK(c);
  // `advance Solution` of the Jacobi sweeps
  if (d.smoother == EXAMG_SMOOTH_JACOBI)
    for (int l = 1; l < a.nlev; ++l) {
      Field &fu = c->fields[sol[l]];
      fu.cur = (fu.cur + d.nu_pre + d.nu_post) % fu.nslots;
    }
  return EXAMG_OK;
}

}  // namespace examg
