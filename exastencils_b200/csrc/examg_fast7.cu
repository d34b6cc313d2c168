// examg_fast7.cu -- 3D 7-point constant-coefficient fast path: S red-black half-sweeps
// (S = 2: one full sweep, S = 4: two sweeps) per pass over HBM.
//
// Reference loops being replaced (two masked full-grid passes per sweep):
//   color with { (i0+i1+i2)%2, communicate; loop over Solution { Solution += omega/diag(A) *
//   (RHS - A*Solution) }; apply bc }         Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4:204-213
//   colour order / duplication               C/s/e/baseExt/l4/L4_ColorLoops.scala:44-70
//   temporal blocking precedent              C/s/e/baseExt/ir/IR_ContractingLoop.scala:48-90,131-195
//
// Design (B200), "register-window wavefront":
//   * out of place: reads Solution, writes the shadow buffer; the two are swapped afterwards
//   * a CTA (512 threads, one per SM, persistent) owns an x-y tile of 64 x 32 nodes (S-deep halo
//     included) and marches in z.  TMA (cp.async.bulk.tensor.3d) stages plane t of Solution and
//     RHS in a 4-deep shared-memory ring (mbarrier completion, three planes in flight).
//   * every thread owns a 2 x 2 block of the plane and keeps a window of S+2 planes of it in
//     registers.  Half-sweep s (colour (s-1)&1) works on plane t-s, one plane behind half-sweep
//     s-1: the z neighbours and two of the four in-plane neighbours are the thread's own
//     registers, the other two come from the neighbouring threads through a colour-compressed
//     exchange buffer in shared memory that half-sweep s-1 filled one step earlier (double
//     buffered: one __syncthreads per plane).  The update is Gauss-Seidel in place -- in the
//     register window.
//   * plane t-S is final after half-sweep S and is written with 128-bit stores
//   * schedule: full waves in which CTA b marches through the whole z column of tile
//     wave*grid + b -- adjacent tiles advance side by side, so the halo two tiles share is read
//     from HBM once and found in the 126 MB L2 by the other (measured: 89% of the copy roofline
//     against 60% with a stream-K split) -- then the remaining tiles*planes units are
//     split evenly over the grid; a segment start costs 2S planes
// HBM traffic per point (S = 2) with the halo served from L2: 8 + 8 + 8 = 24 B, the algorithmic
// minimum; without L2 sharing 8/0.82 * 2 + 8 = 27.5 B.  S = 4: 24 B for 48 B of algorithmic work.
// Tried and dropped (profiles/r01_notes.md): L2 prefetch ahead of the ring (evicts the shared
// halo lines), a loose lockstep between neighbouring CTAs (no gain over free-running).
//
// Arithmetic is the contract of examg_kernels.cu (bit-identical to the unfused order).
#include <cuda.h>

#include <type_traits>

#include "examg_device.cuh"

namespace examg {

constexpr int RW = 64;                 // tile width (x), = 2 * warp size
constexpr int RH = 32;                 // tile height (y), = 2 * warps per CTA
constexpr int RNT = 512;               // threads per CTA
constexpr int RD = 4;                  // depth of the TMA ring (planes)
constexpr int RPLANE = RW * RH;        // doubles per staged plane
constexpr int RXB = (RH + 2) * 32 + 32;  // doubles per exchange buffer: rows -1..RH of 32, 16 pad each end
constexpr int ROFF_RING = 1024;        // bytes in front of the ring: mbarriers + landing zone of row -1 reads

// PRO adds a pair of exchange buffers: the corrected plane is published for half-sweep 1 (the staged plane in the ring
// holds the uncorrected values)
// PRO also stages the coarse planes under the tile (33 x 17 nodes, box 34 x 17) in a 4-deep TMA ring of its own
constexpr int CBW = RW / 2 + 4, CBH = RH / 2 + 1;  // coarse box: 33 x 17 nodes, one more on the left when the corner is odd
constexpr int CPLS = 640;                          // doubles per coarse ring slot (128-byte multiple >= CBW * CBH)
constexpr int CRD = 4;                             // coarse ring depth
static_assert(CBW * CBH <= CPLS && (CPLS * 8) % 128 == 0, "coarse ring slot");
// X0 (without PRO): half-sweep 1 also reads its neighbours from an exchange buffer that the load stage fills, instead
// of from the staged plane in the ring, whose x neighbours sit two doubles apart per lane (2-way bank conflicts)
template <int S, bool PRO = false, bool X0 = false>
constexpr size_t rbw_smem() {
  return ROFF_RING + sizeof(double) * (2 * RD * RPLANE + (S - 1 + (PRO || X0 ? 1 : 0)) * 2 * RXB + (PRO ? CRD * CPLS : 0));
}

struct RbwArgs {
  int n[3];
  int ulo[3], uhi[3];  // update bounds (local node coordinates) of the last half-sweep
  int elo[3], ehi[3];  // 1: neighbour block on that side; half-sweep s also updates S-s layers beyond
  int pitch;
  long long plane;
  int tiles_x, tiles;
  int nzp;             // output planes 0..nzp-1
  int par;
  int pf;              // planes ahead of the TMA loads that are prefetched into L2 (0: none)
  double coef[7];
  double w;            // omega / diag
  // PRO: `Solution += Prolongation@coarser * Solution@coarser` applied to every plane as it enters the register window
  int plo[3], phi[3];  // fine nodes that take the correction (the loop range, S layers wider where a neighbour block is)
  int cpitch;
  long long cplane;
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void tma_load_plane(void *dst, const CUtensorMap *tm, int x, int y, int z, uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      ::"r"(smem_u32(dst)), "l"((uint64_t)tm), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
      : "memory");
}

__device__ __forceinline__ void tma_prefetch_plane(const CUtensorMap *tm, int x, int y, int z) {
  asm volatile("cp.async.bulk.prefetch.tensor.3d.L2.global.tile [%0, {%1, %2, %3}];"
               ::"l"((uint64_t)tm), "r"(x), "r"(y), "r"(z) : "memory");
}

// One point update.  ORD = declaration order of the stencil entries, which fixes the
// floating-point association: 0 = centre,-x,+x,-y,+y,-z,+z (Benchmark/Poisson3D, Examples/Poisson),
// 1 = centre,+x,-x,+y,-y,+z,-z (Testing/Smoothers, Testing/SISC, Testing/Opts).
template <int ORD>
__device__ __forceinline__ double rb_update(double c, double xm, double xp, double ym, double yp, double zm, double zp,
                                            double f, const double (&coef)[7], double w) {
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
  const double r = f - acc;
  return fma(w, r, c);
}

// PRO = true: the correction step of mgCycle (`Solution += NodeProlongation@coarser * Solution@coarser`,
// Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4:233-236) is folded into the load stage of the post-smoothing pass: a
// plane that enters the register window first takes its interpolated coarse values -- the sums and weights of
// prolong_point (examg_point_ops.cuh), so the result is bit-identical to the separate loop -- and is smoothed from
// there.  Saves one read and one write of the fine Solution per cycle.  uc = node (0,0,0) of Solution@coarser.
template <int S, int ORD, bool PRO, bool X0 = false>
__global__ void __launch_bounds__(RNT, 1)
k_rbw(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_f,
      const __grid_constant__ RbwArgs a, double *__restrict__ uout, const __grid_constant__ CUtensorMap tm_c) {
  constexpr int NW = S + 2;  // planes in the register window: t-S-1 .. t
  static_assert(S == 2 || S == 4, "S half-sweeps");
  extern __shared__ __align__(1024) unsigned char smem[];
  uint64_t *full = reinterpret_cast<uint64_t *>(smem);
  double *ring_u = reinterpret_cast<double *>(smem + ROFF_RING);
  double *ring_f = ring_u + RD * RPLANE;
  double *xbuf = ring_f + RD * RPLANE;  // [(s-1)*2 + plane parity][RXB], s = 1..S-1
  uint64_t *cfull = full + 8;           // PRO: mbarriers of the coarse ring
  double *cring = xbuf + S * 2 * RXB;   // PRO: [CRD][CPLS] coarse planes (behind S pairs of exchange buffers)

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int x0 = 2 * lane, y0 = 2 * warp;  // the thread's 2x2 block inside the tile
  if (tid == 0) {
    for (int i = 0; i < RD; ++i) mbar_init(&full[i], 1);
    if (PRO)
      for (int i = 0; i < CRD; ++i) mbar_init(&cfull[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  double coef[7];
#pragma unroll
  for (int k = 0; k < 7; ++k) coef[k] = a.coef[k];
  const double w = a.w;
  const long long plane8 = a.plane * 8;
  const long long pitch8 = (long long)a.pitch * 8;
  const int ld_off = y0 * RW + x0;           // (x0, y0) inside a staged plane
  const int x_off = 16 + (y0 + 1) * 32 + lane;  // (row y0, column pair `lane`) inside an exchange buffer

  // Schedule: `nfull` waves in which every CTA marches through the whole z column of one tile
  // (tile = wave * gridDim + blockIdx: neighbouring tiles run side by side at the same z, so the
  // halo they share is read from HBM once and found in L2 by the other), then the remaining
  // tiles * planes units are split evenly over the grid.
  const int G = gridDim.x, bid = blockIdx.x;
  const int nfull = a.tiles / G;
  const long long rem_units = (long long)(a.tiles - nfull * G) * a.nzp;
  const long long u1 = rem_units * (bid + 1) / G;
  long long ub = rem_units * bid / G;
  uint32_t kt = 0;  // planes loaded so far, modulo 2*RD: ring slot kt & 3, mbarrier phase kt >> 2
  uint32_t kcn0 = 0;  // PRO: the same count for the coarse ring, at the first coarse plane of the tile
  for (int seg = 0;; ++seg) {
    int tile, za, zb;
    if (seg < nfull) {
      tile = seg * G + bid;
      za = 0;
      zb = a.nzp;
    } else {
      if (ub >= u1) break;
      const int rt = (int)(ub / a.nzp);
      tile = nfull * G + rt;
      za = (int)(ub - (long long)rt * a.nzp);
      const long long seg_end = min(u1, (long long)(rt + 1) * a.nzp);
      zb = za + (int)(seg_end - ub);
      ub = seg_end;
    }
    const int ox = (RW - 2 * S) * (tile % a.tiles_x), oy = (RH - 2 * S) * (tile / a.tiles_x);  // first output node
    const int gx = ox - S + x0, gy = oy - S + y0;  // local node coordinates of the thread's block
    // Colour of node (x,y,z) is (x+y+z+par)&1; x0, y0 are even, so the colour pattern of the 2x2
    // block depends on the plane only.  Steps are numbered K = t - t0 with (t0 + bp) even: at odd K
    // every half-sweep updates the diagonal (0,0),(1,1) of its plane, at even K the anti-diagonal.
    const int bp = (ox + oy + a.par) & 1;
    int t0 = za - S;
    t0 -= (t0 + bp) & 1;
    const int tl = zb - 1 + S;  // last plane loaded = last step
    // array coordinates of the tile corner (TMA zero-fills outside the allocation)
    const int cx = ox - S + XPAD, cy = oy - S + GHOST;

    // update masks: bit 4*(s-1)+j, j = jx + 2*jy
    uint32_t mk = 0;
#pragma unroll
    for (int s = 1; s <= S; ++s) {
      const int e = S - s;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int xx = gx + (j & 1), yy = gy + (j >> 1);
        const bool ok = xx >= a.ulo[0] - e * a.elo[0] && xx < a.uhi[0] + e * a.ehi[0] &&
                        yy >= a.ulo[1] - e * a.elo[1] && yy < a.uhi[1] + e * a.ehi[1];
        mk |= (ok ? 1u : 0u) << (4 * (s - 1) + j);
      }
    }
    // PRO: which of the block's four points take the correction (x, y), and the coarse cell under the block
    uint32_t pm = 0;
    if (PRO) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int xx = gx + (j & 1), yy = gy + (j >> 1);
        pm |= (xx >= a.plo[0] && xx < a.phi[0] && yy >= a.plo[1] && yy < a.phi[1] ? 1u : 0u) << j;
      }
    }
    // PRO: coarse planes kc0 .. kce lie under the fine planes t0 .. tl of this segment; plane kc is the
    // (kcn0 + kc - kc0)-th load of the coarse ring
    const int kc0 = t0 >> 1, kce = (tl >> 1) + (tl & 1);
    // ox - S, oy - S are even.  A TMA box has to start on a 16-byte boundary in global memory (an odd x of doubles is
    // an "illegal instruction" on B200, tools/tma_probe.cu): an odd corner (S = 2) is loaded from one node further left
    const int cxo = ((ox - S) >> 1) & 1;
    const int ccx = ((ox - S) >> 1) - cxo + XPAD, ccy = ((oy - S) >> 1) + GHOST;
    auto cissue = [&](int kc) {
      const uint32_t n = (kcn0 + (uint32_t)(kc - kc0)) & (2 * CRD - 1);
      const int sl = n & (CRD - 1);
      mbar_expect_tx(&cfull[sl], CBW * CBH * 8);
      tma_load_plane(cring + sl * CPLS, &tm_c, ccx, ccy, kc + GHOST, &cfull[sl]);
    };
    const bool xst = x0 >= S && x0 < RW - S && gx < a.n[0];
    const bool st0 = xst && y0 >= S && y0 < RH - S && gy < a.n[1];
    const bool st1 = xst && y0 >= S && y0 < RH - S && gy + 1 < a.n[1];
    char *const dst = reinterpret_cast<char *>(uout + gx) + pitch8 * gy;

    auto issue = [&](int z, uint32_t k) {
      const int sl = k & (RD - 1);
      mbar_expect_tx(&full[sl], 2 * RPLANE * 8);
      tma_load_plane(ring_u + sl * RPLANE, &tm_u, cx, cy, z + GHOST, &full[sl]);
      tma_load_plane(ring_f + sl * RPLANE, &tm_f, cx, cy, z + GHOST, &full[sl]);
    };
    if (tid == 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      for (int i = 0; i < RD - 1; ++i)
        if (t0 + i <= tl) issue(t0 + i, kt + i);
      if (PRO)
        for (int i = 0; i < CRD; ++i)
          if (kc0 + i <= kce) cissue(kc0 + i);
      for (int i = RD - 1; i < RD - 1 + a.pf; ++i)
        if (t0 + i <= tl) {
          tma_prefetch_plane(&tm_u, cx, cy, t0 + i + GHOST);
          tma_prefetch_plane(&tm_f, cx, cy, t0 + i + GHOST);
        }
    }

    if (PRO) mbar_wait(&cfull[kcn0 & (CRD - 1)], (kcn0 & (2 * CRD - 1)) >> 2);  // coarse plane kc0
    double u[NW][4], f[NW][4];  // [plane slot][jx + 2*jy]
#pragma unroll
    for (int q = 0; q < NW; ++q)
#pragma unroll
      for (int j = 0; j < 4; ++j) u[q][j] = f[q][j] = 0.0;

    auto step = [&](auto kc, const int t) {
      constexpr int K = decltype(kc)::value;
      constexpr bool DIAG = (K & 1) != 0;
      constexpr int P0 = DIAG ? 0 : 1, P1 = DIAG ? 3 : 2;  // updated points in row y0 / row y0+1
      constexpr int Q0 = DIAG ? 1 : 0, Q1 = DIAG ? 2 : 3;  // the other point of row y0 / row y0+1
      const int slot_t = kt & (RD - 1);
      mbar_wait(&full[slot_t], kt >> 2);
      {
        const double2 *pu = reinterpret_cast<const double2 *>(ring_u + slot_t * RPLANE + ld_off);
        const double2 *pf = reinterpret_cast<const double2 *>(ring_f + slot_t * RPLANE + ld_off);
        const double2 a0 = pu[0], a1 = pu[RW / 2], b0 = pf[0], b1 = pf[RW / 2];
        u[K][0] = a0.x; u[K][1] = a0.y; u[K][2] = a1.x; u[K][3] = a1.y;
        f[K][0] = b0.x; f[K][1] = b0.y; f[K][2] = b1.x; f[K][3] = b1.y;
      }
      if (PRO) {
        // Every coarse plane is waited for exactly once by every thread: plane kc0 at the start of the segment (below),
        // plane (t>>1)+1 at the odd fine plane t that first reads it.
        const uint32_t nq = (kcn0 + (uint32_t)((t >> 1) - kc0)) & (2 * CRD - 1);
        if (t & 1) mbar_wait(&cfull[(nq + 1) & (CRD - 1)], ((nq + 1) & (2 * CRD - 1)) >> 2);
        if (pm != 0u && t >= a.plo[2] && t < a.phi[2]) {
          // c<dy><dx>: coarse plane t>>1; d<dy><dx>: the plane above it (odd fine planes)
          const double *q = cring + (nq & (CRD - 1)) * CPLS + warp * CBW + lane + cxo;
          const double c00 = q[0], c01 = q[1], c10 = q[CBW], c11 = q[CBW + 1];
          double v0, v1, v2, v3;
          if ((t & 1) == 0) {
            v0 = fma(1.0, c00, u[K][0]);
            v1 = fma(0.5, c00 + c01, u[K][1]);
            v2 = fma(0.5, c00 + c10, u[K][2]);
            v3 = fma(0.25, ((c00 + c01) + c10) + c11, u[K][3]);
          } else {
            const double *qd = cring + ((nq + 1) & (CRD - 1)) * CPLS + warp * CBW + lane + cxo;
            const double d00 = qd[0], d01 = qd[1], d10 = qd[CBW], d11 = qd[CBW + 1];
            v0 = fma(0.5, c00 + d00, u[K][0]);
            v1 = fma(0.25, ((c00 + c01) + d00) + d01, u[K][1]);
            v2 = fma(0.25, ((c00 + c10) + d00) + d10, u[K][2]);
            v3 = fma(0.125, ((((((c00 + c01) + c10) + c11) + d00) + d01) + d10) + d11, u[K][3]);
          }
          if (pm == 15u) {  // interior of the block: all four points
            u[K][0] = v0; u[K][1] = v1; u[K][2] = v2; u[K][3] = v3;
          } else {
            if (pm & 1u) u[K][0] = v0;
            if (pm & 2u) u[K][1] = v1;
            if (pm & 4u) u[K][2] = v2;
            if (pm & 8u) u[K][3] = v3;
          }
        }
      }
      if (PRO || X0) {
        // the colour half-sweep 1 will read next step (it then updates the other one): this step's P0 / P1
        double *xw = xbuf + ((S - 1) * 2 + (K & 1)) * RXB + x_off;
        xw[0] = u[K][P0];
        xw[32] = u[K][P1];
      }
#pragma unroll
      for (int s = 1; s <= S; ++s) {
        const int c = (K + NW - s) % NW, cm = (K + NW - s - 1) % NW, cp = (K + NW - s + 1) % NW;
        const int pp = (K + NW - s) & 1;
        const int p = t - s, e = S - s;
        const bool zok = p >= a.ulo[2] - e * a.elo[2] && p < a.uhi[2] + e * a.ehi[2];
        if (zok) {
          // neighbours owned by other threads: row y0 -> (ex0 in x, ey0 at y0-1), row y0+1 -> (ex1, ey1 at y0+2)
          double ex0, ey0, ex1, ey1;
          if (s == 1 && (PRO || X0)) {  // the (corrected) plane t-1, published one step ago (buffer pair S-1)
            const double *xb = xbuf + ((S - 1) * 2 + pp) * RXB + x_off;
            ex0 = xb[DIAG ? -1 : 1];
            ey0 = xb[-32];
            ex1 = xb[32 + (DIAG ? 1 : -1)];
            ey1 = xb[64];
          } else if (s == 1) {  // original values: the staged plane t-1
            const double *pl = ring_u + ((slot_t + RD - 1) & (RD - 1)) * RPLANE + ld_off;
            ex0 = pl[DIAG ? -1 : 2];
            ey0 = pl[-RW + (DIAG ? 0 : 1)];
            ex1 = pl[RW + (DIAG ? 2 : -1)];
            ey1 = pl[2 * RW + (DIAG ? 1 : 0)];
          } else {  // what half-sweep s-1 published for this plane one step ago
            const double *xb = xbuf + ((s - 2) * 2 + pp) * RXB + x_off;
            ex0 = xb[DIAG ? -1 : 1];
            ey0 = xb[-32];
            ex1 = xb[32 + (DIAG ? 1 : -1)];
            ey1 = xb[64];
          }
          const double n0 = rb_update<ORD>(u[c][P0], DIAG ? ex0 : u[c][Q0], DIAG ? u[c][Q0] : ex0, ey0, u[c][Q1],
                                           u[cm][P0], u[cp][P0], f[c][P0], coef, w);
          const double n1 = rb_update<ORD>(u[c][P1], DIAG ? u[c][Q1] : ex1, DIAG ? ex1 : u[c][Q1], u[c][Q0], ey1,
                                           u[cm][P1], u[cp][P1], f[c][P1], coef, w);
          if ((mk >> (4 * (s - 1) + P0)) & 1u) u[c][P0] = n0;
          if ((mk >> (4 * (s - 1) + P1)) & 1u) u[c][P1] = n1;
        }
        if (s < S) {  // publish this colour of the plane for half-sweep s+1 (also where nothing was updated)
          double *xw = xbuf + ((s - 1) * 2 + pp) * RXB + x_off;
          xw[0] = u[c][P0];
          xw[32] = u[c][P1];
        }
      }
      // plane t-S is final
      const int zf = t - S;
      if (zf >= za && zf < zb) {
        const int c = (K + NW - S) % NW;
        char *d = dst + plane8 * zf;
        if (st0) *reinterpret_cast<double2 *>(d) = make_double2(u[c][0], u[c][1]);
        if (st1) *reinterpret_cast<double2 *>(d + pitch8) = make_double2(u[c][2], u[c][3]);
      }
      __syncthreads();
      // The slot of plane t-1 is free now: prefetch plane t+3 into it.  Every generic-proxy read of
      // the slot has returned its data before its thread arrived at the barrier above (the values
      // were consumed by this step's arithmetic), so no proxy fence is needed -- a fence here would
      // also make the TMA issue wait for this thread's outstanding global accesses.
      // PRO: after an odd fine plane the coarse plane under it is done with; its slot takes the plane CRD further up
      if (PRO && tid == 0 && (t & 1) && (t >> 1) + CRD <= kce) cissue((t >> 1) + CRD);
      if (tid == 0 && t + RD - 1 <= tl) {
        issue(t + RD - 1, kt + RD - 1);
        if (a.pf > 0 && t + RD - 1 + a.pf <= tl) {
          tma_prefetch_plane(&tm_u, cx, cy, t + RD - 1 + a.pf + GHOST);
          tma_prefetch_plane(&tm_f, cx, cy, t + RD - 1 + a.pf + GHOST);
        }
      }
      kt = (kt + 1) & (2 * RD - 1);
    };

    for (int t = t0; t <= tl; t += NW) {
      step(std::integral_constant<int, 0>{}, t);
      if (t + 1 > tl) break;
      step(std::integral_constant<int, 1>{}, t + 1);
      if (t + 2 > tl) break;
      step(std::integral_constant<int, 2>{}, t + 2);
      if (t + 3 > tl) break;
      step(std::integral_constant<int, 3>{}, t + 3);
      if (NW > 4) {
        if (t + 4 > tl) break;
        step(std::integral_constant<int, 4 % NW>{}, t + 4);
        if (t + 5 > tl) break;
        step(std::integral_constant<int, 5 % NW>{}, t + 5);
      }
    }
    if (PRO) kcn0 = (kcn0 + (uint32_t)(kce - kc0 + 1)) & (2 * CRD - 1);
  }
}

}  // namespace examg

#include "examg_jac7.cuh"
#include "examg_jacv.cuh"

namespace examg {

// ---- host side -------------------------------------------------------------------

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

static int env_flag(const char *name, int dflt) {
  const char *e = getenv(name);
  return e && e[0] ? atoi(e) : dflt;
}

// L2 promotion of the TMA loads: 0 none, 1 64 B, 2 128 B, 3 256 B
static int l2_promotion() {
  static int v = -1;
  if (v < 0) v = env_flag("EXAMG_TMA_L2PROMO", 1) & 3;
  return v;
}

// planes prefetched into L2 ahead of the shared-memory ring
static int tma_prefetch() {
  static int v = -1;
  if (v < 0) v = env_flag("EXAMG_TMA_PREFETCH", 0);
  return v;
}

static int make_plane_map(const LevelGeom &g, const double *base, CUtensorMap *tm) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return set_error(EXAMG_ECUDA, "cuTensorMapEncodeTiled entry point not available");
  const cuuint64_t dims[3] = {(cuuint64_t)g.pitch, (cuuint64_t)g.ay, (cuuint64_t)g.az};
  const cuuint64_t strides[2] = {(cuuint64_t)g.pitch * 8, (cuuint64_t)g.plane * 8};
  const cuuint32_t box[3] = {RW, RH, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<double *>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)l2_promotion(),
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(EXAMG_ECUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
  return EXAMG_OK;
}

// 0 / 1: one of the two known declaration orders; -1: something else (generic kernels)
int stencil_order(const Stencil &s) {
  static const int dx[2][7] = {{0, -1, 1, 0, 0, 0, 0}, {0, 1, -1, 0, 0, 0, 0}};
  static const int dy[2][7] = {{0, 0, 0, -1, 1, 0, 0}, {0, 0, 0, 1, -1, 0, 0}};
  static const int dz[2][7] = {{0, 0, 0, 0, 0, -1, 1}, {0, 0, 0, 0, 0, 1, -1}};
  if (s.nst != 7) return -1;
  for (int o = 0; o < 2; ++o) {
    bool same = true;
    for (int k = 0; k < 7; ++k)
      same = same && s.off[k][0] == dx[o][k] && s.off[k][1] == dy[o][k] && s.off[k][2] == dz[o][k];
    if (same) return o;
  }
  return -1;
}

static bool fused_enabled() {
  static int v = -1;
  if (v < 0) v = env_flag("EXAMG_NO_FUSED", 0) ? 0 : 1;
  return v == 1;
}

// largest number of half-sweeps fused into one pass (EXAMG_FUSE_HALF_SWEEPS = 2 or 4)
static int max_half_sweeps() {
  static int v = -1;
  if (v < 0) v = env_flag("EXAMG_FUSE_HALF_SWEEPS", 4) >= 4 ? 4 : 2;
  return v;
}

// smallest level extent for which the tiled kernel pays off (below: a few hundred KB per sweep)
constexpr int FUSED_MIN_N = 65;

bool fused_rb_usable(examg_ctx *c, const Stencil &s, const Field &fu) {
  if (!fused_enabled() || !s.is7 || c->cfg.dim != 3 || fu.ncomp != 1 || stencil_order(s) < 0) return false;
  if (c->periodic_any || fu.loc != 0) return false;  // periodic dimensions and cell fields run on the generic kernels
  const LevelGeom &g = *geom(c, s.level);
  return g.n[0] >= FUSED_MIN_N && g.n[1] >= FUSED_MIN_N && g.n[2] >= FUSED_MIN_N;
}

// Allocates the second buffer of a field (not capturable: call before graph capture).
int ensure_shadow(examg_ctx *c, Field &fu) {
  EXAMG_ONCE_PER_DEVICE(c, {
    EXAMG_CUDA(cudaFuncSetAttribute(k_rbw<2, 0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rbw_smem<2>()));
    EXAMG_CUDA(cudaFuncSetAttribute(k_rbw<2, 1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rbw_smem<2>()));
    EXAMG_CUDA(cudaFuncSetAttribute(k_rbw<4, 0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rbw_smem<4>()));
    EXAMG_CUDA(cudaFuncSetAttribute(k_rbw<4, 1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rbw_smem<4>()));
    EXAMG_CUDA(cudaFuncSetAttribute(k_rbw<2, 0, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rbw_smem<2, false, true>()));
    EXAMG_CUDA(cudaFuncSetAttribute(k_rbw<2, 1, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rbw_smem<2, false, true>()));
    EXAMG_CUDA(cudaFuncSetAttribute(k_rbw<4, 0, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rbw_smem<4, false, true>()));
    EXAMG_CUDA(cudaFuncSetAttribute(k_rbw<4, 1, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rbw_smem<4, false, true>()));
    EXAMG_CUDA(cudaFuncSetAttribute(k_rbw<2, 0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rbw_smem<2, true>()));
    EXAMG_CUDA(cudaFuncSetAttribute(k_rbw<2, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rbw_smem<2, true>()));
    EXAMG_CUDA(cudaFuncSetAttribute(k_rbw<4, 0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rbw_smem<4, true>()));
    EXAMG_CUDA(cudaFuncSetAttribute(k_rbw<4, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rbw_smem<4, true>()));
  });
  if (fu.shadow) return EXAMG_OK;
  const LevelGeom &g = *geom(c, fu.level);
  const size_t bytes = sizeof(double) * (size_t)g.vol;
  fu.shadow = fu.orig[4] = dev_alloc(c, bytes);
  if (!fu.shadow) return set_error(EXAMG_ENOMEM, "shadow buffer of %s@%d: cudaMalloc of %zu bytes failed", fu.name.c_str(), fu.level, bytes);
  EXAMG_SYNC(c);
  return EXAMG_OK;
}

// nhalf/2 full sweeps (colour 0, colour 1, ...) in one pass.  Swaps fu.slot[cur] with the shadow buffer.
static int make_box_map(const LevelGeom &g, const double *base, int ncomp, int bw, int bh, CUtensorMap *tm);

// half-sweep 1 reads its neighbours from an exchange buffer (EXAMG_RBW_X0, default on: measured)
static bool rbw_x0() {
  static int v = -1;
  if (v < 0) v = env_flag("EXAMG_RBW_X0", 1) ? 1 : 0;
  return v == 1;
}

// uc != nullptr: node (0,0,0) of Solution@coarser, whose interpolation is added to Solution on the way in (PRO)
template <int S>
static int launch_rbw(examg_ctx *c, const Stencil &s, double omega, Field &fu, const double *rhs,
                      const double *uc = nullptr) {
  const LevelGeom &g = *geom(c, s.level);
  CUtensorMap tm_u, tm_f;
  int rc = make_plane_map(g, fu.slot[fu.cur], &tm_u);
  if (rc) return rc;
  if ((rc = make_plane_map(g, rhs - g.origin, &tm_f))) return rc;
  RbwArgs a;
  for (int d = 0; d < 3; ++d) {
    a.n[d] = g.n[d];
    a.ulo[d] = g.lo[d];
    a.uhi[d] = g.hi[d];
    a.elo[d] = c->nb[d][0] ? 1 : 0;
    a.ehi[d] = c->nb[d][1] ? 1 : 0;
    a.plo[d] = g.lo[d] - S * a.elo[d];
    a.phi[d] = g.hi[d] + S * a.ehi[d];
  }
  a.cpitch = 0;
  a.cplane = 0;
  CUtensorMap tm_c = tm_u;
  if (uc) {
    const LevelGeom &gc = *geom(c, s.level - 1);
    a.cpitch = gc.pitch;
    a.cplane = gc.plane;
    if ((rc = make_box_map(gc, uc - gc.origin, 1, CBW, CBH, &tm_c))) return rc;
  }
  a.pitch = g.pitch;
  a.plane = g.plane;
  a.tiles_x = (g.n[0] + (RW - 2 * S) - 1) / (RW - 2 * S);
  const int tiles_y = (g.n[1] + (RH - 2 * S) - 1) / (RH - 2 * S);
  a.nzp = g.n[2];
  a.tiles = a.tiles_x * tiles_y;
  const long long units = (long long)a.tiles * a.nzp;
  a.par = g.par;
  a.pf = tma_prefetch();
  for (int k = 0; k < 7; ++k) a.coef[k] = s.coef[k];
  const int ord = stencil_order(s);
  a.w = omega / s.coef[s.centre];
  int dev = 0, nsm = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
  long long grid = nsm;
  // at least this many planes per CTA (a segment start costs 2S planes of warm-up): small levels are latency-bound, so
  // rather more CTAs with short segments than few with long ones (level 6: 30 -> 15 us per launch)
  const long long min_units = 2 * S;
  if (grid > units / min_units) grid = units / min_units;
  if (grid < 1) grid = 1;
  if (uc) {
    if (ord == 0)
      k_rbw<S, 0, true><<<(int)grid, RNT, rbw_smem<S, true>(), c->stream>>>(tm_u, tm_f, a, fu.shadow + g.origin, tm_c);
    else
      k_rbw<S, 1, true><<<(int)grid, RNT, rbw_smem<S, true>(), c->stream>>>(tm_u, tm_f, a, fu.shadow + g.origin, tm_c);
  } else if (rbw_x0()) {
    if (ord == 0)
      k_rbw<S, 0, false, true><<<(int)grid, RNT, rbw_smem<S, false, true>(), c->stream>>>(tm_u, tm_f, a, fu.shadow + g.origin, tm_c);
    else
      k_rbw<S, 1, false, true><<<(int)grid, RNT, rbw_smem<S, false, true>(), c->stream>>>(tm_u, tm_f, a, fu.shadow + g.origin, tm_c);
  } else if (ord == 0) {
    k_rbw<S, 0, false><<<(int)grid, RNT, rbw_smem<S>(), c->stream>>>(tm_u, tm_f, a, fu.shadow + g.origin, tm_c);
  } else {
    k_rbw<S, 1, false><<<(int)grid, RNT, rbw_smem<S>(), c->stream>>>(tm_u, tm_f, a, fu.shadow + g.origin, tm_c);
  }
  EXAMG_LAUNCH_CHECK(c);
  double *t = fu.slot[fu.cur];
  fu.slot[fu.cur] = fu.shadow;
  fu.shadow = t;
  fu.swapped = !fu.swapped;
  return EXAMG_OK;
}

// sweeps the next launch will fuse (1 or 2)
int rb_fused_plan(int nsweeps) { return nsweeps >= 2 && max_half_sweeps() >= 4 ? 2 : 1; }

// nsweeps full sweeps; returns the number of sweeps done by one launch through *done.  uc: see launch_rbw
int launch_rb_fused(examg_ctx *c, const Stencil &s, double omega, Field &fu, const double *rhs, int nsweeps,
                    int *done, const double *uc) {
  int rc = ensure_shadow(c, fu);
  if (rc) return rc;
  if (rb_fused_plan(nsweeps) == 2) {
    *done = 2;
    return launch_rbw<4>(c, s, omega, fu, rhs, uc);
  }
  *done = 1;
  return launch_rbw<2>(c, s, omega, fu, rhs, uc);
}

// the correction can ride on the first post-smoothing pass (EXAMG_NO_FUSED_PROLONG=1: separate loop)
bool fused_prolong_enabled() {
  static int v = -1;
  if (v < 0) v = env_flag("EXAMG_NO_FUSED_PROLONG", 0) ? 0 : 1;
  return v == 1;
}

// ---- Jacobi: S sweeps per pass (examg_jac7.cuh) ------------------------------------

// largest number of Jacobi sweeps fused into one pass (EXAMG_JAC_FUSE_SWEEPS = 1..3; 0: generic kernel)
static int max_jac_sweeps() {
  static int v = -1;
  if (v < 0) {
    v = env_flag("EXAMG_JAC_FUSE_SWEEPS", 3);
    v = v < 0 ? 0 : v > 3 ? 3 : v;
  }
  return v;
}

bool fused_jac_usable(examg_ctx *c, const Stencil &s, const Field &fu) {
  return max_jac_sweeps() > 0 && fu.nslots == 2 && fused_rb_usable(c, s, fu);
}

// sweeps the next launch will fuse
int jac_fused_plan(int nsweeps) {
  const int m = max_jac_sweeps();
  if (nsweeps <= m) return nsweeps;
  return nsweeps == m + 1 && m == 3 ? 2 : m;  // 4 = 2 + 2 rather than 3 + 1
}

template <int S>
static int launch_jacw(examg_ctx *c, const Stencil &s, double omega, Field &fu, const double *rhs) {
  EXAMG_ONCE_PER_DEVICE(c, {
    EXAMG_CUDA(cudaFuncSetAttribute(k_jacw<S, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)jacw_smem<S>()));
    EXAMG_CUDA(cudaFuncSetAttribute(k_jacw<S, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)jacw_smem<S>()));
  });
  const LevelGeom &g = *geom(c, s.level);
  const int nxt = (fu.cur + 1) % 2;
  CUtensorMap tm_u, tm_f;
  int rc = make_plane_map(g, fu.slot[fu.cur], &tm_u);
  if (rc) return rc;
  if ((rc = make_plane_map(g, rhs - g.origin, &tm_f))) return rc;
  RbwArgs a;
  for (int d = 0; d < 3; ++d) {
    a.n[d] = g.n[d];
    a.ulo[d] = g.lo[d];
    a.uhi[d] = g.hi[d];
    a.elo[d] = c->nb[d][0] ? 1 : 0;
    a.ehi[d] = c->nb[d][1] ? 1 : 0;
  }
  a.pitch = g.pitch;
  a.plane = g.plane;
  a.tiles_x = (g.n[0] + jacw_iw<S>() - 1) / jacw_iw<S>();
  const int tiles_y = (g.n[1] + jacw_ih<S>() - 1) / jacw_ih<S>();
  a.nzp = g.n[2];
  a.tiles = a.tiles_x * tiles_y;
  const long long units = (long long)a.tiles * a.nzp;
  a.par = g.par;
  a.pf = 0;
  for (int k = 0; k < 7; ++k) a.coef[k] = s.coef[k];
  const int ord = stencil_order(s);
  a.w = omega / s.coef[s.centre];
  int dev = 0, nsm = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
  long long grid = nsm;
  const long long min_units = 8 * S;  // at least this many planes per CTA (2S are restart overhead)
  if (grid > units / min_units) grid = units / min_units;
  if (grid < 1) grid = 1;
  if (ord == 0)
    k_jacw<S, 0><<<(int)grid, RNT, jacw_smem<S>(), c->stream>>>(tm_u, tm_f, a, fu.slot[nxt] + g.origin);
  else
    k_jacw<S, 1><<<(int)grid, RNT, jacw_smem<S>(), c->stream>>>(tm_u, tm_f, a, fu.slot[nxt] + g.origin);
  EXAMG_LAUNCH_CHECK(c);
  if (S & 1) {
    fu.cur = nxt;  // advance Solution an odd number of times
  } else {
    // an even number of `advance`s: the active slot stays, the result sits in the other buffer.
    // Exchange the two buffers (Solution<next> then holds the input of the pass, not the last-but-one sweep).
    double *t = fu.slot[fu.cur];
    fu.slot[fu.cur] = fu.slot[nxt];
    fu.slot[nxt] = t;
  }
  return EXAMG_OK;
}

// up to nsweeps Jacobi sweeps in one launch; *done = sweeps done
int launch_jac_fused(examg_ctx *c, const Stencil &s, double omega, Field &fu, const double *rhs, int nsweeps,
                     int *done) {
  const int plan = jac_fused_plan(nsweeps);
  *done = plan;
  switch (plan) {
    case 1: return launch_jacw<1>(c, s, omega, fu, rhs);
    case 2: return launch_jacw<2>(c, s, omega, fu, rhs);
    default: return launch_jacw<3>(c, s, omega, fu, rhs);
  }
}

// ---- stencil-field Jacobi: two sweeps per pass, coefficients read once (examg_jacv.cuh) ----------

// 3D view of a field with `ncomp` components stacked in z (component k, plane z = plane k * az + z), box bw x bh x 1
static int make_box_map(const LevelGeom &g, const double *base, int ncomp, int bw, int bh, CUtensorMap *tm) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return set_error(EXAMG_ECUDA, "cuTensorMapEncodeTiled entry point not available");
  const cuuint64_t dims[3] = {(cuuint64_t)g.pitch, (cuuint64_t)g.ay, (cuuint64_t)g.az * ncomp};
  const cuuint64_t strides[2] = {(cuuint64_t)g.pitch * 8, (cuuint64_t)g.plane * 8};
  const cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<double *>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)l2_promotion(),
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(EXAMG_ECUDA, "cuTensorMapEncodeTiled (box %d x %d) failed (%d)", bw, bh, (int)r);
  return EXAMG_OK;
}

// 27 lexicographic (dx fastest): 2; 7-point in declaration order 0 / 1: 0 / 1; anything else: -1
static int jacv_shape(const Stencil &s) {
  if (s.nst == 7) return stencil_order(s);
  if (s.nst != 27) return -1;
  for (int k = 0; k < 27; ++k)
    if (s.off[k][0] != k % 3 - 1 || s.off[k][1] != (k / 3) % 3 - 1 || s.off[k][2] != k / 9 - 1) return -1;
  return 2;
}

// EXAMG_JACV_FUSE: 0 = generic kernels only, 1 = 27-point (default), 2 = 27- and 7-point
static int jacv_mode() {
  static int v = -1;
  if (v < 0) v = env_flag("EXAMG_JACV_FUSE", 1);
  return v;
}

constexpr int JACV_MIN_N = 33;

bool fused_jacv_usable(examg_ctx *c, const Stencil &s, const Field &fu) {
  if (!fused_enabled() || jacv_mode() <= 0 || c->cfg.dim != 3 || s.coeff_field < 0 || fu.ncomp != 1 ||
      fu.nslots != 2 || c->periodic_any || fu.loc != 0)
    return false;
  const int shape = jacv_shape(s);
  if (shape < 0 || (shape < 2 && jacv_mode() < 2)) return false;
  const LevelGeom &g = *geom(c, s.level);
  return g.n[0] >= JACV_MIN_N && g.n[1] >= JACV_MIN_N && g.n[2] >= JACV_MIN_N;
}

template <int NST, int ORD>
static int launch_jacv(examg_ctx *c, const Stencil &s, double omega, Field &fu, const double *rhs) {
  EXAMG_ONCE_PER_DEVICE(c, {
    EXAMG_CUDA(cudaFuncSetAttribute(k_jacv2<NST, ORD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)jacv_smem<NST>()));
  });
  const LevelGeom &g = *geom(c, s.level);
  const Field &cf = c->fields[s.coeff_field];
  const int nxt = (fu.cur + 1) % 2;
  CUtensorMap tm_u, tm_c;
  int rc = make_box_map(g, fu.slot[fu.cur], 1, VUW, VUH, &tm_u);
  if (rc) return rc;
  if ((rc = make_box_map(g, cf.slot[0], NST, VCW, VH, &tm_c))) return rc;
  RbwArgs a;
  for (int d = 0; d < 3; ++d) {
    a.n[d] = g.n[d];
    a.ulo[d] = g.lo[d];
    a.uhi[d] = g.hi[d];
    a.elo[d] = c->nb[d][0] ? 1 : 0;  // sweep 1 also runs on the first layer of the neighbour blocks' nodes
    a.ehi[d] = c->nb[d][1] ? 1 : 0;
  }
  a.pitch = g.pitch;
  a.plane = g.plane;
  a.tiles_x = (g.n[0] + VIW - 1) / VIW;
  const int tiles_y = (g.n[1] + VIH - 1) / VIH;
  a.nzp = g.n[2];
  a.tiles = a.tiles_x * tiles_y;
  const long long units = (long long)a.tiles * a.nzp;
  a.par = g.az;  // planes per component of the stacked view
  a.pf = 0;
  for (int k = 0; k < 7; ++k) a.coef[k] = 0.0;
  a.w = omega;  // the diagonal varies: omega / diag is formed per point
  int dev = 0, nsm = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
  long long grid = nsm;
  const long long min_units = 12;  // at least this many planes per CTA (2 are restart overhead)
  if (grid > units / min_units) grid = units / min_units;
  if (grid < 1) grid = 1;
  k_jacv2<NST, ORD><<<(int)grid, VNT, jacv_smem<NST>(), c->stream>>>(tm_u, tm_c, a, rhs, fu.slot[nxt] + g.origin);
  EXAMG_LAUNCH_CHECK(c);
  // two `advance`s: the active slot stays, the result sits in the other buffer -> exchange the two buffers
  double *t = fu.slot[fu.cur];
  fu.slot[fu.cur] = fu.slot[nxt];
  fu.slot[nxt] = t;
  return EXAMG_OK;
}

// exactly two Jacobi sweeps with stencil-field coefficients in one launch
int launch_jacv_fused(examg_ctx *c, const Stencil &s, double omega, Field &fu, const double *rhs) {
  switch (jacv_shape(s)) {
    case 2: return launch_jacv<27, 0>(c, s, omega, fu, rhs);
    case 0: return launch_jacv<7, 0>(c, s, omega, fu, rhs);
    case 1: return launch_jacv<7, 1>(c, s, omega, fu, rhs);
    default: return set_error(EXAMG_EUNSUP, "launch_jacv_fused: stencil shape not covered");
  }
}

}  // namespace examg
