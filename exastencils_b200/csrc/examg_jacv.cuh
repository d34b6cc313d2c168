// examg_jacv.cuh -- damped Jacobi with stencil-field (variable) coefficients, TWO sweeps per pass over HBM,
// for the 27-point and the 7-point operator (included by examg_fast7.cu: TMA / mbarrier helpers).
//
// Reference loops being replaced:
//   Solution<next> = Solution<active> + omega / diag(A) * (RHS - A * Solution<active>); advance
//                                 Testing/SISC/3D_VarCoeff.exa4 `Smoother`, stencil-field convolution
//                                 C/s/e/stencil/ir/IR_StencilConvolution.scala:73-95
//   repeat nu times with contraction [1,1,1]   C/s/e/baseExt/ir/IR_ContractingLoop.scala:131-195
// A sweep reads NST coefficients + Solution + RHS and writes one value per point: 240 B (27-point) / 80 B
// (7-point), and the generic kernel streams that at 97-99 % of the copy roofline.  The only way to go faster is
// to read the coefficients ONCE for two sweeps.  A first attempt (profiles/r01_notes.md, `k_jacvw`) re-read them
// through L2 one plane-step later and lost: the reuse distance was larger than L2.  Here the second use comes
// out of REGISTERS:
//
//   * a persistent CTA (512 threads, one per SM) owns an x-y tile of 32 x 16 nodes and marches in z; a thread
//     owns one node.  The tile is the region of sweep 1; sweep 2 (the output) is its 30 x 14 interior.
//   * per plane NST TMA boxes of 34 x 16 (117 KB for NST = 27; the stencil field is one 3D tensor with the
//     components stacked in z; a single 4D box raised "illegal instruction" on B200) bring the
//     coefficients into a shared-memory landing buffer; every thread copies its NST values into registers,
//     the CTA synchronises and the box of the next plane is issued at once -- it is in flight during the whole
//     arithmetic of this plane (110 KB per SM outstanding, independent of occupancy).  Solution planes
//     (34 x 18, halo for sweep 1) come through a 4-deep TMA ring.
//   * step t: sweep 1 on plane t from the staged Solution planes t-1, t, t+1 -> shared-memory ring of sweep-1
//     planes; barrier; sweep 2 on plane t-1 is FINISHED and stored, sweep 2 on plane t is BEGUN:
//     A*u is a left fold over the entries in declaration order, so the entries that only touch sweep-1 planes
//     t-1 and t (a prefix: KS = 18 of 27 for the lexicographic 27-point order) are folded now, while the
//     plane's coefficients are still in registers, and only the partial sum, the remaining R = NST - KS
//     coefficients, omega/diag, RHS and the centre value (R + 4 doubles) are kept for the next step.
//     Same association as the generic kernel => bit-identical to two separate sweeps.
// HBM traffic per output point and pass: 8 NST / 0.82 (tile halo; less where the neighbouring tile's box is
// still in L2) + 8 * 1.46 (Solution) + 8 / 0.82 (RHS) + 8 (store) = 293 B for 27 points against 2 * 240 B.
#pragma once

namespace examg {

constexpr int VW = 32, VH = 16, VNT = VW * VH;  // sweep-1 region of a tile = thread grid
constexpr int VIW = VW - 2, VIH = VH - 2;       // output region
constexpr int VUW = VW + 2, VUH = VH + 2;       // staged Solution plane
constexpr int VUPL = 640;                       // doubles per ring slot (34 * 18 = 612, rounded to 128 B)
constexpr int VUD = 4;                          // Solution ring depth
constexpr int VOFF_U = 1024;                    // bytes; mbarriers in front
constexpr int VOFF_X = VOFF_U + VUD * VUPL * 8; // sweep-1 ring: 3 planes of 32 x 16
constexpr int VOFF_C = VOFF_X + 3 * VNT * 8;    // coefficient landing buffer [NST][16][34]
// A TMA box must start on a 16-byte boundary of the row (a box at an odd x raised "illegal instruction" on B200):
// the coefficient boxes start one node left of the sweep-1 region, like the Solution planes, and are 34 wide.
constexpr int VCW = VW + 2, VCPL = VCW * VH;    // doubles per component in the landing buffer
template <int NST>
constexpr size_t jacv_smem() { return VOFF_C + (size_t)NST * VCPL * 8; }

// stencil shapes: NST = 27 lexicographic (dx fastest, dz slowest); NST = 7 in the two declaration orders of rb_update
template <int NST, int ORD>
struct JvShape {
  static constexpr int centre = NST == 27 ? 13 : 0;
  // first entry that touches plane p+1: everything before it can be folded one step early
  static constexpr int ksplit = NST == 27 ? 18 : (ORD == 0 ? 6 : 5);
  __host__ __device__ static constexpr int dx(int k) {
    return NST == 27 ? k % 3 - 1 : (k >= 1 && k <= 2 ? sgn(k) : 0);
  }
  __host__ __device__ static constexpr int dy(int k) {
    return NST == 27 ? (k / 3) % 3 - 1 : (k >= 3 && k <= 4 ? sgn(k) : 0);
  }
  __host__ __device__ static constexpr int dz(int k) {
    return NST == 27 ? k / 9 - 1 : (k >= 5 ? sgn(k) : 0);
  }
  __host__ __device__ static constexpr int sgn(int k) { return (((k - 1) & 1) == 0) == (ORD == 0) ? -1 : 1; }
};

template <int NST, int ORD>
__global__ void __launch_bounds__(VNT, 1)
k_jacv2(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_c,
        const __grid_constant__ RbwArgs a, const double *__restrict__ rhs, double *__restrict__ uout) {
  using SH = JvShape<NST, ORD>;
  constexpr int KS = SH::ksplit, R = NST - KS, CEN = SH::centre;
  constexpr uint32_t CBYTES = NST * VCPL * 8, UBYTES = VUW * VUH * 8;
  extern __shared__ __align__(1024) unsigned char smem[];
  uint64_t *bar = reinterpret_cast<uint64_t *>(smem);  // two, alternating per step
  double *ring_u = reinterpret_cast<double *>(smem + VOFF_U);
  double *x1 = reinterpret_cast<double *>(smem + VOFF_X);
  double *land = reinterpret_cast<double *>(smem + VOFF_C);

  const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5, lane = tx, warp = ty;
  const int az = a.par;  // allocated planes per component: component k, plane z is plane k * az + z of the 3D view
  if (tid == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const double omega = a.w;
  const bool interior = tx >= 1 && tx < VW - 1 && ty >= 1 && ty < VH - 1;
  const int uoff = (ty + 1) * VUW + tx + 1;  // own node inside a staged Solution plane
  const int xoff = ty * VW + tx;             // own node inside a sweep-1 plane
  const int coff = ty * VCW + tx + 1;        // own node inside a component of the landing buffer

  const int G = gridDim.x, bid = blockIdx.x;
  const int nfull = a.tiles / G;
  const long long rem_units = (long long)(a.tiles - nfull * G) * a.nzp;
  const long long u1 = rem_units * (bid + 1) / G;
  long long ub = rem_units * bid / G;
  uint32_t kt = 0;  // steps so far modulo 4: barrier kt & 1, phase (kt >> 1) & 1

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
    const int ox = VIW * (tile % a.tiles_x), oy = VIH * (tile / a.tiles_x);  // first output node
    const int gx = ox - 1 + tx, gy = oy - 1 + ty;
    const int t0 = za - 1, tl = zb;  // steps: sweep 1 on planes za-1 .. zb
    const int cxu = ox - 2 + XPAD, cyu = oy - 2 + GHOST;  // array coordinates of the box corners
    const int cxc = ox - 2 + XPAD, cyc = oy - 1 + GHOST;  // (TMA zero-fills outside the allocation)
    const bool m1 = gx >= a.ulo[0] - a.elo[0] && gx < a.uhi[0] + a.ehi[0] && gy >= a.ulo[1] - a.elo[1] &&
                    gy < a.uhi[1] + a.ehi[1];
    const bool m2 = interior && gx >= a.ulo[0] && gx < a.uhi[0] && gy >= a.ulo[1] && gy < a.uhi[1];
    const bool st = interior && gx < a.n[0] && gy < a.n[1];
    const long long pofs = gx + (long long)a.pitch * gy;  // dereferenced under m1 / st only
    const int z1lo = a.ulo[2] - a.elo[2], z1hi = a.uhi[2] + a.ehi[2];

    if (tid == 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      uint64_t *b = &bar[kt & 1];
      mbar_expect_tx(b, CBYTES + 3 * UBYTES);
      for (int p = t0 - 1; p <= t0 + 1; ++p)
        tma_load_plane(ring_u + ((p + 2) & (VUD - 1)) * VUPL, &tm_u, cxu, cyu, p + GHOST, b);
    }
    __syncthreads();  // the other warps' loads must follow the fence and the expect_tx
    if (lane == 0)
      for (int k = warp; k < NST; k += VH)
        tma_load_plane(land + k * VCPL, &tm_c, cxc, cyc, k * az + t0 + GHOST, &bar[kt & 1]);
    double fn = (m1 && t0 >= z1lo && t0 < z1hi) ? rhs[pofs + a.plane * t0] : 0.0;

    // sweep 2 of the previous plane, half done: partial sum, omega/diag, RHS, centre value, last R coefficients
    double acc2 = 0.0, wprev = 0.0, fprev = 0.0, u1prev = 0.0, cr[R];
#pragma unroll
    for (int i = 0; i < R; ++i) cr[i] = 0.0;

    for (int t = t0; t <= tl; ++t) {
      mbar_wait(&bar[kt & 1], (kt >> 1) & 1);
      double c[NST];
#pragma unroll
      for (int k = 0; k < NST; ++k) c[k] = land[k * VCPL + coff];
      const double fcur = fn;
      __syncthreads();  // the landing buffer is free: the box of plane t+1 flies during this plane's arithmetic
      if (tid == 0 && t < tl) {
        uint64_t *b = &bar[(kt + 1) & 1];
        mbar_expect_tx(b, CBYTES + UBYTES);
        tma_load_plane(ring_u + ((t + 4) & (VUD - 1)) * VUPL, &tm_u, cxu, cyu, t + 2 + GHOST, b);
      }
      if (lane == 0 && t < tl)  // one warp brings at most two components: the issue cost is spread over the CTA
        for (int k = warp; k < NST; k += VH)
          tma_load_plane(land + k * VCPL, &tm_c, cxc, cyc, k * az + t + 1 + GHOST, &bar[(kt + 1) & 1]);
      if (t < tl) fn = (m1 && t + 1 >= z1lo && t + 1 < z1hi) ? rhs[pofs + a.plane * (t + 1)] : 0.0;

      // sweep 1 on plane t
      const double *pu[3] = {ring_u + ((t + 1) & (VUD - 1)) * VUPL + uoff, ring_u + ((t + 2) & (VUD - 1)) * VUPL + uoff,
                             ring_u + ((t + 3) & (VUD - 1)) * VUPL + uoff};
      const double wcur = omega / c[CEN];
      double un;
      {
        double acc = c[0] * pu[SH::dz(0) + 1][SH::dy(0) * VUW + SH::dx(0)];
#pragma unroll
        for (int k = 1; k < NST; ++k) acc = fma(c[k], pu[SH::dz(k) + 1][SH::dy(k) * VUW + SH::dx(k)], acc);
        const double uc = pu[1][0];
        un = (m1 && t >= z1lo && t < z1hi) ? fma(wcur, fcur - acc, uc) : uc;
      }
      const int s_m = (t + 2) % 3, s_c = (t + 3) % 3;  // sweep-1 ring slots of planes t-1, t  (t >= -1)
      x1[s_c * VNT + xoff] = un;
      __syncthreads();
      if (interior) {
        const double *px[3] = {x1 + ((t + 1) % 3) * VNT + xoff, x1 + s_m * VNT + xoff, x1 + s_c * VNT + xoff};
        {  // finish sweep 2 on plane t-1 (planes t-2, t-1, t of sweep 1 are px[0..2])
          const int p = t - 1;
          double acc = acc2;
#pragma unroll
          for (int k = KS; k < NST; ++k) acc = fma(cr[k - KS], px[SH::dz(k) + 1][SH::dy(k) * VW + SH::dx(k)], acc);
          const double out = (m2 && p >= a.ulo[2] && p < a.uhi[2]) ? fma(wprev, fprev - acc, u1prev) : u1prev;
          if (st && p >= za && p < zb) uout[pofs + a.plane * p] = out;
        }
        {  // begin sweep 2 on plane t (planes t-1, t of sweep 1 are px[1], px[2])
          double acc = c[0] * px[SH::dz(0) + 2][SH::dy(0) * VW + SH::dx(0)];
#pragma unroll
          for (int k = 1; k < KS; ++k) acc = fma(c[k], px[SH::dz(k) + 2][SH::dy(k) * VW + SH::dx(k)], acc);
          acc2 = acc;
#pragma unroll
          for (int i = 0; i < R; ++i) cr[i] = c[KS + i];
          wprev = wcur;
          fprev = fcur;
          u1prev = un;
        }
      }
      kt = (kt + 1) & 3;
    }
  }
}

}  // namespace examg
