// examg_jac7.cuh -- 3D 7-point constant-coefficient damped Jacobi, S sweeps per pass over HBM
// (included by examg_fast7.cu, which provides the TMA / mbarrier helpers and the tile constants).
//
// Reference loops being replaced:
//   repeat nu times with contraction [1,1,1] { loop over Solution { Solution<next> = Solution<active> +
//   omega/diag(A) * (RHS - A*Solution<active>) }; advance Solution }
//                                            Testing/Opts/tempBlock.exa4:144-153, Testing/SISC/3D_ConstCoeff.exa4
//   C/s/e/baseExt/ir/IR_ContractingLoop.scala:48-90,131-195 (the reference widens the bounds per sweep but
//   still makes nu passes over memory; here nu sweeps are ONE pass)
//   slots / advance                          C/s/e/field/ir/IR_Slot.scala:33-62
//
// Same register-window wavefront as k_rbw: a persistent CTA owns a 64 x 32 tile (S-deep halo included)
// and marches in z; TMA stages plane t of Solution<active> and RHS; a thread owns a 2 x 2 block.  Sweep s
// works on plane t-s and needs sweep s-1 on planes t-s-1..t-s+1: the z neighbours are the thread's own
// registers (3-plane window per sweep), two in-plane neighbours per point are its own block, the other two
// come from a full-plane exchange buffer in shared memory that sweep s-1 filled one step earlier
// (double-buffered, one __syncthreads per plane; sweep 1 reads the staged plane itself).
// Jacobi is out of place per sweep, so every sweep produces all four points of the block.
// Arithmetic: the contract of examg_kernels.cu (u' = fma(omega/diag, rhs - A u, u), A u in declaration
// order), i.e. bit-identical to S separate sweeps.
#pragma once

namespace examg {

constexpr int JXB = (RH + 2) * RW + 2 * RW;  // doubles per exchange buffer: rows -1..RH of 64, one pad row each end

template <int S>
constexpr size_t jacw_smem() { return ROFF_RING + sizeof(double) * (2 * RD * RPLANE + (S - 1) * 2 * JXB); }
// x halo rounded up to even so that the 128-bit stores stay aligned; interior (= output) extent of a tile
template <int S> __host__ __device__ constexpr int jacw_hx() { return S + (S & 1); }
template <int S> __host__ __device__ constexpr int jacw_iw() { return RW - 2 * jacw_hx<S>(); }
template <int S> __host__ __device__ constexpr int jacw_ih() { return RH - 2 * S; }

// fn(integral_constant<K>) for K = 0, 1, ... until it returns false
template <int... Ks, class F>
__device__ __forceinline__ bool jacw_steps(std::integer_sequence<int, Ks...>, F &&fn) {
  return (fn(std::integral_constant<int, Ks>{}) && ...);
}

template <int S, int ORD>
__global__ void __launch_bounds__(RNT, 1)
k_jacw(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_f,
       const __grid_constant__ RbwArgs a, double *__restrict__ uout) {
  static_assert(S >= 1 && S <= 3, "S sweeps");
  constexpr int FWN = S == 3 ? 4 : 3;       // planes in the RHS register window (>= S+1)
  constexpr int PER = S == 3 ? 12 : 3;      // steps unrolled (lcm of the window sizes): compile-time register slots
  constexpr int HX = jacw_hx<S>(), IW = jacw_iw<S>(), IH = jacw_ih<S>();
  extern __shared__ __align__(1024) unsigned char smem[];
  uint64_t *full = reinterpret_cast<uint64_t *>(smem);
  double *ring_u = reinterpret_cast<double *>(smem + ROFF_RING);
  double *ring_f = ring_u + RD * RPLANE;
  double *xbuf = ring_f + RD * RPLANE;  // [(s-1)*2 + plane parity][JXB], s = 1..S-1

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int x0 = 2 * lane, y0 = 2 * warp;
  if (tid == 0) {
    for (int i = 0; i < RD; ++i) mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  double coef[7];
#pragma unroll
  for (int k = 0; k < 7; ++k) coef[k] = a.coef[k];
  const double w = a.w;
  const long long plane8 = a.plane * 8;
  const long long pitch8 = (long long)a.pitch * 8;
  const int ld_off = y0 * RW + x0;                 // (x0, y0) inside a staged plane
  const int x_off = 2 * RW + y0 * RW + x0;         // (x0, y0) inside an exchange buffer (row -1 at RW)

  const int G = gridDim.x, bid = blockIdx.x;
  const int nfull = a.tiles / G;
  const long long rem_units = (long long)(a.tiles - nfull * G) * a.nzp;
  const long long u1 = rem_units * (bid + 1) / G;
  long long ub = rem_units * bid / G;
  uint32_t kt = 0;

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
    const int ox = IW * (tile % a.tiles_x), oy = IH * (tile / a.tiles_x);  // first output node
    const int gx = ox - HX + x0, gy = oy - S + y0;
    const int t0 = za - S, tl = zb - 1 + S;
    const int cx = ox - HX + XPAD, cy = oy - S + GHOST;

    // update masks: bit 4*(s-1)+j, j = jx + 2*jy; sweep s also updates S-s layers of the neighbour blocks' nodes
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
    const bool xst = x0 >= HX && x0 < RW - HX && gx < a.n[0];
    const bool st0 = xst && y0 >= S && y0 < RH - S && gy < a.n[1];
    const bool st1 = xst && y0 + 1 >= S && y0 + 1 < RH - S && gy + 1 < a.n[1];
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
    }

    // v[s][q][j]: result of sweep s (0 = input) on plane slot q (3-plane window); f: RHS window
    double v[S][3][4], f[FWN][4];
#pragma unroll
    for (int s = 0; s < S; ++s)
#pragma unroll
      for (int q = 0; q < 3; ++q)
#pragma unroll
        for (int j = 0; j < 4; ++j) v[s][q][j] = 0.0;
#pragma unroll
    for (int q = 0; q < FWN; ++q)
#pragma unroll
      for (int j = 0; j < 4; ++j) f[q][j] = 0.0;

    auto step = [&](auto kc, const int t) {
      constexpr int K = decltype(kc)::value;  // (t - t0) mod PER
      const int slot_t = kt & (RD - 1);
      mbar_wait(&full[slot_t], kt >> 2);
      {
        const double2 *pu = reinterpret_cast<const double2 *>(ring_u + slot_t * RPLANE + ld_off);
        const double2 *pf = reinterpret_cast<const double2 *>(ring_f + slot_t * RPLANE + ld_off);
        const double2 a0 = pu[0], a1 = pu[RW / 2], b0 = pf[0], b1 = pf[RW / 2];
        v[0][K % 3][0] = a0.x; v[0][K % 3][1] = a0.y; v[0][K % 3][2] = a1.x; v[0][K % 3][3] = a1.y;
        f[K % FWN][0] = b0.x; f[K % FWN][1] = b0.y; f[K % FWN][2] = b1.x; f[K % FWN][3] = b1.y;
      }
      double outv[4];
#pragma unroll
      for (int s = 1; s <= S; ++s) {
        // plane p = t - s of sweep s-1 sits in slot (K-s) mod 3 of v[s-1], its z neighbours next to it
        const int c = (K + 3 * PER - s) % 3, cm = (K + 3 * PER - s - 1) % 3, cp = (K + 3 * PER - s + 1) % 3;
        const int fs = (K + FWN * 4 - s) % FWN;
        const int p = t - s, e = S - s;
        const int par = (PER & 1) ? (p & 1) : ((K + 2 * PER - s) & 1);  // exchange buffer of plane p
        const bool zok = p >= a.ulo[2] - e * a.elo[2] && p < a.uhi[2] + e * a.ehi[2];
        const double (&in)[4] = v[s - 1][c];
        double nv[4] = {in[0], in[1], in[2], in[3]};
        if (zok) {
          // in-plane neighbours owned by other threads, from the plane as sweep s-1 left it
          const double *pl = s == 1 ? ring_u + ((slot_t + RD - 1) & (RD - 1)) * RPLANE + ld_off
                                    : xbuf + ((s - 2) * 2 + par) * JXB + x_off;
          const double xm0 = pl[-1], xp1 = pl[2], xm2 = pl[RW - 1], xp3 = pl[RW + 2];
          const double2 ym = *reinterpret_cast<const double2 *>(pl - RW);      // (x0, y0-1), (x0+1, y0-1)
          const double2 yp = *reinterpret_cast<const double2 *>(pl + 2 * RW);  // (x0, y0+2), (x0+1, y0+2)
          const double (&zm)[4] = v[s - 1][cm];
          const double (&zp)[4] = v[s - 1][cp];
          const double n0 = rb_update<ORD>(in[0], xm0, in[1], ym.x, in[2], zm[0], zp[0], f[fs][0], coef, w);
          const double n1 = rb_update<ORD>(in[1], in[0], xp1, ym.y, in[3], zm[1], zp[1], f[fs][1], coef, w);
          const double n2 = rb_update<ORD>(in[2], xm2, in[3], in[0], yp.x, zm[2], zp[2], f[fs][2], coef, w);
          const double n3 = rb_update<ORD>(in[3], in[2], xp3, in[1], yp.y, zm[3], zp[3], f[fs][3], coef, w);
          if ((mk >> (4 * (s - 1) + 0)) & 1u) nv[0] = n0;
          if ((mk >> (4 * (s - 1) + 1)) & 1u) nv[1] = n1;
          if ((mk >> (4 * (s - 1) + 2)) & 1u) nv[2] = n2;
          if ((mk >> (4 * (s - 1) + 3)) & 1u) nv[3] = n3;
        }
        if (s < S) {
          // keep for sweep s+1 (slot of plane p in v[s]) and publish the whole block for the neighbours
#pragma unroll
          for (int j = 0; j < 4; ++j) v[s][c][j] = nv[j];
          double *xw = xbuf + ((s - 1) * 2 + par) * JXB + x_off;
          *reinterpret_cast<double2 *>(xw) = make_double2(nv[0], nv[1]);
          *reinterpret_cast<double2 *>(xw + RW) = make_double2(nv[2], nv[3]);
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) outv[j] = nv[j];
        }
      }
      const int zf = t - S;
      if (zf >= za && zf < zb) {
        char *d = dst + plane8 * zf;
        if (st0) *reinterpret_cast<double2 *>(d) = make_double2(outv[0], outv[1]);
        if (st1) *reinterpret_cast<double2 *>(d + pitch8) = make_double2(outv[2], outv[3]);
      }
      __syncthreads();
      if (tid == 0 && t + RD - 1 <= tl) issue(t + RD - 1, kt + RD - 1);  // no proxy fence needed, see k_rbw
      kt = (kt + 1) & (2 * RD - 1);
    };

    for (int t = t0; t <= tl; t += PER) {
      const bool more = jacw_steps(std::make_integer_sequence<int, PER>{}, [&](auto kc) {
        constexpr int K = decltype(kc)::value;
        if (t + K > tl) return false;
        step(kc, t + K);
        return true;
      });
      if (!more) break;
    }
  }
}

}  // namespace examg
