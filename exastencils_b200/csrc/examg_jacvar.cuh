// examg_jacvar.cuh -- damped Jacobi with stencil-field (variable) coefficients, TWO sweeps per pass over HBM,
// for the 7-point and the 27-point operator (included by examg_fast7.cu: TMA / mbarrier helpers, tile constants).
//
// Reference loops being replaced:
//   Solution<next> = Solution<active> + omega / diag(A) * (RHS - A * Solution<active>); advance
//                                 Testing/SISC/3D_VarCoeff.exa4 `Smoother`, stencil-field convolution
//                                 C/s/e/stencil/ir/IR_StencilConvolution.scala:73-95
//   repeat nu times with contraction [1,1,1]   C/s/e/baseExt/ir/IR_ContractingLoop.scala:131-195
// A sweep reads NST coefficients + Solution + RHS and writes one value per point: 80 B (7-point) / 240 B
// (27-point); the generic kernel already streams that at 97-99 % of the copy roofline, so the only way to go
// faster is to read the coefficients once for two sweeps.
//
// Design: the tile / TMA ring / wave schedule of k_rbw (64 x 32 nodes with a 2-deep halo, z-marching, one
// persistent CTA per SM).  At step t sweep 1 runs on plane t-1 from the staged planes t-2, t-1, t of
// Solution (all neighbours from shared memory), writes its plane into a 3-plane ring of sweep-1 results;
// after a barrier sweep 2 runs on plane t-2 from that ring and stores to the other slot.  Coefficients are
// read straight from global memory with 128-bit loads by both sweeps, one plane-step apart: the second
// read is served by L2 (reuse distance = one plane of all tiles, about 65 MB of 126 MB).
// HBM traffic per point and pass: (8 NST + 16) / 0.82 + 8 bytes for two sweeps.
// Arithmetic: the contract of examg_kernels.cu -- acc = c0*u0, fma(ck, uk, acc) in declaration order,
// u' = fma(omega / c_centre, rhs - acc, u) -- so results are bit-identical to two separate sweeps.
#pragma once

namespace examg {

constexpr int JV_S = 2;          // sweeps per pass
constexpr int JV_XPL = RPLANE;   // doubles per plane of the sweep-1 ring

constexpr size_t jacv_smem() { return ROFF_RING + sizeof(double) * (2 * RD * RPLANE + 3 * JV_XPL) + 1024; }

struct JvArgs {
  RbwArgs r;
  const double *sf;   // coefficient field, component k at sf + k * sfvol, node (0,0,0) at sf
  long long sfvol;
  int centre;
};

// offsets in declaration order; ORD as in rb_update for NST = 7, lexicographic (dx fastest) for NST = 27
template <int NST, int ORD>
__device__ __forceinline__ void jv_offset(int k, int &dx, int &dy, int &dz) {
  if (NST == 27) {
    dx = k % 3 - 1; dy = (k / 3) % 3 - 1; dz = k / 9 - 1;
  } else {
    const int axis = (k - 1) >> 1, minus_first = ORD == 0;
    const int sgn = k == 0 ? 0 : ((((k - 1) & 1) == 0) == (minus_first != 0) ? -1 : 1);
    dx = axis == 0 ? sgn : 0; dy = axis == 1 ? sgn : 0; dz = axis == 2 ? sgn : 0;
    if (k == 0) dx = dy = dz = 0;
  }
}

// L2 residency control: the first read of a coefficient plane asks L2 to keep the lines (evict_last), the
// second -- one plane-step later -- marks them for early eviction.  Without it the 65 MB between the two
// reads push the coefficients out again (measured: L2 hit rate 2 %, 650 B of DRAM traffic per point and pass).
__device__ __forceinline__ uint64_t jv_policy(bool keep) {
  uint64_t p;
  if (keep) asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  else asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ double2 jv_ldg(const double *ptr, uint64_t pol) {
  double2 v;
  asm("ld.global.nc.L2::cache_hint.v2.f64 {%0, %1}, [%2], %3;" : "=d"(v.x), "=d"(v.y) : "l"(ptr), "l"(pol));
  return v;
}

// One sweep on the thread's 2x2 block of plane p.  pm/pc/pp: shared-memory planes p-1, p, p+1 of the input
// (row stride RW), pointing at the block's first node; cf: coefficients at the block's first node of plane p.
template <int NST, int ORD>
__device__ __forceinline__ void jv_sweep(const double *pm, const double *pc, const double *pp, const double *cf,
                                         long long sfvol, int pitch, uint64_t pol, double omega, const double (&f)[4],
                                         double (&out)[4]) {
  constexpr int centre = NST == 27 ? 13 : 0;
  double acc[4];
  double ctr[4];  // centre coefficients
#pragma unroll
  for (int k = 0; k < NST; ++k) {
    int dx, dy, dz;
    jv_offset<NST, ORD>(k, dx, dy, dz);
    const double *pl = dz < 0 ? pm : (dz > 0 ? pp : pc);
    const double2 c0 = jv_ldg(cf + k * sfvol, pol);
    const double2 c1 = jv_ldg(cf + k * sfvol + pitch, pol);
    // neighbours of the four points: rows y0+dy and y0+1+dy, columns x0+dx and x0+1+dx
    const double *r0 = pl + dy * RW + dx, *r1 = r0 + RW;
    double u00, u10, u01, u11;
    if (dx == 0) {
      const double2 a = *reinterpret_cast<const double2 *>(r0), b = *reinterpret_cast<const double2 *>(r1);
      u00 = a.x; u10 = a.y; u01 = b.x; u11 = b.y;
    } else {
      u00 = r0[0]; u10 = r0[1]; u01 = r1[0]; u11 = r1[1];
    }
    if (k == 0) {
      acc[0] = c0.x * u00; acc[1] = c0.y * u10; acc[2] = c1.x * u01; acc[3] = c1.y * u11;
    } else {
      acc[0] = fma(c0.x, u00, acc[0]); acc[1] = fma(c0.y, u10, acc[1]);
      acc[2] = fma(c1.x, u01, acc[2]); acc[3] = fma(c1.y, u11, acc[3]);
    }
    if (k == centre) { ctr[0] = c0.x; ctr[1] = c0.y; ctr[2] = c1.x; ctr[3] = c1.y; }
  }
  const double2 a = *reinterpret_cast<const double2 *>(pc), b = *reinterpret_cast<const double2 *>(pc + RW);
  const double u[4] = {a.x, a.y, b.x, b.y};
#pragma unroll
  for (int j = 0; j < 4; ++j) out[j] = fma(omega / ctr[j], f[j] - acc[j], u[j]);
}

template <int NST, int ORD>
__global__ void __launch_bounds__(RNT, 1)
k_jacvw(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_f,
        const __grid_constant__ JvArgs av, double *__restrict__ uout) {
  constexpr int S = JV_S;
  const RbwArgs &a = av.r;
  extern __shared__ __align__(1024) unsigned char smem[];
  uint64_t *full = reinterpret_cast<uint64_t *>(smem);
  // one spare row in front of the ring and behind the sweep-1 planes: reads of row -1 / row RH land there
  double *ring_u = reinterpret_cast<double *>(smem + ROFF_RING);
  double *ring_f = ring_u + RD * RPLANE;
  double *x1 = ring_f + RD * RPLANE;  // [plane mod 3][RPLANE]: results of sweep 1

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int x0 = 2 * lane, y0 = 2 * warp;
  if (tid == 0) {
    for (int i = 0; i < RD; ++i) mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const double omega = a.w;  // here: omega itself (the diagonal varies)
  const uint64_t pol_keep = jv_policy(true), pol_drop = jv_policy(false);
  const long long plane8 = a.plane * 8, pitch8 = (long long)a.pitch * 8;
  const int ld_off = y0 * RW + x0;

  const int G = gridDim.x, bid = blockIdx.x;
  const int nfull = a.tiles / G;
  const long long rem_units = (long long)(a.tiles - nfull * G) * a.nzp;
  const long long u1 = rem_units * (bid + 1) / G;
  long long ub = rem_units * bid / G;
  uint32_t kt = 0;

  for (int seg = 0;; ++seg) {
    int tile, za, zb;
    if (seg < nfull) {
      tile = seg * G + bid; za = 0; zb = a.nzp;
    } else {
      if (ub >= u1) break;
      const int rt = (int)(ub / a.nzp);
      tile = nfull * G + rt;
      za = (int)(ub - (long long)rt * a.nzp);
      const long long seg_end = min(u1, (long long)(rt + 1) * a.nzp);
      zb = za + (int)(seg_end - ub);
      ub = seg_end;
    }
    const int ox = (RW - 2 * S) * (tile % a.tiles_x), oy = (RH - 2 * S) * (tile / a.tiles_x);
    const int gx = ox - S + x0, gy = oy - S + y0;
    const int t0 = za - S, tl = zb - 1 + S;
    const int cx = ox - S + XPAD, cy = oy - S + GHOST;
    // Threads on the rim of the tile read neighbours beyond the staged plane (the spare rows around the
    // buffers): their results are garbage, but so far out in the halo that nothing stored depends on them.
    // Coefficients come by plain loads, so the block must lie inside the allocation of the field.
    const bool inner = gx >= -GHOST && gx + 1 < a.n[0] + GHOST && gy >= -GHOST && gy + 1 < a.n[1] + GHOST;
    uint32_t mk = 0;  // update masks, bit 4*(s-1)+j
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
    const bool xst = x0 >= S && x0 < RW - S && gx < a.n[0];
    const bool st0 = xst && y0 >= S && y0 < RH - S && gy < a.n[1];
    const bool st1 = xst && y0 >= S && y0 < RH - S && gy + 1 < a.n[1];
    char *const dst = reinterpret_cast<char *>(uout + gx) + pitch8 * gy;
    const double *const cf0 = av.sf + gx + (long long)a.pitch * gy;

    auto issue = [&](int z, uint32_t k) {
      const int sl = k & (RD - 1);
      mbar_expect_tx(&full[sl], 2 * RPLANE * 8);
      tma_load_plane(ring_u + sl * RPLANE, &tm_u, cx, cy, z + GHOST, &full[sl]);
      tma_load_plane(ring_f + sl * RPLANE, &tm_f, cx, cy, z + GHOST, &full[sl]);
    };
    if (tid == 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      for (int i = 0; i < RD - 2; ++i)
        if (t0 + i <= tl) issue(t0 + i, kt + i);
    }

    for (int t = t0; t <= tl; ++t) {
      const int slot_t = kt & (RD - 1);
      mbar_wait(&full[slot_t], (kt >> 2) & 1);
      // sweep 1 on plane t-1 from the staged planes t-2, t-1, t
      {
        const int p = t - 1;
        const double *pc = ring_u + ((slot_t + RD - 1) & (RD - 1)) * RPLANE + ld_off;
        double nv[4];
        {
          const double2 c = *reinterpret_cast<const double2 *>(pc), d = *reinterpret_cast<const double2 *>(pc + RW);
          nv[0] = c.x; nv[1] = c.y; nv[2] = d.x; nv[3] = d.y;
        }
        const bool zok = p >= a.ulo[2] - a.elo[2] && p < a.uhi[2] + a.ehi[2] && p > t0;
        if (zok && inner && (mk & 0xfu)) {
          const double *pm = ring_u + ((slot_t + RD - 2) & (RD - 1)) * RPLANE + ld_off;
          const double *pp = ring_u + slot_t * RPLANE + ld_off;
          const double *pf = ring_f + ((slot_t + RD - 1) & (RD - 1)) * RPLANE + ld_off;
          const double2 fa = *reinterpret_cast<const double2 *>(pf), fb = *reinterpret_cast<const double2 *>(pf + RW);
          const double f[4] = {fa.x, fa.y, fb.x, fb.y};
          double o[4];
          jv_sweep<NST, ORD>(pm, pc, pp, cf0 + a.plane * p, av.sfvol, a.pitch, pol_keep, omega, f, o);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if ((mk >> j) & 1u) nv[j] = o[j];
        }
        double *xw = x1 + ((p + 3 * 1024) % 3) * JV_XPL + ld_off;
        *reinterpret_cast<double2 *>(xw) = make_double2(nv[0], nv[1]);
        *reinterpret_cast<double2 *>(xw + RW) = make_double2(nv[2], nv[3]);
      }
      __syncthreads();
      // sweep 2 on plane t-2 from the sweep-1 planes t-3, t-2, t-1
      {
        const int p = t - 2;
        const double *pc = x1 + ((p + 3 * 1024) % 3) * JV_XPL + ld_off;
        double nv[4];
        {
          const double2 c = *reinterpret_cast<const double2 *>(pc), d = *reinterpret_cast<const double2 *>(pc + RW);
          nv[0] = c.x; nv[1] = c.y; nv[2] = d.x; nv[3] = d.y;
        }
        const bool zok = p >= a.ulo[2] && p < a.uhi[2];
        if (zok && inner && (mk & 0xf0u) && p >= za && p < zb) {
          const double *pm = x1 + ((p - 1 + 3 * 1024) % 3) * JV_XPL + ld_off;
          const double *pp = x1 + ((p + 1 + 3 * 1024) % 3) * JV_XPL + ld_off;
          const double *pf = ring_f + ((slot_t + RD - 2) & (RD - 1)) * RPLANE + ld_off;
          const double2 fa = *reinterpret_cast<const double2 *>(pf), fb = *reinterpret_cast<const double2 *>(pf + RW);
          const double f[4] = {fa.x, fa.y, fb.x, fb.y};
          double o[4];
          jv_sweep<NST, ORD>(pm, pc, pp, cf0 + a.plane * p, av.sfvol, a.pitch, pol_drop, omega, f, o);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if ((mk >> (4 + j)) & 1u) nv[j] = o[j];
        }
        if (p >= za && p < zb) {
          char *d = dst + plane8 * p;
          if (st0) *reinterpret_cast<double2 *>(d) = make_double2(nv[0], nv[1]);
          if (st1) *reinterpret_cast<double2 *>(d + pitch8) = make_double2(nv[2], nv[3]);
        }
      }
      __syncthreads();
      if (tid == 0 && t + RD - 2 <= tl) issue(t + RD - 2, kt + RD - 2);
      kt = (kt + 1) & (2 * RD - 1);
    }
  }
}

}  // namespace examg
