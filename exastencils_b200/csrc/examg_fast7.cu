// examg_fast7.cu -- 3D 7-point constant-coefficient fast path: one full red-black
// Gauss-Seidel sweep (colour 0 then colour 1) per pass over HBM.
//
// Reference loop being replaced (two passes per sweep, each a masked full-grid loop):
//   color with { (i0+i1+i2)%2, communicate; loop over Solution { Solution += omega/diag(A) *
//   (RHS - A*Solution) }; apply bc }         Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4:204-213
//   colour order / duplication               C/s/e/baseExt/l4/L4_ColorLoops.scala:44-70
//   temporal blocking precedent              C/s/e/baseExt/ir/IR_ContractingLoop.scala:48-90,131-195
//
// Design (B200):
//   * out of place: reads Solution, writes the shadow buffer, buffers are swapped afterwards
//     (so a tile may read its 2-deep halo from neighbours that are being rewritten)
//   * each CTA owns an x-y tile (60 x 32 outputs, 64 x 36 with the halo) and marches in z;
//     TMA (cp.async.bulk.tensor.3d) brings one 64x36 plane per step into a ring of 6 planes
//     in shared memory, completion tracked by mbarriers, two planes prefetched ahead
//   * 2.5-D wavefront: at step t colour 0 is applied in place (in shared memory) to plane
//     t-1 and colour 1 to plane t-3; the skew of two planes makes the stages of one step
//     independent, so there is one __syncthreads per plane; plane t-4 is final and is
//     written back with coalesced stores
//   * lanes 0-15 / 16-31 of a warp take the colour points of two adjacent rows: the 64-bit
//     shared-memory accesses of a warp then cover complementary 8-byte slots -> no bank conflict
//   * work = (tile, plane) pairs split evenly over a persistent grid of 2 CTAs per SM
//     (stream-K style), so there is no tail wave; restart cost is 4 planes per segment
// HBM traffic per point and sweep: 8*1.2 (Solution incl. halo) + 8*1.1 (RHS) + 8 (write)
// = 26.4 B against 24 B algorithmic (32 B for two element-granular colour passes, and
// 48 B measured for them because a 32-byte sector holds both colours).
//
// Arithmetic is the contract of examg_kernels.cu (bit-identical to the unfused order).
#include <cuda.h>

#include <map>

#include "examg_device.cuh"

namespace examg {

constexpr int FS = 2;                 // fused half-sweeps (colour 0, colour 1)
constexpr int FW = 64;                // tile width in shared memory (x)
constexpr int FH = 36;                // tile height in shared memory (y)
constexpr int FR = 2 * FS + 2;        // ring depth (planes)
constexpr int FOW = FW - 2 * FS;      // 60 output columns
constexpr int FOH = FH - 2 * FS;      // 32 output rows
constexpr int FNT = 256;              // threads per CTA
constexpr int FPLANE = FW * FH;       // doubles per plane
constexpr int FITEMS = ((FH - 2) / 2) * (FW / 32);  // warp items per stage: row pairs x 32-wide chunks
constexpr size_t FSMEM = sizeof(double) * FPLANE * FR + 128;

struct FusedArgs {
  int n[3];
  int ulo[FS][3], uhi[FS][3];  // update bounds per stage (local node coordinates)
  int pitch;
  long long plane;
  int tiles_x, tiles_y;
  int nzp;                     // output planes 0..nzp-1
  long long units;             // tiles * nzp
  int par;
  int soff[7];                 // in-plane shared-memory offset of stencil entry k
  int sdz[7];                  // dz of entry k
  int centre;
  double coef[7];
  double w;                    // omega / diag
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

__global__ void __launch_bounds__(FNT, 2)
k_rb_fused(const __grid_constant__ CUtensorMap tm_in, const __grid_constant__ FusedArgs a,
           const double *__restrict__ rhs, double *__restrict__ uout) {
  extern __shared__ unsigned char smem_raw[];
  double *ring = reinterpret_cast<double *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  __shared__ uint64_t full[FR];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    for (int i = 0; i < FR; ++i) mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // this CTA's share of the (tile, plane) units
  const long long u0 = a.units * blockIdx.x / gridDim.x;
  const long long u1 = a.units * (blockIdx.x + 1) / gridDim.x;
  long long loads = 0;  // planes loaded so far by this CTA (ring position / mbarrier phase)

  for (long long ub = u0; ub < u1;) {
    const int tile = (int)(ub / a.nzp);
    const int za = (int)(ub - (long long)tile * a.nzp);
    const long long seg_end = min(u1, (long long)(tile + 1) * a.nzp);
    const int zb = za + (int)(seg_end - ub);
    ub = seg_end;
    const int ox = FOW * (tile % a.tiles_x), oy = FOH * (tile / a.tiles_x);  // first output node of the tile
    const int t0 = za - FS, tl = zb - 1 + FS;  // first / last plane loaded
    const long long kbase = loads - t0;       // ring position of plane z is kbase + z
    // array coordinates of the tile corner (TMA zero-fills outside the allocation)
    const int cx = ox - FS + XPAD, cy = oy - FS + GHOST;

    if (tid == 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      for (int z = t0; z <= min(t0 + 1, tl); ++z) {
        const int sl = (int)((kbase + z) % FR);
        mbar_expect_tx(&full[sl], FPLANE * 8);
        tma_load_plane(ring + sl * FPLANE, &tm_in, cx, cy, z + GHOST, &full[sl]);
      }
    }
    for (int t = t0; t <= zb - 1 + 2 * FS; ++t) {
      if (t <= tl) {
        const long long k = kbase + t;
        mbar_wait(&full[k % FR], (uint32_t)((k / FR) & 1));
      }
      // ---- stages: colour s-1 on plane t-(2s-1) -------------------------------------
#pragma unroll
      for (int s = 1; s <= FS; ++s) {
        const int zs = t - (2 * s - 1);
        if (zs < za - (FS - s) || zs >= zb + (FS - s)) continue;
        if (zs < a.ulo[s - 1][2] || zs >= a.uhi[s - 1][2]) continue;
        const int color = (s - 1) & 1;
        const int pc = (int)((kbase + zs) % FR) * FPLANE;
        const int dm = (int)((kbase + zs - 1) % FR) * FPLANE - pc;
        const int dp = (int)((kbase + zs + 1) % FR) * FPLANE - pc;
        int nb[7];
#pragma unroll
        for (int k = 0; k < 7; ++k) nb[k] = a.soff[k] + (a.sdz[k] > 0 ? dp : (a.sdz[k] < 0 ? dm : 0));
        const double *rz = rhs + a.plane * zs;
        for (int item = warp; item < FITEMS; item += FNT / 32) {
          const int rp = item / (FW / 32), ch = item % (FW / 32);
          const int y = 1 + 2 * rp + (lane >> 4);
          const int xb = 32 * ch + 2 * (lane & 15);
          const int ly = oy - FS + y;
          const int q = (ox - FS + xb + ly + zs + a.par + color) & 1;
          const int x = xb + q, lx = ox - FS + x;
          const bool ok = x >= 1 && x <= FW - 2 && lx >= a.ulo[s - 1][0] && lx < a.uhi[s - 1][0] &&
                          ly >= a.ulo[s - 1][1] && ly < a.uhi[s - 1][1];
          if (ok) {
            double *pt = ring + pc + y * FW + x;
            const double f = __ldg(rz + lx + (long long)a.pitch * ly);
            double acc = a.coef[0] * pt[nb[0]];
#pragma unroll
            for (int k = 1; k < 7; ++k) acc = fma(a.coef[k], pt[nb[k]], acc);
            const double r = f - acc;
            *pt = fma(a.w, r, *pt);
          }
        }
      }
      // ---- plane t-2S is final: write the tile's 60x32 outputs -------------------------
      const int zf = t - 2 * FS;
      if (zf >= za && zf < zb) {
        const double *src = ring + (int)((kbase + zf) % FR) * FPLANE;
        double *dst = uout + a.plane * zf;
        for (int e = tid; e < FOW * FOH; e += FNT) {
          const int r = e / FOW, col = e - r * FOW;
          const int lx = ox + col, ly = oy + r;
          if (lx < a.n[0] && ly < a.n[1]) dst[lx + (long long)a.pitch * ly] = src[(r + FS) * FW + col + FS];
        }
      }
      __syncthreads();
      // the slot of plane t-2S is free now: prefetch plane t+2 into it
      if (tid == 0 && t + 2 <= tl) {
        const long long k = kbase + t + 2;
        const int sl = (int)(k % FR);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(&full[sl], FPLANE * 8);
        tma_load_plane(ring + sl * FPLANE, &tm_in, cx, cy, t + 2 + GHOST, &full[sl]);
      }
    }
    loads += tl - t0 + 1;
  }
}

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

static int make_plane_map(const LevelGeom &g, double *base, CUtensorMap *tm) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return set_error(EXAMG_ECUDA, "cuTensorMapEncodeTiled entry point not available");
  const cuuint64_t dims[3] = {(cuuint64_t)g.pitch, (cuuint64_t)g.ay, (cuuint64_t)g.az};
  const cuuint64_t strides[2] = {(cuuint64_t)g.pitch * 8, (cuuint64_t)g.plane * 8};
  const cuuint32_t box[3] = {FW, FH, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(EXAMG_ECUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
  return EXAMG_OK;
}

static bool fused_enabled() {
  static int v = -1;
  if (v < 0) {
    const char *e = getenv("EXAMG_NO_FUSED");
    v = (e && e[0] == '1') ? 0 : 1;
  }
  return v == 1;
}

// smallest level extent for which the tiled kernel pays off (below: a few hundred KB per sweep)
constexpr int FUSED_MIN_N = 65;

bool fused_rb_usable(examg_ctx *c, const Stencil &s, const Field &fu) {
  if (!fused_enabled() || !s.is7 || c->cfg.dim != 3 || fu.ncomp != 1) return false;
  if (c->world > 1) return false;  // multi-block: colour sweeps + exchange (examg_comm.cu)
  const LevelGeom &g = *geom(c, s.level);
  return g.n[0] >= FUSED_MIN_N && g.n[1] >= FUSED_MIN_N && g.n[2] >= FUSED_MIN_N;
}

// One full sweep: colour 0 then colour 1.  Swaps fu.slot[cur] with the shadow buffer.
int launch_rb_fused(examg_ctx *c, const Stencil &s, double omega, Field &fu, const double *rhs) {
  const LevelGeom &g = *geom(c, s.level);
  if (!fu.shadow) {
    const size_t bytes = sizeof(double) * (size_t)g.vol;
    cudaError_t e = cudaMalloc(&fu.shadow, bytes);
    if (e != cudaSuccess) return set_error(EXAMG_ENOMEM, "shadow buffer of %s@%d: %s", fu.name.c_str(), fu.level, cudaGetErrorString(e));
    EXAMG_CUDA(cudaMemsetAsync(fu.shadow, 0, bytes, c->stream));
    c->dev_bytes += (long long)bytes;
  }
  static bool attr_set = false;
  if (!attr_set) {
    EXAMG_CUDA(cudaFuncSetAttribute(k_rb_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FSMEM));
    attr_set = true;
  }
  CUtensorMap tm;
  int rc = make_plane_map(g, fu.slot[fu.cur], &tm);
  if (rc) return rc;
  FusedArgs a;
  for (int d = 0; d < 3; ++d) {
    a.n[d] = g.n[d];
    for (int st = 0; st < FS; ++st) { a.ulo[st][d] = g.lo[d]; a.uhi[st][d] = g.hi[d]; }
  }
  a.pitch = g.pitch;
  a.plane = g.plane;
  a.tiles_x = (g.n[0] + FOW - 1) / FOW;
  a.tiles_y = (g.n[1] + FOH - 1) / FOH;
  a.nzp = g.n[2];
  a.units = (long long)a.tiles_x * a.tiles_y * a.nzp;
  a.par = g.par;
  a.centre = s.centre;
  for (int k = 0; k < 7; ++k) {
    a.soff[k] = s.off[k][0] + FW * s.off[k][1];
    a.sdz[k] = s.off[k][2];
    a.coef[k] = s.coef[k];
  }
  a.w = omega / s.coef[s.centre];
  int dev = 0, nsm = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
  long long grid = 2LL * nsm;
  const long long min_units = 24;  // at least this many planes per CTA (4 are restart overhead)
  if (grid > a.units / min_units) grid = a.units / min_units;
  if (grid < 1) grid = 1;
  k_rb_fused<<<(int)grid, FNT, FSMEM, c->stream>>>(tm, a, rhs, fu.shadow + g.origin);
  EXAMG_LAUNCH_CHECK(c);
  double *t = fu.slot[fu.cur];
  fu.slot[fu.cur] = fu.shadow;
  fu.shadow = t;
  fu.swapped = !fu.swapped;
  return EXAMG_OK;
}

bool fast7_usable(examg_ctx *c, const Stencil &s) { (void)c; (void)s; return false; }
int fast7_residual(examg_ctx *c, const Stencil &s, const double *u, const double *rhs, double *res, int red_slot) {
  (void)c; (void)s; (void)u; (void)rhs; (void)res; (void)red_slot;
  return set_error(EXAMG_EUNSUP, "fast7_residual");
}
int fast7_rb(examg_ctx *c, const Stencil &s, double omega, double *u, const double *rhs, int color) {
  (void)c; (void)s; (void)omega; (void)u; (void)rhs; (void)color;
  return set_error(EXAMG_EUNSUP, "fast7_rb");
}

}  // namespace examg
