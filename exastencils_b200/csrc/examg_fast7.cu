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
//   * each CTA owns an x-y tile (60 x 30 outputs, 64 x 34 with the halo) and marches in z;
//     TMA (cp.async.bulk.tensor.3d) brings one 64x34 plane per step into a ring of 6 planes
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
#include <type_traits>

#include "examg_device.cuh"

namespace examg {

constexpr int FS = 2;                 // fused half-sweeps (colour 0, colour 1)
constexpr int FW = 64;                // tile width in shared memory (x)
constexpr int FH = 34;                // tile height in shared memory (y)
constexpr int FR = 2 * FS + 2;        // ring depth (planes)
constexpr int FOW = FW - 2 * FS;      // 60 output columns
constexpr int FOH = FH - 2 * FS;      // 30 output rows
constexpr int FNT = 256;              // threads per CTA
constexpr int FPLANE = FW * FH;       // doubles per plane
constexpr int FITEMS = ((FH - 2) / 2) * (FW / 32);  // warp items per stage: row pairs x 32-wide chunks
constexpr int FNI = (FITEMS + FNT / 32 - 1) / (FNT / 32);  // items per warp and stage (4)
constexpr size_t FSMEM = sizeof(double) * FPLANE * FR + 8 * FR;

struct FusedArgs {
  int n[3];
  int ulo[3], uhi[3];  // update bounds (local node coordinates), same for both colours
  int pitch;
  long long plane;
  int tiles_x, tiles_y;
  int nzp;                     // output planes 0..nzp-1
  long long units;             // tiles * nzp
  int par;
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

// Declaration orders of the 7-point stencils in the reference's programs (the order fixes the
// floating-point association): 0 = centre,-x,+x,-y,+y,-z,+z (Benchmark/Poisson3D, Examples/Poisson),
// 1 = centre,+x,-x,+y,-y,+z,-z (Testing/Smoothers, Testing/SISC, Testing/Opts).
__device__ constexpr int ORD_DX[2][7] = {{0, -1, 1, 0, 0, 0, 0}, {0, 1, -1, 0, 0, 0, 0}};
__device__ constexpr int ORD_DY[2][7] = {{0, 0, 0, -1, 1, 0, 0}, {0, 0, 0, 1, -1, 0, 0}};
__device__ constexpr int ORD_DZ[2][7] = {{0, 0, 0, 0, 0, -1, 1}, {0, 0, 0, 0, 0, 1, -1}};

template <int ORD>
__global__ void __launch_bounds__(FNT, 2)
k_rb_fused(const __grid_constant__ CUtensorMap tm_in, const __grid_constant__ FusedArgs a,
           const double *__restrict__ rhs, double *__restrict__ uout) {
  // dynamic shared memory only (starts 1024-byte aligned): ring of FR planes, then the mbarriers
  extern __shared__ __align__(1024) double ring[];
  uint64_t *full = reinterpret_cast<uint64_t *>(ring + FPLANE * FR);
  char *ringb = reinterpret_cast<char *>(ring);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    for (int i = 0; i < FR; ++i) mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  double coef[7];
#pragma unroll
  for (int k = 0; k < 7; ++k) coef[k] = a.coef[k];
  const double w = a.w;
  const long long plane8 = a.plane * 8;
  const long long pitch8 = (long long)a.pitch * 8;

  // this CTA's share of the (tile, plane) units
  const long long u0 = a.units * blockIdx.x / gridDim.x;
  const long long u1 = a.units * (blockIdx.x + 1) / gridDim.x;
  int loads = 0;  // planes loaded so far by this CTA, modulo 2*FR (ring position and mbarrier phase)

  for (long long ub = u0; ub < u1;) {
    const int tile = (int)(ub / a.nzp);
    const int za = (int)(ub - (long long)tile * a.nzp);
    const long long seg_end = min(u1, (long long)(tile + 1) * a.nzp);
    const int zb = za + (int)(seg_end - ub);
    ub = seg_end;
    const int ox = FOW * (tile % a.tiles_x), oy = FOH * (tile / a.tiles_x);  // first output node of the tile
    const int t0 = za - FS, tl = zb - 1 + FS;  // first / last plane loaded
    // array coordinates of the tile corner (TMA zero-fills outside the allocation)
    const int cx = ox - FS + XPAD, cy = oy - FS + GHOST;

    // ---- per-thread item geometry, fixed for the segment ------------------------------
    // A warp item is 2 rows x 32 columns; the lane owns the column pair (xb, xb+1) of one row.
    // 64-bit shared-memory accesses are served per half-warp: lanes 0-7 take row y, lanes 8-15
    // row y+1 of one 16-column group; their colour points sit in complementary 8-byte bank
    // pairs, so every access is conflict-free.
    // Colour of (x,y,z) is (x+y+z+par)&1.  At step t = t0+rel stage 2 (colour 1, plane t-3)
    // updates x = xb + (e ^ rel&1) and stage 1 (colour 0, plane t-1) the other column, with
    // e = (xb+y+par+t0)&1.  "A" = column xb+e, "B" = column xb+(e^1).
    uint32_t offA[FNI], offB[FNI];    // byte offset of the point inside a plane of the ring
    uint32_t roffA[FNI], roffB[FNI];  // byte offset of the point inside a plane of rhs
    uint32_t okmask = 0;              // bit 2*it: A may be updated, bit 2*it+1: B
#pragma unroll
    for (int it = 0; it < FNI; ++it) {
      const int item = warp + it * (FNT / 32);
      const int rp = item / (FW / 32), ch = item % (FW / 32);
      const int y = 1 + 2 * rp + ((lane >> 3) & 1);
      const int xb = 32 * ch + 16 * (lane >> 4) + 2 * (lane & 7);
      const int lx = ox - FS + xb, ly = oy - FS + y;
      const int e = (lx + ly + a.par + t0) & 1;
      const bool rowok = ly >= a.ulo[1] && ly < a.uhi[1];
      const bool v0 = rowok && xb >= 1 && lx >= a.ulo[0] && lx < a.uhi[0];
      const bool v1 = rowok && xb + 1 <= FW - 2 && lx + 1 >= a.ulo[0] && lx + 1 < a.uhi[0];
      offA[it] = (uint32_t)(y * FW + xb + e) * 8u;
      offB[it] = (uint32_t)(y * FW + xb + (e ^ 1)) * 8u;
      const int g = lx + a.pitch * ly;
      roffA[it] = (uint32_t)(g + e) * 8u;
      roffB[it] = (uint32_t)(g + (e ^ 1)) * 8u;
      okmask |= ((e ? v1 : v0) ? 1u : 0u) << (2 * it);
      okmask |= ((e ? v0 : v1) ? 1u : 0u) << (2 * it + 1);
    }
    // output rows of this warp: r = warp, warp+8, ...; two 32-column chunks
    const uint32_t so_off = (uint32_t)((FS + warp) * FW + FS + lane) * 8u;
    const int wlim = min(FOW, a.n[0] - ox), hlim = min(FOH, a.n[1] - oy);
    char *const dst0 = reinterpret_cast<char *>(uout + ox + lane) + pitch8 * (oy + warp);

    if (tid == 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      for (int z = t0; z <= min(t0 + 1, tl); ++z) {
        const int k = (loads + z - t0) % (2 * FR), sl = k % FR;
        mbar_expect_tx(&full[sl], FPLANE * 8);
        tma_load_plane(ring + sl * FPLANE, &tm_in, cx, cy, z + GHOST, &full[sl]);
      }
    }
    // rhs values are fetched one step ahead into registers: step parity 0 consumes rq0 and fills
    // rq1, parity 1 the other way round, so no copies are needed
    double rq0[FS][FNI], rq1[FS][FNI];
#pragma unroll
    for (int s = 0; s < FS; ++s)
#pragma unroll
      for (int it = 0; it < FNI; ++it) rq0[s][it] = rq1[s][it] = 0.0;

    int kt = loads;  // ring counter of plane t (mod 2*FR)

    auto step = [&](auto parity, const int t) {
      constexpr int P = decltype(parity)::value;
      double(&rcur)[FS][FNI] = P ? rq1 : rq0;
      double(&rnxt)[FS][FNI] = P ? rq0 : rq1;
      // ---- (1) rhs prefetch for step t+1 (relative parity P^1) --------------------------
#pragma unroll
      for (int s = 1; s <= FS; ++s) {
        const int zs = t + 1 - (2 * s - 1);
        const bool zok = zs >= za - (FS - s) && zs < zb + (FS - s) && zs >= a.ulo[2] && zs < a.uhi[2];
        if (zok) {
          const bool useA = (s == 2) == ((P ^ 1) == 0);
          const char *rz = reinterpret_cast<const char *>(rhs) + plane8 * zs;  // uniform
#pragma unroll
          for (int it = 0; it < FNI; ++it) {
            const bool ok = (okmask >> (2 * it + (useA ? 0 : 1))) & 1u;
            const uint32_t ro = useA ? roffA[it] : roffB[it];
            rnxt[s - 1][it] = ok ? __ldg(reinterpret_cast<const double *>(rz + ro)) : 0.0;
          }
        }
      }
      const int slot_t = kt >= FR ? kt - FR : kt;
      if (t <= tl) mbar_wait(&full[slot_t], (uint32_t)(kt >= FR));
      // ---- (2) stages: colour s-1 on plane t-(2s-1); all loads first, then all stores ----
      double nv[FS][FNI];
      uint32_t sa[FS][FNI];
      bool act[FS];
#pragma unroll
      for (int s = 1; s <= FS; ++s) {
        const int zs = t - (2 * s - 1);
        act[s - 1] = zs >= za - (FS - s) && zs < zb + (FS - s) && zs >= a.ulo[2] && zs < a.uhi[2];
        if (act[s - 1]) {
          const bool useA = (s == 2) == (P == 0);
          int sc = slot_t - (2 * s - 1);  // ring slots of planes zs, zs-1, zs+1
          if (sc < 0) sc += FR;
          const int sm = sc == 0 ? FR - 1 : sc - 1, sp = sc == FR - 1 ? 0 : sc + 1;
          const char *pcb = ringb + sc * (FPLANE * 8);
          const int dm = (sm - sc) * (FPLANE * 8), dp = (sp - sc) * (FPLANE * 8);
#pragma unroll
          for (int it = 0; it < FNI; ++it) {
            const uint32_t off = useA ? offA[it] : offB[it];
            const char *pt = pcb + off;  // all seven neighbours are inside the ring for every lane
            double v[7];
#pragma unroll
            for (int k = 0; k < 7; ++k) {
              const int d = (ORD_DX[ORD][k] + FW * ORD_DY[ORD][k]) * 8;
              const char *q = ORD_DZ[ORD][k] == 0 ? pt : (ORD_DZ[ORD][k] < 0 ? pt + dm : pt + dp);
              v[k] = *reinterpret_cast<const double *>(q + d);
            }
            double acc = coef[0] * v[0];
#pragma unroll
            for (int k = 1; k < 7; ++k) acc = fma(coef[k], v[k], acc);
            const double r = rcur[s - 1][it] - acc;
            nv[s - 1][it] = fma(w, r, v[0]);
            sa[s - 1][it] = (uint32_t)(sc * (FPLANE * 8)) + off;
          }
        }
      }
#pragma unroll
      for (int s = 0; s < FS; ++s)
        if (act[s]) {
          const bool useA = (s == 1) == (P == 0);
#pragma unroll
          for (int it = 0; it < FNI; ++it)
            if ((okmask >> (2 * it + (useA ? 0 : 1))) & 1u) *reinterpret_cast<double *>(ringb + sa[s][it]) = nv[s][it];
        }
      // ---- (3) plane t-2S is final: write the tile's outputs -----------------------------
      const int zf = t - 2 * FS;
      if (zf >= za && zf < zb) {
        int sf = slot_t - 2 * FS;
        if (sf < 0) sf += FR;
        const char *src = ringb + sf * (FPLANE * 8) + so_off;
        char *dst = dst0 + plane8 * zf;
#pragma unroll
        for (int i = 0; i < (FOH + FNT / 32 - 1) / (FNT / 32); ++i) {
          const int r = warp + i * (FNT / 32);
          if (r < hlim) {
            const double *sr = reinterpret_cast<const double *>(src + i * (FNT / 32) * FW * 8);
            double *dr = reinterpret_cast<double *>(dst + pitch8 * (i * (FNT / 32)));
            if (lane < wlim) dr[0] = sr[0];
            if (lane + 32 < wlim) dr[32] = sr[32];
          }
        }
      }
      __syncthreads();
      // the slot of plane t-2S is free now: prefetch plane t+2 into it
      if (tid == 0 && t + 2 <= tl) {
        const int sl = slot_t + 2 >= FR ? slot_t + 2 - FR : slot_t + 2;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(&full[sl], FPLANE * 8);
        tma_load_plane(ring + sl * FPLANE, &tm_in, cx, cy, t + 2 + GHOST, &full[sl]);
      }
      kt = kt + 1 == 2 * FR ? 0 : kt + 1;
    };

    const int tend = zb - 1 + 2 * FS;
    for (int t = t0; t <= tend; t += 2) {
      step(std::integral_constant<int, 0>{}, t);
      if (t + 1 <= tend) step(std::integral_constant<int, 1>{}, t + 1);
    }
    loads = (loads + tl - t0 + 1) % (2 * FR);
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

// 0 / 1: one of the two known declaration orders; -1: something else (generic kernels)
static int stencil_order(const Stencil &s) {
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
  if (v < 0) {
    const char *e = getenv("EXAMG_NO_FUSED");
    v = (e && e[0] == '1') ? 0 : 1;
  }
  return v == 1;
}

// smallest level extent for which the tiled kernel pays off (below: a few hundred KB per sweep)
constexpr int FUSED_MIN_N = 65;

bool fused_rb_usable(examg_ctx *c, const Stencil &s, const Field &fu) {
  if (!fused_enabled() || !s.is7 || c->cfg.dim != 3 || fu.ncomp != 1 || stencil_order(s) < 0) return false;
  if (c->world > 1) return false;  // multi-block: colour sweeps + exchange (examg_comm.cu)
  const LevelGeom &g = *geom(c, s.level);
  return g.n[0] >= FUSED_MIN_N && g.n[1] >= FUSED_MIN_N && g.n[2] >= FUSED_MIN_N;
}

// Allocates the second buffer of a field (not capturable: call before graph capture).
int ensure_shadow(examg_ctx *c, Field &fu) {
  static bool attr_set = false;
  if (!attr_set) {
    EXAMG_CUDA(cudaFuncSetAttribute(k_rb_fused<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FSMEM));
    EXAMG_CUDA(cudaFuncSetAttribute(k_rb_fused<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FSMEM));
    attr_set = true;
  }
  if (fu.shadow) return EXAMG_OK;
  const LevelGeom &g = *geom(c, fu.level);
  const size_t bytes = sizeof(double) * (size_t)g.vol;
  cudaError_t e = cudaMalloc(&fu.shadow, bytes);
  if (e != cudaSuccess) return set_error(EXAMG_ENOMEM, "shadow buffer of %s@%d: %s", fu.name.c_str(), fu.level, cudaGetErrorString(e));
  EXAMG_CUDA(cudaMemset(fu.shadow, 0, bytes));
  c->dev_bytes += (long long)bytes;
  return EXAMG_OK;
}

// One full sweep: colour 0 then colour 1.  Swaps fu.slot[cur] with the shadow buffer.
int launch_rb_fused(examg_ctx *c, const Stencil &s, double omega, Field &fu, const double *rhs) {
  const LevelGeom &g = *geom(c, s.level);
  int rc0 = ensure_shadow(c, fu);
  if (rc0) return rc0;
  CUtensorMap tm;
  int rc = make_plane_map(g, fu.slot[fu.cur], &tm);
  if (rc) return rc;
  FusedArgs a;
  for (int d = 0; d < 3; ++d) {
    a.n[d] = g.n[d];
    a.ulo[d] = g.lo[d];
    a.uhi[d] = g.hi[d];
  }
  a.pitch = g.pitch;
  a.plane = g.plane;
  a.tiles_x = (g.n[0] + FOW - 1) / FOW;
  a.tiles_y = (g.n[1] + FOH - 1) / FOH;
  a.nzp = g.n[2];
  a.units = (long long)a.tiles_x * a.tiles_y * a.nzp;
  a.par = g.par;
  for (int k = 0; k < 7; ++k) a.coef[k] = s.coef[k];
  const int ord = stencil_order(s);
  a.w = omega / s.coef[s.centre];
  int dev = 0, nsm = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
  long long grid = 2LL * nsm;
  const long long min_units = 24;  // at least this many planes per CTA (4 are restart overhead)
  if (grid > a.units / min_units) grid = a.units / min_units;
  if (grid < 1) grid = 1;
  if (ord == 0)
    k_rb_fused<0><<<(int)grid, FNT, FSMEM, c->stream>>>(tm, a, rhs, fu.shadow + g.origin);
  else
    k_rb_fused<1><<<(int)grid, FNT, FSMEM, c->stream>>>(tm, a, rhs, fu.shadow + g.origin);
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
