// examg_device.cuh -- device helpers shared by the kernel translation units.
#pragma once

#include "examg_internal.h"

namespace examg {

// ---- double-double accumulation (order-insensitive sums) ------
// Knuth TwoSum; compiled with --fmad=false so nothing here is contracted.
__device__ __forceinline__ void dd_add(dd &a, double x) {
  double s = a.hi + x;
  double bb = s - a.hi;
  double err = (a.hi - (s - bb)) + (x - bb);
  a.hi = s;
  a.lo += err;
}
__device__ __forceinline__ void dd_merge(dd &a, const dd &b) {
  dd_add(a, b.hi);
  a.lo += b.lo;
}

__device__ __forceinline__ dd dd_shfl_down(const dd &v, int delta) {
  dd r;
  r.hi = __shfl_down_sync(0xffffffffu, v.hi, delta);
  r.lo = __shfl_down_sync(0xffffffffu, v.lo, delta);
  return r;
}

// Block-wide merge of per-thread partials in a fixed order (warp shuffles, then
// one value per warp through shared memory).  Result valid in thread 0.
__device__ __forceinline__ dd dd_block_reduce(dd v) {
  __shared__ dd warp_part[32];
  const int tid = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
  const int nthreads = blockDim.x * blockDim.y * blockDim.z;
  const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    dd o = dd_shfl_down(v, d);
    dd_merge(v, o);
  }
  if (lane == 0) warp_part[warp] = v;
  __syncthreads();
  const int nwarps = (nthreads + 31) >> 5;
  if (warp == 0) {
    v = lane < nwarps ? warp_part[lane] : dd{0.0, 0.0};
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      dd o = dd_shfl_down(v, d);
      dd_merge(v, o);
    }
  }
  __syncthreads();
  return v;
}

__device__ __forceinline__ double max_block_reduce(double v) {
  __shared__ double warp_part[32];
  const int tid = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
  const int nthreads = blockDim.x * blockDim.y * blockDim.z;
  const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v = fmax(v, __shfl_down_sync(0xffffffffu, v, d));
  if (lane == 0) warp_part[warp] = v;
  __syncthreads();
  const int nwarps = (nthreads + 31) >> 5;
  if (warp == 0) {
    v = lane < nwarps ? warp_part[lane] : 0.0;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = fmax(v, __shfl_down_sync(0xffffffffu, v, d));
  }
  __syncthreads();
  return v;
}

// How a loop nest is spread over the grid: thread (tx,ty) of block (bx,by,bz)
// owns x = x0 + bx*BX + tx (times xstep), y = y0 + by*BY + ty (+ m*gridDim.y*BY,
// grid-stride) and marches over kz consecutive z planes starting at z0 + bz*kz.
struct Shape {
  dim3 grid, block;
  int kz;
  int nblocks;
};

// ex/ey/ez: extents of the loop nest (x already divided by the colour stride)
inline Shape make_shape(int ex, int ey, int ez, int max_blocks = 16384) {
  Shape s;
  int bx = 32;
  while (bx < 128 && bx < ex) bx <<= 1;
  int by = 256 / bx;
  while (by > 1 && by / 2 >= ey) by >>= 1;
  s.block = dim3(bx, by, 1);
  int gx = (ex + bx - 1) / bx, gy = (ey + by - 1) / by;
  if (gx < 1) gx = 1;
  if (gy < 1) gy = 1;
  int kz = 1;
  while ((long long)gx * gy * ((ez + kz - 1) / kz) > max_blocks && kz < ez) kz <<= 1;
  // still too many (large 2D grids): rows are then covered by a grid-stride loop in y
  if ((long long)gx * gy * ((ez + kz - 1) / kz) > max_blocks) gy = max_blocks / (gx * ((ez + kz - 1) / kz));
  if (gy < 1) gy = 1;
  int gz = (ez + kz - 1) / kz;
  if (gz < 1) gz = 1;
  s.grid = dim3(gx, gy, gz);
  s.kz = kz;
  s.nblocks = gx * gy * gz;
  return s;
}

}  // namespace examg
