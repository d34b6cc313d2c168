// examg_noalloc.cu -- examg_dot_dense_noalloc: the dot-product reduction of examg_dot (k_reduce<R_DOT> +
// k_final_sum, examg_kernels.cu) for callers that may not allocate: code under CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// The per-block partial sums live in a file-scope device array instead of the context's cudaMalloc'd scratch
// (examg_ctx::partials).  Every GPU that loads the library holds the array (NOALLOC_SCRATCH_BYTES) until exit.  All
// calls share it and it keeps the last call's contents: the ticket counter is cleared at the start of every call, the
// partial slots a call reads are exactly the ones it wrote.  Two calls that overlap on the GPU (different streams
// without ordering, or two host threads) race on the array -- callers serialise.
//
// The entry point calls no cudaMalloc / cudaFree / allocating library routine: cudaGetDevice, one
// cudaGetSymbolAddress per device (first call), cudaMemsetAsync on 4 bytes of the array, one kernel launch.
#include "examg_device.cuh"

namespace examg {

constexpr long long NOALLOC_CHUNK = 1 << 16;       // elements per thread block
constexpr int NOALLOC_MAX_BLOCKS = 4096;           // => EXAMG_NOALLOC_MAX_N = 2^28 elements
constexpr int NOALLOC_HEADER = 256;                // ticket counter, padded to the array's alignment
constexpr int NOALLOC_SCRATCH_BYTES = NOALLOC_HEADER + NOALLOC_MAX_BLOCKS * (int)sizeof(dd);  // 65 792 bytes
constexpr int NOALLOC_MAX_DEVICES = 64;

static_assert(NOALLOC_CHUNK * NOALLOC_MAX_BLOCKS == EXAMG_NOALLOC_MAX_N, "header and kernel disagree on the maximum");

// In host code this name is a host-side stand-in, not a device address: the kernel gets the address
// cudaGetSymbolAddress returns for the current device (noalloc_scratch below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

static void *g_scratch_addr[NOALLOC_MAX_DEVICES];  // per device, filled on first use with that device current

// Block b sums a[i]*b[i] over its chunk in double-double (the accumulation of k_reduce<R_DOT>), publishes the partial
// and takes a ticket; the block that draws the last ticket merges all partials in index order (k_final_sum's order)
// and leaves the counter at zero.
__global__ void __launch_bounds__(256)
k_dot_dense_noalloc(const double *__restrict__ a, const double *__restrict__ b, long long n, unsigned char *scratch,
                    double *__restrict__ out) {
  unsigned int *ticket = reinterpret_cast<unsigned int *>(scratch);
  dd *partials = reinterpret_cast<dd *>(scratch + NOALLOC_HEADER);
  const long long i0 = (long long)blockIdx.x * NOALLOC_CHUNK;
  const long long i1 = min(i0 + NOALLOC_CHUNK, n);
  dd part = {0.0, 0.0};
  for (long long i = i0 + threadIdx.x; i < i1; i += blockDim.x) dd_add(part, a[i] * b[i]);
  part = dd_block_reduce(part);
  __shared__ bool last;
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = part;
    __threadfence();
    last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  dd v = {0.0, 0.0};
  for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) {
    const volatile dd *p = partials + i;
    dd q = {p->hi, p->lo};
    dd_merge(v, q);
  }
  v = dd_block_reduce(v);
  if (threadIdx.x == 0) {
    *out = v.hi + v.lo;
    *ticket = 0u;
  }
}

// device address of the scratch array on the current device
static int noalloc_scratch(unsigned char **ptr) {
  int dev = -1;
  EXAMG_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES)
    return set_error(EXAMG_EUNSUP, "examg_dot_dense_noalloc: device ordinal %d beyond %d", dev, NOALLOC_MAX_DEVICES - 1);
  if (!g_scratch_addr[dev]) {
    void *p = nullptr;
    EXAMG_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (!p) return set_error(EXAMG_ECUDA, "examg_dot_dense_noalloc: cudaGetSymbolAddress returned a null address");
    g_scratch_addr[dev] = p;
  }
  *ptr = static_cast<unsigned char *>(g_scratch_addr[dev]);
  return 0;
}

}  // namespace examg

using namespace examg;

extern "C" int examg_dot_dense_noalloc(const double *a_dev, const double *b_dev, long long n, double *out_dev,
                                       void *cuda_stream) {
  if (!out_dev || (n > 0 && (!a_dev || !b_dev))) return set_error(EXAMG_EINVAL, "examg_dot_dense_noalloc: null pointer");
  if (n < 0) return set_error(EXAMG_EINVAL, "examg_dot_dense_noalloc: negative length %lld", n);
  if (n > EXAMG_NOALLOC_MAX_N)
    return set_error(EXAMG_EINVAL,
                     "examg_dot_dense_noalloc: %lld elements exceed the maximum of %lld the static scratch array "
                     "covers (EXAMG_NOALLOC_MAX_N); use examg_dot",
                     n, (long long)EXAMG_NOALLOC_MAX_N);
  unsigned char *scratch = nullptr;
  if (int rc = noalloc_scratch(&scratch)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  // the ticket counter must start at zero whatever an earlier (possibly aborted) call left behind
  EXAMG_CUDA(cudaMemsetAsync(scratch, 0, sizeof(unsigned int), st));
  int nblocks = (int)((n + NOALLOC_CHUNK - 1) / NOALLOC_CHUNK);
  if (nblocks < 1) nblocks = 1;  // n == 0: one block writes 0
  k_dot_dense_noalloc<<<nblocks, 256, 0, st>>>(a_dev, b_dev, n, scratch, out_dev);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error(EXAMG_ECUDA, "examg_dot_dense_noalloc launch: %s", cudaGetErrorString(e));
  return 0;
}
