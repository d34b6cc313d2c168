// examg_comm.cu -- `communicate F` (halo exchange), global reductions and the gather/scatter of
// the agglomerated coarse levels, one process per GPU, over NVLink peer memory.
//
// Reference: C/s/e/communication/ir/IR_CommunicateFunction.scala:412-470 (one dimension after the
// other; ghost ranges IR_PackInfoGhost.scala:13-60, duplicate sync IR_PackInfoDuplicate.scala:15-39),
// MPI point-to-point C/s/e/communication/ir/IR_RemoteSend.scala:38-58 and MPI_Allreduce after
// reduction loops C/s/e/parallelization/api/mpi/MPI_Reduction.scala:92-127.
//
// Design (B200 / NVSwitch):
//   * wiring: every rank exports one CUDA IPC handle per device block (Block, examg_internal.h)
//     plus (field, buffer, block, offset) entries and its flag page; the host process group
//     all-gathers the blobs (examg_comm_export / examg_comm_import); a rank maps the blocks of its
//     face neighbours and of rank 0, and the flag pages of everybody.
//   * one `communicate` = ONE kernel and ONE handshake (k_exchange_all): tell the neighbours "ready to
//     receive" (flag in their memory), wait for theirs, store the boundary layers straight into the ghost
//     layers of the face, edge and corner neighbours (remote stores over NVLink, no pack buffer, no host),
//     fence, tell them "done", wait for theirs.  The reference runs the dimensions one after the other and
//     forwards received ghosts so that edges and corners become valid (three dependent message rounds);
//     behind an NVSwitch every peer is one hop away, so the diagonal neighbours are written directly.
//     `depth` ghost layers + the duplicate layer travel (+dir block's node 0 := -dir block's node n-1, as
//     IR_PackInfoDuplicate).
//   * epochs are counted on the device, so the same kernels replay inside a CUDA graph.
//   * sums: every rank stores its double-double partial into a mailbox of every rank and adds the
//     mailboxes in rank order -- deterministic and identical on all ranks.
//   * every wait for a peer is bounded (EXAMG_PEER_TIMEOUT_S, default 60 s, 0 = for ever as MPI would):
//     on a timeout the operation is ABANDONED -- no remote store, no signal, no epoch bump, partial sums
//     untouched -- the failure is sticky (later communication kernels return at once) and every host
//     synchronisation of the context reports EXAMG_ESTATE (comm_poll_error through a mapped pinned word).
//     Ranks must therefore reach matching calls within that time of one another.
// Kernels of different ranks wait for one another: ranks must sit on different GPUs
// (examg_comm_import refuses two ranks on one device).
#include <array>
#include <cstring>

#include "examg_device.cuh"

using namespace examg;

namespace {

constexpr int MAXW = 64;                 // ranks
constexpr int FLAG_WORDS = 1024;         // 8 KiB flag page
// word indices in the flag page
constexpr int F_ERROR = 15;              // set locally when a wait timed out (sticky: every later kernel returns at once)
constexpr int F_EPOCH = 16;              // local: [0] exchange, [3] reduction, [4] gather, [5] scatter
constexpr int F_TICKET = 24;             // local: [0] block counter of the exchange kernel, [4] gather, [5] scatter
constexpr int F_AGG_IN = 32;             // on rank 0: [rank] gather epoch
constexpr int F_AGG_OUT = 96;            // on every rank: scatter epoch, written by rank 0
constexpr int F_MAIL = 128;              // [parity][rank]{hi, lo, epoch}
constexpr int F_READY = 512;             // [direction code]  written by the neighbour in that direction
constexpr int F_DONE = 544;              // [direction code]
constexpr int NPEER = 26;                // face, edge and corner neighbours

constexpr unsigned BLOB_MAGIC = 0x45584d47u;  // "EXMG"
constexpr int BLOB_MAX_BLOCKS = 96, BLOB_MAX_ENTRIES = 512;

struct BlobEntry { int field, buf, block; int pad; long long offset; };
struct Blob {
  unsigned magic;
  int rank, device, nblocks, nentries, flag_block;
  long long flag_offset;
  long long agg_in_off, agg_out_off;  // staging buffers of the agglomeration (rank 0), -1: none
  int agg_in_block, agg_out_block;
  // receive staging of the x faces (side 0: from the - neighbour, 1: from the + neighbour), -1: none
  int xst_block[2];
  long long xst_off[2];
  long long xst_cap;                  // doubles per field component
  cudaIpcMemHandle_t handles[BLOB_MAX_BLOCKS];
  BlobEntry entries[BLOB_MAX_ENTRIES];
};

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// What every kernel of this file gets about waiting: the local flag page, the host-visible error word and the
// longest time a wait for a peer may take (0 = wait for ever, as MPI would).
struct WaitCtl {
  unsigned long long *flags;
  int *err_host;                  // mapped pinned memory: the host sees a failure without a copy
  unsigned long long timeout_ns;
};
// wait until *p >= want; false on timeout
__device__ __forceinline__ bool spin_ge(const unsigned long long *p, unsigned long long want, const WaitCtl &w) {
  if (ld_acquire_sys(p) >= want) return true;
  const unsigned long long t0 = global_ns();
  while (ld_acquire_sys(p) < want) {
    if (w.timeout_ns && global_ns() - t0 > w.timeout_ns) return false;
    __nanosleep(100);
  }
  return true;
}
// A wait timed out: the operation is abandoned (no remote stores, no signals, no epoch bump) and the failure is
// sticky -- every later communication kernel of this context returns at once, the host reports EXAMG_ESTATE at its
// next synchronisation (comm_poll_error).
__device__ __forceinline__ void comm_fail(const WaitCtl &w) {
  *reinterpret_cast<volatile unsigned long long *>(w.flags + F_ERROR) = 1ULL;
  *reinterpret_cast<volatile int *>(w.err_host) = 1;
  __threadfence_system();
}
__device__ __forceinline__ bool comm_dead(const WaitCtl &w) {
  return *reinterpret_cast<volatile unsigned long long *>(w.flags + F_ERROR) != 0ULL;
}

constexpr int EXF = 4;  // field components per exchange kernel

// `communicate` of up to EXF field components with ONE handshake: every block stores its boundary layers straight
// into the ghost layers of its face, edge AND corner neighbours (the reference reaches edges and corners by running
// the dimensions one after the other and forwarding ghosts, IR_CommunicateFunction.scala:412-470: three dependent
// message rounds; over NVSwitch every peer is one hop away, so the diagonal neighbours are written directly).
// Direction v of a peer, per dimension: -1: my layers 1..G -> its n..n+G-1;  +1: my layers n-1-G..n-1 -> its -G..0
// (ghosts + the duplicate layer, IR_PackInfoDuplicate.scala:15-39);  0: my nodes 0..n-1 -> the same nodes.
// Nodes shared by several blocks are written more than once with bit-identical values (arithmetic contract).
struct ExchField {
  const double *src;  // node (0,0,0) of the local buffer
  int n[3];           // own points per dimension: nodes (duplicate layer included) or cells
  int pitch;
  long long plane;
  int depth;
  int dup;            // 1: node-localised (the blocks share their boundary nodes), 0: cell-localised -- nothing is shared:
                      // my cells 0..G-1 -> its n..n+G-1 and my cells n-G..n-1 -> its -G..-1 (IR_PackInfoGhost.scala:13-60)
};
struct ExchPeer {
  int v[3];
  int code;                    // (vx+1) + 3 (vy+1) + 9 (vz+1): index of the flag this peer writes here
  unsigned long long *pflags;  // its flag page
  double *dst[EXF];            // node (0,0,0) of its buffer per field
  double *stage;               // pure x faces: its receive staging for this block (nullptr: store in place)
};
struct ExchArgs {
  int nf, np;
  long long stage_cap;         // doubles per field component in a staging array
  ExchField f[EXF];
  ExchPeer p[NPEER];
  WaitCtl w;
};

__global__ void __launch_bounds__(256) k_exchange_all(const __grid_constant__ ExchArgs a) {
  unsigned long long *fl = a.w.flags;
  if (comm_dead(a.w)) return;
  const unsigned long long e = *reinterpret_cast<volatile unsigned long long *>(fl + F_EPOCH) + 1;
  __shared__ int ok;
  if (threadIdx.x == 0) ok = 1;
  __syncthreads();
  // ready-to-receive: this rank no longer reads the ghost layers of the previous epoch
  if ((int)threadIdx.x < a.np) {
    const ExchPeer &p = a.p[threadIdx.x];
    if (blockIdx.x == 0) st_release_sys(p.pflags + F_READY + (26 - p.code), e);
    if (!spin_ge(fl + F_READY + p.code, e, a.w)) ok = 0;
  }
  __syncthreads();
  if (!ok) {
    if (threadIdx.x == 0) comm_fail(a.w);
    return;
  }
  const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (int pi = 0; pi < a.np; ++pi) {
    const ExchPeer &p = a.p[pi];
    for (int q = 0; q < a.nf; ++q) {
      const ExchField &f = a.f[q];
      const int G = f.depth;
      int slo[3], dlo[3], ext[3];
#pragma unroll
      for (int w = 0; w < 3; ++w) {
        if (p.v[w] == 0) { slo[w] = 0; dlo[w] = 0; ext[w] = f.n[w]; }
        else if (p.v[w] < 0) { slo[w] = f.dup; dlo[w] = f.n[w]; ext[w] = G; }
        else { slo[w] = f.n[w] - f.dup - G; dlo[w] = -G; ext[w] = G + f.dup; }
      }
      const double *sb = f.src + slo[0] + (long long)f.pitch * slo[1] + f.plane * slo[2];
      double *db = p.dst[q] + dlo[0] + (long long)f.pitch * dlo[1] + f.plane * dlo[2];
      if (p.v[0] == 0) {
        // whole rows of the block's own nodes: node 0 of a row is 128-byte aligned on both sides -> 128-bit accesses
        // (the odd last node of a row travels alone), four independent loads per thread in flight
        const int h = (ext[0] + 1) >> 1;
        const long long total = (long long)h * ext[1] * ext[2];
        for (long long t0 = tid; t0 < total; t0 += 4 * stride) {
          double2 val[4];
          long long off[4];
          bool pair[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const long long t = t0 + u * stride;
            if (t < total) {
              const int i2 = (int)(t % h);
              const long long r = t / h;
              const int j = (int)(r % ext[1]), k = (int)(r / ext[1]);
              off[u] = 2 * i2 + (long long)f.pitch * j + f.plane * k;
              pair[u] = 2 * i2 + 1 < ext[0];
              if (pair[u]) val[u] = *reinterpret_cast<const double2 *>(sb + off[u]);
              else val[u].x = sb[off[u]];
            }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u)
            if (t0 + u * stride < total) {
              if (pair[u]) *reinterpret_cast<double2 *>(db + off[u]) = val[u];
              else db[off[u]] = val[u].x;
            }
        }
      } else {
        // x faces, and edges / corners with an x component: G (G+1) nodes per row; consecutive threads take
        // consecutive nodes, so a warp moves whole 32-byte sectors of 32/G rows.  A pure x face has one such
        // 32-byte piece per row of the block -- a million NVLink packets at 1024^3, measured 2.7 ms against 0.09 ms
        // for a z face of the same size -- so it travels PACKED: the same threads store their values contiguously
        // (256 bytes per warp) into a staging array on the receiver, which scatters them into its ghost columns
        // itself (k_unpack_x, a local kernel behind this one on the receiver's stream).
        const long long total = (long long)ext[0] * ext[1] * ext[2];
        if (p.stage && p.v[1] == 0 && p.v[2] == 0) {
          double *stg = p.stage + q * a.stage_cap;
          for (long long t0 = tid; t0 < total; t0 += 4 * stride) {
            double val[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const long long t = t0 + u * stride;
              if (t < total) {
                const int i = (int)(t % ext[0]);
                const long long r = t / ext[0];
                const int j = (int)(r % ext[1]), k = (int)(r / ext[1]);
                val[u] = sb[i + (long long)f.pitch * j + f.plane * k];
              }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
              if (t0 + u * stride < total) stg[t0 + u * stride] = val[u];
          }
          continue;
        }
        for (long long t0 = tid; t0 < total; t0 += 4 * stride) {
          double val[4];
          long long off[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const long long t = t0 + u * stride;
            if (t < total) {
              const int i = (int)(t % ext[0]);
              const long long r = t / ext[0];
              const int j = (int)(r % ext[1]), k = (int)(r / ext[1]);
              off[u] = i + (long long)f.pitch * j + f.plane * k;
              val[u] = sb[off[u]];
            }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u)
            if (t0 + u * stride < total) db[off[u]] = val[u];
        }
      }
    }
  }
  __threadfence_system();
  __syncthreads();
  __shared__ int last;
  if (threadIdx.x == 0) last = atomicAdd(fl + F_TICKET, 1ULL) == gridDim.x - 1;
  __syncthreads();
  if (!last) return;
  __threadfence_system();
  if ((int)threadIdx.x < a.np) {
    const ExchPeer &p = a.p[threadIdx.x];
    st_release_sys(p.pflags + F_DONE + (26 - p.code), e);
    if (!spin_ge(fl + F_DONE + p.code, e, a.w)) ok = 0;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (!ok) { comm_fail(a.w); return; }
    fl[F_TICKET] = 0;
    *reinterpret_cast<volatile unsigned long long *>(fl + F_EPOCH) = e;
  }
}

// Receiver side of the packed x faces: staging (written by the x neighbours during k_exchange_all, complete once
// that kernel has seen their DONE) -> ghost columns.  From the - neighbour: its layers n-1-G..n-1 -> my -G..0; from
// the + neighbour: its layers 1..G -> my n..n+G-1; rows y in [0,n1), planes z in [0,n2) (edges and corners arrive in
// place).
struct UnpackArgs {
  int nf;
  int has[2];
  const double *stage[2];
  long long stage_cap;
  double *dst[EXF];  // node (0,0,0) of the local buffers
  int n0[EXF];
  int pitch[EXF];
  long long plane[EXF];
  int depth[EXF];
  int n1[EXF], n2[EXF];
  int dup[EXF];
};

__global__ void __launch_bounds__(256) k_unpack_x(const __grid_constant__ UnpackArgs a) {
  const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (int s = 0; s < 2; ++s) {
    if (!a.has[s]) continue;
    for (int q = 0; q < a.nf; ++q) {
      const int G = a.depth[q];
      const int ext0 = s == 0 ? G + a.dup[q] : G, dlo0 = s == 0 ? -G : a.n0[q];
      const long long total = (long long)ext0 * a.n1[q] * a.n2[q];
      const double *stg = a.stage[s] + q * a.stage_cap;
      double *db = a.dst[q] + dlo0;
      for (long long t0 = tid; t0 < total; t0 += 4 * stride) {
        double val[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (t0 + u * stride < total) val[u] = stg[t0 + u * stride];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const long long t = t0 + u * stride;
          if (t < total) {
            const int i = (int)(t % ext0);
            const long long r = t / ext0;
            const int j = (int)(r % a.n1[q]), k = (int)(r / a.n1[q]);
            db[i + (long long)a.pitch[q] * j + a.plane[q] * k] = val[u];
          }
        }
      }
    }
  }
}

struct RedArgs {
  int world, rank;
  int op;  // 0: double-double sum, 1: maximum of .hi
  WaitCtl w;
  unsigned long long *pflags[MAXW];
  dd *value;  // in: this rank's partial, out: the global result
};

__global__ void __launch_bounds__(MAXW) k_allreduce_dd(const __grid_constant__ RedArgs a) {
  unsigned long long *fl = a.w.flags;
  if (comm_dead(a.w)) return;
  const unsigned long long e = *reinterpret_cast<volatile unsigned long long *>(fl + F_EPOCH + 3) + 1;
  const int par = (int)(e & 1);
  const int r = threadIdx.x;
  __shared__ int ok;
  if (r == 0) ok = 1;
  __syncthreads();
  if (r < a.world) {
    unsigned long long *m = a.pflags[r] + F_MAIL + 3 * (par * MAXW + a.rank);
    const dd v = *a.value;
    m[0] = (unsigned long long)__double_as_longlong(v.hi);
    m[1] = (unsigned long long)__double_as_longlong(v.lo);
    __threadfence_system();
    st_release_sys(m + 2, e);
    if (!spin_ge(fl + F_MAIL + 3 * (par * MAXW + r) + 2, e, a.w)) ok = 0;
  }
  __syncthreads();
  if (r == 0) {
    if (!ok) { comm_fail(a.w); return; }  // the partial stays: no stale mailbox is added
    dd s = {0.0, 0.0};
    for (int q = 0; q < a.world; ++q) {
      const unsigned long long *m = fl + F_MAIL + 3 * (par * MAXW + q);
      dd v;
      v.hi = __longlong_as_double((long long)ld_acquire_sys(m));
      v.lo = __longlong_as_double((long long)ld_acquire_sys(m + 1));
      if (a.op == 0) dd_merge(s, v);
      else s.hi = q == 0 ? v.hi : fmax(s.hi, v.hi);
    }
    *a.value = s;
    *reinterpret_cast<volatile unsigned long long *>(fl + F_EPOCH + 3) = e;
  }
}

// dense box copy between a field layout and a dense [nz][ny][nx] staging array (gather / scatter)
struct AggArgs {
  int n[3];        // local nodes copied
  int goff[3];     // their offset in the global dense array
  int gn[3];       // global dense extents
  int pitch;
  long long plane;
  const double *fsrc;  // gather: local field (node 0,0,0)
  double *fdst;        // scatter: local field
  double *dense;       // staging array on rank 0 (peer pointer on the other ranks)
  int rank, world;
  WaitCtl w;                           // local flag page
  unsigned long long *flags0;          // rank 0's page
  unsigned long long *pflags[MAXW];    // scatter signal: rank 0 -> everybody
};

// every rank: field -> rank 0's dense array, then signal; rank 0 additionally waits for everybody
__global__ void __launch_bounds__(256) k_agg_gather(const __grid_constant__ AggArgs a) {
  unsigned long long *fl = a.w.flags;
  if (comm_dead(a.w)) return;
  const unsigned long long e = *reinterpret_cast<volatile unsigned long long *>(fl + F_EPOCH + 4) + 1;
  const long long total = (long long)a.n[0] * a.n[1] * a.n[2];
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(t % a.n[0]);
    const long long r = t / a.n[0];
    const int j = (int)(r % a.n[1]), k = (int)(r / a.n[1]);
    a.dense[(a.goff[0] + i) + (long long)a.gn[0] * ((a.goff[1] + j) + (long long)a.gn[1] * (a.goff[2] + k))] =
        a.fsrc[i + (long long)a.pitch * j + a.plane * k];
  }
  __threadfence_system();
  __syncthreads();
  __shared__ int last;
  if (threadIdx.x == 0) last = atomicAdd(fl + F_TICKET + 4, 1ULL) == gridDim.x - 1;
  __syncthreads();
  if (!last) return;
  __threadfence_system();
  __shared__ int ok;
  if (threadIdx.x == 0) {
    ok = 1;
    st_release_sys(a.flags0 + F_AGG_IN + a.rank, e);
  }
  __syncthreads();
  if (a.rank == 0 && (int)threadIdx.x < a.world && !spin_ge(fl + F_AGG_IN + threadIdx.x, e, a.w)) ok = 0;
  __syncthreads();
  if (threadIdx.x == 0) {
    if (!ok) { comm_fail(a.w); return; }
    fl[F_TICKET + 4] = 0;
    *reinterpret_cast<volatile unsigned long long *>(fl + F_EPOCH + 4) = e;
  }
}

// rank 0: signal "dense solution ready" to everybody (one block)
__global__ void __launch_bounds__(MAXW) k_agg_signal(const __grid_constant__ AggArgs a) {
  unsigned long long *fl = a.w.flags;
  if (comm_dead(a.w)) return;
  const unsigned long long e = *reinterpret_cast<volatile unsigned long long *>(fl + F_EPOCH + 5) + 1;
  __threadfence_system();
  if ((int)threadIdx.x < a.world) st_release_sys(a.pflags[threadIdx.x] + F_AGG_OUT, e);
}

// every rank: wait for the signal, pull the block out of rank 0's dense array
__global__ void __launch_bounds__(256) k_agg_scatter(const __grid_constant__ AggArgs a) {
  unsigned long long *fl = a.w.flags;
  if (comm_dead(a.w)) return;
  const unsigned long long e = *reinterpret_cast<volatile unsigned long long *>(fl + F_EPOCH + 5) + 1;
  __shared__ int ok;
  if (threadIdx.x == 0) {
    ok = spin_ge(fl + F_AGG_OUT, e, a.w) ? 1 : 0;
    if (!ok) comm_fail(a.w);
  }
  __syncthreads();
  if (!ok) return;
  const long long total = (long long)a.n[0] * a.n[1] * a.n[2];
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(t % a.n[0]);
    const long long r = t / a.n[0];
    const int j = (int)(r % a.n[1]), k = (int)(r / a.n[1]);
    a.fdst[i + (long long)a.pitch * j + a.plane * k] =
        a.dense[(a.goff[0] + i) + (long long)a.gn[0] * ((a.goff[1] + j) + (long long)a.gn[1] * (a.goff[2] + k))];
  }
  __syncthreads();
  __shared__ int last;
  if (threadIdx.x == 0) last = atomicAdd(fl + F_TICKET + 5, 1ULL) == gridDim.x - 1;
  __syncthreads();
  if (last && threadIdx.x == 0) {
    fl[F_TICKET + 5] = 0;
    *reinterpret_cast<volatile unsigned long long *>(fl + F_EPOCH + 5) = e;
  }
}

}  // namespace

struct examg_comm {
  bool wired = false;
  int world = 1, rank = 0;
  unsigned long long *flags = nullptr;              // local flag page
  std::vector<unsigned long long *> pflags;         // [rank]
  std::vector<std::vector<char *>> pblock;          // [rank][block] mapped base (nullptr: not mapped)
  std::vector<Blob> blobs;                          // [rank]
  std::vector<void *> opened;
  int nbr[3][2];                                    // face neighbour rank per dim/side, -1: none
  // face, edge and corner neighbours (`communicate` writes all of them directly)
  struct Peer { int v[3], code, rank; };
  std::vector<Peer> peers;
  double *xstage[2] = {nullptr, nullptr};           // local receive staging of packed x faces (side 0: from -x, 1: from +x)
  long long xstage_cap = 0;                         // doubles per field component
  double *pxstage[2] = {nullptr, nullptr};          // the x neighbours' staging for this block (side of the neighbour)
  int *err_host = nullptr, *err_dev = nullptr;      // mapped pinned word set by a kernel whose wait for a peer timed out
  unsigned long long timeout_ns = 60ULL * 1000000000ULL;  // EXAMG_PEER_TIMEOUT_S (0: wait for ever)
  // agglomeration staging (dense global arrays of one level), allocated by examg_comm_agg_setup
  double *agg_in = nullptr, *agg_out = nullptr;     // local pointers on rank 0
  double *agg_in0 = nullptr, *agg_out0 = nullptr;   // rank 0's arrays as seen from this rank
  int agg_level = -1;
  int agg_gn[3] = {0, 0, 0};
  // cell-localised staging (comm_agg_setup_cells): agg_in / agg_out hold several dense arrays of agg_tot cells each
  bool agg_cells = false;
  long long agg_tot = 0;
  int agg_nin = 0, agg_nout = 0;
};

namespace examg {

void comm_destroy(examg_ctx *c) {
  if (!c->comm) return;
  for (void *p : c->comm->opened) cudaIpcCloseMemHandle(p);
  if (c->comm->err_host) cudaFreeHost(c->comm->err_host);
  delete c->comm;
  c->comm = nullptr;
}

static examg_comm *comm_get(examg_ctx *c) {
  if (!c->comm) {
    c->comm = new examg_comm();
    c->comm->world = c->world;
    c->comm->rank = c->cfg.rank;
    if (const char *e = getenv("EXAMG_PEER_TIMEOUT_S")) {
      const double sec = atof(e);
      c->comm->timeout_ns = sec > 0.0 ? (unsigned long long)(sec * 1e9) : 0ULL;
    }
  }
  return c->comm;
}

static WaitCtl wait_ctl(const examg_comm *cm) {
  WaitCtl w;
  w.flags = cm->flags;
  w.err_host = cm->err_dev;
  w.timeout_ns = cm->timeout_ns;
  return w;
}

int comm_poll_error(examg_ctx *c) {
  if (!c->comm || !c->comm->err_host || !*reinterpret_cast<volatile int *>(c->comm->err_host)) return EXAMG_OK;
  return set_error(EXAMG_ESTATE, "a wait for a peer rank timed out after %.0f s (lost rank, mismatched call sequence, or ranks "
                   "further apart than EXAMG_PEER_TIMEOUT_S); the operation was abandoned and the multi-GPU state of this "
                   "context is no longer usable", (double)c->comm->timeout_ns * 1e-9);
}

static int find_block(examg_ctx *c, const void *p, long long *off) {
  const char *q = static_cast<const char *>(p);
  for (size_t b = 0; b < c->blocks.size(); ++b)
    if (q >= c->blocks[b].base && q < c->blocks[b].base + c->blocks[b].bytes) {
      *off = q - c->blocks[b].base;
      return (int)b;
    }
  return -1;
}

// the neighbour's pointer to the same buffer as local pointer p (a slot or the shadow of field f)
static double *peer_buffer(examg_ctx *c, int rank, int field, const double *p) {
  examg_comm *cm = c->comm;
  const Field &f = c->fields[field];
  int buf = -1;
  for (int i = 0; i < 5; ++i)
    if (f.orig[i] == p) buf = i;
  if (buf < 0) return nullptr;
  const Blob &bl = cm->blobs[rank];
  for (int e = 0; e < bl.nentries; ++e)
    if (bl.entries[e].field == field && bl.entries[e].buf == buf) {
      char *base = cm->pblock[rank][bl.entries[e].block];
      return base ? reinterpret_cast<double *>(base + bl.entries[e].offset) : nullptr;
    }
  return nullptr;
}

// One `communicate` of up to two fields (e.g. Solution and RHS in front of a fused sweep): one
// kernel and one handshake per dimension for all of them.
// `communicate` across a periodic boundary of a single block: the block is its own neighbour, so the face copies of
// k_exchange_dim become copies inside one array -- duplicate + ghost layers on the low side from the high end
// (the + neighbour's layers n-1-G..n-1 -> -G..0), ghost layers on the high side from the low end (1..G -> n..n+G-1),
// over the orthogonal extents including ghost layers, one dimension after the other.
struct SelfExchArgs {
  int n[3];
  int d, depth, gw[3];
  int pitch;
  long long plane;
};

__global__ void __launch_bounds__(256) k_self_exchange(const SelfExchArgs a, double *f) {
  const int G = a.depth, d = a.d;
  const int da = d == 0 ? 1 : 0, db = d == 2 ? 1 : 2;
  const int ea = a.n[da] + 2 * a.gw[da], eb = a.n[db] + 2 * a.gw[db];
  const int layers = 2 * G + 1;
  const long long total = (long long)layers * ea * eb;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    // the fastest index runs along x whenever x is not the exchanged dimension
    int l, ia, ib;
    if (d == 0) { l = (int)(t % layers); ia = (int)((t / layers) % ea); ib = (int)(t / ((long long)layers * ea)); }
    else { ia = (int)(t % ea); l = (int)((t / ea) % layers); ib = (int)(t / ((long long)ea * layers)); }
    int idx[3], src[3];
    idx[da] = src[da] = ia - a.gw[da];
    idx[db] = src[db] = ib - a.gw[db];
    if (l <= G) { idx[d] = l - G; src[d] = a.n[d] - 1 - G + l; }  // -G..0  <-  n-1-G..n-1
    else { idx[d] = a.n[d] + (l - G - 1); src[d] = 1 + (l - G - 1); }  // n..n+G-1  <-  1..G
    f[idx[0] + (long long)a.pitch * idx[1] + a.plane * idx[2]] = f[src[0] + (long long)a.pitch * src[1] + a.plane * src[2]];
  }
}

static int self_exchange(examg_ctx *c, Field *const *fs, const int *slots, const int *depths, int nfields) {
  for (int d = 0; d < c->cfg.dim; ++d) {
    if (!c->per[d]) continue;
    for (int q = 0; q < nfields; ++q) {
      Field &f = *fs[q];
      const LevelGeom &g = *geom(c, f.level);
      const int G = depths[q];
      if (G < 1 || G > GHOST) return set_error(EXAMG_EINVAL, "examg_exchange: depth %d", G);
      if (g.n[d] - 1 <= 2 * G) return set_error(EXAMG_EUNSUP, "examg_exchange: periodic level with %d cells, depth %d", g.n[d] - 1, G);
      SelfExchArgs a;
      for (int w = 0; w < 3; ++w) {
        a.n[w] = g.n[w];
        a.gw[w] = w < c->cfg.dim ? G : 0;
      }
      a.d = d;
      a.depth = G;
      a.pitch = g.pitch;
      a.plane = g.plane;
      const int da = d == 0 ? 1 : 0, db = d == 2 ? 1 : 2;
      const long long total = (long long)(2 * G + 1) * (g.n[da] + 2 * a.gw[da]) * (g.n[db] + 2 * a.gw[db]);
      int grid = (int)((total + 255) / 256);
      if (grid > 592) grid = 592;
      for (int comp = 0; comp < f.ncomp; ++comp) {
        k_self_exchange<<<grid, 256, 0, c->stream>>>(a, f.slot[slots[q]] + (long long)comp * g.vol + g.origin);
        EXAMG_LAUNCH_CHECK(c);
      }
    }
  }
  for (int q = 0; q < nfields; ++q) fs[q]->halo_depth = depths[q];
  return EXAMG_OK;
}

int do_exchange_many(examg_ctx *c, Field *const *fs, const int *slots, const int *depths, int nfields) {
  if (c->world == 1) return c->periodic_any ? self_exchange(c, fs, slots, depths, nfields) : EXAMG_OK;
  examg_comm *cm = c->comm;
  if (!cm || !cm->wired) return set_error(EXAMG_ESTATE, "examg_exchange: peers not wired (examg_comm_import)");
  if (cm->peers.empty()) return EXAMG_OK;
  ExchArgs a;
  a.w = wait_ctl(cm);
  a.np = (int)cm->peers.size();
  for (int pi = 0; pi < a.np; ++pi) {
    const examg_comm::Peer &pr = cm->peers[pi];
    for (int w = 0; w < 3; ++w) a.p[pi].v[w] = pr.v[w];
    a.p[pi].code = pr.code;
    a.p[pi].pflags = cm->pflags[pr.rank];
    a.p[pi].stage = nullptr;
    if (pr.v[0] != 0 && pr.v[1] == 0 && pr.v[2] == 0) a.p[pi].stage = cm->pxstage[pr.v[0] < 0 ? 0 : 1];
  }
  a.stage_cap = cm->xstage_cap;
  UnpackArgs ua;
  ua.has[0] = cm->nbr[0][0] >= 0 && cm->xstage[0];
  ua.has[1] = cm->nbr[0][1] >= 0 && cm->xstage[1];
  ua.stage[0] = cm->xstage[0];
  ua.stage[1] = cm->xstage[1];
  ua.stage_cap = cm->xstage_cap;
  a.nf = 0;
  long long work = 0;  // threads' worth of work (four accesses per thread)
  auto flush = [&]() -> int {
    if (a.nf == 0) return EXAMG_OK;
    int grid = (int)((work + 255) / 256);
    if (grid > 592) grid = 592;
    if (grid < 1) grid = 1;
    k_exchange_all<<<grid, 256, 0, c->stream>>>(a);
    EXAMG_LAUNCH_CHECK(c);
    if (ua.has[0] || ua.has[1]) {  // packed x faces: scatter what the x neighbours left in the staging arrays
      ua.nf = a.nf;
      long long uw = 0;
      for (int q = 0; q < a.nf; ++q) uw += (long long)(2 * ua.depth[q] + 1) * ua.n1[q] * ua.n2[q] / 4 + 1;
      int ug = (int)((uw + 255) / 256);
      if (ug > 1184) ug = 1184;
      k_unpack_x<<<ug, 256, 0, c->stream>>>(ua);
      EXAMG_LAUNCH_CHECK(c);
    }
    a.nf = 0;
    work = 0;
    return EXAMG_OK;
  };
  for (int q = 0; q < nfields; ++q) {
    Field &f = *fs[q];
    const int field = (int)(&f - c->fields.data());
    const LevelGeom &g = *geom(c, f.level);
    const int G = depths[q];
    if (G < 1 || G > GHOST) return set_error(EXAMG_EINVAL, "examg_exchange: depth %d", G);
    const int dup = f.loc == 0 ? 1 : 0;
    int np[3];  // own points: nodes, or the n - 1 cells between them
    for (int w = 0; w < 3; ++w) np[w] = (w < c->cfg.dim && !dup) ? g.n[w] - 1 : g.n[w];
    for (int w = 0; w < c->cfg.dim; ++w)
      if (g.n[w] - 1 < G)
        return set_error(EXAMG_EUNSUP, "examg_exchange: %d ghost layers of a level with %d cells per block", G, g.n[w] - 1);
    double *buf = f.slot[slots[q]];
    for (int comp = 0; comp < f.ncomp; ++comp) {
      if (a.nf == EXF) {
        int rc = flush();
        if (rc) return rc;
      }
      ExchField &ef = a.f[a.nf];
      ef.src = buf + (long long)comp * g.vol + g.origin;
      for (int w = 0; w < 3; ++w) ef.n[w] = np[w];
      ef.pitch = g.pitch;
      ef.plane = g.plane;
      ef.depth = G;
      ef.dup = dup;
      ua.dst[a.nf] = buf + (long long)comp * g.vol + g.origin;
      ua.n0[a.nf] = np[0];
      ua.pitch[a.nf] = g.pitch;
      ua.plane[a.nf] = g.plane;
      ua.depth[a.nf] = G;
      ua.n1[a.nf] = np[1];
      ua.n2[a.nf] = np[2];
      ua.dup[a.nf] = dup;
      if ((ua.has[0] || ua.has[1]) && (long long)(G + 1) * g.n[1] * g.n[2] > cm->xstage_cap)
        return set_error(EXAMG_ESTATE, "examg_exchange: x face of %s@%d exceeds the staging arrays", f.name.c_str(), f.level);
      for (int pi = 0; pi < a.np; ++pi) {
        const examg_comm::Peer &pr = cm->peers[pi];
        double *pb = peer_buffer(c, pr.rank, field, buf);
        if (!pb) return set_error(EXAMG_ESTATE, "examg_exchange: field %s@%d is not mapped on rank %d (created after "
                                  "examg_comm_export?)", f.name.c_str(), f.level, pr.rank);
        a.p[pi].dst[a.nf] = pb + (long long)comp * g.vol + g.origin;
        long long cnt = 1;
        for (int w = 0; w < 3; ++w) cnt *= pr.v[w] == 0 ? np[w] : (pr.v[w] < 0 ? G : G + dup);
        work += (pr.v[0] == 0 ? cnt / 2 : cnt) / 4 + 1;
      }
      ++a.nf;
    }
  }
  int rc = flush();
  if (rc) return rc;
  for (int q = 0; q < nfields; ++q) fs[q]->halo_depth = depths[q];
  return EXAMG_OK;
}

int do_exchange(examg_ctx *c, Field &f, int slot, int depth) {
  Field *fp = &f;
  return do_exchange_many(c, &fp, &slot, &depth, 1);
}

static int finish_reduce(examg_ctx *c, int slot, int op, double *out) {
  if (c->world > 1) {
    examg_comm *cm = c->comm;
    if (!cm || !cm->wired) return set_error(EXAMG_ESTATE, "global reduction: peers not wired (examg_comm_import)");
    RedArgs a;
    a.world = c->world;
    a.rank = c->cfg.rank;
    a.op = op;
    a.w = wait_ctl(cm);
    for (int r = 0; r < c->world; ++r) a.pflags[r] = cm->pflags[r];
    a.value = c->red_out + slot;
    k_allreduce_dd<<<1, MAXW, 0, c->stream>>>(a);
    EXAMG_LAUNCH_CHECK(c);
  }
  EXAMG_CUDA(cudaMemcpyAsync(c->red_host + slot, c->red_out + slot, sizeof(dd), cudaMemcpyDeviceToHost, c->stream));
  EXAMG_SYNC(c);
  *out = op == 0 ? c->red_host[slot].hi + c->red_host[slot].lo : c->red_host[slot].hi;
  return EXAMG_OK;
}

int finish_sum(examg_ctx *c, int slot, double *out) { return finish_reduce(c, slot, 0, out); }

// n global sums at once (red_out[first .. first+n-1]): one all-reduce launch per slot, ONE host synchronisation
int finish_sums(examg_ctx *c, int first, int n, double *out) {
  if (first < 0 || n < 1 || first + n > RED_SLOTS) return set_error(EXAMG_EINVAL, "finish_sums: slots %d..%d", first, first + n - 1);
  if (c->world > 1) {
    examg_comm *cm = c->comm;
    if (!cm || !cm->wired) return set_error(EXAMG_ESTATE, "global reduction: peers not wired (examg_comm_import)");
    for (int k = 0; k < n; ++k) {
      RedArgs a;
      a.world = c->world;
      a.rank = c->cfg.rank;
      a.op = 0;
      a.w = wait_ctl(cm);
      for (int r = 0; r < c->world; ++r) a.pflags[r] = cm->pflags[r];
      a.value = c->red_out + first + k;
      k_allreduce_dd<<<1, MAXW, 0, c->stream>>>(a);
      EXAMG_LAUNCH_CHECK(c);
    }
  }
  EXAMG_CUDA(cudaMemcpyAsync(c->red_host + first, c->red_out + first, sizeof(dd) * n, cudaMemcpyDeviceToHost, c->stream));
  EXAMG_SYNC(c);
  for (int k = 0; k < n; ++k) out[k] = c->red_host[first + k].hi + c->red_host[first + k].lo;
  return EXAMG_OK;
}
int finish_max(examg_ctx *c, int slot, double *out) { return finish_reduce(c, slot, 1, out); }

// ---- agglomeration staging -------------------------------------------------------------------

// Dense global arrays of `level` (RHS in, Solution out) on rank 0.  Must run on every rank before
// examg_comm_export.  gn = global node counts.
int comm_agg_setup(examg_ctx *c, int level) {
  examg_comm *cm = comm_get(c);
  const LevelGeom &g = *geom(c, level);
  cm->agg_level = level;
  size_t tot = 1;
  for (int d = 0; d < 3; ++d) {
    cm->agg_gn[d] = d < c->cfg.dim ? (g.n[d] - 1) * c->cfg.num_blocks[d] + 1 : 1;
    tot *= (size_t)cm->agg_gn[d];
  }
  if (c->cfg.rank == 0 && !cm->agg_in) {
    cm->agg_in = dev_alloc(c, tot * sizeof(double));
    cm->agg_out = dev_alloc(c, tot * sizeof(double));
    if (!cm->agg_in || !cm->agg_out) return set_error(EXAMG_ENOMEM, "agglomeration staging of level %d", level);
  }
  return EXAMG_OK;
}

// The same for cell-localised fields (nothing shared between blocks): `nin` dense arrays in, `nout` out and `extra`
// more doubles behind them, on rank 0.  Before examg_comm_export, on every rank.
int comm_agg_setup_cells(examg_ctx *c, int level, int nin, int nout, int extra) {
  examg_comm *cm = comm_get(c);
  if (cm->agg_level >= 0) return set_error(EXAMG_ESTATE, "agglomeration staging is already set up (level %d)", cm->agg_level);
  const LevelGeom &g = *geom(c, level);
  cm->agg_level = level;
  cm->agg_cells = true;
  cm->agg_nin = nin;
  cm->agg_nout = nout;
  long long tot = 1;
  for (int d = 0; d < 3; ++d) {
    cm->agg_gn[d] = d < c->cfg.dim ? (g.n[d] - 1) * c->cfg.num_blocks[d] : 1;
    tot *= cm->agg_gn[d];
  }
  cm->agg_tot = tot;
  if (c->cfg.rank == 0) {
    cm->agg_in = dev_alloc(c, (size_t)nin * tot * sizeof(double));
    cm->agg_out = dev_alloc(c, ((size_t)nout * tot + extra) * sizeof(double));
    if (!cm->agg_in || !cm->agg_out) return set_error(EXAMG_ENOMEM, "agglomeration staging of level %d", level);
  }
  return EXAMG_OK;
}

// level of the cell staging (-1: none), global cell counts, cells per dense array
int comm_agg_cells(examg_ctx *c, int gn[3], long long *tot) {
  examg_comm *cm = c->comm;
  if (!cm || !cm->agg_cells) return -1;
  for (int d = 0; d < 3; ++d) gn[d] = cm->agg_gn[d];
  *tot = cm->agg_tot;
  return cm->agg_level;
}

// rank 0's dense "out" arrays as mapped into this rank
const double *comm_agg_out_peer(examg_ctx *c) { return c->comm ? c->comm->agg_out0 : nullptr; }

double *comm_agg_in(examg_ctx *c) { return c->comm ? c->comm->agg_in : nullptr; }
double *comm_agg_out(examg_ctx *c) { return c->comm ? c->comm->agg_out : nullptr; }

static int agg_args(examg_ctx *c, const Field &f, AggArgs *a) {
  examg_comm *cm = c->comm;
  if (!cm || !cm->wired) return set_error(EXAMG_ESTATE, "agglomeration: peers not wired (examg_comm_import)");
  if (f.level != cm->agg_level) return set_error(EXAMG_EINVAL, "agglomeration: field level %d, staging level %d", f.level, cm->agg_level);
  if (cm->agg_cells != (f.loc == 1))
    return set_error(EXAMG_EINVAL, "agglomeration: the staging of level %d is for %s fields", f.level, cm->agg_cells ? "cell" : "node");
  const LevelGeom &g = *geom(c, f.level);
  for (int d = 0; d < 3; ++d) {
    a->n[d] = cm->agg_cells && d < c->cfg.dim ? g.n[d] - 1 : g.n[d];
    a->goff[d] = g.goff[d];
    a->gn[d] = cm->agg_gn[d];
  }
  a->pitch = g.pitch;
  a->plane = g.plane;
  a->rank = c->cfg.rank;
  a->world = c->world;
  a->w = wait_ctl(cm);
  a->flags0 = cm->pflags[0];
  for (int r = 0; r < c->world; ++r) a->pflags[r] = cm->pflags[r];
  a->fsrc = nullptr;
  a->fdst = nullptr;
  a->dense = nullptr;
  return EXAMG_OK;
}

static int agg_grid(const AggArgs &a) {
  const long long total = (long long)a.n[0] * a.n[1] * a.n[2];
  long long grid = (total + 2047) / 2048;
  return (int)(grid > 128 ? 128 : (grid < 1 ? 1 : grid));
}

// all ranks: field (active slot) -> rank 0's dense "in" array number `comp`; rank 0 returns when all blocks arrived
int comm_agg_gather(examg_ctx *c, Field &f, int comp) {
  AggArgs a;
  int rc = agg_args(c, f, &a);
  if (rc) return rc;
  if (comp < 0 || comp >= (c->comm->agg_cells ? c->comm->agg_nin : 1)) return set_error(EXAMG_EINVAL, "agglomeration: in array %d", comp);
  a.fsrc = field_ptr(c, &f, -1);
  a.dense = c->comm->agg_in0 + comp * c->comm->agg_tot;
  k_agg_gather<<<agg_grid(a), 256, 0, c->stream>>>(a);
  EXAMG_LAUNCH_CHECK(c);
  return EXAMG_OK;
}

// rank 0 (after filling the dense "out" array): signal; all ranks: dense "out" -> field (active slot)
int comm_agg_scatter(examg_ctx *c, Field &f, int comp) {
  AggArgs a;
  int rc = agg_args(c, f, &a);
  if (rc) return rc;
  if (comp < 0 || comp >= (c->comm->agg_cells ? c->comm->agg_nout : 1)) return set_error(EXAMG_EINVAL, "agglomeration: out array %d", comp);
  a.fdst = field_ptr(c, &f, -1);
  a.dense = c->comm->agg_out0 + comp * c->comm->agg_tot;
  if (c->cfg.rank == 0) {
    k_agg_signal<<<1, MAXW, 0, c->stream>>>(a);
    EXAMG_LAUNCH_CHECK(c);
  }
  k_agg_scatter<<<agg_grid(a), 256, 0, c->stream>>>(a);
  EXAMG_LAUNCH_CHECK(c);
  f.halo_depth = 0;
  return EXAMG_OK;
}

}  // namespace examg

extern "C" {

int examg_comm_blob_size(examg_ctx *c) {
  (void)c;
  return (int)sizeof(Blob);
}

int examg_comm_export(examg_ctx *c, void *blob) {
  if (!c || !blob) return set_error(EXAMG_EINVAL, "examg_comm_export: null argument");
  examg_comm *cm = comm_get(c);
  if (!cm->flags) {
    cm->flags = reinterpret_cast<unsigned long long *>(dev_alloc(c, FLAG_WORDS * sizeof(unsigned long long)));
    if (!cm->flags) return set_error(EXAMG_ENOMEM, "examg_comm_export: flag page");
  }
  if (!cm->err_host) {
    EXAMG_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&cm->err_host), sizeof(int), cudaHostAllocMapped));
    *cm->err_host = 0;
    EXAMG_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void **>(&cm->err_dev), cm->err_host, 0));
  }
  EXAMG_SYNC(c);
  Blob *b = static_cast<Blob *>(blob);
  memset(b, 0, sizeof(Blob));
  b->magic = BLOB_MAGIC;
  b->rank = c->cfg.rank;
  b->device = c->device;
  // packed x faces: receive staging for the finest level, one array per x neighbour (EXF components each)
  if (c->cfg.num_blocks[0] > 1 && !cm->xstage_cap) {
    const LevelGeom &gf = *geom(c, c->cfg.max_level);
    cm->xstage_cap = (long long)(GHOST + 1) * gf.n[1] * gf.n[2];
    for (int sd = 0; sd < 2; ++sd) {
      if (!c->nb[0][sd]) continue;
      cm->xstage[sd] = dev_alloc(c, sizeof(double) * (size_t)cm->xstage_cap * EXF);
      if (!cm->xstage[sd]) return set_error(EXAMG_ENOMEM, "examg_comm_export: staging of the x faces");
    }
  }
  b->flag_block = find_block(c, cm->flags, &b->flag_offset);
  b->agg_in_block = b->agg_out_block = -1;
  b->xst_cap = cm->xstage_cap;
  for (int sd = 0; sd < 2; ++sd) {
    b->xst_block[sd] = -1;
    b->xst_off[sd] = 0;
  }
  if (cm->agg_in) {
    b->agg_in_block = find_block(c, cm->agg_in, &b->agg_in_off);
    b->agg_out_block = find_block(c, cm->agg_out, &b->agg_out_off);
  }
  // (after the allocations above: the block list is final now)
  b->nblocks = (int)c->blocks.size();
  if (b->nblocks > BLOB_MAX_BLOCKS) return set_error(EXAMG_EUNSUP, "examg_comm_export: %d device blocks", b->nblocks);
  for (int i = 0; i < b->nblocks; ++i) EXAMG_CUDA(cudaIpcGetMemHandle(&b->handles[i], c->blocks[i].base));
  for (int sd = 0; sd < 2; ++sd)
    if (cm->xstage[sd]) b->xst_block[sd] = find_block(c, cm->xstage[sd], &b->xst_off[sd]);
  int ne = 0;
  for (size_t fi = 0; fi < c->fields.size(); ++fi) {
    const Field &f = c->fields[fi];
    if (!f.alive) continue;
    for (int i = 0; i < 5; ++i) {
      if (!f.orig[i]) continue;
      if (ne >= BLOB_MAX_ENTRIES) return set_error(EXAMG_EUNSUP, "examg_comm_export: more than %d field buffers", BLOB_MAX_ENTRIES);
      BlobEntry &e = b->entries[ne++];
      e.field = (int)fi;
      e.buf = i;
      e.block = find_block(c, f.orig[i], &e.offset);
    }
  }
  b->nentries = ne;
  return EXAMG_OK;
}

int examg_comm_import(examg_ctx *c, const void *all_blobs, int world) {
  if (!c || !all_blobs) return set_error(EXAMG_EINVAL, "examg_comm_import: null argument");
  if (world != c->world) return set_error(EXAMG_EINVAL, "examg_comm_import: %d blobs for %d blocks", world, c->world);
  if (world > MAXW) return set_error(EXAMG_EUNSUP, "examg_comm_import: more than %d ranks", MAXW);
  examg_comm *cm = comm_get(c);
  if (!cm->flags) return set_error(EXAMG_ESTATE, "examg_comm_import before examg_comm_export");
  if (cm->wired) return set_error(EXAMG_ESTATE, "examg_comm_import: already wired");
  const Blob *bl = static_cast<const Blob *>(all_blobs);
  const int me = c->cfg.rank;
  for (int r = 0; r < world; ++r) {
    if (bl[r].magic != BLOB_MAGIC || bl[r].rank != r) return set_error(EXAMG_EINVAL, "examg_comm_import: blob %d is malformed", r);
    if (r != me && bl[r].device == c->device)
      return set_error(EXAMG_EUNSUP, "examg_comm_import: ranks %d and %d share device %d; ranks wait for each other inside "
                       "kernels and need one GPU each", me, r, c->device);
  }
  cm->blobs.assign(bl, bl + world);
  // neighbours (rank -> block x fastest, C/s/e/domain/ir/IR_InitGeneratedDomain.scala:41-47)
  const int *nb = c->cfg.num_blocks;
  for (int d = 0; d < 3; ++d)
    for (int s = 0; s < 2; ++s) {
      cm->nbr[d][s] = -1;
      if (d >= c->cfg.dim || !c->nb[d][s]) continue;
      int idx[3] = {c->bidx[0], c->bidx[1], c->bidx[2]};
      idx[d] += s ? 1 : -1;
      cm->nbr[d][s] = idx[0] + nb[0] * (idx[1] + (c->cfg.dim > 1 ? nb[1] : 1) * idx[2]);
    }
  cm->peers.clear();
  for (int vz = -1; vz <= 1; ++vz)
    for (int vy = -1; vy <= 1; ++vy)
      for (int vx = -1; vx <= 1; ++vx) {
        const int v[3] = {vx, vy, vz};
        if (!vx && !vy && !vz) continue;
        bool ok = true;
        int idx[3];
        for (int d = 0; d < 3; ++d) {
          idx[d] = c->bidx[d] + v[d];
          if (v[d] && (d >= c->cfg.dim || idx[d] < 0 || idx[d] >= nb[d])) ok = false;
        }
        if (!ok) continue;
        examg_comm::Peer pr;
        for (int d = 0; d < 3; ++d) pr.v[d] = v[d];
        pr.code = (vx + 1) + 3 * (vy + 1) + 9 * (vz + 1);
        pr.rank = idx[0] + nb[0] * (idx[1] + (c->cfg.dim > 1 ? nb[1] : 1) * idx[2]);
        cm->peers.push_back(pr);
      }
  cm->pblock.assign(world, std::vector<char *>());
  cm->pflags.assign(world, nullptr);
  auto need_all = [&](int r) {
    if (r == 0) return true;  // agglomeration staging lives on rank 0
    for (const auto &pr : cm->peers)
      if (pr.rank == r) return true;
    return false;
  };
  for (int r = 0; r < world; ++r) {
    cm->pblock[r].assign(bl[r].nblocks, nullptr);
    if (r == me) {
      for (int b = 0; b < bl[r].nblocks; ++b) cm->pblock[r][b] = c->blocks[b].base;
      cm->pflags[r] = cm->flags;
      continue;
    }
    const bool all = need_all(r);
    for (int b = 0; b < bl[r].nblocks; ++b) {
      if (!all && b != bl[r].flag_block) continue;
      void *p = nullptr;
      cudaError_t e = cudaIpcOpenMemHandle(&p, bl[r].handles[b], cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) {
        cudaGetLastError();
        return set_error(EXAMG_ECUDA, "examg_comm_import: cudaIpcOpenMemHandle(rank %d, block %d): %s", r, b, cudaGetErrorString(e));
      }
      cm->opened.push_back(p);
      cm->pblock[r][b] = static_cast<char *>(p);
    }
    cm->pflags[r] = reinterpret_cast<unsigned long long *>(cm->pblock[r][bl[r].flag_block] + bl[r].flag_offset);
  }
  for (int sd = 0; sd < 2; ++sd) {
    cm->pxstage[sd] = nullptr;
    const int r = cm->nbr[0][sd];
    if (r < 0) continue;
    // this block sits on side 1-sd of its neighbour on side sd
    if (bl[r].xst_block[1 - sd] < 0 || bl[r].xst_cap != cm->xstage_cap)
      return set_error(EXAMG_ESTATE, "examg_comm_import: rank %d exported no staging for its x face", r);
    cm->pxstage[sd] = reinterpret_cast<double *>(cm->pblock[r][bl[r].xst_block[1 - sd]] + bl[r].xst_off[1 - sd]);
  }
  if (bl[0].agg_in_block >= 0) {
    cm->agg_in0 = reinterpret_cast<double *>(cm->pblock[0][bl[0].agg_in_block] + bl[0].agg_in_off);
    cm->agg_out0 = reinterpret_cast<double *>(cm->pblock[0][bl[0].agg_out_block] + bl[0].agg_out_off);
  }
  cm->wired = true;
  return EXAMG_OK;
}

}  // extern "C"
