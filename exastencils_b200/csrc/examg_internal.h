// examg_internal.h -- data structures shared by the translation units of libexamg.
// Not part of the ABI (include/examg.h is).
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <map>
#include <string>
#include <vector>

#include "../../include/examg.h"

namespace examg {

// ---------------------------------------------------------------------------
// HBM layout of one scalar component of a node field on one level
// (replaces [pad|ghost|dup|inner|dup|ghost|pad]^d of IR_FieldLayout.scala:30-129):
//
//   element(i,j,k) = origin + i + pitch*(j + ay*k)       i,j,k in [-1, n_d]
//
//   x: XPAD = 16 doubles in front of node 0, so node 0 of every row sits on a
//      128-byte boundary (the left ghost is element 15 of the pad); pitch is a
//      multiple of 16 doubles -> rows are 128-byte aligned, which is what TMA
//      (16 B) and 128-bit vector accesses need.
//   y,z: GHOST ghost rows/planes per side (z only in 3D).  Four layers, because two
//      fused red+black sweeps (4 half-sweeps) need a 4-deep halo: one exchange per
//      two sweeps instead of one per colour; the B200 form of the reference's
//      contraction loops, C/s/e/baseExt/ir/IR_ContractingLoop.scala:48-90.
// ---------------------------------------------------------------------------
constexpr int XPAD = 16;
constexpr int GHOST = 4;

struct LevelGeom {
  int n[3];        // local nodes per dim incl. duplicate/boundary nodes (n[2] = 1 in 2D)
  int lo[3], hi[3];  // loop range [lo,hi): physical-boundary nodes excluded
  int rlo[3];      // reduction range lower bound (shared nodes counted once)
  int goff[3];     // global node index of local node 0
  int pitch, ay, az;
  long long plane;  // pitch*ay
  long long vol;    // pitch*ay*az  (elements per component)
  long long origin; // element offset of node (0,0,0)
  double h[3];
  int par;         // (goff[0]+goff[1]+goff[2]) & 1 : local -> global colour
};

// Passed by value to kernels.
struct GridArgs {
  int n[3];
  int lo[3], hi[3];
  int rlo[3], rhi[3];  // reduction range (shared nodes counted once, no redundantly computed ghost nodes)
  int pitch;
  long long plane;
  int par;
};

struct StencilArgs {
  int nst;
  int centre;
  long long lin[EXAMG_MAX_STENCIL];
  double coef[EXAMG_MAX_STENCIL];
};

struct Field {
  bool alive = false;
  std::string name;
  int level = 0, nslots = 1, ncomp = 1, cur = 0;
  double *slot[4] = {nullptr, nullptr, nullptr, nullptr};  // one allocation of ncomp*vol doubles per slot
  double *shadow = nullptr;  // second buffer of the active slot for out-of-place fused sweeps (lazy)
  bool swapped = false;      // slot[cur] and shadow have been exchanged an odd number of times
  // The allocations in creation order (slots 0..3, shadow = 4).  slot[]/shadow are permuted by the
  // fused sweeps; orig[] is not, and identifies a buffer across ranks (examg_comm.cu).
  double *orig[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  int loc = 0;               // 0: node-localised, 1: cell-localised (n-1 cells per dim, ghost cells at -1 and n-1)
  int halo_depth = 0;        // this many ghost layers hold the neighbour blocks' current values
  int bc_kind = EXAMG_BC_NONE;
  double bc_value = 0.0;
  std::vector<examg_bc_op> bc_prog;
  // per-face overrides (cell fields): kind < 0 = the field's condition
  int bc_face_kind[6] = {-1, -1, -1, -1, -1, -1};
  double bc_face_value[6] = {0, 0, 0, 0, 0, 0};
  std::vector<examg_bc_op> bc_face_prog[6];
  // examg_field_eval programs applied to slot 0 since the last upload / fill, in order: a field whose every component
  // was produced this way can be re-created on another grid (the agglomerated coarse levels of a multi-block run)
  std::vector<std::pair<int, std::vector<examg_bc_op>>> evals;
  long long eval_gen = 0;  // bumped whenever the contents change by eval, upload or fill
};

struct Stencil {
  int level = 0, nst = 0, centre = 0;
  int off[EXAMG_MAX_STENCIL][3];
  double coef[EXAMG_MAX_STENCIL];
  int coeff_field = -1;  // -1: constant coefficients
  bool is7 = false;      // 3D 7-point constant stencil (fast path)
  // red-black sweeps over stencil-field coefficients: a copy of the coefficient field with every row split into its
  // even and its odd elements (examg_kernels.cu, k_rbv), valid for coefficient generation rbc_gen
  double *rbc = nullptr;
  long long rbc_gen = -1;
};

// [c_uu + a2*L, c_uv; c_uv, c_vv + a2*L] on two cell fields (the optical-flow operator, examg_optflow.cu)
struct System2 {
  int level = 0;
  double a2 = 1.0;
  int cuu = -1, cuv = -1, cvv = -1;
  double L[5];
  // several blocks, level agglomerated for the coarse solve: 13 padded global arrays on rank 0 (examg_optflow.cu, of_cg_agg)
  double *agg_pad = nullptr;
};

// startTimer / stopTimer of the L4 programs, measured on the device (CUDA events on the context stream)
struct Timer {
  cudaEvent_t start = nullptr;                           // open measurement
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> done; // closed, not yet folded into total_ms
  double total_ms = 0.0;
  long long count = 0;
  int depth = 0;                                         // nested startTimer calls (reference: entries counter)
};

struct dd { double hi, lo; };

// A cudaMalloc'd region.  Buffers of >= 1 MiB get a block of their own, smaller ones are carved
// out of shared arena blocks, so that every buffer is (block, offset) -- which is what is exported
// to the other ranks through CUDA IPC (one handle per block).
struct Block {
  char *base = nullptr;
  size_t bytes = 0, used = 0;
  bool arena = false;
};

}  // namespace examg

struct examg_ctx {
  examg_config cfg;
  int device = 0;
  int bidx[3] = {0, 0, 0};  // block index of this rank
  int world = 1;
  bool nb[3][2];  // neighbour present on the -/+ side of dim d (a periodic dimension: the block itself)
  bool per[3] = {false, false, false};  // periodic dimension (its only block is its own neighbour)
  bool periodic_any = false;
  int loop_loc = 0;  // localisation of the fields of the entry point being executed (0 node, 1 cell)
  std::vector<examg::LevelGeom> levels;  // index level - min_level
  std::vector<examg::Field> fields;
  std::vector<examg::Stencil> stencils;
  std::vector<examg::System2> systems;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  long long launches = 0;
  long long dev_bytes = 0;
  int nvtx_depth = 0;  // open benchmarkStart ranges
  // reduction scratch: per-block partials + final result
  examg::dd *partials = nullptr;  // PARTIAL_CAP
  examg::dd *red_out = nullptr;   // RED_SLOTS results on device
  examg::dd *red_host = nullptr;  // pinned mirror
  int *int_out = nullptr;         // device ints (cg iterations)
  int *int_host = nullptr;        // pinned
  struct examg_comm *comm = nullptr;
  std::vector<examg::Block> blocks;
  std::map<std::string, examg::Timer> timers;
  double *stage = nullptr;        // dense staging array of host <-> field transfers
  size_t stage_doubles = 0;
  // pipelined upload (examg_field_upload_begin / _finish): the H2D copy runs on its own stream into one of two
  // staging arrays while the main stream computes; only the device-side scatter is ordered with the kernels
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t up_h2d[2] = {nullptr, nullptr}, up_d2d[2] = {nullptr, nullptr};
  double *up_stage[2] = {nullptr, nullptr};
  size_t up_doubles[2] = {0, 0};
  int up_turn = 0, up_field = -1, up_slot = -1;
  bool up_pending = false;
};

namespace examg {

constexpr int PARTIAL_CAP = 1 << 15;
constexpr int RED_SLOTS = 8;

int set_error(int code, const char *fmt, ...);
// device memory of the context (zero-filled); see Block
double *dev_alloc(examg_ctx *c, size_t bytes);
void dev_free(examg_ctx *c, double *p);
void comm_destroy(examg_ctx *c);
const LevelGeom *geom(examg_ctx *c, int level);
Field *get_field(examg_ctx *c, int h);
Stencil *get_stencil(examg_ctx *c, int h);
inline double *field_ptr(examg_ctx *c, const Field *f, int slot, int comp = 0) {
  const LevelGeom &g = c->levels[f->level - c->cfg.min_level];
  if (slot < 0) slot = f->cur;
  return f->slot[slot] + (long long)comp * g.vol + g.origin;
}
GridArgs grid_args(const LevelGeom &g);
// loop ranges of the localisation the current entry point works on (examg_ctx::loop_loc): node fields as grid_args,
// cell fields all cells 0 .. n-2 of every dimension (`loop over F` of a cell field, reductions alike)
GridArgs grid_args_loc(examg_ctx *c, const LevelGeom &g);
// sets examg_ctx::loop_loc for the duration of one entry point
struct LocGuard {
  examg_ctx *c;
  LocGuard(examg_ctx *ctx, int loc);
  ~LocGuard();
};
StencilArgs stencil_args(const LevelGeom &g, const Stencil &s);

// global (all-rank) sum of a device dd partial written to c->red_out[slot]; returns hi+lo
int finish_sum(examg_ctx *c, int slot, double *out);
int finish_sums(examg_ctx *c, int first, int n, double *out);  // n sums, one host synchronisation
// the same for a maximum (red_out[slot].hi)
int finish_max(examg_ctx *c, int slot, double *out);
// EXAMG_ESTATE if a device-side wait for a peer rank timed out since the peers were wired (sticky), else 0
int comm_poll_error(examg_ctx *c);

#define EXAMG_CUDA(expr)                                                              \
  do {                                                                                \
    cudaError_t e__ = (expr);                                                         \
    if (e__ != cudaSuccess)                                                           \
      return examg::set_error(EXAMG_ECUDA, "%s:%d %s: %s", __FILE__, __LINE__, #expr, \
                              cudaGetErrorString(e__));                               \
  } while (0)

// cudaFuncSetAttribute is per device: `static unsigned long long done` bit = device ordinal (< 64)
#define EXAMG_ONCE_PER_DEVICE(c, body)                                 \
  do {                                                                 \
    static unsigned long long done__ = 0ULL;                           \
    const unsigned long long bit__ = 1ULL << ((c)->device & 63);       \
    if (!(done__ & bit__)) {                                           \
      body;                                                            \
      done__ |= bit__;                                                 \
    }                                                                  \
  } while (0)

// host synchronisation of the context stream; reports a peer wait that timed out (examg_comm.cu) as EXAMG_ESTATE
#define EXAMG_SYNC(c)                                                  \
  do {                                                                 \
    EXAMG_CUDA(cudaStreamSynchronize((c)->stream));                    \
    if (int rc__ = examg::comm_poll_error(c)) return rc__;             \
  } while (0)

#define EXAMG_LAUNCH_CHECK(c)                                                        \
  do {                                                                               \
    (c)->launches++;                                                                 \
    cudaError_t e__ = cudaGetLastError();                                            \
    if (e__ != cudaSuccess)                                                          \
      return examg::set_error(EXAMG_ECUDA, "%s:%d launch: %s", __FILE__, __LINE__,   \
                              cudaGetErrorString(e__));                              \
  } while (0)

// ---- internal launchers (defined in the kernel translation units) -----------
int launch_apply(examg_ctx *c, const Stencil &s, const double *u, double *out);
int launch_residual(examg_ctx *c, const Stencil &s, const double *u, const double *rhs, double *res,
                    int red_slot /* <0: no norm */, int extend = 0);
// Residual = RHS - A*Solution incl. `communicate Solution`; extend > 0: also on that many layers of the
// neighbour blocks' nodes (then Residual needs no exchange before the restriction)
int residual_op(examg_ctx *c, int A, int u, int u_slot, int rhs, int res, double *norm2_out, int extend);
// global sum of (RHS - A*Solution)^2 without writing Residual: same loop, same sums, 16 instead of 24 bytes per point
int residual_norm2_op(examg_ctx *c, int A, int u, int rhs, double *norm2_out);
// fused residual + restriction (examg_resrestrict.cu); 1 = not covered, run the two loops
int residual_restrict_op(examg_ctx *c, int A, int u, int rhs, int rhs_coarse, double scale);
int launch_jacobi(examg_ctx *c, const Stencil &s, double omega, const double *u, const double *rhs,
                  double *unext);
int launch_rb(examg_ctx *c, const Stencil &s, double omega, double *u, const double *rhs, int color);
// brings the split-row coefficient copy of a stencil-field operator up to date (allocates on first use: call it
// outside graph capture; launch_rb falls back to the strided reads while there is no valid copy)
int rb_compact_prepare(examg_ctx *c, Stencil &s);
int launch_color_generic(examg_ctx *c, const Stencil &s, double omega, double *u, const double *rhs, int ncol, int cx,
                         int cy, int cz);
int launch_restrict(examg_ctx *c, int lvl_fine, double scale, const double *rf, double *rhsc);
int launch_prolong(examg_ctx *c, int lvl_fine, const double *uc, double *uf);
int launch_dot(examg_ctx *c, int level, const double *a, const double *b, int red_slot);
int launch_maxdiff(examg_ctx *c, int level, const double *a, const double *b, int red_slot);
enum { STREAM_SET = 0, STREAM_COPY = 1, STREAM_AXPY = 2, STREAM_XPAY = 3 };
int launch_stream_op(examg_ctx *c, int level, int op, double alpha, const double *x, double *y);
int launch_bc(examg_ctx *c, const Field &f, int slot);
int launch_field_eval(examg_ctx *c, const Field &f, int slot, int comp, const examg_bc_op *prog, int n);
int launch_loop_eval(examg_ctx *c, int level, const double *const *src, int nsrc, double *dst, const examg_bc_op *prog,
                     int n);
int launch_bc_cell(examg_ctx *c, const Field &f, int slot);  // 2D Neumann ghost cells incl. corners (examg_optflow.cu)
int launch_bc_cell_generic(examg_ctx *c, const Field &f, int slot);  // Dirichlet / Neumann faces, 2D and 3D
int launch_cg(examg_ctx *c, const Stencil &s, double *u, const double *rhs, double *res, double *p,
              double *ap, int max_it, double rel_tol);
int launch_cr(examg_ctx *c, const Stencil &s, double *u, const double *rhs, double *res, double *const tmp[3], int max_it,
              double rel_tol);
int launch_minres(examg_ctx *c, const Stencil &s, double *u, const double *rhs, double *res, double *const tmp[7],
                  int max_it, double rel_tol);
int launch_smoother_solve(examg_ctx *c, const Stencil &s, double *u, const double *rhs, double *res, double *other, int kind,
                          double omega, int max_it, double rel_tol);
int launch_bicgstab(examg_ctx *c, const Stencil &s, double *u, const double *rhs, double *res, double *const tmp[5],
                    int max_it, double rel_tol);
// communicate: `depth` ghost layers (+ the duplicate layer) per side, 1 <= depth <= GHOST
int do_exchange(examg_ctx *c, Field &f, int slot, int depth = 1);
// several fields in one handshake per dimension
int do_exchange_many(examg_ctx *c, Field *const *fs, const int *slots, const int *depths, int nfields);
// agglomeration of the coarse levels onto rank 0 (examg_comm.cu)
int copy_field_device(examg_ctx *c, int field, int slot, double *dense, bool to_field);
int comm_agg_setup(examg_ctx *c, int level);
double *comm_agg_in(examg_ctx *c);
double *comm_agg_out(examg_ctx *c);
int comm_agg_gather(examg_ctx *c, Field &f, int comp = 0);
int comm_agg_scatter(examg_ctx *c, Field &f, int comp = 0);
// the same staging for cell-localised fields, several dense arrays in and out (coarse solve of a coupled system)
int comm_agg_setup_cells(examg_ctx *c, int level, int nin, int nout, int extra);
int comm_agg_cells(examg_ctx *c, int gn[3], long long *tot);
const double *comm_agg_out_peer(examg_ctx *c);

}  // namespace examg
