/*
 * examg.h -- C ABI of libexamg, the B200-native runtime behind the geometric-
 * multigrid V-cycle that the ExaStencils generator emits.
 *
 * The reference has no FFI for this path: its generator prints every loop nest
 * inline.  The conventions this boundary mirrors (paths relative to the
 * reference root, C/s/e = Compiler/src/exastencils):
 *   - kernel wrappers are `extern "C"` functions taking plain pointers, ints
 *     and scalars            C/s/e/parallelization/api/cuda/CUDA_Kernel.scala:546-632
 *   - generateCInterface wraps all generated functions in extern "C"
 *                            C/s/e/baseExt/ir/IR_FunctionCollection.scala:52-66
 *   - get<name>/set<name> copy loops between a caller-owned dense array and
 *     the internal field     C/s/e/interfacing/ir/IR_CopyToExternalField.scala:30-90
 *   - errors abort (IR_Assert -> exit(1); CUDA_CheckError)
 *                            C/s/e/parallelization/api/cuda/CUDA_Error.scala:39-68
 * Here every entry point returns 0 on success and a negative EXAMG_E* code on
 * failure; examg_last_error() holds the message.  The emitted host code wraps
 * calls in a check-and-exit(1) macro, which reproduces the reference behaviour
 * (INTEGRATION.md).
 *
 * Ownership: the library owns all device memory (layout, padding and ghost
 * layers are its business); callers hold integer handles.  Host pointers passed
 * to upload/download/norm outputs stay caller-owned and are never retained.
 * Threading: single caller thread per context; calls are stream-ordered on the
 * context's stream and return before completion unless they have host outputs.
 *
 * One context = one block (= one fragment) of the rectangular domain on one
 * GPU:  domain_rect_numBlocks_{x,y,z} blocks, rank -> block index x-fastest
 * (C/s/e/domain/ir/IR_InitGeneratedDomain.scala:41-47).
 */
#ifndef EXAMG_H
#define EXAMG_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EXAMG_ABI_VERSION 1
#define EXAMG_MAX_STENCIL 27

/* error codes */
#define EXAMG_OK 0
#define EXAMG_EINVAL (-1)  /* bad argument / handle                       */
#define EXAMG_ECUDA (-2)   /* CUDA runtime error (message has the string) */
#define EXAMG_ENOMEM (-3)
#define EXAMG_ESTATE (-4)  /* call not valid in this state                */
#define EXAMG_EUNSUP (-5)  /* combination not supported                   */

typedef struct examg_ctx examg_ctx;

/* ---- lifecycle: initGlobals + initDomain + setupBuffers / destroyGlobals ----
 * (C/s/e/domain/ir/IR_InitGeneratedDomain.scala:78-104,
 *  C/s/e/globals/ir/IR_AllocateDataFunction.scala:64-72) */
typedef struct {
  int dim;              /* Knowledge.dimensionality: 2 or 3                        */
  int min_level;        /* Knowledge.minLevel                                      */
  int max_level;        /* Knowledge.maxLevel                                      */
  int block_cells[3];   /* cells per block and dim at level 0 =
                           domain_rect_numFragsPerBlock_d * domain_fragmentLength_d */
  int num_blocks[3];    /* domain_rect_numBlocks_{x,y,z}                           */
  int rank;             /* this block: mpi rank, x-fastest over num_blocks         */
  int device;           /* CUDA device ordinal (cuda_deviceId)                     */
  double domain_lo[3];  /* Domain global< lo to hi >                               */
  double domain_hi[3];
  int periodic[3];      /* domain_rect_periodic_{x,y,z}: the first and the last node of the dimension are the
                           same point (one block in that dimension so far); loops cover both copies, reductions
                           count it once, `communicate` continues the field across the boundary             */
} examg_config;

int examg_abi_version(void);
const char *examg_last_error(void);
examg_ctx *examg_init(const examg_config *cfg); /* NULL on failure */
int examg_finalize(examg_ctx *ctx);
/* Use an existing cudaStream_t (passed as void*) for all work; NULL = own stream. */
int examg_set_stream(examg_ctx *ctx, void *cuda_stream);
int examg_sync(examg_ctx *ctx);
/* local node counts n[d] (duplicate layers included, ghosts excluded) of a level */
int examg_level_dims(examg_ctx *ctx, int level, int n[3]);
/* grid width vf_gridWidth_d of a level (C/s/e/domain/ir/IR_DomainFromAABB.scala:31-40) */
int examg_level_width(examg_ctx *ctx, int level, double h[3]);

/* ---- fields (Field Name< domain, Layout, BC >[slots]@level) -----------------
 * Node-localised, one duplicate layer, one ghost layer per side
 * (Layout NodeWithComm; C/s/e/field/ir/IR_FieldLayout.scala:30-129).
 * ncomp > 1 gives the SoA component dimension of stencil-field coefficients. */
int examg_field_create(examg_ctx *ctx, const char *name, int level, int num_slots, int ncomp);
/* Cell-localised field (Layout< Real, Cell >: duplicateLayers 0, ghostLayers 1): n-1 cells per
 * dimension where the level has n nodes; upload/download move the dense cell array, no ghosts.
 * 2D and 3D; several blocks: `communicate` fills the ghost cells at block faces with the neighbour's cells (nothing is
 * shared between blocks, C/s/e/communication/ir/IR_PackInfoGhost.scala:13-60), `apply bc` the ones at the physical
 * boundary; no periodic dimensions.  The loops of the entry points below (apply, residual, smooth_jacobi, dot,
 * max_abs_diff, set, copy, axpy, xpay) run over all cells when their fields are cell-localised; boundary
 * conditions write the ghost cells: Dirichlet value / program g -> 2*g - inner cell
 * (C/s/e/boundary/ir/IR_DirichletBC.scala:42-50, the program sees vf_boundaryCoord: the face position in the
 * face's dimension, cell centres in the others), Neumann -> inner cell.  Programs: Testing/CellBased. */
int examg_field_create_cell(examg_ctx *ctx, const char *name, int level, int num_slots, int ncomp);
int examg_level_cells(examg_ctx *ctx, int level, int n[3]);
int examg_field_destroy(examg_ctx *ctx, int field);
int examg_field_num_slots(examg_ctx *ctx, int field);
int examg_field_active_slot(examg_ctx *ctx, int field);          /* currentSlot              */
int examg_field_advance(examg_ctx *ctx, int field);              /* `advance F` IR_Slot.scala:33-62 */

/* Boundary conditions (third template argument of the field declaration).
 * Dirichlet on node fields writes the value into the boundary nodes (the
 * duplicate layer of blocks without a neighbour; IR_DirichletBC.scala:37-40). */
enum {
  EXAMG_BC_NONE = 0,
  EXAMG_BC_DIRICHLET_CONST = 1, /* value                                             */
  EXAMG_BC_DIRICHLET_EXPR = 2,  /* postfix program over vf_boundaryPosition_{x,y,z}  */
  EXAMG_BC_DIRICHLET_HOST = 3,  /* boundary nodes keep what upload() wrote           */
  EXAMG_BC_NEUMANN = 4          /* cell fields, order 1 (Neumann(1)): ghost cell = adjacent inner cell
                                   (C/s/e/boundary/ir/IR_NeumannBC.scala:70-80)       */
};
/* postfix (RPN) program for boundary expressions -- what the kernel printer
 * lowers `vf_boundaryPosition_x ** 2 - 0.5 * ...` to.  Evaluated in double with
 * exactly the association of the program (no contraction). */
enum {
  EXAMG_OP_CONST = 0, /* push value */
  EXAMG_OP_X = 1, EXAMG_OP_Y = 2, EXAMG_OP_Z = 3, /* push node position */
  EXAMG_OP_ADD = 4, EXAMG_OP_SUB = 5, EXAMG_OP_MUL = 6, EXAMG_OP_DIV = 7, EXAMG_OP_NEG = 8,
  /* math functions of the L4 language (C/s/e/base/ir/IR_MathFunctions.scala): the device's libm, which can
   * differ from the host's in the last bit */
  EXAMG_OP_EXP = 9, EXAMG_OP_SIN = 10, EXAMG_OP_COS = 11, EXAMG_OP_SINH = 12, EXAMG_OP_SQRT = 13,
  /* examg_field_eval only: push component `value` of the field being written, at the same node */
  EXAMG_OP_COMP = 14,
  /* examg_loop_eval only: push source field `value` (index into its src list) at the same node */
  EXAMG_OP_FIELD = 15
};
typedef struct { int op; double value; } examg_bc_op;
int examg_field_set_bc(examg_ctx *ctx, int field, int kind, double value);
int examg_field_set_bc_expr(examg_ctx *ctx, int field, const examg_bc_op *prog, int n_ops);
/* A boundary condition of its own for ONE face (face = 2*d + side) of a cell-localised field: what a user-written
 * boundary function with `loop over F only ghost [..] on boundary` per face does
 * (Benchmark/FivePointStencil/5pt_Jac_Cell.exa4 `applyBC_solution`).  kind: EXAMG_BC_NONE (leave the face alone),
 * _DIRICHLET_CONST (value), _DIRICHLET_EXPR (prog), _NEUMANN; the other faces keep the field's condition.      */
int examg_field_set_bc_face(examg_ctx *ctx, int field, int face, int kind, double value, const examg_bc_op *prog,
                            int n_ops);
/* loop over F { F[comp] = <expr> } with <expr> a postfix program over vf_nodePosition_{x,y,z}: the initialisation
 * loops of the L4 programs (InitRHS, InitLaplace of Testing/SISC/3D_VarCoeff.exa4:212-240).  Runs on every local node
 * and on the first ghost layer (node positions continue beyond the block), so stencil-field coefficients need no
 * exchange.  EXAMG_OP_COMP reads components written by earlier calls (e.g. centre = -(sum of the others)). */
int examg_field_eval(examg_ctx *ctx, int field, int slot, int comp, const examg_bc_op *prog, int n_ops);
/* The pointwise loop no pattern covers -- `loop over dst { dst<dst_slot> = expr }` with expr a program over the node
 * position and up to eight node fields of the same level read at the same node (EXAMG_OP_FIELD; slot -1 = active;
 * a source may be dst itself).  Runs on dst's loop points (Dirichlet boundary nodes excluded, duplicates included).
 * This is what the nonlinear terms of Testing/NonLinear/FAS_2D_Basic.exa4 go through (Smoother, UpResidual,
 * Restriction: gam * exp(Solution) * Solution next to a stencil convolution done by examg_apply).               */
int examg_loop_eval(examg_ctx *ctx, int dst, int dst_slot, const int *src, const int *src_slot, int nsrc,
                    const examg_bc_op *prog, int n_ops);
/* applyBCs<F>_<lvl>(slot)  (C/s/e/boundary/ir/IR_ApplyBCFunction.scala:53-105); slot < 0 = active */
int examg_apply_bc(examg_ctx *ctx, int field, int slot);

/* writeField ( "file", F ) / readField ( "file", F ) with the "lock" interface in binary mode
 * (C/s/e/field/ir/IR_WriteField.scala:26-45, IR_ReadField.scala:27-46, io/ir/IR_FileAccess_Locking.scala:87-200): the
 * values only, raw doubles [comp][z][y][x] (x fastest), ghost layers excluded.  append != 0 appends (several blocks: the
 * ranks write one after the other, as MPI_Sequential does); offset_doubles = where this block starts in the file.
 * Synchronise.                                                                                                */
int examg_field_write(examg_ctx *ctx, int field, int slot, const char *filename, int append);
int examg_field_read(examg_ctx *ctx, int field, int slot, const char *filename, long long offset_doubles);

/* set<name>/get<name>: dense caller array [comp][nz][ny][nx], x fastest, local
 * nodes incl. duplicate/boundary nodes, no ghosts.  Synchronous. */
int examg_field_upload(examg_ctx *ctx, int field, int slot, const double *host);
int examg_field_download(examg_ctx *ctx, int field, int slot, double *host);
/* same, asynchronous on the context stream (host must be pinned to overlap) */
int examg_field_upload_async(examg_ctx *ctx, int field, int slot, const double *host);
int examg_field_download_async(examg_ctx *ctx, int field, int slot, double *host);
/* Pipelined upload for solve loops that take a new input every step: _begin starts the host -> device
 * copy on the library's own copy stream (host must be pinned and stay valid until _finish returns),
 * _finish orders the scatter into the field with the kernels of the context stream.  Kernels issued
 * between the two calls overlap the PCIe transfer.  One upload in flight at a time.              */
int examg_field_upload_begin(examg_ctx *ctx, int field, int slot, const double *host);
int examg_field_upload_finish(examg_ctx *ctx);
/* whole allocation (ghosts, padding, all slots) := value (initFieldsWithZero) */
int examg_field_fill(examg_ctx *ctx, int field, double value);

/* printField ( "file", F ) of the L4 programs in its ascii form (io interface "lock", C/s/e/field/ir/IR_PrintField.scala:49-100):
 * one line per point of the block in lexicographic order (x fastest) -- the position (vf_nodePosition of node fields,
 * vf_cellCenter of cell fields) in every dimension, then the components, all in std::defaultfloat (6 significant
 * digits), joined by `separator` (" " unless given).  Ghost layers are not written.  Synchronises.            */
int examg_field_print(examg_ctx *ctx, int field, int slot, const char *filename, const char *separator);

/* ---- stencils ---------------------------------------------------------------
 * Stencil Name@lvl { [dx,dy,dz] => coeff ... }: entries in declaration order.
 * Stencil fields take their coefficients from a field with ncomp = n_entries
 * (StencilField; C/s/e/operator/ir/IR_StencilField.scala). */
int examg_stencil_const(examg_ctx *ctx, int level, int n_entries, const int (*offsets)[3],
                        const double *coeffs);
int examg_stencil_field(examg_ctx *ctx, int level, int n_entries, const int (*offsets)[3],
                        int coeff_field);

/* ---- hot path: one entry point per `loop over` shape -------------------------
 * Iteration space of every loop: all local nodes except physical-boundary nodes
 * (C/s/e/baseExt/ir/IR_LoopOverPointsInOneFragment.scala:86-101); reductions
 * count nodes shared between blocks once (:116-125).  slot < 0 = active slot. */

/* out = A * u */
int examg_apply(examg_ctx *ctx, int stencil, int u, int u_slot, int out);
/* Residual = RHS - A * Solution; if norm2_out != NULL also returns the global
 * sum of Residual^2 (fused reduction; synchronises)                         */
int examg_residual(examg_ctx *ctx, int stencil, int u, int u_slot, int rhs, int res,
                   double *norm2_out);
/* nsweeps x { communicate; Solution<next> = Solution<active> + omega/diag(A) *
 * (RHS - A*Solution<active>); advance }   (needs 2 slots)                  */
int examg_smooth_jacobi(examg_ctx *ctx, int stencil, int u, int rhs, double omega, int nsweeps);
/* nsweeps x color with {(i0+i1+i2)%2}: colour 0 then colour 1, each
 * { communicate; Solution += omega/diag(A)*(RHS - A*Solution); apply bc }
 * (C/s/e/baseExt/l4/L4_ColorLoops.scala:44-70).  The library may fuse colours
 * and sweeps; results are identical to the unfused order.                   */
int examg_smooth_rbgs(examg_ctx *ctx, int stencil, int u, int rhs, double omega, int nsweeps);
/* a single colour half-sweep (the unfused form) */
int examg_smooth_rb_color(examg_ctx *ctx, int stencil, int u, int rhs, double omega, int color);
/* nsweeps x Gauss-Seidel with an n^dim-way colouring, n = colors_per_dim = 2 or 3: solver_smoother_coloring "4-way" /
 * "9-way" (2D), "8-way" / "27-way" (3D) of the layer-3 solver generator
 * (C/s/e/solver/l3/L3_VankaForEquation.scala:154-183).  Colours run with the last dimension slowest
 * (C/s/e/baseExt/l4/L4_ColorLoops.scala:44-70), each { communicate; loop over Solution where i_d % n == c_d
 * { Solution += omega/diag(A) * (RHS - A*Solution) }; apply bc }.  For compact stencils with diagonal entries
 * (9- and 27-point), which two colours do not decouple.                                            */
int examg_smooth_colored(examg_ctx *ctx, int stencil, int u, int rhs, double omega, int colors_per_dim, int nsweeps);
/* RHS@coarser = scale * NodeRestriction * Residual
 * (C/s/e/operator/l4/L4_DefaultRestriction.scala:29-90)                     */
int examg_restrict(examg_ctx *ctx, int res_fine, int rhs_coarse, double scale);
/* UpResidual + Restriction of mgCycle in one call: RHS@coarser = scale * NodeRestriction * (RHS - A * Solution).
 * On large 3D constant-coefficient levels this is ONE pass over the fine grid and Residual is left untouched
 * (inside mgCycle it only feeds the restriction, Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4:216-224);
 * otherwise the two loops run one after the other and Residual is written.                       */
int examg_residual_restrict(examg_ctx *ctx, int stencil, int u, int rhs, int res, int rhs_coarse, double scale);
/* Solution += NodeProlongation@coarser * Solution@coarser
 * (C/s/e/operator/l4/L4_DefaultProlongation.scala:30-45)                    */
int examg_prolong_add(examg_ctx *ctx, int u_coarse, int u_coarse_slot, int u_fine, int u_fine_slot);
/* Correction + post-smoothing of mgCycle in one call (Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4:232-248):
 *   Solution += NodeProlongation@coarser * Solution@coarser; apply bc; nsweeps x { red-black sweep }.
 * On large 3D constant-coefficient levels the interpolation is folded into the load stage of the first smoothing pass
 * (one read and one write of Solution less per cycle); otherwise the two loops run one after the other.  The result
 * is bit-identical to examg_prolong_add followed by examg_smooth_rbgs.                                          */
int examg_prolong_add_smooth_rbgs(examg_ctx *ctx, int stencil, int u_coarse, int u, int rhs, double omega, int nsweeps);
/* loop over F with reduction(+: x) { x += a * b }; global over all blocks */
int examg_dot(examg_ctx *ctx, int a, int a_slot, int b, int b_slot, double *out);
/* No counterpart in the reference.  The dot-product reduction of
 * examg_dot for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc): *out_dev = sum a_dev[i] * b_dev[i], i < n, double-double accumulation as examg_dot.  All four
 * pointers are plain device addresses of the caller (dense arrays, no context, no handles); the per-block partial
 * sums live in a 65 792-byte file-scope device array that every GPU loading the library holds until exit.  n is
 * limited to EXAMG_NOALLOC_MAX_N elements (larger: EXAMG_EINVAL with a message).  Calls no cudaMalloc / cudaFree or
 * allocating routine; asynchronous on cuda_stream (a cudaStream_t, NULL = default stream).  All calls share the
 * array: calls that overlap on the GPU (unordered streams, several host threads) race -- serialise them.        */
#define EXAMG_NOALLOC_MAX_N (1LL << 28)
int examg_dot_dense_noalloc(const double *a_dev, const double *b_dev, long long n, double *out_dev, void *cuda_stream);
/* max |a - b| over the loop points (PrintError / NormError); global over all blocks, as examg_dot */
int examg_max_abs_diff(examg_ctx *ctx, int a, int a_slot, int b, int b_slot, double *out);
/* streaming loops on the loop points */
int examg_set(examg_ctx *ctx, int f, int slot, double value);
int examg_copy(examg_ctx *ctx, int src, int src_slot, int dst, int dst_slot);
int examg_axpy(examg_ctx *ctx, double alpha, int x, int x_slot, int y, int y_slot); /* y += alpha*x */
int examg_xpay(examg_ctx *ctx, int x, int x_slot, double beta, int y, int y_slot);  /* y = x + beta*y */
/* exch<F>_<lvl>(slot): duplicate-layer sync then ghost-layer exchange, one
 * dimension after the other (C/s/e/communication/ir/IR_CommunicateFunction.scala:412-470) */
int examg_exchange(examg_ctx *ctx, int field, int slot);
/* same with `depth` ghost layers per side (1..4): what a sweep that fuses depth/2 red-black sweeps consumes
 * (the contraction extents of C/s/e/baseExt/ir/IR_ContractingLoop.scala:48-90) */
int examg_exchange_depth(examg_ctx *ctx, int field, int slot, int depth);
/* mgCycle@coarsest: un-preconditioned CG, stops at ||r|| <= rel_tol*||r0|| or
 * max_it (Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4:152-201).
 * Runs entirely on the device; *iters_out (nullable) synchronises.          */
int examg_cg_coarse(examg_ctx *ctx, int stencil, int u, int rhs, int res, int p, int ap,
                    int max_it, double rel_tol, int *iters_out);

/* The same with BiCGStab (solver_cgs = "BiCGStab", C/s/e/solver/l3/L3_BiCGStabForEquation.scala:36-198) for operators
 * that are not symmetric positive definite; tmp = { v, p, s, t, resHat } (fields of the level).               */
int examg_bicgstab_coarse(examg_ctx *ctx, int stencil, int u, int rhs, int res, const int tmp[5], int max_it,
                          double rel_tol, int *iters_out);
/* Conjugate residuals (solver_cgs = "ConjugateResidual", C/s/e/solver/l3/L3_ConjugateResidualForEquation.scala:36-170) for
 * symmetric, possibly indefinite operators; tmp = { p, ap, ar }.                                              */
int examg_cr_coarse(examg_ctx *ctx, int stencil, int u, int rhs, int res, const int tmp[3], int max_it, double rel_tol,
                    int *iters_out);
/* MINRES (solver_cgs = "MinRes", C/s/e/solver/l3/L3_MinResForEquation.scala:36-230); tmp = seven work fields (three
 * Lanczos vectors, three search directions, A v).  Residual is the initial residual; the norm is tracked by the
 * recurrence, as in the generated code.                                                                      */
int examg_minres_coarse(examg_ctx *ctx, int stencil, int u, int rhs, int res, const int tmp[7], int max_it,
                        double rel_tol, int *iters_out);

/* The smoother as coarse-grid solver (solver_cgs = "Smoother", C/s/e/solver/l3/L3_VankaAsSolverForEquation.scala:31-94):
 * residual and norm, then up to max_it times { ten smoother steps; residual; norm; stop at rel_tol * initial norm }.
 * smoother = EXAMG_SMOOTH_JACOBI (second buffer: Solution<next> of a two-slot field, else the work field tmp) or
 * EXAMG_SMOOTH_RBGS (tmp ignored).  One block, no periodic dimensions.  *iters_out = outer steps (nullable).   */
int examg_smoother_coarse(examg_ctx *ctx, int stencil, int u, int rhs, int res, int smoother, double omega, int tmp,
                          int max_it, double rel_tol, int *iters_out);

/* ---- whole cycle: Solve@finest / mgCycle@lvl as one library object -----------
 * (C/s/e/solver/l3/L3_SolverForEquation.scala:259-395).  Creates the field set
 * of the L4 programs on every level (Solution[slots], RHS, Residual and, on the
 * coarsest level, cgTmp0/cgTmp1) and runs the level loop without host
 * round-trips (captured into a CUDA graph).                                  */
enum { EXAMG_SMOOTH_JACOBI = 0, EXAMG_SMOOTH_RBGS = 1,
       /* Gauss-Seidel with an n^dim-way colouring, n = 2 / 3 (examg_smooth_colored): solver_smoother_coloring
        * "4-way" / "8-way" and "9-way" / "27-way"                                                          */
       EXAMG_SMOOTH_COLORED_2 = 2, EXAMG_SMOOTH_COLORED_3 = 3 };
typedef struct {
  int smoother;          /* EXAMG_SMOOTH_*                                       */
  double omega;          /* solver_smoother_damping                              */
  int nu_pre, nu_post;   /* solver_smoother_numPre / numPost                     */
  double restrict_scale; /* 1, or 4 for `4.0 * RestrictionStencil` programs      */
  int cg_max_it;         /* solver_cgs_maxNumIts                                 */
  double cg_rel_tol;     /* solver_cgs_targetResReduction                        */
  const int *stencils;   /* stencil handle per level, index level - min_level   */
  int coarse_solver;     /* solver_cgs: EXAMG_CGS_CG (0, the default), _BICGSTAB, _CR, _MINRES, _SMOOTHER   */
  int use_fas;           /* solver_useFAS: full approximation scheme -- per level Approx@coarser = Restriction * Solution;
                          * Solution@coarser = Approx@coarser; RHS@coarser = Restriction * Residual + A@coarser *
                          * Solution@coarser; mgCycle@coarser; Solution@coarser -= Approx@coarser; correction
                          * (C/s/e/solver/l3/L3_SolverForEquation.scala:401-452).  One block.  For the linear operators
                          * a stencil handle describes this equals the correction scheme up to rounding; nonlinear
                          * terms are host-driven: examg_loop_eval, tests/fas_program.py                              */
} examg_cycle_desc;
enum { EXAMG_CGS_CG = 0, EXAMG_CGS_BICGSTAB = 1, EXAMG_CGS_CR = 2, EXAMG_CGS_MINRES = 3, EXAMG_CGS_SMOOTHER = 4 };
/* `generate solver for ... with { key = value ... }` / the solver_* keys of a Knowledge file -> cycle descriptor
 * (C/s/e/solver/l3/L3_SolverForEquation.scala:84-94 sets Knowledge fields from the option list; defaults
 * C/s/e/config/Knowledge.scala:203-262).  options: "key = value" pairs separated by newlines, ';' or ','; values may be
 * quoted.  Understood: solver_smoother_numPre / numPost / damping / jacobiType / coloring, solver_cgs,
 * solver_cgs_maxNumIts, solver_cgs_targetResReduction.  desc->stencils and restrict_scale are left alone.
 * *colors_per_dim (nullable) = 0 for Jacobi or red-black, 2 / 3 for the n^d-way colourings (desc->smoother is then
 * EXAMG_SMOOTH_COLORED_2 / _3).  Keys of the outer Solve loop (solver_targetResReduction, solver_maxNumIts, ...) are
 * solver_useFAS.  Keys of the outer loop are
 * accepted and ignored: that loop is host code.  Unknown solver_* keys, coloring "None" with Gauss-Seidel and
 * unsupported values are errors (EXAMG_EUNSUP / EXAMG_EINVAL with the offending key in the message).            */
int examg_cycle_desc_from_options(const char *options, int dim, examg_cycle_desc *desc, int *colors_per_dim);
typedef struct examg_solver examg_solver;
examg_solver *examg_solver_create(examg_ctx *ctx, const examg_cycle_desc *desc);
int examg_solver_destroy(examg_solver *s);
/* field handles of a level: which = 0 Solution, 1 RHS, 2 Residual, 3 cgTmp0, 4 cgTmp1, 5 Approximation */
enum { EXAMG_F_SOLUTION = 0, EXAMG_F_RHS = 1, EXAMG_F_RESIDUAL = 2, EXAMG_F_CGTMP0 = 3, EXAMG_F_CGTMP1 = 4,
       EXAMG_F_APPROX = 5 /* gen_approx (use_fas, all levels but the finest) */ };
int examg_solver_field(examg_solver *s, int level, int which);
/* mgCycle@level (level < 0: finest); asynchronous */
int examg_solver_vcycle(examg_solver *s, int level);
/* n x mgCycle@finest followed by the residual norm of the loop body of Solve@finest:
 * *norm_out = ||RHS - A*Solution||_2 (nullable).  The field Residual@finest is NOT written by this call -- nothing reads
 * it before the next cycle -- which makes the pass 16 instead of 24 bytes per point; examg_solver_residual_norm writes it. */
int examg_solver_cycles(examg_solver *s, int n, double *norm_out);
/* Residual@finest = RHS - A*Solution; *norm_out = sqrt(global sum r^2) */
int examg_solver_residual_norm(examg_solver *s, double *norm_out);
/* CG iterations of the most recent coarse solve (synchronises) */
int examg_solver_last_cg_iters(examg_solver *s, int *iters_out);

/* ---- coupled two-unknown systems on cell fields: the optical-flow programs -----------------
 * (Benchmark/OptFlow2D/2D_FD_OptFlow.exa4, Testing/Application/OpticalFlow2D.exa4)
 *   A = [ c_uu + scale*L   c_uv          ]   L = constant 5-point stencil (declared centre,-x,+x,-y,+y),
 *       [ c_uv             c_vv + scale*L ]   c_* = cell fields (uuSten, uvSten = vuSten, vvSten)
 * All loops run over all cells; Neumann ghost cells of the written fields are refreshed by the same
 * kernel (`apply bc` costs no launch).
 * Several blocks (contexts wired with examg_comm_export / _import): the entry points below `communicate`
 * what they read (the smoother: u, v, the right-hand sides and the coefficients, as many layers as the sweeps
 * of one pass need, in one handshake); results do not depend on the block grid.                  */
typedef struct {
  int laplace;          /* stencil handle; fixes the level */
  double scale;         /* alpha ** 2 */
  int c_uu, c_uv, c_vv; /* cell-field handles */
} examg_system2_desc;
int examg_system2_create(examg_ctx *ctx, const examg_system2_desc *desc);
/* residual_u = rhs_u - A_u(u,v); residual_v alike; apply bc; *norm2_out (nullable) = sum of both squares */
int examg_system2_residual(examg_ctx *ctx, int sys, int u, int v, int rhs_u, int rhs_v, int res_u, int res_v,
                           double *norm2_out);
/* nsweeps x { colour 0, colour 1: loop over u { solve locally relax r { u => A_u == rhs_u; v => A_v == rhs_v } };
 * apply bc }  (C/s/e/solver/ir/IR_LocalSolve.scala:88-236, IR_LocalDirectInvert.scala:44-140)   */
int examg_system2_smooth_rb(examg_ctx *ctx, int sys, int u, int v, int rhs_u, int rhs_v, double relax, int nsweeps);
int examg_system2_smooth_rb_color(examg_ctx *ctx, int sys, int u, int v, int rhs_u, int rhs_v, double relax,
                                  int color);
/* 2D and 3D.  coarse[k] = RestrictionCell * fine[k] (mean of the 2^d children), k < nfields <= 3
 * (C/s/e/operator/l4/L4_DefaultRestriction.scala:40-47)                                         */
int examg_restrict_cell(examg_ctx *ctx, const int *fine, const int *coarse, int nfields);
/* fine[k] += ProlongationCell * coarse[k] (piecewise constant); apply bc
 * (C/s/e/operator/l4/L4_DefaultProlongation.scala:30-45)                                        */
int examg_prolong_add_cell(examg_ctx *ctx, const int *coarse, const int *fine, int nfields);
/* mgCycle@coarsest of the optical-flow programs: CG on both unknowns.  fields = { u, v, rhs_u, rhs_v,
 * residual_u, residual_v, cgTmp0_u, cgTmp0_v, cgTmp1_u, cgTmp1_v }; runs on the device.  Several blocks: on the
 * level examg_optflow_create set up for it, rank 0 solves on the global grid (u, v come back to every block, the
 * other fields stay as they are); on any other level the loops run on the blocks with global sums, which
 * synchronises with the host at every reduction point.                                            */
int examg_system2_cg_coarse(examg_ctx *ctx, int sys, const int *fields, int max_it, double rel_tol, int *iters_out);

/* whole program: InitFields / mgCycle@all / doSolve@finest as one library object.  Several blocks: create it on
 * every rank BEFORE examg_comm_export (it allocates the staging of the coarsest level on rank 0; one such object per
 * context); blocks must start on even cells for the fused sweeps (otherwise the per-colour kernels run). */
typedef struct {
  double alpha;          /* Globals { Val alpha }              */
  double relax;          /* solve locally relax ...            */
  int nu_pre, nu_post;   /* repeat 3 times                     */
  int cg_max_it;         /* 128                                */
  double cg_rel_tol;     /* 0.001                              */
} examg_optflow_desc;
typedef struct examg_optflow examg_optflow;
examg_optflow *examg_optflow_create(examg_ctx *ctx, const examg_optflow_desc *desc);
int examg_optflow_destroy(examg_optflow *o);
enum { EXAMG_OF_U = 0, EXAMG_OF_V = 1, EXAMG_OF_RHS_U = 2, EXAMG_OF_RHS_V = 3, EXAMG_OF_RES_U = 4, EXAMG_OF_RES_V = 5,
       EXAMG_OF_IXIX = 6, EXAMG_OF_IXIY = 7, EXAMG_OF_IYIY = 8, EXAMG_OF_IMG0 = 9, EXAMG_OF_IMG1 = 10 };
int examg_optflow_field(examg_optflow *o, int level, int which);
/* after img0/img1 were uploaded: apply bc to the images, initGradients@all, rhs@finest */
int examg_optflow_init_fields(examg_optflow *o);
int examg_optflow_vcycle(examg_optflow *o, int level);              /* mgCycle@level (< 0: finest) */
int examg_optflow_cycles(examg_optflow *o, int n, double *norm_out); /* n cycles + residual norm    */
int examg_optflow_residual_norm(examg_optflow *o, double *norm_out);
int examg_optflow_last_cg_iters(examg_optflow *o, int *iters_out);

/* ---- staggered grids: face-centred velocities, cell-centred pressure (Stokes, 2D and 3D) -----------
 * (Testing/Application/ExaStokes_2D.exa4 / ExaStokes_3D.exa4, Examples/Stokes/{2D,3D}_FD_Stokes_fromL4.exa4)
 *   Field u < Face_x >, v < Face_y > [, w < Face_z >], p < Cell >; rhs_* and residual_* alike.
 *   Smoother: color with { i0 % 3, i1 % 3 [, i2 % 3] } loop over p { solve locally { the two faces of the cell in every
 *   dimension, p@[0,0] } } -- a 5 x 5 (7 x 7) local system per cell over three (four) localisations
 *   (C/s/e/solver/ir/IR_LocalSolve.scala:88-236, IR_LocalDirectInvert.scala:44-150; wall faces are identity rows,
 *   IR_IsOnBoundary.scala:53-97) -- then apply bc.
 * One block; dimensionality, level range and block_cells (the same in every dimension) of the context.  Every field of a
 * level is an (n+3)^dim array, entry (i, j, k) at (i+1) + (n+3)(j+1) + (n+3)^2 (k+1), one ghost layer around: velocity
 * component d on the faces 0..n in dimension d and the cells 0..n-1 in the others, p on the cells.              */
typedef struct {
  int nu_pre, nu_post;   /* repeat 3 times { Smoother }                                   */
  int nu_coarse;         /* VCycle@coarsest: repeat 10 times { Smoother }                 */
  double top_vel;        /* ApplyBC_u@finest: lid velocity (0.02 in 2D, 0.2 in 3D)        */
} examg_stokes_desc;
typedef struct examg_stokes examg_stokes;
/* field ids: EXAMG_ST_VEL + d, EXAMG_ST_RHS + d, EXAMG_ST_RES + d (d = 0 u, 1 v, 2 w), then p, rhs_p, residual_p */
enum { EXAMG_ST_VEL = 0, EXAMG_ST_RHS = 3, EXAMG_ST_RES = 6, EXAMG_ST_P = 9, EXAMG_ST_RHS_P = 10, EXAMG_ST_RES_P = 11,
       EXAMG_ST_NFIELDS = 12 };
examg_stokes *examg_stokes_create(examg_ctx *ctx, const examg_stokes_desc *desc);
int examg_stokes_destroy(examg_stokes *s);
int examg_stokes_cells(examg_stokes *s, int level);                     /* n (level < 0: finest)          */
int examg_stokes_upload(examg_stokes *s, int level, int which, const double *host);   /* (n+3)^dim doubles */
int examg_stokes_download(examg_stokes *s, int level, int which, double *host);
/* apply bc to the velocities (p: None): u = ApplyBC_u on the finest level (lid: ghost rows in y = 2 top - inner above,
 * - inner below, then 0 on the wall faces in x), Dirichlet 0.0 otherwise -- the value on the wall faces of a face field's
 * own dimension, ghost = 2 g - inner in the others, faces in the order -x,+x,-y,+y,-z,+z
 * (C/s/e/boundary/ir/IR_BoundaryCondition.scala:38-45, IR_ApplyBCFunction.scala:55-88)                         */
int examg_stokes_apply_bc(examg_stokes *s, int level);
int examg_stokes_smooth_color(examg_stokes *s, int level, int c0, int c1, int c2);  /* one of the 3^dim colours (2D: c2 = 0) */
int examg_stokes_smooth(examg_stokes *s, int level, int nsweeps);  /* nsweeps x { all colours, i0 fastest; apply bc } */
/* UpdateResidual_*; norms_out (nullable, dim + 1 doubles) = NormResidual_{u,v[,w],p}: sqrt of the sums of squares
 * over the loop points (synchronises) */
int examg_stokes_residual(examg_stokes *s, int level, double *norms_out);
int examg_stokes_restrict(examg_stokes *s, int level);      /* rhs_*@coarser = Restriction<loc> * residual_*           */
int examg_stokes_set_solution(examg_stokes *s, int level, double value);  /* loop points and one layer around */
int examg_stokes_correct(examg_stokes *s, int level);       /* unknowns += Correction<loc> * unknowns@coarser; apply bc */
int examg_stokes_vcycle(examg_stokes *s, int level);        /* VCycle@level (< 0: finest; the finest as one CUDA graph) */

/* ---- the 3D optical-flow program: a Vec3 unknown on cell centres, 3 x 3 block inverse per cell ------------
 * (Testing/Application/OpticalFlow3D.exa4; golden OpticalFlow3D.results)
 *   combinedOp = alpha**2 * Laplace (x) I_3 + gradContribution,  gradContribution = the symmetric block of the
 *   cell fields IxIx, IxIy, IxIz, IyIy, IyIz, IzIz at the centre entry;
 *   smoother: repeat with { colour 0, colour 1 } { flow += inverse ( diag ( combinedOp ) ) * ( rhs - combinedOp * flow );
 *   apply bc } -- diag() = the centre block, inverted per cell with the cofactor formulas of the generator's
 *   compile-time branch (C/s/e/baseExt/ir/IR_MatOperations/IR_CompiletimeMatOps.scala:804-826).
 * One block, 3D; level range and block_cells of the context.  Every array of a level is dense (nz+2) x (ny+2) x (nx+2)
 * doubles, x fastest, one ghost cell per side (Neumann ghost cells are kept up to date by every kernel that writes).
 * The components of the Vec3 fields are separate arrays.                                                       */
typedef struct {
  double alpha;          /* Globals { Val alpha }                                          */
  int nu_pre, nu_post;   /* repeat 3 times                                                 */
  int cg_max_it;         /* 128                                                            */
  double cg_rel_tol;     /* 0.001                                                          */
} examg_optflow3_desc;
typedef struct examg_optflow3 examg_optflow3;
/* array ids of a level: EXAMG_OF3_FLOW + c, EXAMG_OF3_RHS + c, EXAMG_OF3_RES + c (c = 0..2), EXAMG_OF3_G + 0..5 =
 * xx, xy, xz, yy, yz, zz; the images live on the finest level */
enum { EXAMG_OF3_FLOW = 0, EXAMG_OF3_RHS = 3, EXAMG_OF3_RES = 6, EXAMG_OF3_G = 9, EXAMG_OF3_IMG0 = 15, EXAMG_OF3_IMG1 = 16 };
examg_optflow3 *examg_optflow3_create(examg_ctx *ctx, const examg_optflow3_desc *desc);
int examg_optflow3_destroy(examg_optflow3 *o);
int examg_optflow3_cells(examg_optflow3 *o, int level, int n[3]);       /* level < 0: finest */
int examg_optflow3_upload(examg_optflow3 *o, int level, int which, const double *host);
int examg_optflow3_download(examg_optflow3 *o, int level, int which, double *host);
/* apply bc to the images; initGradients@all; rhs@finest (`InitFields` after the images are uploaded) */
int examg_optflow3_init_fields(examg_optflow3 *o);
int examg_optflow3_smooth_color(examg_optflow3 *o, int level, int color);
int examg_optflow3_smooth(examg_optflow3 *o, int level, int nsweeps);    /* nsweeps x { colour 0, colour 1 } */
/* residual = rhs - combinedOp * flow; apply bc; *norm_out (nullable) = resNorm: sqrt of the sum of dot ( residual, residual ) */
int examg_optflow3_residual(examg_optflow3 *o, int level, double *norm_out);
int examg_optflow3_restrict(examg_optflow3 *o, int level);   /* rhs@coarser = RestrictionCell * residual; flow@coarser = 0 */
int examg_optflow3_correct(examg_optflow3 *o, int level);    /* flow += ProlongationCell@coarser * flow@coarser; apply bc */
int examg_optflow3_cg(examg_optflow3 *o, int *iters_out);    /* mgCycle@coarsest: CG on the Vec3 unknown, one kernel */
int examg_optflow3_last_cg_iters(examg_optflow3 *o, int *iters_out);
int examg_optflow3_vcycle(examg_optflow3 *o, int level);     /* mgCycle@level (< 0: finest; the finest as one CUDA graph) */
int examg_optflow3_cycles(examg_optflow3 *o, int n, double *norm_out);  /* n cycles, then the residual norm (nullable) */

/* ---- multi-GPU plumbing (one process per GPU) ---------------------------------
 * Peers are wired by exchanging opaque blobs through the host's process group
 * (torch.distributed / MPI): every rank exports a blob, all blobs are gathered
 * and imported.  After that `communicate` writes faces directly into the
 * neighbours' ghost layers over NVLink (face, edge and corner neighbours, one handshake per call).
 * Ranks wait for each other inside kernels: they must reach matching calls within EXAMG_PEER_TIMEOUT_S
 * seconds (environment, default 60; 0 = wait for ever) of one another.  A wait that times out abandons
 * the operation (nothing is stored or summed), and every later synchronising call on the context
 * (examg_sync, downloads, reductions, solver cycles) returns EXAMG_ESTATE.                            */
int examg_comm_blob_size(examg_ctx *ctx);
int examg_comm_export(examg_ctx *ctx, void *blob);
int examg_comm_import(examg_ctx *ctx, const void *all_blobs, int world);

/* ---- observability ------------------------------------------------------------ */
/* startTimer("name") / stopTimer("name") / getTotalTime("name") / printAllTimers() of the L4 programs
 * (C/s/e/timing/ir/IR_GetTime.scala:33-63, IR_PrintAllTimers.scala:38-70).  Measured with CUDA events on the
 * context stream: asynchronous; examg_timer_total_ms waits for the measurements it reports.  Do not start or
 * stop a timer inside a region that the library captures into a CUDA graph (examg_solver_vcycle).        */
/* benchmarkStart ( "name" ) / benchmarkStop ( "name" ): the reference lowers them to LIKWID / TALP region markers
 * (C/s/e/util/ir/IR_ResolveBenchmarkFunctions.scala:63-120); here an NVTX range (visible to nsys / ncu --nvtx) plus a
 * device timer of the same name (examg_timer_total_ms).  Ranges nest.                                          */
int examg_benchmark_start(examg_ctx *ctx, const char *name);
int examg_benchmark_stop(examg_ctx *ctx, const char *name);
int examg_timer_start(examg_ctx *ctx, const char *name);
int examg_timer_stop(examg_ctx *ctx, const char *name);
int examg_timer_total_ms(examg_ctx *ctx, const char *name, double *total_ms, long long *count);
int examg_timer_print_all(examg_ctx *ctx);
/* kernels launched by this context since init (bench.py's gpu_launches) */
long long examg_launch_count(examg_ctx *ctx);
/* bytes of device memory held by this context */
long long examg_device_bytes(examg_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* EXAMG_H */
