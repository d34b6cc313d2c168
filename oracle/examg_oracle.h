/*
 * examg_oracle.h -- CPU restatement of the geometric-multigrid V-cycle that the
 * ExaStencils generator emits (TEST INFRASTRUCTURE ONLY).
 *
 * This is the parity oracle for libexamg.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may use it; the product
 * path (exastencils_b200/) never links or calls anything in oracle/.
 *
 * The reference (lssfau/ExaStencils) ships no native code: the V-cycle exists
 * only as text printed by its Scala generator, and no JVM exists on this box.
 * The oracle therefore restates the ExaSlang-4 programs and the generator rules
 * cited per function below (paths relative to the reference root; C/s/e =
 * Compiler/src/exastencils).  Parity is PINNED against the reference's own
 * golden vectors (Testing/ ** / *.results, 4 significant digits) by
 * tests/test_oracle_golden.py.
 *
 * Global-grid formulation: a run on B blocks x F fragments is arithmetically a
 * run on one global node grid with N_d = F_d * 2^level cells per dimension
 * (C/s/e/domain/ir/IR_DomainFromAABB.scala:31-40); duplicate (shared) nodes
 * are computed redundantly from identical inputs by both owners
 * (C/s/e/baseExt/ir/IR_LoopOverPointsInOneFragment.scala:86-101), so one array
 * per field and level reproduces any decomposition.
 *
 * Arrays are dense, x fastest: index = i + n[0]*(j + n[1]*k), n[d] = N_d + 1
 * nodes (n[2] = 1 in 2D).  Node 0 and node n[d]-1 are physical boundary nodes;
 * every `loop over` runs on 1..n[d]-2 (iteration offsets, IR_IterationOffset).
 *
 * Canonical floating-point association (shared with the CUDA kernels so that
 * field results are bit-identical; see DESIGN.md "Arithmetic contract"):
 *   A*u      : acc = c0*u0; acc = fma(ck, uk, acc) for k = 1..nst-1 in stencil
 *              declaration order
 *   residual : r = rhs - acc
 *   smoother : u' = fma(omega/diag, rhs - acc, u)
 *   sums     : double-double (TwoSum) accumulation of the rounded products,
 *              final value = hi + lo  (order-insensitive to ~1e-23 relative)
 */
#ifndef EXAMG_ORACLE_H
#define EXAMG_ORACLE_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OG_MAX_ST 27

/* A constant-coefficient stencil (Stencil Laplace@lvl {...}) or the offset
 * pattern of a stencil field (StencilField Laplace< Coeff => Stencil >). */
typedef struct {
  int nst;
  int off[OG_MAX_ST][3];
  double coef[OG_MAX_ST]; /* ignored when a coefficient field is passed */
} og_stencil;

/* ---- per-loop kernels (one per `loop over` of the L4 programs) ---------- */

/* out = A*u on interior nodes.  sf == NULL: constant coefficients st->coef;
 * else sf[k*size + p] is entry k at node p (ColumnVector<Real,N> SoA layout,
 * C/s/e/field/ir/IR_FieldLayout.scala:30-129). */
void og_apply(const int n[3], const og_stencil *st, const double *sf,
              const double *u, double *out);

/* Residual = RHS - A*Solution  (Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4:123-125) */
void og_residual(const int n[3], const og_stencil *st, const double *sf,
                 const double *u, const double *rhs, double *res);

/* Solution<next> = Solution<active> + omega/diag(A)*(RHS - A*Solution<active>)
 * (Testing/SISC/3D_VarCoeff.exa4 Smoother) */
void og_smooth_jacobi(const int n[3], const og_stencil *st, const double *sf,
                      double omega, const double *u, const double *rhs,
                      double *unext);

/* One colour half-sweep of `color with {(i0+i1+i2)%2, ...}`
 * (Benchmark/Poisson3D/...exa4:204-213; C/s/e/baseExt/l4/L4_ColorLoops.scala:44-70) */
void og_smooth_rb(const int n[3], const og_stencil *st, const double *sf,
                  double omega, double *u, const double *rhs, int color);

/* loop over F with reduction(+:x) { x += a*b }: interior nodes, double-double.
 * Returns hi+lo.  (…exa4:113-119; IR_LoopOverPointsInOneFragment.scala:116-125) */
double og_dot(const int n[3], const double *a, const double *b);

/* max |a - b| over interior nodes (PrintError, Examples/Poisson/2D_FD_Poisson_fromL4.exa4) */
double og_max_abs_diff(const int n[3], const double *a, const double *b);
void og_smooth_color(const int n[3], const og_stencil *st, const double *sf, double omega, double *u,
                     const double *rhs, int ncol, int cx, int cy, int cz);
int og_cr(const int n[3], const og_stencil *st, const double *sf, double *u, const double *rhs, double *res, int max_it,
          double rel_tol);
int og_minres(const int n[3], const og_stencil *st, const double *sf, double *u, const double *rhs, double *res,
              int max_it, double rel_tol);
int og_bicgstab(const int n[3], const og_stencil *st, const double *sf, double *u, const double *rhs, double *res,
                int max_it, double rel_tol);
/* cell-centred fields: padded arrays (cells + 2 per dimension), see examg_oracle.c */
void og_cell_apply_bc(int dim, const int n[3], double *u, int kind, const double *const g[6]);
void og_cell_restrict(int dim, const int nf[3], const int nc[3], const double *rf, double *rc);
void og_cell_prolong_add(int dim, const int nf[3], const int nc[3], const double *uc, double *uf);
/* domain_rect_periodic_{x,y,z}: process-wide switch, see examg_oracle.c */
void og_set_periodic(const int per[3]);
/* out[i] = (double)rand()/RAND_MAX, i = 0..n-1, after srand(1): what `native("((double)std::rand()/RAND_MAX)")` in a
 * `loop ... sequentially` yields in a fresh process (Testing/Opts/base.exa4:167-169; glibc's rand) */
void og_glibc_rand_fill(double *out, long long n);

/* RHS@coarser = scale * NodeRestriction * Residual  (…exa4:49-77,221-224;
 * C/s/e/operator/l4/L4_DefaultRestriction.scala:29-90) */
void og_restrict(int dim, const int nf[3], const int nc[3], double scale,
                 const double *rf, double *rhsc);

/* Solution += NodeProlongation@coarser * Solution@coarser  (…exa4:79-107,233-236;
 * C/s/e/operator/l4/L4_DefaultProlongation.scala:30-45) */
void og_prolong_add(int dim, const int nf[3], const int nc[3], const double *uc,
                    double *uf);

/* interior-only streaming loops (…exa4:160-162,181-198,226-228) */
/* `loop over dst { dst = expr }`: a postfix program (the EXAMG_OP_* codes of include/examg.h: 0 const, 1-3 x/y/z, 4-7
 * + - * /, 8 neg, 9 exp, 10 sin, 11 cos, 12 sinh, 13 sqrt, 15 source field vals[q] at the same node) evaluated on the
 * loop points; node position = i * h + x0 (C/s/e/grid/ir/IR_VF_NodePosition.scala:109-111).  The pointwise loops of
 * Testing/NonLinear/FAS_2D_Basic.exa4 (Smoother, UpResidual, Restriction). */
void og_loop_eval(const int n[3], const double h[3], const double x0[3], int nsrc, const double *const *src,
                  double *dst, const int *ops, const double *vals, int nops);
void og_set_interior(const int n[3], double *a, double v);
void og_copy_interior(const int n[3], const double *src, double *dst);
void og_axpy(const int n[3], double alpha, const double *x, double *y);  /* y = fma(alpha,x,y) */
void og_xpay(const int n[3], const double *x, double beta, double *y);   /* y = fma(beta,y,x)  */

/* ---- whole solver (Solve / mgCycle of the L4 programs) ------------------ */

enum { OG_SMOOTH_JACOBI = 0, OG_SMOOTH_RBGS = 1 };
enum { OG_OP_LAPLACE_SCALED = 0,   /* 2/h^2.., -1/h^2 per level (Benchmark, Examples, SISC ConstCoeff) */
       OG_OP_LAPLACE_UNSCALED = 1, /* [2d, -1 ...] on every level (Testing/Smoothers, Testing/Opts)    */
       OG_OP_VARCOEFF = 2,         /* 7-pt stencil field (Testing/SISC/3D_VarCoeff.exa4)                */
       OG_OP_VARCOEFF27 = 3 };     /* 27-pt stencil field of BASELINE config 3 (the reference ships none: SURVEY 8h);
                                      defined here, see og_init_varcoeff27                                */
enum { OG_ORDER_BENCH = 0,   /* centre,-x,+x,-y,+y,-z,+z  (Benchmark/Poisson3D, Examples/Poisson) */
       OG_ORDER_TESTING = 1 }; /* centre,+x,-x,+y,-y,+z,-z  (Testing/Smoothers, SISC, Opts)      */

/* solver_cgs = "Smoother" (C/s/e/solver/l3/L3_VankaAsSolverForEquation.scala:31-94): kind 0 Jacobi, 1 red-black GS; ten
 * smoother steps per residual check; returns the number of outer steps */
int og_smoother_solve(const int n[3], const og_stencil *st, const double *sf, double *u, const double *rhs, double *res,
                      int kind, double omega, int max_it, double rel_tol);

typedef struct {
  int dim;              /* 2 or 3 */
  int F[3];             /* global cells per dim at level 0 = blocks*frags*fragLen */
  int min_level, max_level;
  int op;               /* OG_OP_* */
  int order;            /* OG_ORDER_* */
  double kappa;         /* OG_OP_VARCOEFF, OG_OP_VARCOEFF27 */
  int smoother;         /* OG_SMOOTH_* */
  double omega;
  int nu_pre, nu_post;
  double restrict_scale; /* 1 or 4 (Testing/Smoothers: 4.0 * RestrictionStencil) */
  int cg_max_it;
  double cg_rel_tol;     /* 1e-3 */
  double domain_hi[3];   /* Domain global< [0,0,0] to hi >; 0 = 1.0 (the unit cube of every shipped program).  bench.py's
                            weak-scaling runs grow the domain with the block grid so that the mesh width stays put */
  int use_fas;           /* solver_useFAS: full approximation scheme (C/s/e/solver/l3/L3_SolverForEquation.scala:401-452) */
} og_config;

/* OpenMP team size of the `parallel for` loops (omp_numThreads of the Knowledge file,
 * C/s/e/parallelization/api/omp/OMP_Loop.scala:101-110).  A launcher such as torchrun exports OMP_NUM_THREADS=1;
 * the timed CPU baseline sets the count explicitly and reports what OpenMP really uses. */
void og_set_num_threads(int n);
int og_get_max_threads(void);
/* threads of an actual parallel region (what a loop gets, after OMP_DYNAMIC / cgroup limits) */
int og_count_threads(void);

typedef struct og_solver og_solver;

og_solver *og_solver_create(const og_config *cfg);
void og_solver_destroy(og_solver *s);
/* Logical dense arrays of level lvl (slot = active slot for Solution). */
double *og_solver_solution(og_solver *s, int lvl);
double *og_solver_rhs(og_solver *s, int lvl);
double *og_solver_residual(og_solver *s, int lvl);
void og_solver_dims(const og_solver *s, int lvl, int n[3]);
/* stencil-field coefficients [nst][size] of level lvl (NULL for constant stencils) */
double *og_solver_coeff(og_solver *s, int lvl);
void og_solver_stencil(const og_solver *s, int lvl, og_stencil *out);
/* copy the active Solution slot (incl. Dirichlet nodes) into the other slot */
void og_solver_sync_slots(og_solver *s);
/* UpResidual@lvl + ResNorm@lvl */
double og_solver_update_residual_norm(og_solver *s, int lvl);
/* mgCycle@lvl (recursive down to min_level, CG there) */
void og_solver_vcycle(og_solver *s, int lvl);
/* number of CG iterations done by the most recent coarse solve */
int og_solver_last_cg_iters(const og_solver *s);


/* ---- optical flow: 2D cell-centred, two coupled unknowns, 2x2 block smoother ------------------
 * (examg_oracle_optflow.c; Benchmark/OptFlow2D/2D_FD_OptFlow.exa4, Testing/Application/OpticalFlow2D.exa4).
 * Dense arrays: (ny+2) x (nx+2), one ghost cell per side. */
void ogf_neumann(int nx, int ny, double *a);
void ogf_apply(int nx, int ny, const double L[5], double a2, const double *cuu, const double *cuv,
               const double *cvv, const double *u, const double *v, double *ou, double *ov);
void ogf_residual(int nx, int ny, const double L[5], double a2, const double *cuu, const double *cuv,
                  const double *cvv, const double *u, const double *v, const double *fu, const double *fv,
                  double *ru, double *rv);
void ogf_smooth_color(int nx, int ny, const double L[5], double a2, const double *cuu, const double *cuv,
                      const double *cvv, double *u, double *v, const double *fu, const double *fv,
                      double relax, int color);
void ogf_restrict(int nxc, int nyc, const double *fine, double *coarse);
void ogf_prolong_add(int nxc, int nyc, const double *coarse, double *fine);
double ogf_dot2(int nx, int ny, const double *a, const double *b, const double *c, const double *d);

typedef struct {
  int F[2];              /* cells per dim at level 0 */
  int min_level, max_level;
  double alpha;          /* Globals { Val alpha } */
  double relax;          /* solve locally relax ... */
  int nu_pre, nu_post;
  int cg_max_it;
  double cg_rel_tol;
} ogf_config;
enum { OGF_U = 0, OGF_V, OGF_RHS_U, OGF_RHS_V, OGF_RES_U, OGF_RES_V, OGF_CUU, OGF_CUV, OGF_CVV, OGF_IMG0, OGF_IMG1 };
typedef struct ogf_solver ogf_solver;
ogf_solver *ogf_solver_create(const ogf_config *cfg);
void ogf_solver_destroy(ogf_solver *s);
void ogf_solver_dims(const ogf_solver *s, int lvl, int n[2]);
double *ogf_solver_field(ogf_solver *s, int lvl, int which);
void ogf_solver_init_images(ogf_solver *s);   /* initImgs@finest */
void ogf_solver_init_fields(ogf_solver *s);   /* apply bc to the images, initGradients@all, rhs */
double ogf_solver_update_residual_norm(ogf_solver *s, int lvl);
void ogf_solver_vcycle(ogf_solver *s, int lvl);
int ogf_solver_last_cg_iters(const ogf_solver *s);

/* ---- 3D optical flow: Vec3 unknown, 3x3 block inverse per cell (examg_oracle_optflow3.c;
 *      Testing/Application/OpticalFlow3D.exa4) ------------------------------------------------ */
typedef struct {
  int F[3];               /* cells per dimension at level 0 */
  int min_level, max_level;
  double alpha;
  int nu_pre, nu_post;
  int cg_max_it;
  double cg_rel_tol;
} og3_config;
/* fields of a level: flow 0..2, rhs 3..5, residual 6..8, gradContribution xx,xy,xz,yy,yz,zz 9..14 */
enum { OG3_FLOW = 0, OG3_RHS = 3, OG3_RES = 6, OG3_G = 9, OG3_IMG0 = 15, OG3_IMG1 = 16 };
typedef struct og3_solver og3_solver;
og3_solver *og3_solver_create(const og3_config *cfg);
void og3_solver_destroy(og3_solver *s);
void og3_solver_dims(const og3_solver *s, int lvl, int n[3]);
double *og3_solver_field(og3_solver *s, int lvl, int which);
void og3_solver_init_images(og3_solver *s);
void og3_solver_init_fields(og3_solver *s);
double og3_solver_update_residual_norm(og3_solver *s, int lvl);
void og3_solver_smooth_color(og3_solver *s, int lvl, int color);
void og3_solver_smooth(og3_solver *s, int lvl, int n);
void og3_solver_restrict(og3_solver *s, int lvl);
void og3_solver_correct(og3_solver *s, int lvl);
void og3_solver_vcycle(og3_solver *s, int lvl);
int og3_solver_last_cg_iters(const og3_solver *s);

/* printWithReducedPrec (C/s/e/util/ir/IR_ResolvePrintWithReducedPrec.scala:33-78):
 * formats like std::cout with precision p (default float notation). */
void og_print_reduced_prec(double v, char *buf, size_t buflen);

#ifdef __cplusplus
}
#endif
#endif
