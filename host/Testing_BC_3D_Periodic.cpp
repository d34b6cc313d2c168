// host/Testing_BC_3D_Periodic.cpp -- what the libexamg back-end of the generator prints for
// Testing/BC/3D_Periodic.exa4 (+ 3D_Periodic.knowledge): Dirichlet data x^2 on the x faces, periodic in y and z
// (domain_rect_periodic_y/z), RHS = -2, scaled 7-point stencil, omega = 0.8 Jacobi V(3,3), CG on the coarsest level,
// NormError_0 printed after every cycle.  One C++ function per L4 function, C ABI only; stdout must equal
// Testing/BC/3D_Periodic.results.
//
// Knowledge: dimensionality 3, minLevel 0, maxLevel 6, 2x2x2 blocks of 2x2x2 fragments = one global grid of
// 4 * 2^level cells per dimension, here one block on one GPU (block_cells = 4).
#include <cmath>
#include <cstdlib>
#include <iostream>

#include "examg.h"

#define EXAMG_CHECK(call)                                                        \
  do {                                                                           \
    if ((call) < 0) {                                                            \
      std::cerr << "examg: " << examg_last_error() << std::endl;                 \
      exit(1); /* reference convention: IR_Assert / CUDA_CHECK -> exit(1) */     \
    }                                                                            \
  } while (0)

// ---- Global/Global.h ------------------------------------------------------------------------
static const int minLevel = 0, maxLevel = 6;
static examg_ctx *ctx = nullptr;
static int f_Solution[maxLevel + 1], f_Residual[maxLevel + 1], f_RHS[maxLevel + 1];
static int f_VecP, f_VecGradP;    // @coarsest
static int f_Exact;               // vf_nodePosition_x ** 2 at the finest level (operand of NormError_0)
static int Laplace[maxLevel + 1];  // Stencil Laplace@all

static void initGlobals() {}

static void initDomain() {
  examg_config cfg = {};
  cfg.dim = 3;
  cfg.min_level = minLevel;
  cfg.max_level = maxLevel;
  for (int d = 0; d < 3; ++d) {
    cfg.block_cells[d] = 4;  // domain_rect_numBlocks_d * domain_rect_numFragsPerBlock_d
    cfg.num_blocks[d] = 1;
    cfg.domain_lo[d] = 0.0;
    cfg.domain_hi[d] = 1.0;  // Domain global< [0,0,0] to [1,1,1] >
  }
  cfg.periodic[1] = 1;  // domain_rect_periodic_y = true
  cfg.periodic[2] = 1;  // domain_rect_periodic_z = true
  ctx = examg_init(&cfg);
  if (!ctx) {
    std::cerr << "examg: " << examg_last_error() << std::endl;
    exit(1);
  }
}

// setupBuffers(): Field / Stencil declarations (fields are created zero-filled = initFieldsWithZero)
static void setupBuffers() {
  static const int offsets[7][3] = {{0, 0, 0}, {1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
  for (int lvl = minLevel; lvl <= maxLevel; ++lvl) {
    EXAMG_CHECK(f_Solution[lvl] = examg_field_create(ctx, "Solution", lvl, 2, 1));
    EXAMG_CHECK(f_Residual[lvl] = examg_field_create(ctx, "Residual", lvl, 1, 1));
    EXAMG_CHECK(f_RHS[lvl] = examg_field_create(ctx, "RHS", lvl, 1, 1));
    EXAMG_CHECK(examg_field_set_bc(ctx, f_Solution[lvl], EXAMG_BC_DIRICHLET_CONST, 0.0));
    EXAMG_CHECK(examg_field_set_bc(ctx, f_Residual[lvl], EXAMG_BC_DIRICHLET_CONST, 0.0));
    // [0,0,0] => ((2.0/(hx*hx)) + (2.0/(hy*hy))) + (2.0/(hz*hz)); [+-1,0,0] => -1.0/(hx*hx); ...
    double h[3];
    EXAMG_CHECK(examg_level_width(ctx, lvl, h));
    const double coeffs[7] = {((2.0 / (h[0] * h[0])) + (2.0 / (h[1] * h[1]))) + (2.0 / (h[2] * h[2])),
                              -1.0 / (h[0] * h[0]), -1.0 / (h[0] * h[0]), -1.0 / (h[1] * h[1]), -1.0 / (h[1] * h[1]),
                              -1.0 / (h[2] * h[2]), -1.0 / (h[2] * h[2])};
    EXAMG_CHECK(Laplace[lvl] = examg_stencil_const(ctx, lvl, 7, offsets, coeffs));
  }
  // Field Solution< global, CommFullTempBlockable, ( vf_boundaryCoord_x * vf_boundaryCoord_x ) >[2]@finest
  static const examg_bc_op xx[] = {{EXAMG_OP_X, 0}, {EXAMG_OP_X, 0}, {EXAMG_OP_MUL, 0}};
  EXAMG_CHECK(examg_field_set_bc_expr(ctx, f_Solution[maxLevel], xx, 3));
  EXAMG_CHECK(f_VecP = examg_field_create(ctx, "VecP", minLevel, 1, 1));
  EXAMG_CHECK(f_VecGradP = examg_field_create(ctx, "VecGradP", minLevel, 1, 1));
  EXAMG_CHECK(examg_field_set_bc(ctx, f_VecP, EXAMG_BC_DIRICHLET_CONST, 0.0));
  EXAMG_CHECK(f_Exact = examg_field_create(ctx, "Exact", maxLevel, 1, 1));
  EXAMG_CHECK(examg_field_eval(ctx, f_Exact, 0, 0, xx, 3));
}

static void destroyGlobals() { EXAMG_CHECK(examg_finalize(ctx)); }

// ---- User/User_<fct>_<lvl>.cpp: one function per leveled L4 function ---------------------------

// Function UpResidual@all { communicate Solution<active>; loop over Residual { Residual = RHS - Laplace * Solution } }
static void UpResidual(int lvl) {
  EXAMG_CHECK(examg_residual(ctx, Laplace[lvl], f_Solution[lvl], -1, f_RHS[lvl], f_Residual[lvl], NULL));
}

// Function NormResidual_0@(finest, coarsest): reduction over the loop points, the periodic node counted once
static double NormResidual_0(int lvl) {
  double res = 0.0;
  EXAMG_CHECK(examg_dot(ctx, f_Residual[lvl], -1, f_Residual[lvl], -1, &res));
  return std::sqrt(res);
}

// Function NormError_0@finest: loop over Solution with reduction(max : err) { fabs(Solution<active> - x*x) }
static double NormError_0() {
  double err = 0.0;
  EXAMG_CHECK(examg_max_abs_diff(ctx, f_Solution[maxLevel], -1, f_Exact, -1, &err));
  return err;
}

// Function Smoother@...: communicate ghost of Solution<active>; Solution<next> = Solution<active> + ((1/diag)*0.8)*(...);
// advance -- `repeat 3 times` folded into the sweep count
static void Smoother(int lvl, int n) {
  EXAMG_CHECK(examg_smooth_jacobi(ctx, Laplace[lvl], f_Solution[lvl], f_RHS[lvl], 0.8, n));
}

// Function Restriction@...: communicate ghost of Residual; RHS@coarser = RestrictionStencil * Residual
static void Restriction(int lvl) { EXAMG_CHECK(examg_restrict(ctx, f_Residual[lvl], f_RHS[lvl - 1], 1.0)); }

// Function Correction@...: communicate ghost of Solution@coarser; Solution += CorrectionStencil * Solution@coarser
static void Correction(int lvl) {
  EXAMG_CHECK(examg_exchange(ctx, f_Solution[lvl - 1], -1));
  EXAMG_CHECK(examg_prolong_add(ctx, f_Solution[lvl - 1], -1, f_Solution[lvl], -1));
}

static void SetSolution(int lvl, double value) { EXAMG_CHECK(examg_set(ctx, f_Solution[lvl], -1, value)); }

// Function VCycle_0@coarsest: the CG of the program text (VecP, VecGradP), 512 steps, 0.001 * initialRes
static void VCycle_0() {
  EXAMG_CHECK(examg_cg_coarse(ctx, Laplace[minLevel], f_Solution[minLevel], f_RHS[minLevel], f_Residual[minLevel], f_VecP,
                              f_VecGradP, 512, 0.001, NULL));
}

static void VCycle(int lvl) {
  if (lvl == minLevel) {
    VCycle_0();
    return;
  }
  Smoother(lvl, 3);
  UpResidual(lvl);
  Restriction(lvl);
  SetSolution(lvl - 1, 0);
  VCycle(lvl - 1);
  Correction(lvl);
  Smoother(lvl, 3);
}

static void InitSolution() { EXAMG_CHECK(examg_set(ctx, f_Solution[maxLevel], -1, 0)); }

static void InitRHS() { EXAMG_CHECK(examg_set(ctx, f_RHS[maxLevel], -1, -2.0)); }

// the program's reduced-precision printing (native std::cout.precision(...) blocks of Solve)
static void printReduced(double v) {
  if (v <= 1.0E-12) {
    std::cout << "EFFECTIVELY ZERO" << std::endl;
    return;
  }
  std::streamsize oldPrec = std::cout.precision();
  std::cout.precision(v <= 1.0E-11 ? 1 : (v <= 9.999999999999999E-11 ? 2 : (v <= 9.999999999999999E-10 ? 3 : 4)));
  std::cout << v << std::endl;
  std::cout.precision(oldPrec);
}

static void Solve() {
  UpResidual(maxLevel);
  double resStart_0 = NormResidual_0(maxLevel);
  double res_0 = resStart_0;
  double resOld_0 = 0;
  printReduced(resStart_0);
  int numIt = 0;
  while (!((res_0 < (1.0E-5 * resStart_0)) || (numIt >= 100))) {
    numIt += 1;
    VCycle(maxLevel);
    UpResidual(maxLevel);
    resOld_0 = res_0;
    res_0 = NormResidual_0(maxLevel);
    double curError_0 = NormError_0();
    printReduced(curError_0);
  }
  (void)resOld_0;
  std::cout << numIt << std::endl;
}

// Function Application
int main() {
  initGlobals();
  initDomain();
  setupBuffers();  // incl. initFieldsWithZero
  InitRHS();
  InitSolution();
  for (int lvl = minLevel; lvl <= maxLevel; ++lvl) {  // apply bc to Solution<0>@lvl; apply bc to Solution<1>@lvl
    EXAMG_CHECK(examg_apply_bc(ctx, f_Solution[lvl], 0));
    EXAMG_CHECK(examg_apply_bc(ctx, f_Solution[lvl], 1));
  }
  EXAMG_CHECK(examg_apply_bc(ctx, f_VecP, -1));
  Solve();
  destroyGlobals();
  return 0;
}
