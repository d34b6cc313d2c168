// host/Testing_Smoothers_RBGS.cpp -- what the libexamg back-end of the generator prints for
// Testing/Smoothers/RBGS.exa4 (+ RBGS.knowledge): the reference's leveled user functions with every loop
// nest replaced by a call into the C ABI (include/examg.h).  One C++ function per L4 function, same names,
// same call order, same printing (std::cout with the precision the program sets), so its stdout can be
// compared with the reference's own golden file Testing/Smoothers/RBGS.results.
//
// Build: make -C host   (g++ against libexamg.so; no CUDA toolkit needed)
//
// Knowledge: dimensionality 3, minLevel 0, maxLevel 6, 3x3x3 blocks of 3x3x3 fragments.  The 27 x 27 fragments
// of the reference form one global grid of 9 * 2^level cells per dimension; here that grid is one block on
// one GPU (block_cells = 9), which is arithmetically the same program (SURVEY.md 8(g) "global-grid equivalence").
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
static int Laplace[maxLevel + 1];  // Stencil Laplace@all

static void initGlobals() {}

static void initDomain() {
  examg_config cfg = {};
  cfg.dim = 3;
  cfg.min_level = minLevel;
  cfg.max_level = maxLevel;
  for (int d = 0; d < 3; ++d) {
    cfg.block_cells[d] = 9;  // domain_rect_numBlocks_d * domain_rect_numFragsPerBlock_d
    cfg.num_blocks[d] = 1;
    cfg.domain_lo[d] = 0.0;
    cfg.domain_hi[d] = 1.0;  // Domain global< [0,0,0] to [1,1,1] >
  }
  ctx = examg_init(&cfg);
  if (!ctx) {
    std::cerr << "examg: " << examg_last_error() << std::endl;
    exit(1);
  }
}

// setupBuffers(): Field / Stencil declarations of the program (fields are created zero-filled = initFieldsWithZero)
static void setupBuffers() {
  static const int offsets[7][3] = {{0, 0, 0}, {1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
  static const double coeffs[7] = {6.0, -1.0, -1.0, -1.0, -1.0, -1.0, -1.0};
  for (int lvl = minLevel; lvl <= maxLevel; ++lvl) {
    EXAMG_CHECK(f_Solution[lvl] = examg_field_create(ctx, "Solution", lvl, 1, 1));
    EXAMG_CHECK(f_Residual[lvl] = examg_field_create(ctx, "Residual", lvl, 1, 1));
    EXAMG_CHECK(f_RHS[lvl] = examg_field_create(ctx, "RHS", lvl, 1, 1));
    EXAMG_CHECK(examg_field_set_bc(ctx, f_Solution[lvl], EXAMG_BC_DIRICHLET_CONST, 0.0));
    EXAMG_CHECK(examg_field_set_bc(ctx, f_Residual[lvl], EXAMG_BC_DIRICHLET_CONST, 0.0));
    EXAMG_CHECK(Laplace[lvl] = examg_stencil_const(ctx, lvl, 7, offsets, coeffs));
  }
  // Field Solution< global, CommFullTempBlockable, ((x*x) - ((0.5*y)*y)) - ((0.5*z)*z) >@finest, x = vf_boundaryCoord_x
  static const examg_bc_op bc[] = {
      {EXAMG_OP_X, 0}, {EXAMG_OP_X, 0}, {EXAMG_OP_MUL, 0},
      {EXAMG_OP_CONST, 0.5}, {EXAMG_OP_Y, 0}, {EXAMG_OP_MUL, 0}, {EXAMG_OP_Y, 0}, {EXAMG_OP_MUL, 0}, {EXAMG_OP_SUB, 0},
      {EXAMG_OP_CONST, 0.5}, {EXAMG_OP_Z, 0}, {EXAMG_OP_MUL, 0}, {EXAMG_OP_Z, 0}, {EXAMG_OP_MUL, 0}, {EXAMG_OP_SUB, 0}};
  EXAMG_CHECK(examg_field_set_bc_expr(ctx, f_Solution[maxLevel], bc, (int)(sizeof(bc) / sizeof(bc[0]))));
  EXAMG_CHECK(f_VecP = examg_field_create(ctx, "VecP", minLevel, 1, 1));
  EXAMG_CHECK(f_VecGradP = examg_field_create(ctx, "VecGradP", minLevel, 1, 1));
  EXAMG_CHECK(examg_field_set_bc(ctx, f_VecP, EXAMG_BC_DIRICHLET_CONST, 0.0));
}

static void destroyGlobals() { EXAMG_CHECK(examg_finalize(ctx)); }

// ---- User/User_<fct>_<lvl>.cpp: one function per leveled L4 function ---------------------------

// Function UpResidual@all { communicate Solution; loop over Residual { Residual = RHS - Laplace * Solution } }
static void UpResidual(int lvl) {
  EXAMG_CHECK(examg_residual(ctx, Laplace[lvl], f_Solution[lvl], -1, f_RHS[lvl], f_Residual[lvl], NULL));
}

// Function NormResidual_0@(finest, coarsest): loop over Residual where (i0>0 && i1>0 && i2>0) with reduction(+ : res)
static double NormResidual_0(int lvl) {
  double res = 0.0;
  EXAMG_CHECK(examg_dot(ctx, f_Residual[lvl], -1, f_Residual[lvl], -1, &res));
  return std::sqrt(res);
}

// Function Smoother@((coarsest + 1) to finest): the two colour loops with their `communicate`s
static void Smoother(int lvl) {
  EXAMG_CHECK(examg_smooth_rbgs(ctx, Laplace[lvl], f_Solution[lvl], f_RHS[lvl], 1.0, 1));
}

// Function Restriction@...: communicate ghost of Residual; RHS@coarser = (4.0 * RestrictionStencil) * Residual
static void Restriction(int lvl) { EXAMG_CHECK(examg_restrict(ctx, f_Residual[lvl], f_RHS[lvl - 1], 4.0)); }

// Function Correction@...: communicate ghost of Solution@coarser; Solution += CorrectionStencil * Solution@coarser
static void Correction(int lvl) {
  EXAMG_CHECK(examg_exchange(ctx, f_Solution[lvl - 1], -1));
  EXAMG_CHECK(examg_prolong_add(ctx, f_Solution[lvl - 1], -1, f_Solution[lvl], -1));
}

// Function SetSolution@all ( value : Real )
static void SetSolution(int lvl, double value) { EXAMG_CHECK(examg_set(ctx, f_Solution[lvl], -1, value)); }

// Function VCycle_0@coarsest: the CG of the program text (VecP, VecGradP), 512 steps, 0.001 * initialRes
static void VCycle_0() {
  EXAMG_CHECK(examg_cg_coarse(ctx, Laplace[minLevel], f_Solution[minLevel], f_RHS[minLevel], f_Residual[minLevel], f_VecP,
                              f_VecGradP, 512, 0.001, NULL));
}

static void VCycle(int lvl) {
  if (lvl == minLevel) {  // Function VCycle@coarsest { VCycle_0@current ( ) }
    VCycle_0();
    return;
  }
  // `repeat 3 times { Smoother@current ( ) }` is folded into the sweep count of one call (the library fuses sweeps)
  EXAMG_CHECK(examg_smooth_rbgs(ctx, Laplace[lvl], f_Solution[lvl], f_RHS[lvl], 1.0, 3));
  UpResidual(lvl);
  Restriction(lvl);
  SetSolution(lvl - 1, 0);
  VCycle(lvl - 1);
  Correction(lvl);
  EXAMG_CHECK(examg_smooth_rbgs(ctx, Laplace[lvl], f_Solution[lvl], f_RHS[lvl], 1.0, 3));
}

static void InitSolution() {
  EXAMG_CHECK(examg_set(ctx, f_Solution[maxLevel], -1, 0));
  EXAMG_CHECK(examg_apply_bc(ctx, f_Solution[maxLevel], -1));  // apply bc to Solution@finest
}

static void InitRHS() { EXAMG_CHECK(examg_set(ctx, f_RHS[maxLevel], -1, 0.0)); }

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
    printReduced(res_0);
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
  (void)Smoother;
  Solve();
  destroyGlobals();
  return 0;
}
