// host/Benchmark_Poisson3D.cpp -- what the libexamg back-end of the generator prints for
// Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4 (+ 3D_FD_Poisson_fromL4_CUDA.knowledge: minLevel 2, maxLevel 9): the
// north-star program.  One C++ function per L4 function, C ABI only, same printing
// ("Starting residual: ...", "Residual after k iterations is ... --- convergence factor is ..."), timers through
// examg_timer_*, results.json as printJSON writes it.
//
//   host/_build/Benchmark_Poisson3D [maxLevel = 9] [nuPre = nuPost = 3]
//
// The shipped program text says `repeat 3 times` (V(3,3)); BASELINE.json's north star measures V(2,2): the second
// argument selects it.  The level loop runs either statement by statement (mgCycle below, the form the back-end prints)
// or -- EXAMG_HOST_LIBRARY_CYCLE=1 -- through examg_solver_* (one CUDA graph per cycle), which is bit-identical.
#include <cmath>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <vector>

#include "examg.h"

#define EXAMG_CHECK(call)                                                        \
  do {                                                                           \
    if ((call) < 0) {                                                            \
      std::cerr << "examg: " << examg_last_error() << std::endl;                 \
      exit(1); /* reference convention: IR_Assert / CUDA_CHECK -> exit(1) */     \
    }                                                                            \
  } while (0)

// ---- Global/Global.h ------------------------------------------------------------------------
static const int minLevel = 2;
static int maxLevel = 9, nuPre = 3, nuPost = 3;
static examg_ctx *ctx = nullptr;
static examg_solver *solver = nullptr;  // owns Solution, RHS, Residual, cgTmp0, cgTmp1 of every level
static std::vector<int> f_Solution, f_RHS, f_Residual, Laplace;
static int f_cgTmp0, f_cgTmp1;
static bool libraryCycle = false;

static int lv(int lvl) { return lvl - minLevel; }

static void initGlobals() {}

static void initDomain() {
  examg_config cfg = {};
  cfg.dim = 3;
  cfg.min_level = minLevel;
  cfg.max_level = maxLevel;
  for (int d = 0; d < 3; ++d) {
    cfg.block_cells[d] = 1;  // one fragment of fragmentLength 1 per block
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

// setupBuffers() + initFieldsWithZero(): Stencil Laplace@all, the fields of the program
static void setupBuffers() {
  // [0,0,0] => 2/hx^2 + 2/hy^2 + 2/hz^2; [-1,0,0] => -1/hx^2; [1,0,0]; [0,-1,0]; [0,1,0]; [0,0,-1]; [0,0,1]
  static const int offsets[7][3] = {{0, 0, 0}, {-1, 0, 0}, {1, 0, 0}, {0, -1, 0}, {0, 1, 0}, {0, 0, -1}, {0, 0, 1}};
  for (int lvl = minLevel; lvl <= maxLevel; ++lvl) {
    double h[3];
    EXAMG_CHECK(examg_level_width(ctx, lvl, h));
    const double coeffs[7] = {2.0 / (h[0] * h[0]) + 2.0 / (h[1] * h[1]) + 2.0 / (h[2] * h[2]),
                              -1.0 / (h[0] * h[0]), -1.0 / (h[0] * h[0]), -1.0 / (h[1] * h[1]), -1.0 / (h[1] * h[1]),
                              -1.0 / (h[2] * h[2]), -1.0 / (h[2] * h[2])};
    int A;
    EXAMG_CHECK(A = examg_stencil_const(ctx, lvl, 7, offsets, coeffs));
    Laplace.push_back(A);
  }
  examg_cycle_desc desc = {};
  desc.smoother = EXAMG_SMOOTH_RBGS;
  desc.omega = 0.8;
  desc.nu_pre = nuPre;
  desc.nu_post = nuPost;
  desc.restrict_scale = 1.0;
  desc.cg_max_it = 128;
  desc.cg_rel_tol = 0.001;
  desc.stencils = Laplace.data();
  solver = examg_solver_create(ctx, &desc);
  if (!solver) {
    std::cerr << "examg: " << examg_last_error() << std::endl;
    exit(1);
  }
  for (int lvl = minLevel; lvl <= maxLevel; ++lvl) {
    f_Solution.push_back(examg_solver_field(solver, lvl, EXAMG_F_SOLUTION));
    f_RHS.push_back(examg_solver_field(solver, lvl, EXAMG_F_RHS));
    f_Residual.push_back(examg_solver_field(solver, lvl, EXAMG_F_RESIDUAL));
  }
  f_cgTmp0 = examg_solver_field(solver, minLevel, EXAMG_F_CGTMP0);
  f_cgTmp1 = examg_solver_field(solver, minLevel, EXAMG_F_CGTMP1);
  // Field Solution< global, NodeWithComm, vf_boundaryPosition_x ** 2 - 0.5 * vf_boundaryPosition_y ** 2 - 0.5 * ... >@finest
  static const examg_bc_op bc[] = {
      {EXAMG_OP_X, 0}, {EXAMG_OP_X, 0}, {EXAMG_OP_MUL, 0},
      {EXAMG_OP_CONST, 0.5}, {EXAMG_OP_Y, 0}, {EXAMG_OP_MUL, 0}, {EXAMG_OP_Y, 0}, {EXAMG_OP_MUL, 0}, {EXAMG_OP_SUB, 0},
      {EXAMG_OP_CONST, 0.5}, {EXAMG_OP_Z, 0}, {EXAMG_OP_MUL, 0}, {EXAMG_OP_Z, 0}, {EXAMG_OP_MUL, 0}, {EXAMG_OP_SUB, 0}};
  EXAMG_CHECK(examg_field_set_bc_expr(ctx, f_Solution[lv(maxLevel)], bc, (int)(sizeof(bc) / sizeof(bc[0]))));
}

static void destroyGlobals() {
  EXAMG_CHECK(examg_solver_destroy(solver));
  EXAMG_CHECK(examg_finalize(ctx));
}

// ---- User/User_<fct>_<lvl>.cpp ------------------------------------------------------------------

// Function ResNorm@(coarsest and finest)
static double ResNorm(int lvl) {
  double norm = 0.0;
  EXAMG_CHECK(examg_dot(ctx, f_Residual[lv(lvl)], -1, f_Residual[lv(lvl)], -1, &norm));
  return std::sqrt(norm);
}

// communicate Solution; loop over Residual { Residual = RHS - Laplace * Solution }; apply bc to Residual
static void UpResidual(int lvl) {
  EXAMG_CHECK(examg_residual(ctx, Laplace[lv(lvl)], f_Solution[lv(lvl)], -1, f_RHS[lv(lvl)], f_Residual[lv(lvl)], NULL));
  EXAMG_CHECK(examg_apply_bc(ctx, f_Residual[lv(lvl)], -1));
}

// Function mgCycle@coarsest: the CG of the program text (:152-201)
static void mgCycle_coarsest() {
  EXAMG_CHECK(examg_cg_coarse(ctx, Laplace[0], f_Solution[0], f_RHS[0], f_Residual[0], f_cgTmp0, f_cgTmp1, 128, 0.001, NULL));
}

// Function mgCycle@(all but coarsest) (:203-249)
static void mgCycle(int lvl) {
  if (lvl == minLevel) {
    mgCycle_coarsest();
    return;
  }
  const int l = lv(lvl);
  // repeat nuPre times { color with { (i0+i1+i2) % 2, communicate Solution; loop over Solution {...}; apply bc to Solution } }
  EXAMG_CHECK(examg_smooth_rbgs(ctx, Laplace[l], f_Solution[l], f_RHS[l], 0.8, nuPre));
  // communicate Solution; Residual = RHS - Laplace * Solution; apply bc to Residual; communicate Residual;
  // RHS@coarser = NodeRestriction * Residual   -- Residual has no other reader: one call, one pass on large levels
  EXAMG_CHECK(examg_residual_restrict(ctx, Laplace[l], f_Solution[l], f_RHS[l], f_Residual[l], f_RHS[l - 1], 1.0));
  EXAMG_CHECK(examg_set(ctx, f_Solution[l - 1], -1, 0.0));  // Solution@coarser = 0.0
  EXAMG_CHECK(examg_apply_bc(ctx, f_Solution[l - 1], -1));
  mgCycle(lvl - 1);
  EXAMG_CHECK(examg_exchange(ctx, f_Solution[l - 1], -1));  // communicate Solution@coarser
  EXAMG_CHECK(examg_prolong_add(ctx, f_Solution[l - 1], -1, f_Solution[l], -1));
  EXAMG_CHECK(examg_apply_bc(ctx, f_Solution[l], -1));
  EXAMG_CHECK(examg_smooth_rbgs(ctx, Laplace[l], f_Solution[l], f_RHS[l], 0.8, nuPost));
}

// Function Solve@finest (:121-150)
static int Solve() {
  UpResidual(maxLevel);
  double initRes = ResNorm(maxLevel);
  double curRes = initRes;
  double prevRes = curRes;
  std::cout << "Starting residual: " << initRes << std::endl;
  int curIt = 0;
  while (!(curIt >= 100 || curRes <= 1.0E-6 * initRes)) {
    curIt += 1;
    if (libraryCycle)
      EXAMG_CHECK(examg_solver_vcycle(solver, maxLevel));
    else
      mgCycle(maxLevel);
    UpResidual(maxLevel);
    prevRes = curRes;
    curRes = ResNorm(maxLevel);
    std::cout << "Residual after " << curIt << " iterations is " << curRes << " --- convergence factor is "
              << curRes / prevRes << std::endl;
  }
  return curIt;
}

// Function Application
int main(int argc, char **argv) {
  if (argc > 1) maxLevel = atoi(argv[1]);
  if (argc > 2) nuPre = nuPost = atoi(argv[2]);
  const char *e = getenv("EXAMG_HOST_LIBRARY_CYCLE");
  libraryCycle = e && e[0] == '1';
  // init
  initGlobals();
  initDomain();
  EXAMG_CHECK(examg_timer_start(ctx, "setup"));
  setupBuffers();  // incl. initFieldsWithZero, initGeometry
  EXAMG_CHECK(examg_apply_bc(ctx, f_Solution[lv(maxLevel)], -1));  // apply bc to Solution@finest
  EXAMG_CHECK(examg_timer_stop(ctx, "setup"));
  // solve
  EXAMG_CHECK(examg_timer_start(ctx, "solve"));
  EXAMG_CHECK(examg_benchmark_start(ctx, "solvePhase"));  // benchmarkStart ( "solvePhase" ): an NVTX range on B200
  const int its = Solve();
  EXAMG_CHECK(examg_benchmark_stop(ctx, "solvePhase"));
  EXAMG_CHECK(examg_timer_stop(ctx, "solve"));
  // de-init
  EXAMG_CHECK(examg_timer_print_all(ctx));
  double tSolve = 0.0, tSetup = 0.0;
  long long n = 0;
  EXAMG_CHECK(examg_timer_total_ms(ctx, "solve", &tSolve, &n));
  EXAMG_CHECK(examg_timer_total_ms(ctx, "setup", &tSetup, &n));
  const double cells = std::pow(2.0, maxLevel), dof = (cells - 1.0) * (cells - 1.0) * (cells - 1.0);
  std::cout << "solve phase: " << its << " cycles, " << dof * its / (tSolve * 1e-3) / 1e9 << " GDoF-cycles/s" << std::endl;
  destroyGlobals();
  // printJSON ( "results.json", 'totalTimeSolve', getTotalTime ( 'solve' ), 'totalSetupTime', getTotalTime ( 'setup' ) )
  std::ofstream js("results.json");
  js << "{ \"totalTimeSolve\": " << tSolve << ", \"totalSetupTime\": " << tSetup << " }" << std::endl;
  return 0;
}
