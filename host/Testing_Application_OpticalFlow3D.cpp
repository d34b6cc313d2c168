// host/Testing_Application_OpticalFlow3D.cpp -- host program for Testing/Application/OpticalFlow3D.exa4
// (+ OpticalFlow3D.knowledge: 3D, minLevel 2, maxLevel 6) on libexamg: `Application`, `InitFields`, `doSolve@finest`,
// `mgCycle@all` statement by statement through the examg_optflow3_* entry points (EXAMG_HOST_LIBRARY_CYCLE=1: the
// cycle as the library's CUDA graph).  Prints what the reference prints (`printVal`), so stdout can be compared with
// Testing/Application/OpticalFlow3D.results.  Optional argument: maxLevel (default 6).
#include <cmath>
#include <cstdlib>
#include <iostream>
#include <vector>

#include "examg.h"

#define EXAMG_CHECK(call)                                        \
  do {                                                           \
    if ((call) < 0) {                                            \
      std::cerr << "examg: " << examg_last_error() << std::endl; \
      exit(1);                                                   \
    }                                                            \
  } while (0)

static examg_ctx *ctx = nullptr;
static examg_optflow3 *flow = nullptr;
static int maxLevel = 6;
static const int minLevel = 2;
static bool libraryCycle = false;

// Function printVal ( toPrint : Real )
static void printVal(double toPrint) {
  if (toPrint <= 1e-12) {
    std::cout << "EFFECTIVELY ZERO" << std::endl;
    return;
  }
  std::streamsize oldPrec = std::cout.precision();
  std::cout.precision(toPrint <= 1e-11 ? 1 : (toPrint <= 1e-10 ? 2 : (toPrint <= 1e-9 ? 3 : 4)));
  std::cout << toPrint << std::endl;
  std::cout.precision(oldPrec);
}

// Function initImgs@finest: img0 = sin ( x ) + cos ( y ) + sin ( z ) at the cell centres, img1 shifted
static void initImgs() {
  int n[3];
  EXAMG_CHECK(examg_optflow3_cells(flow, maxLevel, n));
  const double h[3] = {1.0 / n[0], 1.0 / n[1], 1.0 / n[2]};
  const size_t sy = (size_t)n[0] + 2, sz = sy * ((size_t)n[1] + 2);
  std::vector<double> img0(sz * ((size_t)n[2] + 2), 0.0), img1(img0.size(), 0.0);  // one ghost cell per side
  for (int k = 0; k < n[2]; ++k)
    for (int j = 0; j < n[1]; ++j)
      for (int i = 0; i < n[0]; ++i) {
        const double x = ((double)i + 0.5) * h[0], y = ((double)j + 0.5) * h[1], z = ((double)k + 0.5) * h[2];
        const size_t p = (size_t)(i + 1) + sy * (size_t)(j + 1) + sz * (size_t)(k + 1);
        img0[p] = std::sin(x) + std::cos(y) + std::sin(z);
        img1[p] = std::sin(x - 0.01) + std::cos(y + 0.01) + std::sin(z - 0.005);
      }
  EXAMG_CHECK(examg_optflow3_upload(flow, maxLevel, EXAMG_OF3_IMG0, img0.data()));
  EXAMG_CHECK(examg_optflow3_upload(flow, maxLevel, EXAMG_OF3_IMG1, img1.data()));
}

// Function InitFields: initImgs@finest; initGradients@finest (recursive); rhs@finest
static void InitFields() {
  initImgs();
  EXAMG_CHECK(examg_optflow3_init_fields(flow));  // apply bc to the images, gradients, coefficient hierarchy, rhs
}

// Function mgCycle@(all but coarsest) / mgCycle@coarsest
static void mgCycle(int level) {
  if (level == minLevel) {
    EXAMG_CHECK(examg_optflow3_cg(flow, nullptr));  // the CG loop of mgCycle@coarsest
    return;
  }
  EXAMG_CHECK(examg_optflow3_smooth(flow, level, 3));           // repeat 3 times { repeat with { colour 0, colour 1 } }
  EXAMG_CHECK(examg_optflow3_residual(flow, level, nullptr));   // communicate flow; residual = rhs - combinedOp * flow; apply bc
  EXAMG_CHECK(examg_optflow3_restrict(flow, level));            // rhs@coarser = RestrictionCell * residual; flow@coarser = 0
  mgCycle(level - 1);
  EXAMG_CHECK(examg_optflow3_correct(flow, level));             // flow += ProlongationCell@coarser * flow@coarser; apply bc
  EXAMG_CHECK(examg_optflow3_smooth(flow, level, 3));
}

// Function doSolve@finest
static void doSolve() {
  double initRes = 0.0;
  EXAMG_CHECK(examg_optflow3_residual(flow, maxLevel, &initRes));  // update residual; resNorm
  double curRes = initRes, prevRes = curRes;
  printVal(initRes);
  int curIt = 0;
  while (!((curIt >= 128) || (curRes <= (1.0E-5 * initRes)))) {
    curIt += 1;
    if (libraryCycle)
      EXAMG_CHECK(examg_optflow3_vcycle(flow, maxLevel));
    else
      mgCycle(maxLevel);
    prevRes = curRes;
    EXAMG_CHECK(examg_optflow3_residual(flow, maxLevel, &curRes));
    printVal(curRes);
  }
  (void)prevRes;
}

// Function Application
int main(int argc, char **argv) {
  if (argc > 1) maxLevel = atoi(argv[1]);
  const char *lc = getenv("EXAMG_HOST_LIBRARY_CYCLE");
  libraryCycle = lc && lc[0] == '1';
  examg_config cfg = {};
  cfg.dim = 3;
  cfg.min_level = minLevel;
  cfg.max_level = maxLevel;
  for (int d = 0; d < 3; ++d) {
    cfg.block_cells[d] = 1;
    cfg.num_blocks[d] = 1;
    cfg.domain_lo[d] = 0.0;
    cfg.domain_hi[d] = 1.0;
  }
  ctx = examg_init(&cfg);  // initGlobals, initDomain
  if (!ctx) {
    std::cerr << "examg: " << examg_last_error() << std::endl;
    return 1;
  }
  examg_optflow3_desc desc;
  desc.alpha = 1.0;  // Globals { Val alpha : Real = 1 }
  desc.nu_pre = desc.nu_post = 3;
  desc.cg_max_it = 128;
  desc.cg_rel_tol = 0.001;
  flow = examg_optflow3_create(ctx, &desc);  // setupBuffers, initFieldsWithZero
  if (!flow) {
    std::cerr << "examg: " << examg_last_error() << std::endl;
    return 1;
  }
  InitFields();
  doSolve();
  EXAMG_CHECK(examg_optflow3_destroy(flow));
  EXAMG_CHECK(examg_finalize(ctx));  // destroyGlobals
  return 0;
}
