// host/Testing_Application_OpticalFlow2D.cpp -- host program for Testing/Application/OpticalFlow2D.exa4
// (+ OpticalFlow2D.knowledge: 2D, minLevel 2, maxLevel 8) on libexamg: `Application`, `InitFields`, `doSolve@finest`
// with the cycle as the library's whole-program object (examg_optflow_*).  Prints what the reference prints
// (`printVal`), so stdout can be compared with Testing/Application/OpticalFlow2D.results.
// Optional argument: maxLevel (default 8; 13 = the 8192^2 configuration of BASELINE.json).
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
static examg_optflow *flow = nullptr;
static int maxLevel = 8;
static const int minLevel = 2;

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

// Function initImgs@finest: img0 = sin(vf_cellCenter_x) + cos(vf_cellCenter_y), img1 shifted by 0.01
static void initImgs() {
  int n[3];
  EXAMG_CHECK(examg_level_cells(ctx, maxLevel, n));
  double h[3];
  EXAMG_CHECK(examg_level_width(ctx, maxLevel, h));
  std::vector<double> img0((size_t)n[0] * n[1]), img1((size_t)n[0] * n[1]);
  for (int j = 0; j < n[1]; ++j)
    for (int i = 0; i < n[0]; ++i) {
      const double x = ((double)i + 0.5) * h[0], y = ((double)j + 0.5) * h[1];
      img0[(size_t)j * n[0] + i] = std::sin(x) + std::cos(y);
      img1[(size_t)j * n[0] + i] = std::sin(x - 0.01) + std::cos(y + 0.01);
    }
  EXAMG_CHECK(examg_field_upload(ctx, examg_optflow_field(flow, maxLevel, EXAMG_OF_IMG0), -1, img0.data()));
  EXAMG_CHECK(examg_field_upload(ctx, examg_optflow_field(flow, maxLevel, EXAMG_OF_IMG1), -1, img1.data()));
}

// Function InitFields: initImgs@finest; initGradients@finest (recursive); rhs@finest
static void InitFields() {
  initImgs();
  EXAMG_CHECK(examg_optflow_init_fields(flow));  // apply bc to the images, gradients, coefficient hierarchy, rhs
}

// Function doSolve@finest
static void doSolve() {
  double initRes = 0.0;
  EXAMG_CHECK(examg_optflow_residual_norm(flow, &initRes));  // residual = rhs - combinedOp * flow; resNorm
  double curRes = initRes, prevRes = curRes;
  printVal(initRes);
  int curIt = 0;
  while (!((curIt >= 128) || (curRes <= (1.0E-5 * initRes)))) {
    curIt += 1;
    prevRes = curRes;
    EXAMG_CHECK(examg_optflow_cycles(flow, 1, &curRes));  // mgCycle ( ); residual update; resNorm
    printVal(curRes);
  }
  (void)prevRes;
}

// Function Application
int main(int argc, char **argv) {
  if (argc > 1) maxLevel = atoi(argv[1]);
  examg_config cfg = {};
  cfg.dim = 2;
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
  examg_optflow_desc desc;
  desc.alpha = 1.0;  // Globals { Val alpha : Real = 1 }
  desc.relax = 1.0;
  desc.nu_pre = desc.nu_post = 3;  // repeat 3 times
  desc.cg_max_it = 128;
  desc.cg_rel_tol = 0.001;
  flow = examg_optflow_create(ctx, &desc);  // setupBuffers, initFieldsWithZero
  if (!flow) {
    std::cerr << "examg: " << examg_last_error() << std::endl;
    return 1;
  }
  InitFields();
  doSolve();
  EXAMG_CHECK(examg_optflow_destroy(flow));
  EXAMG_CHECK(examg_finalize(ctx));  // destroyGlobals
  return 0;
}
