// examg_fast7.cu -- 3D 7-point constant-coefficient fast path (placeholder until the
// tiled kernels land; dispatch falls through to the generic kernels).
#include "examg_internal.h"
namespace examg {
bool fast7_usable(examg_ctx *c, const Stencil &s) { (void)c; (void)s; return false; }
int fast7_residual(examg_ctx *c, const Stencil &s, const double *u, const double *rhs, double *res, int red_slot) {
  (void)c; (void)s; (void)u; (void)rhs; (void)res; (void)red_slot; return set_error(EXAMG_EUNSUP, "fast7");
}
int fast7_rb(examg_ctx *c, const Stencil &s, double omega, double *u, const double *rhs, int color) {
  (void)c; (void)s; (void)omega; (void)u; (void)rhs; (void)color; return set_error(EXAMG_EUNSUP, "fast7");
}
}
