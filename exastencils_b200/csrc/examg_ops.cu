// examg_ops.cu -- C-ABI entry points of the hot path (argument checking, slot
// resolution, dispatch to the kernel launchers).  One entry per `loop over`
// shape of the L4 programs; see include/examg.h for the reference lines.
#include <cmath>

#include "examg_internal.h"

using namespace examg;

namespace examg {
// generic launchers (examg_kernels.cu, examg_transfer.cu)
int launch_residual_generic(examg_ctx *c, const Stencil &s, const double *u, const double *rhs, double *res, int red_slot, int extend);
int launch_jacobi_generic(examg_ctx *c, const Stencil &s, double omega, const double *u, const double *rhs, double *unext);
int launch_rb_generic(examg_ctx *c, const Stencil &s, double omega, double *u, const double *rhs, int color);
int launch_restrict_generic(examg_ctx *c, int lvl_fine, double scale, const double *rf, double *rhsc);
int launch_prolong_generic(examg_ctx *c, int lvl_fine, const double *uc, double *uf);
bool fused_rb_usable(examg_ctx *c, const Stencil &s, const Field &fu);
int launch_rb_fused(examg_ctx *c, const Stencil &s, double omega, Field &fu, const double *rhs, int nsweeps, int *done,
                    const double *uc = nullptr);
bool fused_prolong_enabled();
bool fused_rb2d_usable(examg_ctx *c, const Stencil &s, const Field &fu);
int launch_rb2d_fused(examg_ctx *c, const Stencil &s, double omega, Field &fu, const double *rhs, int nsweeps, int *done);
int rb_fused_plan(int nsweeps);
bool fused_jac_usable(examg_ctx *c, const Stencil &s, const Field &fu);
int launch_jac_fused(examg_ctx *c, const Stencil &s, double omega, Field &fu, const double *rhs, int nsweeps, int *done);
int jac_fused_plan(int nsweeps);
bool fused_jacv_usable(examg_ctx *c, const Stencil &s, const Field &fu);
int launch_jacv_fused(examg_ctx *c, const Stencil &s, double omega, Field &fu, const double *rhs);
bool residual_restrict_usable(examg_ctx *c, const Stencil &s);
int launch_residual_restrict(examg_ctx *c, const Stencil &s, const double *u, const double *rhs, double *rhsc, double scale);

int launch_apply_generic(examg_ctx *c, const Stencil &s, const double *u, double *out);
// full compact constant-coefficient stencils in lexicographic order: register-window kernels (examg_fast27.cu)
bool lex_usable(examg_ctx *c, const Stencil &s);
int launch_lex(examg_ctx *c, const Stencil &s, int mode, const double *u, const double *rhs, double *out, double omega,
               int extend);

// 3D 7-point constant coefficients: the residual loop with the z and x neighbours out of registers
bool star7_usable(examg_ctx *c, const Stencil &s);
int launch_star7(examg_ctx *c, const Stencil &s, const double *u, const double *rhs, double *res, int extend);


int launch_apply(examg_ctx *c, const Stencil &s, const double *u, double *out) {
  if (lex_usable(c, s)) return launch_lex(c, s, 0, u, nullptr, out, 0.0, 0);
  return launch_apply_generic(c, s, u, out);
}
int launch_residual(examg_ctx *c, const Stencil &s, const double *u, const double *rhs, double *res, int red_slot,
                    int extend) {
  if (red_slot < 0 && res && lex_usable(c, s)) return launch_lex(c, s, 1, u, rhs, res, 0.0, extend);
  if (red_slot < 0 && res && star7_usable(c, s)) return launch_star7(c, s, u, rhs, res, extend);
  return launch_residual_generic(c, s, u, rhs, res, red_slot, extend);
}
int launch_jacobi(examg_ctx *c, const Stencil &s, double omega, const double *u, const double *rhs, double *unext) {
  if (lex_usable(c, s)) return launch_lex(c, s, 2, u, rhs, unext, omega, 0);
  return launch_jacobi_generic(c, s, omega, u, rhs, unext);
}
int launch_rb(examg_ctx *c, const Stencil &s, double omega, double *u, const double *rhs, int color) {
  return launch_rb_generic(c, s, omega, u, rhs, color);
}
int launch_restrict(examg_ctx *c, int lvl_fine, double scale, const double *rf, double *rhsc) {
  return launch_restrict_generic(c, lvl_fine, scale, rf, rhsc);
}
int launch_prolong(examg_ctx *c, int lvl_fine, const double *uc, double *uf) {
  return launch_prolong_generic(c, lvl_fine, uc, uf);
}
}  // namespace examg

#define GET_FIELD(var, h)                                                   \
  Field *var = get_field(c, h);                                             \
  if (!var) return set_error(EXAMG_EINVAL, "%s: bad field handle %d", __func__, h)
#define GET_STENCIL(var, h)                                                 \
  Stencil *var = get_stencil(c, h);                                         \
  if (!var) return set_error(EXAMG_EINVAL, "%s: bad stencil handle %d", __func__, h)
#define CHECK_SLOT(f, slot) \
  if ((slot) >= (f)->nslots) return set_error(EXAMG_EINVAL, "%s: slot %d of %d", __func__, slot, (f)->nslots)
#define CHECK_LEVEL(f, lvl) \
  if ((f)->level != (lvl)) return set_error(EXAMG_EINVAL, "%s: field %s lives on level %d, not %d", __func__, (f)->name.c_str(), (f)->level, lvl)

int examg::residual_op(examg_ctx *c, int A, int u, int u_slot, int rhs, int res, double *norm2_out, int extend) {
  GET_STENCIL(s, A); GET_FIELD(fu, u); GET_FIELD(fr, rhs); GET_FIELD(fo, res);
  CHECK_SLOT(fu, u_slot); CHECK_LEVEL(fu, s->level); CHECK_LEVEL(fr, s->level); CHECK_LEVEL(fo, s->level);
  if (fu == fo) return set_error(EXAMG_EINVAL, "examg_residual: Residual aliases Solution");
  if (fu->loc != fr->loc || fu->loc != fo->loc) return set_error(EXAMG_EINVAL, "examg_residual: fields of different localisation");
  LocGuard lg(c, fu->loc);
  if (c->world == 1) extend = 0;
  const int us = u_slot < 0 ? fu->cur : u_slot;
  int rc;
  if (extend > 0 && s->coeff_field >= 0 && c->fields[s->coeff_field].halo_depth < extend) {
    if (extend > 1) return set_error(EXAMG_EUNSUP, "examg_residual: stencil fields carry one valid ghost layer");
    if ((rc = do_exchange(c, c->fields[s->coeff_field], 0, 1))) return rc;  // static: once
  }
  if (extend > 0 && fr->halo_depth < extend) {  // communicate Solution (and RHS where its ghost layers are stale)
    Field *fs[2] = {fu, fr};
    const int slots[2] = {us, fr->cur}, depths[2] = {extend + 1, extend};
    rc = do_exchange_many(c, fs, slots, depths, 2);
  } else {
    rc = do_exchange(c, *fu, us, extend + 1);
  }
  if (rc) return rc;
  rc = launch_residual(c, *s, field_ptr(c, fu, u_slot), field_ptr(c, fr, -1), field_ptr(c, fo, -1),
                       norm2_out ? 0 : -1, extend);
  fo->halo_depth = extend;
  if (rc || !norm2_out) return rc;
  return finish_sum(c, 0, norm2_out);
}

int examg::residual_norm2_op(examg_ctx *c, int A, int u, int rhs, double *norm2_out) {
  GET_STENCIL(s, A); GET_FIELD(fu, u); GET_FIELD(fr, rhs);
  CHECK_LEVEL(fu, s->level); CHECK_LEVEL(fr, s->level);
  if (!norm2_out) return set_error(EXAMG_EINVAL, "residual norm: null output");
  if (fu->loc != fr->loc) return set_error(EXAMG_EINVAL, "residual norm: fields of different localisation");
  LocGuard lg(c, fu->loc);
  int rc = do_exchange(c, *fu, fu->cur, 1);
  if (rc) return rc;
  if ((rc = launch_residual(c, *s, field_ptr(c, fu, -1), field_ptr(c, fr, -1), nullptr, 0, 0))) return rc;
  return finish_sum(c, 0, norm2_out);
}

// RHS@coarser = scale * Restriction * (RHS - A*Solution) without materialising Residual (examg_resrestrict.cu);
// returns 1 if the combination is not covered (the caller then runs the two separate loops)
int examg::residual_restrict_op(examg_ctx *c, int A, int u, int rhs, int rhs_coarse, double scale) {
  GET_STENCIL(s, A); GET_FIELD(fu, u); GET_FIELD(fr, rhs); GET_FIELD(fc, rhs_coarse);
  CHECK_LEVEL(fu, s->level); CHECK_LEVEL(fr, s->level); CHECK_LEVEL(fc, s->level - 1);
  if (!residual_restrict_usable(c, *s) || fu->ncomp != 1 || fu->loc != 0 || fr->loc != 0 || fc->loc != 0) return 1;
  int rc;
  if (c->world > 1) {  // the residual is needed one layer beyond the block: Solution 2 layers, RHS 1
    if (fr->halo_depth < 1) {
      Field *fs[2] = {fu, fr};
      const int slots[2] = {fu->cur, fr->cur}, depths[2] = {2, 1};
      rc = do_exchange_many(c, fs, slots, depths, 2);
    } else {
      rc = do_exchange(c, *fu, fu->cur, 2);
    }
    if (rc) return rc;
  }
  fc->halo_depth = 0;
  return launch_residual_restrict(c, *s, field_ptr(c, fu, -1), field_ptr(c, fr, -1), field_ptr(c, fc, -1), scale);
}

extern "C" {

int examg_apply(examg_ctx *c, int A, int u, int u_slot, int out) {
  GET_STENCIL(s, A); GET_FIELD(fu, u); GET_FIELD(fo, out);
  CHECK_SLOT(fu, u_slot); CHECK_LEVEL(fu, s->level); CHECK_LEVEL(fo, s->level);
  if (fu == fo) return set_error(EXAMG_EINVAL, "examg_apply: in-place application");
  LocGuard lg(c, fu->loc);
  int rc = do_exchange(c, *fu, u_slot < 0 ? fu->cur : u_slot);
  if (rc) return rc;
  fo->halo_depth = 0;
  return launch_apply(c, *s, field_ptr(c, fu, u_slot), field_ptr(c, fo, -1));
}

int examg_residual(examg_ctx *c, int A, int u, int u_slot, int rhs, int res, double *norm2_out) {
  return residual_op(c, A, u, u_slot, rhs, res, norm2_out, 0);
}

int examg_residual_restrict(examg_ctx *c, int A, int u, int rhs, int res, int rhs_coarse, double scale) {
  int rc = residual_restrict_op(c, A, u, rhs, rhs_coarse, scale);
  if (rc != 1) return rc;
  if ((rc = residual_op(c, A, u, -1, rhs, res, nullptr, 1))) return rc;
  return examg_restrict(c, res, rhs_coarse, scale);
}

int examg_smooth_jacobi(examg_ctx *c, int A, int u, int rhs, double omega, int nsweeps) {
  GET_STENCIL(s, A); GET_FIELD(fu, u); GET_FIELD(fr, rhs);
  CHECK_LEVEL(fu, s->level); CHECK_LEVEL(fr, s->level);
  if (fu->nslots < 2) return set_error(EXAMG_EINVAL, "examg_smooth_jacobi: Solution needs 2 slots");
  if (fu->loc != fr->loc) return set_error(EXAMG_EINVAL, "examg_smooth_jacobi: fields of different localisation");
  LocGuard lg(c, fu->loc);
  if (fused_jac_usable(c, *s, *fu)) {  // up to three sweeps in one pass over HBM (examg_jac7.cuh)
    for (int it = 0; it < nsweeps;) {
      // `plan` sweeps need that many ghost layers of Solution and one less of RHS (redundant updates there)
      const int plan = jac_fused_plan(nsweeps - it);
      int rc, done = 0;
      if (c->world > 1 && plan > 1 && fr->halo_depth < plan - 1) {
        Field *fs[2] = {fu, fr};
        const int slots[2] = {fu->cur, fr->cur}, depths[2] = {plan, plan - 1};
        rc = do_exchange_many(c, fs, slots, depths, 2);
      } else {
        rc = do_exchange(c, *fu, fu->cur, plan);
      }
      if (rc) return rc;
      if ((rc = launch_jac_fused(c, *s, omega, *fu, field_ptr(c, fr, -1), nsweeps - it, &done))) return rc;
      it += done;
      fu->halo_depth = 0;
    }
    return EXAMG_OK;
  }
  int it0 = 0;
  if (fused_jacv_usable(c, *s, *fu)) {  // stencil-field coefficients: pairs of sweeps read them once (examg_jacv.cuh)
    for (; it0 + 2 <= nsweeps; it0 += 2) {
      int rc = EXAMG_OK;
      if (c->world > 1) {
        // sweep 1 is computed redundantly on the first ghost layer: two layers of Solution, one of RHS and one of
        // the coefficients (static: exchanged once unless examg_field_eval already filled the layer)
        Field &cf = c->fields[s->coeff_field];
        if (cf.halo_depth < 1 && (rc = do_exchange(c, cf, 0, 1))) return rc;
        if (fr->halo_depth < 1) {
          Field *fs[2] = {fu, fr};
          const int slots[2] = {fu->cur, fr->cur}, depths[2] = {2, 1};
          rc = do_exchange_many(c, fs, slots, depths, 2);
        } else {
          rc = do_exchange(c, *fu, fu->cur, 2);
        }
        if (rc) return rc;
      }
      rc = launch_jacv_fused(c, *s, omega, *fu, field_ptr(c, fr, -1));
      if (rc) return rc;
      fu->halo_depth = 0;
    }
  }
  for (int it = it0; it < nsweeps; ++it) {
    int rc = do_exchange(c, *fu, fu->cur);
    if (rc) return rc;
    const int nxt = (fu->cur + 1) % fu->nslots;
    rc = launch_jacobi(c, *s, omega, field_ptr(c, fu, fu->cur), field_ptr(c, fr, -1), field_ptr(c, fu, nxt));
    if (rc) return rc;
    fu->cur = nxt;  // advance Solution
    fu->halo_depth = 0;
    // cell-localised fields: the ghost cells are functions of the inner cells (Dirichlet 2g - u, Neumann u), so the
    // next sweep needs `apply bc` (the loop bodies of Testing/CellBased/*.exa4 Smoother); node fields keep their
    // Dirichlet nodes (never written by the loop)
    if (fu->loc == 1 && (rc = examg_apply_bc(c, u, -1))) return rc;
  }
  return EXAMG_OK;
}

int examg_smooth_rb_color(examg_ctx *c, int A, int u, int rhs, double omega, int color) {
  GET_STENCIL(s, A); GET_FIELD(fu, u); GET_FIELD(fr, rhs);
  CHECK_LEVEL(fu, s->level); CHECK_LEVEL(fr, s->level);
  if (color != 0 && color != 1) return set_error(EXAMG_EINVAL, "examg_smooth_rb_color: colour %d", color);
  if (fu->loc != fr->loc) return set_error(EXAMG_EINVAL, "examg_smooth_rb_color: fields of different localisation");
  LocGuard lg(c, fu->loc);
  for (int k = 0; k < s->nst; ++k)
    if (k != s->centre && ((s->off[k][0] + s->off[k][1] + s->off[k][2]) & 1) == 0)
      return set_error(EXAMG_EUNSUP, "examg_smooth_rb: two colours do not decouple this stencil (entry [%d,%d,%d] "
                       "couples points of one colour); use a star stencil or Jacobi", s->off[k][0], s->off[k][1], s->off[k][2]);
  int rc = do_exchange(c, *fu, fu->cur);
  if (rc) return rc;
  fu->halo_depth = 0;
  if ((rc = launch_rb(c, *s, omega, field_ptr(c, fu, -1), field_ptr(c, fr, -1), color))) return rc;
  // { communicate; update; apply bc } per colour: ghost cells of a cell-localised field follow the inner cells
  return fu->loc == 1 ? examg_apply_bc(c, u, -1) : EXAMG_OK;
}

int examg_smooth_colored(examg_ctx *c, int A, int u, int rhs, double omega, int ncol, int nsweeps) {
  GET_STENCIL(s, A); GET_FIELD(fu, u); GET_FIELD(fr, rhs);
  CHECK_LEVEL(fu, s->level); CHECK_LEVEL(fr, s->level);
  if (ncol != 2 && ncol != 3) return set_error(EXAMG_EINVAL, "examg_smooth_colored: %d colours per dimension (2 or 3)", ncol);
  if (fu->loc != fr->loc) return set_error(EXAMG_EINVAL, "examg_smooth_colored: fields of different localisation");
  for (int k = 0; k < s->nst; ++k)
    for (int d = 0; d < 3; ++d)
      if (s->off[k][d] < -1 || s->off[k][d] > 1)
        return set_error(EXAMG_EUNSUP, "examg_smooth_colored: compact stencils only (entry [%d,%d,%d])", s->off[k][0], s->off[k][1], s->off[k][2]);
  LocGuard lg(c, fu->loc);
  const int nz = c->cfg.dim == 3 ? ncol : 1;
  for (int it = 0; it < nsweeps; ++it)
    for (int cz = 0; cz < nz; ++cz)
      for (int cy = 0; cy < ncol; ++cy)
        for (int cx = 0; cx < ncol; ++cx) {
          int rc = do_exchange(c, *fu, fu->cur);
          if (rc) return rc;
          fu->halo_depth = 0;
          if ((rc = launch_color_generic(c, *s, omega, field_ptr(c, fu, -1), field_ptr(c, fr, -1), ncol, cx, cy, cz))) return rc;
          if ((rc = examg_apply_bc(c, u, -1))) return rc;
        }
  return EXAMG_OK;
}

// nsweeps red-black sweeps through the fused kernel; fc != nullptr: Solution += Prolongation * fc rides on the first pass
static int rbgs_fused(examg_ctx *c, Stencil *s, Field *fu, Field *fr, double omega, int nsweeps, Field *fc) {
  for (int it = 0; it < nsweeps;) {
    // one launch does `plan` sweeps = 2*plan half-sweeps: that many ghost layers of Solution and one
    // less of RHS (the redundant updates in the ghost layers read RHS there); the correction of the ghost layers
    // needs plan + 1 layers of Solution@coarser (fine layer -2 plan sits on coarse layer -plan, odd layers between two)
    const int plan = rb_fused_plan(nsweeps - it);
    int rc, done = 0;
    Field *fs[3];
    int slots[3], depths[3], nf = 0;
    fs[nf] = fu; slots[nf] = fu->cur; depths[nf++] = 2 * plan;
    if (c->world > 1 && fr->halo_depth < 2 * plan - 1) {  // Solution and RHS in one handshake
      fs[nf] = fr; slots[nf] = fr->cur; depths[nf++] = 2 * plan - 1;
    }
    if (c->world > 1 && fc && it == 0 && fc->halo_depth < plan + 1) {
      fs[nf] = fc; slots[nf] = fc->cur; depths[nf++] = plan + 1;
    }
    if ((rc = do_exchange_many(c, fs, slots, depths, nf))) return rc;
    const double *uc = fc && it == 0 ? field_ptr(c, fc, -1) : nullptr;
    if ((rc = launch_rb_fused(c, *s, omega, *fu, field_ptr(c, fr, -1), nsweeps - it, &done, uc))) return rc;
    it += done;
    fu->halo_depth = 0;
  }
  return EXAMG_OK;
}

int examg_prolong_add_smooth_rbgs(examg_ctx *c, int A, int u_coarse, int u, int rhs, double omega, int nsweeps) {
  {
    GET_STENCIL(s, A); GET_FIELD(fu, u); GET_FIELD(fr, rhs); GET_FIELD(fc, u_coarse);
    CHECK_LEVEL(fu, s->level); CHECK_LEVEL(fr, s->level); CHECK_LEVEL(fc, s->level - 1);
    if (nsweeps >= 1 && fused_prolong_enabled() && fused_rb_usable(c, *s, *fu) && fc->loc == 0 && fc->ncomp == 1)
      return rbgs_fused(c, s, fu, fr, omega, nsweeps, fc);
  }
  int rc = examg_prolong_add(c, u_coarse, -1, u, -1);
  if (rc) return rc;
  return examg_smooth_rbgs(c, A, u, rhs, omega, nsweeps);
}

int examg_smooth_rbgs(examg_ctx *c, int A, int u, int rhs, double omega, int nsweeps) {
  {
    GET_STENCIL(s, A); GET_FIELD(fu, u); GET_FIELD(fr, rhs);
    CHECK_LEVEL(fu, s->level); CHECK_LEVEL(fr, s->level);
    // both colours of one or two sweeps in one pass over HBM
    if (fused_rb_usable(c, *s, *fu)) return rbgs_fused(c, s, fu, fr, omega, nsweeps, nullptr);
    if (fused_rb2d_usable(c, *s, *fu) && fr->loc == 0) {  // 2D 5-point: up to three sweeps per launch (examg_fast5.cu)
      for (int it = 0; it < nsweeps;) {
        int done = 0;
        int rc = launch_rb2d_fused(c, *s, omega, *fu, field_ptr(c, fr, -1), nsweeps - it, &done);
        if (rc) return rc;
        it += done;
        fu->halo_depth = 0;
      }
      return EXAMG_OK;
    }
  }
  for (int it = 0; it < nsweeps; ++it)
    for (int color = 0; color < 2; ++color) {
      int rc = examg_smooth_rb_color(c, A, u, rhs, omega, color);
      if (rc) return rc;
    }
  return EXAMG_OK;
}

int examg_restrict(examg_ctx *c, int res_fine, int rhs_coarse, double scale) {
  GET_FIELD(ff, res_fine); GET_FIELD(fc, rhs_coarse);
  if (fc->level != ff->level - 1) return set_error(EXAMG_EINVAL, "examg_restrict: levels %d -> %d", ff->level, fc->level);
  if (ff->loc != 0 || fc->loc != 0) return set_error(EXAMG_EUNSUP, "examg_restrict: node-localised fields (cell fields: examg_restrict_cell)");
  int rc = ff->halo_depth >= 1 ? EXAMG_OK : do_exchange(c, *ff, ff->cur);  // communicate Residual
  if (rc) return rc;
  fc->halo_depth = 0;
  return launch_restrict(c, ff->level, scale, field_ptr(c, ff, -1), field_ptr(c, fc, -1));
}

int examg_prolong_add(examg_ctx *c, int u_coarse, int uc_slot, int u_fine, int uf_slot) {
  GET_FIELD(fc, u_coarse); GET_FIELD(ff, u_fine);
  CHECK_SLOT(fc, uc_slot); CHECK_SLOT(ff, uf_slot);
  if (fc->level != ff->level - 1) return set_error(EXAMG_EINVAL, "examg_prolong_add: levels %d -> %d", fc->level, ff->level);
  if (ff->loc != 0 || fc->loc != 0) return set_error(EXAMG_EUNSUP, "examg_prolong_add: node-localised fields (cell fields: examg_prolong_add_cell)");
  ff->halo_depth = 0;
  return launch_prolong(c, ff->level, field_ptr(c, fc, uc_slot), field_ptr(c, ff, uf_slot));
}

int examg_dot(examg_ctx *c, int a, int a_slot, int b, int b_slot, double *out) {
  GET_FIELD(fa, a); GET_FIELD(fb, b);
  CHECK_SLOT(fa, a_slot); CHECK_SLOT(fb, b_slot); CHECK_LEVEL(fb, fa->level);
  if (!out) return set_error(EXAMG_EINVAL, "examg_dot: null output");
  if (fa->loc != fb->loc) return set_error(EXAMG_EINVAL, "examg_dot: fields of different localisation");
  LocGuard lg(c, fa->loc);
  int rc = launch_dot(c, fa->level, field_ptr(c, fa, a_slot), field_ptr(c, fb, b_slot), 0);
  if (rc) return rc;
  return finish_sum(c, 0, out);
}

int examg_max_abs_diff(examg_ctx *c, int a, int a_slot, int b, int b_slot, double *out) {
  GET_FIELD(fa, a); GET_FIELD(fb, b);
  CHECK_SLOT(fa, a_slot); CHECK_SLOT(fb, b_slot); CHECK_LEVEL(fb, fa->level);
  if (!out) return set_error(EXAMG_EINVAL, "examg_max_abs_diff: null output");
  if (fa->loc != fb->loc) return set_error(EXAMG_EINVAL, "examg_max_abs_diff: fields of different localisation");
  LocGuard lg(c, fa->loc);
  int rc = launch_maxdiff(c, fa->level, field_ptr(c, fa, a_slot), field_ptr(c, fb, b_slot), 1);
  if (rc) return rc;
  return finish_max(c, 1, out);  // global over all blocks, as examg_dot
}

int examg_set(examg_ctx *c, int f, int slot, double value) {
  GET_FIELD(ff, f); CHECK_SLOT(ff, slot);
  ff->halo_depth = 0;
  LocGuard lg(c, ff->loc);
  return launch_stream_op(c, ff->level, STREAM_SET, value, nullptr, field_ptr(c, ff, slot));
}
int examg_copy(examg_ctx *c, int src, int src_slot, int dst, int dst_slot) {
  GET_FIELD(fs, src); GET_FIELD(fd, dst);
  CHECK_SLOT(fs, src_slot); CHECK_SLOT(fd, dst_slot); CHECK_LEVEL(fd, fs->level);
  if (fs->loc != fd->loc) return set_error(EXAMG_EINVAL, "examg_copy: fields of different localisation");
  fd->halo_depth = 0;
  LocGuard lg(c, fs->loc);
  return launch_stream_op(c, fs->level, STREAM_COPY, 0.0, field_ptr(c, fs, src_slot), field_ptr(c, fd, dst_slot));
}
int examg_axpy(examg_ctx *c, double alpha, int x, int x_slot, int y, int y_slot) {
  GET_FIELD(fx, x); GET_FIELD(fy, y);
  CHECK_SLOT(fx, x_slot); CHECK_SLOT(fy, y_slot); CHECK_LEVEL(fy, fx->level);
  if (fx->loc != fy->loc) return set_error(EXAMG_EINVAL, "examg_axpy: fields of different localisation");
  if (fx->loc != fy->loc) return set_error(EXAMG_EINVAL, "examg_xpay: fields of different localisation");
  fy->halo_depth = 0;
  LocGuard lg(c, fx->loc);
  return launch_stream_op(c, fx->level, STREAM_AXPY, alpha, field_ptr(c, fx, x_slot), field_ptr(c, fy, y_slot));
}
int examg_xpay(examg_ctx *c, int x, int x_slot, double beta, int y, int y_slot) {
  GET_FIELD(fx, x); GET_FIELD(fy, y);
  CHECK_SLOT(fx, x_slot); CHECK_SLOT(fy, y_slot); CHECK_LEVEL(fy, fx->level);
  fy->halo_depth = 0;
  LocGuard lg(c, fx->loc);
  return launch_stream_op(c, fx->level, STREAM_XPAY, beta, field_ptr(c, fx, x_slot), field_ptr(c, fy, y_slot));
}

int examg_exchange(examg_ctx *c, int field, int slot) {
  GET_FIELD(f, field); CHECK_SLOT(f, slot);
  return do_exchange(c, *f, slot < 0 ? f->cur : slot);
}

int examg_exchange_depth(examg_ctx *c, int field, int slot, int depth) {
  GET_FIELD(f, field); CHECK_SLOT(f, slot);
  return do_exchange(c, *f, slot < 0 ? f->cur : slot, depth);
}

int examg_cr_coarse(examg_ctx *c, int A, int u, int rhs, int res, const int tmp[3], int max_it, double rel_tol,
                    int *iters_out) {
  GET_STENCIL(s, A); GET_FIELD(fu, u); GET_FIELD(fr, rhs); GET_FIELD(fres, res);
  CHECK_LEVEL(fu, s->level); CHECK_LEVEL(fr, s->level); CHECK_LEVEL(fres, s->level);
  if (!tmp) return set_error(EXAMG_EINVAL, "examg_cr_coarse: null tmp");
  if (fu->loc != 0) return set_error(EXAMG_EUNSUP, "examg_cr_coarse: node-localised fields");
  double *tp[3];
  for (int k = 0; k < 3; ++k) {
    Field *ft = get_field(c, tmp[k]);
    if (!ft || ft->level != s->level || ft->loc != 0) return set_error(EXAMG_EINVAL, "examg_cr_coarse: tmp[%d]", k);
    for (int j = 0; j < k; ++j)
      if (tmp[j] == tmp[k]) return set_error(EXAMG_EINVAL, "examg_cr_coarse: tmp fields must be distinct");
    tp[k] = field_ptr(c, ft, -1);
    ft->halo_depth = 0;
  }
  fu->halo_depth = 0;
  fres->halo_depth = 0;
  int rc = launch_cr(c, *s, field_ptr(c, fu, -1), field_ptr(c, fr, -1), field_ptr(c, fres, -1), tp, max_it, rel_tol);
  if (rc || !iters_out) return rc;
  EXAMG_CUDA(cudaMemcpyAsync(c->int_host, c->int_out, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  EXAMG_SYNC(c);
  *iters_out = c->int_host[0];
  return EXAMG_OK;
}

int examg_minres_coarse(examg_ctx *c, int A, int u, int rhs, int res, const int tmp[7], int max_it, double rel_tol,
                        int *iters_out) {
  GET_STENCIL(s, A); GET_FIELD(fu, u); GET_FIELD(fr, rhs); GET_FIELD(fres, res);
  CHECK_LEVEL(fu, s->level); CHECK_LEVEL(fr, s->level); CHECK_LEVEL(fres, s->level);
  if (!tmp) return set_error(EXAMG_EINVAL, "examg_minres_coarse: null tmp");
  if (fu->loc != 0) return set_error(EXAMG_EUNSUP, "examg_minres_coarse: node-localised fields");
  double *tp[7];
  for (int k = 0; k < 7; ++k) {
    Field *ft = get_field(c, tmp[k]);
    if (!ft || ft->level != s->level || ft->loc != 0) return set_error(EXAMG_EINVAL, "examg_minres_coarse: tmp[%d]", k);
    for (int j = 0; j < k; ++j)
      if (tmp[j] == tmp[k]) return set_error(EXAMG_EINVAL, "examg_minres_coarse: tmp fields must be distinct");
    tp[k] = field_ptr(c, ft, -1);
    ft->halo_depth = 0;
  }
  fu->halo_depth = 0;
  fres->halo_depth = 0;
  int rc = launch_minres(c, *s, field_ptr(c, fu, -1), field_ptr(c, fr, -1), field_ptr(c, fres, -1), tp, max_it, rel_tol);
  if (rc || !iters_out) return rc;
  EXAMG_CUDA(cudaMemcpyAsync(c->int_host, c->int_out, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  EXAMG_SYNC(c);
  *iters_out = c->int_host[0];
  return EXAMG_OK;
}

int examg_bicgstab_coarse(examg_ctx *c, int A, int u, int rhs, int res, const int tmp[5], int max_it, double rel_tol,
                          int *iters_out) {
  GET_STENCIL(s, A); GET_FIELD(fu, u); GET_FIELD(fr, rhs); GET_FIELD(fres, res);
  CHECK_LEVEL(fu, s->level); CHECK_LEVEL(fr, s->level); CHECK_LEVEL(fres, s->level);
  if (!tmp) return set_error(EXAMG_EINVAL, "examg_bicgstab_coarse: null tmp");
  if (fu->loc != 0) return set_error(EXAMG_EUNSUP, "examg_bicgstab_coarse: node-localised fields");
  double *tp[5];
  for (int k = 0; k < 5; ++k) {
    Field *ft = get_field(c, tmp[k]);
    if (!ft || ft->level != s->level || ft->loc != 0) return set_error(EXAMG_EINVAL, "examg_bicgstab_coarse: tmp[%d]", k);
    for (int j = 0; j < k; ++j)
      if (tmp[j] == tmp[k]) return set_error(EXAMG_EINVAL, "examg_bicgstab_coarse: tmp fields must be distinct");
    tp[k] = field_ptr(c, ft, -1);
    ft->halo_depth = 0;
  }
  fu->halo_depth = 0;
  fres->halo_depth = 0;
  int rc = launch_bicgstab(c, *s, field_ptr(c, fu, -1), field_ptr(c, fr, -1), field_ptr(c, fres, -1), tp, max_it, rel_tol);
  if (rc || !iters_out) return rc;
  EXAMG_CUDA(cudaMemcpyAsync(c->int_host, c->int_out, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  EXAMG_SYNC(c);
  *iters_out = c->int_host[0];
  return EXAMG_OK;
}

int examg_smoother_coarse(examg_ctx *c, int A, int u, int rhs, int res, int smoother, double omega, int tmp, int max_it,
                          double rel_tol, int *iters_out) {
  GET_STENCIL(s, A); GET_FIELD(fu, u); GET_FIELD(fr, rhs); GET_FIELD(fres, res);
  CHECK_LEVEL(fu, s->level); CHECK_LEVEL(fr, s->level); CHECK_LEVEL(fres, s->level);
  if (fu->loc != 0 || fr->loc != 0 || fres->loc != 0) return set_error(EXAMG_EUNSUP, "examg_smoother_coarse: node-localised fields");
  if (u == rhs || u == res || rhs == res) return set_error(EXAMG_EINVAL, "examg_smoother_coarse: u, rhs and res must be distinct fields");
  if (smoother != EXAMG_SMOOTH_JACOBI && smoother != EXAMG_SMOOTH_RBGS) return set_error(EXAMG_EINVAL, "examg_smoother_coarse: smoother %d", smoother);
  double *other = nullptr;
  if (smoother == EXAMG_SMOOTH_JACOBI) {
    // the second buffer: Solution<next> of a two-slot field, else the work field `tmp`
    if (fu->nslots >= 2) {
      other = field_ptr(c, fu, (fu->cur + 1) % fu->nslots);
    } else {
      Field *ft = get_field(c, tmp);
      if (!ft || ft->level != s->level || ft->loc != 0 || tmp == u || tmp == rhs || tmp == res)
        return set_error(EXAMG_EINVAL, "examg_smoother_coarse: Jacobi on a one-slot Solution needs a work field (tmp)");
      other = field_ptr(c, ft, -1);
      ft->halo_depth = 0;
    }
  } else {
    for (int k = 0; k < s->nst; ++k)
      if (k != s->centre && ((s->off[k][0] + s->off[k][1] + s->off[k][2]) & 1) == 0)
        return set_error(EXAMG_EUNSUP, "examg_smoother_coarse: two colours do not decouple this stencil; use Jacobi");
  }
  fu->halo_depth = fres->halo_depth = 0;
  int rc = launch_smoother_solve(c, *s, field_ptr(c, fu, -1), field_ptr(c, fr, -1), field_ptr(c, fres, -1), other, smoother,
                                 omega, max_it, rel_tol);
  if (rc || !iters_out) return rc;
  EXAMG_CUDA(cudaMemcpyAsync(c->int_host, c->int_out, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  EXAMG_SYNC(c);
  *iters_out = c->int_host[0];
  return EXAMG_OK;
}

int examg_cg_coarse(examg_ctx *c, int A, int u, int rhs, int res, int p, int ap, int max_it, double rel_tol,
                    int *iters_out) {
  GET_STENCIL(s, A); GET_FIELD(fu, u); GET_FIELD(fr, rhs); GET_FIELD(fres, res); GET_FIELD(fp, p); GET_FIELD(fap, ap);
  CHECK_LEVEL(fu, s->level); CHECK_LEVEL(fr, s->level); CHECK_LEVEL(fres, s->level);
  CHECK_LEVEL(fp, s->level); CHECK_LEVEL(fap, s->level);
  if (fu->loc != 0) return set_error(EXAMG_EUNSUP, "examg_cg_coarse: node-localised fields");
  {
    const int hs[5] = {u, rhs, res, p, ap};
    for (int k = 0; k < 5; ++k)
      for (int j = 0; j < k; ++j)
        if (hs[j] == hs[k]) return set_error(EXAMG_EINVAL, "examg_cg_coarse: u, rhs, res, p and ap must be distinct fields");
  }
  fu->halo_depth = fres->halo_depth = fp->halo_depth = fap->halo_depth = 0;
  int rc = launch_cg(c, *s, field_ptr(c, fu, -1), field_ptr(c, fr, -1), field_ptr(c, fres, -1), field_ptr(c, fp, -1),
                     field_ptr(c, fap, -1), max_it, rel_tol);
  if (rc || !iters_out) return rc;
  EXAMG_CUDA(cudaMemcpyAsync(c->int_host, c->int_out, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  EXAMG_SYNC(c);
  *iters_out = c->int_host[0];
  return EXAMG_OK;
}

}  // extern "C"
