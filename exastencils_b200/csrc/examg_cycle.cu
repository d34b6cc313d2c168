// examg_cycle.cu -- Solve@finest / mgCycle@lvl as a library object: the level
// loop and coarse-grid handling of the generated solver, without host round-trips.
//
// Reference: Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4:121-150 (Solve),
// :152-201 (mgCycle@coarsest), :203-249 (mgCycle@(all but coarsest)); generator
// C/s/e/solver/l3/L3_SolverForEquation.scala:259-395.  The reference prints one
// C++ function per level and synchronises the device after every kernel; here one
// mgCycle@finest is captured once into a CUDA graph and replayed.
//
// `apply bc` after each smoother colour / residual of the reference cycle is not
// replayed: loops never write physical-boundary nodes (iteration offsets,
// IR_LoopOverPointsInOneFragment.scala:86-101), so Dirichlet values set at setup
// stay valid; examg_apply_bc remains available as its own entry point.
#include <cmath>
#include <cstdlib>
#include <cctype>
#include <map>
#include <string>

#include "examg_internal.h"

using namespace examg;

struct examg_solver {
  examg_ctx *c = nullptr;
  examg_cycle_desc desc;
  std::vector<int> stencils;          // per level
  std::vector<int> sol, rhs, res;     // field handles per level
  std::vector<int> approx;            // gen_approx_* per level (solver_useFAS; -1 on the finest)
  int cg_p = -1, cg_ap = -1;
  int cgs_tmp[7] = {-1, -1, -1, -1, -1, -1, -1};  // work fields of the other coarse solvers (solver_cgs)
  int coarse_top = -1;                // mgCycle@level for level <= coarse_top is one single-CTA kernel
  // several blocks: mgCycle@agg_level runs on rank 0 on the agglomerated (global, single-block) grid
  int agg_level = -1;
  examg_ctx *agg_ctx = nullptr;       // rank 0 only
  examg_solver *agg = nullptr;
  // stencil fields on the agglomerated levels (rank 0): coefficient field of this context -> its twin on the
  // agglomerated grid, re-created there by replaying the examg_field_eval programs of the source
  struct AggCoef { int src, dst; long long gen; };
  std::vector<AggCoef> agg_coef;
  bool use_graph = true;
  struct Graph { cudaGraphExec_t exec; long long launches; std::vector<int> exit_slots; };
  std::map<std::vector<int>, Graph> graphs;  // keyed by the active slots of Solution on entry
};

namespace examg {
bool fused_rb_usable(examg_ctx *c, const Stencil &s, const Field &fu);
bool fused_rb2d_usable(examg_ctx *c, const Stencil &s, const Field &fu);
int ensure_shadow(examg_ctx *c, Field &fu);
// single-CTA coarse sub-cycle (examg_coarse.cu)
int coarse_cycle_top(examg_ctx *c, const examg_cycle_desc &d, const int *stencils, const int *sol, int max_top);
int launch_coarse_cycle(examg_ctx *c, const examg_cycle_desc &d, int top, const int *stencils, const int *sol,
                        const int *rhs, const int *res, int cg_p, int cg_ap);
}

// State of a Solution field as recorded in the graph keys: the active slot and which allocation
// (index into Field::orig) every slot and the shadow currently point at -- the fused sweeps
// exchange buffers instead of copying (RBGS: slot <-> shadow, Jacobi with an even number of sweeps:
// slot <-> slot).
static int field_state(const Field &f) {
  auto index_of = [&](const double *p) {
    for (int i = 0; i < 5; ++i)
      if (p && f.orig[i] == p) return i;
    return 0;
  };
  int code = f.cur, mult = 4;
  for (int i = 0; i < 4; ++i, mult *= 5) code += mult * index_of(f.slot[i]);
  return code + mult * index_of(f.shadow);
}

static void restore_state(Field &f, int state) {
  f.cur = state % 4;
  state /= 4;
  for (int i = 0; i < 4; ++i, state /= 5)
    if (f.slot[i]) f.slot[i] = f.orig[state % 5];
  if (f.shadow) f.shadow = f.orig[state % 5];
  f.swapped = f.shadow && f.shadow != f.orig[4];
}

static int lv(examg_solver *s, int level) { return level - s->c->cfg.min_level; }

static int smooth(examg_solver *s, int level, int n) {
  examg_ctx *c = s->c;
  const int A = s->stencils[lv(s, level)], u = s->sol[lv(s, level)], f = s->rhs[lv(s, level)];
  if (s->desc.smoother == EXAMG_SMOOTH_RBGS) return examg_smooth_rbgs(c, A, u, f, s->desc.omega, n);
  if (s->desc.smoother == EXAMG_SMOOTH_COLORED_2) return examg_smooth_colored(c, A, u, f, s->desc.omega, 2, n);
  if (s->desc.smoother == EXAMG_SMOOTH_COLORED_3) return examg_smooth_colored(c, A, u, f, s->desc.omega, 3, n);
  return examg_smooth_jacobi(c, A, u, f, s->desc.omega, n);
}

static int vcycle_rec(examg_solver *s, int level);

// mgCycle@agg_level of a multi-block run: RHS blocks -> rank 0, cycle on the global grid there,
// Solution blocks back.  Arithmetically the same global-grid cycle (duplicate nodes carry identical
// values, the oracle's formulation), executed by one GPU because these levels are latency bound.
static int agg_cycle(examg_solver *s) {
  examg_ctx *c = s->c;
  const int T = s->agg_level, l = lv(s, T);
  int rc = comm_agg_gather(c, c->fields[s->rhs[l]]);
  if (rc) return rc;
  if (c->cfg.rank == 0) {
    examg_solver *a = s->agg;
    examg_ctx *ac = s->agg_ctx;
    const int al = lv(a, T);
    if ((rc = copy_field_device(ac, a->rhs[al], -1, comm_agg_in(c), true))) return rc;
    if ((rc = examg_set(ac, a->sol[al], -1, 0.0))) return rc;
    if ((rc = vcycle_rec(a, T))) return rc;
    if ((rc = copy_field_device(ac, a->sol[al], -1, comm_agg_out(c), false))) return rc;
    c->launches += ac->launches;
    ac->launches = 0;
  }
  return comm_agg_scatter(c, c->fields[s->sol[l]]);
}

static int vcycle_rec(examg_solver *s, int level) {
  examg_ctx *c = s->c;
  const int l = lv(s, level);
  if (c->world > 1 && level <= s->agg_level)
    return set_error(EXAMG_EUNSUP, "mgCycle@%d of a multi-block run: levels <= %d live on the agglomerated grid", level, s->agg_level);
  if (level == c->cfg.min_level)
    switch (s->desc.coarse_solver) {
      case EXAMG_CGS_BICGSTAB:
        return examg_bicgstab_coarse(c, s->stencils[l], s->sol[l], s->rhs[l], s->res[l], s->cgs_tmp, s->desc.cg_max_it,
                                     s->desc.cg_rel_tol, nullptr);
      case EXAMG_CGS_CR:
        return examg_cr_coarse(c, s->stencils[l], s->sol[l], s->rhs[l], s->res[l], s->cgs_tmp, s->desc.cg_max_it,
                               s->desc.cg_rel_tol, nullptr);
      case EXAMG_CGS_SMOOTHER:  // Jacobi on a two-slot Solution uses Solution<next> as its second buffer
        return examg_smoother_coarse(c, s->stencils[l], s->sol[l], s->rhs[l], s->res[l], s->desc.smoother, s->desc.omega,
                                     s->cg_p, s->desc.cg_max_it, s->desc.cg_rel_tol, nullptr);
      case EXAMG_CGS_MINRES:
        return examg_minres_coarse(c, s->stencils[l], s->sol[l], s->rhs[l], s->res[l], s->cgs_tmp, s->desc.cg_max_it,
                                   s->desc.cg_rel_tol, nullptr);
      default:
        return examg_cg_coarse(c, s->stencils[l], s->sol[l], s->rhs[l], s->res[l], s->cg_p, s->cg_ap,
                               s->desc.cg_max_it, s->desc.cg_rel_tol, nullptr);
    }
  if (level <= s->coarse_top)
    return launch_coarse_cycle(c, s->desc, level, s->stencils.data(), s->sol.data(), s->rhs.data(), s->res.data(),
                               s->cg_p, s->cg_ap);
  int rc;
  if ((rc = smooth(s, level, s->desc.nu_pre))) return rc;
  if (s->desc.use_fas) {
    // Full approximation scheme (C/s/e/solver/l3/L3_SolverForEquation.scala:401-452): the coarse level carries the
    // restricted approximation instead of starting from zero, its right-hand side gains A@coarser * approximation and
    // only the difference is interpolated back.  Residual@coarser is the scratch of A * approximation.
    if ((rc = residual_op(c, s->stencils[l], s->sol[l], -1, s->rhs[l], s->res[l], nullptr, 1))) return rc;
    if ((rc = examg_restrict(c, s->res[l], s->rhs[l - 1], s->desc.restrict_scale))) return rc;
    if ((rc = examg_exchange(c, s->sol[l], -1))) return rc;
    if ((rc = examg_restrict(c, s->sol[l], s->approx[l - 1], 1.0))) return rc;
    if ((rc = examg_copy(c, s->approx[l - 1], -1, s->sol[l - 1], -1))) return rc;
    if ((rc = examg_apply(c, s->stencils[l - 1], s->sol[l - 1], -1, s->res[l - 1]))) return rc;
    if ((rc = examg_axpy(c, 1.0, s->res[l - 1], -1, s->rhs[l - 1], -1))) return rc;
    if ((rc = vcycle_rec(s, level - 1))) return rc;
    if ((rc = examg_axpy(c, -1.0, s->approx[l - 1], -1, s->sol[l - 1], -1))) return rc;
    if ((rc = examg_prolong_add(c, s->sol[l - 1], -1, s->sol[l], -1))) return rc;
    return smooth(s, level, s->desc.nu_post);
  }
  // several blocks: the residual of the first ghost layer is computed here too, so the restriction needs no exchange
  // (on large 3D constant-coefficient levels residual and restriction are one pass and Residual is not materialised)
  rc = residual_restrict_op(c, s->stencils[l], s->sol[l], s->rhs[l], s->rhs[l - 1], s->desc.restrict_scale);
  if (rc < 0) return rc;
  if (rc == 1) {
    if ((rc = residual_op(c, s->stencils[l], s->sol[l], -1, s->rhs[l], s->res[l], nullptr, 1))) return rc;
    if ((rc = examg_restrict(c, s->res[l], s->rhs[l - 1], s->desc.restrict_scale))) return rc;
  }
  if (c->world > 1 && level - 1 == s->agg_level) {
    if ((rc = agg_cycle(s))) return rc;
  } else {
    if ((rc = examg_set(c, s->sol[l - 1], -1, 0.0))) return rc;
    if ((rc = vcycle_rec(s, level - 1))) return rc;
  }
  // Correction + post-smoothing: with red-black sweeps the interpolation rides on the first smoothing pass
  if (s->desc.smoother == EXAMG_SMOOTH_RBGS && s->desc.nu_post >= 1)
    return examg_prolong_add_smooth_rbgs(c, s->stencils[l], s->sol[l - 1], s->sol[l], s->rhs[l], s->desc.omega, s->desc.nu_post);
  if ((rc = examg_prolong_add(c, s->sol[l - 1], -1, s->sol[l], -1))) return rc;
  return smooth(s, level, s->desc.nu_post);
}

// ---- `generate solver for ... with { ... }` options / solver_* Knowledge keys -> cycle descriptor -------------------

static std::string trim_token(const std::string &t) {
  size_t a = 0, b = t.size();
  while (a < b && (isspace((unsigned char)t[a]) || t[a] == '"' || t[a] == '\'')) ++a;
  while (b > a && (isspace((unsigned char)t[b - 1]) || t[b - 1] == '"' || t[b - 1] == '\'')) --b;
  return t.substr(a, b - a);
}

static std::string lower(std::string v) {
  for (char &ch : v) ch = (char)tolower((unsigned char)ch);
  return v;
}

extern "C" {

int examg_cycle_desc_from_options(const char *options, int dim, examg_cycle_desc *d, int *colors_per_dim) {
  if (!d) return set_error(EXAMG_EINVAL, "examg_cycle_desc_from_options: null descriptor");
  if (dim != 2 && dim != 3) return set_error(EXAMG_EINVAL, "examg_cycle_desc_from_options: dim %d", dim);
  // defaults: C/s/e/config/Knowledge.scala:203-262
  int num_pre = 3, num_post = 3, cgs_max = 512;
  double damping = 1.0, cgs_tol = 0.001;
  bool jacobi = false, use_fas = false;
  std::string coloring = "none", cgs = "cg";
  const std::string text = options ? options : "";
  size_t pos = 0;
  while (pos <= text.size()) {
    size_t end = text.find_first_of("\n;,", pos);
    if (end == std::string::npos) end = text.size();
    const std::string item = text.substr(pos, end - pos);
    pos = end + 1;
    const size_t eq = item.find('=');
    if (eq == std::string::npos) {
      if (!trim_token(item).empty()) return set_error(EXAMG_EINVAL, "examg_cycle_desc_from_options: '%s' is not key = value", trim_token(item).c_str());
      continue;
    }
    const std::string key = trim_token(item.substr(0, eq)), val = trim_token(item.substr(eq + 1)), lval = lower(val);
    auto as_bool = [&](bool *out) {
      if (lval == "true") { *out = true; return true; }
      if (lval == "false") { *out = false; return true; }
      return false;
    };
    char *rest = nullptr;
    if (key == "solver_smoother_numPre") { num_pre = (int)strtol(val.c_str(), &rest, 10); }
    else if (key == "solver_smoother_numPost") { num_post = (int)strtol(val.c_str(), &rest, 10); }
    else if (key == "solver_smoother_damping") { damping = strtod(val.c_str(), &rest); }
    else if (key == "solver_cgs_maxNumIts") { cgs_max = (int)strtol(val.c_str(), &rest, 10); }
    else if (key == "solver_cgs_targetResReduction") { cgs_tol = strtod(val.c_str(), &rest); }
    else if (key == "solver_smoother_jacobiType") { if (!as_bool(&jacobi)) return set_error(EXAMG_EINVAL, "%s = %s: true or false", key.c_str(), val.c_str()); }
    else if (key == "solver_smoother_coloring") { coloring = lval; }
    else if (key == "solver_cgs") { cgs = lval; }
    else if (key == "solver_useFAS") {
      bool fas = false;
      if (!as_bool(&fas)) return set_error(EXAMG_EINVAL, "%s = %s: true or false", key.c_str(), val.c_str());
      use_fas = fas;
    } else if (key == "solver_cgs_restart") {
      bool r = false;
      if (!as_bool(&r) || r) return set_error(EXAMG_EUNSUP, "solver_cgs_restart = %s: restarts of the coarse solver are not offered", val.c_str());
    } else if (key == "solver_cgs_absResThreshold") {
      if (strtod(val.c_str(), &rest) != 0.0) return set_error(EXAMG_EUNSUP, "solver_cgs_absResThreshold = %s: only the relative criterion", val.c_str());
    } else if (key == "solver_targetResReduction" || key == "solver_maxNumIts" || key == "solver_absResThreshold" ||
               key == "solver_silent" || key == "solver_generateDbgOutputs" || key == "solver_cgs_restartAfter" ||
               key == "solver_eliminateRHS" || key == "solver_overwriteSolutionFields" || key == "solver_splitLocalSolveLoops") {
      // outer Solve loop / code-generation details: host side
    } else {
      return set_error(EXAMG_EUNSUP, "examg_cycle_desc_from_options: unknown option %s", key.c_str());
    }
    if (rest && *rest != '\0') return set_error(EXAMG_EINVAL, "%s = %s: not a number", key.c_str(), val.c_str());
  }
  if (num_pre < 0 || num_post < 0 || cgs_max < 1) return set_error(EXAMG_EINVAL, "examg_cycle_desc_from_options: negative sweep or iteration count");
  int ncol = 0;
  if (jacobi) {
    if (coloring != "none") return set_error(EXAMG_EUNSUP, "solver_smoother_jacobiType = true with solver_smoother_coloring = %s", coloring.c_str());
    d->smoother = EXAMG_SMOOTH_JACOBI;
  } else if (coloring == "red-black" || coloring == "rb" || coloring == "2-way") {
    d->smoother = EXAMG_SMOOTH_RBGS;
  } else if ((coloring == "4-way" && dim == 2) || (coloring == "8-way" && dim == 3)) {
    d->smoother = EXAMG_SMOOTH_COLORED_2;
    ncol = 2;
  } else if ((coloring == "9-way" && dim == 2) || (coloring == "27-way" && dim == 3)) {
    d->smoother = EXAMG_SMOOTH_COLORED_3;
    ncol = 3;
  } else if (coloring == "none") {
    return set_error(EXAMG_EUNSUP, "solver_smoother_coloring = \"None\" with Gauss-Seidel: a lexicographic sweep depends on the "
                     "traversal order and the decomposition; choose a colouring or solver_smoother_jacobiType = true");
  } else {
    return set_error(EXAMG_EUNSUP, "solver_smoother_coloring = %s in %dD", coloring.c_str(), dim);
  }
  if (cgs == "cg") d->coarse_solver = EXAMG_CGS_CG;
  else if (cgs == "bicgstab") d->coarse_solver = EXAMG_CGS_BICGSTAB;
  else if (cgs == "conjugateresidual" || cgs == "cr") d->coarse_solver = EXAMG_CGS_CR;
  else if (cgs == "minres") d->coarse_solver = EXAMG_CGS_MINRES;
  else if (cgs == "smoother") d->coarse_solver = EXAMG_CGS_SMOOTHER;
  else return set_error(EXAMG_EUNSUP, "solver_cgs = %s", cgs.c_str());
  if (d->coarse_solver == EXAMG_CGS_SMOOTHER && ncol != 0)
    return set_error(EXAMG_EUNSUP, "solver_cgs = \"Smoother\" with an n^d-way colouring");
  d->omega = damping;
  d->nu_pre = num_pre;
  d->nu_post = num_post;
  d->cg_max_it = cgs_max;
  d->cg_rel_tol = cgs_tol;
  d->use_fas = use_fas ? 1 : 0;
  if (colors_per_dim) *colors_per_dim = ncol;
  return EXAMG_OK;
}

examg_solver *examg_solver_create(examg_ctx *c, const examg_cycle_desc *d) {
  if (!c || !d || !d->stencils) { set_error(EXAMG_EINVAL, "examg_solver_create: null argument"); return nullptr; }
  if (d->smoother < EXAMG_SMOOTH_JACOBI || d->smoother > EXAMG_SMOOTH_COLORED_3) {
    set_error(EXAMG_EINVAL, "examg_solver_create: smoother %d", d->smoother);
    return nullptr;
  }
  examg_solver *s = new examg_solver();
  s->c = c;
  s->desc = *d;
  const int nl = c->cfg.max_level - c->cfg.min_level + 1;
  for (int l = 0; l < nl; ++l) {
    Stencil *st = get_stencil(c, d->stencils[l]);
    if (!st || st->level != c->cfg.min_level + l) {
      set_error(EXAMG_EINVAL, "examg_solver_create: stencil for level %d missing", c->cfg.min_level + l);
      delete s;
      return nullptr;
    }
    s->stencils.push_back(d->stencils[l]);
  }
  s->desc.stencils = nullptr;
  const int slots = d->smoother == EXAMG_SMOOTH_JACOBI ? 2 : 1;
  for (int l = 0; l < nl; ++l) {
    const int level = c->cfg.min_level + l;
    int u = examg_field_create(c, "Solution", level, slots, 1);
    int f = examg_field_create(c, "RHS", level, 1, 1);
    int r = examg_field_create(c, "Residual", level, 1, 1);
    if (u < 0 || f < 0 || r < 0) { delete s; return nullptr; }
    // Field Solution< global, NodeWithComm, 0.0 >@(all but finest); the finest BC is set by the caller
    examg_field_set_bc(c, u, EXAMG_BC_DIRICHLET_CONST, 0.0);
    examg_field_set_bc(c, r, EXAMG_BC_DIRICHLET_CONST, 0.0);
    s->sol.push_back(u); s->rhs.push_back(f); s->res.push_back(r);
    int a = -1;
    if (d->use_fas && level != c->cfg.max_level) {
      if ((a = examg_field_create(c, "Approximation", level, 1, 1)) < 0) { delete s; return nullptr; }
      examg_field_set_bc(c, a, EXAMG_BC_DIRICHLET_CONST, 0.0);
    }
    s->approx.push_back(a);
  }
  if (d->use_fas && c->world > 1) {
    set_error(EXAMG_EUNSUP, "examg_solver_create: solver_useFAS on several blocks (the agglomerated levels start from zero)");
    delete s;
    return nullptr;
  }
  s->cg_p = examg_field_create(c, "cgTmp0", c->cfg.min_level, 1, 1);
  s->cg_ap = examg_field_create(c, "cgTmp1", c->cfg.min_level, 1, 1);
  if (s->cg_p < 0 || s->cg_ap < 0) { delete s; return nullptr; }
  if (d->coarse_solver < EXAMG_CGS_CG || d->coarse_solver > EXAMG_CGS_SMOOTHER) {
    set_error(EXAMG_EINVAL, "examg_solver_create: coarse_solver %d", d->coarse_solver);
    delete s;
    return nullptr;
  }
  {
    const int ntmp = d->coarse_solver == EXAMG_CGS_BICGSTAB ? 5 : d->coarse_solver == EXAMG_CGS_CR ? 3
                     : d->coarse_solver == EXAMG_CGS_MINRES ? 7 : 0;
    for (int k = 0; k < ntmp; ++k) {
      s->cgs_tmp[k] = examg_field_create(c, "cgsTmp", c->cfg.min_level, 1, 1);
      if (s->cgs_tmp[k] < 0) { delete s; return nullptr; }
    }
  }
  examg_field_set_bc(c, s->cg_p, EXAMG_BC_DIRICHLET_CONST, 0.0);
  // second buffers of the fused sweeps: now, so that they are part of what examg_comm_export publishes
  if (d->smoother == EXAMG_SMOOTH_RBGS)
    for (int l = 0; l < nl; ++l)
      if ((fused_rb_usable(c, *get_stencil(c, s->stencils[l]), c->fields[s->sol[l]]) ||
           fused_rb2d_usable(c, *get_stencil(c, s->stencils[l]), c->fields[s->sol[l]])) &&
          ensure_shadow(c, c->fields[s->sol[l]])) {
        examg_solver_destroy(s);
        return nullptr;
      }
  if (c->world > 1) {
    // agglomeration level: the largest one whose global grid has <= 256 cells per dimension and <= 2^21 cells in all
    // (measured at 8 x 512^3, blocks 1x2x4: gathering level 6 = 64 x 128 x 256 cells onto rank 0 gives 4.66 ms per
    // step against 4.72 ms with level 5 and 5.41 ms with level 7, profiles/r01_notes.md)
    int T = c->cfg.min_level;
    for (int lvl = c->cfg.min_level; lvl < c->cfg.max_level; ++lvl) {
      bool ok = true;
      long long cells = 1;
      for (int dd = 0; dd < c->cfg.dim; ++dd) {
        const long long g = (long long)c->cfg.block_cells[dd] * c->cfg.num_blocks[dd] << lvl;
        ok = ok && g <= 256;
        cells *= g;
      }
      if (ok && cells <= (1LL << 21)) T = lvl;
    }
    if (const char *e = getenv("EXAMG_AGG_LEVEL")) T = atoi(e);
    if (T < c->cfg.min_level || T >= c->cfg.max_level) {
      set_error(EXAMG_EUNSUP, "multi-block solver: agglomeration level %d outside [minLevel, maxLevel)", T);
      examg_solver_destroy(s);
      return nullptr;
    }
    s->agg_level = T;
    if (comm_agg_setup(c, T)) { examg_solver_destroy(s); return nullptr; }
    if (c->cfg.rank == 0) {
      examg_config ac = c->cfg;
      for (int dd = 0; dd < 3; ++dd) {
        ac.block_cells[dd] = c->cfg.block_cells[dd] * (dd < c->cfg.dim ? c->cfg.num_blocks[dd] : 1);
        ac.num_blocks[dd] = 1;
      }
      ac.rank = 0;
      ac.max_level = T;
      s->agg_ctx = examg_init(&ac);
      if (!s->agg_ctx) { examg_solver_destroy(s); return nullptr; }
      examg_set_stream(s->agg_ctx, c->stream);
      std::vector<int> ast;
      for (int lvl = c->cfg.min_level; lvl <= T; ++lvl) {
        const Stencil &st = *get_stencil(c, s->stencils[lvl - c->cfg.min_level]);
        if (st.coeff_field >= 0) {
          const Field &cf = c->fields[st.coeff_field];
          const int dst = examg_field_create(s->agg_ctx, cf.name.c_str(), lvl, 1, cf.ncomp);
          if (dst < 0) { examg_solver_destroy(s); return nullptr; }
          s->agg_coef.push_back({st.coeff_field, dst, -1});
          ast.push_back(examg_stencil_field(s->agg_ctx, lvl, st.nst, st.off, dst));
          continue;
        }
        ast.push_back(examg_stencil_const(s->agg_ctx, lvl, st.nst, st.off, st.coef));
      }
      examg_cycle_desc ad = *d;
      ad.stencils = ast.data();
      s->agg = examg_solver_create(s->agg_ctx, &ad);
      if (!s->agg) { examg_solver_destroy(s); return nullptr; }
      s->agg->use_graph = false;
    }
  }
  s->coarse_top = coarse_cycle_top(c, s->desc, s->stencils.data(), s->sol.data(), c->cfg.max_level);
  const char *ng = getenv("EXAMG_NO_GRAPH");
  s->use_graph = !(ng && ng[0] == '1');
  return s;
}

int examg_solver_destroy(examg_solver *s) {
  if (!s) return set_error(EXAMG_EINVAL, "examg_solver_destroy: null");
  cudaStreamSynchronize(s->c->stream);
  if (s->agg) examg_solver_destroy(s->agg);
  if (s->agg_ctx) {
    s->agg_ctx->stream = nullptr;  // borrowed from the parent context
    s->agg_ctx->own_stream = false;
    examg_finalize(s->agg_ctx);
  }
  for (auto &kv : s->graphs) cudaGraphExecDestroy(kv.second.exec);
  for (int h : s->sol) examg_field_destroy(s->c, h);
  for (int h : s->rhs) examg_field_destroy(s->c, h);
  for (int h : s->res) examg_field_destroy(s->c, h);
  examg_field_destroy(s->c, s->cg_p);
  examg_field_destroy(s->c, s->cg_ap);
  for (int k = 0; k < 7; ++k)
    if (s->cgs_tmp[k] >= 0) examg_field_destroy(s->c, s->cgs_tmp[k]);
  delete s;
  return EXAMG_OK;
}

int examg_solver_field(examg_solver *s, int level, int which) {
  if (!s) return set_error(EXAMG_EINVAL, "null solver");
  examg_ctx *c = s->c;
  if (level < c->cfg.min_level || level > c->cfg.max_level) return set_error(EXAMG_EINVAL, "examg_solver_field: level %d", level);
  switch (which) {
    case EXAMG_F_SOLUTION: return s->sol[lv(s, level)];
    case EXAMG_F_RHS: return s->rhs[lv(s, level)];
    case EXAMG_F_RESIDUAL: return s->res[lv(s, level)];
    case EXAMG_F_CGTMP0: if (level == c->cfg.min_level) return s->cg_p; break;
    case EXAMG_F_CGTMP1: if (level == c->cfg.min_level) return s->cg_ap; break;
    case EXAMG_F_APPROX: if (s->desc.use_fas && level != c->cfg.max_level) return s->approx[lv(s, level)]; break;
  }
  return set_error(EXAMG_EINVAL, "examg_solver_field: no field %d on level %d", which, level);
}

// State that decides which kernels a cycle launches and on which buffers: (active slot, buffer
// permutation) of every Solution field -- of the agglomerated solver too -- and how many ghost layers
// of every RHS field are valid (a cycle skips the RHS exchange when they already are).
static void cycle_state(examg_solver *s, std::vector<int> &key) {
  examg_ctx *c = s->c;
  for (int h : s->sol) key.push_back(field_state(c->fields[h]));
  for (int h : s->rhs) key.push_back(c->fields[h].halo_depth);
  for (int h : s->sol) key.push_back(c->fields[h].halo_depth);
  for (int h : s->res) key.push_back(c->fields[h].halo_depth);
  if (s->agg) cycle_state(s->agg, key);
}

static void restore_cycle_state(examg_solver *s, const std::vector<int> &key, size_t &pos) {
  examg_ctx *c = s->c;
  for (int h : s->sol) restore_state(c->fields[h], key[pos++]);
  for (int h : s->rhs) c->fields[h].halo_depth = key[pos++];
  for (int h : s->sol) c->fields[h].halo_depth = key[pos++];
  for (int h : s->res) c->fields[h].halo_depth = key[pos++];
  if (s->agg) restore_cycle_state(s->agg, key, pos);
}

int examg_solver_vcycle(examg_solver *s, int level) {
  if (!s) return set_error(EXAMG_EINVAL, "null solver");
  examg_ctx *c = s->c;
  if (level < 0) level = c->cfg.max_level;
  if (level < c->cfg.min_level || level > c->cfg.max_level) return set_error(EXAMG_EINVAL, "examg_solver_vcycle: level %d", level);
  // static operands: ghost layers of stencil-field coefficients are exchanged once, outside the captured cycle
  // (the fused stencil-field Jacobi reads them, examg_jacv.cuh); the twins on the agglomerated grid are brought
  // up to date
  if (s->desc.smoother == EXAMG_SMOOTH_RBGS || s->desc.smoother == EXAMG_SMOOTH_COLORED_2)  // split-row copies of stencil-field coefficients (k_rbv): allocate / refresh now
    for (int h : s->stencils) {
      Stencil *st = get_stencil(c, h);
      int rc = st ? rb_compact_prepare(c, *st) : EXAMG_OK;
      if (rc) return rc;
    }
  if (c->world > 1) {
    for (int l = lv(s, s->agg_level) + 1; l < (int)s->stencils.size(); ++l) {
      const Stencil *st = get_stencil(c, s->stencils[l]);
      if (st && st->coeff_field >= 0 && c->fields[st->coeff_field].halo_depth < 1) {
        int rc = do_exchange(c, c->fields[st->coeff_field], 0, 1);
        if (rc) return rc;
      }
    }
    for (auto &ac : s->agg_coef) {
      const Field &src = c->fields[ac.src];
      if (ac.gen == src.eval_gen) continue;
      std::vector<char> seen(src.ncomp, 0);
      for (const auto &e : src.evals) seen[e.first] = 1;
      for (char v : seen)
        if (!v)
          return set_error(EXAMG_EUNSUP, "multi-block solver: the stencil field %s@%d of an agglomerated level was not "
                           "produced by examg_field_eval (uploaded coefficients cannot be re-created on rank 0)",
                           src.name.c_str(), src.level);
      for (const auto &e : src.evals) {
        int rc = examg_field_eval(s->agg_ctx, ac.dst, 0, e.first, e.second.data(), (int)e.second.size());
        if (rc) return rc;
      }
      ac.gen = src.eval_gen;
    }
  }
  if (!s->use_graph || level != c->cfg.max_level) return vcycle_rec(s, level);

  std::vector<int> key;
  cycle_state(s, key);
  auto it = s->graphs.find(key);
  if (it == s->graphs.end()) {
    // capture mgCycle@finest for this state (allocations cannot be captured: the shadow buffers of the
    // fused sweeps were created with the solver)
    const long long l0 = c->launches;
    cudaGraph_t graph = nullptr;
    EXAMG_CUDA(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    int rc = vcycle_rec(s, level);
    cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
    // nothing ran: on any failure the host bookkeeping (active slots, buffer permutation, valid ghost layers, launch
    // count) goes back to the entry state
    auto undo = [&]() {
      size_t pos0 = 0;
      restore_cycle_state(s, key, pos0);
      c->launches = l0;
    };
    if (rc) { if (graph) cudaGraphDestroy(graph); undo(); return rc; }
    if (e != cudaSuccess) { undo(); return set_error(EXAMG_ECUDA, "graph capture of mgCycle failed: %s", cudaGetErrorString(e)); }
    examg_solver::Graph g;
    e = cudaGraphInstantiate(&g.exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) { undo(); return set_error(EXAMG_ECUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(e)); }
    g.launches = c->launches - l0;
    c->launches = l0;  // nothing ran yet
    cycle_state(s, g.exit_slots);
    size_t pos = 0;
    restore_cycle_state(s, key, pos);
    it = s->graphs.emplace(key, g).first;
  }
  EXAMG_CUDA(cudaGraphLaunch(it->second.exec, c->stream));
  c->launches += it->second.launches;
  size_t pos = 0;
  restore_cycle_state(s, it->second.exit_slots, pos);
  return EXAMG_OK;
}

int examg_solver_residual_norm(examg_solver *s, double *norm_out) {
  if (!s || !norm_out) return set_error(EXAMG_EINVAL, "examg_solver_residual_norm: null argument");
  const int l = lv(s, s->c->cfg.max_level);
  double n2 = 0.0;
  int rc = examg_residual(s->c, s->stencils[l], s->sol[l], -1, s->rhs[l], s->res[l], &n2);
  if (rc) return rc;
  *norm_out = std::sqrt(n2);
  return EXAMG_OK;
}

int examg_solver_cycles(examg_solver *s, int n, double *norm_out) {
  if (!s) return set_error(EXAMG_EINVAL, "null solver");
  for (int i = 0; i < n; ++i) {
    int rc = examg_solver_vcycle(s, -1);
    if (rc) return rc;
  }
  if (!norm_out) return EXAMG_OK;
  // the loop body of Solve@finest asks for the norm only; Residual@finest has no reader until the next cycle overwrites
  // it, so it is not written here (examg_solver_residual_norm does write it).  EXAMG_CYCLES_WRITE_RESIDUAL=1: write it
  static const bool write_res = getenv("EXAMG_CYCLES_WRITE_RESIDUAL") && getenv("EXAMG_CYCLES_WRITE_RESIDUAL")[0] == '1';
  if (write_res) return examg_solver_residual_norm(s, norm_out);
  const int l = lv(s, s->c->cfg.max_level);
  double n2 = 0.0;
  int rc = residual_norm2_op(s->c, s->stencils[l], s->sol[l], s->rhs[l], &n2);
  if (rc) return rc;
  *norm_out = std::sqrt(n2);
  return EXAMG_OK;
}

int examg_solver_last_cg_iters(examg_solver *s, int *iters_out) {
  if (!s || !iters_out) return set_error(EXAMG_EINVAL, "examg_solver_last_cg_iters: null argument");
  examg_ctx *c = s->c;
  if (c->world > 1) {  // the coarse solve ran on rank 0's agglomerated grid
    if (!s->agg) { *iters_out = 0; return EXAMG_OK; }
    return examg_solver_last_cg_iters(s->agg, iters_out);
  }
  EXAMG_CUDA(cudaMemcpyAsync(c->int_host, c->int_out, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  EXAMG_SYNC(c);
  *iters_out = c->int_host[0];
  return EXAMG_OK;
}

}  // extern "C"
