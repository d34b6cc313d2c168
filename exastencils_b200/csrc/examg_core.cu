// examg_core.cu -- lifecycle, fields, stencils, host<->device copies of libexamg.
//
// Replaces what the reference generator prints into Global/Global_initGlobals.cpp,
// Domain/Domain_initDomain.cpp and setupBuffers():
//   C/s/e/domain/ir/IR_InitGeneratedDomain.scala:40-104   (rank -> block, positions)
//   C/s/e/domain/ir/IR_ConnectFragments.scala:60-93        (neighbours, iteration offsets)
//   C/s/e/globals/ir/IR_AllocateDataFunction.scala:64-72   (allocation)
//   C/s/e/field/ir/IR_FieldLayout.scala:30-129             (layout; redesigned, see examg_internal.h)
#include <cstring>

#include <nvtx3/nvToolsExt.h>

#include "examg_internal.h"

using namespace examg;

static thread_local char g_err[512] = "";

int examg::set_error(int code, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

double *examg::dev_alloc(examg_ctx *c, size_t bytes) {
  constexpr size_t ARENA = 64u << 20, BIG = 1u << 20;
  bytes = (bytes + 255) & ~(size_t)255;
  if (bytes < BIG) {
    for (Block &b : c->blocks)
      if (b.arena && b.used + bytes <= b.bytes) {
        double *p = reinterpret_cast<double *>(b.base + b.used);
        b.used += bytes;
        return p;  // arena blocks are zero-filled when created and never recycled
      }
  }
  Block b;
  b.arena = bytes < BIG;
  b.bytes = b.arena ? ARENA : bytes;
  if (cudaMalloc(&b.base, b.bytes) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  cudaMemsetAsync(b.base, 0, b.bytes, c->stream);
  b.used = bytes;
  c->dev_bytes += (long long)b.bytes;
  c->blocks.push_back(b);
  return reinterpret_cast<double *>(b.base);
}

void examg::dev_free(examg_ctx *c, double *p) {
  for (size_t i = 0; i < c->blocks.size(); ++i)
    if (!c->blocks[i].arena && reinterpret_cast<double *>(c->blocks[i].base) == p) {
      cudaFree(c->blocks[i].base);
      c->dev_bytes -= (long long)c->blocks[i].bytes;
      c->blocks.erase(c->blocks.begin() + i);
      return;
    }
  // arena memory stays with the context until examg_finalize
}

const LevelGeom *examg::geom(examg_ctx *c, int level) {
  if (!c || level < c->cfg.min_level || level > c->cfg.max_level) return nullptr;
  return &c->levels[level - c->cfg.min_level];
}

Field *examg::get_field(examg_ctx *c, int h) {
  if (!c || h < 0 || h >= (int)c->fields.size() || !c->fields[h].alive) return nullptr;
  return &c->fields[h];
}

Stencil *examg::get_stencil(examg_ctx *c, int h) {
  if (!c || h < 0 || h >= (int)c->stencils.size()) return nullptr;
  return &c->stencils[h];
}

GridArgs examg::grid_args(const LevelGeom &g) {
  GridArgs a;
  for (int d = 0; d < 3; ++d) { a.n[d] = g.n[d]; a.lo[d] = g.lo[d]; a.hi[d] = g.hi[d]; a.rlo[d] = g.rlo[d]; a.rhi[d] = g.hi[d]; }
  a.pitch = g.pitch;
  a.plane = g.plane;
  a.par = g.par;
  return a;
}

GridArgs examg::grid_args_loc(examg_ctx *c, const LevelGeom &g) {
  GridArgs a = grid_args(g);
  if (c->loop_loc == 1)
    for (int d = 0; d < c->cfg.dim; ++d) { a.lo[d] = 0; a.hi[d] = g.n[d] - 1; a.rlo[d] = 0; a.rhi[d] = g.n[d] - 1; }
  return a;
}

examg::LocGuard::LocGuard(examg_ctx *ctx, int loc) : c(ctx) { c->loop_loc = loc; }
examg::LocGuard::~LocGuard() { c->loop_loc = 0; }

StencilArgs examg::stencil_args(const LevelGeom &g, const Stencil &s) {
  StencilArgs a;
  a.nst = s.nst;
  a.centre = s.centre;
  for (int k = 0; k < EXAMG_MAX_STENCIL; ++k) { a.lin[k] = 0; a.coef[k] = 0.0; }
  for (int k = 0; k < s.nst; ++k) {
    a.lin[k] = (long long)s.off[k][0] + (long long)g.pitch * s.off[k][1] + g.plane * s.off[k][2];
    a.coef[k] = s.coef[k];
  }
  return a;
}

extern "C" {

int examg_abi_version(void) { return EXAMG_ABI_VERSION; }
const char *examg_last_error(void) { return g_err; }

examg_ctx *examg_init(const examg_config *cfg) {
  if (!cfg || (cfg->dim != 2 && cfg->dim != 3) || cfg->min_level < 0 || cfg->max_level < cfg->min_level ||
      cfg->max_level > 14) {
    set_error(EXAMG_EINVAL, "examg_init: bad dimensionality or level range");
    return nullptr;
  }
  int world = 1;
  for (int d = 0; d < cfg->dim; ++d) {
    if (cfg->block_cells[d] < 1 || cfg->num_blocks[d] < 1) {
      set_error(EXAMG_EINVAL, "examg_init: block_cells/num_blocks must be >= 1");
      return nullptr;
    }
    world *= cfg->num_blocks[d];
  }
  if (cfg->rank < 0 || cfg->rank >= world) {
    // reference: "Invalid number of MPI processes" assert, IR_InitGeneratedDomain.scala:83-85
    set_error(EXAMG_EINVAL, "examg_init: rank %d outside the %d blocks of the domain", cfg->rank, world);
    return nullptr;
  }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev < 1) {
    set_error(EXAMG_ECUDA, "examg_init: no CUDA device (%s); libexamg has no CPU path",
              e == cudaSuccess ? "count 0" : cudaGetErrorString(e));
    return nullptr;
  }
  if (cfg->device < 0 || cfg->device >= ndev) {
    set_error(EXAMG_EINVAL, "examg_init: device %d of %d", cfg->device, ndev);
    return nullptr;
  }
  if (cudaSetDevice(cfg->device) != cudaSuccess) {
    set_error(EXAMG_ECUDA, "examg_init: cudaSetDevice failed");
    return nullptr;
  }
  examg_ctx *c = new examg_ctx();
  c->cfg = *cfg;
  c->device = cfg->device;
  c->world = world;
  int r = cfg->rank;
  for (int d = 0; d < 3; ++d) {
    int nbk = d < cfg->dim ? cfg->num_blocks[d] : 1;
    c->bidx[d] = r % nbk;
    r /= nbk;
    c->nb[d][0] = c->bidx[d] > 0;
    c->nb[d][1] = c->bidx[d] < nbk - 1;
    if (d < cfg->dim && cfg->periodic[d]) {
      if (nbk != 1) {
        set_error(EXAMG_EUNSUP, "examg_init: a periodic dimension takes one block so far (num_blocks[%d] = %d)", d, nbk);
        delete c;
        return nullptr;
      }
      c->nb[d][0] = c->nb[d][1] = true;  // iteration offsets 0 on both sides, C/s/e/domain/ir/IR_ConnectFragments.scala:60-93
      c->per[d] = true;
      c->periodic_any = true;
    }
  }
  for (int lvl = cfg->min_level; lvl <= cfg->max_level; ++lvl) {
    LevelGeom g;
    for (int d = 0; d < 3; ++d) {
      if (d < cfg->dim) {
        int cells = cfg->block_cells[d] << lvl;
        g.n[d] = cells + 1;
        g.lo[d] = c->nb[d][0] ? 0 : 1;
        g.hi[d] = c->nb[d][1] ? g.n[d] : g.n[d] - 1;
        g.rlo[d] = 1;
        g.goff[d] = c->bidx[d] * cells;
        g.h[d] = (cfg->domain_hi[d] - cfg->domain_lo[d]) / (double)(cells * cfg->num_blocks[d]);
      } else {
        g.n[d] = 1; g.lo[d] = 0; g.hi[d] = 1; g.rlo[d] = 0; g.goff[d] = 0; g.h[d] = 1.0;
      }
    }
    g.pitch = ((XPAD + g.n[0] + GHOST) + 15) / 16 * 16;
    g.ay = g.n[1] + 2 * GHOST;
    g.az = cfg->dim == 3 ? g.n[2] + 2 * GHOST : 1;
    g.plane = (long long)g.pitch * g.ay;
    g.vol = g.plane * g.az;
    g.origin = XPAD + (long long)g.pitch * GHOST + (cfg->dim == 3 ? g.plane * GHOST : 0);
    g.par = (g.goff[0] + g.goff[1] + g.goff[2]) & 1;
    c->levels.push_back(g);
  }
  if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
    set_error(EXAMG_ECUDA, "examg_init: cudaStreamCreate failed");
    delete c;
    return nullptr;
  }
  c->own_stream = true;
  bool ok = cudaMalloc(&c->partials, sizeof(dd) * PARTIAL_CAP) == cudaSuccess &&
            cudaMalloc(&c->red_out, sizeof(dd) * RED_SLOTS) == cudaSuccess &&
            cudaMallocHost(&c->red_host, sizeof(dd) * RED_SLOTS) == cudaSuccess &&
            cudaMalloc(&c->int_out, sizeof(int) * 8) == cudaSuccess &&
            cudaMallocHost(&c->int_host, sizeof(int) * 8) == cudaSuccess;
  if (!ok) {
    set_error(EXAMG_ENOMEM, "examg_init: scratch allocation failed");
    examg_finalize(c);
    return nullptr;
  }
  cudaMemset(c->red_out, 0, sizeof(dd) * RED_SLOTS);
  cudaMemset(c->int_out, 0, sizeof(int) * 8);
  c->dev_bytes += sizeof(dd) * (PARTIAL_CAP + RED_SLOTS);
  return c;
}

static int fold_timer(examg::Timer &t, bool wait);

int examg_finalize(examg_ctx *c) {
  if (!c) return set_error(EXAMG_EINVAL, "examg_finalize: null context");
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  comm_destroy(c);
  for (auto &b : c->blocks) cudaFree(b.base);
  c->blocks.clear();
  if (c->partials) cudaFree(c->partials);
  if (c->red_out) cudaFree(c->red_out);
  if (c->red_host) cudaFreeHost(c->red_host);
  if (c->int_out) cudaFree(c->int_out);
  if (c->int_host) cudaFreeHost(c->int_host);
  if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
  for (auto &kv : c->timers) {
    fold_timer(kv.second, false);
    if (kv.second.start) cudaEventDestroy(kv.second.start);
  }
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  for (int b = 0; b < 2; ++b) {
    if (c->up_h2d[b]) cudaEventDestroy(c->up_h2d[b]);
    if (c->up_d2d[b]) cudaEventDestroy(c->up_d2d[b]);
  }
  delete c;
  return EXAMG_OK;
}

int examg_set_stream(examg_ctx *c, void *s) {
  if (!c) return set_error(EXAMG_EINVAL, "null context");
  EXAMG_SYNC(c);
  if (s) {
    if (c->own_stream) cudaStreamDestroy(c->stream);
    c->stream = (cudaStream_t)s;
    c->own_stream = false;
  } else if (!c->own_stream) {
    EXAMG_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    c->own_stream = true;
  }
  return EXAMG_OK;
}

int examg_sync(examg_ctx *c) {
  if (!c) return set_error(EXAMG_EINVAL, "null context");
  EXAMG_SYNC(c);
  return EXAMG_OK;
}

int examg_level_dims(examg_ctx *c, int level, int n[3]) {
  const LevelGeom *g = geom(c, level);
  if (!g) return set_error(EXAMG_EINVAL, "examg_level_dims: level %d", level);
  for (int d = 0; d < 3; ++d) n[d] = g->n[d];
  return EXAMG_OK;
}

int examg_level_width(examg_ctx *c, int level, double h[3]) {
  const LevelGeom *g = geom(c, level);
  if (!g) return set_error(EXAMG_EINVAL, "examg_level_width: level %d", level);
  for (int d = 0; d < 3; ++d) h[d] = g->h[d];
  return EXAMG_OK;
}

// ---- timers: startTimer("x") / stopTimer("x") / getTotalTime("x") / printAllTimers() ------------
// (C/s/e/timing/ir/IR_GetTime.scala:33-63, IR_PrintAllTimers.scala:38-70).  The reference stops a host stopwatch
// after a device synchronisation; here both ends are CUDA events on the context stream, so timing neither
// synchronises nor measures host overhead.  Nested starts of one name count once, as in the reference.
static int fold_timer(Timer &t, bool wait) {
  size_t keep = 0;
  for (size_t i = 0; i < t.done.size(); ++i) {
    if (wait) cudaEventSynchronize(t.done[i].second);
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, t.done[i].first, t.done[i].second) == cudaSuccess) {
      t.total_ms += ms;
      cudaEventDestroy(t.done[i].first);
      cudaEventDestroy(t.done[i].second);
    } else {
      cudaGetLastError();  // not finished yet
      t.done[keep++] = t.done[i];
    }
  }
  t.done.resize(keep);
  return EXAMG_OK;
}

// benchmarkStart("name") / benchmarkStop("name") of the L4 programs: the reference lowers them to LIKWID / TALP region
// markers (C/s/e/util/ir/IR_ResolveBenchmarkFunctions.scala:63-120).  The B200 analogue is an NVTX range (nsys / ncu
// --nvtx pick it up; without a tool attached the calls cost nothing), plus a device timer of the same name.
int examg_benchmark_start(examg_ctx *c, const char *name) {
  if (!c || !name) return set_error(EXAMG_EINVAL, "examg_benchmark_start: null argument");
  nvtxRangePushA(name);
  c->nvtx_depth++;
  return examg_timer_start(c, name);
}

int examg_benchmark_stop(examg_ctx *c, const char *name) {
  if (!c || !name) return set_error(EXAMG_EINVAL, "examg_benchmark_stop: null argument");
  if (c->nvtx_depth <= 0) return set_error(EXAMG_ESTATE, "benchmarkStop(\"%s\") without benchmarkStart", name);
  int rc = examg_timer_stop(c, name);
  nvtxRangePop();
  c->nvtx_depth--;
  return rc;
}

int examg_timer_start(examg_ctx *c, const char *name) {
  if (!c || !name) return set_error(EXAMG_EINVAL, "examg_timer_start: null argument");
  Timer &t = c->timers[name];
  if (t.depth++ > 0) return EXAMG_OK;
  EXAMG_CUDA(cudaEventCreate(&t.start));
  EXAMG_CUDA(cudaEventRecord(t.start, c->stream));
  return EXAMG_OK;
}

int examg_timer_stop(examg_ctx *c, const char *name) {
  if (!c || !name) return set_error(EXAMG_EINVAL, "examg_timer_stop: null argument");
  auto it = c->timers.find(name);
  if (it == c->timers.end() || it->second.depth == 0) return set_error(EXAMG_ESTATE, "stopTimer(\"%s\") without startTimer", name);
  Timer &t = it->second;
  if (--t.depth > 0) return EXAMG_OK;
  cudaEvent_t stop;
  EXAMG_CUDA(cudaEventCreate(&stop));
  EXAMG_CUDA(cudaEventRecord(stop, c->stream));
  t.done.push_back({t.start, stop});
  t.start = nullptr;
  t.count++;
  if (t.done.size() > 256) fold_timer(t, false);
  return EXAMG_OK;
}

int examg_timer_total_ms(examg_ctx *c, const char *name, double *total_ms, long long *count) {
  if (!c || !name || !total_ms) return set_error(EXAMG_EINVAL, "examg_timer_total_ms: null argument");
  auto it = c->timers.find(name);
  if (it == c->timers.end()) return set_error(EXAMG_EINVAL, "getTotalTime(\"%s\"): unknown timer", name);
  fold_timer(it->second, true);
  *total_ms = it->second.total_ms;
  if (count) *count = it->second.count;
  return EXAMG_OK;
}

int examg_timer_print_all(examg_ctx *c) {
  if (!c) return set_error(EXAMG_EINVAL, "null context");
  for (auto &kv : c->timers) {
    fold_timer(kv.second, true);
    // the reference prints "Mean mean total time for Timer <name>: <ms>" per timer (one rank: mean = total)
    printf("Mean mean total time for Timer %s: %g\n", kv.first.c_str(), kv.second.total_ms);
  }
  fflush(stdout);
  return EXAMG_OK;
}

long long examg_launch_count(examg_ctx *c) { return c ? c->launches : -1; }
long long examg_device_bytes(examg_ctx *c) { return c ? c->dev_bytes : -1; }

// ---- fields -------------------------------------------------------------------

static int field_create(examg_ctx *c, const char *name, int level, int num_slots, int ncomp, int loc);

int examg_field_create(examg_ctx *c, const char *name, int level, int num_slots, int ncomp) {
  return field_create(c, name, level, num_slots, ncomp, 0);
}

int examg_field_create_cell(examg_ctx *c, const char *name, int level, int num_slots, int ncomp) {
  if (c && c->periodic_any)
    return set_error(EXAMG_EUNSUP, "examg_field_create_cell: cell fields are non-periodic so far");
  return field_create(c, name, level, num_slots, ncomp, 1);
}

int examg_level_cells(examg_ctx *c, int level, int n[3]) {
  const LevelGeom *g = geom(c, level);
  if (!g || !n) return set_error(EXAMG_EINVAL, "examg_level_cells: level %d", level);
  for (int d = 0; d < 3; ++d) n[d] = d < c->cfg.dim ? g->n[d] - 1 : 1;
  return EXAMG_OK;
}

static int field_create(examg_ctx *c, const char *name, int level, int num_slots, int ncomp, int loc) {
  const LevelGeom *g = geom(c, level);
  if (!g) return set_error(EXAMG_EINVAL, "examg_field_create: level %d", level);
  if (num_slots < 1 || num_slots > 4 || ncomp < 1 || ncomp > EXAMG_MAX_STENCIL)
    return set_error(EXAMG_EINVAL, "examg_field_create: slots %d comps %d", num_slots, ncomp);
  Field f;
  f.alive = true;
  f.name = name ? name : "";
  f.level = level;
  f.nslots = num_slots;
  f.ncomp = ncomp;
  f.loc = loc;
  size_t bytes = sizeof(double) * (size_t)ncomp * g->vol;
  for (int sl = 0; sl < num_slots; ++sl) {
    f.slot[sl] = f.orig[sl] = dev_alloc(c, bytes);  // zero-filled: initFieldsWithZero
    if (!f.slot[sl]) {
      for (int q = 0; q < sl; ++q) dev_free(c, f.slot[q]);
      return set_error(EXAMG_ENOMEM, "examg_field_create(%s@%d): cudaMalloc of %zu bytes failed", f.name.c_str(),
                       level, bytes);
    }
  }
  c->fields.push_back(f);
  return (int)c->fields.size() - 1;
}

int examg_field_destroy(examg_ctx *c, int h) {
  Field *f = get_field(c, h);
  if (!f) return set_error(EXAMG_EINVAL, "examg_field_destroy: handle %d", h);
  EXAMG_SYNC(c);
  for (int i = 0; i < 5; ++i)
    if (f->orig[i]) dev_free(c, f->orig[i]);
  for (int sl = 0; sl < 4; ++sl) f->slot[sl] = nullptr;
  f->shadow = nullptr;
  f->alive = false;
  return EXAMG_OK;
}

int examg_field_num_slots(examg_ctx *c, int h) {
  Field *f = get_field(c, h);
  return f ? f->nslots : set_error(EXAMG_EINVAL, "bad field %d", h);
}
int examg_field_active_slot(examg_ctx *c, int h) {
  Field *f = get_field(c, h);
  return f ? f->cur : set_error(EXAMG_EINVAL, "bad field %d", h);
}
int examg_field_advance(examg_ctx *c, int h) {
  Field *f = get_field(c, h);
  if (!f) return set_error(EXAMG_EINVAL, "bad field %d", h);
  f->cur = (f->cur + 1) % f->nslots;
  return EXAMG_OK;
}

int examg_field_set_bc(examg_ctx *c, int h, int kind, double value) {
  Field *f = get_field(c, h);
  if (!f) return set_error(EXAMG_EINVAL, "bad field %d", h);
  if (kind == EXAMG_BC_NEUMANN && f->loc != 1)
    return set_error(EXAMG_EUNSUP, "examg_field_set_bc: Neumann is implemented for cell-localised fields");
  if (f->loc == 1 && kind == EXAMG_BC_DIRICHLET_HOST)
    return set_error(EXAMG_EUNSUP, "examg_field_set_bc: cell-localised fields take None, Neumann or a Dirichlet value / program "
                     "(the ghost cell is 2 * value - inner cell)");
  if (kind != EXAMG_BC_NONE && kind != EXAMG_BC_DIRICHLET_CONST && kind != EXAMG_BC_DIRICHLET_HOST &&
      kind != EXAMG_BC_NEUMANN)
    return set_error(EXAMG_EINVAL, "examg_field_set_bc: kind %d (use examg_field_set_bc_expr for programs)", kind);
  f->bc_kind = kind;
  f->bc_value = value;
  f->bc_prog.clear();
  return EXAMG_OK;
}

// validates a postfix program on the host so the device never has to; ncomp > 0 allows EXAMG_OP_COMP
static int check_program(const char *who, const examg_bc_op *prog, int n, int ncomp, int nsrc = 0) {
  if (!prog || n < 1 || n > 64) return set_error(EXAMG_EINVAL, "%s: 1..64 ops", who);
  int depth = 0;
  for (int i = 0; i < n; ++i) {
    switch (prog[i].op) {
      case EXAMG_OP_COMP:
        if (ncomp <= 0 || prog[i].value < 0 || prog[i].value >= ncomp || prog[i].value != (int)prog[i].value)
          return set_error(EXAMG_EINVAL, "%s: op %d reads component %g of %d", who, i, prog[i].value, ncomp);
        ++depth; break;
      case EXAMG_OP_FIELD:
        if (prog[i].value < 0 || prog[i].value >= nsrc || prog[i].value != (int)prog[i].value)
          return set_error(EXAMG_EINVAL, "%s: op %d reads source field %g of %d", who, i, prog[i].value, nsrc);
        ++depth; break;
      case EXAMG_OP_CONST: case EXAMG_OP_X: case EXAMG_OP_Y: case EXAMG_OP_Z: ++depth; break;
      case EXAMG_OP_ADD: case EXAMG_OP_SUB: case EXAMG_OP_MUL: case EXAMG_OP_DIV:
        if (depth < 2) return set_error(EXAMG_EINVAL, "%s: stack underflow at op %d", who, i);
        --depth; break;
      case EXAMG_OP_NEG: case EXAMG_OP_EXP: case EXAMG_OP_SIN: case EXAMG_OP_COS: case EXAMG_OP_SINH:
      case EXAMG_OP_SQRT:
        if (depth < 1) return set_error(EXAMG_EINVAL, "%s: stack underflow at op %d", who, i);
        break;
      default: return set_error(EXAMG_EINVAL, "%s: unknown op %d", who, prog[i].op);
    }
    if (depth > 16) return set_error(EXAMG_EINVAL, "%s: stack deeper than 16", who);
  }
  if (depth != 1) return set_error(EXAMG_EINVAL, "%s: program leaves %d values", who, depth);
  return EXAMG_OK;
}

int examg_field_set_bc_expr(examg_ctx *c, int h, const examg_bc_op *prog, int n) {
  Field *f = get_field(c, h);
  if (!f) return set_error(EXAMG_EINVAL, "bad field %d", h);
  int rc = check_program("bc program", prog, n, 0);
  if (rc) return rc;
  f->bc_kind = EXAMG_BC_DIRICHLET_EXPR;
  f->bc_prog.assign(prog, prog + n);
  return EXAMG_OK;
}

int examg_field_eval(examg_ctx *c, int h, int slot, int comp, const examg_bc_op *prog, int n) {
  Field *f = get_field(c, h);
  if (!f) return set_error(EXAMG_EINVAL, "bad field %d", h);
  if (slot >= f->nslots) return set_error(EXAMG_EINVAL, "examg_field_eval: slot %d of %d", slot, f->nslots);
  if (comp < 0 || comp >= f->ncomp) return set_error(EXAMG_EINVAL, "examg_field_eval: component %d of %d", comp, f->ncomp);
  if (f->loc != 0) return set_error(EXAMG_EUNSUP, "examg_field_eval: node-localised fields only");
  int rc = check_program("examg_field_eval", prog, n, f->ncomp);
  if (rc) return rc;
  if ((rc = launch_field_eval(c, *f, slot, comp, prog, n))) return rc;
  if (slot <= 0 || f->nslots == 1) {
    // a later program for the same component supersedes the earlier one unless something in between read it
    bool read_since = false;
    for (size_t i = f->evals.size(); i-- > 0;) {
      if (f->evals[i].first == comp) {
        if (!read_since) f->evals.erase(f->evals.begin() + i);
        break;
      }
      for (const examg_bc_op &o : f->evals[i].second)
        read_since = read_since || (o.op == EXAMG_OP_COMP && (int)o.value == comp);
    }
    if (f->evals.size() >= 1024) f->evals.clear();  // a log this long is not an assembly recipe any more
    f->evals.emplace_back(comp, std::vector<examg_bc_op>(prog, prog + n));
    f->eval_gen++;
    // the loop also ran on the first ghost layer (global node positions): once every component was produced this
    // way the layer holds what the neighbour blocks computed for their own nodes
    std::vector<char> seen(f->ncomp, 0);
    for (const auto &e : f->evals) seen[e.first] = 1;
    bool all = true;
    for (char v : seen) all = all && v;
    if (all && f->halo_depth < 1) f->halo_depth = 1;
  } else {
    f->evals.clear();
    f->eval_gen++;
  }
  return EXAMG_OK;
}

int examg_loop_eval(examg_ctx *c, int dst, int dst_slot, const int *src, const int *src_slot, int nsrc,
                    const examg_bc_op *prog, int n) {
  Field *fd = get_field(c, dst);
  if (!fd) return set_error(EXAMG_EINVAL, "bad field %d", dst);
  if (dst_slot >= fd->nslots) return set_error(EXAMG_EINVAL, "examg_loop_eval: slot %d of %d", dst_slot, fd->nslots);
  if (fd->loc != 0 || fd->ncomp != 1) return set_error(EXAMG_EUNSUP, "examg_loop_eval: node-localised scalar fields");
  if (nsrc < 0 || nsrc > 8 || (nsrc > 0 && !src)) return set_error(EXAMG_EINVAL, "examg_loop_eval: 0..8 source fields");
  const double *ptr[8] = {nullptr};
  for (int k = 0; k < nsrc; ++k) {
    Field *fs = get_field(c, src[k]);
    if (!fs) return set_error(EXAMG_EINVAL, "bad field %d", src[k]);
    const int sl = src_slot ? src_slot[k] : -1;
    if (sl >= fs->nslots) return set_error(EXAMG_EINVAL, "examg_loop_eval: source %d: slot %d of %d", k, sl, fs->nslots);
    if (fs->level != fd->level || fs->loc != 0 || fs->ncomp != 1)
      return set_error(EXAMG_EINVAL, "examg_loop_eval: source %d is not a node-localised scalar field of level %d", k, fd->level);
    ptr[k] = field_ptr(c, fs, sl);
  }
  int rc = check_program("examg_loop_eval", prog, n, 0, nsrc);
  if (rc) return rc;
  fd->halo_depth = 0;
  if (dst_slot <= 0 || fd->nslots == 1) { fd->evals.clear(); fd->eval_gen++; }
  return launch_loop_eval(c, fd->level, ptr, nsrc, field_ptr(c, fd, dst_slot), prog, n);
}

int examg_field_set_bc_face(examg_ctx *c, int h, int face, int kind, double value, const examg_bc_op *prog, int n) {
  Field *f = get_field(c, h);
  if (!f) return set_error(EXAMG_EINVAL, "bad field %d", h);
  if (f->loc != 1) return set_error(EXAMG_EUNSUP, "examg_field_set_bc_face: cell-localised fields");
  if (face < 0 || face >= 2 * c->cfg.dim) return set_error(EXAMG_EINVAL, "examg_field_set_bc_face: face %d", face);
  if (kind != EXAMG_BC_NONE && kind != EXAMG_BC_DIRICHLET_CONST && kind != EXAMG_BC_DIRICHLET_EXPR && kind != EXAMG_BC_NEUMANN)
    return set_error(EXAMG_EINVAL, "examg_field_set_bc_face: kind %d", kind);
  f->bc_face_prog[face].clear();
  if (kind == EXAMG_BC_DIRICHLET_EXPR) {
    int rc = check_program("bc program", prog, n, 0);
    if (rc) return rc;
    f->bc_face_prog[face].assign(prog, prog + n);
  }
  f->bc_face_kind[face] = kind;
  f->bc_face_value[face] = value;
  if (f->bc_kind == EXAMG_BC_NONE) f->bc_kind = EXAMG_BC_DIRICHLET_CONST;  // so that apply bc reaches the faces
  return EXAMG_OK;
}

int examg_apply_bc(examg_ctx *c, int h, int slot) {
  Field *f = get_field(c, h);
  if (!f) return set_error(EXAMG_EINVAL, "bad field %d", h);
  if (slot >= f->nslots) return set_error(EXAMG_EINVAL, "bad slot %d", slot);
  if (f->bc_kind == EXAMG_BC_NONE || f->bc_kind == EXAMG_BC_DIRICHLET_HOST) return EXAMG_OK;
  if (f->loc == 1) {
    bool face_own = false;
    for (int k = 0; k < 6; ++k) face_own = face_own || f->bc_face_kind[k] >= 0;
    return f->bc_kind == EXAMG_BC_NEUMANN && c->cfg.dim == 2 && !face_own && c->world == 1 ? launch_bc_cell(c, *f, slot)
                                                                          : launch_bc_cell_generic(c, *f, slot);
  }
  return launch_bc(c, *f, slot);
}

static int copy_field(examg_ctx *c, int h, int slot, double *host, bool up, bool async, bool device_side = false) {
  Field *f = get_field(c, h);
  if (!f) return set_error(EXAMG_EINVAL, "bad field %d", h);
  if (!host) return set_error(EXAMG_EINVAL, "null host pointer");
  if (slot >= f->nslots) return set_error(EXAMG_EINVAL, "bad slot %d", slot);
  const LevelGeom &g = *geom(c, f->level);
  if (up) { f->halo_depth = 0; f->evals.clear(); f->eval_gen++; }
  int e[3];  // dense extents: nodes, or cells of a cell-localised field
  for (int d = 0; d < 3; ++d) e[d] = g.n[d] - (f->loc == 1 && d < c->cfg.dim ? 1 : 0);
  const size_t dense = (size_t)e[0] * e[1] * e[2];
  // Large host transfers go through a dense device staging array: one flat copy over PCIe at full
  // rate (a pitched 3D copy of 4 KB rows reaches ~60% of it), then a device-side 3D copy into /
  // out of the padded layout.
  double *stage = nullptr;
  if (!device_side && dense * sizeof(double) >= (1u << 20)) {
    if (c->stage_doubles < dense) {
      EXAMG_SYNC(c);
      if (c->stage) dev_free(c, c->stage);
      c->stage = dev_alloc(c, dense * sizeof(double));
      c->stage_doubles = c->stage ? dense : 0;
    }
    stage = c->stage;  // nullptr (allocation failed): fall back to the direct pitched copy
  }
  for (int comp = 0; comp < f->ncomp; ++comp) {
    double *dev = field_ptr(c, f, slot, comp);
    double *hp = host + (size_t)comp * dense;
    if (stage && up) EXAMG_CUDA(cudaMemcpyAsync(stage, hp, dense * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    double *other = stage ? stage : hp;
    cudaMemcpy3DParms p;
    memset(&p, 0, sizeof(p));
    // device: rows of `pitch` doubles, `ay` rows per plane; node (0,0,0) at dev
    cudaPitchedPtr dp = make_cudaPitchedPtr(dev, sizeof(double) * g.pitch, e[0], g.ay);
    cudaPitchedPtr hpp = make_cudaPitchedPtr(other, sizeof(double) * e[0], e[0], e[1]);
    p.srcPtr = up ? hpp : dp;
    p.dstPtr = up ? dp : hpp;
    p.extent = make_cudaExtent(sizeof(double) * e[0], e[1], e[2]);
    p.kind = (device_side || stage) ? cudaMemcpyDeviceToDevice : (up ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost);
    EXAMG_CUDA(cudaMemcpy3DAsync(&p, c->stream));
    if (stage && !up) EXAMG_CUDA(cudaMemcpyAsync(hp, stage, dense * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  }
  if (!async) EXAMG_SYNC(c);
  return EXAMG_OK;
}

}  // extern "C"

// dense device array <-> field (agglomeration staging, examg_cycle.cu); asynchronous, capturable
int examg::copy_field_device(examg_ctx *c, int h, int slot, double *dense, bool to_field) {
  return copy_field(c, h, slot, dense, to_field, true, true);
}

extern "C" {

int examg_field_upload(examg_ctx *c, int h, int slot, const double *host) {
  return copy_field(c, h, slot, const_cast<double *>(host), true, false);
}
int examg_field_download(examg_ctx *c, int h, int slot, double *host) {
  return copy_field(c, h, slot, host, false, false);
}
int examg_field_upload_async(examg_ctx *c, int h, int slot, const double *host) {
  return copy_field(c, h, slot, const_cast<double *>(host), true, true);
}
int examg_field_download_async(examg_ctx *c, int h, int slot, double *host) {
  return copy_field(c, h, slot, host, false, true);
}

// Pipelined upload: _begin starts the host -> device copy of the NEXT input on the library's copy stream (the
// host array must be pinned and stay untouched until _finish), _finish orders the device-side scatter into
// the field with the kernels of the main stream.  Between the two calls the main stream can compute.
int examg_field_upload_begin(examg_ctx *c, int h, int slot, const double *host) {
  Field *f = get_field(c, h);
  if (!f || !host) return set_error(EXAMG_EINVAL, "examg_field_upload_begin: bad field %d or null host pointer", h);
  if (slot >= f->nslots) return set_error(EXAMG_EINVAL, "bad slot %d", slot);
  if (c->up_pending) return set_error(EXAMG_ESTATE, "examg_field_upload_begin: an upload is already in flight");
  const LevelGeom &g = *geom(c, f->level);
  size_t dense = 1;
  for (int d = 0; d < 3; ++d) dense *= (size_t)(g.n[d] - (f->loc == 1 && d < c->cfg.dim ? 1 : 0));
  const size_t need = dense * f->ncomp;
  if (!c->copy_stream) {
    EXAMG_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
      EXAMG_CUDA(cudaEventCreateWithFlags(&c->up_h2d[b], cudaEventDisableTiming));
      EXAMG_CUDA(cudaEventCreateWithFlags(&c->up_d2d[b], cudaEventDisableTiming));
    }
  }
  const int b = c->up_turn;
  if (c->up_doubles[b] < need) {
    EXAMG_SYNC(c);
    EXAMG_CUDA(cudaStreamSynchronize(c->copy_stream));
    if (c->up_stage[b]) dev_free(c, c->up_stage[b]);
    c->up_stage[b] = dev_alloc(c, need * sizeof(double));
    c->up_doubles[b] = c->up_stage[b] ? need : 0;
    if (!c->up_stage[b]) return set_error(EXAMG_ENOMEM, "examg_field_upload_begin: staging array of %zu bytes", need * sizeof(double));
  } else {
    EXAMG_CUDA(cudaStreamWaitEvent(c->copy_stream, c->up_d2d[b], 0));  // the scatter that last read this array
  }
  EXAMG_CUDA(cudaMemcpyAsync(c->up_stage[b], host, need * sizeof(double), cudaMemcpyHostToDevice, c->copy_stream));
  EXAMG_CUDA(cudaEventRecord(c->up_h2d[b], c->copy_stream));
  c->up_field = h;
  c->up_slot = slot;
  c->up_pending = true;
  return EXAMG_OK;
}

int examg_field_upload_finish(examg_ctx *c) {
  if (!c || !c->up_pending) return set_error(EXAMG_ESTATE, "examg_field_upload_finish: no upload in flight");
  Field *f = get_field(c, c->up_field);
  if (!f) return set_error(EXAMG_EINVAL, "examg_field_upload_finish: field %d vanished", c->up_field);
  const int b = c->up_turn;
  const LevelGeom &g = *geom(c, f->level);
  int e[3];
  for (int d = 0; d < 3; ++d) e[d] = g.n[d] - (f->loc == 1 && d < c->cfg.dim ? 1 : 0);
  const size_t dense = (size_t)e[0] * e[1] * e[2];
  EXAMG_CUDA(cudaStreamWaitEvent(c->stream, c->up_h2d[b], 0));
  for (int comp = 0; comp < f->ncomp; ++comp) {
    cudaMemcpy3DParms p;
    memset(&p, 0, sizeof(p));
    p.srcPtr = make_cudaPitchedPtr(c->up_stage[b] + (size_t)comp * dense, sizeof(double) * e[0], e[0], e[1]);
    p.dstPtr = make_cudaPitchedPtr(field_ptr(c, f, c->up_slot, comp), sizeof(double) * g.pitch, e[0], g.ay);
    p.extent = make_cudaExtent(sizeof(double) * e[0], e[1], e[2]);
    p.kind = cudaMemcpyDeviceToDevice;
    EXAMG_CUDA(cudaMemcpy3DAsync(&p, c->stream));
  }
  EXAMG_CUDA(cudaEventRecord(c->up_d2d[b], c->stream));
  f->halo_depth = 0;
  f->evals.clear();
  f->eval_gen++;
  c->up_turn ^= 1;
  c->up_pending = false;
  return EXAMG_OK;
}

int examg_field_print(examg_ctx *c, int h, int slot, const char *filename, const char *separator) {
  Field *f = get_field(c, h);
  if (!f) return set_error(EXAMG_EINVAL, "bad field %d", h);
  if (!filename) return set_error(EXAMG_EINVAL, "examg_field_print: null file name");
  if (slot >= f->nslots) return set_error(EXAMG_EINVAL, "bad slot %d", slot);
  const LevelGeom &g = *geom(c, f->level);
  const char *sep = separator ? separator : " ";
  int e[3];
  for (int d = 0; d < 3; ++d) e[d] = g.n[d] - (f->loc == 1 && d < c->cfg.dim ? 1 : 0);
  const size_t dense = (size_t)e[0] * e[1] * e[2];
  std::vector<double> host((size_t)f->ncomp * dense);
  int rc = examg_field_download(c, h, slot, host.data());
  if (rc) return rc;
  FILE *fp = fopen(filename, "w");
  if (!fp) return set_error(EXAMG_EINVAL, "examg_field_print: cannot open %s", filename);
  const double off = f->loc == 1 ? 0.5 : 0.0;
  for (int k = 0; k < e[2]; ++k)
    for (int j = 0; j < e[1]; ++j)
      for (int i = 0; i < e[0]; ++i) {
        const int idx[3] = {i, j, k};
        for (int d = 0; d < c->cfg.dim; ++d) {
          // vf_nodePosition = i * h + lo; vf_cellCenter = vf_nodePosition + 0.5 * vf_cellWidth
          const double pos = ((double)(g.goff[d] + idx[d]) * g.h[d] + c->cfg.domain_lo[d]) + off * g.h[d];
          fprintf(fp, "%g%s", pos, sep);
        }
        const size_t p = (size_t)i + (size_t)e[0] * ((size_t)j + (size_t)e[1] * (size_t)k);
        for (int comp = 0; comp < f->ncomp; ++comp)
          fprintf(fp, comp + 1 < f->ncomp ? "%g%s" : "%g\n", host[(size_t)comp * dense + p], sep);
      }
  fclose(fp);
  return EXAMG_OK;
}

// writeField / readField with the "lock" interface in binary mode (C/s/e/field/ir/IR_WriteField.scala:26-45,
// IR_ReadField.scala:27-46, io/ir/IR_FileAccess_Locking.scala:87-200): the values only, raw doubles in the layout of the
// get<name>/set<name> arrays -- [comp][z][y][x], x fastest, ghost layers excluded (includeGhostLayers = false).  With
// several blocks the reference appends block after block in rank order (MPI_Sequential): `append` != 0 does that for
// this block; the caller sequences the ranks.
int examg_field_write(examg_ctx *c, int h, int slot, const char *filename, int append) {
  Field *f = get_field(c, h);
  if (!f) return set_error(EXAMG_EINVAL, "bad field %d", h);
  if (!filename) return set_error(EXAMG_EINVAL, "examg_field_write: null file name");
  if (slot >= f->nslots) return set_error(EXAMG_EINVAL, "bad slot %d", slot);
  const LevelGeom &g = *geom(c, f->level);
  size_t dense = 1;
  for (int d = 0; d < 3; ++d) dense *= (size_t)(g.n[d] - (f->loc == 1 && d < c->cfg.dim ? 1 : 0));
  std::vector<double> host((size_t)f->ncomp * dense);
  int rc = examg_field_download(c, h, slot, host.data());
  if (rc) return rc;
  FILE *fp = fopen(filename, append ? "ab" : "wb");
  if (!fp) return set_error(EXAMG_EINVAL, "examg_field_write: cannot open %s", filename);
  const size_t nw = fwrite(host.data(), sizeof(double), host.size(), fp);
  fclose(fp);
  if (nw != host.size()) return set_error(EXAMG_EINVAL, "examg_field_write: short write to %s", filename);
  return EXAMG_OK;
}

// offset_doubles: where this block's values start in the file (0 for one block; rank * values per block for the
// rank-ordered file examg_field_write produces)
int examg_field_read(examg_ctx *c, int h, int slot, const char *filename, long long offset_doubles) {
  Field *f = get_field(c, h);
  if (!f) return set_error(EXAMG_EINVAL, "bad field %d", h);
  if (!filename) return set_error(EXAMG_EINVAL, "examg_field_read: null file name");
  if (slot >= f->nslots) return set_error(EXAMG_EINVAL, "bad slot %d", slot);
  if (offset_doubles < 0) return set_error(EXAMG_EINVAL, "examg_field_read: negative offset");
  const LevelGeom &g = *geom(c, f->level);
  size_t dense = 1;
  for (int d = 0; d < 3; ++d) dense *= (size_t)(g.n[d] - (f->loc == 1 && d < c->cfg.dim ? 1 : 0));
  std::vector<double> host((size_t)f->ncomp * dense);
  FILE *fp = fopen(filename, "rb");
  if (!fp) return set_error(EXAMG_EINVAL, "examg_field_read: cannot open %s", filename);
  size_t nr = 0;
  if (fseek(fp, (long)(offset_doubles * (long long)sizeof(double)), SEEK_SET) == 0)
    nr = fread(host.data(), sizeof(double), host.size(), fp);
  fclose(fp);
  if (nr != host.size())
    return set_error(EXAMG_EINVAL, "examg_field_read: %s holds %zu of the %zu values of %s@%d", filename, nr, host.size(),
                     f->name.c_str(), f->level);
  return examg_field_upload(c, h, slot, host.data());
}

int examg_field_fill(examg_ctx *c, int h, double value) {
  Field *f = get_field(c, h);
  if (!f) return set_error(EXAMG_EINVAL, "bad field %d", h);
  const LevelGeom &g = *geom(c, f->level);
  const size_t count = (size_t)f->ncomp * g.vol;
  f->halo_depth = 0;
  f->evals.clear();
  f->eval_gen++;
  std::vector<double> row(g.pitch, value);
  for (int sl = 0; sl < f->nslots; ++sl) {
    double *base = f->slot[sl];
    if (value == 0.0) {
      EXAMG_CUDA(cudaMemsetAsync(base, 0, sizeof(double) * count, c->stream));
      continue;
    }
    // no 64-bit memset in the runtime API: stage one row, then doubling device copies
    EXAMG_CUDA(cudaMemcpyAsync(base, row.data(), sizeof(double) * g.pitch, cudaMemcpyHostToDevice, c->stream));
    EXAMG_SYNC(c);
    for (size_t done = g.pitch; done < count;) {
      size_t chunk = done < count - done ? done : count - done;
      EXAMG_CUDA(cudaMemcpyAsync(base + done, base, sizeof(double) * chunk, cudaMemcpyDeviceToDevice, c->stream));
      done += chunk;
    }
  }
  return EXAMG_OK;
}

// ---- stencils -----------------------------------------------------------------

static int add_stencil(examg_ctx *c, int level, int n, const int (*off)[3], const double *coef, int cf) {
  const LevelGeom *g = geom(c, level);
  if (!g) return set_error(EXAMG_EINVAL, "stencil: level %d", level);
  if (n < 1 || n > EXAMG_MAX_STENCIL || !off) return set_error(EXAMG_EINVAL, "stencil: %d entries", n);
  Stencil s;
  s.level = level;
  s.nst = n;
  s.centre = -1;
  s.coeff_field = cf;
  for (int k = 0; k < n; ++k) {
    for (int d = 0; d < 3; ++d) {
      s.off[k][d] = off[k][d];
      if (off[k][d] < -1 || off[k][d] > 1 || (d >= c->cfg.dim && off[k][d] != 0))
        return set_error(EXAMG_EUNSUP, "stencil: offset %d of entry %d outside the compact 3^d box", off[k][d], k);
    }
    s.coef[k] = coef ? coef[k] : 0.0;
    if (!off[k][0] && !off[k][1] && !off[k][2]) s.centre = k;
  }
  if (s.centre < 0) return set_error(EXAMG_EINVAL, "stencil without centre entry (diag() undefined)");
  s.is7 = false;
  if (cf < 0 && c->cfg.dim == 3 && n == 7) {
    int seen = 0;
    for (int k = 0; k < n; ++k) {
      int nz = (off[k][0] != 0) + (off[k][1] != 0) + (off[k][2] != 0);
      if (nz <= 1) ++seen;
    }
    s.is7 = (seen == 7);
  }
  c->stencils.push_back(s);
  return (int)c->stencils.size() - 1;
}

int examg_stencil_const(examg_ctx *c, int level, int n, const int (*off)[3], const double *coef) {
  if (!coef) return set_error(EXAMG_EINVAL, "examg_stencil_const: null coefficients");
  return add_stencil(c, level, n, off, coef, -1);
}

int examg_stencil_field(examg_ctx *c, int level, int n, const int (*off)[3], int cf) {
  Field *f = get_field(c, cf);
  if (!f || f->level != level || f->ncomp != n)
    return set_error(EXAMG_EINVAL, "examg_stencil_field: coefficient field must live on level %d with %d components",
                     level, n);
  return add_stencil(c, level, n, off, nullptr, cf);
}

}  // extern "C"
