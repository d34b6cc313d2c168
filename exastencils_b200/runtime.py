"""ctypes binding of libexamg.so (include/examg.h) -- the only way the host side
reaches the GPU.  There is no CPU path: if the library is missing or no CUDA device
is present, calls raise.

The binding is the Python analogue of the host C++ the new generator back-end
emits: every call is checked and a failure raises ExaMGError with
examg_last_error(), which is the `IR_Assert -> exit(1)` convention of the reference
(C/s/e/parallelization/api/cuda/CUDA_Error.scala:39-68) turned into an exception.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "lib", "libexamg.so")
HEADER_PATH = os.path.join(os.path.dirname(_PKG), "include", "examg.h")

MAX_STENCIL = 27
BC_NONE, BC_DIRICHLET_CONST, BC_DIRICHLET_EXPR, BC_DIRICHLET_HOST = 0, 1, 2, 3
OP_CONST, OP_X, OP_Y, OP_Z, OP_ADD, OP_SUB, OP_MUL, OP_DIV, OP_NEG = range(9)
OP_EXP, OP_SIN, OP_COS, OP_SINH, OP_SQRT, OP_COMP, OP_FIELD = range(9, 16)
SMOOTH_JACOBI, SMOOTH_RBGS, SMOOTH_COLORED_2, SMOOTH_COLORED_3 = 0, 1, 2, 3
CGS_CG, CGS_BICGSTAB, CGS_CR, CGS_MINRES, CGS_SMOOTHER = range(5)
BC_NEUMANN = 4
F_SOLUTION, F_RHS, F_RESIDUAL, F_CGTMP0, F_CGTMP1, F_APPROX = range(6)


class ExaMGError(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [("dim", C.c_int), ("min_level", C.c_int), ("max_level", C.c_int),
                ("block_cells", C.c_int * 3), ("num_blocks", C.c_int * 3), ("rank", C.c_int),
                ("device", C.c_int), ("domain_lo", C.c_double * 3), ("domain_hi", C.c_double * 3),
                ("periodic", C.c_int * 3)]


class BcOp(C.Structure):
    _fields_ = [("op", C.c_int), ("value", C.c_double)]


class CycleDesc(C.Structure):
    _fields_ = [("smoother", C.c_int), ("omega", C.c_double), ("nu_pre", C.c_int), ("nu_post", C.c_int),
                ("restrict_scale", C.c_double), ("cg_max_it", C.c_int), ("cg_rel_tol", C.c_double),
                ("stencils", C.POINTER(C.c_int)), ("coarse_solver", C.c_int), ("use_fas", C.c_int)]


class System2Desc(C.Structure):
    _fields_ = [("laplace", C.c_int), ("scale", C.c_double), ("c_uu", C.c_int), ("c_uv", C.c_int), ("c_vv", C.c_int)]


class OptFlow3Desc(C.Structure):
    _fields_ = [("alpha", C.c_double), ("nu_pre", C.c_int), ("nu_post", C.c_int), ("cg_max_it", C.c_int),
                ("cg_rel_tol", C.c_double)]


class StokesDesc(C.Structure):
    _fields_ = [("nu_pre", C.c_int), ("nu_post", C.c_int), ("nu_coarse", C.c_int), ("top_vel", C.c_double)]


class OptFlowDesc(C.Structure):
    _fields_ = [("alpha", C.c_double), ("relax", C.c_double), ("nu_pre", C.c_int), ("nu_post", C.c_int),
                ("cg_max_it", C.c_int), ("cg_rel_tol", C.c_double)]


def build(force=False, verbose=False):
    """Compile libexamg.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    csrc = os.path.join(_PKG, "csrc")
    if force:
        subprocess.check_call(["make", "-C", csrc, "clean"], stdout=subprocess.DEVNULL)
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-C", csrc, "-j8"], stdout=out)
    return LIB_PATH


_lib = None
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_off_t = C.POINTER(C.c_int * 3)

# name -> (restype, argtypes); every symbol declared in include/examg.h
SIGNATURES = {
    "examg_abi_version": (C.c_int, []),
    "examg_last_error": (C.c_char_p, []),
    "examg_init": (C.c_void_p, [C.POINTER(Config)]),
    "examg_finalize": (C.c_int, [C.c_void_p]),
    "examg_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "examg_sync": (C.c_int, [C.c_void_p]),
    "examg_level_dims": (C.c_int, [C.c_void_p, C.c_int, _ip]),
    "examg_level_width": (C.c_int, [C.c_void_p, C.c_int, _dp]),
    "examg_field_create": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.c_int]),
    "examg_field_destroy": (C.c_int, [C.c_void_p, C.c_int]),
    "examg_field_num_slots": (C.c_int, [C.c_void_p, C.c_int]),
    "examg_field_active_slot": (C.c_int, [C.c_void_p, C.c_int]),
    "examg_field_advance": (C.c_int, [C.c_void_p, C.c_int]),
    "examg_field_set_bc": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_double]),
    "examg_field_set_bc_expr": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(BcOp), C.c_int]),
    "examg_cr_coarse": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_double,
                                  C.POINTER(C.c_int)]),
    "examg_minres_coarse": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int,
                                      C.c_double, C.POINTER(C.c_int)]),
    "examg_bicgstab_coarse": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int,
                                        C.c_double, C.POINTER(C.c_int)]),
    "examg_smooth_colored": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int]),
    "examg_field_print": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_char_p, C.c_char_p]),
    "examg_field_set_bc_face": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.POINTER(BcOp), C.c_int]),
    "examg_field_eval": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(BcOp), C.c_int]),
    "examg_loop_eval": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _ip, _ip, C.c_int, C.POINTER(BcOp), C.c_int]),
    "examg_apply_bc": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "examg_field_upload": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "examg_field_download": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "examg_field_upload_async": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "examg_field_download_async": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "examg_field_upload_begin": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "examg_field_upload_finish": (C.c_int, [C.c_void_p]),
    "examg_field_fill": (C.c_int, [C.c_void_p, C.c_int, C.c_double]),
    "examg_stencil_const": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _off_t, _dp]),
    "examg_stencil_field": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _off_t, C.c_int]),
    "examg_apply": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]),
    "examg_residual": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _dp]),
    "examg_smooth_jacobi": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int]),
    "examg_smooth_rbgs": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int]),
    "examg_smooth_rb_color": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int]),
    "examg_restrict": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_double]),
    "examg_residual_restrict": (C.c_int, [C.c_void_p] + [C.c_int] * 5 + [C.c_double]),
    "examg_prolong_add": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]),
    "examg_prolong_add_smooth_rbgs": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int]),
    "examg_dot": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, _dp]),
    "examg_dot_dense_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p]),
    "examg_max_abs_diff": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, _dp]),
    "examg_set": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_double]),
    "examg_copy": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]),
    "examg_axpy": (C.c_int, [C.c_void_p, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int]),
    "examg_xpay": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int]),
    "examg_exchange": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "examg_exchange_depth": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "examg_smoother_coarse": (C.c_int, [C.c_void_p] + [C.c_int] * 5 + [C.c_double, C.c_int, C.c_int, C.c_double, _ip]),
    "examg_cycle_desc_from_options": (C.c_int, [C.c_char_p, C.c_int, C.c_void_p, _ip]),
    "examg_cg_coarse": (C.c_int, [C.c_void_p] + [C.c_int] * 7 + [C.c_double, _ip]),
    "examg_solver_create": (C.c_void_p, [C.c_void_p, C.POINTER(CycleDesc)]),
    "examg_solver_destroy": (C.c_int, [C.c_void_p]),
    "examg_solver_field": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "examg_solver_vcycle": (C.c_int, [C.c_void_p, C.c_int]),
    "examg_solver_cycles": (C.c_int, [C.c_void_p, C.c_int, _dp]),
    "examg_solver_residual_norm": (C.c_int, [C.c_void_p, _dp]),
    "examg_solver_last_cg_iters": (C.c_int, [C.c_void_p, _ip]),
    "examg_field_create_cell": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.c_int]),
    "examg_level_cells": (C.c_int, [C.c_void_p, C.c_int, _ip]),
    "examg_system2_create": (C.c_int, [C.c_void_p, C.POINTER(System2Desc)]),
    "examg_system2_residual": (C.c_int, [C.c_void_p] + [C.c_int] * 7 + [_dp]),
    "examg_system2_smooth_rb": (C.c_int, [C.c_void_p] + [C.c_int] * 5 + [C.c_double, C.c_int]),
    "examg_system2_smooth_rb_color": (C.c_int, [C.c_void_p] + [C.c_int] * 5 + [C.c_double, C.c_int]),
    "examg_restrict_cell": (C.c_int, [C.c_void_p, _ip, _ip, C.c_int]),
    "examg_prolong_add_cell": (C.c_int, [C.c_void_p, _ip, _ip, C.c_int]),
    "examg_system2_cg_coarse": (C.c_int, [C.c_void_p, C.c_int, _ip, C.c_int, C.c_double, _ip]),
    "examg_optflow_create": (C.c_void_p, [C.c_void_p, C.POINTER(OptFlowDesc)]),
    "examg_optflow_destroy": (C.c_int, [C.c_void_p]),
    "examg_optflow_field": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "examg_optflow_init_fields": (C.c_int, [C.c_void_p]),
    "examg_optflow_vcycle": (C.c_int, [C.c_void_p, C.c_int]),
    "examg_optflow_cycles": (C.c_int, [C.c_void_p, C.c_int, _dp]),
    "examg_optflow_residual_norm": (C.c_int, [C.c_void_p, _dp]),
    "examg_optflow_last_cg_iters": (C.c_int, [C.c_void_p, _ip]),
    "examg_stokes_create": (C.c_void_p, [C.c_void_p, C.POINTER(StokesDesc)]),
    "examg_stokes_destroy": (C.c_int, [C.c_void_p]),
    "examg_stokes_cells": (C.c_int, [C.c_void_p, C.c_int]),
    "examg_stokes_upload": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "examg_stokes_download": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "examg_stokes_apply_bc": (C.c_int, [C.c_void_p, C.c_int]),
    "examg_stokes_smooth_color": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]),
    "examg_stokes_smooth": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "examg_stokes_residual": (C.c_int, [C.c_void_p, C.c_int, _dp]),
    "examg_stokes_restrict": (C.c_int, [C.c_void_p, C.c_int]),
    "examg_stokes_set_solution": (C.c_int, [C.c_void_p, C.c_int, C.c_double]),
    "examg_stokes_correct": (C.c_int, [C.c_void_p, C.c_int]),
    "examg_stokes_vcycle": (C.c_int, [C.c_void_p, C.c_int]),
    "examg_optflow3_create": (C.c_void_p, [C.c_void_p, C.POINTER(OptFlow3Desc)]),
    "examg_optflow3_destroy": (C.c_int, [C.c_void_p]),
    "examg_optflow3_cells": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_int)]),
    "examg_optflow3_upload": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "examg_optflow3_download": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "examg_optflow3_init_fields": (C.c_int, [C.c_void_p]),
    "examg_optflow3_smooth_color": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "examg_optflow3_smooth": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "examg_optflow3_residual": (C.c_int, [C.c_void_p, C.c_int, _dp]),
    "examg_optflow3_restrict": (C.c_int, [C.c_void_p, C.c_int]),
    "examg_optflow3_correct": (C.c_int, [C.c_void_p, C.c_int]),
    "examg_optflow3_cg": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "examg_optflow3_last_cg_iters": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "examg_optflow3_vcycle": (C.c_int, [C.c_void_p, C.c_int]),
    "examg_optflow3_cycles": (C.c_int, [C.c_void_p, C.c_int, _dp]),
    "examg_comm_blob_size": (C.c_int, [C.c_void_p]),
    "examg_comm_export": (C.c_int, [C.c_void_p, C.c_void_p]),
    "examg_comm_import": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "examg_benchmark_start": (C.c_int, [C.c_void_p, C.c_char_p]),
    "examg_benchmark_stop": (C.c_int, [C.c_void_p, C.c_char_p]),
    "examg_field_write": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_char_p, C.c_int]),
    "examg_field_read": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_char_p, C.c_longlong]),
    "examg_timer_start": (C.c_int, [C.c_void_p, C.c_char_p]),
    "examg_timer_stop": (C.c_int, [C.c_void_p, C.c_char_p]),
    "examg_timer_total_ms": (C.c_int, [C.c_void_p, C.c_char_p, _dp, C.POINTER(C.c_longlong)]),
    "examg_timer_print_all": (C.c_int, [C.c_void_p]),
    "examg_launch_count": (C.c_longlong, [C.c_void_p]),
    "examg_device_bytes": (C.c_longlong, [C.c_void_p]),
}


def lib():
    """Load libexamg.so.  Fails loudly if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ExaMGError("libexamg.so not built (%s); run `python -c 'import __graft_entry__ as g; g.build()'` "
                         "-- there is no CPU fallback" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    if L.examg_abi_version() != 1:
        raise ExaMGError("libexamg ABI version mismatch")
    _lib = L
    return L


def cycle_desc_from_options(options, dim):
    """`generate solver ... with { ... }` options / solver_* Knowledge keys -> (CycleDesc, colours per dimension)."""
    desc, ncol = CycleDesc(), C.c_int()
    desc.restrict_scale = 1.0
    check(lib().examg_cycle_desc_from_options(options.encode(), dim, C.byref(desc), C.byref(ncol)))
    return desc, ncol.value


def check(rc):
    if rc < 0:
        raise ExaMGError("libexamg error %d: %s" % (rc, lib().examg_last_error().decode()))
    return rc


def _ptr(a):
    assert isinstance(a, np.ndarray) and a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


class Context:
    """One block of the domain on one GPU (initGlobals + initDomain + setupBuffers)."""

    def __init__(self, dim, min_level, max_level, block_cells=(1, 1, 1), num_blocks=(1, 1, 1), rank=0, device=0,
                 domain_lo=(0.0, 0.0, 0.0), domain_hi=(1.0, 1.0, 1.0), periodic=(0, 0, 0)):
        cfg = Config()
        cfg.dim, cfg.min_level, cfg.max_level, cfg.rank, cfg.device = dim, min_level, max_level, rank, device
        bc = list(block_cells) + [1] * (3 - len(block_cells))
        nb = list(num_blocks) + [1] * (3 - len(num_blocks))
        for d in range(3):
            cfg.block_cells[d], cfg.num_blocks[d] = bc[d], nb[d]
            cfg.domain_lo[d], cfg.domain_hi[d] = domain_lo[d], domain_hi[d]
            cfg.periodic[d] = int(bool(periodic[d])) if d < len(periodic) else 0
        self.cfg = cfg
        self.dim, self.min_level, self.max_level = dim, min_level, max_level
        self.L = lib()
        self.h = self.L.examg_init(C.byref(cfg))
        if not self.h:
            raise ExaMGError("examg_init: " + self.L.examg_last_error().decode())

    def close(self):
        if getattr(self, "h", None):
            self.L.examg_finalize(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- geometry
    def dims(self, level):
        n = (C.c_int * 3)()
        check(self.L.examg_level_dims(self.h, level, n))
        return n[0], n[1], n[2]

    def shape(self, level):
        nx, ny, nz = self.dims(level)
        return (nz, ny, nx)

    def width(self, level):
        h = (C.c_double * 3)()
        check(self.L.examg_level_width(self.h, level, h))
        return h[0], h[1], h[2]

    def set_stream(self, cuda_stream):
        check(self.L.examg_set_stream(self.h, C.c_void_p(cuda_stream)))

    def sync(self):
        check(self.L.examg_sync(self.h))

    def timer_start(self, name):
        check(self.L.examg_timer_start(self.h, name.encode()))

    def timer_stop(self, name):
        check(self.L.examg_timer_stop(self.h, name.encode()))

    def timer_total_ms(self, name):
        ms, n = C.c_double(), C.c_longlong()
        check(self.L.examg_timer_total_ms(self.h, name.encode(), C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def launch_count(self):
        return self.L.examg_launch_count(self.h)

    def device_bytes(self):
        return self.L.examg_device_bytes(self.h)

    # -- fields
    def field_create(self, name, level, num_slots=1, ncomp=1):
        return check(self.L.examg_field_create(self.h, name.encode(), level, num_slots, ncomp))

    def field_create_cell(self, name, level, num_slots=1, ncomp=1):
        return check(self.L.examg_field_create_cell(self.h, name.encode(), level, num_slots, ncomp))

    def cells(self, level):
        n = (C.c_int * 3)()
        check(self.L.examg_level_cells(self.h, level, n))
        return n[0], n[1], n[2]

    def cell_shape(self, level):
        nx, ny, nz = self.cells(level)
        return (nz, ny, nx)

    def download_cells(self, f, level, slot=-1):
        out = np.empty(self.cell_shape(level), dtype=np.float64)
        check(self.L.examg_field_download(self.h, f, slot, _ptr(out)))
        return out

    def field_destroy(self, f):
        check(self.L.examg_field_destroy(self.h, f))

    def field_active_slot(self, f):
        return check(self.L.examg_field_active_slot(self.h, f))

    def field_advance(self, f):
        check(self.L.examg_field_advance(self.h, f))

    def field_set_bc(self, f, kind, value=0.0):
        check(self.L.examg_field_set_bc(self.h, f, kind, value))

    def field_set_bc_expr(self, f, prog):
        """prog: list of (op, value) in postfix order."""
        arr = (BcOp * len(prog))()
        for i, (op, val) in enumerate(prog):
            arr[i].op, arr[i].value = op, val
        check(self.L.examg_field_set_bc_expr(self.h, f, arr, len(prog)))

    def field_set_bc_face(self, f, face, kind, value=0.0, prog=None):
        """A condition of its own for face 2*d + side of a cell field."""
        n = len(prog) if prog else 0
        arr = (BcOp * max(n, 1))()
        for i, (op, val) in enumerate(prog or []):
            arr[i].op, arr[i].value = op, val
        check(self.L.examg_field_set_bc_face(self.h, f, face, kind, value, arr, n))

    def field_eval(self, f, prog, comp=0, slot=-1):
        """loop over F { F[comp] = <postfix program over the node position> } (local nodes + one ghost layer)."""
        arr = (BcOp * len(prog))()
        for i, (op, val) in enumerate(prog):
            arr[i].op, arr[i].value = op, val
        check(self.L.examg_field_eval(self.h, f, slot, comp, arr, len(prog)))

    def loop_eval(self, dst, srcs, prog, dst_slot=-1, src_slots=None):
        """loop over dst { dst<dst_slot> = <postfix program over the node position and srcs at the same node> }
        (OP_FIELD k pushes srcs[k]); slots default to the active ones."""
        arr = (BcOp * len(prog))()
        for i, (op, val) in enumerate(prog):
            arr[i].op, arr[i].value = int(op), float(val)
        n = len(srcs)
        hs = (C.c_int * max(1, n))(*srcs)
        sl = (C.c_int * max(1, n))(*(src_slots if src_slots is not None else [-1] * n))
        check(self.L.examg_loop_eval(self.h, dst, dst_slot, hs, sl, n, arr, len(prog)))

    def field_print(self, f, filename, slot=-1, separator=None):
        """printField ( "file", F ): ascii, one line per point (position, value[s])."""
        check(self.L.examg_field_print(self.h, f, slot, filename.encode(), separator.encode() if separator else None))

    def field_write(self, f, filename, slot=-1, append=False):
        check(self.L.examg_field_write(self.h, f, slot, str(filename).encode(), 1 if append else 0))

    def field_read(self, f, filename, slot=-1, offset_doubles=0):
        check(self.L.examg_field_read(self.h, f, slot, str(filename).encode(), offset_doubles))

    def benchmark_start(self, name):
        check(self.L.examg_benchmark_start(self.h, name.encode()))

    def benchmark_stop(self, name):
        check(self.L.examg_benchmark_stop(self.h, name.encode()))

    def apply_bc(self, f, slot=-1):
        check(self.L.examg_apply_bc(self.h, f, slot))

    def upload(self, f, host, slot=-1):
        check(self.L.examg_field_upload(self.h, f, slot, _ptr(host)))

    def download(self, f, level, slot=-1, ncomp=1, out=None):
        shp = self.shape(level)
        if ncomp > 1:
            shp = (ncomp,) + shp
        if out is None:
            out = np.empty(shp, dtype=np.float64)
        check(self.L.examg_field_download(self.h, f, slot, _ptr(out)))
        return out

    def upload_async(self, f, host_ptr, slot=-1):
        check(self.L.examg_field_upload_async(self.h, f, slot, C.c_void_p(host_ptr)))

    def download_async(self, f, host_ptr, slot=-1):
        check(self.L.examg_field_download_async(self.h, f, slot, C.c_void_p(host_ptr)))

    def upload_begin(self, f, host_ptr, slot=-1):
        check(self.L.examg_field_upload_begin(self.h, f, slot, C.c_void_p(host_ptr)))

    def upload_finish(self):
        check(self.L.examg_field_upload_finish(self.h))

    def fill(self, f, value):
        check(self.L.examg_field_fill(self.h, f, value))

    # -- stencils
    @staticmethod
    def _offsets(offsets):
        arr = ((C.c_int * 3) * len(offsets))()
        for k, o in enumerate(offsets):
            o = list(o) + [0] * (3 - len(o))
            for d in range(3):
                arr[k][d] = int(o[d])
        return arr

    def stencil_const(self, level, offsets, coeffs):
        co = (C.c_double * len(coeffs))(*[float(v) for v in coeffs])
        return check(self.L.examg_stencil_const(self.h, level, len(offsets), self._offsets(offsets), co))

    def stencil_field(self, level, offsets, coeff_field):
        return check(self.L.examg_stencil_field(self.h, level, len(offsets), self._offsets(offsets), coeff_field))

    # -- hot path
    def apply(self, A, u, out, u_slot=-1):
        check(self.L.examg_apply(self.h, A, u, u_slot, out))

    def residual(self, A, u, rhs, res, u_slot=-1, norm=False):
        if norm:
            n2 = C.c_double()
            check(self.L.examg_residual(self.h, A, u, u_slot, rhs, res, C.byref(n2)))
            return n2.value
        check(self.L.examg_residual(self.h, A, u, u_slot, rhs, res, None))

    def smooth_jacobi(self, A, u, rhs, omega, nsweeps=1):
        check(self.L.examg_smooth_jacobi(self.h, A, u, rhs, omega, nsweeps))

    def smooth_rbgs(self, A, u, rhs, omega, nsweeps=1):
        check(self.L.examg_smooth_rbgs(self.h, A, u, rhs, omega, nsweeps))

    def smooth_colored(self, A, u, rhs, omega, colors_per_dim, nsweeps=1):
        check(self.L.examg_smooth_colored(self.h, A, u, rhs, omega, colors_per_dim, nsweeps))

    def smooth_rb_color(self, A, u, rhs, omega, color):
        check(self.L.examg_smooth_rb_color(self.h, A, u, rhs, omega, color))

    def restrict(self, res_fine, rhs_coarse, scale=1.0):
        check(self.L.examg_restrict(self.h, res_fine, rhs_coarse, scale))

    def residual_restrict(self, A, u, rhs, res, rhs_coarse, scale=1.0):
        check(self.L.examg_residual_restrict(self.h, A, u, rhs, res, rhs_coarse, scale))

    def prolong_add(self, u_coarse, u_fine, uc_slot=-1, uf_slot=-1):
        check(self.L.examg_prolong_add(self.h, u_coarse, uc_slot, u_fine, uf_slot))

    def prolong_add_smooth_rbgs(self, A, u_coarse, u, rhs, omega, nsweeps=1):
        check(self.L.examg_prolong_add_smooth_rbgs(self.h, A, u_coarse, u, rhs, omega, nsweeps))

    def dot(self, a, b, a_slot=-1, b_slot=-1):
        out = C.c_double()
        check(self.L.examg_dot(self.h, a, a_slot, b, b_slot, C.byref(out)))
        return out.value

    def max_abs_diff(self, a, b, a_slot=-1, b_slot=-1):
        out = C.c_double()
        check(self.L.examg_max_abs_diff(self.h, a, a_slot, b, b_slot, C.byref(out)))
        return out.value

    def set(self, f, value, slot=-1):
        check(self.L.examg_set(self.h, f, slot, value))

    def copy(self, src, dst, src_slot=-1, dst_slot=-1):
        check(self.L.examg_copy(self.h, src, src_slot, dst, dst_slot))

    def axpy(self, alpha, x, y, x_slot=-1, y_slot=-1):
        check(self.L.examg_axpy(self.h, alpha, x, x_slot, y, y_slot))

    def xpay(self, x, beta, y, x_slot=-1, y_slot=-1):
        check(self.L.examg_xpay(self.h, x, x_slot, beta, y, y_slot))

    def exchange(self, f, slot=-1, depth=1):
        check(self.L.examg_exchange_depth(self.h, f, slot, depth))

    def cg_coarse(self, A, u, rhs, res, p, ap, max_it, rel_tol, want_iters=True):
        its = C.c_int()
        check(self.L.examg_cg_coarse(self.h, A, u, rhs, res, p, ap, max_it, rel_tol,
                                     C.byref(its) if want_iters else None))
        return its.value

    # -- coupled two-unknown systems on cell fields (optical flow)
    def smoother_coarse(self, A, u, rhs, res, smoother, omega, max_it, rel_tol, tmp=-1):
        it = C.c_int()
        check(self.L.examg_smoother_coarse(self.h, A, u, rhs, res, smoother, omega, tmp, max_it, rel_tol, C.byref(it)))
        return it.value

    def cr_coarse(self, A, u, rhs, res, tmp, max_it, rel_tol):
        """tmp = (p, ap, ar); returns the number of steps."""
        it = C.c_int()
        check(self.L.examg_cr_coarse(self.h, A, u, rhs, res, (C.c_int * 3)(*tmp), max_it, rel_tol, C.byref(it)))
        return it.value

    def minres_coarse(self, A, u, rhs, res, tmp, max_it, rel_tol):
        """tmp = seven work fields; returns the number of steps."""
        it = C.c_int()
        check(self.L.examg_minres_coarse(self.h, A, u, rhs, res, (C.c_int * 7)(*tmp), max_it, rel_tol, C.byref(it)))
        return it.value

    def bicgstab_coarse(self, A, u, rhs, res, tmp, max_it, rel_tol):
        """tmp = (v, p, s, t, resHat); returns the number of steps."""
        it = C.c_int()
        check(self.L.examg_bicgstab_coarse(self.h, A, u, rhs, res, (C.c_int * 5)(*tmp), max_it, rel_tol, C.byref(it)))
        return it.value

    @staticmethod
    def _ints(v):
        return (C.c_int * len(v))(*[int(x) for x in v])

    def system2_create(self, laplace, scale, c_uu, c_uv, c_vv):
        d = System2Desc(laplace, scale, c_uu, c_uv, c_vv)
        return check(self.L.examg_system2_create(self.h, C.byref(d)))

    def system2_residual(self, sys, u, v, rhs_u, rhs_v, res_u, res_v, norm=False):
        if norm:
            n2 = C.c_double()
            check(self.L.examg_system2_residual(self.h, sys, u, v, rhs_u, rhs_v, res_u, res_v, C.byref(n2)))
            return n2.value
        check(self.L.examg_system2_residual(self.h, sys, u, v, rhs_u, rhs_v, res_u, res_v, None))

    def system2_smooth_rb(self, sys, u, v, rhs_u, rhs_v, relax=1.0, nsweeps=1):
        check(self.L.examg_system2_smooth_rb(self.h, sys, u, v, rhs_u, rhs_v, relax, nsweeps))

    def system2_smooth_rb_color(self, sys, u, v, rhs_u, rhs_v, relax, color):
        check(self.L.examg_system2_smooth_rb_color(self.h, sys, u, v, rhs_u, rhs_v, relax, color))

    def restrict_cell(self, fine, coarse):
        check(self.L.examg_restrict_cell(self.h, self._ints(fine), self._ints(coarse), len(fine)))

    def prolong_add_cell(self, coarse, fine):
        check(self.L.examg_prolong_add_cell(self.h, self._ints(coarse), self._ints(fine), len(fine)))

    def system2_cg_coarse(self, sys, fields, max_it, rel_tol, want_iters=True):
        its = C.c_int()
        check(self.L.examg_system2_cg_coarse(self.h, sys, self._ints(fields), max_it, rel_tol,
                                             C.byref(its) if want_iters else None))
        return its.value
