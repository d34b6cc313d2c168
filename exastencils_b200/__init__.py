"""exastencils_b200 -- host side of libexamg, the B200-native runtime behind the
geometric-multigrid V-cycle the ExaStencils generator emits.

  runtime.py  ctypes binding of the C ABI (include/examg.h)
  program.py  the generated solver's leveled functions (Solve, mgCycle, ResNorm ...)
  optflow.py  the optical-flow programs (cell fields, Neumann, 2x2 block smoother)
  stokes.py   the staggered-grid Stokes program (face-centred velocities, 5x5 block smoother)
  optflow3.py the 3D optical-flow program (Vec3 unknown, 3x3 block inverse per cell)
  csrc/       CUDA sources of libexamg.so (sm_100a), built in-tree into lib/

There is no CPU implementation in this package; the CPU oracle lives in oracle/ and is
test infrastructure only.
"""
from .runtime import ExaMGError, Context, build, lib  # noqa: F401
from .program import MultigridProgram, laplace_stencil  # noqa: F401
from .optflow import OpticalFlowProgram  # noqa: F401
from .stokes import StokesProgram  # noqa: F401
from .optflow3 import OpticalFlow3DProgram  # noqa: F401
