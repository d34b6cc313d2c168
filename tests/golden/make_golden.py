#!/usr/bin/env python
"""Copies the reference's own golden stdout vectors (Testing/**/*.results) into
tests/golden/.  Run in the build container (where /root/reference is mounted);
the fixtures are committed because /root/reference does not exist on the GPU box.

These files are data (expected residual / error histories printed with
printWithReducedPrec, 4 significant digits), not source code.
"""
import os
import shutil
import sys

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))

FILES = {
    "Poisson_2D_FD_Poisson_fromL4": "Testing/Poisson/2D_FD_Poisson_fromL4.results",
    "Smoothers_RBGS": "Testing/Smoothers/RBGS.results",
    "Smoothers_Jac": "Testing/Smoothers/Jac.results",
    "SISC_3D_ConstCoeff": "Testing/SISC/3D_ConstCoeff.results",
    "SISC_3D_VarCoeff": "Testing/SISC/3D_VarCoeff.results",
    "SISC_2D_ConstCoeff": "Testing/SISC/2D_ConstCoeff.results",
    "SISC_2D_VarCoeff": "Testing/SISC/2D_VarCoeff.results",
    "Opts_seq": "Testing/Opts/seq.results",
    "Opts_par": "Testing/Opts/par.results",
    "CommBasic_2D": "Testing/CommBasic/2D.results",
    "CommBasic_HybridCell": "Testing/CommBasic/HybridCell.results",
    "CommBasic_Strategy26Cell": "Testing/CommBasic/Strategy26Cell.results",
    "CellBased_2D_Basic": "Testing/CellBased/2D_Basic.results",
    "CellBased_2D_Neumann": "Testing/CellBased/2D_Neumann.results",
    "CellBased_3D_Basic": "Testing/CellBased/3D_Basic.results",
    "CellBased_3D_Neumann": "Testing/CellBased/3D_Neumann.results",
    "CommBasic_PureMPI": "Testing/CommBasic/PureMPI.results",
    "CommBasic_PureOMP": "Testing/CommBasic/PureOMP.results",
    "CommBasic_Hybrid": "Testing/CommBasic/Hybrid.results",
    "CommBasic_Strategy26": "Testing/CommBasic/Strategy26.results",
    "CommBasic_Summarize": "Testing/CommBasic/Summarize.results",
    "Poisson_2D_FD_Poisson_fromL2": "Testing/Poisson/2D_FD_Poisson_fromL2.results",
    "Poisson_2D_FD_Poisson_fromL3": "Testing/Poisson/2D_FD_Poisson_fromL3.results",
    "Poisson_2D_FV_Poisson_fromL4": "Testing/Poisson/2D_FV_Poisson_fromL4.results",
    "Poisson_2D_FV_Poisson_fromL2": "Testing/Poisson/2D_FV_Poisson_fromL2.results",
    "CUDA_2D_VarCoeff": "Testing/CUDA/2D_VarCoeff.results",
    "FMG_2D_ConstCoeff": "Testing/FMG/2D_ConstCoeff.results",
    "FMG_2D_Polynomial": "Testing/FMG/2D_Polynomial.results",
    "FMG_3D_Trigonometric": "Testing/FMG/3D_Trigonometric.results",
    "FMG_3D_VarCoeff": "Testing/FMG/3D_VarCoeff.results",
    "Large_Hybrid": "Testing/Large/Hybrid.results",
    "Application_OpticalFlow2D": "Testing/Application/OpticalFlow2D.results",
    "Application_OpticalFlow3D": "Testing/Application/OpticalFlow3D.results",
    "BC_2D_Polynomial": "Testing/BC/2D_Polynomial.results",
    "BC_2D_Trigonometric": "Testing/BC/2D_Trigonometric.results",
    "BC_3D_Polynomial": "Testing/BC/3D_Polynomial.results",
    "BC_2D_Periodic": "Testing/BC/2D_Periodic.results",
    "BC_3D_Periodic": "Testing/BC/3D_Periodic.results",
    "BC_3D_Trigonometric": "Testing/BC/3D_Trigonometric.results",
    "NonLinear_FAS_2D_Basic": "Testing/NonLinear/FAS_2D_Basic.results",
    "Application_ExaStokes_2D": "Testing/Application/ExaStokes_2D.results",
    "Application_ExaStokes_3D": "Testing/Application/ExaStokes_3D.results",
}

if __name__ == "__main__":
    for name, rel in FILES.items():
        shutil.copyfile(os.path.join(REF, rel), os.path.join(HERE, name + ".results"))
        print("copied", rel)
