//=============================================================================
//
//  libexamg back-end for the ExaStencils generator (overlay: copy this tree over
//  the reference's Compiler/ directory, apply Compiler/patches/*.patch, run sbt).
//
//  Written against the reference's IR as of the revision surveyed in SURVEY.md;
//  not compiled in the repository that ships it (no JVM there) -- see
//  Compiler/README.md for what is and is not verified.
//
//=============================================================================

package exastencils.parallelization.api.examg

import scala.collection.mutable.ListBuffer

import exastencils.base.ir._
import exastencils.prettyprinting.PpStream

/// ExaMG_Symbols

/** Every entry point of include/examg.h this back-end prints.  tests/test_scala_backend.py parses this table and
  * requires (a) every name to be declared in the header and (b) every hot-path entry point of the header to be here. */
object ExaMG_Symbols {
  // lifecycle
  val init = "examg_init"
  val finalizeCtx = "examg_finalize"
  val sync = "examg_sync"
  val lastError = "examg_last_error"
  // fields and operators
  val fieldCreate = "examg_field_create"
  val fieldCreateCell = "examg_field_create_cell"
  val fieldSetBC = "examg_field_set_bc"
  val fieldSetBCExpr = "examg_field_set_bc_expr"
  val fieldSetBCFace = "examg_field_set_bc_face"
  val fieldEval = "examg_field_eval"
  val loopEval = "examg_loop_eval"
  val fieldFill = "examg_field_fill"
  val fieldUpload = "examg_field_upload"
  val fieldDownload = "examg_field_download"
  val fieldAdvance = "examg_field_advance"
  val fieldPrint = "examg_field_print"
  val fieldWrite = "examg_field_write"
  val fieldRead = "examg_field_read"
  val stencilConst = "examg_stencil_const"
  val stencilField = "examg_stencil_field"
  // loops
  val apply = "examg_apply"
  val residual = "examg_residual"
  val residualRestrict = "examg_residual_restrict"
  val smoothJacobi = "examg_smooth_jacobi"
  val smoothRBGS = "examg_smooth_rbgs"
  val smoothRBColor = "examg_smooth_rb_color"
  val smoothColored = "examg_smooth_colored"
  val restrict = "examg_restrict"
  val prolongAdd = "examg_prolong_add"
  val prolongAddSmoothRBGS = "examg_prolong_add_smooth_rbgs"
  val restrictCell = "examg_restrict_cell"
  val prolongAddCell = "examg_prolong_add_cell"
  val dot = "examg_dot"
  val maxAbsDiff = "examg_max_abs_diff"
  val set = "examg_set"
  val copy = "examg_copy"
  val axpy = "examg_axpy"
  val xpay = "examg_xpay"
  val exchange = "examg_exchange"
  val applyBC = "examg_apply_bc"
  // coarse-grid solvers
  val cgCoarse = "examg_cg_coarse"
  val bicgstabCoarse = "examg_bicgstab_coarse"
  val crCoarse = "examg_cr_coarse"
  val minresCoarse = "examg_minres_coarse"
  val smootherCoarse = "examg_smoother_coarse"
  // coupled two-unknown systems (solve locally)
  val system2Create = "examg_system2_create"
  val system2Residual = "examg_system2_residual"
  val system2SmoothRB = "examg_system2_smooth_rb"
  val system2SmoothRBColor = "examg_system2_smooth_rb_color"
  val system2CGCoarse = "examg_system2_cg_coarse"
  // staggered-grid Stokes (ExaMG_Stokes)
  val stokesCreate = "examg_stokes_create"
  val stokesSmoothColor = "examg_stokes_smooth_color"
  val stokesSmooth = "examg_stokes_smooth"
  val stokesResidual = "examg_stokes_residual"
  val stokesRestrict = "examg_stokes_restrict"
  val stokesSetSolution = "examg_stokes_set_solution"
  val stokesCorrect = "examg_stokes_correct"
  val stokesApplyBC = "examg_stokes_apply_bc"
  val stokesVCycle = "examg_stokes_vcycle"
  // whole cycle
  val cycleDescFromOptions = "examg_cycle_desc_from_options"
  val solverCreate = "examg_solver_create"
  val solverCycles = "examg_solver_cycles"
  val solverVCycle = "examg_solver_vcycle"
  val solverField = "examg_solver_field"
  val solverResidualNorm = "examg_solver_residual_norm"
  // multi-GPU wiring
  val commBlobSize = "examg_comm_blob_size"
  val commExport = "examg_comm_export"
  val commImport = "examg_comm_import"
  // timers
  val timerStart = "examg_timer_start"
  val timerStop = "examg_timer_stop"
  val timerTotalMs = "examg_timer_total_ms"
  val timerPrintAll = "examg_timer_print_all"
  val benchmarkStart = "examg_benchmark_start"
  val benchmarkStop = "examg_benchmark_stop"
}

/// ExaMG_Context

/** The `examg_ctx *ctx` global of Global/Global.h. */
object ExaMG_Context {
  def name = "examgCtx"
  def datatype = IR_PointerDatatype(IR_SpecialDatatype("examg_ctx"))
  def access = IR_VariableAccess(name, datatype)
  def declaration = IR_VariableDeclaration(datatype, name, IR_VariableAccess("NULL", datatype))
}

/// ExaMG_Call

/** One call into libexamg.  Printed as `EXAMG_CHECK(<name>(examgCtx, <args>));` -- the macro (ExaMG_Globals) prints
  * examg_last_error() and exits, which is the reference's own error convention for generated code
  * (IR_Assert -> exit(1), parallelization/api/cuda/CUDA_Error.scala:39-68). */
case class ExaMG_Call(var name : String, var arguments : ListBuffer[IR_Expression], var withContext : Boolean = true) extends IR_Statement {
  override def prettyprint(out : PpStream) : Unit = {
    out << "EXAMG_CHECK(" << name << '('
    if (withContext) {
      out << ExaMG_Context.name
      if (arguments.nonEmpty) out << ", "
    }
    out <<< (arguments, ", ") << "));"
  }
}

object ExaMG_Call {
  def apply(name : String, args : IR_Expression*) = new ExaMG_Call(name, args.to[ListBuffer])
}

/// ExaMG_HandleAccess

/** Integer handle of a field (`f_<name>_<level>`) or operator (`op_<name>_<level>`) created in setupBuffers. */
case class ExaMG_HandleAccess(var kind : String, var objName : String, var level : Int) extends IR_Expression {
  def handleName = s"${ kind }_${ objName }_$level"
  override def datatype = IR_IntegerDatatype
  override def prettyprint(out : PpStream) : Unit = out << handleName
  def declaration = IR_VariableDeclaration(IR_IntegerDatatype, handleName, IR_IntegerConstant(-1))
}

object ExaMG_FieldHandle {
  def apply(field : exastencils.fieldlike.ir.IR_FieldLike) = ExaMG_HandleAccess("f", field.name, field.level)
}

object ExaMG_StencilHandle {
  def apply(name : String, level : Int) = ExaMG_HandleAccess("op", name, level)
}

/// ExaMG_AddressOf

/** `&x` of a host scalar that receives a reduction result. */
case class ExaMG_AddressOf(var target : IR_Expression) extends IR_Expression {
  override def datatype = IR_PointerDatatype(target.datatype)
  override def prettyprint(out : PpStream) : Unit = out << "&(" << target << ')'
}
