package exastencils.parallelization.api.examg

import scala.collection.mutable.ListBuffer

import exastencils.base.ir._
import exastencils.fieldlike.ir.IR_FieldLikeAccess
import exastencils.grid.ir._
import exastencils.solver.ir.IR_LocalSolve

/// ExaMG_Stokes
//
// `color with { i0 % 3, i1 % 3 [, i2 % 3] } loop over p { solve locally { u@[0,0] => ..., u@[1,0] => ..., v@[0,0] => ...,
//  v@[0,1] => ..., [w@[0,0,0] => ..., w@[0,0,1] => ...,] p@[0,0] => ... } }`
// (Testing/Application/ExaStokes_2D.exa4:228-303, ExaStokes_3D.exa4:272-381, Examples/Stokes/*_FD_Stokes_fromL4.exa4): the
// block smoother of the staggered-grid Stokes programs -- per dimension d the two faces of the cell in a Face_d field, and
// the cell's value of a Cell field.  One colour becomes examg_stokes_smooth_color(st, level, c0, c1, c2); the descriptor
// `st` (examg_stokes_create) is created once in setupBuffers (ExaMG_Globals) from the sweep counts of the cycle functions.
// The component-wise loops of the program (UpdateResidual_*, Restriction_*, Correction_*, SetSolution_*, apply bc) map to
// examg_stokes_residual / _restrict / _correct / _set_solution / _apply_bc: all components in one call.

object ExaMG_Stokes {

  case class Recognised(level : Int, dim : Int, velocities : ListBuffer[IR_FieldLikeAccess], pressure : IR_FieldLikeAccess) {
    def handle = IR_VariableAccess("stokes_", IR_SpecialDatatype("examg_stokes*"))
  }

  var found : Option[Recognised] = None

  private def offsetOf(fa : IR_FieldLikeAccess, dim : Int) : Array[Long] =
    (0 until dim).map(d => fa.offset.map(_.indices(d)).getOrElse(0L)).toArray

  /** 2 dim + 1 unknowns: for d = 0 .. dim-1 a Face_d field at the cell and one face further in dimension d, then a Cell field */
  def recognise(ls : IR_LocalSolve) : Option[Recognised] = {
    val dim = ls.unknowns.head.field.layout.numDimsGrid
    if (ls.unknowns.length != 2 * dim + 1 || ls.equations.length != 2 * dim + 1 || ls.jacobiType || ls.relax.isDefined) return None
    val vel = ListBuffer[IR_FieldLikeAccess]()
    for (d <- 0 until dim) {
      val (a, b) = (ls.unknowns(2 * d), ls.unknowns(2 * d + 1))
      if (a.field != b.field || a.field.layout.localization != IR_AtFaceCenter(d)) return None
      val (oa, ob) = (offsetOf(a, dim), offsetOf(b, dim))
      if (oa.exists(_ != 0) || (0 until dim).exists(e => ob(e) != (if (e == d) 1 else 0))) return None
      vel += a
    }
    val p = ls.unknowns.last
    if (p.field.layout.localization != IR_AtCellCenter || offsetOf(p, dim).exists(_ != 0)) return None
    val r = Recognised(p.field.level, dim, vel, p)
    if (found.isEmpty) found = Some(r)
    Some(r)
  }

  /** one colour of the 3^dim-way colouring around a recognised local solve */
  def lower(ls : IR_LocalSolve, colouring : Option[ExaMG_Patterns.Colouring]) : Option[IR_Statement] =
    (colouring, recognise(ls)) match {
      case (Some(ExaMG_Patterns.PerDimension(3, colour)), Some(st)) =>
        val c = (0 until 3).map(d => IR_IntegerConstant(if (d < st.dim) colour(d) else 0))
        Some(ExaMG_Call(ExaMG_Symbols.stokesSmoothColor, st.handle, IR_IntegerConstant(st.level), c(0), c(1), c(2)))
      case _                                                        => None
    }
}
