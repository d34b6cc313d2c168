package exastencils.parallelization.api.examg

import scala.collection.mutable.ListBuffer

import exastencils.base.ir._
import exastencils.fieldlike.ir.IR_FieldLikeAccess
import exastencils.solver.ir.IR_LocalSolve

/// ExaMG_System2
//
// `solve locally relax r { u => c_uu * u + c_uv * v + a * ( L * u ) == rhs_u; v => c_uv * u + c_vv * v + a * ( L * v ) == rhs_v }`
// (Benchmark/OptFlow2D/2D_FD_OptFlow.exa4 mgCycle; generator side solver/ir/IR_LocalSolve.scala:88-236): the coupled
// two-unknown operator of examg_system2_create.  The descriptor is created once per level in setupBuffers
// (ExaMG_Globals) and referred to by handle `sys_<level>`.

object ExaMG_System2 {

  case class Recognised(level : Int, laplace : ExaMG_Patterns.Operator, scale : IR_Expression,
      cuu : IR_FieldLikeAccess, cuv : IR_FieldLikeAccess, cvv : IR_FieldLikeAccess, rhsU : IR_Expression, rhsV : IR_Expression) {
    def handle = IR_VariableAccess(s"sys_$level", IR_IntegerDatatype)
  }

  val found = ListBuffer[Recognised]()

  private def terms(e : IR_Expression) : ListBuffer[IR_Expression] = e match {
    case IR_Addition(s) => s.flatMap(terms)
    case other          => ListBuffer(other)
  }

  /** (coefficient field, unknown) of a term `c * u`, or (scale, operator, unknown) of `a * ( L * u )` */
  private def split(t : IR_Expression) : Option[Either[(IR_FieldLikeAccess, IR_FieldLikeAccess), (IR_Expression, ExaMG_Patterns.Operator, IR_FieldLikeAccess)]] = t match {
    case IR_Multiplication(ListBuffer(c : IR_FieldLikeAccess, u : IR_FieldLikeAccess)) => Some(Left((c, u)))
    case IR_Multiplication(ListBuffer(a, conv)) if ExaMG_Patterns.asConvolution(conv).isDefined =>
      val c = ExaMG_Patterns.asConvolution(conv).get
      Some(Right((a, c.op, c.u)))
    case conv if ExaMG_Patterns.asConvolution(conv).isDefined                             =>
      val c = ExaMG_Patterns.asConvolution(conv).get
      Some(Right((IR_RealConstant(1.0), c.op, c.u)))
    case _                                                                                => None
  }

  def recognise(ls : IR_LocalSolve) : Option[Recognised] = {
    if (ls.unknowns.length != 2 || ls.equations.length != 2) return None
    val (u, v) = (ls.unknowns(0), ls.unknowns(1))
    var lap : Option[(IR_Expression, ExaMG_Patterns.Operator)] = None
    val coef = Array.ofDim[Option[IR_FieldLikeAccess]](2, 2)
    for (i <- 0 until 2; j <- 0 until 2) coef(i)(j) = None
    val rhs = Array.ofDim[IR_Expression](2)
    for ((eq, row) <- ls.equations.zipWithIndex) {
      rhs(row) = eq.rhs
      for (t <- terms(eq.lhs)) split(t) match {
        case Some(Left((c, x)))      =>
          val col = if (x.field == u.field) 0 else if (x.field == v.field) 1 else return None
          coef(row)(col) = Some(c)
        case Some(Right((a, op, x))) =>
          if (x.field != ls.unknowns(row).field) return None // the Laplacian couples an unknown with itself only
          if (lap.exists(l => l._2.name != op.name)) return None
          lap = Some((a, op))
        case None                    => return None
      }
    }
    if (lap.isEmpty || coef.flatten.exists(_.isEmpty)) return None
    if (coef(0)(1).get.field != coef(1)(0).get.field) return None // symmetric coupling (uvSten = vuSten)
    val r = Recognised(u.field.level, lap.get._2, lap.get._1, coef(0)(0).get, coef(0)(1).get, coef(1)(1).get,
      rhs(0) match { case f : IR_FieldLikeAccess => ExaMG_FieldHandle(f.field); case _ => return None },
      rhs(1) match { case f : IR_FieldLikeAccess => ExaMG_FieldHandle(f.field); case _ => return None })
    if (!found.exists(_.level == r.level)) found += r
    Some(r)
  }
}
