package exastencils.parallelization.api.examg

import scala.collection.mutable.ListBuffer

import exastencils.base.ir._
import exastencils.communication.ir.IR_Communicate
import exastencils.datastructures._

/// ExaMG_MergeCalls
//
// Peephole pass over statement lists after ExaMG_ResolveCalls.  The L4 text spells the smoother out colour by colour and
// sweep by sweep (`repeat nu times { color with { ..., communicate; loop; apply bc } }`, already unrolled by
// L4_ResolveColorLoops and IR_ResolveRepeat at this point); the library wants to see the whole smoothing step, because
// it fuses colours and sweeps into one pass over HBM (k_rbw: 2 sweeps = 4 half-sweeps per pass, one 4-deep exchange).
//   colour 0, colour 1                         -> examg_smooth_rbgs(A, u, f, omega, 1)
//   k x examg_smooth_rbgs(.., 1)               -> examg_smooth_rbgs(.., k)          (same for Jacobi sweeps)
//   all n^d colours in canonical order         -> examg_smooth_colored(A, u, f, omega, n, 1)
//   prolongation, (apply bc), red-black sweeps -> examg_prolong_add_smooth_rbgs     (the interpolation rides on the first pass)
//   residual, (apply bc), restriction          -> examg_residual_restrict           (Residual has no other reader)
//   residual, (apply bc), norm reduction       -> examg_residual(..., &norm)        (fused sum of squares)
//   one transfer loop per unknown (u, v)       -> one examg_restrict_cell / examg_prolong_add_cell with nfields = 2
// The `communicate` and `apply bc` statements between the merged calls belong to them and are dropped
// (include/examg.h documents each entry point as { communicate; update; apply bc }).

object ExaMG_MergeCalls extends DefaultStrategy("Merge libexamg calls") {

  import ExaMG_Symbols._

  /** `communicate F` in front of a call that reads F: done by the call itself */
  def impliedBy(comm : IR_Communicate, call : IR_Statement) : Boolean = call match {
    case ExaMG_Call(name, args, _) if Set(smoothRBColor, smoothRBGS, smoothJacobi, smoothColored, residual, residualRestrict, apply,
      restrict, system2SmoothRB, system2SmoothRBColor, system2Residual).contains(name) =>
      args.contains(ExaMG_FieldHandle(comm.field))
    case j : ExaMG_JacobiSweep                                                         => j.u == ExaMG_FieldHandle(comm.field)
    case cs : ExaMG_ColourStep                                                         => cs.u == ExaMG_FieldHandle(comm.field)
    case _                                                                             => false
  }

  private def isGlue(s : IR_Statement, field : IR_Expression) : Boolean = s match {
    case ExaMG_Call(`exchange`, args, _) => args.head == field
    case ExaMG_Call(`applyBC`, args, _)  => args.head == field
    case IR_NullStatement                => true
    case _                               => false
  }

  private def sameSmoother(a : ListBuffer[IR_Expression], b : ListBuffer[IR_Expression]) = a.take(4) == b.take(4)

  /** one left-to-right pass; returns true if something was merged */
  private def mergeOnce(stmts : ListBuffer[IR_Statement]) : Boolean = {
    var i = 0
    while (i < stmts.length) {
      // index of the next statement that is not glue of `field`
      def next(from : Int, field : IR_Expression) : Int = {
        var j = from
        while (j < stmts.length && isGlue(stmts(j), field)) j += 1
        j
      }
      stmts(i) match {
        // red + black = one sweep
        case ExaMG_Call(`smoothRBColor`, a, _) if a(4) == IR_IntegerConstant(0) =>
          val j = next(i + 1, a(1))
          if (j < stmts.length) stmts(j) match {
            case ExaMG_Call(`smoothRBColor`, b, _) if sameSmoother(a, b) && b(4) == IR_IntegerConstant(1) =>
              val end = next(j + 1, a(1))
              stmts.remove(i, end - i)
              stmts.insert(i, ExaMG_Call(smoothRBGS, a(0), a(1), a(2), a(3), IR_IntegerConstant(1)))
              return true
            case _                                                                                        =>
          }
        // k sweeps + 1 sweep
        case ExaMG_Call(name, a, _) if name == smoothRBGS || name == smoothJacobi || name == smoothColored =>
          val j = next(i + 1, a(1))
          if (j < stmts.length) stmts(j) match {
            case ExaMG_Call(`name`, b, _) if sameSmoother(a, b) && a.length == b.length && a.drop(4).dropRight(1) == b.drop(4).dropRight(1) =>
              (a.last, b.last) match {
                case (IR_IntegerConstant(k), IR_IntegerConstant(m)) =>
                  stmts.remove(i + 1, j - i)
                  a(a.length - 1) = IR_IntegerConstant(k + m)
                  return true
                case _                                              =>
              }
            case _                                                                                                                        =>
          }
        // Jacobi sweep + advance
        case js : ExaMG_JacobiSweep =>
          val j = next(i + 1, js.u)
          if (j < stmts.length) stmts(j) match {
            case ExaMG_Call(`fieldAdvance`, adv, _) if adv.head == js.u =>
              stmts.remove(i, j - i + 1)
              stmts.insert(i, ExaMG_Call(smoothJacobi, js.op, js.u, js.rhs, js.omega, IR_IntegerConstant(1)))
              return true
            case _                                                      =>
          }
        // n^d colours in canonical order (last dimension slowest, baseExt/l4/L4_ColorLoops.scala:44-70)
        case first : ExaMG_ColourStep if first.colour.forall(_ == 0) =>
          val total = math.pow(first.n, first.colour.length).toInt
          var j = i
          var count = 0
          var ok = true
          while (ok && count < total && j < stmts.length) stmts(j) match {
            case cs : ExaMG_ColourStep if cs.op == first.op && cs.u == first.u && cs.n == first.n =>
              val want = (0 until cs.colour.length).map(d => (count / math.pow(first.n, d).toInt) % first.n)
              ok = cs.colour.toSeq == want
              if (ok) { count += 1; j = next(j + 1, first.u) }
            case _                                                                                 => ok = false
          }
          if (ok && count == total) {
            stmts.remove(i, j - i)
            stmts.insert(i, ExaMG_Call(smoothColored, first.op, first.u, first.rhs, first.omega, IR_IntegerConstant(first.n), IR_IntegerConstant(1)))
            return true
          }
        // correction followed by the post-smoothing of the same field (after the sweeps were merged into one call)
        case ExaMG_Call(`prolongAdd`, pa, _) if pa(1) == IR_IntegerConstant(-1) && pa(3) == IR_IntegerConstant(-1) =>
          val j = next(i + 1, pa(2))
          if (j < stmts.length) stmts(j) match {
            case ExaMG_Call(`smoothRBGS`, sm, _) if sm(1) == pa(2) =>
              // wait until the sweeps behind it are merged: the merged call takes the sweep count
              val jj = next(j + 1, pa(2))
              val more = jj < stmts.length && (stmts(jj) match {
                case ExaMG_Call(`smoothRBGS`, b, _)    => sameSmoother(sm, b)
                case ExaMG_Call(`smoothRBColor`, b, _) => sameSmoother(sm, b)
                case _                                 => false
              })
              if (!more) {
                stmts.remove(i, j - i + 1)
                stmts.insert(i, ExaMG_Call(prolongAddSmoothRBGS, sm(0), pa(0), pa(2), sm(2), sm(3), sm(4)))
                return true
              }
            case _                                                 =>
          }
        // residual (+ apply bc) followed by the restriction of the same field: one pass
        case ExaMG_Call(`residual`, r, _) if r.last.isInstanceOf[IR_VariableAccess] /* NULL norm */ =>
          val j = next(i + 1, r(4))
          if (j < stmts.length) stmts(j) match {
            case ExaMG_Call(`restrict`, t, _) if t(0) == r(4) =>
              stmts.remove(i, j - i + 1)
              stmts.insert(i, ExaMG_Call(residualRestrict, r(0), r(1), r(3), r(4), t(1), t(2)))
              return true
            case ExaMG_Call(`dot`, d, _) if d(0) == r(4) && d(2) == r(4) =>
              stmts.remove(i, j - i + 1)
              stmts.insert(i, ExaMG_Call(residual, r(0), r(1), r(2), r(3), r(4), d(4)))
              return true
            case _                                                         =>
          }
        // one cell-transfer loop per unknown -> one call with both
        case a : ExaMG_CellTransfer if i + 1 < stmts.length =>
          stmts(i + 1) match {
            case b : ExaMG_CellTransfer if b.name == a.name =>
              stmts.remove(i, 2)
              stmts.insert(i, ExaMG_CellTransfer2(a.name, a.from, b.from, a.to, b.to))
              return true
            case _                                          =>
          }
        case _ =>
      }
      i += 1
    }
    false
  }

  private def merge(stmts : ListBuffer[IR_Statement]) : ListBuffer[IR_Statement] = {
    while (mergeOnce(stmts)) {}
    stmts
  }

  this += new Transformation("Function bodies and nested scopes", {
    case fct : IR_Function           => merge(fct.body); fct
    case scope : IR_Scope            => merge(scope.body); scope
    case loop : IR_ForLoop           => merge(loop.body); loop
    case loop : IR_WhileLoop         => merge(loop.body); loop
    case cond : IR_IfCondition       => merge(cond.trueBody); merge(cond.falseBody); cond
  })
}

case class ExaMG_CellTransfer2(var name : String, var from0 : IR_Expression, var from1 : IR_Expression, var to0 : IR_Expression, var to1 : IR_Expression) extends IR_Statement {
  override def prettyprint(out : exastencils.prettyprinting.PpStream) : Unit = {
    out << "{ const int from_[2] = { " << from0 << ", " << from1 << " }, to_[2] = { " << to0 << ", " << to1 << " }; "
    out << ExaMG_Call(name, IR_VariableAccess("from_", IR_IntegerDatatype), IR_VariableAccess("to_", IR_IntegerDatatype), IR_IntegerConstant(2)) << " }"
  }
}
