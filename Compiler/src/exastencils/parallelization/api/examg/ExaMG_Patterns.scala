package exastencils.parallelization.api.examg

import scala.collection.mutable.ListBuffer

import exastencils.base.ir._
import exastencils.baseExt.ir.IR_FieldIteratorAccess
import exastencils.baseExt.ir.IR_LoopOverPoints
import exastencils.field.ir._
import exastencils.fieldlike.ir.IR_FieldLikeAccess
import exastencils.operator.ir._

/// ExaMG_Patterns
//
// Shape recognition on the IR *before* IR_ResolveLoopOverPoints (app/ir/IR_LayerHandler.scala:143): loops are still
// IR_LoopOverPoints with their field, condition, reduction and attached communicate statements; `A * u` is still an
// IR_Multiplication of an operator access and a field access (IR_StencilConvolutionWrapper runs later, :152), and
// `diag ( A )` is still a function call (IR_ResolveStencilFunction, :154).  After :208 only `for` loops are left.

object ExaMG_Patterns {

  /// operators

  sealed trait Operator { def name : String; def level : Int; def stencil : IR_Stencil }
  case class ConstOperator(acc : IR_StencilAccess) extends Operator {
    def name = acc.target.name; def level = acc.target.level; def stencil = acc.target
  }
  case class FieldOperator(acc : IR_StencilFieldAccess) extends Operator {
    def name = acc.stencilField.name; def level = acc.stencilField.level; def stencil = acc.stencilField.stencil
  }

  def asOperator(e : IR_Expression) : Option[Operator] = e match {
    case sa : IR_StencilAccess if sa.offset.isEmpty       => Some(ConstOperator(sa))
    case sfa : IR_StencilFieldAccess if sfa.offset.isEmpty => Some(FieldOperator(sfa))
    case _                                                 => None
  }

  def isCentered(fa : IR_FieldLikeAccess) : Boolean =
    fa.offset.forall(_.indices.forall(_ == 0)) && fa.matIndex.isEmpty

  /// inter-grid operators: recognised by their entries (operator/l4/L4_DefaultRestriction.scala:29-90,
  /// L4_DefaultProlongation.scala:30-45), not by name

  /** colStride 2 in every dimension = rows on the coarse, columns on the fine grid: a restriction */
  def isRestriction(st : IR_Stencil) : Boolean = st.colStride.forall(_ == 2.0)
  def isProlongation(st : IR_Stencil) : Boolean = st.colStride.forall(_ == 0.5)

  /// `A * u` with a centred access to u

  case class Convolution(op : Operator, u : IR_FieldLikeAccess)

  def asConvolution(e : IR_Expression) : Option[Convolution] = e match {
    case IR_Multiplication(ListBuffer(a, u : IR_FieldLikeAccess)) if isCentered(u) => asOperator(a).map(Convolution(_, u))
    case _                                                                         => None
  }

  /// `f - A * u`

  case class ResidualExpr(rhs : IR_FieldLikeAccess, op : Operator, u : IR_FieldLikeAccess)

  def asResidual(e : IR_Expression) : Option[ResidualExpr] = e match {
    case IR_Subtraction(f : IR_FieldLikeAccess, prod) if isCentered(f) =>
      asConvolution(prod).map(c => ResidualExpr(f, c.op, c.u))
    case _                                                             => None
  }

  /// `omega / diag ( A ) * ( f - A * u )`, factors in any order and nesting

  case class Relaxation(omega : IR_Expression, res : ResidualExpr)

  private def flattenProduct(e : IR_Expression, num : ListBuffer[IR_Expression], den : ListBuffer[IR_Expression]) : Unit = e match {
    case IR_Multiplication(fs) => fs.foreach(flattenProduct(_, num, den))
    case IR_Division(l, r)     => flattenProduct(l, num, den); flattenProduct(r, den, num)
    case other                 => num += other
  }

  private def isDiagOf(e : IR_Expression, op : Operator) : Boolean = e match {
    case IR_FunctionCall(fct, ListBuffer(arg)) if fct.name == "diag" => asOperator(arg).exists(o => o.name == op.name && o.level == op.level)
    case _                                                             => false
  }

  def asRelaxation(e : IR_Expression) : Option[Relaxation] = {
    val num, den = ListBuffer[IR_Expression]()
    flattenProduct(e, num, den)
    val residuals = num.flatMap(f => asResidual(f).map((f, _)))
    if (residuals.length != 1) return None
    val (resNode, res) = residuals.head
    val diags = den.filter(isDiagOf(_, res.op))
    if (diags.length != 1) return None
    val restNum = num.filterNot(_ eq resNode)
    val restDen = den.filterNot(_ eq diags.head)
    // what is left is the damping factor: constants and scalar variables only
    def scalar(x : IR_Expression) = x match {
      case _ : IR_RealConstant | _ : IR_IntegerConstant | _ : IR_FloatConstant | _ : IR_VariableAccess => true
      case _                                                                                          => false
    }
    if (!(restNum ++ restDen).forall(scalar)) return None
    var omega : IR_Expression = if (restNum.isEmpty) IR_RealConstant(1.0) else IR_Multiplication(restNum)
    if (restDen.nonEmpty) omega = IR_Division(omega, IR_Multiplication(restDen))
    Some(Relaxation(omega, res))
  }

  /// colour conditions: `( i0 + i1 + i2 ) % 2 == c` (red-black) and `i0 % n == cx && i1 % n == cy ...`
  /// (what L4_ColorLoops.toRepeatLoops, baseExt/l4/L4_ColorLoops.scala:44-70, leaves in the loop condition)

  sealed trait Colouring
  case class RedBlack(colour : Int) extends Colouring
  case class PerDimension(n : Int, colour : Array[Int]) extends Colouring

  private def iteratorSum(e : IR_Expression, numDims : Int) : Boolean = e match {
    case IR_Addition(summands) =>
      summands.length == numDims && (0 until numDims).forall(d => summands.contains(IR_FieldIteratorAccess(d)))
    case _                     => false
  }

  private def conjuncts(e : IR_Expression) : List[IR_Expression] = e match {
    case IR_AndAnd(l, r) => conjuncts(l) ++ conjuncts(r)
    case other           => List(other)
  }

  def asColouring(cond : IR_Expression, numDims : Int) : Option[Colouring] = cond match {
    case IR_EqEq(IR_Modulo(sum, IR_IntegerConstant(2)), IR_IntegerConstant(c)) if iteratorSum(sum, numDims) =>
      Some(RedBlack(c.toInt))
    case _                                                                                                  =>
      val cs = conjuncts(cond)
      if (cs.length != numDims) return None
      val colour = Array.fill(numDims)(-1)
      var n = -1
      for (term <- cs) term match {
        case IR_EqEq(IR_Modulo(it : IR_VariableAccess, IR_IntegerConstant(m)), IR_IntegerConstant(c)) =>
          val d = (0 until numDims).indexWhere(IR_FieldIteratorAccess(_) == it)
          if (d < 0 || (n >= 0 && n != m.toInt)) return None
          n = m.toInt
          colour(d) = c.toInt
        case _                                                                                        => return None
      }
      if (colour.contains(-1) || (n != 2 && n != 3)) None else Some(PerDimension(n, colour))
  }

  /// postfix programs (examg_bc_op) of boundary and initialisation expressions

  /** Lowers an expression over vf_nodePosition / vf_boundaryPosition, constants and the L4 math functions to the
    * postfix program of examg_field_set_bc_expr / examg_field_eval; None if something else occurs in it. */
  def toPostfix(e : IR_Expression, fields : Option[ListBuffer[IR_FieldLikeAccess]] = None) : Option[ListBuffer[(String, Double)]] = {
    val prog = ListBuffer[(String, Double)]()
    def emit(x : IR_Expression) : Boolean = x match {
      // examg_loop_eval: a field read at the loop's own point becomes EXAMG_OP_FIELD <index into the source list>
      case fa : IR_FieldLikeAccess if fields.isDefined && isCentered(fa) =>
        val known = fields.get.indexWhere(o => o.field == fa.field && o.slot == fa.slot)
        val idx = if (known >= 0) known else { fields.get += fa; fields.get.length - 1 }
        prog += (("EXAMG_OP_FIELD", idx.toDouble)); idx < 8
      case IR_RealConstant(v)                                            => prog += (("EXAMG_OP_CONST", v)); true
      case IR_IntegerConstant(v)                                         => prog += (("EXAMG_OP_CONST", v.toDouble)); true
      case va : exastencils.grid.ir.IR_VirtualFieldAccess
        if va.target.name.startsWith("vf_nodePosition") || va.target.name.startsWith("vf_boundaryPosition")
          || va.target.name.startsWith("vf_cellCenter")                  =>
        val d = va.target.name.last match { case 'x' => 1; case 'y' => 2; case 'z' => 3; case _ => return false }
        prog += ((List("", "EXAMG_OP_X", "EXAMG_OP_Y", "EXAMG_OP_Z")(d), 0.0)); true
      case IR_Addition(s)                                                =>
        s.nonEmpty && emit(s.head) && s.tail.forall(t => emit(t) && { prog += (("EXAMG_OP_ADD", 0.0)); true })
      case IR_Multiplication(f)                                          =>
        f.nonEmpty && emit(f.head) && f.tail.forall(t => emit(t) && { prog += (("EXAMG_OP_MUL", 0.0)); true })
      case IR_Subtraction(l, r)                                          => emit(l) && emit(r) && { prog += (("EXAMG_OP_SUB", 0.0)); true }
      case IR_Division(l, r)                                             => emit(l) && emit(r) && { prog += (("EXAMG_OP_DIV", 0.0)); true }
      case IR_Negative(v)                                                => emit(v) && { prog += (("EXAMG_OP_NEG", 0.0)); true }
      case IR_Power(b, IR_IntegerConstant(2))                            => emit(b) && emit(b) && { prog += (("EXAMG_OP_MUL", 0.0)); true }
      case IR_FunctionCall(fct, ListBuffer(arg))
        if Set("exp", "sin", "cos", "sinh", "sqrt").contains(fct.name) =>
        emit(arg) && { prog += (("EXAMG_OP_" + fct.name.toUpperCase, 0.0)); true }
      case _                                                             => false
    }
    if (emit(e)) Some(prog) else None
  }

  /// the one statement of a loop body, if there is exactly one

  def singleStatement(loop : IR_LoopOverPoints) : Option[IR_Statement] =
    if (loop.body.length == 1) Some(loop.body.head) else None
}
