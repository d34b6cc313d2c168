package exastencils.parallelization.api.examg

import scala.collection.mutable.ListBuffer

import exastencils.base.ir._
import exastencils.baseExt.ir._
import exastencils.boundary.ir.IR_ApplyBC
import exastencils.communication.ir.IR_Communicate
import exastencils.config.Knowledge
import exastencils.datastructures._
import exastencils.field.ir._
import exastencils.fieldlike.ir._
import exastencils.logger.Logger
import exastencils.solver.ir.IR_LocalSolve
import exastencils.util.ir.IR_Print

/// ExaMG_ResolveCalls
//
// Lowers the loop nests of the hot path to calls into libexamg.  Runs right before IR_ResolveLoopOverPoints
// (app/ir/IR_LayerHandler.scala:143).  One case per row of INTEGRATION.md section 2; a loop that matches no case is left
// to the reference's own loop printer and reported (ExaMG_ReportLeftovers) -- it then works on host copies obtained
// with examg_field_download, which is slow and meant to be noticed.

object ExaMG_ResolveCalls extends DefaultStrategy("Lower loops of the multigrid hot path to libexamg calls") {

  import ExaMG_Patterns._
  import ExaMG_Symbols._

  private def slotOf(fa : IR_FieldLikeAccess) : IR_Expression = fa.slot match {
    case IR_SlotAccess(_, 0)   => IR_IntegerConstant(-1) // active slot
    case IR_SlotAccess(iv, o)  => (IR_IV_ActiveSlot(iv.field) + o) Mod iv.field.numSlots
    case IR_IntegerConstant(s) => IR_IntegerConstant(s)
    case other                 => other
  }

  private def handle(fa : IR_FieldLikeAccess) = ExaMG_FieldHandle(fa.field)
  private def handle(op : Operator) = ExaMG_StencilHandle(op.name, op.level)
  private def nullPtr = IR_VariableAccess("NULL", IR_PointerDatatype(IR_RealDatatype))
  private def isNextSlot(fa : IR_FieldLikeAccess) = fa.slot match {
    case IR_SlotAccess(_, 1) => true
    case _                   => false
  }

  /** the statement a loop over `field` with this single-statement body lowers to */
  private def lower(loop : IR_LoopOverPoints) : Option[IR_Statement] = {
    val numDims = loop.field.layout.numDimsGrid
    val colouring = loop.condition.flatMap(asColouring(_, numDims))
    if (loop.condition.isDefined && colouring.isEmpty && loop.parallelization.reduction.isEmpty) return None
    if (loop.region.isDefined) return None // `only ghost/dup ... on boundary` loops: user-written boundary functions

    singleStatement(loop) match {
      // ---- smoothers: u += omega / diag ( A ) * ( f - A * u ) ------------------------------------------------
      case Some(IR_Assignment(dst : IR_FieldLikeAccess, src, "+=")) if isCentered(dst) && asRelaxation(src).exists(_.res.u.field == dst.field) =>
        val r = asRelaxation(src).get
        colouring match {
          case Some(RedBlack(c))         => Some(ExaMG_Call(smoothRBColor, handle(r.res.op), handle(dst), handle(r.res.rhs), r.omega, c))
          case Some(PerDimension(n, cs)) => Some(ExaMG_ColourStep(handle(r.res.op), handle(dst), handle(r.res.rhs), r.omega, n, cs))
          case None                      => None // lexicographic Gauss-Seidel: depends on the traversal order, not offered
        }

      // ---- Jacobi: u<next> = u<active> + omega / diag ( A ) * ( f - A * u<active> ) --------------------------
      case Some(IR_Assignment(dst : IR_FieldLikeAccess, IR_Addition(ListBuffer(u0 : IR_FieldLikeAccess, upd)), "=")) if colouring.isEmpty && isNextSlot(dst) && u0.field == dst.field && asRelaxation(upd).exists(_.res.u.field == dst.field) =>
        val r = asRelaxation(upd).get
        Some(ExaMG_JacobiSweep(handle(r.res.op), handle(dst), handle(r.res.rhs), r.omega))

      // ---- prolongation + correction: u += P@coarser * u@coarser ----------------------------------------------
      case Some(IR_Assignment(dst : IR_FieldLikeAccess, src, "+=")) if asConvolution(src).exists(c => isProlongation(c.op.stencil)) =>
        val c = asConvolution(src).get
        val cell = dst.field.layout.localization == exastencils.grid.ir.IR_AtCellCenter
        if (cell) Some(ExaMG_CellTransfer(prolongAddCell, handle(c.u), handle(dst)))
        else Some(ExaMG_Call(prolongAdd, handle(c.u), slotOf(c.u), handle(dst), slotOf(dst)))

      // ---- residual / restriction / operator application: r = f - A * u, f@coarser = R * r, v = A * u ----------
      case Some(IR_Assignment(dst : IR_FieldLikeAccess, src, "=")) if colouring.isEmpty && isCentered(dst) =>
        asResidual(src) match {
          case Some(res) =>
            Some(ExaMG_Call(residual, handle(res.op), handle(res.u), slotOf(res.u), handle(res.rhs), handle(dst), nullPtr))
          case None      => asConvolution(src) match {
            case Some(c) if isRestriction(c.op.stencil) =>
              val cell = dst.field.layout.localization == exastencils.grid.ir.IR_AtCellCenter
              if (cell) Some(ExaMG_CellTransfer(restrictCell, handle(c.u), handle(dst)))
              else Some(ExaMG_Call(restrict, handle(c.u), handle(dst), IR_RealConstant(1.0)))
            case Some(c)                                =>
              Some(ExaMG_Call(apply, handle(c.op), handle(c.u), slotOf(c.u), handle(dst)))
            case None                                   => lowerStreaming(dst, src)
          }
        }

      // ---- scaled restriction: f@coarser = 4.0 * ( R * r )  (Testing/Smoothers) --------------------------------
      // handled by lowerStreaming's caller below when the product has a scalar factor

      // ---- axpy family with compound assignment -------------------------------------------------------------------
      case Some(IR_Assignment(dst : IR_FieldLikeAccess, src, "+=")) if colouring.isEmpty && isCentered(dst) =>
        src match {
          case IR_Multiplication(ListBuffer(alpha, x : IR_FieldLikeAccess)) if isCentered(x) && !alpha.isInstanceOf[IR_FieldLikeAccess] =>
            Some(ExaMG_Call(axpy, alpha, handle(x), slotOf(x), handle(dst), slotOf(dst)))
          case x : IR_FieldLikeAccess if isCentered(x)                                                                                   =>
            Some(ExaMG_Call(axpy, IR_RealConstant(1.0), handle(x), slotOf(x), handle(dst), slotOf(dst)))
          case _                                                                                                                         => None
        }

      // ---- reductions: x += a * b --------------------------------------------------------------------------------
      case Some(IR_Assignment(acc : IR_VariableAccess, IR_Multiplication(ListBuffer(a : IR_FieldLikeAccess, b : IR_FieldLikeAccess)), "+="))
        if loop.parallelization.reduction.exists(r => r.op == "+" && r.targetName == acc.name) && isCentered(a) && isCentered(b) =>
        // reductions count nodes shared between blocks once (baseExt/ir/IR_LoopOverPointsInOneFragment.scala:116-125);
        // a `where i0 > 0 && ...` condition that says the same thing (Testing/SISC NormResidual) is absorbed
        Some(ExaMG_Call(dot, handle(a), slotOf(a), handle(b), slotOf(b), ExaMG_AddressOf(acc)))

      // ---- 2x2 block smoother: solve locally at the cell ------------------------------------------------------------
      case Some(ls : IR_LocalSolve) if ls.unknowns.length == 2 && !ls.jacobiType =>
        (colouring, ExaMG_System2.recognise(ls)) match {
          case (Some(RedBlack(c)), Some(sys)) =>
            Some(ExaMG_Call(system2SmoothRBColor, sys.handle, handle(ls.unknowns(0)), handle(ls.unknowns(1)), sys.rhsU, sys.rhsV,
              ls.relax.getOrElse(IR_RealConstant(1.0)), c))
          case _                              => None
        }

      // ---- (2 dim + 1)^2 block smoother of the staggered-grid Stokes programs ---------------------------------------
      case Some(ls : IR_LocalSolve) if ls.unknowns.length == 2 * numDims + 1 => ExaMG_Stokes.lower(ls, colouring)

      case _ => None
    }
  }

  /** u = c, y = x, y = x + beta * y */
  private def lowerStreaming(dst : IR_FieldLikeAccess, src : IR_Expression) : Option[IR_Statement] = src match {
    case c @ (_ : IR_RealConstant | _ : IR_IntegerConstant | _ : IR_VariableAccess)                                             =>
      Some(ExaMG_Call(set, handle(dst), slotOf(dst), c))
    case x : IR_FieldLikeAccess if isCentered(x) && x.field != dst.field                                                         =>
      Some(ExaMG_Call(copy, handle(x), slotOf(x), handle(dst), slotOf(dst)))
    case IR_Addition(ListBuffer(x : IR_FieldLikeAccess, IR_Multiplication(ListBuffer(beta, y : IR_FieldLikeAccess))))
      if y.field == dst.field && isCentered(x) && isCentered(y)                                                                  =>
      Some(ExaMG_Call(xpay, handle(x), slotOf(x), beta, handle(dst), slotOf(dst)))
    case IR_Multiplication(ListBuffer(scale, conv)) if asConvolution(conv).exists(c => isRestriction(c.op.stencil))              =>
      val c = asConvolution(conv).get
      Some(ExaMG_Call(restrict, handle(c.u), handle(dst), scale)) // `4.0 * ( Restriction * Residual )`
    case _                                                                                                                       =>
      // initialisation loops F = <expr>(vf_nodePosition): lowered to the postfix program of examg_field_eval
      toPostfix(src).map(prog => ExaMG_FieldEval(handle(dst), slotOf(dst), 0, prog) : IR_Statement).orElse {
        // any other pointwise expression over fields of the level read at the loop's own point (the nonlinear terms of
        // Testing/NonLinear/FAS_2D_Basic.exa4 once `Laplace * Solution` sits in a work field): examg_loop_eval
        val srcs = ListBuffer[IR_FieldLikeAccess]()
        toPostfix(src, Some(srcs)).filter(_ => srcs.forall(_.field.level == dst.field.level))
          .map(prog => ExaMG_LoopEval(handle(dst), slotOf(dst), srcs.map(handle), srcs.map(slotOf), prog))
      }
  }

  this += new Transformation("Loops over points", {
    case loop : IR_LoopOverPoints =>
      lower(loop) match {
        case Some(call) =>
          // with experimental_splitLoopsForAsyncComm the loop carries its communicate statements; the library calls
          // exchange what they consume themselves (ghost depth per fused pass), so pre-communicates of the field a
          // call reads are dropped; post-communicates stay
          val out = ListBuffer[IR_Statement]()
          out ++= loop.preComms.filterNot(ExaMG_MergeCalls.impliedBy(_, call))
          out += call
          out ++= loop.postComms
          out
        case None       =>
          ExaMG_ReportLeftovers.note(loop)
          loop
      }
  }, false)

  this += new Transformation("communicate / apply bc / advance", {
    case c : IR_Communicate if c.targets.forall(t => t.begin.isEmpty && t.end.isEmpty) && c.condition.isEmpty =>
      // `communicate [begin|finish] F`: one exchange per statement pair; the library runs its handshake in `finish`
      if (c.op == "begin") IR_NullStatement
      else ExaMG_Call(exchange, ExaMG_FieldHandle(c.field), c.slot match {
        case IR_SlotAccess(_, 0) => IR_IntegerConstant(-1)
        case other               => other
      })
    case bc : IR_ApplyBC                                                                                      =>
      ExaMG_Call(applyBC, ExaMG_FieldHandle(bc.field), bc.slot match {
        case IR_SlotAccess(_, 0) => IR_IntegerConstant(-1)
        case other               => other
      })
    case IR_AdvanceSlot(iv, 1)                                                                                =>
      ExaMG_Call(fieldAdvance, ExaMG_FieldHandle(iv.field))
  })
}

/// intermediate nodes that ExaMG_MergeCalls folds

/** one Jacobi sweep (`advance` follows); consecutive sweeps merge into examg_smooth_jacobi(..., nSweeps) */
case class ExaMG_JacobiSweep(var op : IR_Expression, var u : IR_Expression, var rhs : IR_Expression, var omega : IR_Expression) extends IR_Statement {
  override def prettyprint(out : exastencils.prettyprinting.PpStream) : Unit =
    out << ExaMG_Call(ExaMG_Symbols.smoothJacobi, op, u, rhs, omega, IR_IntegerConstant(1))
}

/** one colour of an n^d-way colouring; the full set in canonical order merges into examg_smooth_colored */
case class ExaMG_ColourStep(var op : IR_Expression, var u : IR_Expression, var rhs : IR_Expression, var omega : IR_Expression,
    var n : Int, var colour : Array[Int]) extends IR_Statement {
  override def prettyprint(out : exastencils.prettyprinting.PpStream) : Unit =
    Logger.error("examg back-end: a single colour of an n^d-way colouring outside a complete `color with` block")
}

/** examg_restrict_cell / examg_prolong_add_cell take arrays of handles (one loop per unknown merges into one call) */
case class ExaMG_CellTransfer(var name : String, var from : IR_Expression, var to : IR_Expression) extends IR_Statement {
  override def prettyprint(out : exastencils.prettyprinting.PpStream) : Unit = {
    out << "{ const int from_[1] = { " << from << " }, to_[1] = { " << to << " }; "
    out << ExaMG_Call(name, IR_VariableAccess("from_", IR_IntegerDatatype), IR_VariableAccess("to_", IR_IntegerDatatype), IR_IntegerConstant(1)) << " }"
  }
}

/** examg_field_eval with its program as a static array */
case class ExaMG_FieldEval(var field : IR_Expression, var slot : IR_Expression, var comp : Int, var prog : ListBuffer[(String, Double)]) extends IR_Statement {
  override def prettyprint(out : exastencils.prettyprinting.PpStream) : Unit = {
    out << "{ static const examg_bc_op prog_[] = { "
    out << prog.map(p => s"{ ${ p._1 }, ${ java.lang.Double.toString(p._2) } }").mkString(", ")
    out << " }; "
    out << ExaMG_Call(ExaMG_Symbols.fieldEval, field, slot, IR_IntegerConstant(comp), IR_VariableAccess("prog_", IR_IntegerDatatype), IR_IntegerConstant(prog.length)) << " }"
  }
}

/** examg_loop_eval: destination, source handles / slots and the program as static arrays */
case class ExaMG_LoopEval(var field : IR_Expression, var slot : IR_Expression, var srcs : ListBuffer[IR_Expression],
    var srcSlots : ListBuffer[IR_Expression], var prog : ListBuffer[(String, Double)]) extends IR_Statement {
  override def prettyprint(out : exastencils.prettyprinting.PpStream) : Unit = {
    out << "{ static const examg_bc_op prog_[] = { "
    out << prog.map(p => s"{ ${ p._1 }, ${ java.lang.Double.toString(p._2) } }").mkString(", ")
    out << " }; const int src_[] = { " <<< (srcs, ", ") << " }, slot_[] = { " <<< (srcSlots, ", ") << " }; "
    out << ExaMG_Call(ExaMG_Symbols.loopEval, field, slot, IR_VariableAccess("src_", IR_IntegerDatatype),
      IR_VariableAccess("slot_", IR_IntegerDatatype), IR_IntegerConstant(srcs.length), IR_VariableAccess("prog_", IR_IntegerDatatype),
      IR_IntegerConstant(prog.length)) << " }"
  }
}

/// ExaMG_ReportLeftovers

object ExaMG_ReportLeftovers {
  val loops = ListBuffer[String]()
  def note(loop : IR_LoopOverPoints) : Unit = {
    val what = s"loop over ${ loop.field.name }@${ loop.field.level }"
    loops += what
    Logger.warn(s"examg back-end: $what matches no libexamg entry point; it stays a host loop on downloaded data")
  }
}
