package exastencils.parallelization.api.examg

import scala.collection.mutable.ListBuffer

import exastencils.base.ir._
import exastencils.boundary.ir._
import exastencils.config._
import exastencils.datastructures._
import exastencils.domain.ir.IR_DomainCollection
import exastencils.field.ir._
import exastencils.globals.ir._
import exastencils.logger.Logger
import exastencils.operator.ir._
import exastencils.parallelization.api.mpi.MPI_IV_MpiRank
import exastencils.prettyprinting.PpStream
import exastencils.scheduling.NoStrategyWrapper

/// ExaMG_SetupFunction
//
// setupBuffers() of the generated program (globals/ir/IR_AllocateDataFunction.scala:64-72 allocates one array per field,
// level, fragment and slot there): with this back-end it creates the context, every field and every operator inside
// libexamg and keeps their integer handles in Global/Global.h.  The library owns the device memory (layout, padding
// and ghost layers are its business, DESIGN.md section 3).

case class ExaMG_SetupFunction(var fields : ListBuffer[IR_Field], var stencils : ListBuffer[IR_Stencil],
    var stencilFields : ListBuffer[IR_StencilField]) extends IR_FuturePlainFunction {

  import ExaMG_Symbols._

  override var name = IR_AllocateDataFunction.fctName
  override def prettyprint_decl() : String = prettyprint

  private def intArray(values : Seq[Int]) = values.mkString("{ ", ", ", " }")

  private def configInit : IR_Statement = {
    val d = Knowledge.dimensionality
    val cells = (0 until 3).map(i => if (i < d) Knowledge.domain_rect_numFragsPerBlockAsVec(i) * Knowledge.domain_fragmentLengthAsVec(i) else 1)
    val blocks = (0 until 3).map(i => if (i < d) Knowledge.domain_rect_numBlocksAsVec(i) else 1)
    val periodic = (0 until 3).map(i => if (i < d && Knowledge.domain_rect_periodicAsVec(i)) 1 else 0)
    val dom = IR_DomainCollection.getByIdentifier("global").get // Domain global< lo to hi >
    val (lo, hi) = dom.asInstanceOf[exastencils.domain.ir.IR_DomainFromAABB].aabb match {
      case aabb => ((0 until 3).map(i => if (i < d) aabb.lower(i) else 0.0), (0 until 3).map(i => if (i < d) aabb.upper(i) else 1.0))
    }
    // rank -> block x fastest (domain/ir/IR_InitGeneratedDomain.scala:41-47); one fragment per block: the library
    // fuses what the reference splits into fragments (numFragsPerBlock x fragmentLength cells at level 0)
    val rank : String = if (Knowledge.mpi_enabled) MPI_IV_MpiRank.prettyprint() else "0"
    ExaMG_NativeBlock(
      s"""examg_config cfg_ = { $d, ${ Knowledge.minLevel }, ${ Knowledge.maxLevel }, ${ intArray(cells) }, ${ intArray(blocks) }, $rank,
         |    ($rank) % examgNumDevices(), ${ lo.mkString("{ ", ", ", " }") }, ${ hi.mkString("{ ", ", ", " }") }, ${ intArray(periodic) } };
         |${ ExaMG_Context.name } = $init(&cfg_);
         |if (!${ ExaMG_Context.name }) { std::cerr << "examg: " << $lastError() << std::endl; exit(1); }""".stripMargin)
  }

  private def boundary(field : IR_Field) : ListBuffer[IR_Statement] = {
    val h = ExaMG_FieldHandle(field)
    field.boundary match {
      case IR_NoBC                                                   => ListBuffer()
      case IR_NeumannBC(_)                                           => ListBuffer(ExaMG_Call(fieldSetBC, h, IR_VariableAccess("EXAMG_BC_NEUMANN", IR_IntegerDatatype), IR_RealConstant(0.0)))
      case IR_DirichletBC(v @ (_ : IR_RealConstant | _ : IR_IntegerConstant), _) =>
        ListBuffer(ExaMG_Call(fieldSetBC, h, IR_VariableAccess("EXAMG_BC_DIRICHLET_CONST", IR_IntegerDatatype), v))
      case IR_DirichletBC(expr, _)                                   =>
        ExaMG_Patterns.toPostfix(expr) match {
          case Some(prog) => ListBuffer(ExaMG_BCProgram(h, prog))
          case None       =>
            Logger.warn(s"examg back-end: boundary expression of ${ field.name }@${ field.level } is not a position expression; " +
              "boundary values stay with the host (EXAMG_BC_DIRICHLET_HOST, written by set<name>)")
            ListBuffer(ExaMG_Call(fieldSetBC, h, IR_VariableAccess("EXAMG_BC_DIRICHLET_HOST", IR_IntegerDatatype), IR_RealConstant(0.0)))
        }
      case other                                                     =>
        Logger.warn(s"examg back-end: boundary condition $other of ${ field.name } is not supported"); ListBuffer()
    }
  }

  override def generateFct() = {
    val body = ListBuffer[IR_Statement]()
    body += configInit
    for (field <- fields.sortBy(f => (f.level, f.name))) {
      val cell = field.layout.localization == exastencils.grid.ir.IR_AtCellCenter
      val ncomp = field.layout.datatype.resolveFlattendSize
      body += IR_Assignment(ExaMG_FieldHandle(field),
        IR_FunctionCall(IR_ExternalFunctionReference(if (cell) fieldCreateCell else fieldCreate, IR_IntegerDatatype),
          ExaMG_Context.access, IR_StringConstant(field.name), IR_IntegerConstant(field.level), IR_IntegerConstant(field.numSlots), IR_IntegerConstant(ncomp)))
      body += ExaMG_CheckHandle(ExaMG_FieldHandle(field))
      body ++= boundary(field)
    }
    // operators: entries in declaration order -- it fixes the floating-point association of A * u
    for (st <- stencils.filter(st => st.colStride.forall(_ == 1.0)).sortBy(s => (s.level, s.name)))
      body += ExaMG_StencilCreate(st, None)
    for (sf <- stencilFields.sortBy(s => (s.level, s.name)))
      body += ExaMG_StencilCreate(sf.stencil, Some(sf))
    for (sys <- ExaMG_System2.found)
      body += ExaMG_NativeBlock(
        s"""{ examg_system2_desc d_ = { ${ ExaMG_StencilHandle(sys.laplace.name, sys.laplace.level).handleName }, ${ sys.scale.prettyprint() },
           |    ${ ExaMG_FieldHandle(sys.cuu.field).handleName }, ${ ExaMG_FieldHandle(sys.cuv.field).handleName }, ${ ExaMG_FieldHandle(sys.cvv.field).handleName } };
           |  ${ sys.handle.name } = $system2Create(${ ExaMG_Context.name }, &d_); if (${ sys.handle.name } < 0) examgFail(); }""".stripMargin)
    // several blocks: wire the peers (INTEGRATION.md section 5)
    if (Knowledge.mpi_enabled)
      body += ExaMG_NativeBlock(
        s"""{ const int nb_ = $commBlobSize(${ ExaMG_Context.name });
           |  std::vector<char> mine_(nb_), all_((size_t)nb_ * mpiSize);
           |  EXAMG_CHECK($commExport(${ ExaMG_Context.name }, mine_.data()));
           |  MPI_Allgather(mine_.data(), nb_, MPI_BYTE, all_.data(), nb_, MPI_BYTE, mpiCommunicator);
           |  EXAMG_CHECK($commImport(${ ExaMG_Context.name }, all_.data(), mpiSize)); }""".stripMargin)
    IR_PlainFunction(name, IR_UnitDatatype, body)
  }
}

/** verbatim host C++ (several statements) */
case class ExaMG_NativeBlock(var code : String) extends IR_Statement {
  override def prettyprint(out : PpStream) : Unit = out << code
}

case class ExaMG_CheckHandle(var h : ExaMG_HandleAccess) extends IR_Statement {
  override def prettyprint(out : PpStream) : Unit = out << "if (" << h << " < 0) examgFail();"
}

case class ExaMG_BCProgram(var field : IR_Expression, var prog : ListBuffer[(String, Double)]) extends IR_Statement {
  override def prettyprint(out : PpStream) : Unit = {
    out << "{ static const examg_bc_op prog_[] = { " << prog.map(p => s"{ ${ p._1 }, ${ java.lang.Double.toString(p._2) } }").mkString(", ") << " }; "
    out << ExaMG_Call(ExaMG_Symbols.fieldSetBCExpr, field, IR_VariableAccess("prog_", IR_IntegerDatatype), IR_IntegerConstant(prog.length)) << " }"
  }
}

case class ExaMG_StencilCreate(var stencil : IR_Stencil, var stencilField : Option[IR_StencilField]) extends IR_Statement {
  override def prettyprint(out : PpStream) : Unit = {
    val name = stencilField.map(_.name).getOrElse(stencil.name)
    val h = ExaMG_StencilHandle(name, stencil.level)
    val entries = stencil.entries.map(_.asStencilOffsetEntry)
    val offs = entries.map(e => (0 until 3).map(d => if (d < e.offset.length) e.offset(d) else 0).mkString("{ ", ", ", " }"))
    out << "{ static const int off_[][3] = { " << offs.mkString(", ") << " }; "
    stencilField match {
      case None     =>
        // coefficients are generation-time constants (2 / h^2 ...): printed as expressions, evaluated by the C++ compiler
        out << "const double coef_[] = { " <<< (entries.map(_.coefficient), ", ") << " }; "
        out << h << " = " << ExaMG_Symbols.stencilConst << '(' << ExaMG_Context.name << ", " << stencil.level << ", " << entries.length << ", off_, coef_); "
      case Some(sf) =>
        out << h << " = " << ExaMG_Symbols.stencilField << '(' << ExaMG_Context.name << ", " << stencil.level << ", " << entries.length << ", off_, " << ExaMG_FieldHandle(sf.field) << "); "
    }
    out << "if (" << h << " < 0) examgFail(); }"
  }
}

/** number of values one block holds of a field (all components, ghost layers excluded) */
object ExaMG_BlockValues {
  def apply(field : IR_Field) : IR_Expression = {
    val layout = field.layout
    var n : Long = field.layout.datatype.resolveFlattendSize
    for (d <- 0 until layout.numDimsGrid) n *= (layout.idxById("DRE", d) - layout.idxById("DLB", d)) match {
      case IR_IntegerConstant(v) => v
      case _                     => Logger.error("examg back-end: field extents are not compile-time constants"); 1L
    }
    IR_IntegerConstant(n)
  }
}

/// ExaMG_AddGlobals

/** Handles, the context, the check macro and the (de)initialisation calls in Global/Global.h (+ initGlobals / destroyGlobals). */
object ExaMG_AddGlobals extends NoStrategyWrapper {
  override def callback : () => Unit = () => {
    val globals = IR_GlobalCollection.get
    if (!Settings.additionalIncludes.contains("examg.h")) Settings.additionalIncludes += "examg.h"
    if (!Settings.additionalLibs.contains("examg")) Settings.additionalLibs += "examg"
    Settings.additionalMacros +=
      """#define EXAMG_CHECK(call) do { if ((call) < 0) examgFail(); } while (0)
        |static inline void examgFail() { std::cerr << "examg: " << examg_last_error() << std::endl; exit(1); }
        |static inline int examgNumDevices() { const char *e = getenv("EXAMG_NUM_DEVICES"); return e ? atoi(e) : 8; }""".stripMargin
    globals.variables += ExaMG_Context.declaration
    for (field <- IR_FieldCollection.objects) globals.variables += ExaMG_FieldHandle(field).declaration
    for (st <- IR_StencilCollection.objects if st.colStride.forall(_ == 1.0)) globals.variables += ExaMG_StencilHandle(st.name, st.level).declaration
    for (sf <- IR_StencilFieldCollection.objects) globals.variables += ExaMG_StencilHandle(sf.name, sf.level).declaration
    for (lvl <- Knowledge.minLevel to Knowledge.maxLevel) globals.variables += IR_VariableDeclaration(IR_IntegerDatatype, s"sys_$lvl", IR_IntegerConstant(-1))

    // setupBuffers: replace the reference's allocation function
    globals.functions = globals.functions.filterNot(_.name == IR_AllocateDataFunction.fctName)
    globals += ExaMG_SetupFunction(IR_FieldCollection.objects, IR_StencilCollection.objects, IR_StencilFieldCollection.objects)

    // destroyGlobals: examg_finalize
    globals.functions.find(_.name == "destroyGlobals").foreach {
      case fct : IR_PlainFunction => fct.body += ExaMG_Call(ExaMG_Symbols.finalizeCtx, ListBuffer[IR_Expression](ExaMG_Context.access), withContext = false)
      case _                      =>
    }
  }
}

/// ExaMG_LibraryCycle

/** Knowledge.experimental_examg_libraryCycle: the L3-generated `gen_mgCycle@finest` (solver/l3/L3_SolverForEquation.scala:341-395)
  * becomes ONE library object -- the solver_* Knowledge values travel as text through examg_cycle_desc_from_options
  * (smoother, colouring, damping, sweep counts, solver_cgs incl. "Smoother" = examg_smoother_coarse at the bottom), the
  * cycle runs as a CUDA graph (examg_solver_vcycle).  The statement-by-statement lowering stays the default. */
object ExaMG_LibraryCycle extends DefaultStrategy("Replace the generated multigrid cycle by the library's solver object") {
  def options : String = Seq(
    s"solver_smoother_numPre = ${ Knowledge.solver_smoother_numPre }", s"solver_smoother_numPost = ${ Knowledge.solver_smoother_numPost }",
    s"solver_smoother_damping = ${ Knowledge.solver_smoother_damping }", s"solver_smoother_jacobiType = ${ Knowledge.solver_smoother_jacobiType }",
    s"solver_smoother_coloring = ${ Knowledge.solver_smoother_coloring }", s"solver_cgs = ${ Knowledge.solver_cgs }",
    s"solver_cgs_maxNumIts = ${ Knowledge.solver_cgs_maxNumIts }", s"solver_cgs_targetResReduction = ${ Knowledge.solver_cgs_targetResReduction }",
    s"solver_useFAS = ${ Knowledge.solver_useFAS }").mkString("; ")

  /** printed once into setupBuffers, after the operators exist; `stencilNames` = handle names per level, coarsest first */
  def setup(stencilNames : Seq[String]) : IR_Statement = ExaMG_NativeBlock(
    s"""{ static const int st_[] = { ${ stencilNames.mkString(", ") } };
       |  examg_cycle_desc d_; d_.restrict_scale = 1.0; d_.stencils = st_;
       |  EXAMG_CHECK(${ ExaMG_Symbols.cycleDescFromOptions }("$options", ${ Knowledge.dimensionality }, &d_, NULL));
       |  examgSolver = ${ ExaMG_Symbols.solverCreate }(${ ExaMG_Context.name }, &d_); if (!examgSolver) examgFail(); }""".stripMargin)

  this += new Transformation("gen_mgCycle@finest", {
    case fct : IR_LeveledFunction if fct.baseName == "gen_mgCycle" && fct.level == Knowledge.maxLevel =>
      fct.body = ListBuffer[IR_Statement](ExaMG_NativeBlock(
        s"EXAMG_CHECK(${ ExaMG_Symbols.solverVCycle }(examgSolver, ${ Knowledge.maxLevel }));"))
      fct
  })
}

/// ExaMG_ResolveHostFunctions

/** timers, printField, external fields: the observable host-side functions keep their names and forward */
object ExaMG_ResolveHostFunctions extends DefaultStrategy("Forward timers, field printing and external field copies to libexamg") {

  import ExaMG_Symbols._

  this += new Transformation("Timers", {
    case IR_ExpressionStatement(IR_FunctionCall(fct, ListBuffer(name : IR_StringConstant))) if fct.name == "startTimer" =>
      ExaMG_Call(timerStart, name)
    case IR_ExpressionStatement(IR_FunctionCall(fct, ListBuffer(name : IR_StringConstant))) if fct.name == "stopTimer"  =>
      ExaMG_Call(timerStop, name)
    case IR_ExpressionStatement(IR_FunctionCall(fct, args)) if fct.name == "printAllTimers" && args.isEmpty            =>
      ExaMG_Call(timerPrintAll)
    // benchmarkStart / benchmarkStop: LIKWID / TALP markers in the reference (util/ir/IR_ResolveBenchmarkFunctions.scala:107-116),
    // NVTX ranges here; must run before IR_ResolveBenchmarkFunctions turns them into IR_NullStatement
    case IR_ExpressionStatement(IR_FunctionCall(fct, ListBuffer(name : IR_StringConstant))) if fct.name == "benchmarkStart" =>
      ExaMG_Call(benchmarkStart, name)
    case IR_ExpressionStatement(IR_FunctionCall(fct, ListBuffer(name : IR_StringConstant))) if fct.name == "benchmarkStop"  =>
      ExaMG_Call(benchmarkStop, name)
  })

  this += new Transformation("printField / writeField / readField", {
    case pf : IR_PrintField if !pf.useBinary =>
      ExaMG_Call(fieldPrint, ExaMG_FieldHandle(pf.field), IR_IntegerConstant(-1), pf.filename, pf.separator)
    // binary dumps through the "lock" interface: block after block in rank order (MPI_Sequential in the reference)
    case wf : IR_WriteField if wf.useBinary && !wf.includeGhostLayers =>
      if (Knowledge.mpi_enabled)
        ExaMG_NativeBlock(
          s"""for (int r_ = 0; r_ < mpiSize; ++r_) {
             |  if (r_ == mpiRank) EXAMG_CHECK($fieldWrite(${ ExaMG_Context.name }, ${ ExaMG_FieldHandle(wf.field).handleName }, -1, ${ wf.filename.prettyprint() }, r_ > 0));
             |  MPI_Barrier(mpiCommunicator);
             |}""".stripMargin)
      else
        ExaMG_Call(fieldWrite, ExaMG_FieldHandle(wf.field), IR_IntegerConstant(-1), wf.filename, IR_IntegerConstant(0))
    case rf : IR_ReadField if rf.useBinary && !rf.includeGhostLayers  =>
      ExaMG_Call(fieldRead, ExaMG_FieldHandle(rf.field), IR_IntegerConstant(-1), rf.filename,
        if (Knowledge.mpi_enabled) IR_VariableAccess("mpiRank", IR_IntegerDatatype) * ExaMG_BlockValues(rf.field) else IR_IntegerConstant(0))
  })

  this += new Transformation("external fields", {
    // set<name>(double* src, int slot) / get<name>(double* dest, int slot): interfacing/ir/IR_CopyToExternalField.scala:30-90
    case fct : exastencils.interfacing.ir.IR_CopyFromExternalField =>
      IR_PlainFunction(fct.name, IR_UnitDatatype,
        ListBuffer(IR_FunctionArgument("src", IR_PointerDatatype(IR_RealDatatype)), IR_FunctionArgument("slot", IR_IntegerDatatype)),
        ListBuffer[IR_Statement](ExaMG_Call(fieldUpload, ExaMG_FieldHandle(fct.dest), IR_VariableAccess("slot", IR_IntegerDatatype), IR_VariableAccess("src", IR_PointerDatatype(IR_RealDatatype)))))
    case fct : exastencils.interfacing.ir.IR_CopyToExternalField   =>
      IR_PlainFunction(fct.name, IR_UnitDatatype,
        ListBuffer(IR_FunctionArgument("dest", IR_PointerDatatype(IR_RealDatatype)), IR_FunctionArgument("slot", IR_IntegerDatatype)),
        ListBuffer[IR_Statement](ExaMG_Call(fieldDownload, ExaMG_FieldHandle(fct.src), IR_VariableAccess("slot", IR_IntegerDatatype), IR_VariableAccess("dest", IR_PointerDatatype(IR_RealDatatype)))))
  })
}
