package exastencils.app.ir

import exastencils.baseExt.ir.IR_ResolveLoopOverPoints
import exastencils.config.Knowledge
import exastencils.globals.ir.IR_SetupAllocateDataFunctionWrapper
import exastencils.parallelization.api.examg._
import exastencils.scheduling._
import exastencils.util.ir.IR_ResolveBenchmarkFunctions

/// IR_ExaMGLayerHandler
//
// IR schedule of the libexamg back-end (Knowledge.experimental_examg = true).  Precedent and shape:
// app/ir/IR_WaLBerlaLayerHandler.scala -- take the default schedule and splice strategies in at named positions.
//
//   IR_SetupAllocateDataFunctionWrapper (:104)  -> followed by ExaMG_AddGlobals: handles + setupBuffers through the library
//   IR_ResolveBenchmarkFunctions (:127)         -> preceded by ExaMG_ResolveHostFunctions (timers, printField, external fields)
//   IR_ResolveLoopOverPoints (:143)             -> preceded by ExaMG_ResolveCalls + ExaMG_MergeCalls: every loop nest of the hot
//                                                  path becomes a call while it is still an IR_LoopOverPoints
// Everything downstream (loop resolution, polyhedral optimisation, OpenMP / CUDA extraction, vectorisation) then finds
// no loop over points on the hot path and has nothing to do there; host-side control flow (Solve, mgCycle recursion,
// printing) goes through the standard pipeline unchanged.

object IR_ExaMGLayerHandler extends IR_LayerHandler {
  var scheduler : Scheduler = Scheduler()

  override def initialize() : Unit = IR_DefaultLayerHandler.initialize()

  override def schedule() : Unit = {
    IR_DefaultLayerHandler.schedule()
    scheduler.queue ++= IR_DefaultLayerHandler.scheduler.queue

    scheduler.appendToFirstFound(IR_SetupAllocateDataFunctionWrapper, ExaMG_AddGlobals)

    scheduler.prependToFirstFound(IR_ResolveBenchmarkFunctions, ExaMG_ResolveHostFunctions)

    scheduler.prependToFirstFound(IR_ResolveLoopOverPoints,
      ExaMG_ResolveCalls,
      ExaMG_MergeCalls,
      ConditionedSingleStrategyWrapper(Knowledge.experimental_examg_reportLeftovers, ExaMG_PrintLeftovers))
  }

  override def print() : Unit = IR_DefaultLayerHandler.print()

  override def shutdown() : Unit = IR_DefaultLayerHandler.shutdown()
}

object ExaMG_PrintLeftovers extends NoStrategyWrapper {
  override def callback : () => Unit = () => {
    if (ExaMG_ReportLeftovers.loops.nonEmpty)
      exastencils.logger.Logger.warn(s"examg back-end: ${ ExaMG_ReportLeftovers.loops.length } loop(s) stay on the host: " +
        ExaMG_ReportLeftovers.loops.distinct.mkString(", "))
  }
}
