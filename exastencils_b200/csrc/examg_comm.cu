// examg_comm.cu -- `communicate F` (halo exchange) and global reductions.
//
// Reference: C/s/e/communication/ir/IR_CommunicateFunction.scala:412-470 (one
// dimension after the other; ghost ranges IR_PackInfoGhost.scala:13-60), MPI
// point-to-point C/s/e/communication/ir/IR_RemoteSend.scala:38-58 and
// MPI_Allreduce after reduction loops C/s/e/parallelization/api/mpi/MPI_Reduction.scala:92-127.
#include "examg_internal.h"

using namespace examg;

namespace examg {

int do_exchange(examg_ctx *c, const Field &f, int slot) {
  (void)f; (void)slot;
  if (c->world == 1) return EXAMG_OK;  // no neighbour blocks: nothing to exchange
  return set_error(EXAMG_EUNSUP, "examg_exchange: peers not wired (examg_comm_import)");
}

int finish_sum(examg_ctx *c, int slot, double *out) {
  EXAMG_CUDA(cudaMemcpyAsync(c->red_host + slot, c->red_out + slot, sizeof(dd), cudaMemcpyDeviceToHost, c->stream));
  EXAMG_CUDA(cudaStreamSynchronize(c->stream));
  if (c->world > 1) return set_error(EXAMG_EUNSUP, "global reduction: peers not wired (examg_comm_import)");
  *out = c->red_host[slot].hi + c->red_host[slot].lo;
  return EXAMG_OK;
}

}  // namespace examg

extern "C" {
int examg_comm_blob_size(examg_ctx *c) { (void)c; return 64; }
int examg_comm_export(examg_ctx *c, void *blob) { (void)c; (void)blob; return set_error(EXAMG_EUNSUP, "not yet"); }
int examg_comm_import(examg_ctx *c, const void *blobs, int world) { (void)c; (void)blobs; (void)world; return set_error(EXAMG_EUNSUP, "not yet"); }
}
