// tc_engine.cu -- bf16 tensor-core engine (tcgen05 / TMEM / TMA implicit-GEMM convolutions).
// Placeholder until the tcgen05 path lands: the bf16 dtype fails loudly, it never falls back.
#include "engine.h"

namespace rb {

int tc_pack_weights(rb_handle*) { return RB_OK; }
void tc_free(rb_handle*) {}

int tc_run_program(rb_handle*, int, int, const void*, void*, int, int, cudaStream_t) {
  set_error("the bf16 tensor-core engine is not built into this library");
  return RB_ERR_UNSUPPORTED;
}

}  // namespace rb
