// placeholder until the tcgen05 kernel lands (next milestone): the BF16 precision is refused loudly,
// never routed to another implementation.
#include "common.cuh"
#include "internal.cuh"

int tc_prepare_layer(dcv_ctx*, Layer&) { return 0; }
void tc_free_layer(Layer&) {}
int tc_deconv(dcv_ctx* ctx, const Layer&, const float*, int, int, int, float*) {
  return ctx->fail(DCV_ERR_UNSUPPORTED, "DCV_BF16: tensor-core kernel not built into this library yet");
}
