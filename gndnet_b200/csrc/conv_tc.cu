// conv_tc.cu -- tcgen05 / TMEM / TMA implicit-GEMM convolution (placeholder until the kernel lands).
#include "model.cuh"
namespace gnd {
struct TensorMaps { int unused; };
int tc_build_tensor_maps(gnd_ctx *) { return fail(GND_ERR_UNSUPPORTED, "tcgen05 convolution path not built yet"); }
void tc_free_tensor_maps(gnd_ctx *) {}
int tc_conv_run(gnd_ctx *, int, const Act &, const Act &, int, bool, cudaStream_t) {
    return fail(GND_ERR_UNSUPPORTED, "tcgen05 convolution path not built yet");
}
}  // namespace gnd
