// placeholder until the tiled kernel lands
#include "pva_kernels.cuh"
namespace pva {
bool tiled_supported(const KParams&, int, int, const char** why) { *why = "tiled kernel not built"; return false; }
cudaError_t launch_tiled(KParams&, int, int, int, cudaStream_t, TiledConfig*) { return cudaErrorNotSupported; }
}
