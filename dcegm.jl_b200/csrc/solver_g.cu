// solver_g.cu — general model with state-dependent (Markov) income, dc_EGM!(::GModel)
// (src/dc_algo.jl:112-206).  Placeholder until the kernel lands: fails loudly.
#include "device_common.cuh"

namespace dcegm {
size_t solver_g_smem_bytes(int, int, int) { return 0; }
cudaError_t launch_solve_g(const DeviceBatch&, cudaStream_t) { return cudaErrorNotSupported; }
}  // namespace dcegm
