// tcgen05 3xTF32 grouped contraction (placeholder until the TMA/TMEM kernel lands).
#include "gemm.cuh"

namespace porl {

int launch_gemm_tc(porl_ctx*, bool, bool, GemmBatch&, cudaStream_t) { return PORL_ERR_UNSUPPORTED; }

}  // namespace porl
