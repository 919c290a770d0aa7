// mmq_tc.cu — prefill path (tcgen05 int8 GEMM).  Placeholder until the kernel lands: reports "unsupported" so that
// abi.cu routes n > 8 through the column-tiled GEMV.
#include "mm_problem.cuh"

int b200_tc_supported(const b200_ctx *, int, const mm_problem &) { return 0; }
int b200_launch_gemm_tc(b200_ctx *, int, const mm_problem &, int64_t) {
    b200_set_error("tcgen05 GEMM not built");
    return B200_E_UNSUPPORTED;
}
