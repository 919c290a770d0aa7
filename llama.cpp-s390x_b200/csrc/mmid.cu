// mmid.cu — device-side expert gather for MUL_MAT_ID (replaces the serial bucketing on thread 0 in
// ggml_compute_forward_mul_mat_id, /root/reference/ggml/src/ggml-cpu/ggml-cpu.c:1609-1624; CUDA analogue
// mm_ids_helper, ggml/src/ggml-cuda/mmid.cu:26-116).
//
// One CTA per expert e, no inter-CTA dependency: in a single sweep over the (token, slot) pairs the CTA counts how many
// pairs route to an expert < e (that is expert_bounds[e]) and how many to e itself, then a second ordered sweep
// compacts its own pairs, in pair-index order (token-major => stable), to ids_src1[bounds[e] ...].
#include "common.cuh"

namespace {

constexpr int MMID_THREADS = 256;

__global__ void __launch_bounds__(MMID_THREADS)
mm_ids_kernel(const int32_t * __restrict__ ids, const int64_t ids_stride, const int n_expert, const int n_used,
              const int64_t n_pairs, int32_t * __restrict__ ids_src1, int32_t * __restrict__ expert_bounds) {
    __shared__ int s_less[MMID_THREADS / 32], s_eq[MMID_THREADS / 32], s_warp[MMID_THREADS / 32];
    __shared__ int s_base;
    const int e    = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    int less = 0, eq = 0;
    for (int64_t p = threadIdx.x; p < n_pairs; p += MMID_THREADS) {
        const int64_t token = p / n_used;
        const int     id    = __ldg(ids + token * ids_stride + (p - token * n_used));
        less += (id >= 0 && id < e);
        eq   += (id == e);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { less += __shfl_xor_sync(0xffffffffu, less, o); eq += __shfl_xor_sync(0xffffffffu, eq, o); }
    if (lane == 0) { s_less[warp] = less; s_eq[warp] = eq; }
    __syncthreads();
    if (threadIdx.x == 0) {
        int l = 0, q = 0;
        for (int w = 0; w < MMID_THREADS / 32; ++w) { l += s_less[w]; q += s_eq[w]; }
        s_base = l;
        expert_bounds[e] = l;
        if (e == n_expert - 1) expert_bounds[n_expert] = l + q;
    }
    __syncthreads();
    int running = s_base;

    const int64_t rounds = (n_pairs + MMID_THREADS - 1) / MMID_THREADS;
    for (int64_t it = 0; it < rounds; ++it) {
        const int64_t p = it * MMID_THREADS + threadIdx.x;
        bool mine = false;
        if (p < n_pairs) {
            const int64_t token = p / n_used;
            mine = __ldg(ids + token * ids_stride + (p - token * n_used)) == e;
        }
        const unsigned ballot = __ballot_sync(0xffffffffu, mine);
        if (lane == 0) s_warp[warp] = __popc(ballot);
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < MMID_THREADS / 32; ++w) { const int c = s_warp[w]; before += w < warp ? c : 0; total += c; }
        if (mine) ids_src1[running + before + __popc(ballot & ((1u << lane) - 1u))] = (int32_t) p;
        running += total;
        __syncthreads();
    }
}

}  // namespace

int b200_launch_mm_ids(b200_ctx * ctx, const int32_t * ids, int64_t ids_stride, int n_expert, int n_used, int64_t n_tokens,
                       int32_t * ids_src1, int32_t * expert_bounds) {
    if (n_expert <= 0 || n_used <= 0 || n_tokens < 0) { b200_set_error("mm_ids: bad sizes"); return B200_E_ARG; }
    mm_ids_kernel<<<n_expert, MMID_THREADS, 0, ctx->stream>>>(ids, ids_stride, n_expert, n_used, n_tokens * n_used, ids_src1, expert_bounds);
    ctx->launches++;
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}
