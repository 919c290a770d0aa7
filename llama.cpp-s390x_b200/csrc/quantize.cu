// quantize.cu — standalone launches of the Q8_1 activation quantizer (see quantize.cuh for the arithmetic).
// Replaces quantize_row_q8_1 on the reference's path (/root/reference/ggml/src/ggml-quants.c:260-293; the CUDA
// analogue is quantize_row_q8_1_cuda, ggml/src/ggml-cuda/quantize.cu:369-383).
// This translation unit must never be compiled with -use_fast_math.
#include "quantize.cuh"

namespace {

// One 8-lane group per 32-value block; grid-stride over blocks.  HBM-bound: 128 B in, 36 B out per block.
//   AOS : y is block_q8_1[] (36 B each);  !AOS : qs / ds planes
//   VEC : rows are 16-byte aligned -> one 128-bit load per lane
template <bool AOS, bool VEC>
__global__ void __launch_bounds__(256)
quantize_q8_1_kernel(const float * __restrict__ x, const int64_t nbk /* blocks per row */, const int64_t ne1, const int64_t ne2,
                     const int64_t nb1, const int64_t nb2, const int64_t nb3, const int64_t total_blocks,
                     uint8_t * __restrict__ y_aos, int8_t * __restrict__ y_qs, half2 * __restrict__ y_ds) {
    const int     sub     = threadIdx.x & 7;
    const int64_t groups  = (int64_t) gridDim.x * (blockDim.x >> 3);
    const int64_t rounds  = (total_blocks + groups - 1) / groups;       // identical for every thread: shuffles stay converged
    int64_t       blk     = (int64_t) blockIdx.x * (blockDim.x >> 3) + (threadIdx.x >> 3);
    for (int64_t it = 0; it < rounds; ++it, blk += groups) {
        const bool    live = blk < total_blocks;
        const int64_t b    = live ? blk : total_blocks - 1;
        const int64_t col  = b / nbk;
        const int64_t ib   = b - col * nbk;
        const int64_t i1   = col % ne1;
        const int64_t i2   = (col / ne1) % ne2;
        const int64_t i3   = col / (ne1 * ne2);
        const float * p = reinterpret_cast<const float *>(reinterpret_cast<const char *>(x) + i1 * nb1 + i2 * nb2 + i3 * nb3)
                          + ib * 32 + sub * 4;
        float4 v;
        if (VEC) {
            v = __ldg(reinterpret_cast<const float4 *>(p));
        } else {
            v.x = __ldg(p); v.y = __ldg(p + 1); v.z = __ldg(p + 2); v.w = __ldg(p + 3);
        }
        const q8_part q = quantize_q8_1_part(v);
        if (!live) continue;
        if (AOS) {
            uint8_t * o = y_aos + b * SZ_Q8_1;                           // 36 B blocks: 4-byte aligned
            reinterpret_cast<uint32_t *>(o + 4)[sub] = q.packed;
            if (sub == 0) *reinterpret_cast<half2 *>(o) = q8_1_ds(q);
        } else {
            reinterpret_cast<uint32_t *>(y_qs + b * 32)[sub] = q.packed;
            if (sub == 0) y_ds[b] = q8_1_ds(q);
        }
    }
}

template <bool AOS>
int launch(b200_ctx * ctx, const float * x, int64_t k, int64_t ne1, int64_t ne2, int64_t ne3, int64_t nb1, int64_t nb2,
           int64_t nb3, uint8_t * y_aos, int8_t * y_qs, half2 * y_ds) {
    if (k <= 0 || k % 32 != 0 || ne1 < 0 || ne2 < 0 || ne3 < 0) {
        b200_set_error("quantize_q8_1: k=%lld must be a positive multiple of 32", (long long) k);
        return B200_E_ARG;
    }
    const int64_t nbk   = k / 32;
    const int64_t total = nbk * ne1 * ne2 * ne3;
    if (total == 0) return B200_OK;
    const bool vec = (((uintptr_t) x | (uintptr_t) nb1 | (uintptr_t) nb2 | (uintptr_t) nb3) & 15) == 0;
    const int     threads = 256;
    const int64_t want    = (total + (threads / 8) - 1) / (threads / 8);
    const int64_t cap     = (int64_t) ctx->sm_count * 8 * 4;             // 8 resident CTAs per SM, a few waves
    const int     grid    = (int) (want < cap ? want : cap);
    if (vec) quantize_q8_1_kernel<AOS, true ><<<grid, threads, 0, ctx->stream>>>(x, nbk, ne1, ne2, nb1, nb2, nb3, total, y_aos, y_qs, y_ds);
    else     quantize_q8_1_kernel<AOS, false><<<grid, threads, 0, ctx->stream>>>(x, nbk, ne1, ne2, nb1, nb2, nb3, total, y_aos, y_qs, y_ds);
    ctx->launches++;
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

}  // namespace

int b200_launch_quantize_aos(b200_ctx * ctx, const float * x, int64_t k, int64_t ne1, int64_t ne2, int64_t ne3,
                             int64_t nb1, int64_t nb2, int64_t nb3, void * y) {
    return launch<true>(ctx, x, k, ne1, ne2, ne3, nb1, nb2, nb3, (uint8_t *) y, nullptr, nullptr);
}

int b200_launch_quantize_planes(b200_ctx * ctx, const float * x, int64_t k, int64_t ne1, int64_t ne2, int64_t ne3,
                                int64_t nb1, int64_t nb2, int64_t nb3, int8_t * qs, half2 * ds) {
    return launch<false>(ctx, x, k, ne1, ne2, ne3, nb1, nb2, nb3, nullptr, qs, ds);
}
