// quantize.cuh — the Q8_1 activation quantizer, bit-exact against quantize_row_q8_1_ref
// (/root/reference/ggml/src/ggml-quants.c:260-293):
//     amax = max |x_j| over the 32 values;  d = amax / 127;  id = d ? 1/d : 0
//     q_j  = (int8) roundf(x_j * id)            -- round half AWAY from zero
//     y.d  = fp16(d);  y.s = fp16((float) sum(q_j) * d)   -- s uses the fp32 d
// Everything below uses explicit IEEE intrinsics (__fdiv_rn, __fmul_rn) so that neither fast-math nor FMA
// contraction can change a bit.  Note this is NOT what ggml-cuda's quantize_q8_1 computes
// (ggml/src/ggml-cuda/quantize.cu:31-47 stores s = sum of the unquantized x and divides x/d).
//
// Work split: 8 consecutive lanes own one 32-value block, 4 values (one float4) per lane; the max and the
// integer sum are reduced with three xor-shuffles that stay inside the aligned 8-lane group.
#pragma once
#include "common.cuh"

struct q8_part {
    uint32_t packed;   // this lane's 4 int8 values, little-endian (value j of the block at byte j % 4 of lane j / 4)
    float    d;        // fp32 scale of the whole block (identical in the 8 lanes)
    int      sum;      // sum of all 32 quantized values (identical in the 8 lanes)
};

// all 32 lanes of the warp must call this (full-mask shuffles)
__device__ __forceinline__ q8_part quantize_q8_1_part(float4 v) {
    float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    const int q0 = (int) roundf(__fmul_rn(v.x, id));
    const int q1 = (int) roundf(__fmul_rn(v.y, id));
    const int q2 = (int) roundf(__fmul_rn(v.z, id));
    const int q3 = (int) roundf(__fmul_rn(v.w, id));
    int sum = q0 + q1 + q2 + q3;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    q8_part r;
    r.packed = (uint32_t) (q0 & 0xff) | ((uint32_t) (q1 & 0xff) << 8) | ((uint32_t) (q2 & 0xff) << 16) | ((uint32_t) (q3 & 0xff) << 24);
    r.d      = d;
    r.sum    = sum;
    return r;
}

// (d, s) of the block as the two fp16 fields of block_q8_1
__device__ __forceinline__ half2 q8_1_ds(const q8_part & p) {
    return __halves2half2(__float2half_rn(p.d), __float2half_rn(__fmul_rn((float) p.sum, p.d)));
}

// host launchers (quantize.cu)
// standard block_q8_1 output (the public ABI): x is [k, ne1, ne2, ne3] f32 with byte strides nb1..nb3
int b200_launch_quantize_aos(b200_ctx * ctx, const float * x, int64_t k, int64_t ne1, int64_t ne2, int64_t ne3,
                             int64_t nb1, int64_t nb2, int64_t nb3, void * y);
// plane output for the matmul kernels; columns are packed in (i1, i2, i3) order
int b200_launch_quantize_planes(b200_ctx * ctx, const float * x, int64_t k, int64_t ne1, int64_t ne2, int64_t ne3,
                                int64_t nb1, int64_t nb2, int64_t nb3, int8_t * qs, half2 * ds);
