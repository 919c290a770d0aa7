// ops_move.cu — CPY / CONT / DUP, GET_ROWS (float and the four quantized weight types), SET_ROWS and the float MUL_MAT of attention
// (see ops.cu for the reference functions each one restates).
#include <cuda_bf16.h>
#include <math.h>

#include "ops.cuh"
#include "mm_problem.cuh"

int    b200_tc_supported(const b200_ctx * ctx, int type, const mm_problem & P);
size_t b200_gemm_tc_scratch(const b200_ctx * ctx, const mm_problem & P, int64_t n_pos, size_t ws_used);
int    b200_launch_gemm_tc(b200_ctx * ctx, int type, const mm_problem & P, int64_t n_pos, size_t ws_used);
int    b200_ctx_reserve(b200_ctx * ctx, size_t bytes);

namespace {

__device__ __forceinline__ float load_elem(const char * p, const int type) {
    switch (type) {
        case B200_TYPE_F32:  return *(const float *) p;
        case B200_TYPE_F16:  return __half2float(*(const __half *) p);
        default:             return __bfloat162float(*(const __nv_bfloat16 *) p);
    }
}
__device__ __forceinline__ void store_elem(char * p, const int type, const float v) {
    switch (type) {
        case B200_TYPE_F32:  *(float *) p = v; break;
        case B200_TYPE_F16:  *(__half *) p = __float2half_rn(v); break;
        default:             *(__nv_bfloat16 *) p = __float2bfloat16_rn(v); break;
    }
}

__device__ __forceinline__ float to_float(const float v) { return v; }
__device__ __forceinline__ float to_float(const __half v) { return __half2float(v); }
__device__ __forceinline__ float to_float(const __nv_bfloat16 v) { return __bfloat162float(v); }

// four consecutive elements from an address aligned to their size
__device__ __forceinline__ void load4(const float * p, float (&v)[4]) {
    const float4 t = *reinterpret_cast<const float4 *>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void load4(const __half * p, float (&v)[4]) {
    const uint2 t = *reinterpret_cast<const uint2 *>(p);
    const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&t.x)), b = __half22float2(*reinterpret_cast<const __half2 *>(&t.y));
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
__device__ __forceinline__ void load4(const __nv_bfloat16 * p, float (&v)[4]) {
    const uint2 t = *reinterpret_cast<const uint2 *>(p);
    v[0] = __uint_as_float(t.x << 16); v[1] = __uint_as_float(t.x & 0xFFFF0000u); v[2] = __uint_as_float(t.y << 16); v[3] = __uint_as_float(t.y & 0xFFFF0000u);
}

// ---- CPY: element i of src (in src's index order) goes to element i of dst (in dst's index order) -------------------------------
__global__ void __launch_bounds__(OPS_THREADS) cpy_kernel(const tview a, const int ta, const tview d, const int td, const int64_t n) {
    pdl_enter();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t a0, a1, a2, a3, d0, d1, d2, d3;
        split4(i, a.ne, a0, a1, a2, a3);
        split4(i, d.ne, d0, d1, d2, d3);
        const float v = load_elem(a.p + a0 * a.nb[0] + a1 * a.nb[1] + a2 * a.nb[2] + a3 * a.nb[3], ta);
        store_elem(d.p + d0 * d.nb[0] + d1 * d.nb[1] + d2 * d.nb[2] + d3 * d.nb[3], td, v);
    }
}

// the 4-bit code books kvalues_iq4nl / kvalues_mxfp4 (ggml-common.h:1110-1118) as bytes of two 64-bit constants (an indexed local array
// would live in local memory)
__device__ __forceinline__ int tab_iq4nl(const int i) {
    // -127, -104, -83, -65, -49, -35, -22, -10 | 1, 13, 25, 38, 53, 69, 89, 113
    const unsigned long long lo = 0xF6EADDCFBFAD9881ull, hi = 0x7159453526190D01ull;
    return (int) (int8_t) (((i & 8) ? hi : lo) >> (8 * (i & 7)));
}
__device__ __forceinline__ int tab_mxfp4(const int i) {
    // 0, 1, 2, 3, 4, 6, 8, 12 | 0, -1, -2, -3, -4, -6, -8, -12
    const unsigned long long lo = 0x0C08060403020100ull, hi = 0xF4F8FAFCFDFEFF00ull;
    return (int) (int8_t) (((i & 8) ? hi : lo) >> (8 * (i & 7)));
}

// CPY / CONT between tensors of the SAME shape whose rows are contiguous (a permuted view made contiguous, a type conversion): four
// elements per thread, one index split for both sides
__device__ __forceinline__ void store4(float * p, const float (&v)[4]) { *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]); }
__device__ __forceinline__ void store4(__half * p, const float (&v)[4]) {
    const __half2 a = __floats2half2_rn(v[0], v[1]), b = __floats2half2_rn(v[2], v[3]);
    *reinterpret_cast<uint2 *>(p) = make_uint2(*reinterpret_cast<const uint32_t *>(&a), *reinterpret_cast<const uint32_t *>(&b));
}
__device__ __forceinline__ void store4(__nv_bfloat16 * p, const float (&v)[4]) {
    const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
    *reinterpret_cast<uint2 *>(p) = make_uint2(*reinterpret_cast<const uint32_t *>(&a), *reinterpret_cast<const uint32_t *>(&b));
}

template <typename TS, typename TD>
__global__ void __launch_bounds__(OPS_THREADS) cpy_rows_kernel(const tview a, const tview d, const int64_t ne0q, const int64_t n4) {
    pdl_enter();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t) gridDim.x * blockDim.x) {
        const int64_t q = i % ne0q;
        int64_t r = i / ne0q;
        const int64_t i1 = r % a.ne[1]; r /= a.ne[1];
        const int64_t i2 = r % a.ne[2], i3 = r / a.ne[2];
        float v[4];
        load4((const TS *) (a.p + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]) + 4 * q, v);
        store4((TD *) (d.p + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]) + 4 * q, v);
    }
}

template <typename TS>
int launch_cpy_rows(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * d, const int64_t n4) {
    const int g = grid_for(n4, ctx->sm_count);
    switch (d->type) {
        case B200_TYPE_F32: OPS_LAUNCH(ctx, (cpy_rows_kernel<TS, float>), g, OPS_THREADS, view_of(a), view_of(d), a->ne[0] / 4, n4); break;
        case B200_TYPE_F16: OPS_LAUNCH(ctx, (cpy_rows_kernel<TS, __half>), g, OPS_THREADS, view_of(a), view_of(d), a->ne[0] / 4, n4); break;
        default:            OPS_LAUNCH(ctx, (cpy_rows_kernel<TS, __nv_bfloat16>), g, OPS_THREADS, view_of(a), view_of(d), a->ne[0] / 4, n4); break;
    }
    return finish(ctx);
}

// ---- one element of a row of any type GET_ROWS reads (dequantize_row_* of ggml-quants.c, element by element) -------------------
__device__ __forceinline__ float row_elem(const uint8_t * row, const int type, const int64_t i) {
    switch (type) {
        case B200_TYPE_F32: return ((const float *) row)[i];
        case B200_TYPE_F16: return __half2float(((const __half *) row)[i]);
        case B200_TYPE_BF16: return __bfloat162float(((const __nv_bfloat16 *) row)[i]);
        case B200_TYPE_Q4_0: {                                  // ggml-quants.c:314-332
            const uint8_t * b = row + (i >> 5) * SZ_Q4_0;
            const int j = (int) (i & 31);
            const int q = j < 16 ? (b[2 + j] & 0xF) : (b[2 + j - 16] >> 4);
            return (float) (q - 8) * ld_half(b);
        }
        case B200_TYPE_Q8_0: {                                  // ggml-quants.c:392-406
            const uint8_t * b = row + (i >> 5) * SZ_Q8_0;
            return (float) (int8_t) b[2 + (i & 31)] * ld_half(b);
        }
        case B200_TYPE_Q4_1: {                                  // ggml-quants.c:417-436
            const uint8_t * b = row + (i >> 5) * SZ_Q4_1;
            const int j = (int) (i & 31);
            const int q = j < 16 ? (b[4 + j] & 0xF) : (b[4 + j - 16] >> 4);
            return __fadd_rn(__fmul_rn((float) q, ld_half(b)), ld_half(b + 2));
        }
        case B200_TYPE_Q5_0: case B200_TYPE_Q5_1: {             // ggml-quants.c:438-490
            const bool one = type == B200_TYPE_Q5_1;
            const uint8_t * b = row + (i >> 5) * (one ? SZ_Q5_1 : SZ_Q5_0);
            const uint8_t * qhp = b + (one ? 4 : 2);
            const uint32_t qh = (uint32_t) qhp[0] | ((uint32_t) qhp[1] << 8) | ((uint32_t) qhp[2] << 16) | ((uint32_t) qhp[3] << 24);
            const int j = (int) (i & 31);
            const int nib = j < 16 ? (qhp[4 + j] & 0xF) : (qhp[4 + j - 16] >> 4);
            const int q = nib | (int) (((qh >> j) & 1u) << 4);
            return one ? __fadd_rn(__fmul_rn((float) q, ld_half(b)), ld_half(b + 2)) : (float) (q - 16) * ld_half(b);
        }
        case B200_TYPE_Q5_K: {                                  // ggml-quants.c:1669-1697
            const uint8_t * b = row + (i >> 8) * SZ_Q5_K;
            const int e = (int) (i & 255), j = e >> 5, l = e & 31;
            const float d = ld_half(b), dmin = ld_half(b + 2);
            const uint8_t * sc = b + 4;
            int s, m;
            if (j < 4) { s = sc[j] & 63; m = sc[j + 4] & 63; }
            else       { s = (sc[j + 4] & 0xF) | ((sc[j - 4] >> 6) << 4); m = (sc[j + 4] >> 4) | ((sc[j] >> 6) << 4); }
            const uint8_t qb = b[48 + 32 * (j >> 1) + l];
            const int q = ((j & 1) ? (qb >> 4) : (qb & 0xF)) + (((b[16 + l] >> j) & 1) ? 16 : 0);
            return __fsub_rn(__fmul_rn(__fmul_rn(d, (float) s), (float) q), __fmul_rn(dmin, (float) m));
        }
        case B200_TYPE_Q2_K: {                                  // ggml-quants.c:899-930
            const uint8_t * b = row + (i >> 8) * SZ_Q2_K;
            const int e = (int) (i & 255), j = e >> 5, l = e & 31, n = j >> 2, r = j & 3;
            const uint8_t sc = b[8 * n + 2 * r + (l >> 4)];
            const int q = (b[16 + 32 * n + l] >> (2 * r)) & 3;
            return __fsub_rn(__fmul_rn(__fmul_rn(ld_half(b + 80), (float) (sc & 0xF)), (float) q), __fmul_rn(ld_half(b + 82), (float) (sc >> 4)));
        }
        case B200_TYPE_Q3_K: {                                  // ggml-quants.c:1243-1290
            const uint8_t * b = row + (i >> 8) * SZ_Q3_K;
            const int e = (int) (i & 255), j = e >> 5, l = e & 31, n = j >> 2, r = j & 3, is = 2 * j + (l >> 4);
            const uint8_t * raw = b + 96;
            const int sc = (is < 8 ? (raw[is] & 0xF) : (raw[is - 8] >> 4)) | (((raw[8 + (is & 3)] >> (2 * (is >> 2))) & 3) << 4);
            const int q = ((b[32 + 32 * n + l] >> (2 * r)) & 3) - (((b[l] >> j) & 1) ? 0 : 4);
            return __fmul_rn(__fmul_rn(ld_half(b + 108), (float) (sc - 32)), (float) q);
        }
        case B200_TYPE_IQ4_NL: case B200_TYPE_IQ4_XS: case B200_TYPE_MXFP4: {   // ggml-quants.c:2649-2688, 507-525
            const int l = (int) (i & 31);
            if (type == B200_TYPE_IQ4_XS) {
                const uint8_t * b = row + (i >> 8) * SZ_IQ4_XS;
                const int ib = (int) (i & 255) >> 5;
                const int sh = b[2] | (b[3] << 8);
                const int ls = ((b[4 + ib / 2] >> (4 * (ib % 2))) & 0xF) | (((sh >> (2 * ib)) & 3) << 4);
                const uint8_t qb = b[8 + 16 * ib + (l & 15)];
                return __fmul_rn(__fmul_rn(ld_half(b), (float) (ls - 32)), (float) tab_iq4nl(l < 16 ? (qb & 0xF) : (qb >> 4)));
            }
            if (type == B200_TYPE_IQ4_NL) {
                const uint8_t * b = row + (i >> 5) * SZ_IQ4_NL;
                const uint8_t qb = b[2 + (l & 15)];
                return __fmul_rn(ld_half(b), (float) tab_iq4nl(l < 16 ? (qb & 0xF) : (qb >> 4)));
            }
            const uint8_t * b = row + (i >> 5) * SZ_MXFP4;
            const uint32_t e8 = b[0];
            const uint8_t qb = b[1 + (l & 15)];
            return __fmul_rn((float) tab_mxfp4(l < 16 ? (qb & 0xF) : (qb >> 4)), __uint_as_float(e8 < 2 ? 0x00200000u << e8 : (e8 - 1) << 23));
        }
        case B200_TYPE_NVFP4: {                                 // ggml-quants.c:527-550
            const uint8_t * b = row + (i >> 6) * SZ_NVFP4;
            const int e = (int) (i & 63), s = e >> 4, j = e & 15;
            const uint32_t x = b[s];
            const uint32_t ex = (x >> 3) & 0xFu, man = x & 7u;
            const float d = (x == 0 || x == 0x7F) ? 0.0f : (ex == 0 ? (float) man * 0.0009765625f : __uint_as_float(((ex + 119u) << 23) | (man << 20)));
            const uint8_t qb = b[4 + 8 * s + (j & 7)];
            return __fmul_rn((float) tab_mxfp4(j < 8 ? (qb & 0xF) : (qb >> 4)), d);
        }
        case B200_TYPE_Q4_K: {                                  // ggml-quants.c:1467-1489
            const uint8_t * b = row + (i >> 8) * SZ_Q4_K;
            const int e = (int) (i & 255), j = e >> 5, l = e & 31;       // sub-block j, element l
            const float d = ld_half(b), dmin = ld_half(b + 2);
            const uint8_t * sc = b + 4;
            int s, m;
            if (j < 4) { s = sc[j] & 63; m = sc[j + 4] & 63; }
            else       { s = (sc[j + 4] & 0xF) | ((sc[j - 4] >> 6) << 4); m = (sc[j + 4] >> 4) | ((sc[j] >> 6) << 4); }
            const uint8_t qb = b[16 + 32 * (j >> 1) + l];
            const int q = (j & 1) ? (qb >> 4) : (qb & 0xF);
            return __fsub_rn(__fmul_rn(__fmul_rn(d, (float) s), (float) q), __fmul_rn(dmin, (float) m));      // no FMA contraction: the bits of dequantize_row_q4_K
        }
        default: {                                               // Q6_K, ggml-quants.c:1877-1906
            const uint8_t * b = row + (i >> 8) * SZ_Q6_K;
            const int e = (int) (i & 255), n = e >> 7, r = (e & 127) >> 5, l = e & 31;
            const uint8_t ql = b[64 * n + 32 * (r & 1) + l], qh = b[128 + 32 * n + l];
            const int q = (int) (((r >> 1) ? (ql >> 4) : (ql & 0xF)) | (((qh >> (2 * r)) & 3) << 4)) - 32;
            const int8_t sc = (int8_t) b[192 + 8 * n + 2 * r + l / 16];
            return ld_half(b + 208) * (float) sc * (float) q;
        }
    }
}

// dst[i0, i10, i11, i12] = src0[i0, idx[i10, i11, i12], i11, i12]
__global__ void __launch_bounds__(OPS_THREADS) get_rows_kernel(const tview a, const int ta, const tview idx, const tview d, const int64_t n) {
    pdl_enter();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t i0, i10, i11, i12;
        split4(i, d.ne, i0, i10, i11, i12);
        const int64_t r = *(const int32_t *) (idx.p + i10 * idx.nb[0] + i11 * idx.nb[1] + i12 * idx.nb[2]);
        float v = 0.0f;
        if (r >= 0 && r < a.ne[1]) v = row_elem((const uint8_t *) (a.p + r * a.nb[1] + i11 * a.nb[2] + i12 * a.nb[3]), ta, i0);
        *(float *) (d.p + i0 * 4 + i10 * d.nb[1] + i11 * d.nb[2] + i12 * d.nb[3]) = v;
    }
}

// dst[i0, idx[i, i02 % ne11, i03 % ne12], i02, i03] = src0[i0, i, i02, i03]
template <typename IDX>
__global__ void __launch_bounds__(OPS_THREADS) set_rows_kernel(const tview a, const tview idx, const tview d, const int td, const int64_t n) {
    pdl_enter();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t i0, i1, i2, i3;
        split4(i, a.ne, i0, i1, i2, i3);
        const int64_t r = (int64_t) *(const IDX *) (idx.p + i1 * idx.nb[0] + (i2 % idx.ne[1]) * idx.nb[1] + (i3 % idx.ne[2]) * idx.nb[2]);
        if (r < 0 || r >= d.ne[1]) continue;
        const float v = *(const float *) (a.p + i0 * 4 + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
        store_elem(d.p + i0 * d.nb[0] + r * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3], td, v);
    }
}

// ---- float MUL_MAT: dst[i0, i1, i2, i3] = sum_k src0[k, i0, i2 / r2, i3 / r3] * src1[k, i1, i2, i3]; one warp per output element ------
// (the attention products of a decode step: rows of 64-128 values against thousands of keys — HBM / L2 streaming, no reuse to tile for)
template <typename T>
__global__ void __launch_bounds__(OPS_THREADS) mul_mat_f_kernel(const tview a, const tview b, const tview d, const int64_t r2, const int64_t r3,
                                                                const int64_t n_out) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t) gridDim.x * (OPS_THREADS / 32);
    for (int64_t o = (int64_t) blockIdx.x * (OPS_THREADS / 32) + (threadIdx.x >> 5); o < n_out; o += warps) {
        int64_t i0, i1, i2, i3;
        split4(o, d.ne, i0, i1, i2, i3);
        const T *     w = (const T *) (a.p + i0 * a.nb[1] + (i2 / r2) * a.nb[2] + (i3 / r3) * a.nb[3]);
        const float * x = (const float *) (b.p + i1 * b.nb[1] + i2 * b.nb[2] + i3 * b.nb[3]);
        float s = 0.0f;
        const int64_t K = a.ne[0];
        int64_t k_done = 0;
        // four elements per lane and step where the row starts allow it (a decode step's K.q and V.p: 128 / n_kv elements per row — the
        // scalar loop was a chain of 2-byte loads)
        if (((uintptr_t) w % (4 * sizeof(T))) == 0 && ((uintptr_t) x & 15) == 0) {
            const int64_t K4 = K & ~(int64_t) 3;
            for (int64_t k = 4 * lane; k < K4; k += 128) {
                float wv[4];
                load4(w + k, wv);
                const float4 xv = *reinterpret_cast<const float4 *>(x + k);
                s += wv[0] * xv.x + wv[1] * xv.y + wv[2] * xv.z + wv[3] * xv.w;
            }
            k_done = K4;
        }
        for (int64_t k = k_done + lane; k < K; k += 32) s += to_float(w[k]) * x[k];
        s = warp_sum(s);
        if (lane == 0) *(float *) (d.p + i0 * 4 + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]) = s;
    }
}

// the same product for many activation columns (prefill attention: hundreds of query rows against hundreds of keys per head): a 64 x 64
// output tile per CTA, K in steps of 16 through shared memory, 4 x 4 outputs per thread, fp32 accumulation
constexpr int MMF_T = 64, MMF_K = 16;
template <typename T>
__global__ void __launch_bounds__(256) mul_mat_f_tiled_kernel(const tview a, const tview b, const tview d, const int64_t r2, const int64_t r3) {
    pdl_enter();
    __shared__ float sa[MMF_K][MMF_T + 1], sb[MMF_K][MMF_T + 1];
    const int64_t batch = blockIdx.z, i2 = batch % d.ne[2], i3 = batch / d.ne[2];
    const int64_t row0 = (int64_t) blockIdx.x * MMF_T, col0 = (int64_t) blockIdx.y * MMF_T;
    const char * ap = a.p + (i2 / r2) * a.nb[2] + (i3 / r3) * a.nb[3];
    const char * bp = b.p + i2 * b.nb[2] + i3 * b.nb[3];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;           // outputs rows 4 tx .. +3, columns 4 ty .. +3
    const int lk = threadIdx.x & 15, lr = threadIdx.x >> 4;           // loader: element lk of rows lr, lr + 16, lr + 32, lr + 48
    float acc[4][4] = {};
    const int64_t K = a.ne[0];
    for (int64_t k0 = 0; k0 < K; k0 += MMF_K) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int64_t r = row0 + lr + 16 * j, c = col0 + lr + 16 * j, k = k0 + lk;
            sa[lk][lr + 16 * j] = (r < a.ne[1] && k < K) ? to_float(((const T *) (ap + r * a.nb[1]))[k]) : 0.0f;
            sb[lk][lr + 16 * j] = (c < b.ne[1] && k < K) ? ((const float *) (bp + c * b.nb[1]))[k] : 0.0f;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < MMF_K; ++k) {
            float av[4], bv[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) { av[j] = sa[k][4 * tx + j]; bv[j] = sb[k][4 * ty + j]; }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] += av[i] * bv[j];
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int64_t c = col0 + 4 * ty + j;
        if (c >= d.ne[1]) continue;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int64_t r = row0 + 4 * tx + i;
            if (r < d.ne[0]) *(float *) (d.p + r * 4 + c * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]) = acc[i][j];
        }
    }
}

// rows of any quantized type -> fp16, natural order (the prefill route of the weight types the tcgen05 GEMM has no converter for: abi.cu).
// One thread per 8 consecutive elements (always inside one block) and one 16-byte store.
__global__ void __launch_bounds__(OPS_THREADS) dequant_f16_kernel(const uint8_t * __restrict__ W, const int type, const int64_t nb01, const int64_t k8,
                                                                  const int64_t n8, __half * __restrict__ out) {
    pdl_enter();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t) gridDim.x * blockDim.x) {
        const int64_t r = i / k8, c = (i - r * k8) * 8;
        const uint8_t * row = W + r * nb01;
        __half2 h[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) h[j] = __floats2half2_rn(row_elem(row, type, c + 2 * j), row_elem(row, type, c + 2 * j + 1));
        *reinterpret_cast<uint4 *>(out + i * 8) = *reinterpret_cast<const uint4 *>(h);
    }
}

}  // namespace

// rows x k elements of `type` at W (row stride nb01 bytes) -> out[rows][k] fp16; k a multiple of 8, out 16-byte aligned
int b200_launch_dequant_f16(b200_ctx * ctx, int type, const uint8_t * W, int64_t nb01, int64_t rows, int64_t k, void * out) {
    const int64_t n8 = rows * (k / 8);
    if (n8 <= 0) return B200_OK;
    OPS_LAUNCH(ctx, dequant_f16_kernel, grid_for(n8, ctx->sm_count), OPS_THREADS, W, type, nb01, k / 8, n8, (__half *) out);
    return finish(ctx);
}

extern "C" {

static bool cpy_type_ok(int t) { return t == B200_TYPE_F32 || t == B200_TYPE_F16 || t == B200_TYPE_BF16; }

int b200_op_cpy(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * d) {
    if (!ctx || !a || !d) { b200_set_error("b200_op_cpy: null argument"); return B200_E_ARG; }
    if (!cpy_type_ok(a->type) || !cpy_type_ok(d->type) || nelem(a) != nelem(d)) {
        b200_set_error("b200_op_cpy: f32 / f16 / bf16 tensors with the same number of elements"); return B200_E_UNSUPPORTED;
    }
    const int64_t n = nelem(a);
    if (n == 0) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    auto esz = [](int t) { return t == B200_TYPE_F32 ? 4 : 2; };
    auto al = [&](const b200_tensor * t) {                 // four elements at a time: every row starts on a multiple of 4 elements' bytes
        const uintptr_t m = (uintptr_t) (4 * esz(t->type) - 1);
        return t->nb[0] == esz(t->type) && (((uintptr_t) t->data | (uintptr_t) t->nb[1] | (uintptr_t) t->nb[2] | (uintptr_t) t->nb[3]) & m) == 0;
    };
    if (same_shape(a, d) && a->ne[0] % 4 == 0 && al(a) && al(d)) {
        switch (a->type) {
            case B200_TYPE_F32: return launch_cpy_rows<float>(ctx, a, d, n / 4);
            case B200_TYPE_F16: return launch_cpy_rows<__half>(ctx, a, d, n / 4);
            default:            return launch_cpy_rows<__nv_bfloat16>(ctx, a, d, n / 4);
        }
    }
    OPS_LAUNCH(ctx, cpy_kernel, grid_for(n, ctx->sm_count), OPS_THREADS, view_of(a), a->type, view_of(d), d->type, n);
    return finish(ctx);
}

int b200_op_get_rows(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * idx, const b200_tensor * d) {
    if (!ctx || !a || !idx || !d) { b200_set_error("b200_op_get_rows: null argument"); return B200_E_ARG; }
    const bool type_ok = cpy_type_ok(a->type) || a->type == B200_TYPE_Q4_0 || a->type == B200_TYPE_Q8_0 || a->type == B200_TYPE_Q4_K || a->type == B200_TYPE_Q6_K ||
                         a->type == B200_TYPE_Q4_1 || a->type == B200_TYPE_Q5_0 || a->type == B200_TYPE_Q5_1 || a->type == B200_TYPE_Q5_K ||
                         a->type == B200_TYPE_Q2_K || a->type == B200_TYPE_Q3_K || a->type == B200_TYPE_IQ4_NL || a->type == B200_TYPE_IQ4_XS || a->type == B200_TYPE_MXFP4 ||
                         a->type == B200_TYPE_NVFP4;
    if (!type_ok || idx->type != B200_TYPE_I32 || d->type != B200_TYPE_F32 || d->nb[0] != 4 || d->ne[0] != a->ne[0] ||
        d->ne[1] != idx->ne[0] || d->ne[2] != idx->ne[1] || d->ne[3] != idx->ne[2] || a->ne[0] % b200_blck_size(a->type) != 0) {
        b200_set_error("b200_op_get_rows: f32 / f16 / bf16 / q4_0 / q8_0 / q4_K / q6_K rows, i32 indices, f32 result"); return B200_E_UNSUPPORTED;
    }
    const int64_t n = nelem(d);
    if (n == 0) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    OPS_LAUNCH(ctx, get_rows_kernel, grid_for(n, ctx->sm_count), OPS_THREADS, view_of(a), a->type, view_of(idx), view_of(d), n);
    return finish(ctx);
}

int b200_op_set_rows(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * idx, const b200_tensor * d) {
    if (!ctx || !a || !idx || !d) { b200_set_error("b200_op_set_rows: null argument"); return B200_E_ARG; }
    if (a->type != B200_TYPE_F32 || a->nb[0] != 4 || (idx->type != B200_TYPE_I64 && idx->type != B200_TYPE_I32) || !cpy_type_ok(d->type) ||
        d->ne[0] != a->ne[0] || d->ne[2] != a->ne[2] || d->ne[3] != a->ne[3] || idx->ne[0] != a->ne[1] || idx->ne[1] <= 0 || idx->ne[2] <= 0 ||
        a->ne[2] % idx->ne[1] != 0 || a->ne[3] % idx->ne[2] != 0) {
        b200_set_error("b200_op_set_rows: f32 rows, i64 / i32 indices, f32 / f16 / bf16 destination"); return B200_E_UNSUPPORTED;
    }
    const int64_t n = nelem(a);
    if (n == 0) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    const int g = grid_for(n, ctx->sm_count);
    if (idx->type == B200_TYPE_I64) OPS_LAUNCH(ctx, set_rows_kernel<int64_t>, g, OPS_THREADS, view_of(a), view_of(idx), view_of(d), d->type, n);
    else                            OPS_LAUNCH(ctx, set_rows_kernel<int32_t>, g, OPS_THREADS, view_of(a), view_of(idx), view_of(d), d->type, n);
    return finish(ctx);
}

int b200_mul_mat_f_supported(const b200_tensor * a, const b200_tensor * b, const b200_tensor * d) {
    if (!a || !b || !d) return 0;
    if (!cpy_type_ok(a->type) || b->type != B200_TYPE_F32 || d->type != B200_TYPE_F32) return 0;
    if (a->nb[0] != (a->type == B200_TYPE_F32 ? 4 : 2) || b->nb[0] != 4 || d->nb[0] != 4) return 0;
    if (a->ne[0] != b->ne[0] || d->ne[0] != a->ne[1] || d->ne[1] != b->ne[1] || d->ne[2] != b->ne[2] || d->ne[3] != b->ne[3]) return 0;
    if (a->ne[2] <= 0 || a->ne[3] <= 0 || b->ne[2] % a->ne[2] != 0 || b->ne[3] % a->ne[3] != 0) return 0;
    return 1;
}

int b200_mul_mat_f(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * b, const b200_tensor * d) {
    if (!ctx) { b200_set_error("b200_mul_mat_f: null context"); return B200_E_ARG; }
    if (!b200_mul_mat_f_supported(a, b, d)) { b200_set_error("b200_mul_mat_f: f32 / f16 / bf16 src0 rows x f32 src1 rows -> f32"); return B200_E_UNSUPPORTED; }
    const int64_t n_out = nelem(d);
    if (n_out == 0) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    const int64_t r2t = b->ne[2] / a->ne[2], r3t = b->ne[3] / a->ne[3];
    // many columns against fp16 rows — the weights of an F16 model, or K / V of a KV cache at prompt time (a batch of strided views with
    // grouped-query broadcast): the tcgen05 GEMM reads the rows as they are through its tensor map (w_frag<F16>, mmq_tc.cu); src1 goes to
    // normalised fp16 like for the quantized types (ggml-cpu converts src1 to fp16 for this product as well)
    if (a->type == B200_TYPE_F16 && b->ne[1] >= 17 && ctx->use_tc && !ctx->force_generic && a->ne[3] == 1 && b->ne[3] == 1 &&
        a->nb[0] == 2 && b->nb[0] == 4 && contiguous(d, 4)) {
        mm_problem P = {};
        P.W = (const uint8_t *) a->data; P.nb01 = a->nb[1]; P.nb02 = a->ne[2] > 1 ? a->nb[2] : a->nb[1] * a->ne[1]; P.nb03 = P.nb02 * a->ne[2];
        P.m = a->ne[1]; P.k = a->ne[0];
        P.X = (const float *) b->data; P.nb11 = b->nb[1]; P.nb12 = b->ne[2] > 1 ? b->nb[2] : b->nb[1] * b->ne[1]; P.nb13 = P.nb12 * b->ne[2];
        P.ne11 = b->ne[1]; P.ne12 = b->ne[2]; P.ne13 = 1; P.r2 = r2t; P.r3 = 1;
        P.Y = (float *) d->data; P.mode = MM_PLAIN;
        const int64_t n_pos = b->ne[1] * b->ne[2];
        if (n_pos < ((int64_t) 1 << 24) && b200_tc_supported(ctx, B200_TYPE_F16, P)) {
            const int rc = b200_ctx_reserve(ctx, b200_gemm_tc_scratch(ctx, P, n_pos, 0));
            if (rc) return rc;
            ctx->n_pend = 16;
            return b200_launch_gemm_tc(ctx, B200_TYPE_F16, P, n_pos, 0);
        }
    }
    if (b->ne[1] > 8 && d->ne[2] * d->ne[3] <= 65535) {            // many columns: the tiled kernel
        const dim3 grid((unsigned) ((d->ne[0] + MMF_T - 1) / MMF_T), (unsigned) ((d->ne[1] + MMF_T - 1) / MMF_T), (unsigned) (d->ne[2] * d->ne[3]));
        switch (a->type) {
            case B200_TYPE_F32: OPS_LAUNCH(ctx, mul_mat_f_tiled_kernel<float>, grid, 256, view_of(a), view_of(b), view_of(d), r2t, r3t); break;
            case B200_TYPE_F16: OPS_LAUNCH(ctx, mul_mat_f_tiled_kernel<__half>, grid, 256, view_of(a), view_of(b), view_of(d), r2t, r3t); break;
            default:            OPS_LAUNCH(ctx, mul_mat_f_tiled_kernel<__nv_bfloat16>, grid, 256, view_of(a), view_of(b), view_of(d), r2t, r3t); break;
        }
        return finish(ctx);
    }
    const int64_t blocks = (n_out + OPS_THREADS / 32 - 1) / (OPS_THREADS / 32), cap = (int64_t) ctx->sm_count * 16;
    const int g = (int) (blocks < cap ? blocks : cap);
    const int64_t r2 = b->ne[2] / a->ne[2], r3 = b->ne[3] / a->ne[3];
    switch (a->type) {
        case B200_TYPE_F32: OPS_LAUNCH(ctx, mul_mat_f_kernel<float>, g, OPS_THREADS, view_of(a), view_of(b), view_of(d), r2, r3, n_out); break;
        case B200_TYPE_F16: OPS_LAUNCH(ctx, mul_mat_f_kernel<__half>, g, OPS_THREADS, view_of(a), view_of(b), view_of(d), r2, r3, n_out); break;
        default:            OPS_LAUNCH(ctx, mul_mat_f_kernel<__nv_bfloat16>, g, OPS_THREADS, view_of(a), view_of(b), view_of(d), r2, r3, n_out); break;
    }
    return finish(ctx);
}

}  // extern "C"
