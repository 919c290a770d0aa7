// ops_rows.cu — the row-wise ops of a Llama graph: RMS_NORM, ROPE, SOFT_MAX, CPY / CONT / DUP, GET_ROWS, SET_ROWS and the float MUL_MAT of
// attention (see ops.cu for the reference functions each one restates).
#include <cuda_bf16.h>
#include <math.h>

#include "ops.cuh"

namespace {

__device__ __forceinline__ float block_sum(float v, float * red) {          // CTA of OPS_THREADS
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.0f;
#pragma unroll
    for (int w = 0; w < OPS_THREADS / 32; ++w) t += red[w];
    return t;
}
__device__ __forceinline__ float block_max(float v, float * red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = -INFINITY;
#pragma unroll
    for (int w = 0; w < OPS_THREADS / 32; ++w) t = fmaxf(t, red[w]);
    return t;
}

// ---- RMS_NORM: y = x / sqrt(mean(x^2) + eps), one CTA per row -------------------------------------------------------------
// (w.p != NULL: the product with a weight tensor repeated into x's shape — RMS_NORM followed by MUL, as every norm of a Llama graph is)
__global__ void __launch_bounds__(OPS_THREADS) rms_norm_kernel(const tview a, const tview w, const tview d, const float eps) {
    pdl_enter();
    __shared__ float red[OPS_THREADS / 32];
    const int64_t r = blockIdx.x;
    const int64_t i1 = r % a.ne[1], i2 = (r / a.ne[1]) % a.ne[2], i3 = r / (a.ne[1] * a.ne[2]);
    const float * x = (const float *) (a.p + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
    float *       y = (float *) (d.p + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]);
    const int64_t n = a.ne[0];
    float s = 0.0f;
    for (int64_t i = threadIdx.x; i < n; i += OPS_THREADS) { const float v = x[i]; s += v * v; }
    s = block_sum(s, red);
    const float scale = 1.0f / sqrtf(s / (float) n + eps);
    if (w.p) {
        const char * wr = w.p + (i1 % w.ne[1]) * w.nb[1] + (i2 % w.ne[2]) * w.nb[2] + (i3 % w.ne[3]) * w.nb[3];
        if (w.ne[0] == n && w.nb[0] == 4) {           // the usual case: one weight per channel (no 64-bit modulo per element)
            const float * wf = (const float *) wr;
            for (int64_t i = threadIdx.x; i < n; i += OPS_THREADS) y[i] = (x[i] * scale) * wf[i];
        } else {
            for (int64_t i = threadIdx.x; i < n; i += OPS_THREADS) y[i] = (x[i] * scale) * *(const float *) (wr + (i % w.ne[0]) * w.nb[0]);
        }
    } else {
        for (int64_t i = threadIdx.x; i < n; i += OPS_THREADS) y[i] = x[i] * scale;
    }
}

// the same for rows of float4s (16-byte aligned rows, one weight per channel): 1024 threads, the first two float4 of every thread stay in
// registers between the two passes — a 4096-value row is ONE load per thread, against a 16-deep serial load chain per thread in the kernel
// above (a decode step normalises single rows: the latency of that chain is the kernel)
// Many rows at once (a prompt: 512 rows of 4096) take 256 threads per row and four float4 in registers instead: more rows in flight per SM
// and an 8-warp instead of a 32-warp reduction.
constexpr int RMS_VEC_THREADS = 1024;
template <int THREADS, int KEEP>
__global__ void __launch_bounds__(THREADS) rms_norm_vec_kernel(const tview a, const tview w, const tview d, const float eps) {
    pdl_enter();
    __shared__ float red[THREADS / 32];
    const int64_t r = blockIdx.x;
    const int64_t i1 = r % a.ne[1], i2 = (r / a.ne[1]) % a.ne[2], i3 = r / (a.ne[1] * a.ne[2]);
    const float4 * x = (const float4 *) (a.p + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
    float4 *       y = (float4 *) (d.p + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]);
    const float4 * wf = w.p ? (const float4 *) (w.p + (i1 % w.ne[1]) * w.nb[1] + (i2 % w.ne[2]) * w.nb[2] + (i3 % w.ne[3]) * w.nb[3]) : nullptr;
    const int64_t n4 = a.ne[0] >> 2;
    const int tid = threadIdx.x;
    float4 keep[KEEP];
    float s = 0.0f;
#pragma unroll
    for (int j = 0; j < KEEP; ++j) {
        const int64_t i = tid + (int64_t) j * THREADS;
        keep[j] = i < n4 ? x[i] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        s += keep[j].x * keep[j].x + keep[j].y * keep[j].y + keep[j].z * keep[j].z + keep[j].w * keep[j].w;
    }
    for (int64_t i = tid + (int64_t) KEEP * THREADS; i < n4; i += THREADS) { const float4 v = x[i]; s += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w; }
    s = warp_sum(s);
    if ((tid & 31) == 0) red[tid >> 5] = s;
    __syncthreads();
    float t = 0.0f;
#pragma unroll
    for (int k = 0; k < THREADS / 32; ++k) t += red[k];
    const float scale = 1.0f / sqrtf(t / (float) a.ne[0] + eps);
    auto emit = [&](const int64_t i, float4 v) {
        v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
        if (wf) { const float4 g = wf[i]; v.x *= g.x; v.y *= g.y; v.z *= g.z; v.w *= g.w; }
        y[i] = v;
    };
#pragma unroll
    for (int j = 0; j < KEEP; ++j) { const int64_t i = tid + (int64_t) j * THREADS; if (i < n4) emit(i, keep[j]); }
    for (int64_t i = tid + (int64_t) KEEP * THREADS; i < n4; i += THREADS) emit(i, x[i]);
}

// ---- ROPE (normal and NeoX pair layouts, YaRN, frequency factors) ------------------------------------------------------------
struct rope_args {
    int     n_dims, mode;
    float   theta_scale, freq_scale, ext_factor, attn_factor, corr0, corr1;
    const int32_t * pos;
    const float *   freq_factors;
};

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

// one thread per pair of a row; rows = (i1 head, i2 position, i3 batch)
template <typename T, typename IDX>
__global__ void __launch_bounds__(OPS_THREADS) rope_kernel(const tview a, const tview d, const rope_args R, const int64_t n) {
    pdl_enter();
    const int64_t half0 = a.ne[0] / 2;
    // (IDX = uint32_t whenever the pair count fits: the 64-bit divisions of the index split were most of this kernel's instructions)
    const IDX h0 = (IDX) half0, ne1 = (IDX) a.ne[1], ne2 = (IDX) a.ne[2];
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const IDX ii = (IDX) i;
        const int64_t pr = (int64_t) (ii % h0);
        IDX r = ii / h0;
        const int64_t i1 = (int64_t) (r % ne1); r /= ne1;
        const int64_t i2 = (int64_t) (r % ne2), i3 = (int64_t) (r / ne2);
        const T * src = (const T *) (a.p + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
        T *       dst = (T *) (d.p + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]);
        const int64_t i0 = 2 * pr;                      // index into the (cos, sin) sequence
        if (i0 >= R.n_dims) {                           // channels past the rotated ones are copied
            dst[i0] = src[i0]; dst[i0 + 1] = src[i0 + 1];
            continue;
        }
        const float ff = R.freq_factors ? R.freq_factors[pr] : 1.0f;
        const float theta_extrap = (float) R.pos[i2] * powf(R.theta_scale, (float) pr) / ff;
        const float theta_interp = R.freq_scale * theta_extrap;
        float theta = theta_interp, mscale = R.attn_factor;
        if (R.ext_factor != 0.0f) {
            const float y = ((float) pr - R.corr0) / fmaxf(0.001f, R.corr1 - R.corr0);
            const float ramp_mix = (1.0f - fminf(1.0f, fmaxf(0.0f, y))) * R.ext_factor;
            theta = theta_interp * (1.0f - ramp_mix) + theta_extrap * ramp_mix;
            mscale *= 1.0f + 0.1f * logf(1.0f / R.freq_scale);
        }
        const float c = cosf(theta) * mscale, s = sinf(theta) * mscale;
        const int64_t ia = R.mode == 0 ? i0 : pr, ib = R.mode == 0 ? i0 + 1 : pr + R.n_dims / 2;
        const float x0 = to_f32<T>(src[ia]), x1 = to_f32<T>(src[ib]);
        dst[ia] = from_f32<T>(x0 * c - x1 * s);
        dst[ib] = from_f32<T>(x0 * s + x1 * c);
    }
}

// ---- SOFT_MAX: softmax(x * scale + slope * mask) with optional sinks, one CTA per row -----------------------------------------
struct softmax_args {
    float scale, max_bias, m0, m1;
    uint32_t n_head_log2;
    const char * mask; int mask_f16; int64_t mnb1, mnb2, mnb3, mne2, mne3;
    const float * sinks;
};

__global__ void __launch_bounds__(OPS_THREADS) soft_max_kernel(const tview a, const tview d, const softmax_args S) {
    pdl_enter();
    __shared__ float red[OPS_THREADS / 32];
    const int64_t r = blockIdx.x;
    const int64_t i1 = r % a.ne[1], i2 = (r / a.ne[1]) % a.ne[2], i3 = r / (a.ne[1] * a.ne[2]);
    const float * x = (const float *) (a.p + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
    float *       y = (float *) (d.p + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]);
    const int64_t n = a.ne[0];
    float slope = 1.0f;
    if (S.max_bias > 0.0f) {
        const uint32_t h = (uint32_t) i2;
        slope = h < S.n_head_log2 ? powf(S.m0, (float) (h + 1)) : powf(S.m1, (float) (2 * (h - S.n_head_log2) + 1));
    }
    const char * mrow = S.mask ? S.mask + i1 * S.mnb1 + (i2 % S.mne2) * S.mnb2 + (i3 % S.mne3) * S.mnb3 : nullptr;
    auto value = [&](const int64_t i) {
        float v = x[i] * S.scale;
        if (mrow) v += slope * (S.mask_f16 ? __half2float(((const __half *) mrow)[i]) : ((const float *) mrow)[i]);
        return v;
    };
    float mx = -INFINITY;
    for (int64_t i = threadIdx.x; i < n; i += OPS_THREADS) mx = fmaxf(mx, value(i));
    mx = block_max(mx, red);
    if (S.sinks) mx = fmaxf(mx, S.sinks[i2]);
    float sum = 0.0f;
    for (int64_t i = threadIdx.x; i < n; i += OPS_THREADS) {
        const float e = expf(value(i) - mx);        // (-inf - finite max -> 0, as in the reference)
        y[i] = e;
        sum += e;
    }
    sum = block_sum(sum, red);
    if (S.sinks) sum += expf(S.sinks[i2] - mx);
    const float inv = 1.0f / sum;
    for (int64_t i = threadIdx.x; i < n; i += OPS_THREADS) y[i] *= inv;
}

// rows of up to 32 * NPL values (the attention scores of a prompt: 16 384 rows of 512): one warp per row, the row stays in registers
// between the three steps — one read, one write, no block-wide synchronisation.  Same arithmetic as soft_max_kernel.
constexpr int SMW_ROWS = OPS_THREADS / 32;
template <int NPL>
__global__ void __launch_bounds__(OPS_THREADS) soft_max_warp_kernel(const tview a, const tview d, const softmax_args S, const int64_t rows) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t) blockIdx.x * SMW_ROWS + (threadIdx.x >> 5);
    if (r >= rows) return;                                      // (whole warps leave; nothing block-wide follows)
    const int64_t i1 = r % a.ne[1], i2 = (r / a.ne[1]) % a.ne[2], i3 = r / (a.ne[1] * a.ne[2]);
    const float * x = (const float *) (a.p + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
    float *       y = (float *) (d.p + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]);
    const int n = (int) a.ne[0];
    float slope = 1.0f;
    if (S.max_bias > 0.0f) {
        const uint32_t h = (uint32_t) i2;
        slope = h < S.n_head_log2 ? powf(S.m0, (float) (h + 1)) : powf(S.m1, (float) (2 * (h - S.n_head_log2) + 1));
    }
    const char * mrow = S.mask ? S.mask + i1 * S.mnb1 + (i2 % S.mne2) * S.mnb2 + (i3 % S.mne3) * S.mnb3 : nullptr;
    float v[NPL];
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < NPL; ++j) {
        const int i = lane + 32 * j;
        float t = -INFINITY;
        if (i < n) {
            t = x[i] * S.scale;
            if (mrow) t += slope * (S.mask_f16 ? __half2float(((const __half *) mrow)[i]) : ((const float *) mrow)[i]);
        }
        v[j] = t;
        mx = fmaxf(mx, t);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (S.sinks) mx = fmaxf(mx, S.sinks[i2]);
    float sum = 0.0f;
#pragma unroll
    for (int j = 0; j < NPL; ++j) {
        const float e = lane + 32 * j < n ? expf(v[j] - mx) : 0.0f;
        v[j] = e;
        sum += e;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (S.sinks) sum += expf(S.sinks[i2] - mx);
    const float inv = 1.0f / sum;
#pragma unroll
    for (int j = 0; j < NPL; ++j) if (lane + 32 * j < n) y[lane + 32 * j] = v[j] * inv;
}

}  // namespace

extern "C" {

int b200_op_rms_norm(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * d, float eps) { return b200_op_rms_norm_mul(ctx, a, nullptr, d, eps); }

int b200_op_rms_norm_mul(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * w, const b200_tensor * d, float eps) {
    if (!ctx || !a || !d) { b200_set_error("b200_op_rms_norm: null argument"); return B200_E_ARG; }
    if (w) {
        if (w->type != B200_TYPE_F32) { b200_set_error("b200_op_rms_norm_mul: f32 weight"); return B200_E_UNSUPPORTED; }
        for (int i = 0; i < 4; ++i) if (w->ne[i] <= 0 || a->ne[i] % w->ne[i] != 0) { b200_set_error("b200_op_rms_norm_mul: the weight cannot be repeated into src0"); return B200_E_ARG; }
    }
    if (a->type != B200_TYPE_F32 || d->type != B200_TYPE_F32 || a->nb[0] != 4 || d->nb[0] != 4 || !same_shape(a, d)) {
        b200_set_error("b200_op_rms_norm: f32 rows with nb[0] == 4"); return B200_E_UNSUPPORTED;
    }
    const int64_t rows = nrows_of(a);
    if (rows == 0 || a->ne[0] == 0) return B200_OK;
    if (rows >= ((int64_t) 1 << 31)) { b200_set_error("b200_op_rms_norm: too many rows"); return B200_E_UNSUPPORTED; }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    tview vw = {};
    if (w) vw = view_of(w);
    auto al16 = [](const b200_tensor * t) { return (((uintptr_t) t->data | (uintptr_t) t->nb[1] | (uintptr_t) t->nb[2] | (uintptr_t) t->nb[3]) & 15) == 0; };
    const bool vec = a->ne[0] % 4 == 0 && a->ne[0] >= 1024 && al16(a) && al16(d) && (!w || (w->ne[0] == a->ne[0] && w->nb[0] == 4 && al16(w)));
    if (vec && rows >= 64) OPS_LAUNCH(ctx, (rms_norm_vec_kernel<256, 4>), (unsigned) rows, 256, view_of(a), vw, view_of(d), eps);
    else if (vec) OPS_LAUNCH(ctx, (rms_norm_vec_kernel<RMS_VEC_THREADS, 2>), (unsigned) rows, RMS_VEC_THREADS, view_of(a), vw, view_of(d), eps);
    else     OPS_LAUNCH(ctx, rms_norm_kernel, (unsigned) rows, OPS_THREADS, view_of(a), vw, view_of(d), eps);
    return finish(ctx);
}

/* op_params: the 15 int32 words of a GGML_OP_ROPE node ([1] n_dims, [2] mode, [4] n_ctx_orig, [5..10] freq_base, freq_scale, ext_factor,
 * attn_factor, beta_fast, beta_slow as f32 bits) */
int b200_op_rope(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * pos, const b200_tensor * freq_factors, const b200_tensor * d,
                 const int32_t * op_params) {
    if (!ctx || !a || !pos || !d || !op_params) { b200_set_error("b200_op_rope: null argument"); return B200_E_ARG; }
    const int n_dims = op_params[1], mode = op_params[2], n_ctx_orig = op_params[4];
    float freq_base, freq_scale, ext_factor, attn_factor, beta_fast, beta_slow;
    memcpy(&freq_base, op_params + 5, 4); memcpy(&freq_scale, op_params + 6, 4); memcpy(&ext_factor, op_params + 7, 4);
    memcpy(&attn_factor, op_params + 8, 4); memcpy(&beta_fast, op_params + 9, 4); memcpy(&beta_slow, op_params + 10, 4);
    const size_t esz = a->type == B200_TYPE_F16 ? 2 : 4;
    if ((a->type != B200_TYPE_F32 && a->type != B200_TYPE_F16) || d->type != a->type || pos->type != B200_TYPE_I32 || !same_shape(a, d) ||
        a->nb[0] != (int64_t) esz || d->nb[0] != (int64_t) esz || (mode != 0 && mode != 2) || n_dims <= 0 || (n_dims & 1) || n_dims > a->ne[0] || (a->ne[0] & 1) ||
        pos->ne[0] < a->ne[2] || (freq_factors && (freq_factors->type != B200_TYPE_F32 || freq_factors->ne[0] < n_dims / 2))) {
        b200_set_error("b200_op_rope: f32 / f16 rows, i32 positions, mode normal (0) or NeoX (2)"); return B200_E_UNSUPPORTED;
    }
    rope_args R;
    R.n_dims = n_dims; R.mode = mode;
    R.theta_scale = powf(freq_base, -2.0f / (float) n_dims);
    R.freq_scale = freq_scale; R.ext_factor = ext_factor; R.attn_factor = attn_factor;
    // ggml_rope_yarn_corr_dims (ggml.c:4329-4343)
    auto corr_dim = [&](float n_rot) { return (float) n_dims * logf((float) n_ctx_orig / (n_rot * 2.0f * (float) M_PI)) / (2.0f * logf(freq_base)); };
    const float start = floorf(corr_dim(beta_fast)), end = ceilf(corr_dim(beta_slow));
    R.corr0 = start > 0.0f ? start : 0.0f;
    R.corr1 = end < (float) (n_dims - 1) ? end : (float) (n_dims - 1);
    R.pos = (const int32_t *) pos->data;
    R.freq_factors = freq_factors ? (const float *) freq_factors->data : nullptr;
    const int64_t n = nelem(a) / 2;
    if (n == 0) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    const int g = grid_for(n, ctx->sm_count);
    const bool small = n < ((int64_t) 1 << 31);
    if (a->type == B200_TYPE_F32) {
        if (small) OPS_LAUNCH(ctx, (rope_kernel<float, uint32_t>), g, OPS_THREADS, view_of(a), view_of(d), R, n);
        else       OPS_LAUNCH(ctx, (rope_kernel<float, int64_t>), g, OPS_THREADS, view_of(a), view_of(d), R, n);
    } else {
        if (small) OPS_LAUNCH(ctx, (rope_kernel<__half, uint32_t>), g, OPS_THREADS, view_of(a), view_of(d), R, n);
        else       OPS_LAUNCH(ctx, (rope_kernel<__half, int64_t>), g, OPS_THREADS, view_of(a), view_of(d), R, n);
    }
    return finish(ctx);
}

int b200_op_soft_max(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * mask, const b200_tensor * sinks, const b200_tensor * d,
                     float scale, float max_bias) {
    if (!ctx || !a || !d) { b200_set_error("b200_op_soft_max: null argument"); return B200_E_ARG; }
    if (a->type != B200_TYPE_F32 || d->type != B200_TYPE_F32 || a->nb[0] != 4 || !contiguous(d, 4) || !same_shape(a, d) ||
        (mask && ((mask->type != B200_TYPE_F32 && mask->type != B200_TYPE_F16) || mask->ne[0] != a->ne[0] || mask->ne[1] < a->ne[1] ||
                  mask->nb[0] != (mask->type == B200_TYPE_F16 ? 2 : 4))) ||
        (sinks && (sinks->type != B200_TYPE_F32 || sinks->ne[0] < a->ne[2]))) {
        b200_set_error("b200_op_soft_max: f32 rows, f16 / f32 mask rows of the same length"); return B200_E_UNSUPPORTED;
    }
    const int64_t rows = nrows_of(a);
    if (rows == 0 || a->ne[0] == 0) return B200_OK;
    if (rows >= ((int64_t) 1 << 31)) { b200_set_error("b200_op_soft_max: too many rows"); return B200_E_UNSUPPORTED; }
    softmax_args S = {};
    S.scale = scale; S.max_bias = max_bias;
    const uint32_t n_head = (uint32_t) a->ne[2];
    S.n_head_log2 = 1u << (uint32_t) floor(log2((double) (n_head > 0 ? n_head : 1)));
    S.m0 = powf(2.0f, -(max_bias) / (float) S.n_head_log2);
    S.m1 = powf(2.0f, -(max_bias / 2.0f) / (float) S.n_head_log2);
    if (mask) {
        S.mask = (const char *) mask->data; S.mask_f16 = mask->type == B200_TYPE_F16;
        S.mnb1 = mask->nb[1]; S.mnb2 = mask->nb[2]; S.mnb3 = mask->nb[3]; S.mne2 = mask->ne[2]; S.mne3 = mask->ne[3];
    }
    S.sinks = sinks ? (const float *) sinks->data : nullptr;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    if (rows >= 64 && a->ne[0] <= 32 * 32) {                   // many short rows: a warp per row
        const unsigned g = (unsigned) ((rows + SMW_ROWS - 1) / SMW_ROWS);
        if (a->ne[0] <= 32 * 8) OPS_LAUNCH(ctx, soft_max_warp_kernel<8>, g, OPS_THREADS, view_of(a), view_of(d), S, rows);
        else                    OPS_LAUNCH(ctx, soft_max_warp_kernel<32>, g, OPS_THREADS, view_of(a), view_of(d), S, rows);
        return finish(ctx);
    }
    OPS_LAUNCH(ctx, soft_max_kernel, (unsigned) rows, OPS_THREADS, view_of(a), view_of(d), S);
    return finish(ctx);
}

}  // extern "C"
