// ops.cu — the other ops of a Llama graph, so that a cgraph stays on the device between its MUL_MATs:
// element-wise binary ops with ggml broadcasting, ADD_ID, GLU, unary activations, SCALE, RMS_NORM, ROPE, SOFT_MAX, CPY / CONT / DUP,
// GET_ROWS, SET_ROWS and a plain float MUL_MAT (F16 / F32 weights: the K.q and V.p products of attention).  All of them are
// HBM-bound byte movers: one thread per element (or one warp / CTA per row where a row is reduced), coalesced along ne0, byte
// strides taken from the tensors, so views and permuted tensors need no copy.
//
// Each entry point restates the semantics of the reference's CPU forward function (file:line under /root/reference/ggml/src):
//   b200_op_binary    ggml_compute_forward_{add,sub,mul,div}      ggml-cpu/binary-ops.cpp (broadcast: ggml.c ggml_can_repeat)
//   b200_op_repeat    ggml_compute_forward_repeat_f32              ggml-cpu/ops.cpp
//   b200_op_add_id    ggml_compute_forward_add_id_f32             ggml-cpu/ops.cpp:703-752
//   b200_op_glu       ggml_compute_forward_{reglu,geglu,swiglu,swiglu_oai,geglu_erf,geglu_quick}   ggml-cpu/ops.cpp:3133-3345
//   b200_op_unary     ggml_compute_forward_unary                   ggml-cpu/unary-ops.cpp, formulas ggml-cpu/vec.h:981-1070
//   b200_op_scale     ggml_compute_forward_scale_f32               ggml-cpu/ops.cpp
//   b200_op_rms_norm  ggml_compute_forward_rms_norm_f32            ggml-cpu/ops.cpp
//   b200_op_rope      ggml_compute_forward_rope_flt                ggml-cpu/ops.cpp:5647-5925 (+ ggml_rope_yarn_corr_dims ggml.c:4329-4343)
//   b200_op_soft_max  ggml_compute_forward_soft_max_f32            ggml-cpu/ops.cpp:5283-5393
//   b200_op_cpy       ggml_compute_forward_dup                     ggml-cpu/ops.cpp
//   b200_op_get_rows  ggml_compute_forward_get_rows                ggml-cpu/ops.cpp:4820-4940 (dequantize_row_* ggml-quants.c)
//   b200_op_set_rows  ggml_compute_forward_set_rows_f32            ggml-cpu/ops.cpp:4950-4998
//   b200_mul_mat_f    ggml_compute_forward_mul_mat (float src0)    ggml-cpu/ggml-cpu.c:1241-1439
// No CPU fallback: every function launches a kernel on the context's stream or fails.
#include <cuda_bf16.h>
#include <math.h>

#include "ops.cuh"

namespace {

// ---- binary ops with broadcast ------------------------------------------------------------------------------------------
template <int OP>
__global__ void __launch_bounds__(OPS_THREADS) bin_kernel(const tview a, const tview b, const tview d, const int64_t n) {
    pdl_enter();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t i0, i1, i2, i3;
        split4(i, d.ne, i0, i1, i2, i3);
        const float x = *(const float *) (a.p + i0 * a.nb[0] + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
        const float y = *(const float *) (b.p + (i0 % b.ne[0]) * b.nb[0] + (i1 % b.ne[1]) * b.nb[1] + (i2 % b.ne[2]) * b.nb[2] + (i3 % b.ne[3]) * b.nb[3]);
        float r;
        if (OP == B200_OP_ADD) r = x + y; else if (OP == B200_OP_SUB) r = x - y; else if (OP == B200_OP_MUL) r = x * y; else r = x / y;
        *(float *) (d.p + i0 * d.nb[0] + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]) = r;
    }
}

// same-shape, contiguous, 16-byte aligned operands (the residual adds of a prompt: 8 MB each): four elements per thread, no index arithmetic
template <int OP>
__global__ void __launch_bounds__(OPS_THREADS) bin_vec_kernel(const float4 * a, const float4 * b, float4 * d, const int64_t n4) {          // (d may be a or b: in-place adds)
    pdl_enter();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t) gridDim.x * blockDim.x) {
        const float4 x = a[i], y = b[i];
        float4 r;
        if (OP == B200_OP_ADD)      r = make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
        else if (OP == B200_OP_SUB) r = make_float4(x.x - y.x, x.y - y.y, x.z - y.z, x.w - y.w);
        else if (OP == B200_OP_MUL) r = make_float4(x.x * y.x, x.y * y.y, x.z * y.z, x.w * y.w);
        else                        r = make_float4(x.x / y.x, x.y / y.y, x.z / y.z, x.w / y.w);
        d[i] = r;
    }
}

// ---- REPEAT: dst = src0 tiled -------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(OPS_THREADS) repeat_kernel(const tview a, const tview d, const int64_t n) {
    pdl_enter();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t i0, i1, i2, i3;
        split4(i, d.ne, i0, i1, i2, i3);
        *(float *) (d.p + i0 * d.nb[0] + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]) =
            *(const float *) (a.p + (i0 % a.ne[0]) * a.nb[0] + (i1 % a.ne[1]) * a.nb[1] + (i2 % a.ne[2]) * a.nb[2] + (i3 % a.ne[3]) * a.nb[3]);
    }
}

// ---- ADD_ID: dst[:, i1, i2] = a[:, i1, i2] + b[:, ids[i1, i2]] -----------------------------------------------------------
__global__ void __launch_bounds__(OPS_THREADS) add_id_kernel(const tview a, const tview b, const tview ids, const tview d, const int64_t n) {
    pdl_enter();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t i0, i1, i2, i3;
        split4(i, d.ne, i0, i1, i2, i3);
        const int32_t e = *(const int32_t *) (ids.p + i1 * ids.nb[0] + i2 * ids.nb[1]);
        float r = *(const float *) (a.p + i0 * 4 + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
        if (e >= 0 && e < b.ne[1]) r += *(const float *) (b.p + i0 * 4 + (int64_t) e * b.nb[1]);
        *(float *) (d.p + i0 * 4 + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]) = r;
    }
}

// rows: x row r at xp + r * xs, gate row at gp + r * gs, nc values per row
__global__ void __launch_bounds__(OPS_THREADS) glu_kernel(const char * xp, const int64_t xs, const char * gp, const int64_t gs, char * dp, const int64_t ds,
                                                          const int64_t nc, const int64_t n, const int op, const float alpha, const float limit) {
    pdl_enter();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const int64_t r = i / nc, c = i - r * nc;
        const float x = ((const float *) (xp + r * xs))[c], g = ((const float *) (gp + r * gs))[c];
        ((float *) (dp + r * ds))[c] = glu_value(op, x, g, alpha, limit);
    }
}

// the same with four values per thread (rows a multiple of 4 long, 16-byte aligned)
__global__ void __launch_bounds__(OPS_THREADS) glu_vec_kernel(const char * xp, const int64_t xs, const char * gp, const int64_t gs, char * dp, const int64_t ds,
                                                              const int64_t nc4, const int64_t n4, const int op, const float alpha, const float limit) {
    pdl_enter();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t) gridDim.x * blockDim.x) {
        const int64_t r = i / nc4, c = i - r * nc4;
        const float4 x = ((const float4 *) (xp + r * xs))[c], g = ((const float4 *) (gp + r * gs))[c];
        ((float4 *) (dp + r * ds))[c] = make_float4(glu_value(op, x.x, g.x, alpha, limit), glu_value(op, x.y, g.y, alpha, limit),
                                                    glu_value(op, x.z, g.z, alpha, limit), glu_value(op, x.w, g.w, alpha, limit));
    }
}

__global__ void __launch_bounds__(OPS_THREADS) unary_kernel(const float * __restrict__ x, float * __restrict__ y, const int64_t n, const int op) {
    pdl_enter();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const float v = x[i];
        float r;
        switch (op) {
            case B200_UNARY_ABS:        r = fabsf(v); break;
            case B200_UNARY_SGN:        r = v > 0.0f ? 1.0f : (v < 0.0f ? -1.0f : 0.0f); break;
            case B200_UNARY_NEG:        r = -v; break;
            case B200_UNARY_STEP:       r = v > 0.0f ? 1.0f : 0.0f; break;
            case B200_UNARY_TANH:       r = tanhf(v); break;
            case B200_UNARY_RELU:       r = v > 0.0f ? v : 0.0f; break;
            case B200_UNARY_SIGMOID:    r = 1.0f / (1.0f + expf(-v)); break;
            case B200_UNARY_GELU:       r = act_gelu(v); break;
            case B200_UNARY_GELU_QUICK: r = act_gelu_quick(v); break;
            case B200_UNARY_SILU:       r = act_silu(v); break;
            case B200_UNARY_EXP:        r = expf(v); break;
            default:                    r = act_gelu_erf(v); break;       // B200_UNARY_GELU_ERF
        }
        y[i] = r;
    }
}

__global__ void __launch_bounds__(OPS_THREADS) scale_kernel(const float * __restrict__ x, float * __restrict__ y, const int64_t n, const float s, const float b) {
    pdl_enter();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) y[i] = x[i] * s + b;
}

}  // namespace

extern "C" {

int b200_op_binary(b200_ctx * ctx, int op, const b200_tensor * a, const b200_tensor * b, const b200_tensor * d) {
    if (!ctx || !a || !b || !d) { b200_set_error("b200_op_binary: null argument"); return B200_E_ARG; }
    if (a->type != B200_TYPE_F32 || b->type != B200_TYPE_F32 || d->type != B200_TYPE_F32 || !same_shape(a, d)) {
        b200_set_error("b200_op_binary: f32 tensors, dst shaped like src0"); return B200_E_UNSUPPORTED;
    }
    for (int i = 0; i < 4; ++i) if (b->ne[i] <= 0 || a->ne[i] % b->ne[i] != 0) { b200_set_error("b200_op_binary: src1 cannot be repeated into src0"); return B200_E_ARG; }
    const int64_t n = nelem(d);
    if (n == 0) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    if (op >= B200_OP_ADD && op <= B200_OP_DIV && same_shape(a, b) && contiguous(a, 4) && contiguous(b, 4) && contiguous(d, 4) && n % 4 == 0 &&
        (((uintptr_t) a->data | (uintptr_t) b->data | (uintptr_t) d->data) & 15) == 0) {
        const float4 * pa = (const float4 *) a->data, * pb = (const float4 *) b->data;
        float4 * pd = (float4 *) d->data;
        const int gv = grid_for(n / 4, ctx->sm_count);
        switch (op) {
            case B200_OP_ADD: OPS_LAUNCH(ctx, bin_vec_kernel<B200_OP_ADD>, gv, OPS_THREADS, pa, pb, pd, n / 4); break;
            case B200_OP_SUB: OPS_LAUNCH(ctx, bin_vec_kernel<B200_OP_SUB>, gv, OPS_THREADS, pa, pb, pd, n / 4); break;
            case B200_OP_MUL: OPS_LAUNCH(ctx, bin_vec_kernel<B200_OP_MUL>, gv, OPS_THREADS, pa, pb, pd, n / 4); break;
            default:          OPS_LAUNCH(ctx, bin_vec_kernel<B200_OP_DIV>, gv, OPS_THREADS, pa, pb, pd, n / 4); break;
        }
        return finish(ctx);
    }
    const int g = grid_for(n, ctx->sm_count);
    const tview va = view_of(a), vb = view_of(b), vd = view_of(d);
    switch (op) {
        case B200_OP_ADD: OPS_LAUNCH(ctx, bin_kernel<B200_OP_ADD>, g, OPS_THREADS, va, vb, vd, n); break;
        case B200_OP_SUB: OPS_LAUNCH(ctx, bin_kernel<B200_OP_SUB>, g, OPS_THREADS, va, vb, vd, n); break;
        case B200_OP_MUL: OPS_LAUNCH(ctx, bin_kernel<B200_OP_MUL>, g, OPS_THREADS, va, vb, vd, n); break;
        case B200_OP_DIV: OPS_LAUNCH(ctx, bin_kernel<B200_OP_DIV>, g, OPS_THREADS, va, vb, vd, n); break;
        default: b200_set_error("b200_op_binary: unknown op %d", op); return B200_E_ARG;
    }
    return finish(ctx);
}

int b200_op_repeat(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * d) {
    if (!ctx || !a || !d) { b200_set_error("b200_op_repeat: null argument"); return B200_E_ARG; }
    if (a->type != B200_TYPE_F32 || d->type != B200_TYPE_F32) { b200_set_error("b200_op_repeat: f32 tensors"); return B200_E_UNSUPPORTED; }
    for (int i = 0; i < 4; ++i) if (a->ne[i] <= 0 || d->ne[i] % a->ne[i] != 0) { b200_set_error("b200_op_repeat: src0 cannot be repeated into dst"); return B200_E_ARG; }
    const int64_t n = nelem(d);
    if (n == 0) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    OPS_LAUNCH(ctx, repeat_kernel, grid_for(n, ctx->sm_count), OPS_THREADS, view_of(a), view_of(d), n);
    return finish(ctx);
}

int b200_op_add_id(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * b, const b200_tensor * ids, const b200_tensor * d) {
    if (!ctx || !a || !b || !ids || !d) { b200_set_error("b200_op_add_id: null argument"); return B200_E_ARG; }
    if (a->type != B200_TYPE_F32 || b->type != B200_TYPE_F32 || d->type != B200_TYPE_F32 || ids->type != B200_TYPE_I32 || !same_shape(a, d) ||
        a->nb[0] != 4 || b->nb[0] != 4 || d->nb[0] != 4 || b->ne[0] != a->ne[0]) {
        b200_set_error("b200_op_add_id: f32 rows with nb[0] == 4, i32 ids"); return B200_E_UNSUPPORTED;
    }
    const int64_t n = nelem(d);
    if (n == 0) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    OPS_LAUNCH(ctx, add_id_kernel, grid_for(n, ctx->sm_count), OPS_THREADS, view_of(a), view_of(b), view_of(ids), view_of(d), n);
    return finish(ctx);
}

int b200_op_glu(b200_ctx * ctx, int glu_op, const b200_tensor * a, const b200_tensor * b, const b200_tensor * d, int swapped, float alpha, float limit) {
    if (!ctx || !a || !d) { b200_set_error("b200_op_glu: null argument"); return B200_E_ARG; }
    if (glu_op < 0 || glu_op > B200_GLU_GEGLU_QUICK) { b200_set_error("b200_op_glu: unknown op %d", glu_op); return B200_E_ARG; }
    if (a->type != B200_TYPE_F32 || d->type != B200_TYPE_F32 || (b && b->type != B200_TYPE_F32) || !rows_contiguous(a, 4) || !rows_contiguous(d, 4) ||
        (b && (!rows_contiguous(b, 4) || !same_shape(a, b)))) {
        b200_set_error("b200_op_glu: f32 tensors whose rows are contiguous"); return B200_E_UNSUPPORTED;
    }
    const int64_t nc = b ? a->ne[0] : a->ne[0] / 2, nr = nrows_of(a);
    if (d->ne[0] != nc || nrows_of(d) != nr) { b200_set_error("b200_op_glu: dst shape"); return B200_E_ARG; }
    const int64_t n = nc * nr;
    if (n == 0) return B200_OK;
    const char * xp = (const char *) a->data, * gp = b ? (const char *) b->data : (const char *) a->data;
    if (!b) { xp += swapped ? nc * 4 : 0; gp += swapped ? 0 : nc * 4; }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    const int64_t gs = b ? b->nb[1] : a->nb[1];
    if (nc % 4 == 0 && (((uintptr_t) xp | (uintptr_t) gp | (uintptr_t) d->data | (uintptr_t) a->nb[1] | (uintptr_t) gs | (uintptr_t) d->nb[1]) & 15) == 0) {
        OPS_LAUNCH(ctx, glu_vec_kernel, grid_for(n / 4, ctx->sm_count), OPS_THREADS, xp, a->nb[1], gp, gs, (char *) d->data, d->nb[1], nc / 4, n / 4, glu_op, alpha, limit);
        return finish(ctx);
    }
    OPS_LAUNCH(ctx, glu_kernel, grid_for(n, ctx->sm_count), OPS_THREADS, xp, a->nb[1], gp, gs, (char *) d->data, d->nb[1], nc, n, glu_op, alpha, limit);
    return finish(ctx);
}

int b200_op_unary(b200_ctx * ctx, int unary_op, const b200_tensor * a, const b200_tensor * d) {
    if (!ctx || !a || !d) { b200_set_error("b200_op_unary: null argument"); return B200_E_ARG; }
    if (!b200_op_unary_supported(unary_op)) { b200_set_error("b200_op_unary: op %d not implemented", unary_op); return B200_E_UNSUPPORTED; }
    if (a->type != B200_TYPE_F32 || d->type != B200_TYPE_F32 || !contiguous(a, 4) || !contiguous(d, 4) || nelem(a) != nelem(d)) {
        b200_set_error("b200_op_unary: contiguous f32 tensors"); return B200_E_UNSUPPORTED;
    }
    const int64_t n = nelem(d);
    if (n == 0) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    OPS_LAUNCH(ctx, unary_kernel, grid_for(n, ctx->sm_count), OPS_THREADS, (const float *) a->data, (float *) d->data, n, unary_op);
    return finish(ctx);
}

int b200_op_unary_supported(int unary_op) {
    switch (unary_op) {
        case B200_UNARY_ABS: case B200_UNARY_SGN: case B200_UNARY_NEG: case B200_UNARY_STEP: case B200_UNARY_TANH: case B200_UNARY_RELU:
        case B200_UNARY_SIGMOID: case B200_UNARY_GELU: case B200_UNARY_GELU_QUICK: case B200_UNARY_SILU: case B200_UNARY_EXP: case B200_UNARY_GELU_ERF:
            return 1;
        default:
            return 0;
    }
}

int b200_op_scale(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * d, float scale, float bias) {
    if (!ctx || !a || !d) { b200_set_error("b200_op_scale: null argument"); return B200_E_ARG; }
    if (a->type != B200_TYPE_F32 || d->type != B200_TYPE_F32 || !contiguous(a, 4) || !contiguous(d, 4) || nelem(a) != nelem(d)) {
        b200_set_error("b200_op_scale: contiguous f32 tensors"); return B200_E_UNSUPPORTED;
    }
    const int64_t n = nelem(d);
    if (n == 0) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    OPS_LAUNCH(ctx, scale_kernel, grid_for(n, ctx->sm_count), OPS_THREADS, (const float *) a->data, (float *) d->data, n, scale, bias);
    return finish(ctx);
}

}  // extern "C"
