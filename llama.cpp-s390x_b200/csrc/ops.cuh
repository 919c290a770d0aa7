// ops.cuh — helpers shared by the element-wise / row ops (ops.cu, ops_rows.cu)
#pragma once

#include "common.cuh"

namespace {

struct tview {              // a tensor as a kernel sees it
    char *  p;
    int64_t ne[4];
    int64_t nb[4];
};

inline tview view_of(const b200_tensor * t) {
    tview v;
    v.p = (char *) t->data;
    for (int i = 0; i < 4; ++i) { v.ne[i] = t->ne[i]; v.nb[i] = t->nb[i]; }
    return v;
}

inline int64_t nelem(const b200_tensor * t) { return t->ne[0] * t->ne[1] * t->ne[2] * t->ne[3]; }
inline int64_t nrows_of(const b200_tensor * t) { return t->ne[1] * t->ne[2] * t->ne[3]; }

inline bool same_shape(const b200_tensor * a, const b200_tensor * b) {
    for (int i = 0; i < 4; ++i) if (a->ne[i] != b->ne[i]) return false;
    return true;
}

// dims >= 1 packed behind rows of nb[1] bytes (ggml_is_contiguous_1), elements of a row packed
inline bool rows_contiguous(const b200_tensor * t, size_t esz) {
    return t->nb[0] == (int64_t) esz && t->nb[2] == t->nb[1] * t->ne[1] && t->nb[3] == t->nb[2] * t->ne[2];
}
inline bool contiguous(const b200_tensor * t, size_t esz) { return rows_contiguous(t, esz) && t->nb[1] == (int64_t) esz * t->ne[0]; }

constexpr int OPS_THREADS = 256;

inline int grid_for(int64_t n, int sm_count) {
    const int64_t blocks = (n + OPS_THREADS - 1) / OPS_THREADS;
    const int64_t cap = (int64_t) sm_count * 32;           // grid-stride beyond 32 CTAs per SM
    return (int) (blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

// Every kernel of these files starts with pdl_enter() and is launched with the programmatic-serialization attribute (ctx->pdl and
// ctx->ops_pdl): it may become resident while its predecessor in the stream is still running, waits there until that grid has completed
// and its writes are visible (griddepcontrol.wait), and at once lets ITS successor do the same — the launch latency of a chain of
// small kernels overlaps their execution, and a GEMV that follows starts streaming its weights before the small kernels are done
// (the GEMVs fetch weights ahead of their own wait, mmvq.cu).  Every kernel waits before it touches memory, so "my predecessor is
// complete" implies "everything before it is complete".  Without the attribute both instructions are no-ops.
__device__ __forceinline__ void pdl_enter() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

template <typename... KA, typename... A>
inline cudaError_t ops_launch(b200_ctx * ctx, void (*kern)(KA...), const dim3 grid, const dim3 block, A... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = 0; cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = ctx->pdl && ctx->ops_pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KA>(args)...);
}
#define OPS_LAUNCH(ctx, kern, grid, block, ...) do { B200_CUDA_TRY(ops_launch(ctx, kern, dim3(grid), dim3(block), __VA_ARGS__)); } while (0)

inline int finish(b200_ctx * ctx) {
    B200_CUDA_TRY(cudaGetLastError());
    ctx->launches++;
    ctx->n_pend = 16;          // written by a kernel the GEMV write tracker does not know: the next GEMV waits before it reads
    return B200_OK;
}

__device__ __forceinline__ void split4(int64_t i, const int64_t (&ne)[4], int64_t & i0, int64_t & i1, int64_t & i2, int64_t & i3) {
    if (((((uint64_t) i) | (uint64_t) ne[0] | (uint64_t) ne[1] | (uint64_t) ne[2]) >> 32) == 0) {      // (the usual case: 32-bit divisions)
        uint32_t r = (uint32_t) i;
        const uint32_t n0 = (uint32_t) ne[0], n1 = (uint32_t) ne[1], n2 = (uint32_t) ne[2];
        i0 = r % n0; r /= n0;
        i1 = r % n1; r /= n1;
        i2 = r % n2;
        i3 = r / n2;
        return;
    }
    i0 = i % ne[0]; i /= ne[0];
    i1 = i % ne[1]; i /= ne[1];
    i2 = i % ne[2];
    i3 = i / ne[2];
}

// ---- activations ----------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float act_silu(float x)       { return x / (1.0f + expf(-x)); }
__device__ __forceinline__ float act_gelu(float x)       { return 0.5f * x * (1.0f + tanhf(0.79788456080286535587989211986876f * x * (1.0f + 0.044715f * x * x))); }
__device__ __forceinline__ float act_gelu_quick(float x) { return x * (1.0f / (1.0f + expf(-1.702f * x))); }
__device__ __forceinline__ float act_gelu_erf(float x)   { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }

__device__ __forceinline__ float glu_value(const int op, const float x, const float g, const float alpha, const float limit) {
    switch (op) {
        case B200_GLU_REGLU:       return (x > 0.0f ? x : 0.0f) * g;
        case B200_GLU_GEGLU:       return act_gelu(x) * g;
        case B200_GLU_SWIGLU:      return act_silu(x) * g;
        case B200_GLU_GEGLU_ERF:   return act_gelu_erf(x) * g;
        case B200_GLU_GEGLU_QUICK: return act_gelu_quick(x) * g;
        default: {                 // SWIGLU_OAI (ops.cpp:3325-3330)
            const float xc = fminf(x, limit), gc = fminf(fmaxf(g, -limit), limit);
            return xc / (1.0f + expf(alpha * (-xc))) * (gc + 1.0f);
        }
    }
}

}  // namespace
