// oracle/ref_shim.cpp — thin C wrapper over the UNMODIFIED reference (ggml-base + ggml-cpu compiled from
// /root/reference by oracle/Makefile) so that Python tests can call the reference's own functions
// through ctypes.  TEST INFRASTRUCTURE ONLY: built into oracle/_ref/libref_shim.so, never linked or
// loaded by the product path.  All arithmetic here happens inside the reference's code; this file
// only builds graphs and forwards pointers.
#include "ggml.h"
#include "ggml-alloc.h"
#include "ggml-backend.h"
#include "ggml-cpu.h"

#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

extern "C" {
// ggml-cpu/quants.h (generic C kernels and the arch-dispatched ones; exported by libggml-cpu)
void quantize_row_q8_1(const float * x, void * y, int64_t k);
void quantize_row_q8_0(const float * x, void * y, int64_t k);
void quantize_row_q8_K(const float * x, void * y, int64_t k);
void quantize_row_q8_K_ref(const float * x, void * y, int64_t k);   // ggml-quants.h (exported by libggml-base)
void ggml_vec_dot_q4_0_q8_0_generic(int n, float * s, size_t bs, const void * vx, size_t bx, const void * vy, size_t by, int nrc);
void ggml_vec_dot_q8_0_q8_0_generic(int n, float * s, size_t bs, const void * vx, size_t bx, const void * vy, size_t by, int nrc);
void ggml_vec_dot_q4_K_q8_K_generic(int n, float * s, size_t bs, const void * vx, size_t bx, const void * vy, size_t by, int nrc);
void ggml_vec_dot_q6_K_q8_K_generic(int n, float * s, size_t bs, const void * vx, size_t bx, const void * vy, size_t by, int nrc);
}

#define SHIM_API extern "C" __attribute__((visibility("default")))

SHIM_API int     ref_blck_size(int type)            { return (int) ggml_blck_size((ggml_type) type); }
SHIM_API size_t  ref_type_size(int type)            { return ggml_type_size((ggml_type) type); }
SHIM_API size_t  ref_row_size(int type, int64_t k)  { return ggml_row_size((ggml_type) type, k); }
SHIM_API const char * ref_type_name(int type)       { return ggml_type_name((ggml_type) type); }

SHIM_API float    ref_fp16_to_fp32(uint16_t h)      { return ggml_fp16_to_fp32(h); }
SHIM_API uint16_t ref_fp32_to_fp16(float f)         { return ggml_fp32_to_fp16(f); }

// type_traits.from_float_ref — e.g. quantize_row_q8_1_ref for Q8_1 (ggml/src/ggml.c:729-735)
SHIM_API int ref_quantize_row_ref(int type, const float * x, void * y, int64_t k) {
    if (type == GGML_TYPE_Q8_K) { quantize_row_q8_K_ref(x, y, k); return 0; }  // no from_float_ref in the traits table
    const ggml_type_traits * tt = ggml_get_type_traits((ggml_type) type);
    if (!tt || !tt->from_float_ref) return -1;
    tt->from_float_ref(x, y, k);
    return 0;
}

// the arch-dispatched CPU quantizers (x86 AVX path in this build: ggml-cpu/arch/x86/quants.c:400-502)
SHIM_API int ref_quantize_row_cpu(int type, const float * x, void * y, int64_t k) {
    switch ((ggml_type) type) {
        case GGML_TYPE_Q8_0: quantize_row_q8_0(x, y, k); return 0;
        case GGML_TYPE_Q8_1: quantize_row_q8_1(x, y, k); return 0;
        case GGML_TYPE_Q8_K: quantize_row_q8_K(x, y, k); return 0;
        default: return -1;
    }
}

// ggml_quantize_chunk (ggml.h:2764) — how test-backend-ops creates weights (tests/test-backend-ops.cpp:108-139)
SHIM_API size_t ref_quantize_chunk(int type, const float * src, void * dst, int64_t nrows, int64_t n_per_row) {
    return ggml_quantize_chunk((ggml_type) type, src, dst, 0, nrows, n_per_row, nullptr);
}

SHIM_API int ref_dequantize_row(int type, const void * x, float * y, int64_t k) {
    const ggml_type_traits * tt = ggml_get_type_traits((ggml_type) type);
    if (!tt || !tt->to_float) return -1;
    tt->to_float(x, y, k);
    return 0;
}

SHIM_API int ref_validate_row_data(int type, const void * data, size_t nbytes) {
    return ggml_validate_row_data((ggml_type) type, data, nbytes) ? 1 : 0;
}

// generic (scalar C) vec_dot of the reference; act is in the CPU vec_dot_type (Q8_0 / Q8_K)
SHIM_API float ref_vec_dot_generic(int wtype, int64_t k, const void * w, const void * act) {
    float s = 0.0f;
    ggml_cpu_init();   // fills the fp16->fp32 lookup table the CPU kernels read (ggml-cpu.c:3749-3775)
    switch ((ggml_type) wtype) {
        case GGML_TYPE_Q4_0: ggml_vec_dot_q4_0_q8_0_generic((int) k, &s, 0, w, 0, act, 0, 1); break;
        case GGML_TYPE_Q8_0: ggml_vec_dot_q8_0_q8_0_generic((int) k, &s, 0, w, 0, act, 0, 1); break;
        case GGML_TYPE_Q4_K: ggml_vec_dot_q4_K_q8_K_generic((int) k, &s, 0, w, 0, act, 0, 1); break;
        case GGML_TYPE_Q6_K: ggml_vec_dot_q6_K_q8_K_generic((int) k, &s, 0, w, 0, act, 0, 1); break;
        default: return NAN;
    }
    return s;
}

// ---- whole-op execution through the reference's CPU backend -------------------------------------------

namespace {
struct cpu_graph {
    ggml_context *        ctx     = nullptr;
    ggml_backend_t        backend = nullptr;
    ggml_backend_buffer_t buf     = nullptr;
    ggml_cgraph *         gf      = nullptr;
    ~cpu_graph() {
        if (buf)     ggml_backend_buffer_free(buf);
        if (backend) ggml_backend_free(backend);
        if (ctx)     ggml_free(ctx);
    }
    bool init(int n_threads) {
        ggml_init_params ip = { ggml_tensor_overhead() * 64 + ggml_graph_overhead_custom(8192, false), nullptr, true };
        ctx = ggml_init(ip);
        backend = ggml_backend_cpu_init();
        if (!ctx || !backend) return false;
        if (n_threads > 0) ggml_backend_cpu_set_n_threads(backend, n_threads);
        ggml_backend_cpu_set_use_ref(backend, true);   // as tests/test-backend-ops.cpp:9408-9414
        return true;
    }
    bool alloc() { buf = ggml_backend_alloc_ctx_tensors(ctx, backend); return buf != nullptr; }
};
}

// Y[m,n,ne12,ne13] = MUL_MAT(W[k,m,ne02,ne03] (wtype), X[k,n,ne12,ne13] f32) on the reference CPU backend
SHIM_API int ref_mul_mat(int wtype, const void * W, const float * X, float * Y,
                         int64_t m, int64_t n, int64_t k, int64_t ne02, int64_t ne03, int64_t ne12, int64_t ne13,
                         int n_threads) {
    cpu_graph g;
    if (!g.init(n_threads)) return -1;
    ggml_tensor * a = ggml_new_tensor_4d(g.ctx, (ggml_type) wtype, k, m, ne02, ne03);
    ggml_tensor * b = ggml_new_tensor_4d(g.ctx, GGML_TYPE_F32, k, n, ne12, ne13);
    ggml_tensor * c = ggml_mul_mat(g.ctx, a, b);
    g.gf = ggml_new_graph_custom(g.ctx, 16, false);
    ggml_build_forward_expand(g.gf, c);
    if (!g.alloc()) return -2;
    ggml_backend_tensor_set(a, W, 0, ggml_nbytes(a));
    ggml_backend_tensor_set(b, X, 0, ggml_nbytes(b));
    if (ggml_backend_graph_compute(g.backend, g.gf) != GGML_STATUS_SUCCESS) return -3;
    ggml_backend_tensor_get(c, Y, 0, ggml_nbytes(c));
    return 0;
}

// c[m,n_used,n_tokens] = MUL_MAT_ID(as[k,m,n_expert], b[k,b_ne1,n_tokens], ids[n_used,n_tokens]) (ggml.c:3279-3315)
SHIM_API int ref_mul_mat_id(int wtype, const void * as, const float * b, const int32_t * ids, float * c,
                            int64_t m, int64_t k, int64_t n_expert, int64_t n_used, int64_t n_tokens, int64_t b_ne1,
                            int n_threads) {
    cpu_graph g;
    if (!g.init(n_threads)) return -1;
    ggml_tensor * ta  = ggml_new_tensor_3d(g.ctx, (ggml_type) wtype, k, m, n_expert);
    ggml_tensor * tb  = ggml_new_tensor_3d(g.ctx, GGML_TYPE_F32, k, b_ne1, n_tokens);
    ggml_tensor * tid = ggml_new_tensor_2d(g.ctx, GGML_TYPE_I32, n_used, n_tokens);
    ggml_tensor * tc  = ggml_mul_mat_id(g.ctx, ta, tb, tid);
    g.gf = ggml_new_graph_custom(g.ctx, 16, false);
    ggml_build_forward_expand(g.gf, tc);
    if (!g.alloc()) return -2;
    ggml_backend_tensor_set(ta, as, 0, ggml_nbytes(ta));
    ggml_backend_tensor_set(tb, b, 0, ggml_nbytes(tb));
    ggml_backend_tensor_set(tid, ids, 0, ggml_nbytes(tid));
    if (ggml_backend_graph_compute(g.backend, g.gf) != GGML_STATUS_SUCCESS) return -3;
    ggml_backend_tensor_get(tc, c, 0, ggml_nbytes(tc));
    return 0;
}

// Time MUL_MAT on the reference CPU backend the way the reference's own perf harness does (tests/test-backend-ops.cpp:
// 1455-1583, eval_perf): ONE graph holding `iters` MUL_MAT nodes, so the threadpool start-up is paid once per graph and
// not per matmul.  The nodes cycle over `n_mats` distinct weight matrices so the working set can exceed the LLC.
// Returns seconds per single matmul (mean), or <0 on error.  Used by bench.py's cpu_baseline / --impl reference arm.
SHIM_API double ref_time_mul_mat(int wtype, const void * W_one, int64_t m, int64_t n, int64_t k, int n_mats,
                                 int iters, int n_threads) {
    if (iters < 1 || n_mats < 1) return -1.0;
    ggml_init_params ip = { ggml_tensor_overhead() * (size_t) (n_mats + 2 * iters + 16) + ggml_graph_overhead_custom(iters + 16, false),
                            nullptr, true };
    ggml_context * ctx = ggml_init(ip);
    ggml_backend_t backend = ggml_backend_cpu_init();
    if (!ctx || !backend) return -1.0;
    if (n_threads > 0) ggml_backend_cpu_set_n_threads(backend, n_threads);
    std::vector<ggml_tensor *> as(n_mats);
    ggml_tensor * b = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, k, n);
    for (int i = 0; i < n_mats; i++) as[i] = ggml_new_tensor_2d(ctx, (ggml_type) wtype, k, m);
    ggml_cgraph * gf = ggml_new_graph_custom(ctx, iters + 16, false);
    ggml_cgraph * gw = ggml_new_graph_custom(ctx, 16, false);
    for (int i = 0; i < iters; i++) ggml_build_forward_expand(gf, ggml_mul_mat(ctx, as[i % n_mats], b));
    ggml_build_forward_expand(gw, ggml_mul_mat(ctx, as[0], b));
    ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors(ctx, backend);
    double result = -2.0;
    if (buf) {
        for (int i = 0; i < n_mats; i++) ggml_backend_tensor_set(as[i], W_one, 0, ggml_nbytes(as[i]));
        std::vector<float> x((size_t) k * n);
        for (size_t i = 0; i < x.size(); i++) x[i] = (float) ((i * 2654435761u) % 2001) / 1000.0f - 1.0f;
        ggml_backend_tensor_set(b, x.data(), 0, ggml_nbytes(b));
        ggml_backend_graph_compute(backend, gw);   // warm-up
        const auto t0 = std::chrono::steady_clock::now();
        const bool ok = ggml_backend_graph_compute(backend, gf) == GGML_STATUS_SUCCESS;
        const auto t1 = std::chrono::steady_clock::now();
        result = ok ? std::chrono::duration<double>(t1 - t0).count() / iters : -3.0;
        ggml_backend_buffer_free(buf);
    }
    ggml_backend_free(backend);
    ggml_free(ctx);
    return result;
}
