// abi.cu — the extern "C" entry points declared in include/ggml_b200.h: context, argument validation and dispatch
// onto the kernels (quantize.cu, mmvq.cu, mmid.cu, mmq_tc.cu).  No compute happens on the host and there is no CPU
// fallback: if CUDA is unavailable every call returns B200_E_CUDA.
#include "mm_problem.cuh"
#include "quantize.cuh"
#include "stream_jobs.cuh"

#include <cstdarg>
#include <vector>
#include <cstdio>
#include <cstdlib>
#include <cstring>

// kernels' host launchers (other TUs)
int    b200_launch_gemv_generic(b200_ctx * ctx, int type, const mm_problem & P, int nc, int64_t tiles);
int    b200_launch_gemv_q4k(b200_ctx * ctx, const mm_problem & P, int nc, int64_t tiles, bool late_wait);
int    b200_launch_gemv_fused(b200_ctx * ctx, int type, const mm_problem & P, const uint8_t * Wg, const float * bias_up, const int64_t * bu_nb,
                              const float * bias_gate, const int64_t * bg_nb, int glu_op, int64_t batches);
size_t b200_gemv_q4k_smem(int nc, int64_t k);
int    b200_gemv_mma_ok(const b200_ctx * ctx, int type, const mm_problem & P);
int    b200_launch_gemv_mma(b200_ctx * ctx, int type, const mm_problem & P, int64_t tiles, bool late_wait, int64_t live_tiles = 0);
int    b200_launch_mm_ids(b200_ctx * ctx, const int32_t * ids, int64_t ids_stride, int n_expert, int n_used, int64_t n_tokens,
                          int32_t * ids_src1, int32_t * expert_bounds);
int    b200_tc_supported(const b200_ctx * ctx, int type, const mm_problem & P);
int    b200_launch_dequant_f16(b200_ctx * ctx, int type, const uint8_t * W, int64_t nb01, int64_t rows, int64_t k, void * out);
int    b200_launch_gemm_tc(b200_ctx * ctx, int type, const mm_problem & P, int64_t n_pos, size_t ws_used);
int    b200_launch_ll_unpack(b200_ctx * ctx, const uint2 * buf, int64_t cap, uint32_t * state, int n_mats, const int64_t * m_full, float * const * Y);
int    b200_mul_mat_multi_tc(b200_ctx * ctx, int n_mats, const b200_tensor * const * a, const b200_tensor * b, const b200_tensor * const * d);
size_t b200_gemm_tc_scratch(const b200_ctx * ctx, const mm_problem & P, int64_t n_pos, size_t ws_used);
int    b200_mul_mat_ex(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * b, const b200_tensor * d,
                       int n_peers, float * const * peer_Y, int64_t peer_m, int64_t peer_row0, const float * bias = nullptr);

// ---- errors ---------------------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

void b200_set_error(const char * fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" {

int          b200_abi_version(void) { return B200_ABI_VERSION; }
const char * b200_last_error(void)  { return g_err; }

int b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    int ok = 0;
    for (int i = 0; i < n; ++i) {
        int major = 0;
        if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, i) == cudaSuccess && major == 10) ok++;
        else break;    // devices are exposed as a dense prefix 0..ok-1
    }
    return ok;
}

int b200_device_info(int device, char * name, size_t name_len, size_t * free_bytes, size_t * total_bytes, int * cc_major,
                     int * cc_minor, int * sm_count) {
    cudaDeviceProp prop;
    B200_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (name && name_len) { strncpy(name, prop.name, name_len - 1); name[name_len - 1] = 0; }
    if (cc_major) *cc_major = prop.major;
    if (cc_minor) *cc_minor = prop.minor;
    if (sm_count) *sm_count = prop.multiProcessorCount;
    if (free_bytes || total_bytes) {
        size_t f = 0, t = 0;
        B200_CUDA_TRY(cudaSetDevice(device));
        B200_CUDA_TRY(cudaMemGetInfo(&f, &t));
        if (free_bytes) *free_bytes = f;
        if (total_bytes) *total_bytes = t;
    }
    return B200_OK;
}

int    b200_blck_size(int type) { return b200_blck(type); }
size_t b200_type_size(int type) { return (size_t) b200_tsize(type); }
size_t b200_row_size(int type, int64_t k) {
    const int b = b200_blck(type);
    return b ? (size_t) (k / b) * (size_t) b200_tsize(type) : 0;
}

// ---- context ------------------------------------------------------------------------------------------------------
static int env_int(const char * name, int dflt) {
    const char * v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

int b200_ctx_create(int device, void * stream, b200_ctx ** out) {
    if (!out) { b200_set_error("ctx_create: out is NULL"); return B200_E_ARG; }
    *out = nullptr;
    B200_CUDA_TRY(cudaSetDevice(device));
    b200_ctx * ctx = (b200_ctx *) calloc(1, sizeof(b200_ctx));
    if (!ctx) { b200_set_error("ctx_create: out of host memory"); return B200_E_ALLOC; }
    ctx->device = device;
    cudaDeviceProp prop;
    cudaError_t err = cudaGetDeviceProperties(&prop, device);
    if (err != cudaSuccess) { free(ctx); b200_set_error("cudaGetDeviceProperties: %s", cudaGetErrorString(err)); return B200_E_CUDA; }
    if (prop.major != 10) {
        free(ctx);
        b200_set_error("device %d is sm_%d%d; this library carries sm_100a code only", device, prop.major, prop.minor);
        return B200_E_UNSUPPORTED;
    }
    ctx->sm_count   = prop.multiProcessorCount;
    ctx->cc_major   = prop.major;
    ctx->cc_minor   = prop.minor;
    ctx->smem_optin = prop.sharedMemPerBlockOptin;
    if (stream) {
        ctx->stream = (cudaStream_t) stream;
    } else {
        err = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
        if (err != cudaSuccess) { free(ctx); b200_set_error("cudaStreamCreate: %s", cudaGetErrorString(err)); return B200_E_CUDA; }
        ctx->owns_stream = true;
    }
    ctx->force_generic = env_int("GGML_B200_FORCE_GENERIC", 0);
    ctx->use_tc        = env_int("GGML_B200_USE_TC", 1);
    ctx->tc_pair       = env_int("GGML_B200_TC_PAIR", 1);
    ctx->gather_push   = env_int("GGML_B200_GATHER_PUSH", 1);
    ctx->max_gemv_cols = env_int("GGML_B200_MAX_GEMV_COLS", 8);
    ctx->pdl           = env_int("GGML_B200_PDL", 1);
    ctx->fuse_quant    = env_int("GGML_B200_FUSE_QUANT", 1);
    ctx->gemv_cfg      = env_int("GGML_B200_GEMV_CFG", 0);
    ctx->tc_nt         = env_int("GGML_B200_TC_NT", 0);
    ctx->gemv_rpw      = env_int("GGML_B200_GEMV_RPW", 1);
    ctx->gemv_rpw_max_m = env_int("GGML_B200_GEMV_RPW_MAX_M", 4096);
    ctx->tc_pdl        = env_int("GGML_B200_TC_PDL", 0);
    ctx->ops_pdl       = env_int("GGML_B200_OPS_PDL", 1);
    ctx->tc_dequant    = env_int("GGML_B200_TC_DEQUANT", 1);
    ctx->use_mma       = env_int("GGML_B200_USE_MMA", 1);
    ctx->mma_ks        = env_int("GGML_B200_MMA_KS", 0);
    ctx->mma_ch        = env_int("GGML_B200_MMA_CH", 0);
    ctx->mma_min_n     = env_int("GGML_B200_MMA_MIN_N", 1);
    ctx->mma_prequant  = env_int("GGML_B200_MMA_PREQUANT", 1);
    ctx->tc_split_k   = env_int("GGML_B200_TC_SPLIT_K", 1);
    ctx->noop          = env_int("GGML_B200_NOOP", 0);
    ctx->defer_wait    = env_int("GGML_B200_DEFER_WAIT", 1);
    ctx->use_stream    = env_int("GGML_B200_USE_STREAM", 1);
    ctx->stream_min_jobs = env_int("GGML_B200_STREAM_MIN_JOBS", 6);
    ctx->stream_stages = env_int("GGML_B200_STREAM_STAGES", 0);
    ctx->stream_ll     = env_int("GGML_B200_STREAM_LL", 1);
    ctx->stream_grid   = env_int("GGML_B200_STREAM_GRID", 0);
    ctx->stream_chain  = env_int("GGML_B200_STREAM_CHAIN", 0);
    if (ctx->max_gemv_cols < 1) ctx->max_gemv_cols = 1;
    if (ctx->max_gemv_cols > 8) ctx->max_gemv_cols = 8;
    *out = ctx;
    return B200_OK;
}

void b200_ctx_destroy(b200_ctx * ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->ws) cudaFree(ctx->ws);
    if (ctx->gs_state) cudaFree(ctx->gs_state);
    if (ctx->ll_state) cudaFree(ctx->ll_state);
    if (ctx->sq_done) cudaFree(ctx->sq_done);
    if (ctx->sq_ll) cudaFree(ctx->sq_ll);
    if (ctx->sq_img) cudaFree(ctx->sq_img);
    if (ctx->sq_xflag) cudaFree(ctx->sq_xflag);
    if (ctx->sq_emu) cudaFree(ctx->sq_emu);
    if (ctx->sq_trace) cudaFree(ctx->sq_trace);
    if (ctx->dev_x) cudaFree(ctx->dev_x);
    if (ctx->dev_y) cudaFree(ctx->dev_y);
    if (ctx->owns_stream) cudaStreamDestroy(ctx->stream);
    free(ctx);
}

void *  b200_ctx_stream(b200_ctx * ctx) { return ctx ? (void *) ctx->stream : nullptr; }
int64_t b200_ctx_launch_count(const b200_ctx * ctx) { return ctx ? ctx->launches : 0; }

int b200_ctx_synchronize(b200_ctx * ctx) {
    if (!ctx) { b200_set_error("ctx is NULL"); return B200_E_ARG; }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return B200_OK;
}

int b200_ctx_set_option(b200_ctx * ctx, const char * name, int value) {
    if (!ctx || !name) return B200_E_ARG;
    int * slot = nullptr;
    if      (!strcmp(name, "force_generic")) slot = &ctx->force_generic;
    else if (!strcmp(name, "use_tc"))        slot = &ctx->use_tc;
    else if (!strcmp(name, "tc_pair"))       slot = &ctx->tc_pair;
    else if (!strcmp(name, "gather_push"))   slot = &ctx->gather_push;
    else if (!strcmp(name, "max_gemv_cols")) slot = &ctx->max_gemv_cols;
    else if (!strcmp(name, "pdl"))           slot = &ctx->pdl;
    else if (!strcmp(name, "fuse_quant"))    slot = &ctx->fuse_quant;
    else if (!strcmp(name, "gemv_cfg"))      slot = &ctx->gemv_cfg;
    else if (!strcmp(name, "tc_nt"))         slot = &ctx->tc_nt;
    else if (!strcmp(name, "gemv_rpw"))      slot = &ctx->gemv_rpw;
    else if (!strcmp(name, "gemv_rpw_max_m")) slot = &ctx->gemv_rpw_max_m;
    else if (!strcmp(name, "tc_pdl"))        slot = &ctx->tc_pdl;
    else if (!strcmp(name, "ops_pdl"))       slot = &ctx->ops_pdl;
    else if (!strcmp(name, "tc_dequant"))    slot = &ctx->tc_dequant;
    else if (!strcmp(name, "use_mma"))       slot = &ctx->use_mma;
    else if (!strcmp(name, "mma_ks"))        slot = &ctx->mma_ks;
    else if (!strcmp(name, "mma_ch"))        slot = &ctx->mma_ch;
    else if (!strcmp(name, "mma_min_n"))     slot = &ctx->mma_min_n;
    else if (!strcmp(name, "mma_prequant"))  slot = &ctx->mma_prequant;
    else if (!strcmp(name, "tc_split_k"))   slot = &ctx->tc_split_k;
    else if (!strcmp(name, "noop"))          slot = &ctx->noop;
    else if (!strcmp(name, "defer_wait"))    slot = &ctx->defer_wait;
    else if (!strcmp(name, "use_stream"))    slot = &ctx->use_stream;
    else if (!strcmp(name, "stream_min_jobs")) slot = &ctx->stream_min_jobs;
    else if (!strcmp(name, "stream_stages")) slot = &ctx->stream_stages;
    else if (!strcmp(name, "stream_ll"))     slot = &ctx->stream_ll;
    else if (!strcmp(name, "stream_grid"))   slot = &ctx->stream_grid;
    else if (!strcmp(name, "stream_chain"))  slot = &ctx->stream_chain;
    else if (!strcmp(name, "stream_trace"))  slot = &ctx->stream_trace;
    if (!slot) { b200_set_error("unknown option '%s'", name); return B200_E_ARG; }
    const int prev = *slot;
    *slot = value;
    if (ctx->max_gemv_cols < 1) ctx->max_gemv_cols = 1;
    if (ctx->max_gemv_cols > 8) ctx->max_gemv_cols = 8;
    return prev;
}

}  // extern "C"

int b200_ctx_reserve(b200_ctx * ctx, size_t nbytes) {
    if (nbytes <= ctx->ws_size) return B200_OK;
    // the scratch may still be read by kernels in flight on this stream
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    if (ctx->ws) { cudaFree(ctx->ws); ctx->ws = nullptr; ctx->ws_size = 0; }
    ctx->alloc_gen++;
    const size_t want = (nbytes + (size_t) (1 << 20) - 1) & ~((size_t) (1 << 20) - 1);
    cudaError_t err = cudaMalloc(&ctx->ws, want);
    if (err != cudaSuccess) {
        cudaGetLastError();
        b200_set_error("scratch allocation of %zu B failed: %s", want, cudaGetErrorString(err));
        return B200_E_ALLOC;
    }
    ctx->ws_size = want;
    return B200_OK;
}

// ---- validation -----------------------------------------------------------------------------------------------------
static bool is_weight_type(int t) {
    return t == B200_TYPE_Q4_0 || t == B200_TYPE_Q8_0 || t == B200_TYPE_Q4_K || t == B200_TYPE_Q6_K ||
           t == B200_TYPE_Q4_1 || t == B200_TYPE_Q5_0 || t == B200_TYPE_Q5_1 || t == B200_TYPE_Q5_K ||       // (Q5_K: generic GEMV + tcgen05 GEMM)
           t == B200_TYPE_Q2_K || t == B200_TYPE_Q3_K || t == B200_TYPE_IQ4_NL || t == B200_TYPE_IQ4_XS || t == B200_TYPE_MXFP4 ||
           t == B200_TYPE_NVFP4;                                                                                                       // (generic GEMV family only)
}
static int64_t weight_align(int t) { return t == B200_TYPE_MXFP4 ? 1 : 2; }        // 17-byte MXFP4 blocks: rows are only byte-aligned

static bool f32_contiguous(const b200_tensor * t) {
    return t->nb[0] == 4 && t->nb[1] == t->ne[0] * 4 && t->nb[2] == t->nb[1] * t->ne[1] && t->nb[3] == t->nb[2] * t->ne[2];
}

static const char * check_mul_mat(const b200_tensor * a, const b200_tensor * b, const b200_tensor * d) {
    if (!a || !b || !d) return "NULL tensor";
    if (!is_weight_type(a->type)) return "src0 type is not one of Q4_0 / Q4_1 / Q5_0 / Q5_1 / Q8_0 / Q2_K / Q3_K / Q4_K / Q5_K / Q6_K / IQ4_NL / IQ4_XS / MXFP4 / NVFP4";
    if (b->type != B200_TYPE_F32 || d->type != B200_TYPE_F32) return "src1/dst must be F32";
    const int64_t k = a->ne[0];
    if (k <= 0 || k % b200_blck(a->type) != 0) return "ne00 is not a whole number of blocks";
    if (b->ne[0] != k) return "src1->ne[0] != src0->ne[0]";
    if (d->ne[0] != a->ne[1] || d->ne[1] != b->ne[1] || d->ne[2] != b->ne[2] || d->ne[3] != b->ne[3]) return "dst shape mismatch";
    if (a->ne[2] <= 0 || a->ne[3] <= 0 || b->ne[2] % a->ne[2] != 0 || b->ne[3] % a->ne[3] != 0) return "batch dims do not broadcast";
    if (a->nb[0] != (int64_t) b200_tsize(a->type)) return "src0 is transposed/permuted in dim 0";
    if (a->nb[1] % weight_align(a->type) != 0 || a->nb[2] % weight_align(a->type) != 0 || a->nb[3] % weight_align(a->type) != 0) return "src0 strides not 2-byte aligned";
    if (b->nb[0] != 4) return "src1 is not contiguous in dim 0";
    if (b->nb[1] % 4 != 0 || b->nb[2] % 4 != 0 || b->nb[3] % 4 != 0) return "src1 strides not 4-byte aligned";
    if (!f32_contiguous(d)) return "dst is not contiguous";
    if (k > (int64_t) 1 << 24) return "k too large";
    return nullptr;
}

static const char * check_mul_mat_id(const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids, const b200_tensor * c) {
    if (!as || !b || !ids || !c) return "NULL tensor";
    if (!is_weight_type(as->type)) return "as type is not a supported weight type";
    if (b->type != B200_TYPE_F32 || c->type != B200_TYPE_F32 || ids->type != B200_TYPE_I32) return "b/c must be F32 and ids I32";
    const int64_t k = as->ne[0];
    if (k <= 0 || k % b200_blck(as->type) != 0) return "ne00 is not a whole number of blocks";
    if (as->ne[3] != 1 || b->ne[3] != 1 || c->ne[3] != 1 || ids->ne[2] != 1 || ids->ne[3] != 1) return "4-D MUL_MAT_ID";
    if (b->ne[0] != k) return "b->ne[0] != as->ne[0]";
    const int64_t n_used = ids->ne[0], n_tokens = ids->ne[1];
    if (b->ne[2] != n_tokens) return "b->ne[2] != n_tokens";
    if (b->ne[1] <= 0 || n_used % b->ne[1] != 0) return "n_used is not a multiple of b->ne[1]";   // ggml.c:3305
    if (c->ne[0] != as->ne[1] || c->ne[1] != n_used || c->ne[2] != n_tokens) return "c shape mismatch";
    if (as->nb[0] != (int64_t) b200_tsize(as->type) || as->nb[1] % weight_align(as->type) != 0 || as->nb[2] % weight_align(as->type) != 0) return "as strides";
    if (b->nb[0] != 4 || b->nb[1] % 4 != 0 || b->nb[2] % 4 != 0) return "b strides";
    if (ids->nb[0] != 4 || ids->nb[1] % 4 != 0) return "ids strides";
    if (!f32_contiguous(c)) return "c is not contiguous";
    if (n_used > 1024 || as->ne[2] > 65535) return "too many experts";
    return nullptr;
}

// ---- in-flight write tracking for the deferred dependency wait of the fused GEMV (mmvq.cu, flags bit 1) -----------------
// A GEMV launched with the PDL attribute whose inputs are not written by any kernel that may still be running defers its
// griddepcontrol.wait to just before its first store.  "May still be running" = every kernel launched since (and
// including) the last one that waited early: an early waiter proves all of ITS predecessors complete, and because every
// kernel waits before it stores/finishes, completion is transitive.
static void tensor_extent(const b200_tensor * t, const char ** lo, const char ** hi) {
    size_t span = (size_t) b200_tsize(t->type);
    if (b200_blck(t->type) > 1) span = (size_t) t->nb[1];       // quantized rows: one whole row
    for (int i = (b200_blck(t->type) > 1 ? 1 : 0); i < 4; ++i) span += (size_t) (t->ne[i] > 0 ? t->ne[i] - 1 : 0) * (size_t) t->nb[i];
    *lo = (const char *) t->data;
    *hi = *lo + span;
}

static bool pending_overlaps(const b200_ctx * ctx, const b200_tensor * t) {
    const char * lo, * hi;
    tensor_extent(t, &lo, &hi);
    for (int i = 0; i < ctx->n_pend; ++i) if (lo < ctx->pend_hi[i] && ctx->pend_lo[i] < hi) return true;
    return false;
}

// record the output of a kernel about to be launched; `fresh` = this kernel waits early (or is a normal launch), so
// everything before it is known complete by the time it runs
static void pending_push(b200_ctx * ctx, const b200_tensor * dst, bool fresh) {
    if (fresh) ctx->n_pend = 0;
    const char * lo, * hi;
    tensor_extent(dst, &lo, &hi);
    ctx->pend_lo[ctx->n_pend] = lo; ctx->pend_hi[ctx->n_pend] = hi; ctx->n_pend++;
}

static void pending_reset_range(b200_ctx * ctx, const void * ptr, size_t nbytes) {
    ctx->n_pend = 1;
    ctx->pend_lo[0] = (const char *) ptr;
    ctx->pend_hi[0] = (const char *) ptr + nbytes;
}

static bool can_defer_wait(b200_ctx * ctx, const b200_tensor * x, const b200_tensor * ids) {
    if (!ctx->pdl || !ctx->defer_wait || ctx->n_pend >= 16) return false;
    if (pending_overlaps(ctx, x)) return false;
    if (ids && pending_overlaps(ctx, ids)) return false;
    return true;
}

static bool aligned16(const void * p, int64_t a, int64_t b, int64_t c) {
    return (((uintptr_t) p | (uintptr_t) a | (uintptr_t) b | (uintptr_t) c) & 15) == 0;
}

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// quantize src1 [k, ne1, ne2, ne3] into planes inside the scratch (after `reserve_extra` bytes kept for the caller)
static int quantize_into_ws(b200_ctx * ctx, const b200_tensor * b, size_t extra, q8_planes * out, uint8_t ** extra_ptr) {
    const int64_t k = b->ne[0];
    const int64_t ncols = b->ne[1] * b->ne[2] * b->ne[3];
    const size_t qs_bytes = align_up((size_t) ncols * k, 256);
    const size_t ds_bytes = align_up((size_t) ncols * (k / 32) * sizeof(half2), 256);
    const int rc = b200_ctx_reserve(ctx, qs_bytes + ds_bytes + align_up(extra, 256) + 256);
    if (rc) return rc;
    uint8_t * base = (uint8_t *) ctx->ws;
    int8_t * qs = (int8_t *) base;
    half2 *  ds = (half2 *) (base + qs_bytes);
    if (extra_ptr) *extra_ptr = base + qs_bytes + ds_bytes;
    out->qs = qs; out->ds = ds; out->k = k;
    return b200_launch_quantize_planes(ctx, (const float *) b->data, k, b->ne[1], b->ne[2], b->ne[3], b->nb[1], b->nb[2], b->nb[3], qs, ds);
}

extern "C" {

// ---- (1) ----------------------------------------------------------------------------------------------------------
int b200_quantize_row_q8_1(b200_ctx * ctx, const float * x, int64_t x_row_stride_bytes, void * y, int64_t k, int64_t nrows) {
    if (!ctx || !x || !y) { b200_set_error("quantize_row_q8_1: NULL argument"); return B200_E_ARG; }
    if (x_row_stride_bytes % 4 != 0) { b200_set_error("quantize_row_q8_1: row stride must be a multiple of 4"); return B200_E_ARG; }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    pending_reset_range(ctx, y, (size_t) nrows * (size_t) (k / 32) * SZ_Q8_1);
    return b200_launch_quantize_aos(ctx, x, k, nrows, 1, 1, x_row_stride_bytes, x_row_stride_bytes * nrows, x_row_stride_bytes * nrows, y);
}

// ---- (2)(3) -------------------------------------------------------------------------------------------------------
int b200_mul_mat_supported(const b200_tensor * a, const b200_tensor * b, const b200_tensor * d) {
    if (check_mul_mat(a, b, d)) return 0;
    const int64_t batches = b->ne[2] * b->ne[3];
    const int64_t tiles   = batches * ((b->ne[1] + 7) / 8);
    return tiles <= 65535 ? 1 : 0;
}

int b200_mul_mat(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * b, const b200_tensor * d) {
    return b200_mul_mat_ex(ctx, a, b, d, 0, nullptr, 0, 0);
}

}  // extern "C"

// MUL_MAT + ADD in the store of the fast GEMV kernels (n <= 8): dst = src0 x src1 + addend, addend shaped like dst (a bias repeated by the
// caller, or the residual stream).  dst may be the addend's own memory (each element is read and written by one thread); it must not
// overlap src1.
extern "C" int b200_mul_mat_add_supported(const b200_tensor * a, const b200_tensor * b, const b200_tensor * addend, const b200_tensor * d) {
    if (!a || !b || !addend || !d || check_mul_mat(a, b, d)) return 0;
    if (b->ne[1] > 8 || b->ne[2] * b->ne[3] > 65535) return 0;
    if (addend->type != B200_TYPE_F32 || !f32_contiguous(addend)) return 0;
    for (int i = 0; i < 4; ++i) if (addend->ne[i] != d->ne[i]) return 0;
    const char * xl, * xh, * dl, * dh;
    tensor_extent(b, &xl, &xh); tensor_extent(d, &dl, &dh);
    return !(xl < dh && dl < xh);
}

extern "C" int b200_mul_mat_add(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * b, const b200_tensor * addend, const b200_tensor * d) {
    if (!ctx) { b200_set_error("mul_mat_add: ctx is NULL"); return B200_E_ARG; }
    if (!b200_mul_mat_add_supported(a, b, addend, d) || b->ne[1] > ctx->max_gemv_cols) {
        b200_set_error("mul_mat_add: quantized src0, n <= 8, a contiguous F32 addend shaped like dst, dst not overlapping src1"); return B200_E_UNSUPPORTED;
    }
    const int rc = b200_mul_mat_ex(ctx, a, b, d, 0, nullptr, 0, 0, (const float *) addend->data);
    ctx->n_pend = 16;          // the addend may come from a kernel the write tracker does not know
    return rc;
}

// ---- N2: MUL_MAT (+ bias) (+ gate MUL_MAT (+ bias) + GLU), one activation column per batch ---------------------------------------
static const char * check_fused(const b200_tensor * up, const b200_tensor * gate, const b200_tensor * x, const b200_tensor * bu, const b200_tensor * bg,
                                int glu_op, const b200_tensor * d) {
    if (const char * why = check_mul_mat(up, x, d)) return why;
    if (x->ne[1] != 1) return "one activation column per batch";
    if (x->ne[2] * x->ne[3] > 65535) return "too many batches";
    if (gate) {
        if (glu_op < B200_GLU_REGLU || glu_op > B200_GLU_GEGLU_QUICK || glu_op == B200_GLU_SWIGLU_OAI) return "unsupported GLU op";
        if (gate->type != up->type) return "gate and up must have one type";
        for (int i = 0; i < 4; ++i) if (gate->ne[i] != up->ne[i] || gate->nb[i] != up->nb[i]) return "gate and up must have one shape and layout";
    } else if (bg) return "gate bias without a gate";
    for (const b200_tensor * b : {bu, bg}) {
        if (!b) continue;
        if (b->type != B200_TYPE_F32 || b->nb[0] != 4) return "bias must be f32 with nb[0] == 4";
        for (int i = 0; i < 4; ++i) if (b->ne[i] != d->ne[i]) return "bias must be shaped like dst";
    }
    return nullptr;
}

extern "C" int b200_mul_mat_fused_supported(const b200_tensor * up, const b200_tensor * gate, const b200_tensor * x, const b200_tensor * bias_up,
                                            const b200_tensor * bias_gate, int glu_op, const b200_tensor * dst) {
    return up && x && dst && check_fused(up, gate, x, bias_up, bias_gate, glu_op, dst) == nullptr;
}

extern "C" int b200_mul_mat_fused(b200_ctx * ctx, const b200_tensor * up, const b200_tensor * gate, const b200_tensor * x, const b200_tensor * bias_up,
                                  const b200_tensor * bias_gate, int glu_op, const b200_tensor * d) {
    if (!ctx || !up || !x || !d) { b200_set_error("mul_mat_fused: NULL argument"); return B200_E_ARG; }
    if (const char * why = check_fused(up, gate, x, bias_up, bias_gate, glu_op, d)) { b200_set_error("mul_mat_fused: %s", why); return B200_E_UNSUPPORTED; }
    const int64_t batches = x->ne[2] * x->ne[3];
    if (up->ne[1] == 0 || batches == 0) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    mm_problem P = {};
    P.W = (const uint8_t *) up->data; P.nb01 = up->nb[1]; P.nb02 = up->nb[2]; P.nb03 = up->nb[3];
    P.m = up->ne[1]; P.k = up->ne[0];
    P.X = (const float *) x->data; P.nb11 = x->nb[1]; P.nb12 = x->nb[2]; P.nb13 = x->nb[3];
    P.ne11 = 1; P.ne12 = x->ne[2]; P.ne13 = x->ne[3];
    P.r2 = x->ne[2] / up->ne[2]; P.r3 = x->ne[3] / up->ne[3];
    P.Y = (float *) d->data;
    P.mode = MM_PLAIN;
    pending_push(ctx, d, true);
    ctx->n_pend = 16;            // the biases may come from kernels the tracker does not know; nothing is deferred around this launch
    const int rc = quantize_into_ws(ctx, x, 0, &P.act, nullptr);
    if (rc) return rc;
    return b200_launch_gemv_fused(ctx, up->type, P, gate ? (const uint8_t *) gate->data : nullptr, bias_up ? (const float *) bias_up->data : nullptr,
                                  bias_up ? &bias_up->nb[2] : nullptr, bias_gate ? (const float *) bias_gate->data : nullptr,
                                  bias_gate ? &bias_gate->nb[2] : nullptr, glu_op, batches);
}

// MUL_MAT with the row-split all-gather fused into the store: results also go to peer_Y[p][col*peer_m + peer_row0 + row]
int b200_mul_mat_ex(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * b, const b200_tensor * d,
                    int n_peers, float * const * peer_Y, int64_t peer_m, int64_t peer_row0, const float * bias) {
    if (!ctx) { b200_set_error("mul_mat: ctx is NULL"); return B200_E_ARG; }
    if (n_peers < 0 || n_peers > B200_MAX_PEERS) { b200_set_error("mul_mat: n_peers out of range"); return B200_E_ARG; }
    if (const char * why = check_mul_mat(a, b, d)) { b200_set_error("mul_mat: %s", why); return B200_E_ARG; }
    const int64_t k = a->ne[0], m = a->ne[1], n = b->ne[1];
    const int64_t batches = b->ne[2] * b->ne[3];
    if (m == 0 || n == 0 || batches == 0) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));

    mm_problem P = {};
    P.W = (const uint8_t *) a->data; P.nb01 = a->nb[1]; P.nb02 = a->nb[2]; P.nb03 = a->nb[3];
    P.m = m; P.k = k;
    P.X = (const float *) b->data; P.nb11 = b->nb[1]; P.nb12 = b->nb[2]; P.nb13 = b->nb[3];
    P.ne11 = n; P.ne12 = b->ne[2]; P.ne13 = b->ne[3];
    P.r2 = b->ne[2] / a->ne[2]; P.r3 = b->ne[3] / a->ne[3];
    P.Y = (float *) d->data;
    P.bias = bias;
    if (bias && n > ctx->max_gemv_cols) { b200_set_error("mul_mat: an addend is taken by the GEMV kernels only (n <= %d)", ctx->max_gemv_cols); return B200_E_UNSUPPORTED; }
    P.mode = MM_PLAIN;
    P.n_peers = n_peers; P.peer_m = peer_m; P.peer_row0 = peer_row0;
    for (int p = 0; p < n_peers; ++p) P.peer_Y[p] = peer_Y[p];

    if (n <= ctx->max_gemv_cols) {
        if (a->type == B200_TYPE_Q4_K && !ctx->force_generic && ctx->fuse_quant && batches <= 65535 &&
            aligned16(a->data, a->nb[1], a->nb[2], a->nb[3]) && aligned16(b->data, b->nb[1], b->nb[2], b->nb[3]) &&
            b200_gemv_q4k_smem((int) n, k) <= ctx->smem_optin) {
            // 2..8 columns: integer tensor-core GEMV (mmvq_mma.cu).  Measured (us dp4a / mma, CUDA graph, rotating weights): n = 2:
            // 1024x4096 2.9 / 3.3, 4096x4096 7.0 / 6.3, 14336x4096 12.8 / 10.4, output head 74.5 / 63.6; n = 8: 8.3 / 4.6, 20.0 / 8.7,
            // 39.7 / 12.7, 241 / 66.5.  The choice must not depend on m: a row shard has to run the same arithmetic as the whole
            // matrix (row split).  use_mma = 0 keeps the dp4a kernel.
            const bool mma_pays = n >= 2;
            if (n >= 2 && ctx->use_mma && mma_pays && !(n_peers > 0 && ctx->gs_world > 0) && b200_gemv_mma_ok(ctx, B200_TYPE_Q4_K, P)) {
                const bool late = n_peers == 0 && can_defer_wait(ctx, b, nullptr);
                pending_push(ctx, d, !late);
                return b200_launch_gemv_mma(ctx, B200_TYPE_Q4_K, P, batches, late);
            }
            const bool late = n_peers == 0 && can_defer_wait(ctx, b, nullptr);
            pending_push(ctx, d, !late);
            if (n_peers > 0 && ctx->gs_world > 0) {
                P.sync_flags = ctx->gs_flags[ctx->gs_rank];
                for (int p = 0; p < ctx->gs_world; ++p) P.peer_flags[p] = ctx->gs_flags[p];
                P.sync_state = ctx->gs_state; P.sync_rank = ctx->gs_rank; P.sync_world = ctx->gs_world;
            }
            return b200_launch_gemv_q4k(ctx, P, (int) n, batches, late);
        }
        if (n_peers > 0 && ctx->gs_world > 0) { b200_set_error("mul_mat_gather: the completion protocol needs the fused Q4_K GEMV"); return B200_E_UNSUPPORTED; }
        // Q8_0 / Q4_0 / Q6_K: tensor-core GEMV (mmvq_mma.cu) from ctx->mma_min_n columns on (the choice depends on n only, see above)
        if ((a->type == B200_TYPE_Q8_0 || a->type == B200_TYPE_Q4_0 || a->type == B200_TYPE_Q6_K) && ctx->use_mma && !ctx->force_generic && ctx->fuse_quant &&
            n >= ctx->mma_min_n && batches <= 65535 && b200_gemv_mma_ok(ctx, a->type, P)) {
            const bool late = n_peers == 0 && can_defer_wait(ctx, b, nullptr);
            pending_push(ctx, d, !late);
            return b200_launch_gemv_mma(ctx, a->type, P, batches, late);
        }
        // rows too long to keep all n columns in shared memory (k = 14336 at n > 4 / 6): two passes over the weights with half
        // the columns each still beat the CUDA-core kernel (2 x 20 us against 53-76 us)
        if ((a->type == B200_TYPE_Q8_0 || a->type == B200_TYPE_Q6_K) && ctx->use_mma && !ctx->force_generic && ctx->fuse_quant &&
            n > 4 && n >= ctx->mma_min_n && batches == 1 && n_peers == 0) {
            mm_problem P2 = P;
            P2.ne11 = 4;
            if (b200_gemv_mma_ok(ctx, a->type, P2)) {
                const bool late = can_defer_wait(ctx, b, nullptr);
                pending_push(ctx, d, !late);
                for (int64_t c0 = 0; c0 < n; c0 += 4) {
                    P2 = P;
                    P2.X    = reinterpret_cast<const float *>(reinterpret_cast<const char *>(P.X) + c0 * P.nb11);
                    P2.Y    = P.Y + c0 * m;
                    P2.bias = P.bias ? P.bias + c0 * m : nullptr;
                    P2.ne11 = n - c0 < 4 ? n - c0 : 4;
                    const int rc = b200_launch_gemv_mma(ctx, a->type, P2, 1, late && c0 == 0);
                    if (rc) return rc;
                }
                return B200_OK;
            }
        }
        pending_push(ctx, d, true);
        int rc = quantize_into_ws(ctx, b, 0, &P.act, nullptr);
        if (rc) return rc;
        return b200_launch_gemv_generic(ctx, a->type, P, (int) n, batches);
    }
    pending_push(ctx, d, true);
    if (n_peers > 0 && ctx->gs_world > 0) { b200_set_error("mul_mat_gather: the in-kernel completion protocol is for n <= 8; use a barrier for prefill"); return B200_E_UNSUPPORTED; }

    if ((n_peers == 0 || batches == 1) && ctx->use_tc && !ctx->force_generic && b200_tc_supported(ctx, a->type, P)) {
        // (row split: the tcgen05 epilogue stores every tile into all peers' full-height outputs as well)
        const int rc = b200_ctx_reserve(ctx, b200_gemm_tc_scratch(ctx, P, n * batches, 0));
        if (rc) return rc;
        return b200_launch_gemm_tc(ctx, a->type, P, n * batches, 0);
    }
    // a weight type the tcgen05 GEMM has no converter for (Q4_1, Q5_0 / Q5_1, Q2_K, Q3_K, IQ4_NL / XS, MXFP4, NVFP4) and three or more column
    // tiles of the generic GEMV: dequantize the matrix into fp16 scratch (one pass over the blocks, 2 bytes per weight written) and run the
    // tensor-core GEMM on that — the GEMV tiles would walk the blocks ceil(n / 8) times with dp4a
    if (n_peers == 0 && batches == 1 && n >= 17 && ctx->use_tc && ctx->tc_dequant && !ctx->force_generic && a->type != B200_TYPE_F16) {
        mm_problem PF = P;
        PF.W = nullptr; PF.nb01 = P.k * 2; PF.nb02 = PF.nb03 = PF.nb01 * P.m;
        if (b200_tc_supported(ctx, B200_TYPE_F16, PF)) {
            const size_t wbytes = ((size_t) P.m * (size_t) P.k * 2 + 1023) / 1024 * 1024;
            int rc = b200_ctx_reserve(ctx, b200_gemm_tc_scratch(ctx, PF, n, wbytes));
            if (rc) return rc;
            PF.W = (const uint8_t *) ctx->ws;
            rc = b200_launch_dequant_f16(ctx, a->type, P.W, P.nb01, P.m, P.k, ctx->ws);
            if (rc) return rc;
            return b200_launch_gemm_tc(ctx, B200_TYPE_F16, PF, n, wbytes);
        }
    }
    int rc = quantize_into_ws(ctx, b, 0, &P.act, nullptr);
    if (rc) return rc;
    const int64_t tiles = batches * ((n + 7) / 8);
    return b200_launch_gemv_generic(ctx, a->type, P, 8, tiles);
}

extern "C" {

// ---- (4) ----------------------------------------------------------------------------------------------------------
int b200_mul_mat_id_supported(const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids, const b200_tensor * c) {
    return check_mul_mat_id(as, b, ids, c) ? 0 : 1;
}

int b200_mm_ids(b200_ctx * ctx, const int32_t * ids, int64_t ids_nb1, int n_expert, int n_used, int64_t n_tokens,
                int32_t * ids_src1, int32_t * expert_bounds) {
    if (!ctx || !ids || !ids_src1 || !expert_bounds || ids_nb1 % 4 != 0) { b200_set_error("mm_ids: bad argument"); return B200_E_ARG; }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    ctx->n_pend = 2;
    ctx->pend_lo[0] = (const char *) ids_src1;      ctx->pend_hi[0] = (const char *) (ids_src1 + (size_t) n_tokens * n_used);
    ctx->pend_lo[1] = (const char *) expert_bounds; ctx->pend_hi[1] = (const char *) (expert_bounds + n_expert + 1);
    return b200_launch_mm_ids(ctx, ids, ids_nb1 / 4, n_expert, n_used, n_tokens, ids_src1, expert_bounds);
}

int b200_mul_mat_id(b200_ctx * ctx, const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids, const b200_tensor * c) {
    if (!ctx) { b200_set_error("mul_mat_id: ctx is NULL"); return B200_E_ARG; }
    if (const char * why = check_mul_mat_id(as, b, ids, c)) { b200_set_error("mul_mat_id: %s", why); return B200_E_ARG; }
    const int64_t k = as->ne[0], m = as->ne[1];
    const int     n_expert = (int) as->ne[2];
    const int     n_used   = (int) ids->ne[0];
    const int64_t n_tokens = ids->ne[1];
    const int64_t n_pairs  = n_tokens * n_used;
    if (m == 0 || n_pairs == 0) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));

    mm_problem P = {};
    P.W = (const uint8_t *) as->data; P.nb01 = as->nb[1]; P.nb02 = as->nb[2]; P.nb03 = 0;
    P.m = m; P.k = k;
    P.X = (const float *) b->data; P.nb11 = b->nb[1]; P.nb12 = b->nb[2]; P.nb13 = 0;
    P.ne11 = b->ne[1]; P.ne12 = n_tokens; P.ne13 = 1; P.r2 = 1; P.r3 = 1;
    P.Y = (float *) c->data;
    P.ids = (const int32_t *) ids->data; P.ids_stride = ids->nb[1] / 4;
    P.n_used = n_used; P.n_expert = n_expert;

    if (n_tokens >= 2 && n_tokens <= ctx->max_gemv_cols && (n_pairs >= n_expert || ctx->use_mma == 2) && as->type == B200_TYPE_Q4_K &&
        ctx->use_mma && !ctx->force_generic && ctx->fuse_quant) {
        // small batches: gather the pairs by expert, then the tensor-core GEMV takes up to 8 pairs of an expert per pass, so an
        // expert's weights are read once however many of the tokens chose it (the per-pair GEMV below reads them once per pair:
        // faster per byte, 5.1 vs 4.3 TB/s, so it keeps the batches in which few pairs share an expert)
        P.mode = MM_ID_SORTED;
        if (b200_gemv_mma_ok(ctx, B200_TYPE_Q4_K, P)) {
            pending_push(ctx, c, true);
            const size_t ids_bytes = (size_t) (n_pairs + n_expert + 1) * sizeof(int32_t);
            int rc = b200_ctx_reserve(ctx, ids_bytes);
            if (rc) return rc;
            int32_t * ids_src1 = (int32_t *) ctx->ws;
            int32_t * bounds   = ids_src1 + n_pairs;
            rc = b200_launch_mm_ids(ctx, P.ids, P.ids_stride, n_expert, n_used, n_tokens, ids_src1, bounds);
            if (rc) return rc;
            P.ids_src1 = ids_src1; P.expert_bounds = bounds;
            return b200_launch_gemv_mma(ctx, B200_TYPE_Q4_K, P, (n_pairs + 7) / 8 + n_expert, false, n_pairs < n_expert ? n_pairs : n_expert);
        }
    }
    if (n_tokens <= ctx->max_gemv_cols && n_pairs <= 65535) {
        // decode: every (token, slot) pair is a GEMV against the expert the kernel looks up itself
        P.mode = MM_ID_PAIRS;
        if (as->type == B200_TYPE_Q4_K && !ctx->force_generic && ctx->fuse_quant &&
            aligned16(as->data, as->nb[1], as->nb[2], 0) && aligned16(b->data, b->nb[1], b->nb[2], 0) &&
            b200_gemv_q4k_smem(1, k) <= ctx->smem_optin) {
            const bool late = can_defer_wait(ctx, b, ids);
            pending_push(ctx, c, !late);
            return b200_launch_gemv_q4k(ctx, P, 1, n_pairs, late);
        }
        pending_push(ctx, c, true);
        int rc = quantize_into_ws(ctx, b, 0, &P.act, nullptr);
        if (rc) return rc;
        return b200_launch_gemv_generic(ctx, as->type, P, 1, n_pairs);
    }

    // prefill: gather the pairs by expert on the device, then column tiles that never straddle two experts
    pending_push(ctx, c, true);
    uint8_t * extra = nullptr;
    const size_t extra_bytes = (size_t) (n_pairs + n_expert + 1) * sizeof(int32_t);
    P.mode = MM_ID_SORTED;
    if (ctx->use_tc && !ctx->force_generic && b200_tc_supported(ctx, as->type, P)) {
        // scratch: [ids_src1 | expert_bounds | fp16 activations gathered by expert | per-position scale and dst column]
        int rc = b200_ctx_reserve(ctx, b200_gemm_tc_scratch(ctx, P, n_pairs, extra_bytes));
        if (rc) return rc;
        int32_t * ids_src1 = (int32_t *) ctx->ws;
        int32_t * bounds   = ids_src1 + n_pairs;
        rc = b200_launch_mm_ids(ctx, P.ids, P.ids_stride, n_expert, n_used, n_tokens, ids_src1, bounds);
        if (rc) return rc;
        P.ids_src1 = ids_src1; P.expert_bounds = bounds;
        return b200_launch_gemm_tc(ctx, as->type, P, n_pairs, extra_bytes);
    }
    if (ctx->use_tc && ctx->tc_dequant && !ctx->force_generic && n_pairs >= 8 * (int64_t) n_expert && as->nb[2] == as->nb[1] * as->ne[1]) {
        // (as above: every expert's matrix -> fp16 scratch, then the grouped tensor-core GEMM; worth it from ~8 positions per expert on)
        mm_problem PF = P;
        PF.W = nullptr; PF.nb01 = k * 2; PF.nb02 = PF.nb03 = PF.nb01 * P.m;
        if (b200_tc_supported(ctx, B200_TYPE_F16, PF)) {
            const size_t off_w  = (extra_bytes + 1023) / 1024 * 1024;
            const size_t wbytes = ((size_t) n_expert * (size_t) P.m * (size_t) k * 2 + 1023) / 1024 * 1024;
            int rc = b200_ctx_reserve(ctx, b200_gemm_tc_scratch(ctx, PF, n_pairs, off_w + wbytes));
            if (rc) return rc;
            int32_t * ids_src1 = (int32_t *) ctx->ws;
            int32_t * bounds   = ids_src1 + n_pairs;
            rc = b200_launch_mm_ids(ctx, P.ids, P.ids_stride, n_expert, n_used, n_tokens, ids_src1, bounds);
            if (rc) return rc;
            PF.ids_src1 = ids_src1; PF.expert_bounds = bounds;
            PF.W = (const uint8_t *) ctx->ws + off_w;
            rc = b200_launch_dequant_f16(ctx, as->type, P.W, P.nb01, (int64_t) n_expert * P.m, k, (void *) PF.W);
            if (rc) return rc;
            return b200_launch_gemm_tc(ctx, B200_TYPE_F16, PF, n_pairs, off_w + wbytes);
        }
    }
    int rc = quantize_into_ws(ctx, b, extra_bytes, &P.act, &extra);
    if (rc) return rc;
    int32_t * ids_src1 = (int32_t *) extra;
    int32_t * bounds   = ids_src1 + n_pairs;
    rc = b200_launch_mm_ids(ctx, P.ids, P.ids_stride, n_expert, n_used, n_tokens, ids_src1, bounds);
    if (rc) return rc;
    P.ids_src1 = ids_src1; P.expert_bounds = bounds;
    const int64_t tiles = (n_pairs + 7) / 8 + n_expert;
    if (tiles > 65535) { b200_set_error("mul_mat_id: too many column tiles"); return B200_E_UNSUPPORTED; }
    return b200_launch_gemv_generic(ctx, as->type, P, 8, tiles);
}

// ---- raw memory -----------------------------------------------------------------------------------------------------
// ---- stream capture of a whole graph_compute (reference: ggml_backend_cuda_graph_compute's CUDA-graph path, ggml-cuda.cu:4457-4490) --------
// Everything the entry points of this library queue on the context's stream between begin and end is recorded instead of executed; an entry
// point that would have to synchronize or reallocate (a scratch buffer that must grow) fails, which invalidates the capture — the caller
// ends it, runs the work eagerly (which grows the buffers) and captures the next time.  A captured graph holds pointers into the
// context's scratch buffers: b200_ctx_alloc_generation() changes whenever one of them moves, graphs captured before that are stale.
struct b200_graph { cudaGraphExec_t exec; int64_t n_launches; int device; };

int b200_capture_begin(b200_ctx * ctx) {
    if (!ctx) { b200_set_error("capture_begin: ctx is NULL"); return B200_E_ARG; }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed));
    ctx->capture_launches0 = ctx->launches;
    return B200_OK;
}

int b200_capture_end(b200_ctx * ctx, void ** graph_out) {
    if (!ctx || !graph_out) { b200_set_error("capture_end: NULL argument"); return B200_E_ARG; }
    *graph_out = nullptr;
    cudaGraph_t g = nullptr;
    const cudaError_t err = cudaStreamEndCapture(ctx->stream, &g);
    const int64_t n = ctx->launches - ctx->capture_launches0;
    ctx->launches = ctx->capture_launches0;                    // nothing ran
    ctx->n_pend = 16;
    if (err != cudaSuccess || !g) {
        cudaGetLastError();
        if (g) cudaGraphDestroy(g);
        b200_set_error("stream capture failed: %s", cudaGetErrorString(err));
        return B200_E_CUDA;
    }
    cudaGraphExec_t exec = nullptr;
    const cudaError_t ierr = cudaGraphInstantiate(&exec, g, 0);
    cudaGraphDestroy(g);
    if (ierr != cudaSuccess) { cudaGetLastError(); b200_set_error("cudaGraphInstantiate failed: %s", cudaGetErrorString(ierr)); return B200_E_CUDA; }
    *graph_out = new b200_graph{exec, n, ctx->device};
    return B200_OK;
}

int b200_graph_launch(b200_ctx * ctx, void * graph) {
    if (!ctx || !graph) { b200_set_error("graph_launch: NULL argument"); return B200_E_ARG; }
    b200_graph * g = (b200_graph *) graph;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaGraphLaunch(g->exec, ctx->stream));
    ctx->launches += g->n_launches;
    ctx->n_pend = 16;
    return B200_OK;
}

void b200_graph_free(void * graph) {
    if (!graph) return;
    b200_graph * g = (b200_graph *) graph;
    device_guard guard;
    cudaSetDevice(g->device);
    cudaGraphExecDestroy(g->exec);
    delete g;
}

uint64_t b200_ctx_alloc_generation(const b200_ctx * ctx) { return ctx ? ctx->alloc_gen : 0; }

int b200_malloc(int device, size_t nbytes, void ** out) {
    if (!out) return B200_E_ARG;
    *out = nullptr;
    B200_CUDA_TRY(cudaSetDevice(device));
    cudaError_t err = cudaMalloc(out, nbytes);
    if (err != cudaSuccess) {
        cudaGetLastError();
        b200_set_error("cudaMalloc(%zu) on device %d: %s", nbytes, device, cudaGetErrorString(err));
        return B200_E_ALLOC;
    }
    return B200_OK;
}

int b200_free(int device, void * ptr) {
    if (!ptr) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(device));
    B200_CUDA_TRY(cudaFree(ptr));
    return B200_OK;
}

int b200_memcpy_h2d(b200_ctx * ctx, void * dst_dev, const void * src_host, size_t nbytes) {
    if (!ctx) return B200_E_ARG;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaMemcpyAsync(dst_dev, src_host, nbytes, cudaMemcpyHostToDevice, ctx->stream));
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return B200_OK;
}

int b200_memcpy_d2h(b200_ctx * ctx, void * dst_host, const void * src_dev, size_t nbytes) {
    if (!ctx) return B200_E_ARG;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaMemcpyAsync(dst_host, src_dev, nbytes, cudaMemcpyDeviceToHost, ctx->stream));
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return B200_OK;
}

int b200_memcpy_d2d(b200_ctx * ctx, void * dst_dev, const void * src_dev, size_t nbytes) {
    if (!ctx) return B200_E_ARG;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaMemcpyAsync(dst_dev, src_dev, nbytes, cudaMemcpyDefault, ctx->stream));
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return B200_OK;
}

int b200_memset(b200_ctx * ctx, void * dst_dev, int value, size_t nbytes) {
    if (!ctx) return B200_E_ARG;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaMemsetAsync(dst_dev, value, nbytes, ctx->stream));
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return B200_OK;
}

int b200_memcpy_async(b200_ctx * ctx, void * dst, const void * src, size_t nbytes, int kind) {
    if (!ctx || kind < 0 || kind > 2) return B200_E_ARG;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    const cudaMemcpyKind kk = kind == 0 ? cudaMemcpyHostToDevice : kind == 1 ? cudaMemcpyDeviceToHost : cudaMemcpyDefault;
    B200_CUDA_TRY(cudaMemcpyAsync(dst, src, nbytes, kk, ctx->stream));
    return B200_OK;
}

// ---- pinned host memory and events (what the shell needs to report caps.host_buffer / caps.events / caps.async) --------------
int b200_host_malloc(size_t nbytes, void ** out) {
    if (!out) return B200_E_ARG;
    *out = nullptr;
    const cudaError_t err = cudaMallocHost(out, nbytes);
    if (err != cudaSuccess) {
        cudaGetLastError();
        b200_set_error("cudaMallocHost(%zu): %s", nbytes, cudaGetErrorString(err));
        return B200_E_ALLOC;
    }
    return B200_OK;
}

int b200_host_free(void * ptr) {
    if (!ptr) return B200_OK;
    B200_CUDA_TRY(cudaFreeHost(ptr));
    return B200_OK;
}

int b200_event_create(int device, void ** out) {
    if (!out) return B200_E_ARG;
    *out = nullptr;
    B200_CUDA_TRY(cudaSetDevice(device));
    cudaEvent_t ev;
    B200_CUDA_TRY(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    *out = (void *) ev;
    return B200_OK;
}

// an event that keeps its time stamp, and the device time between two of them (waits for `stop`)
int b200_event_create_timed(int device, void ** out) {
    if (!out) return B200_E_ARG;
    *out = nullptr;
    B200_CUDA_TRY(cudaSetDevice(device));
    cudaEvent_t ev;
    B200_CUDA_TRY(cudaEventCreate(&ev));
    *out = (void *) ev;
    return B200_OK;
}

int b200_event_elapsed_ms(void * start, void * stop, float * ms) {
    if (!start || !stop || !ms) return B200_E_ARG;
    B200_CUDA_TRY(cudaEventSynchronize((cudaEvent_t) stop));
    B200_CUDA_TRY(cudaEventElapsedTime(ms, (cudaEvent_t) start, (cudaEvent_t) stop));
    return B200_OK;
}

int b200_event_destroy(int device, void * event) {
    if (!event) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(device));
    B200_CUDA_TRY(cudaEventDestroy((cudaEvent_t) event));
    return B200_OK;
}

int b200_event_record(b200_ctx * ctx, void * event) {
    if (!ctx || !event) return B200_E_ARG;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaEventRecord((cudaEvent_t) event, ctx->stream));
    return B200_OK;
}

int b200_event_wait(b200_ctx * ctx, void * event) {
    if (!ctx || !event) return B200_E_ARG;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaStreamWaitEvent(ctx->stream, (cudaEvent_t) event, 0));
    ctx->n_pend = 16;       // whatever the event orders is not a kernel of this context: the next GEMV waits before it reads
    return B200_OK;
}

int b200_event_synchronize(void * event) {
    if (!event) return B200_E_ARG;
    B200_CUDA_TRY(cudaEventSynchronize((cudaEvent_t) event));
    return B200_OK;
}

// ---- host-buffer convenience ------------------------------------------------------------------------------------------
int b200_weights_upload(b200_ctx * ctx, int type, const void * host_rows, int64_t k, int64_t m, void ** dev_out) {
    if (!ctx || !host_rows || !dev_out || !is_weight_type(type) || k % b200_blck(type) != 0) {
        b200_set_error("weights_upload: bad argument");
        return B200_E_ARG;
    }
    const size_t nbytes = b200_row_size(type, k) * (size_t) m;
    int rc = b200_malloc(ctx->device, nbytes + 512, dev_out);
    if (rc) return rc;
    rc = b200_memcpy_h2d(ctx, *dev_out, host_rows, nbytes);
    if (rc) { b200_free(ctx->device, *dev_out); *dev_out = nullptr; }
    return rc;
}

void b200_weights_free(b200_ctx * ctx, void * dev) { if (ctx) b200_free(ctx->device, dev); }

static int grow(b200_ctx * ctx, void ** buf, size_t * have, size_t want) {
    if (want <= *have) return B200_OK;
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    if (*buf) { cudaFree(*buf); *buf = nullptr; *have = 0; }
    ctx->alloc_gen++;
    cudaError_t err = cudaMalloc(buf, want);
    if (err != cudaSuccess) { cudaGetLastError(); b200_set_error("staging allocation failed: %s", cudaGetErrorString(err)); return B200_E_ALLOC; }
    *have = want;
    return B200_OK;
}

// prefill row split: this rank's [n][rows] block goes into rows [row0, row0 + rows) of every rank's [n][dst_m] output with 16-byte stores
// (NVLink peer mappings).  Round 1 stored from the GEMM epilogue, one 4-byte lane store per value and peer, which made the step slower with
// every GPU added (10.0 -> 20.0 ms at N = 8); a separate pass moves the same bytes at link speed and leaves the GEMM as fast as unsplit.
struct push_dst { float * p[B200_MAX_PEERS]; };

__global__ void __launch_bounds__(256) push_shard_kernel(const float * __restrict__ ysh, const int64_t rows, const int64_t n, const push_dst D,
                                                         const int64_t dst_m, const int64_t row0, const int vec) {
    float * dst = D.p[blockIdx.y];
    if (vec) {
        const int64_t r4 = rows >> 2, total = r4 * n;
        for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t) gridDim.x * blockDim.x) {
            const int64_t col = i / r4, r = (i - col * r4) << 2;
            const uint4 v = *reinterpret_cast<const uint4 *>(ysh + col * rows + r);
            *reinterpret_cast<uint4 *>(dst + col * dst_m + row0 + r) = v;
        }
    } else {
        const int64_t total = rows * n;
        for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t) gridDim.x * blockDim.x) {
            const int64_t col = i / rows, r = i - col * rows;
            dst[col * dst_m + row0 + r] = ysh[i];
        }
    }
}

int b200_mul_mat_gather(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * b, const b200_tensor * d_shard,
                        int n_dst, float * const * dst, int64_t dst_m, int64_t dst_row0) {
    if (!ctx || !b || n_dst < 1 || !dst) { b200_set_error("mul_mat_gather: bad argument"); return B200_E_ARG; }
    // n <= 8: GEMV family, results stored to the peers from the epilogue; n > 8: tcgen05 GEMM into the shard, then one pass that pushes the
    // shard to every rank (option gather_push = 0: from the GEMM epilogue as in round 1); completion is the caller's barrier
    if (b->ne[2] * b->ne[3] != 1) { b200_set_error("mul_mat_gather: no batch dims"); return B200_E_UNSUPPORTED; }
    if (n_dst > B200_MAX_PEERS) { b200_set_error("mul_mat_gather: too many destinations"); return B200_E_ARG; }
    if (b->ne[1] > ctx->max_gemv_cols && ctx->gather_push && d_shard && a) {
        int rc = b200_mul_mat_ex(ctx, a, b, d_shard, 0, nullptr, 0, 0);
        if (rc) return rc;
        const int64_t rows = a->ne[1], n = b->ne[1];
        if (rows > 0 && n > 0) {
            push_dst D = {};
            bool aligned = ((uintptr_t) d_shard->data & 15) == 0 && rows % 4 == 0 && dst_m % 4 == 0 && dst_row0 % 4 == 0;
            for (int p = 0; p < n_dst; ++p) { D.p[p] = dst[p]; aligned = aligned && ((uintptr_t) dst[p] & 15) == 0; }
            const int64_t work = aligned ? (rows / 4) * n : rows * n, blocks = (work + 255) / 256, cap = (int64_t) ctx->sm_count * 8;
            const dim3 grid((unsigned) (blocks < cap ? blocks : cap), (unsigned) n_dst);
            push_shard_kernel<<<grid, 256, 0, ctx->stream>>>((const float *) d_shard->data, rows, n, D, dst_m, dst_row0, aligned ? 1 : 0);
            B200_CUDA_TRY(cudaGetLastError());
            ctx->launches++;
        }
        ctx->n_pend = 16;
        return B200_OK;
    }
    const int rc = b200_mul_mat_ex(ctx, a, b, d_shard, n_dst, dst, dst_m, dst_row0);
    // The in-flight write tracker knows d_shard only; the same kernel also wrote this rank's rows of the full-height outputs
    // (dst[p] through peer mappings), and other ranks write the rest.  A later GEMV that reads such a vector must not defer its
    // dependency wait: force the next launch to wait before it reads.  (Deferral contract: only writes made through this context
    // are tracked — see ggml_b200.h.)
    ctx->n_pend = 16;
    return rc;
}

// Several weight matrices against the same activations (q/k/v projections, gate/up): one launch instead of n_mats.
int b200_mul_mat_multi(b200_ctx * ctx, int n_mats, const b200_tensor * const * a, const b200_tensor * b, const b200_tensor * const * d) {
    if (!ctx || n_mats < 1 || !a || !b || !d) { b200_set_error("mul_mat_multi: bad argument"); return B200_E_ARG; }
    for (int j = 0; j < n_mats; ++j) {
        if (const char * why = check_mul_mat(a[j], b, d[j])) { b200_set_error("mul_mat_multi[%d]: %s", j, why); return B200_E_ARG; }
    }
    bool fusable = n_mats >= 2 && n_mats <= 3 && !ctx->force_generic && b->ne[2] * b->ne[3] == 1;
    for (int j = 0; fusable && j < n_mats; ++j) {
        fusable = a[j]->type == a[0]->type && a[j]->ne[0] == a[0]->ne[0] && a[j]->nb[1] == a[0]->nb[1] && a[j]->ne[2] * a[j]->ne[3] == 1 &&
                  a[j]->ne[1] > 0 && aligned16(a[j]->data, a[j]->nb[1], 0, 0);
    }
    const int64_t k = a[0]->ne[0], n = b->ne[1];
    if (fusable && n == 1 && a[0]->type == B200_TYPE_Q4_K && ctx->fuse_quant && aligned16(b->data, b->nb[1], 0, 0) &&
        b200_gemv_q4k_smem(1, k) <= ctx->smem_optin) {
        B200_CUDA_TRY(cudaSetDevice(ctx->device));
        mm_problem P = {};
        P.W = (const uint8_t *) a[0]->data; P.nb01 = a[0]->nb[1]; P.k = k;
        P.X = (const float *) b->data; P.nb11 = b->nb[1]; P.nb12 = b->nb[2]; P.nb13 = b->nb[3];
        P.ne11 = 1; P.ne12 = 1; P.ne13 = 1; P.r2 = 1; P.r3 = 1;
        P.Y = (float *) d[0]->data; P.mode = MM_PLAIN;
        P.n_mats = n_mats;
        int64_t start = a[0]->ne[1];
        P.m_end[0] = start; P.m_end[1] = INT64_MAX;
        for (int j = 1; j < n_mats; ++j) {
            P.Wv[j - 1] = (const uint8_t *) a[j]->data - start * P.nb01;
            P.Yv[j - 1] = (float *) d[j]->data - start;
            start += a[j]->ne[1];
            if (j == 1) P.m_end[1] = start;
        }
        if (n_mats == 2) { P.Wv[1] = P.Wv[0]; P.Yv[1] = P.Yv[0]; P.m_end[1] = start; }
        P.m = start;
        const bool late = ctx->n_pend + n_mats <= 16 && can_defer_wait(ctx, b, nullptr);
        for (int j = 0; j < n_mats; ++j) pending_push(ctx, d[j], !late && j == 0);
        return b200_launch_gemv_q4k(ctx, P, 1, 1, late);
    }
    if (fusable && n > 8 && ctx->use_tc) {
        for (int j = 0; j < n_mats; ++j) pending_push(ctx, d[j], j == 0);      // normal launches: everything before is complete
        const int rc = b200_mul_mat_multi_tc(ctx, n_mats, a, b, d);
        if (rc != B200_E_UNSUPPORTED) return rc;
    }
    for (int j = 0; j < n_mats; ++j) {
        const int rc = b200_mul_mat(ctx, a[j], b, d[j]);
        if (rc) return rc;
    }
    return B200_OK;
}

// ---- a list of matmuls in one call: what graph_compute sees when a cgraph holds several MUL_MAT nodes ---------------------------
// Runs of Q4_K n = 1 matmuls go through ONE persistent launch (mmvq_stream.cu); the job table, its activation groups and the
// dependencies between its jobs are resolved here, on the host, from the tensors' address ranges.
static bool stream_job_ok(const b200_ctx * ctx, const b200_tensor * a, const b200_tensor * b, const b200_tensor * d) {
    return a->type == B200_TYPE_Q4_K && b->ne[1] == 1 && b->ne[2] * b->ne[3] == 1 && a->ne[2] * a->ne[3] == 1 &&
           a->ne[0] % 256 == 0 && a->ne[0] / 256 <= 65535 && a->ne[1] >= 1 && a->ne[1] < ((int64_t) 1 << 31) && a->nb[1] < ((int64_t) 1 << 32) &&
           aligned16(a->data, a->nb[1], 0, 0) && aligned16(b->data, 0, 0, 0) && ((uintptr_t) d->data & 3) == 0 &&
           b200_gemv_stream_stages(ctx, a->ne[0]) >= 2;
}

// image offsets of the launch's activation groups + the buffers behind them (grow-only: like the other scratch, a list needs one
// un-captured call before it is captured into a CUDA graph)
static int stream_images(b200_ctx * ctx, sq_params & P) {
    size_t total = 0;
    for (int g = 0; g < P.n_groups; ++g) {
        P.gimg[g] = (uint32_t) total;
        total += ((size_t) P.jobs[P.gfirst[g]].nsb * 320 + 127) & ~(size_t) 127;       // k int8 + k / 32 float2 = 320 B per super-block
    }
    // the images' places in the shared-memory ring of 2 x the largest image: bump allocation with wrap-around; a place can be
    // reused once every consumer warp has finished the groups that lay there
    {
        P.stages = b200_gemv_stream_stages(ctx, P.kmax);
        P.xring  = (unsigned) b200_gemv_stream_xring(ctx, P.stages > 0 ? P.stages : 2);
        const size_t ring = P.xring;
        size_t off = 0;
        std::vector<int> live;                              // groups whose place has not been handed on, oldest first
        std::vector<size_t> lo((size_t) P.n_groups), hi((size_t) P.n_groups);
        for (int g = 0; g < P.n_groups; ++g) {
            const size_t sz = ((size_t) P.jobs[P.gfirst[g]].nsb * 320 + 127) & ~(size_t) 127;
            if (off + sz > ring) off = 0;
            lo[g] = off; hi[g] = off + sz;
            int wait = -1;
            std::vector<int> keep;
            for (int q : live) { if (lo[q] < hi[g] && lo[g] < hi[q]) { if (q > wait) wait = q; } else keep.push_back(q); }
            live.swap(keep); live.push_back(g);
            P.gsm[g] = (uint32_t) off; P.gwait[g] = (int16_t) wait;
            off += sz;
        }
    }
    if (total > ctx->sq_img_cap) {
        B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
        if (ctx->sq_img) { cudaFree(ctx->sq_img); ctx->sq_img = nullptr; ctx->sq_img_cap = 0; }
        ctx->alloc_gen++;
        const size_t want = total < ((size_t) 4 << 20) ? ((size_t) 4 << 20) : total;
        if (cudaMalloc((void **) &ctx->sq_img, want) != cudaSuccess) { cudaGetLastError(); b200_set_error("mul_mat_batch: activation image allocation failed"); return B200_E_ALLOC; }
        ctx->sq_img_cap = want;
    }
    if (!ctx->sq_xflag) {
        B200_CUDA_TRY(cudaMalloc((void **) &ctx->sq_xflag, (size_t) SQ_MAX_JOBS * 64 * sizeof(uint32_t)));
        B200_CUDA_TRY(cudaMemsetAsync(ctx->sq_xflag, 0, (size_t) SQ_MAX_JOBS * 64 * sizeof(uint32_t), ctx->stream));
    }
    P.ximg = ctx->sq_img; P.xflag = ctx->sq_xflag;
    return B200_OK;
}

static int stream_run(b200_ctx * ctx, int n, const b200_tensor * const * a, const b200_tensor * const * b, const b200_tensor * const * d) {
    if (!ctx->sq_done) {
        B200_CUDA_TRY(cudaMalloc((void **) &ctx->sq_done, (SQ_MAX_JOBS + 2) * sizeof(uint32_t)));
        B200_CUDA_TRY(cudaMemsetAsync(ctx->sq_done, 0, (SQ_MAX_JOBS + 2) * sizeof(uint32_t), ctx->stream));
    }
    static thread_local sq_params P;                        // ~26 KB: not on the stack of every caller
    memset(&P, 0, offsetof(sq_params, jobs));
    P.n_jobs = n; P.done = ctx->sq_done;
    const uint32_t G = (uint32_t) b200_gemv_stream_grid(ctx);
    std::vector<const char *> xlo((size_t) n), xhi((size_t) n), ylo((size_t) n), yhi((size_t) n);
    std::vector<int> xsrc((size_t) n, -1);                  // job whose exchange entries job j reads instead of X
    int group = -1, group_first = 0, group_size = 0;
    const char * xbox_lo = nullptr, * xbox_hi = nullptr, * ybox_lo = nullptr, * ybox_hi = nullptr;
    size_t ll_entries = 0;
    for (int j = 0; j < n; ++j) {
        sq_job & J = P.jobs[j];
        const uint32_t m = (uint32_t) a[j]->ne[1];
        J.W = (const uint8_t *) a[j]->data; J.X = (const float *) b[j]->data; J.Y = (float *) d[j]->data;
        J.nb01 = (uint32_t) a[j]->nb[1];
        J.base = ((m + 1) / 2) / G; J.rem = (uint16_t) (((m + 1) / 2) % G);          // row pairs
        J.nsb = (uint16_t) (a[j]->ne[0] / 256);
        J.dep = -1; J.xsrc = -1; J.flags = (uint16_t) ((m & 1u) << 1); J.ll_off = SQ_NO_LL;
        if (a[j]->ne[0] > P.kmax) P.kmax = (int) a[j]->ne[0];
        xlo[j] = (const char *) J.X; xhi[j] = xlo[j] + 4 * a[j]->ne[0];
        ylo[j] = (const char *) J.Y; yhi[j] = ylo[j] + 4 * a[j]->ne[1];
        if (xlo[j] < yhi[j] && ylo[j] < xhi[j]) { b200_set_error("mul_mat_batch[%d]: dst overlaps src1", j); return B200_E_ARG; }
        // Hazards against earlier jobs of the list, latest first:
        //  * X is EXACTLY the output of the latest job that wrote into X's range: the dependency travels in the data — that job
        //    also writes {value, epoch} words into the exchange buffer and this group polls them (no counters on that path);
        //  * anything else — X partly written by an earlier job, WAW on Y, WAR against a job that reads the plain tensor — waits
        //    for a completion counter.  Completion is cumulative (every job <= dep is complete), so the LAST such job is enough.
        int dep = -1, src = -1;
        bool raw_open = true;
        // (bounding boxes of everything earlier: the usual list has no overlaps at all and skips the scan)
        const bool maybe = j > 0 && ((xlo[j] < ybox_hi && ybox_lo < xhi[j]) || (ylo[j] < ybox_hi && ybox_lo < yhi[j]) || (ylo[j] < xbox_hi && xbox_lo < yhi[j]));
        for (int i = maybe ? j - 1 : -1; i >= 0 && (dep < 0 || raw_open); --i) {
            bool hazard = false;
            if (raw_open && xlo[j] < yhi[i] && ylo[i] < xhi[j]) {
                raw_open = false;                              // older writers of this range are superseded, or covered by the cumulative counter
                if (ylo[i] == xlo[j] && yhi[i] == xhi[j] && ctx->stream_ll) src = i; else hazard = true;
            }
            if (ylo[j] < yhi[i] && ylo[i] < yhi[j]) hazard = true;                                   // WAW
            if (xsrc[i] < 0 && ylo[j] < xhi[i] && xlo[i] < yhi[j]) hazard = true;                      // WAR (job i reads the tensor itself)
            if (hazard && dep < 0) dep = i;
        }
        xsrc[j] = src;
        if (j == 0 || xlo[j] < xbox_lo) xbox_lo = xlo[j];
        if (j == 0 || xhi[j] > xbox_hi) xbox_hi = xhi[j];
        if (j == 0 || ylo[j] < ybox_lo) ybox_lo = ylo[j];
        if (j == 0 || yhi[j] > ybox_hi) ybox_hi = yhi[j];
        const bool join = j > 0 && J.X == P.jobs[j - 1].X && J.nsb == P.jobs[j - 1].nsb && group_size < SQ_MAX_GROUP && dep < group_first &&
                          src == xsrc[group_first];
        if (join) {
            group_size++;
            if (dep > P.jobs[group_first].dep) P.jobs[group_first].dep = (int16_t) dep;
        } else {
            group++; group_first = j; group_size = 1;
            J.dep = (int16_t) dep; J.xsrc = (int16_t) src;
            P.gfirst[group] = (uint16_t) j;
            if (src >= 0 && P.jobs[src].ll_off == SQ_NO_LL) {
                P.jobs[src].ll_off = (uint32_t) ll_entries;
                ll_entries += ((size_t) a[src]->ne[1] + 3) & ~(size_t) 3;
                P.n_ll++;
            }
        }
        J.xg = (uint16_t) group;
    }
    P.n_groups = group + 1;
    // A list that is mostly a dependency CHAIN (every group waits for the one before it) gains nothing from streaming across matrix
    // boundaries, and a dependency inside the persistent kernel costs more than a kernel boundary (measured on the Llama-3-8B step
    // wired like the model: 1.64 ms in one launch, 1.20 ms with one PDL launch per group): such lists keep the per-launch kernels.
    {
        int waiting = 0;
        for (int g = 0; g < P.n_groups; ++g) waiting += P.jobs[P.gfirst[g]].dep >= 0 || P.jobs[P.gfirst[g]].xsrc >= 0;
        if (!ctx->stream_chain && 2 * waiting > P.n_groups) return 1;      // not taken
    }
    for (int j = 0; j < n; ++j) {
        if (P.jobs[j].dep >= 0 && !(P.jobs[P.jobs[j].dep].flags & 1)) { P.jobs[P.jobs[j].dep].flags |= 1; P.n_flagged++; }
    }
    if (ll_entries > ctx->sq_ll_cap) {
        // grow-only; like the scratch of the other paths this needs one un-captured call before a CUDA-graph capture
        B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
        if (ctx->sq_ll) { cudaFree(ctx->sq_ll); ctx->sq_ll = nullptr; ctx->sq_ll_cap = 0; }
        ctx->alloc_gen++;
        const size_t want = ll_entries < ((size_t) 1 << 20) ? ((size_t) 1 << 20) : ll_entries;
        if (cudaMalloc((void **) &ctx->sq_ll, want * 8) != cudaSuccess) { cudaGetLastError(); b200_set_error("mul_mat_batch: exchange buffer allocation failed"); return B200_E_ALLOC; }
        B200_CUDA_TRY(cudaMemsetAsync(ctx->sq_ll, 0, want * 8, ctx->stream));
        ctx->sq_ll_cap = want;
    }
    P.ll = ctx->sq_ll;
    if (const int rc = stream_images(ctx, P)) return rc;
    if (ctx->stream_trace) {
        if (!ctx->sq_trace) B200_CUDA_TRY(cudaMalloc((void **) &ctx->sq_trace, 2 * SQ_MAX_JOBS * 4 * sizeof(unsigned long long)));
        P.trace = ctx->sq_trace;
    }
    // everything before this launch is complete when its CTAs pass griddepcontrol.wait (no deferred wait here)
    ctx->n_pend = 0;
    return b200_launch_gemv_stream(ctx, P);
}

// debugging aid, deliberately not in ggml_b200.h: copies the coordinator time stamps of the last traced launch (tools/stream_trace.py)
extern "C" __attribute__((visibility("default"))) int b200_dbg_stream_trace(b200_ctx * ctx, unsigned long long * out, int n_words) {
    if (!ctx || !ctx->sq_trace || !out || n_words > 2 * SQ_MAX_JOBS * 4) return B200_E_ARG;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    B200_CUDA_TRY(cudaMemcpy(out, ctx->sq_trace, (size_t) n_words * 8, cudaMemcpyDeviceToHost));
    return B200_OK;
}

// ---- row split of a whole list, one process per GPU ---------------------------------------------------------------------------
extern "C" int b200_stream_comm_init(b200_ctx * ctx, int rank, int world, void * const * bufs, int64_t slot_entries) {
    if (!ctx || world < 0 || world > B200_MAX_PEERS || (world > 0 && (!bufs || rank < 0 || rank >= world || slot_entries < 4 || slot_entries % 4 != 0))) {
        b200_set_error("stream_comm_init: bad argument (slot_entries must be a positive multiple of 4)"); return B200_E_ARG;
    }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    ctx->sq_world = world; ctx->sq_rank = rank; ctx->sq_peer_cap = (size_t) slot_entries;
    for (int p = 0; p < world; ++p) {
        if (!bufs[p] || ((uintptr_t) bufs[p] & 15)) { b200_set_error("stream_comm_init: buffers must be 16-byte aligned"); ctx->sq_world = 0; return B200_E_ARG; }
        ctx->sq_peer[p] = (unsigned long long *) bufs[p];
    }
    if (!ctx->sq_done) {
        B200_CUDA_TRY(cudaMalloc((void **) &ctx->sq_done, (SQ_MAX_JOBS + 2) * sizeof(uint32_t)));
    }
    // all ranks start counting their launches from zero together
    B200_CUDA_TRY(cudaMemsetAsync(ctx->sq_done, 0, (SQ_MAX_JOBS + 2) * sizeof(uint32_t), ctx->stream));
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return B200_OK;
}

// emu_world > 0 (tests): all ranks inside one launch on this GPU — a[j] are the FULL matrices, every rank r works on rows
// [r m / world, (r + 1) m / world) and on its own copy of the vectors (emu_stride bytes apart), see gemv_q4k_stream<2>
static int gather_impl(b200_ctx * ctx, int n_jobs, const b200_tensor * const * a, const b200_tensor * const * b, const b200_tensor * const * d_full,
                       const int64_t * row0, int emu_world, size_t emu_stride, unsigned long long * emu_ll, size_t emu_cap) {
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    const int world = emu_world > 0 ? emu_world : ctx->sq_world;
    const uint32_t G = emu_world > 0 ? (uint32_t) (ctx->sm_count / emu_world) : (uint32_t) b200_gemv_stream_grid(ctx);
    const size_t cap = emu_world > 0 ? emu_cap : ctx->sq_peer_cap;
    // emulation: per-rank counters, flags and images in one zero-initialised block [done | flags | images]
    const size_t emu_done_b = (size_t) 8 * (SQ_MAX_JOBS + 2) * 4, emu_flag_b = (size_t) 8 * SQ_MAX_JOBS * 64 * 4, emu_img_b = (size_t) 4 << 20;
    if (emu_world > 0 && !ctx->sq_emu) {
        B200_CUDA_TRY(cudaMalloc((void **) &ctx->sq_emu, emu_done_b + emu_flag_b + 8 * emu_img_b));
        B200_CUDA_TRY(cudaMemsetAsync(ctx->sq_emu, 0, emu_done_b + emu_flag_b + 8 * emu_img_b, ctx->stream));
    }
    static thread_local sq_params P;
    for (int j0 = 0; j0 < n_jobs; ) {
        // one launch: up to SQ_MAX_JOBS jobs whose exchange entries fit a slot
        memset(&P, 0, offsetof(sq_params, jobs));
        P.done = ctx->sq_done;
        P.world = world; P.rank = emu_world > 0 ? 0 : ctx->sq_rank; P.ll_cap = cap;
        for (int p = 0; p < world; ++p) P.ll_peer[p] = emu_world > 0 ? emu_ll + (size_t) p * 2 * cap : ctx->sq_peer[p];
        P.ll = P.ll_peer[P.rank];
        if (emu_world > 0) { P.emu_stride = emu_stride; P.emu_img_stride = emu_img_b; }
        size_t entries = 0;
        int n = 0, group = -1, group_first = 0, group_size = 0;
        std::vector<int> xsrc;
        for (int j = j0; j < n_jobs && n < SQ_MAX_JOBS; ++j, ++n) {
            const b200_tensor * A = a[j], * B = b[j], * D = d_full[j];
            if (!A || !B || !D || A->type != B200_TYPE_Q4_K || A->ne[0] % 256 != 0 || A->ne[0] / 256 > 65535 || A->ne[2] * A->ne[3] != 1 || A->ne[1] < 0 ||
                (A->ne[1] > 0 && !aligned16(A->data, A->nb[1], 0, 0)) || A->nb[1] >= ((int64_t) 1 << 32) || A->nb[0] != SZ_Q4_K ||
                B->type != B200_TYPE_F32 || B->ne[0] != A->ne[0] || B->ne[1] * B->ne[2] * B->ne[3] != 1 || B->nb[0] != 4 || !aligned16(B->data, 0, 0, 0) ||
                D->type != B200_TYPE_F32 || D->nb[0] != 4 || D->ne[1] * D->ne[2] * D->ne[3] != 1 || (emu_world == 0 && (row0[j] < 0 || row0[j] + A->ne[1] > D->ne[0])) ||
                D->ne[0] >= ((int64_t) 1 << 31) || b200_gemv_stream_stages(ctx, A->ne[0]) < 2) {
                b200_set_error("mul_mat_batch_gather[%d]: Q4_K row shards (16-byte aligned rows, k %% 256 == 0) x one f32 column -> one full-height f32 column", j);
                return B200_E_UNSUPPORTED;
            }
            const size_t need = ((size_t) D->ne[0] + 3) & ~(size_t) 3;
            if (entries + need > cap) {
                if (n == 0) { b200_set_error("mul_mat_batch_gather[%d]: %lld results exceed the exchange slot (%zu entries)", j, (long long) D->ne[0], cap); return B200_E_ARG; }
                break;
            }
            if (emu_world > 0 && (A->ne[1] % emu_world != 0 || A->ne[1] != D->ne[0])) { b200_set_error("emulated row split: m must be a multiple of the world size"); return B200_E_ARG; }
            const int64_t kmax2 = A->ne[0] > P.kmax ? A->ne[0] : P.kmax;
            if (b200_gemv_stream_stages(ctx, kmax2) < 2) break;
            sq_job & J = P.jobs[n];
            const uint32_t m = (uint32_t) (emu_world > 0 ? A->ne[1] / emu_world : A->ne[1]);
            J.W = (const uint8_t *) A->data; J.X = (const float *) B->data; J.Y = (float *) D->data;
            J.nb01 = (uint32_t) A->nb[1]; J.base = ((m + 1) / 2) / G; J.rem = (uint16_t) (((m + 1) / 2) % G); J.nsb = (uint16_t) (A->ne[0] / 256);
            J.row0 = (uint32_t) (emu_world > 0 ? 0 : row0[j]); J.m_full = (uint32_t) D->ne[0];
            J.dep = -1; J.xsrc = -1; J.flags = (uint16_t) ((m & 1u) << 1); J.ll_off = (uint32_t) entries;
            // (the two results of a row pair travel as one aligned 16-byte store: shards start at even rows and, emulated, have an even height)
            if ((J.row0 & 1u) || (emu_world > 0 && (m & 1u))) { b200_set_error("mul_mat_batch_gather[%d]: row shards must start at an even row", j); return B200_E_UNSUPPORTED; }
            entries += need;
            P.kmax = (int) kmax2;
            // X is exactly the full-height output of an earlier job of this launch: polled in the exchange buffer.  Any other overlap
            // between X and an earlier output cannot be ordered across ranks inside the launch: the list is cut there.
            const char * xlo = (const char *) J.X, * xhi = xlo + 4 * A->ne[0];
            int src = -1; bool cut = false;
            for (int i = n - 1; i >= 0; --i) {
                const char * ylo = (const char *) P.jobs[i].Y, * yhi = ylo + 4 * (size_t) P.jobs[i].m_full;
                if (xlo < yhi && ylo < xhi) { if (ylo == xlo && yhi == xhi) src = i; else cut = true; break; }
            }
            if (cut) { if (n == 0) { b200_set_error("mul_mat_batch_gather[%d]: src1 overlaps dst", j); return B200_E_ARG; } entries -= need; break; }
            {
                const char * ylo = (const char *) J.Y, * yhi = ylo + 4 * (size_t) J.m_full;
                if (xlo < yhi && ylo < xhi) { b200_set_error("mul_mat_batch_gather[%d]: dst overlaps src1", j); return B200_E_ARG; }
            }
            xsrc.push_back(src);
            const bool join = n > 0 && J.X == P.jobs[n - 1].X && J.nsb == P.jobs[n - 1].nsb && group_size < SQ_MAX_GROUP && src == xsrc[group_first];
            if (join) group_size++;
            else { group++; group_first = n; group_size = 1; J.xsrc = (int16_t) src; P.gfirst[group] = (uint16_t) n; }
            J.xg = (uint16_t) group;
        }
        P.n_jobs = n; P.n_groups = group + 1; P.n_ll = n;
        if (const int rc = stream_images(ctx, P)) return rc;
        if (emu_world > 0) {
            if ((size_t) P.gimg[P.n_groups - 1] + (size_t) P.jobs[P.gfirst[P.n_groups - 1]].nsb * 320 > emu_img_b) { b200_set_error("emulated row split: images exceed 4 MB"); return B200_E_ARG; }
            P.done  = (uint32_t *) ctx->sq_emu;
            P.xflag = (uint32_t *) ((char *) ctx->sq_emu + emu_done_b);
            P.ximg  = (uint8_t *) ctx->sq_emu + emu_done_b + emu_flag_b;
        }
        ctx->n_pend = 0;
        const int rc = b200_launch_gemv_stream(ctx, P);
        if (rc) return rc;
        j0 += n;
    }
    return B200_OK;
}

extern "C" int b200_mul_mat_batch_gather(b200_ctx * ctx, int n_jobs, const b200_tensor * const * a, const b200_tensor * const * b,
                                         const b200_tensor * const * d_full, const int64_t * row0) {
    if (!ctx || n_jobs < 0 || (n_jobs > 0 && (!a || !b || !d_full || !row0))) { b200_set_error("mul_mat_batch_gather: bad argument"); return B200_E_ARG; }
    if (ctx->sq_world < 2) { b200_set_error("mul_mat_batch_gather: b200_stream_comm_init was not called (world >= 2)"); return B200_E_ARG; }
    return gather_impl(ctx, n_jobs, a, b, d_full, row0, 0, 0, nullptr, 0);
}

// testing aid, deliberately not in ggml_b200.h: b200_mul_mat_batch_gather with `world` ranks emulated inside ONE cooperative launch
// on this GPU.  a[j] = the full matrices; b[j] / d_full[j] = rank 0's vectors, rank r's copies lie r * arena_stride bytes further;
// ll = zero-initialised exchange memory of world * 2 * slot_entries * 8 bytes.
extern "C" __attribute__((visibility("default"))) int b200_dbg_batch_gather_emulated(b200_ctx * ctx, int world, size_t arena_stride, void * ll, int64_t slot_entries,
        int n_jobs, const b200_tensor * const * a, const b200_tensor * const * b, const b200_tensor * const * d_full) {
    if (!ctx || world < 2 || world > B200_MAX_PEERS || !ll || slot_entries < 4 || slot_entries % 4 != 0 || arena_stride % 16 != 0 || arena_stride == 0 ||
        n_jobs < 0 || (n_jobs > 0 && (!a || !b || !d_full))) { b200_set_error("dbg_batch_gather_emulated: bad argument"); return B200_E_ARG; }
    return gather_impl(ctx, n_jobs, a, b, d_full, nullptr, world, arena_stride, (unsigned long long *) ll, (size_t) slot_entries);
}

extern "C" int b200_mul_mat_batch(b200_ctx * ctx, int n_jobs, const b200_tensor * const * a, const b200_tensor * const * b,
                                  const b200_tensor * const * d) {
    if (!ctx || n_jobs < 0 || (n_jobs > 0 && (!a || !b || !d))) { b200_set_error("mul_mat_batch: bad argument"); return B200_E_ARG; }
    for (int j = 0; j < n_jobs; ++j) {
        if (const char * why = check_mul_mat(a[j], b[j], d[j])) { b200_set_error("mul_mat_batch[%d]: %s", j, why); return B200_E_ARG; }
    }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    for (int j = 0; j < n_jobs; ) {
        // a run of jobs the persistent kernel takes
        int e = j;
        if (ctx->use_stream && !ctx->force_generic && ctx->fuse_quant) {
            int64_t kmax = 0;
            while (e < n_jobs && e - j < SQ_MAX_JOBS && a[e]->ne[1] > 0 && stream_job_ok(ctx, a[e], b[e], d[e])) {
                const int64_t k2 = a[e]->ne[0] > kmax ? a[e]->ne[0] : kmax;
                if (b200_gemv_stream_stages(ctx, k2) < 2) break;
                kmax = k2; ++e;
            }
        }
        // (a persistent launch costs ~8 us before its first row — launch, barriers, ring prologue, the first group's image — against
        // ~4 us for a per-matmul kernel: measured break-even at one decoder layer's 7 matmuls, so shorter runs keep the per-launch kernels)
        if (e - j >= (ctx->stream_min_jobs > 1 ? ctx->stream_min_jobs : 1)) {
            const int rc = stream_run(ctx, e - j, a + j, b + j, d + j);
            if (rc < 0) return rc;
            if (rc == 0) { j = e; continue; }
            // (rc == 1: a dependency chain — the per-launch kernels below take it)
        }
        // otherwise: the matrices that share their activations in one launch (q / k / v, gate / up), everything else one by one
        e = j + 1;
        while (e < n_jobs && e - j < 3 && b[e]->data == b[j]->data && b[e]->ne[1] == b[j]->ne[1] && b[e]->nb[1] == b[j]->nb[1] &&
               b[e]->ne[2] * b[e]->ne[3] == 1 && b[j]->ne[2] * b[j]->ne[3] == 1 && a[e]->type == a[j]->type && a[e]->ne[0] == a[j]->ne[0]) ++e;
        const int rc = e - j == 1 ? b200_mul_mat(ctx, a[j], b[j], d[j]) : b200_mul_mat_multi(ctx, e - j, a + j, b[j], d + j);
        if (rc) return rc;
        j = e;
    }
    return B200_OK;
}

// ---- row split, LL exchange ------------------------------------------------------------------------------------------
int b200_gather_ll_init(b200_ctx * ctx, int rank, int world, void * const * bufs, int64_t slot_entries) {
    if (!ctx || world < 0 || world > B200_MAX_PEERS || (world > 0 && (!bufs || rank < 0 || rank >= world || slot_entries < 1))) {
        b200_set_error("gather_ll_init: bad argument"); return B200_E_ARG;
    }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    ctx->ll_world = world; ctx->ll_rank = rank; ctx->ll_cap = slot_entries;
    if (world > 0) {
        if (!ctx->ll_state) B200_CUDA_TRY(cudaMalloc((void **) &ctx->ll_state, 2 * sizeof(uint32_t)));
        B200_CUDA_TRY(cudaMemset(ctx->ll_state, 0, 2 * sizeof(uint32_t)));
    }
    for (int p = 0; p < world; ++p) {
        if (((uintptr_t) bufs[p]) & 7) { b200_set_error("gather_ll_init: buffers must be 8-byte aligned"); ctx->ll_world = 0; return B200_E_ARG; }
        ctx->ll_bufs[p] = (uint2 *) bufs[p];
    }
    return B200_OK;
}

int b200_mul_mat_gather_ll(b200_ctx * ctx, int n_mats, const b200_tensor * const * a, const b200_tensor * b, const b200_tensor * const * d_full,
                           const int64_t * row0) {
    if (!ctx || n_mats < 1 || n_mats > 3 || !a || !b || !d_full || !row0) { b200_set_error("mul_mat_gather_ll: bad argument"); return B200_E_ARG; }
    if (ctx->ll_world < 1) { b200_set_error("mul_mat_gather_ll: b200_gather_ll_init was not called"); return B200_E_ARG; }
    const int64_t k = a[0]->ne[0];
    if (b->type != B200_TYPE_F32 || b->ne[0] != k || b->ne[1] != 1 || b->ne[2] * b->ne[3] != 1 || b->nb[0] != 4 || !b->data) {
        b200_set_error("mul_mat_gather_ll: activations must be one contiguous f32 column of k values"); return B200_E_UNSUPPORTED;
    }
    int64_t total_full = 0, shard_rows = 0;
    const char * xlo, * xhi;
    tensor_extent(b, &xlo, &xhi);
    for (int j = 0; j < n_mats; ++j) {
        const b200_tensor * A = a[j], * D = d_full[j];
        if (!A || !D || (!A->data && A->ne[1] > 0) || !D->data) { b200_set_error("mul_mat_gather_ll[%d]: null tensor", j); return B200_E_ARG; }
        // (a rank may own no rows of a small matrix: ne[1] == 0)
        if (A->type != B200_TYPE_Q4_K || A->ne[0] != k || k % 256 != 0 || A->ne[2] * A->ne[3] != 1 || A->nb[1] != a[0]->nb[1] || A->ne[1] < 0 ||
            !aligned16(A->data, A->nb[1], 0, 0)) {
            b200_set_error("mul_mat_gather_ll[%d]: Q4_K shards of equal row length, 16-byte aligned, no batch dims", j); return B200_E_UNSUPPORTED;
        }
        if (D->type != B200_TYPE_F32 || D->nb[0] != 4 || D->ne[1] * D->ne[2] * D->ne[3] != 1 || row0[j] < 0 || row0[j] + A->ne[1] > D->ne[0]) {
            b200_set_error("mul_mat_gather_ll[%d]: d_full must be one contiguous f32 column that holds rows [row0, row0 + shard rows)", j); return B200_E_ARG;
        }
        const char * lo, * hi;
        tensor_extent(D, &lo, &hi);
        if (lo < xhi && xlo < hi) { b200_set_error("mul_mat_gather_ll[%d]: d_full overlaps the activations", j); return B200_E_ARG; }
        total_full += D->ne[0]; shard_rows += A->ne[1];
    }
    if (total_full > ctx->ll_cap) { b200_set_error("mul_mat_gather_ll: %lld results exceed the exchange slot (%lld)", (long long) total_full, (long long) ctx->ll_cap); return B200_E_ARG; }
    if (!ctx->fuse_quant || !aligned16(b->data, 0, 0, 0) || b200_gemv_q4k_smem(1, k) > ctx->smem_optin) {
        b200_set_error("mul_mat_gather_ll: needs the fused Q4_K GEMV (k = %lld)", (long long) k); return B200_E_UNSUPPORTED;
    }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    mm_problem P = {};
    P.W = (const uint8_t *) a[0]->data; P.nb01 = a[0]->nb[1]; P.k = k;
    P.X = (const float *) b->data; P.nb11 = b->nb[1]; P.nb12 = b->nb[2]; P.nb13 = b->nb[3];
    P.ne11 = 1; P.ne12 = 1; P.ne13 = 1; P.r2 = 1; P.r3 = 1;
    P.Y = (float *) d_full[0]->data; P.mode = MM_PLAIN;
    P.n_mats = n_mats; P.m = shard_rows;
    P.m_end[0] = P.m_end[1] = INT64_MAX;
    // rows are numbered through the shards; entry of (matrix j, full row r) in the exchange slot = (sum of full heights before j) + r
    int64_t start = 0, full_before = 0, m_full[3] = {0, 0, 0};
    float * Y[3] = {nullptr, nullptr, nullptr};
    for (int j = 0; j < n_mats; ++j) {
        if (j > 0) { P.Wv[j - 1] = (const uint8_t *) a[j]->data - start * P.nb01; P.m_end[j - 1] = start; }
        P.ll_off[j] = full_before + row0[j] - start;
        start += a[j]->ne[1]; full_before += d_full[j]->ne[0];
        m_full[j] = d_full[j]->ne[0]; Y[j] = (float *) d_full[j]->data;
    }
    // (boundaries of absent matrices stay at INT64_MAX, so their Wv / ll_off entries are never selected)
    P.ll_world = ctx->ll_world; P.ll_cap = ctx->ll_cap; P.ll_state = ctx->ll_state;
    for (int p = 0; p < ctx->ll_world; ++p) P.ll_dst[p] = ctx->ll_bufs[p];
    // the GEMV waits for everything before it at its start (no deferred wait): the receiver below may become resident while the
    // GEMV runs and writes d_full as soon as the values arrive, which must not overtake older readers of d_full
    for (int j = 0; j < n_mats; ++j) pending_push(ctx, d_full[j], j == 0);
    int rc = b200_launch_gemv_q4k(ctx, P, 1, 1, false);
    if (rc) return rc;
    return b200_launch_ll_unpack(ctx, ctx->ll_bufs[ctx->ll_rank], ctx->ll_cap, ctx->ll_state, n_mats, m_full, Y);
}

int b200_gather_sync_init(b200_ctx * ctx, int rank, int world, uint32_t * const * flags) {
    if (!ctx || world < 0 || world > B200_MAX_PEERS || (world > 0 && (!flags || rank < 0 || rank >= world))) {
        b200_set_error("gather_sync_init: bad argument"); return B200_E_ARG;
    }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    ctx->gs_world = world; ctx->gs_rank = rank;
    for (int p = 0; p < world; ++p) ctx->gs_flags[p] = flags[p];
    if (world > 0) {
        if (!ctx->gs_state) B200_CUDA_TRY(cudaMalloc((void **) &ctx->gs_state, 2 * sizeof(uint32_t)));
        B200_CUDA_TRY(cudaMemset(ctx->gs_state, 0, 2 * sizeof(uint32_t)));
    }
    return B200_OK;
}

int b200_mul_mat_host(b200_ctx * ctx, int type, const void * w_dev, int64_t m, int64_t k, const float * x_host, int64_t n,
                      float * y_host) {
    if (!ctx || !w_dev || !x_host || !y_host || !is_weight_type(type)) { b200_set_error("mul_mat_host: bad argument"); return B200_E_ARG; }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    const size_t xb = (size_t) k * n * 4, yb = (size_t) m * n * 4;
    int rc = grow(ctx, &ctx->dev_x, &ctx->dev_x_size, xb); if (rc) return rc;
    rc     = grow(ctx, &ctx->dev_y, &ctx->dev_y_size, yb); if (rc) return rc;
    B200_CUDA_TRY(cudaMemcpyAsync(ctx->dev_x, x_host, xb, cudaMemcpyHostToDevice, ctx->stream));
    b200_tensor a = {}, b = {}, d = {};
    a.data = (void *) w_dev; a.type = type; a.ne[0] = k; a.ne[1] = m; a.ne[2] = a.ne[3] = 1;
    a.nb[0] = (int64_t) b200_tsize(type); a.nb[1] = (int64_t) b200_row_size(type, k); a.nb[2] = a.nb[3] = a.nb[1] * m;
    b.data = ctx->dev_x; b.type = B200_TYPE_F32; b.ne[0] = k; b.ne[1] = n; b.ne[2] = b.ne[3] = 1;
    b.nb[0] = 4; b.nb[1] = k * 4; b.nb[2] = b.nb[3] = k * 4 * n;
    d.data = ctx->dev_y; d.type = B200_TYPE_F32; d.ne[0] = m; d.ne[1] = n; d.ne[2] = d.ne[3] = 1;
    d.nb[0] = 4; d.nb[1] = m * 4; d.nb[2] = d.nb[3] = m * 4 * n;
    rc = b200_mul_mat(ctx, &a, &b, &d);
    if (rc) return rc;
    B200_CUDA_TRY(cudaMemcpyAsync(y_host, ctx->dev_y, yb, cudaMemcpyDeviceToHost, ctx->stream));
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return B200_OK;
}

// A whole step's worth of matmuls with host activations / results: uploads (contiguous host ranges merged into one copy),
// all launches back to back on the context's stream, downloads (merged likewise), ONE synchronisation.  This is what a
// graph_compute over many MUL_MAT nodes amounts to for the scheduler (ggml-backend.cpp: copy inputs, compute, copy outputs).
int b200_mul_mat_host_batch(b200_ctx * ctx, const b200_host_mm * jobs, int n_jobs) {
    if (!ctx || (!jobs && n_jobs > 0) || n_jobs < 0) { b200_set_error("mul_mat_host_batch: bad argument"); return B200_E_ARG; }
    if (n_jobs == 0) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    // device placement.  Jobs that follow one another and read the SAME host activations (wq / wk / wv, ffn_gate / ffn_up of a
    // decode step) share one uploaded copy — and, below, one launch.
    std::vector<size_t> xdev((size_t) n_jobs), ydev((size_t) n_jobs);
    struct span { const char * host; size_t dev, bytes; };
    std::vector<span> uploads;
    size_t xb = 0, yb = 0;
    for (int j = 0; j < n_jobs; ++j) {
        const b200_host_mm & J = jobs[j];
        if (!J.w_dev || !J.x_host || !J.y_host || !is_weight_type(J.type) || J.m <= 0 || J.k <= 0 || J.n <= 0) {
            b200_set_error("mul_mat_host_batch: bad job %d", j); return B200_E_ARG;
        }
        const size_t bytes = (size_t) J.k * J.n * 4;
        if (j > 0 && J.x_host == jobs[j - 1].x_host && J.k == jobs[j - 1].k && J.n == jobs[j - 1].n) {
            xdev[j] = xdev[j - 1];
        } else {
            xdev[j] = xb;
            // host vectors that follow one another in memory are moved with a single copy
            if (!uploads.empty() && uploads.back().host + uploads.back().bytes == (const char *) J.x_host &&
                uploads.back().dev + uploads.back().bytes == xb) uploads.back().bytes += bytes;
            else uploads.push_back({(const char *) J.x_host, xb, bytes});
            xb += align_up(bytes, 256);
        }
        ydev[j] = yb;
        yb += align_up((size_t) J.m * J.n * 4, 256);
    }
    int rc = grow(ctx, &ctx->dev_x, &ctx->dev_x_size, xb); if (rc) return rc;
    rc     = grow(ctx, &ctx->dev_y, &ctx->dev_y_size, yb); if (rc) return rc;
    char * dx = (char *) ctx->dev_x, * dy = (char *) ctx->dev_y;
    for (const span & u : uploads) B200_CUDA_TRY(cudaMemcpyAsync(dx + u.dev, u.host, u.bytes, cudaMemcpyHostToDevice, ctx->stream));

    // launches
    auto describe = [&](int i, b200_tensor & a, b200_tensor & b, b200_tensor & d) {
        const b200_host_mm & J = jobs[i];
        a = {}; b = {}; d = {};
        a.data = (void *) J.w_dev; a.type = J.type; a.ne[0] = J.k; a.ne[1] = J.m; a.ne[2] = a.ne[3] = 1;
        a.nb[0] = (int64_t) b200_tsize(J.type); a.nb[1] = (int64_t) b200_row_size(J.type, J.k); a.nb[2] = a.nb[3] = a.nb[1] * J.m;
        b.data = dx + xdev[i]; b.type = B200_TYPE_F32; b.ne[0] = J.k; b.ne[1] = J.n; b.ne[2] = b.ne[3] = 1;
        b.nb[0] = 4; b.nb[1] = J.k * 4; b.nb[2] = b.nb[3] = J.k * 4 * J.n;
        d.data = dy + ydev[i]; d.type = B200_TYPE_F32; d.ne[0] = J.m; d.ne[1] = J.n; d.ne[2] = d.ne[3] = 1;
        d.nb[0] = 4; d.nb[1] = J.m * 4; d.nb[2] = d.nb[3] = J.m * 4 * J.n;
    };
    {
        // the whole list in one b200_mul_mat_batch: runs of Q4_K n = 1 jobs become one persistent launch, the rest is grouped there
        std::vector<b200_tensor> ta((size_t) n_jobs), tb((size_t) n_jobs), td((size_t) n_jobs);
        std::vector<const b200_tensor *> pa((size_t) n_jobs), pb((size_t) n_jobs), pd((size_t) n_jobs);
        for (int i = 0; i < n_jobs; ++i) { describe(i, ta[i], tb[i], td[i]); pa[i] = &ta[i]; pb[i] = &tb[i]; pd[i] = &td[i]; }
        rc = b200_mul_mat_batch(ctx, n_jobs, pa.data(), pb.data(), pd.data());
        if (rc) return rc;
    }
    // downloads: merge jobs whose results are adjacent on the host AND on the device (sizes that are multiples of 256 B)
    for (int j = 0; j < n_jobs; ) {
        char * h0 = (char *) jobs[j].y_host;
        size_t run = 0;
        int e = j;
        while (e < n_jobs && (char *) jobs[e].y_host == h0 + run && ydev[e] == ydev[j] + run) { run += (size_t) jobs[e].m * jobs[e].n * 4; ++e; }
        B200_CUDA_TRY(cudaMemcpyAsync(h0, dy + ydev[j], run, cudaMemcpyDeviceToHost, ctx->stream));
        j = e;
    }
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return B200_OK;
}

}  // extern "C"
