// common.cuh — shared device/host helpers for the B200 quantized MUL_MAT path.
//
// Block layouts are the reference's on-disk / in-memory formats and are parsed bit-for-bit
// (/root/reference/ggml/src/ggml-common.h:184-366):
//   block_q4_0 {fp16 d; u8 qs[16]}                                   18 B / 32 weights
//   block_q8_0 {fp16 d; i8 qs[32]}                                   34 B / 32
//   block_q8_1 {fp16 d; fp16 s; i8 qs[32]}                           36 B / 32   (activations)
//   block_q4_K {fp16 d; fp16 dmin; u8 scales[12]; u8 qs[128]}       144 B / 256
//   block_q6_K {u8 ql[128]; u8 qh[64]; i8 scales[16]; fp16 d}       210 B / 256
#pragma once

#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stddef.h>

#include "../../include/ggml_b200.h"

#define B200_WARP 32
#define B200_MAX_PEERS 8

// ---- block geometry -----------------------------------------------------------------------------
#define QK_SMALL 32
#define QK_K     256
#define SZ_Q4_0  18
#define SZ_Q8_0  34
#define SZ_Q8_1  36
#define SZ_Q4_K  144
#define SZ_Q6_K  210
#define SZ_Q4_1  20
#define SZ_Q5_0  22
#define SZ_Q5_1  24
#define SZ_Q5_K  176
#define SZ_Q2_K  84
#define SZ_Q3_K  110
#define SZ_IQ4_NL 18
#define SZ_IQ4_XS 136
#define SZ_MXFP4 17
#define SZ_NVFP4 36

__host__ __device__ constexpr int b200_blck(int type) {
    return (type == B200_TYPE_Q4_K || type == B200_TYPE_Q6_K || type == B200_TYPE_Q5_K || type == B200_TYPE_Q2_K || type == B200_TYPE_Q3_K || type == B200_TYPE_IQ4_XS) ? QK_K
         : (type == B200_TYPE_Q4_0 || type == B200_TYPE_Q8_0 || type == B200_TYPE_Q8_1 || type == B200_TYPE_Q4_1 || type == B200_TYPE_Q5_0 || type == B200_TYPE_Q5_1 ||
            type == B200_TYPE_IQ4_NL || type == B200_TYPE_MXFP4) ? QK_SMALL
         : type == B200_TYPE_NVFP4 ? 64
         : (type == B200_TYPE_F32 || type == B200_TYPE_I32 || type == B200_TYPE_F16 || type == B200_TYPE_BF16 || type == B200_TYPE_I64) ? 1 : 0;
}
__host__ __device__ constexpr int b200_tsize(int type) {
    return type == B200_TYPE_Q4_0 ? SZ_Q4_0 : type == B200_TYPE_Q8_0 ? SZ_Q8_0 : type == B200_TYPE_Q8_1 ? SZ_Q8_1
         : type == B200_TYPE_Q4_K ? SZ_Q4_K : type == B200_TYPE_Q6_K ? SZ_Q6_K
         : type == B200_TYPE_Q4_1 ? SZ_Q4_1 : type == B200_TYPE_Q5_0 ? SZ_Q5_0 : type == B200_TYPE_Q5_1 ? SZ_Q5_1 : type == B200_TYPE_Q5_K ? SZ_Q5_K
         : type == B200_TYPE_Q2_K ? SZ_Q2_K : type == B200_TYPE_Q3_K ? SZ_Q3_K : type == B200_TYPE_IQ4_NL ? SZ_IQ4_NL : type == B200_TYPE_IQ4_XS ? SZ_IQ4_XS : type == B200_TYPE_MXFP4 ? SZ_MXFP4 : type == B200_TYPE_NVFP4 ? SZ_NVFP4
         : (type == B200_TYPE_F32 || type == B200_TYPE_I32) ? 4 : (type == B200_TYPE_F16 || type == B200_TYPE_BF16) ? 2 : type == B200_TYPE_I64 ? 8 : 0;
}

// ---- internal activation format ("q8 planes") -----------------------------------------------------
// The GEMV / GEMM kernels consume activations quantized exactly like block_q8_1 (same d, s, qs values as
// quantize_row_q8_1_ref) but stored as two planes so that 16-byte chunks of qs are 16-byte aligned:
//   qs : int8  [ncols][k]       (column stride k bytes)
//   ds : half2 [ncols][k/32]    (d, s) per 32-block
struct q8_planes {
    const int8_t * qs;
    const half2 *  ds;
    int64_t        k;
};

// ---- host-side error plumbing ---------------------------------------------------------------------
void b200_set_error(const char * fmt, ...);
#define B200_CUDA_TRY(expr)                                                                          \
    do {                                                                                             \
        cudaError_t err__ = (expr);                                                                  \
        if (err__ != cudaSuccess) {                                                                  \
            b200_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(err__), __FILE__, __LINE__); \
            return B200_E_CUDA;                                                                      \
        }                                                                                            \
    } while (0)

// ---- context (definition shared by the TUs; opaque to callers) -------------------------------------
struct b200_ctx {
    int          device;
    cudaStream_t stream;
    bool         owns_stream;
    int          sm_count;
    int          cc_major, cc_minor;
    size_t       smem_optin;      // max dynamic shared memory per block
    // grow-only scratch (activation planes, ids helper arrays, split-K partials)
    void *       ws;
    size_t       ws_size;
    // pinned staging for the host-buffer convenience path
    void *       dev_x;   size_t dev_x_size;
    void *       dev_y;   size_t dev_y_size;
    int64_t      launches;
    uint64_t     alloc_gen;       // bumped whenever a device buffer kernels were given pointers into is reallocated (captured graphs go stale)
    int64_t      capture_launches0; // ctx->launches when the running stream capture began
    // options
    int          force_generic;   // 1: never use the specialised GEMV kernels
    int          use_tc;          // 1: n > max_gemv_cols goes to the tcgen05 GEMM (when the shape allows)
    int          max_gemv_cols;   // columns per batch handled by the GEMV family (<= 8)
    int          pdl;             // programmatic dependent launch for back-to-back GEMVs
    int          fuse_quant;      // 1: GEMV quantizes activations in its prologue (no separate launch)
    int          tc_pdl;         // tcgen05 GEMM: programmatic dependent launch of the conversion kernel and the GEMM
    int          gather_push;    // b200_mul_mat_gather at n > 8: 1 = GEMM into the shard + one pass of 16-byte peer stores (default), 0 = stores from the GEMM epilogue
    int          tc_pair;        // tcgen05 GEMM: CTA pairs (cta_group::2) on 256-row tiles: 0 never, 1 auto (two or more waves of tiles; default), 2 always
    int          tc_split_k;     // tcgen05 GEMM: split long-K matrices with few tiles along K (0 never, 1 auto, 2/4 forced)
    // fused row-split all-gather, one process per GPU (b200_gather_sync_init)
    int          gs_rank, gs_world;
    uint32_t *   gs_flags[8];
    uint32_t *   gs_state;
    // LL exchange (b200_gather_ll_init): every rank's exchange buffer (2 slots of ll_cap entries), epoch of the last launch
    int          ll_rank, ll_world;
    uint2 *      ll_bufs[8];
    int64_t      ll_cap;
    uint32_t *   ll_state;       // device: [0] epoch of the last completed exchange, [1] receiver CTAs done
    int          tc_nt;           // tcgen05 GEMM tile width override (0 = auto, 128, 256)
    int          gemv_rpw_max_m;  // auto rows-per-warp: largest m that gets two rows per warp
    int          gemv_rpw;        // n=1 Q4_K GEMV: rows per warp below which a matrix does not get more CTAs (1 = default)
    int          gemv_cfg;        // CTA shape of the n=1 Q4_K GEMV: 0 = 32 warps x 1 CTA/SM, 1 = 16x2, 2 = 8x4, 3 = 16x1
    // byte ranges written by kernels that may still be in flight behind a deferred dependency wait (see abi.cu)
    const char * pend_lo[16];
    const char * pend_hi[16];
    int          tc_dequant;   // weight types without a tcgen05 converter: dequantize to fp16 scratch + tensor-core GEMM at n >= 17 (GGML_B200_TC_DEQUANT, default 1)
    int          ops_pdl;      // the graph-op kernels (ops*.cu) take part in programmatic dependent launch (GGML_B200_OPS_PDL, default 1)
    int          n_pend;
    int          defer_wait;      // 1: allow the deferred dependency wait
    int          use_mma;         // 1: Q4_K with 2..8 columns runs the tensor-core GEMV (mmvq_mma.cu) instead of the dp4a one
    int          mma_prequant;    // tensor-core GEMV: 1 = activations quantized once per call by a small kernel, 0 = in every CTA's prologue
    int          mma_min_n;       // Q8_0 / Q4_0: columns from which the tensor-core GEMV replaces gemv_generic
    int          mma_ch;          // tensor-core GEMV: super-blocks per ring stage override (0 = auto, 1)
    int          mma_ks;          // tensor-core GEMV: K slices per row block override (0 = auto; 1, 2, 4, 8, 16)
    int          noop;            // debugging: GEMV kernels return immediately (measures the launch floor)
    // persistent job-table GEMV (mmvq_stream.cu, b200_mul_mat_batch)
    int          use_stream;      // 1: runs of Q4_K n = 1 matmuls of a batch go through one persistent launch
    int          stream_min_jobs; // shortest run that takes the persistent kernel
    int          stream_stages;   // ring slots per warp override (0 = auto)
    int          stream_chain;    // 1: lists that are mostly a dependency chain also take the persistent kernel (default: per-launch kernels)
    int          stream_grid;     // CTAs of the persistent kernel (0 = one per SM); smaller grids let several launches be co-resident
    int          stream_ll;       // 1: "X is exactly an earlier job's output" travels as {value, epoch} words (0: completion counters only)
    uint32_t *   sq_done;         // device: [SQ_MAX_JOBS + 2] completion counters, ticket, launch epoch
    unsigned long long * sq_ll;   // device: exchange buffer of the in-launch dependencies (8-byte {epoch, value} entries)
    size_t       sq_ll_cap;       // ... its capacity in entries
    uint8_t *    sq_img;          // device: quantized-activation images of a launch's groups
    size_t       sq_img_cap;
    uint32_t *   sq_xflag;        // device: [SQ_MAX_JOBS][64] unit flags
    void *       sq_emu;          // tests: per-rank counters, flags and images of the emulated row split
    // row split of the persistent kernel, one process per GPU (b200_stream_comm_init)
    int          sq_world, sq_rank;
    unsigned long long * sq_peer[8];  // every rank's exchange buffer (peer-mapped), 2 * sq_peer_cap entries
    size_t       sq_peer_cap;
    int          stream_trace;    // debugging: 1 = record coordinator time stamps (b200_dbg_stream_trace reads them)
    unsigned long long * sq_trace;
};

int b200_ctx_reserve(b200_ctx * ctx, size_t nbytes);   // ensures ctx->ws_size >= nbytes

// ---- device helpers ---------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// 32-bit load from a 2-byte aligned address (fields of block_q4_0/q8_0/q6_K are only 2-byte aligned)
__device__ __forceinline__ uint32_t ld_b2(const uint8_t * p) {
    const uint16_t * q = reinterpret_cast<const uint16_t *>(p);
    return (uint32_t) __ldg(q) | ((uint32_t) __ldg(q + 1) << 16);
}
__device__ __forceinline__ float ld_half(const uint8_t * p) {
    return __half2float(__ushort_as_half(__ldg(reinterpret_cast<const uint16_t *>(p))));
}

// streaming 128-bit load that does not allocate in L1 (weights are read exactly once)
__device__ __forceinline__ uint4 ld_stream_u4(const void * p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// restores the caller's current device on scope exit (the multi-device entry points switch devices internally)
struct device_guard {
    int saved = -1;
    device_guard()  { if (cudaGetDevice(&saved) != cudaSuccess) saved = -1; }
    ~device_guard() { if (saved >= 0) cudaSetDevice(saved); }
};

// get_scale_min_k4 (ggml/src/ggml-quants.c:818-825) on the 12 scale bytes held as three little-endian words
__device__ __forceinline__ void q4k_scale_min(int j, uint32_t s0, uint32_t s1, uint32_t s2, int & sc, int & mn) {
    if (j < 4) {
        sc = (s0 >> (8 * j)) & 63;
        mn = (s1 >> (8 * j)) & 63;
    } else {
        const int jj = j - 4;
        sc = ((s2 >> (8 * jj)) & 0xF)       | (((s0 >> (8 * jj + 6)) & 3) << 4);
        mn = ((s2 >> (8 * jj + 4)) & 0xF)   | (((s1 >> (8 * jj + 6)) & 3) << 4);
    }
}




