/*
 * ggml_b200.h — the C ABI of the B200-native quantized MUL_MAT / MUL_MAT_ID path.
 *
 * This is the drop-in boundary below the ggml backend shell (libggml-b200.so, which implements the
 * reference's ggml_backend_{reg,device,buffer_type,buffer}_i + ggml_backend_i vtables,
 * /root/reference/ggml/src/ggml-backend-impl.h:17-230) and above the hand-written sm_100a kernels.
 * Plain C: POD arguments, raw DEVICE pointers (unless a function says "host"), byte strides, enum
 * ggml_type values passed as int, cudaStream_t passed inside the opaque context.  No ggml, torch or
 * C++ types cross it, so the same library is bound by the C++ shell, by ctypes (tests, bench.py) and
 * by any other FFI.
 *
 * Every entry point returns 0 on success or a negative B200_E* code; b200_last_error() gives the text.
 * There is NO CPU fallback anywhere behind this header: without a CUDA device every compute call
 * fails with B200_E_CUDA.
 *
 * Reference interfaces each entry replaces (file:line under /root/reference):
 *   b200_quantize_row_q8_1   quantize_row_q8_1_ref            ggml/src/ggml-quants.c:260-293
 *                            (= quantize_row_q8_1_generic      ggml/src/ggml-cpu/quants.c:49-51,
 *                             = s390x VXE quantize_row_q8_1    ggml/src/ggml-cpu/arch/s390/quants.c:94-144;
 *                             CUDA analogue quantize_row_q8_1_cuda ggml/src/ggml-cuda/quantize.cu:369-383)
 *   b200_mul_mat             ggml_compute_forward_mul_mat      ggml/src/ggml-cpu/ggml-cpu.c:1241-1439
 *                            (CUDA analogue ggml_cuda_mul_mat  ggml/src/ggml-cuda/ggml-cuda.cu:2537-2620)
 *   b200_mul_mat_id          ggml_compute_forward_mul_mat_id   ggml/src/ggml-cpu/ggml-cpu.c:1521-1694
 *                            (CUDA analogue ggml_cuda_mul_mat_id ggml/src/ggml-cuda/ggml-cuda.cu:2622-2778)
 *   b200_mm_ids              mm_ids_helper                     ggml/src/ggml-cuda/mmid.cu:26-164
 *   b200_rowsplit_*          ggml_backend_cuda_split_buffer_type + ggml_cuda_op_mul_mat row-split driver
 *                                                              ggml/src/ggml-cuda/ggml-cuda.cu:855-1144,1801-2119
 */
#ifndef GGML_B200_H
#define GGML_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#  define B200_API
#else
#  define B200_API __attribute__((visibility("default")))
#endif

#define B200_ABI_VERSION 1

/* error codes */
enum {
    B200_OK            =  0,
    B200_E_ARG         = -1,  /* bad argument / unsupported shape, type or stride        */
    B200_E_CUDA        = -2,  /* a CUDA runtime call failed (text in b200_last_error())  */
    B200_E_ALLOC       = -3,  /* device or host allocation failed                         */
    B200_E_UNSUPPORTED = -4,  /* valid ggml op that this path does not implement          */
    B200_E_NCCL        = -5,  /* NCCL call failed / library unavailable                   */
};

/* enum ggml_type values this path understands (ggml/include/ggml.h:389-432) */
enum {
    B200_TYPE_F32  = 0,
    B200_TYPE_F16  = 1,      /* graph ops only (KV cache, masks); MUL_MAT weights of this type go to b200_mul_mat_f */
    B200_TYPE_Q4_0 = 2,
    B200_TYPE_Q4_1 = 3,      /* Q4_1 / Q5_0 / Q5_1: the generic GEMV family, GET_ROWS; no tensor-core path (Q5_K has one) */
    B200_TYPE_Q5_0 = 6,
    B200_TYPE_Q5_1 = 7,
    B200_TYPE_Q8_0 = 8,
    B200_TYPE_Q8_1 = 9,
    B200_TYPE_Q2_K = 10,     /* Q2_K / Q3_K / IQ4_NL / IQ4_XS / MXFP4 / NVFP4 (N4): the generic GEMV family and GET_ROWS */
    B200_TYPE_Q3_K = 11,
    B200_TYPE_Q4_K = 12,
    B200_TYPE_Q5_K = 13,
    B200_TYPE_Q6_K = 14,
    B200_TYPE_IQ4_NL = 20,
    B200_TYPE_IQ4_XS = 23,
    B200_TYPE_I32  = 26,
    B200_TYPE_I64  = 27,     /* SET_ROWS indices */
    B200_TYPE_BF16 = 30,     /* graph ops only */
    B200_TYPE_MXFP4 = 39,    /* 17-byte blocks: rows are only byte-aligned */
    B200_TYPE_NVFP4 = 40,    /* 64 values: four UE4M3 scales + 32 bytes of e2m1 nibbles */
};

/* A ggml tensor as the kernels see it: ne[] in elements, nb[] in bytes, ne[0] contiguous
 * (ggml/include/ggml.h:660-700).  `data` is a device pointer. */
typedef struct b200_tensor {
    void *  data;
    int32_t type;
    int32_t _pad;
    int64_t ne[4];
    int64_t nb[4];
} b200_tensor;

typedef struct b200_ctx b200_ctx;   /* one device + one stream + scratch; not thread-safe, like a ggml_backend_t */

/* ---- library / device ------------------------------------------------------------------------- */
B200_API int          b200_abi_version(void);
B200_API const char * b200_last_error(void);               /* thread-local text of the last failure        */
B200_API int          b200_device_count(void);             /* number of CUDA devices with CC 10.x; 0 if none */
B200_API int          b200_device_info(int device, char * name, size_t name_len,
                                       size_t * free_bytes, size_t * total_bytes, int * cc_major, int * cc_minor,
                                       int * sm_count);

/* block geometry — ggml_blck_size / ggml_type_size / ggml_row_size (ggml/src/ggml.c:1263-1283) */
B200_API int     b200_blck_size(int type);
B200_API size_t  b200_type_size(int type);
B200_API size_t  b200_row_size(int type, int64_t k);

/* ---- context ----------------------------------------------------------------------------------- */
/* stream == NULL: the context creates and owns a non-blocking stream; otherwise it borrows `stream`
 * (a cudaStream_t) — this is how PyTorch's current stream is shared. */
B200_API int   b200_ctx_create(int device, void * stream, b200_ctx ** out);
B200_API void  b200_ctx_destroy(b200_ctx * ctx);
B200_API void *b200_ctx_stream(b200_ctx * ctx);
B200_API int   b200_ctx_synchronize(b200_ctx * ctx);
/* number of kernels this context has launched since creation (bench.py's gpu_launches) */
B200_API int64_t b200_ctx_launch_count(const b200_ctx * ctx);
/* tuning knobs (also read once from the GGML_B200_* environment): name in
 * {"force_generic","use_tc","max_gemv_cols","pdl","fuse_quant","defer_wait","gemv_cfg","use_mma","mma_min_n","mma_ks","mma_ch",
 *  "mma_prequant","tc_pair","tc_nt","tc_split_k","tc_pdl","tc_dequant" (weight types without a tcgen05 converter go through an fp16 scratch
 *  copy at n >= 17),"ops_pdl" (programmatic dependent launch on the graph-op kernels),"gather_push","noop","use_stream","stream_min_jobs",
 *  "stream_stages","stream_ll","stream_grid","stream_chain","stream_trace","gemv_rpw","gemv_rpw_max_m"}; returns the previous value or B200_E_ARG */
B200_API int   b200_ctx_set_option(b200_ctx * ctx, const char * name, int value);

/* ---- (1) activation quantizer ------------------------------------------------------------------- */
/* x: nrows rows of k f32 (row stride x_row_stride_bytes); y: nrows * (k/32) block_q8_1
 * {fp16 d; fp16 s; int8 qs[32]} = 36 B each (ggml/src/ggml-common.h:248-259), rows packed.
 * Bit-exact against quantize_row_q8_1_ref.  k % 32 == 0. */
B200_API int b200_quantize_row_q8_1(b200_ctx * ctx, const float * x, int64_t x_row_stride_bytes,
                                    void * y, int64_t k, int64_t nrows);

/* ---- (2)(3) MUL_MAT ----------------------------------------------------------------------------- */
/* dst[m, n, ne12, ne13] = src0[k, m, ne02, ne03] (Q4_0|Q8_0|Q4_K|Q6_K) x src1[k, n, ne12, ne13] (F32),
 * broadcast i02 = i12 / (ne12/ne02), i03 = i13 / (ne13/ne03)  (ggml/src/ggml.c:3232-3255).
 * src0 rows must be whole blocks with nb[0] == type_size; src1 needs nb[0] == 4; dst must be contiguous f32.
 * n <= 8 per batch runs the GEMV family, larger n the tensor-core GEMM. */
B200_API int b200_mul_mat(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * src1, const b200_tensor * dst);
/* Ordering contract of the decode kernels: consecutive launches of one context overlap (programmatic dependent launch); a GEMV
 * reads its activations BEFORE its predecessors have finished only when no kernel still in flight on this context writes them.
 * The context tracks the outputs of its OWN launches for that; memory written by anything else — another context or stream, a
 * peer GPU, a copy — must be ordered before the call by the caller (event / synchronize), as with any CUDA stream.  The entry
 * points that write through peer mappings or wait for events (b200_mul_mat_gather*, b200_event_wait, b200_rowsplit_*) make the
 * next launch wait early on their own. */
/* returns 1 if b200_mul_mat accepts this triple, 0 if not (used by supports_op) — no device needed */
B200_API int b200_mul_mat_supported(const b200_tensor * src0, const b200_tensor * src1, const b200_tensor * dst);

/* ---- MUL_MAT without any allocation ---------------------------------------------------------------- */
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, shape rules and results
 * contract as b200_mul_mat, computed by the pre-quantized GEMV route (Q8_1 activation planes -> one integer-dot GEMV), with
 * the planes in a file-scope static __device__ array of B200_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds
 * until exit.  Calls no cudaMalloc / cudaFree / allocating library routine.  Limit: the planes of src1,
 * ncols * k + ncols * (k/32) * 4 bytes (each term rounded up to 256; ncols = ne1 * ne2 * ne3 of src1), must fit the array,
 * else B200_E_ARG — use b200_mul_mat for bigger activations.  Every call on a device shares the one array (it keeps the last
 * call's contents; every entry read is rewritten first, nothing needs clearing): calls that overlap on the GPU race. */
#define B200_NOALLOC_SCRATCH_BYTES (1u << 20)
B200_API int b200_mul_mat_noalloc(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * src1, const b200_tensor * dst);

/* ---- (4) MUL_MAT_ID ----------------------------------------------------------------------------- */
/* as[k, m, n_expert], b[k, n_used|1, n_tokens] F32, ids[n_used, n_tokens] I32 (may be a strided view),
 * c[m, n_used, n_tokens] F32:  c[:, e, t] = as[:, :, ids[e,t]] . b[:, e % b.ne[1], t]  (ggml/src/ggml.c:3279-3315) */
B200_API int b200_mul_mat_id(b200_ctx * ctx, const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids,
                             const b200_tensor * c);
B200_API int b200_mul_mat_id_supported(const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids,
                                       const b200_tensor * c);

/* expert gather helper: from ids[n_used, n_tokens] (row stride ids_nb1 bytes) build
 *   ids_src1[n_used*n_tokens]  : for sorted position p, the flat (token*n_used + slot) pair index
 *   expert_bounds[n_expert+1]  : sorted positions [bounds[e], bounds[e+1]) belong to expert e
 * Stable within an expert (token-major order), like mm_ids_helper (ggml/src/ggml-cuda/mmid.cu:26-116). */
B200_API int b200_mm_ids(b200_ctx * ctx, const int32_t * ids, int64_t ids_nb1, int n_expert, int n_used, int64_t n_tokens,
                         int32_t * ids_src1, int32_t * expert_bounds);

/* ---- host-buffer convenience (the e2e path: H2D + compute + D2H inside the call) ---------------- */
/* Device-resident weights handle: uploads `nbytes` of quantized rows once. */
B200_API int  b200_weights_upload(b200_ctx * ctx, int type, const void * host_rows, int64_t k, int64_t m, void ** dev_out);
B200_API void b200_weights_free(b200_ctx * ctx, void * dev);
/* y_host[m, n] = W (device, from b200_weights_upload) x x_host[k, n]; copies x H2D and y D2H on the
 * context's stream and synchronizes.  x_host / y_host should be pinned for full speed. */
B200_API int  b200_mul_mat_host(b200_ctx * ctx, int type, const void * w_dev, int64_t m, int64_t k,
                                const float * x_host, int64_t n, float * y_host);

/* The same for many matmuls at once (one decode / prefill step): all uploads, all launches, all downloads on the context's
 * stream and ONE synchronisation — what ggml_backend_sched does around a graph_compute of several MUL_MAT nodes
 * (ggml/src/ggml-backend.cpp: copy inputs -> compute -> copy outputs).  Host vectors that follow one another in memory are
 * moved with a single copy. */
typedef struct b200_host_mm {
    int32_t type, _pad;
    const void *  w_dev;     /* from b200_weights_upload */
    int64_t       m, k, n;
    const float * x_host;    /* [k, n] */
    float *       y_host;    /* [m, n] */
} b200_host_mm;
B200_API int  b200_mul_mat_host_batch(b200_ctx * ctx, const b200_host_mm * jobs, int n_jobs);

/* n_mats (1..3 fused; more are run one by one) weight matrices against the SAME activations `b`: d[j] = a[j] x b.  What a
 * transformer layer does for the q/k/v projections and for ffn gate/up: the reference issues one MUL_MAT node per matrix
 * (src/llama-graph.cpp build_attn / build_ffn -> ggml_mul_mat per weight); here matrices of equal type and row length share
 * one launch — a single activation prologue, one grid — with results bit-identical to n_mats b200_mul_mat calls. */
B200_API int  b200_mul_mat_multi(b200_ctx * ctx, int n_mats, const b200_tensor * const * a, const b200_tensor * b,
                                 const b200_tensor * const * d);

/* A list of MUL_MATs in one call, executed as if in list order (a later job may read what an earlier one wrote): what
 * ggml_backend_graph_compute hands a backend when a cgraph holds several MUL_MAT nodes (ggml/src/ggml-backend.cpp:444-448;
 * test-backend-ops' perf mode builds such graphs, tests/test-backend-ops.cpp:1455-1583) and what the reference's CUDA backend
 * replays as one CUDA graph (ggml/src/ggml-cuda/ggml-cuda.cu:4457-4490).  Runs of Q4_K matmuls with one activation column go
 * through ONE persistent launch: a grid of one CTA per SM walks the list with per-warp shared-memory rings fed by bulk async
 * copies, so the weight stream does not drain between matrices; consecutive jobs with the same src1 share one quantized
 * copy of it; dependencies between jobs are found from the tensors' address ranges and waited for inside the kernel.
 * Results are bit-identical to n_jobs b200_mul_mat calls.  Everything else in the list runs as b200_mul_mat[_multi]. */
B200_API int  b200_mul_mat_batch(b200_ctx * ctx, int n_jobs, const b200_tensor * const * a, const b200_tensor * const * b,
                                 const b200_tensor * const * d);

/* ---- row split, one process per GPU -------------------------------------------------------------- */
/* MUL_MAT on this rank's row shard `a` (rows [dst_row0, dst_row0 + a->ne[1]) of the full matrix) with the all-gather of
 * the output fused into the kernel's stores: every result also goes to dst[p][col * dst_m + dst_row0 + row] for the
 * n_dst (<= 8) full-height outputs `dst` — peer-mapped pointers (CUDA IPC / symmetric memory over NVLink), normally one
 * per rank including this one.  n <= 8 columns: GEMV family; n > 8 (prefill): the tcgen05 GEMM's epilogue stores its tiles
 * to the peers (completion = the caller's barrier; the in-kernel protocols below are for decode).  The multi-process counterpart of b200_rowsplit_mul_mat mode 1;
 * replaces the per-device dst copies of the reference's row-split driver (ggml/src/ggml-cuda/ggml-cuda.cu:2030-2119). */
B200_API int  b200_mul_mat_gather(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * b, const b200_tensor * d_shard,
                                  int n_dst, float * const * dst, int64_t dst_m, int64_t dst_row0);

/* Completion protocol for b200_mul_mat_gather (replaces a cross-GPU barrier after every matmul): flags[p] is rank p's
 * array of `world` zero-initialised uint32 in peer-mapped memory.  The last CTA of every gather launch publishes the
 * launch's epoch into flags[p][rank] for all p; the next gather launch waits, before reading its activations, until all
 * ranks have published the previous launch.  All ranks must issue the same sequence of gather calls.  world = 0 turns
 * the protocol off.  (ggml-cuda orders the same dependency with cross-device events, ggml-cuda.cu:2098-2119.) */
B200_API int  b200_gather_sync_init(b200_ctx * ctx, int rank, int world, uint32_t * const * flags);

/* The low-latency ("LL") form of the same exchange for decode (n = 1, Q4_K): the all-gather AND its completion travel in the
 * data.  Every result is sent as one indivisible 8-byte store {f32 bits, epoch} into an exchange slot on every rank (own
 * rank included); the receiver — a second small kernel launched by the same call — polls its local slot until every entry
 * carries the epoch and writes plain full-height f32 vectors.  No fence, no flag, no barrier kernel, no NCCL call: the
 * cost per matmul is one NVLink store latency.  (The reference's row split instead issues per-device peer copies and
 * cross-device events per matmul, ggml/src/ggml-cuda/ggml-cuda.cu:2030-2119.)
 *
 * b200_gather_ll_init: bufs[p] = rank p's exchange buffer, peer-mapped (CUDA IPC / symmetric memory), zero-initialised,
 *   2 * slot_entries * 8 bytes (two slots used alternately).  world = 0 turns it off.
 * b200_mul_mat_gather_ll: n_mats (1..3) row shards a[j] = rows [row0[j], row0[j] + a[j]->ne[1]) of matrices that share the
 *   activations b (one f32 column); d_full[j] = this rank's full-height f32 result vector of matrix j, complete (all ranks'
 *   rows) when the call's work is done on the stream.  All ranks issue the same sequence of calls; the sum of the full
 *   heights of one call must fit a slot.  d_full must not overlap b. */
B200_API int  b200_gather_ll_init(b200_ctx * ctx, int rank, int world, void * const * bufs, int64_t slot_entries);
B200_API int  b200_mul_mat_gather_ll(b200_ctx * ctx, int n_mats, const b200_tensor * const * a_shard, const b200_tensor * b,
                                     const b200_tensor * const * d_full, const int64_t * row0);

/* The row split of a whole LIST of matmuls, one process per GPU (decode): the persistent job-table kernel of b200_mul_mat_batch with
 * the all-gather folded into it.  Every rank passes its row shards a[j] = rows [row0[j], row0[j] + a[j]->ne[1]) and gets the
 * FULL-height results d_full[j].  Inside the one launch every result travels as one 8-byte {value, epoch} word into the exchange
 * buffer of every rank (NVLink peer stores, one lane per destination rank); a later job of the list whose src1 IS an earlier job's
 * full-height output polls those words while it quantizes its activations — there is no receiver kernel, no barrier, no NCCL call
 * and no kernel boundary between dependent matmuls.  The plain d_full vectors are written at the end of the launch.
 * b200_stream_comm_init: bufs[p] = rank p's exchange buffer, peer-mapped (CUDA IPC / symmetric memory), zero-initialised,
 *   2 * slot_entries * 8 bytes (two slots alternate between launches; a rank can be at most one launch ahead of the slowest).
 * All ranks issue the same sequence of b200_mul_mat_batch_gather calls.  (The reference orders the same exchange with per-matmul
 * peer copies and cross-device events, ggml/src/ggml-cuda/ggml-cuda.cu:2030-2119.) */
B200_API int  b200_stream_comm_init(b200_ctx * ctx, int rank, int world, void * const * bufs, int64_t slot_entries);
B200_API int  b200_mul_mat_batch_gather(b200_ctx * ctx, int n_jobs, const b200_tensor * const * a_shard, const b200_tensor * const * b,
                                        const b200_tensor * const * d_full, const int64_t * row0);

/* ---- raw device memory (used by the backend shell's buffers and by ctypes callers without torch) -- */
B200_API int  b200_malloc(int device, size_t nbytes, void ** out);
B200_API int  b200_free(int device, void * ptr);
B200_API int  b200_memcpy_h2d(b200_ctx * ctx, void * dst_dev, const void * src_host, size_t nbytes);   /* synchronous */
B200_API int  b200_memcpy_d2h(b200_ctx * ctx, void * dst_host, const void * src_dev, size_t nbytes);   /* synchronous */
B200_API int  b200_memcpy_d2d(b200_ctx * ctx, void * dst_dev, const void * src_dev, size_t nbytes);    /* synchronous; any two devices */
B200_API int  b200_memset(b200_ctx * ctx, void * dst_dev, int value, size_t nbytes);                   /* synchronous */
/* asynchronous on the context's stream; kind: 0 = host->device, 1 = device->host, 2 = device->device */
B200_API int  b200_memcpy_async(b200_ctx * ctx, void * dst, const void * src, size_t nbytes, int kind);

/* ---- pinned host memory, events ------------------------------------------------------------------- */
/* Page-locked host memory for the shell's host buffer type (the reference: ggml_cuda_host_malloc + the "CUDA_Host" buffer
 * type, ggml/src/ggml-cuda/ggml-cuda.cu:1493-1542), which is what makes llama.cpp's asynchronous weight upload and the
 * scheduler's asynchronous input copies truly asynchronous (src/llama-model-loader.cpp:1461 asks for caps.async &&
 * caps.host_buffer && caps.events). */
B200_API int  b200_host_malloc(size_t nbytes, void ** out);
B200_API int  b200_host_free(void * ptr);
/* Events on a context's stream (ggml_backend_cuda_event_record / _wait, ggml-cuda.cu:4498-4520; device event_new / _free /
 * _synchronize, :5430-5460).  An event is created for a device and may be waited for by any context of any device. */
B200_API int  b200_event_create(int device, void ** event_out);
B200_API int  b200_event_create_timed(int device, void ** event_out);                 /* keeps its time stamp (the plug-in times a graph run eagerly and replayed) */
B200_API int  b200_event_elapsed_ms(void * start, void * stop, float * ms);           /* device time between two timed events; waits for `stop` */
B200_API int  b200_event_destroy(int device, void * event);
B200_API int  b200_event_record(b200_ctx * ctx, void * event);        /* after everything queued on ctx's stream so far      */
B200_API int  b200_event_wait(b200_ctx * ctx, void * event);          /* ctx's stream waits (device side) for the event      */
B200_API int  b200_event_synchronize(void * event);                   /* the host waits                                      */

/* ---- row-split across GPUs (single process, one NCCL communicator per device) -------------------- */
typedef struct b200_rowsplit b200_rowsplit;
/* devices[n_dev]; split[n_dev] cumulative-free fractions (NULL = even).  Cut points are rounded down to
 * `B200_ROWSPLIT_ROUND` rows like get_row_split (ggml/src/ggml-cuda/ggml-cuda.cu:855-881). */
#define B200_ROWSPLIT_ROUND 128
B200_API int  b200_rowsplit_create(const int * devices, int n_dev, const float * split, b200_rowsplit ** out);
B200_API void b200_rowsplit_destroy(b200_rowsplit * rs);
B200_API int  b200_rowsplit_n_dev(const b200_rowsplit * rs);
B200_API int  b200_rowsplit_rows(const b200_rowsplit * rs, int64_t m, int idx, int64_t * row_lo, int64_t * row_hi);
/* scatter host rows of one [k, m] weight matrix to the devices; returns a handle */
typedef struct b200_split_weights b200_split_weights;
B200_API int  b200_rowsplit_upload(b200_rowsplit * rs, int type, const void * host_rows, int64_t k, int64_t m,
                                   b200_split_weights ** out);
B200_API void b200_rowsplit_weights_free(b200_rowsplit * rs, b200_split_weights * w);
/* y[m, n] on every device (y_dev[i] is device i's full [m, n] f32 output) = W x x, where x_dev[i] is device i's
 * replica of x[k, n] f32.  Each device computes its row shard, then an all-gather of the row slices
 * (NCCL over NVLink, or direct peer stores when `mode` == 1) fills every y_dev[i].  Asynchronous on the
 * per-device streams; b200_rowsplit_synchronize waits for all. */
B200_API int  b200_rowsplit_mul_mat(b200_rowsplit * rs, const b200_split_weights * w, float * const * x_dev, int64_t n,
                                    float * const * y_dev, int mode);
/* The split-buffer-type form (what the shell's "ggml_backend_split_buffer_type" needs; reference: ggml_cuda_op_mul_mat's row-split
 * driver, ggml/src/ggml-cuda/ggml-cuda.cu:1801-2119): src1 [k, n] and dst [m, n] live on ONE device — the one of main_ctx, on whose
 * stream the op is ordered — and shards[i] holds rows [row_lo[i], row_hi[i]) of the [k, m] weight matrix on rs's device i.  The
 * other devices fetch src1 over NVLink, run the single-GPU kernel on their shard and store their rows straight into dst (peer
 * stores from the kernel epilogue); per-row results are bit-identical to the unsplit b200_mul_mat. */
B200_API int  b200_rowsplit_mul_mat_main(b200_rowsplit * rs, b200_ctx * main_ctx, int type, int64_t k, int64_t m, void * const * shards,
                                         const int64_t * row_lo, const int64_t * row_hi, const b200_tensor * src1, const b200_tensor * dst);
B200_API int  b200_rowsplit_synchronize(b200_rowsplit * rs);
B200_API b200_ctx * b200_rowsplit_ctx(b200_rowsplit * rs, int idx);


/* ---- all-reduce between the contexts of one process ---------------------------------------------------
 * What the shell answers "ggml_backend_comm_init / _free / _allreduce_tensor" with (typedefs ggml/include/ggml-backend.h:205-208; the
 * meta / tensor-parallel backend is the caller, ggml-backend-meta.cpp:1522-1553; reference: ggml-cuda.cu:1184-1312).  data[i] is
 * participant i's packed F32 tensor of ne elements (16-byte aligned) on ctxs[i]'s device; afterwards every participant holds the
 * element-wise sum, bit-identical everywhere (summed in participant order).  zero[i] != 0 (or zero == NULL: none): participant i
 * contributes zeros.  Peer memory over NVLink + events, no library collective; asynchronous on every participant's stream. */
typedef struct b200_comm b200_comm;
B200_API int  b200_comm_create(b200_ctx * const * ctxs, int n, b200_comm ** out);       /* n <= 8; enables peer access between the devices */
B200_API void b200_comm_destroy(b200_comm * comm);
B200_API int  b200_comm_size(const b200_comm * comm);
B200_API int  b200_comm_allreduce(b200_comm * comm, void * const * data, const int * zero, int64_t ne);

/* ---- stream capture (reference: the CUDA-graph path of ggml_backend_cuda_graph_compute, ggml/src/ggml-cuda/ggml-cuda.cu:4457-4490) -------
 * Between b200_capture_begin and b200_capture_end every entry point of this header RECORDS its work on the context's stream instead of
 * running it; b200_capture_end returns a graph that b200_graph_launch replays with one call.  An entry point that would have to
 * synchronize or reallocate a scratch buffer fails during capture and invalidates it (end then returns an error): run the same work
 * eagerly once — that grows the buffers — and capture the next time.  A graph holds pointers into the context's scratch buffers and into
 * every tensor it was recorded with; it is stale once b200_ctx_alloc_generation() has changed. */
B200_API int      b200_capture_begin(b200_ctx * ctx);
B200_API int      b200_capture_end(b200_ctx * ctx, void ** graph_out);
B200_API int      b200_graph_launch(b200_ctx * ctx, void * graph);
B200_API void     b200_graph_free(void * graph);
B200_API uint64_t b200_ctx_alloc_generation(const b200_ctx * ctx);

/* ---- the other ops of a Llama graph ---------------------------------------------------------
 * So that a cgraph stays on the device between its MUL_MATs.  Each call restates the reference's CPU forward function of the op
 * (ggml/src/ggml-cpu/ops.cpp, binary-ops.cpp, unary-ops.cpp; the exact lines are cited in csrc/ops.cu) on device tensors with
 * ggml's shapes, byte strides and broadcasting rules, queued on the context's stream.  F32 data unless said otherwise; a form a
 * function does not implement returns B200_E_UNSUPPORTED (the shell's supports_op() answers false for it, ggml then runs the node on
 * another backend). */
enum { B200_OP_ADD = 0, B200_OP_SUB = 1, B200_OP_MUL = 2, B200_OP_DIV = 3 };
/* enum ggml_glu_op (ggml/include/ggml.h:616-625) */
enum { B200_GLU_REGLU = 0, B200_GLU_GEGLU = 1, B200_GLU_SWIGLU = 2, B200_GLU_SWIGLU_OAI = 3, B200_GLU_GEGLU_ERF = 4, B200_GLU_GEGLU_QUICK = 5 };
/* enum ggml_unary_op (ggml/include/ggml.h:589-614), the members implemented here */
enum { B200_UNARY_ABS = 0, B200_UNARY_SGN = 1, B200_UNARY_NEG = 2, B200_UNARY_STEP = 3, B200_UNARY_TANH = 4, B200_UNARY_RELU = 6,
       B200_UNARY_SIGMOID = 7, B200_UNARY_GELU = 8, B200_UNARY_GELU_QUICK = 9, B200_UNARY_SILU = 10, B200_UNARY_EXP = 13, B200_UNARY_GELU_ERF = 16 };
/* dst = src0 (op) src1, src1 repeated into src0's shape (ggml_can_repeat); any strides */
B200_API int b200_op_binary(b200_ctx * ctx, int op, const b200_tensor * src0, const b200_tensor * src1, const b200_tensor * dst);
/* GGML_OP_REPEAT: dst = src0 tiled into dst's shape */
B200_API int b200_op_repeat(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * dst);
/* GGML_OP_ADD_ID: dst[:, i1, i2] = src0[:, i1, i2] + src1[:, ids[i1, i2]] (expert biases of MUL_MAT_ID) */
B200_API int b200_op_add_id(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * src1, const b200_tensor * ids, const b200_tensor * dst);
/* GGML_OP_GLU: dst = act(x) * g; src1 != NULL: x = src0, g = src1 (ggml_glu_split); src1 == NULL: the two halves of src0's rows
 * (`swapped` exchanges them).  alpha / limit: SWIGLU_OAI only. */
B200_API int b200_op_glu(b200_ctx * ctx, int glu_op, const b200_tensor * src0, const b200_tensor * src1, const b200_tensor * dst,
                         int swapped, float alpha, float limit);
B200_API int b200_op_unary(b200_ctx * ctx, int unary_op, const b200_tensor * src0, const b200_tensor * dst);      /* contiguous tensors */
B200_API int b200_op_unary_supported(int unary_op);
B200_API int b200_op_scale(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * dst, float scale, float bias);   /* dst = src0 * scale + bias */
B200_API int b200_op_rms_norm(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * dst, float eps);
/* RMS_NORM followed by MUL with a weight repeated into src0's shape, in one pass: dst = (src0 / rms) * weight (the shell fuses the two nodes) */
B200_API int b200_op_rms_norm_mul(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * weight, const b200_tensor * dst, float eps);
/* GGML_OP_ROPE, modes normal (0) and NeoX (2), F32 or F16 rows; op_params = the node's 15 int32 parameter words */
B200_API int b200_op_rope(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * pos, const b200_tensor * freq_factors /* or NULL */,
                          const b200_tensor * dst, const int32_t * op_params);
/* GGML_OP_SOFT_MAX: softmax(src0 * scale + slope(head) * mask) over ne0, mask F16 or F32 broadcast over heads / batches, optional sinks */
B200_API int b200_op_soft_max(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * mask /* or NULL */, const b200_tensor * sinks /* or NULL */,
                              const b200_tensor * dst, float scale, float max_bias);
/* GGML_OP_CPY / CONT / DUP between F32 / F16 / BF16 tensors with the same number of elements, any strides */
B200_API int b200_op_cpy(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * dst);
/* GGML_OP_GET_ROWS: rows of an F32 / F16 / BF16 / Q4_0 / Q8_0 / Q4_K / Q6_K matrix by I32 index -> F32 (token embeddings) */
B200_API int b200_op_get_rows(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * idx, const b200_tensor * dst);
/* GGML_OP_SET_ROWS: F32 rows scattered by I64 / I32 index into an F32 / F16 / BF16 tensor (KV-cache store) */
B200_API int b200_op_set_rows(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * idx, const b200_tensor * dst);
/* MUL_MAT with F32 / F16 / BF16 src0 (attention: K.q, V.p), src0 rows contiguous, batch broadcast like b200_mul_mat */
B200_API int b200_mul_mat_f(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * src1, const b200_tensor * dst);
B200_API int b200_mul_mat_f_supported(const b200_tensor * src0, const b200_tensor * src1, const b200_tensor * dst);
/* N2 — MUL_MAT(+ bias)(+ gate MUL_MAT (+ bias) + GLU) for one activation column per batch in ONE launch (reference:
 * ggml_cuda_should_fuse_mul_mat_vec_q, ggml/src/ggml-cuda/ggml-cuda.cu:2499-2535, kernel mmvq.cu:440-477):
 *   gate == NULL: dst = up x + bias_up
 *   gate != NULL: dst = glu(gate x + bias_gate, up x + bias_up)        (act on the gate product, as ggml_glu_split(gate, up))
 * up / gate: quantized matrices of one type and shape, x: F32 with ne[1] == 1, biases: F32 shaped like dst or NULL. */
/* dst = src0 x src1 + addend in the store of the fast GEMV kernels (n <= 8; addend: contiguous F32 shaped like dst — a bias, or the residual
 * stream; dst may be the addend's memory, not src1's).  What the shell uses for MUL_MAT + ADD at any matrix size. */
B200_API int b200_mul_mat_add(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * src1, const b200_tensor * addend, const b200_tensor * dst);
B200_API int b200_mul_mat_add_supported(const b200_tensor * src0, const b200_tensor * src1, const b200_tensor * addend, const b200_tensor * dst);
B200_API int b200_mul_mat_fused(b200_ctx * ctx, const b200_tensor * up, const b200_tensor * gate, const b200_tensor * x,
                                const b200_tensor * bias_up, const b200_tensor * bias_gate, int glu_op, const b200_tensor * dst);
B200_API int b200_mul_mat_fused_supported(const b200_tensor * up, const b200_tensor * gate, const b200_tensor * x,
                                          const b200_tensor * bias_up, const b200_tensor * bias_gate, int glu_op, const b200_tensor * dst);

#ifdef __cplusplus
}
#endif
#endif /* GGML_B200_H */
