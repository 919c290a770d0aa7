// stream_jobs.cuh — the job table of the persistent decode GEMV (mmvq_stream.cu), shared with its host-side builder (abi.cu).
//
// One launch = one list of Q4_K n = 1 MUL_MATs executed "as if in list order": what ggml_backend_graph_compute hands a backend
// when a cgraph holds several MUL_MAT nodes (/root/reference/ggml/src/ggml-backend.cpp:444-448 -> graph_compute; the perf mode
// of test-backend-ops builds exactly such graphs, tests/test-backend-ops.cpp:1455-1583) and what the reference's CUDA backend
// turns into ONE CUDA-graph launch (ggml/src/ggml-cuda/ggml-cuda.cu:4457-4490).  Here the list is a kernel PARAMETER
// (__grid_constant__, <= 32 KB), so a launch needs no table upload and is capturable as is.
#pragma once
#include "common.cuh"

#define SQ_MAX_JOBS    448      // per launch (longer lists are cut into several launches; bounded by the 32 KB of kernel parameters)
#define SQ_MAX_GROUP   64       // jobs that may share one quantized activation vector (bounds the live window of the CTA counters)

#define SQ_NO_LL 0xffffffffu

struct sq_job {                 // 56 bytes
    const uint8_t * W;          // first row of the [m, k] Q4_K matrix (or row shard), 16-byte aligned
    const float *   X;          // k f32, 16-byte aligned
    float *         Y;          // m f32
    uint32_t nb01;              // row stride in bytes (multiple of 16)
    uint32_t base;              // row PAIRS / grid: CTA b owns pairs [b * base + min(b, rem), ... + base + (b < rem)); pair p = rows 2p, 2p + 1
    uint32_t row0;              // row split (world > 1): W holds rows [row0, row0 + m) of a matrix of m_full rows; Y is the FULL-height vector
    uint32_t m_full;
    uint32_t ll_off;            // SQ_NO_LL, or: every result also goes, as {f32 bits, launch epoch} in one 8-byte word, to entry
                                // ll_off + row of the exchange buffer — a later group of this launch reads its activations from there
    uint16_t rem;               // row pairs % grid
    uint16_t nsb;               // k / 256
    uint16_t xg;                // index of the job's activation group (consecutive jobs with the same X share one quantized copy)
    int16_t  dep;               // first job of a group: index of the LAST earlier job of this launch whose completion the group
                                // has to wait for (hazards other than "X is exactly an earlier job's output"); -1 = none
    int16_t  xsrc;              // first job of a group: the earlier job whose exchange entries ARE this group's activations; -1 = read X
    uint16_t flags;             // bit 0: a later group waits for this job -> completion is counted in done[job]; bit 1: odd number of rows
};

struct sq_params {
    int        n_jobs, n_groups;
    int        kmax;            // largest k of the launch: an activation buffer holds kmax int8 + kmax / 32 float2
    int        stages;          // weight-ring slots per consumer warp (2 or 3)
    unsigned   xring;           // bytes of the shared-memory ring of activation images
    int        n_flagged;       // jobs with flags & 1
    int        _pad;
    int        n_ll;            // jobs with an exchange region
    uint32_t * done;            // [SQ_MAX_JOBS] completion counters (CTAs done with job j), [SQ_MAX_JOBS] ticket, [SQ_MAX_JOBS + 1] launch
                                // epoch; counters and ticket are zero between launches, the epoch counts the launches that used it
    unsigned long long * trace; // debugging (ctx option stream_trace): %globaltimer stamps of the coordinator of CTA 0 and of the last CTA,
                                // [2][SQ_MAX_JOBS][4] = {group start, opened, own units done, all units done}; NULL = off
    uint8_t *  ximg;            // quantized-activation images of the launch's groups (gimg[g] = byte offset of group g's image)
    uint32_t * xflag;           // [SQ_MAX_JOBS][64] unit flags: == launch epoch once unit u of group g is in the image
    unsigned long long * ll;    // exchange buffer (entries of 8 bytes: epoch << 32 | f32 bits), zero-initialised
    // row split, one process per GPU (b200_mul_mat_batch_gather): every result goes as one {value, epoch} word into the exchange
    // buffer of EVERY rank (peer-mapped, own rank included), at entry (epoch & 1) * ll_cap + ll_off + row0 + row; the plain Y
    // vectors are written from the local exchange buffer at the end of the launch
    int        world, rank;
    unsigned long long   ll_cap;
    unsigned long long * ll_peer[8];
    // tests only: all ranks inside one launch on one GPU (gemv_q4k_stream<2>); 0 = off
    unsigned long long   emu_stride;      // bytes between the activation / result vectors of consecutive ranks
    unsigned long long   emu_img_stride;  // bytes between the ranks' image buffers
    uint16_t   gfirst[SQ_MAX_JOBS];   // first job of activation group g
    uint32_t   gimg[SQ_MAX_JOBS];     // byte offset of group g's image in ximg
    uint32_t   gsm[SQ_MAX_JOBS];      // byte offset of group g's image in the shared-memory image ring (2 x the largest image of the launch)
    int16_t    gwait[SQ_MAX_JOBS];    // the latest earlier group whose image lay in that place (-1: none): it must be finished first
    sq_job     jobs[SQ_MAX_JOBS];
};

// host launcher (mmvq_stream.cu); P.jobs[0 .. n_jobs) filled, groups / deps / flags resolved by the caller
int    b200_launch_gemv_stream(b200_ctx * ctx, sq_params & P);
// weight-ring slots per warp for a launch whose longest activation vector has kmax values (0: does not fit), and the bytes of
// the image ring that go with it
int    b200_gemv_stream_stages(const b200_ctx * ctx, int64_t kmax);
size_t b200_gemv_stream_xring(const b200_ctx * ctx, int stages);
// CTAs of a launch on this context (the job table's base / rem fields are computed for this grid)
int    b200_gemv_stream_grid(const b200_ctx * ctx);
