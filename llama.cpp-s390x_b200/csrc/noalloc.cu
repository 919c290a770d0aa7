// noalloc.cu — b200_mul_mat_noalloc: MUL_MAT for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
// serving loops, harnesses that forbid it).  No counterpart in the reference.
//
// It is the pre-quantized GEMV route of b200_mul_mat (quantize the activations into q8 planes, then gemv_generic reads the
// planes — the route abi.cu takes with ctx->ws as scratch), with the planes in a file-scope static __device__ array instead of
// the context's grow-only cudaMalloc scratch.  The array costs B200_NOALLOC_SCRATCH_BYTES on every GPU that loads the library,
// until the process exits.
//
//   * The array's NAME is a host-side stand-in in host code, not a device address; the device address is fetched once per
//     device with cudaGetSymbolAddress (that device current) and that pointer is what the kernels receive.
//   * Every call on a device shares the one array and it keeps the previous call's contents.  Nothing here needs to start at
//     zero: the quantizer writes every plane entry the GEMV reads, so stale bytes are never consumed.
//   * Calls that overlap on the GPU (two streams / two contexts on one device) RACE on the array.  Serialize them.
//   * Calls neither cudaMalloc / cudaFree nor any allocating library routine.
#include "mm_problem.cuh"
#include "quantize.cuh"

#include <mutex>

int b200_launch_gemv_generic(b200_ctx * ctx, int type, const mm_problem & P, int nc, int64_t tiles);   // mmvq.cu

namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[B200_NOALLOC_SCRATCH_BYTES];

constexpr int MAX_DEV = 64;
std::mutex      g_mu;
unsigned char * g_scratch_dev[MAX_DEV] = {};     // device address of g_noalloc_scratch on device i (0 = not fetched yet)

// device address of the scratch on `device`, which must be the current device
int scratch_address(int device, unsigned char ** out) {
    if (device < 0 || device >= MAX_DEV) { b200_set_error("mul_mat_noalloc: device %d out of range", device); return B200_E_ARG; }
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_dev[device]) {
        unsigned char * p = nullptr;
        const cudaError_t err = cudaGetSymbolAddress((void **) &p, g_noalloc_scratch);
        if (err != cudaSuccess || !p) {
            cudaGetLastError();
            b200_set_error("mul_mat_noalloc: cudaGetSymbolAddress on device %d failed: %s", device, cudaGetErrorString(err));
            return B200_E_CUDA;
        }
        g_scratch_dev[device] = p;
    }
    *out = g_scratch_dev[device];
    return B200_OK;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// bytes of q8 planes for the activations b: int8 qs[ncols][k] + half2 (d, s)[ncols][k/32], each plane 256-byte aligned
size_t scratch_bytes(const b200_tensor * b) {
    const size_t ncols = (size_t) b->ne[1] * (size_t) b->ne[2] * (size_t) b->ne[3];
    return align_up(ncols * (size_t) b->ne[0], 256) + align_up(ncols * (size_t) (b->ne[0] / 32) * sizeof(half2), 256);
}

}  // namespace

extern "C" {

int b200_mul_mat_noalloc(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * b, const b200_tensor * d) {
    if (!ctx) { b200_set_error("mul_mat_noalloc: ctx is NULL"); return B200_E_ARG; }
    if (!a || !b || !d || !b200_mul_mat_supported(a, b, d)) {
        b200_set_error("mul_mat_noalloc: not a MUL_MAT this path accepts (same rules as b200_mul_mat)");
        return B200_E_ARG;
    }
    const int64_t k = a->ne[0], m = a->ne[1], n = b->ne[1];
    const int64_t batches = b->ne[2] * b->ne[3];
    if (m == 0 || n == 0 || batches == 0) return B200_OK;
    const size_t need = scratch_bytes(b);
    if (need > (size_t) B200_NOALLOC_SCRATCH_BYTES) {
        b200_set_error("mul_mat_noalloc: %lld activation columns of k = %lld need %zu B of scratch, the static maximum is %u B "
                       "(use b200_mul_mat)", (long long) (n * batches), (long long) k, need, (unsigned) B200_NOALLOC_SCRATCH_BYTES);
        return B200_E_ARG;
    }
    const int     nc    = n < 8 ? (int) n : 8;                   // columns per tile, like b200_mul_mat's generic route
    const int64_t tiles = batches * ((n + nc - 1) / nc);
    if (tiles > 65535) { b200_set_error("mul_mat_noalloc: %lld column tiles exceed gridDim.y", (long long) tiles); return B200_E_ARG; }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    unsigned char * scratch = nullptr;
    int rc = scratch_address(ctx->device, &scratch);
    if (rc) return rc;

    const size_t qs_bytes = align_up((size_t) n * batches * k, 256);
    int8_t * qs = (int8_t *) scratch;
    half2 *  ds = (half2 *) (scratch + qs_bytes);

    mm_problem P = {};
    P.W = (const uint8_t *) a->data; P.nb01 = a->nb[1]; P.nb02 = a->nb[2]; P.nb03 = a->nb[3];
    P.m = m; P.k = k;
    P.X = (const float *) b->data; P.nb11 = b->nb[1]; P.nb12 = b->nb[2]; P.nb13 = b->nb[3];
    P.ne11 = n; P.ne12 = b->ne[2]; P.ne13 = b->ne[3];
    P.r2 = b->ne[2] / a->ne[2]; P.r3 = b->ne[3] / a->ne[3];
    P.Y = (float *) d->data;
    P.mode = MM_PLAIN;
    P.act.qs = qs; P.act.ds = ds; P.act.k = k;

    // both kernels are ordinary launches (they wait for everything before them on the stream): the in-flight write tracker of
    // the deferred-wait GEMVs (abi.cu) restarts with this call's output
    ctx->n_pend = 1;
    ctx->pend_lo[0] = (const char *) d->data;
    ctx->pend_hi[0] = (const char *) d->data + (size_t) m * n * batches * sizeof(float);

    rc = b200_launch_quantize_planes(ctx, (const float *) b->data, k, b->ne[1], b->ne[2], b->ne[3], b->nb[1], b->nb[2], b->nb[3], qs, ds);
    if (rc) return rc;
    return b200_launch_gemv_generic(ctx, a->type, P, nc, tiles);
}

}  // extern "C"
