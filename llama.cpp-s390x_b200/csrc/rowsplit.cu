// rowsplit.cu — weight rows sharded over the GPUs of one node, output row slices all-gathered over NVLink.
//
// Replaces the reference's split buffer type + row-split driver (ggml_backend_cuda_split_buffer_type and
// ggml_cuda_op_mul_mat, /root/reference/ggml/src/ggml-cuda/ggml-cuda.cu:855-1144, 1801-2119), which copies dst slices
// back to the main GPU with cudaMemcpy3DPeerAsync.  Here every device ends up with the full output (all-gather), by one
// of three exchange modes:
//   mode 1 (default, decode): FUSED — the GEMV epilogue stores each result straight into every GPU's output through
//           peer mappings over NVLink/NVSwitch (mm_store in mm_problem.cuh); no collective launch at all, only events.
//   mode 0: NCCL — shard results go to a staging block, then a group of ncclBroadcast (= all-gather with unequal counts)
//           over one communicator per device (ncclCommInitAll, single process), then an unstage kernel for n > 1.
//   mode 2: copy engine — cudaMemcpy2DAsync of each staging block into every peer's output (the reference's way).
// Per-row arithmetic is the single-GPU kernel on the shard, so results are bit-identical to the 1-GPU result.
// NCCL is dlopen'ed (libnccl.so.2) so that the library has no link-time dependency on it.
#include "common.cuh"

#include <dlfcn.h>
#include <cstdlib>
#include <cstring>

int b200_mul_mat_ex(b200_ctx * ctx, const b200_tensor * a, const b200_tensor * b, const b200_tensor * d,
                    int n_peers, float * const * peer_Y, int64_t peer_m, int64_t peer_row0, const float * bias = nullptr);

namespace {

// ---- minimal NCCL surface, resolved at run time -----------------------------------------------------------------------------
typedef struct ncclComm * nccl_comm_t;
struct nccl_api {
    void * handle;
    int (*CommInitAll)(nccl_comm_t *, int, const int *);
    int (*CommDestroy)(nccl_comm_t);
    int (*GroupStart)(void);
    int (*GroupEnd)(void);
    int (*Broadcast)(const void *, void *, size_t, int /*ncclDataType_t*/, int, nccl_comm_t, cudaStream_t);
    const char * (*GetErrorString)(int);
};
constexpr int NCCL_FLOAT = 7;   // ncclFloat32 (nccl.h)

bool load_nccl(nccl_api & api) {
    const char * names[] = { "libnccl.so.2", "libnccl.so" };
    for (const char * nm : names) {
        api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (api.handle) break;
    }
    if (!api.handle) return false;
    api.CommInitAll    = (int (*)(nccl_comm_t *, int, const int *)) dlsym(api.handle, "ncclCommInitAll");
    api.CommDestroy    = (int (*)(nccl_comm_t)) dlsym(api.handle, "ncclCommDestroy");
    api.GroupStart     = (int (*)(void)) dlsym(api.handle, "ncclGroupStart");
    api.GroupEnd       = (int (*)(void)) dlsym(api.handle, "ncclGroupEnd");
    api.Broadcast      = (int (*)(const void *, void *, size_t, int, int, nccl_comm_t, cudaStream_t)) dlsym(api.handle, "ncclBroadcast");
    api.GetErrorString = (const char * (*)(int)) dlsym(api.handle, "ncclGetErrorString");
    return api.CommInitAll && api.CommDestroy && api.GroupStart && api.GroupEnd && api.Broadcast;
}

// gathered[concat over g of (n x rows_g)] -> y[n][m]
__global__ void unstage_kernel(const float * __restrict__ gathered, float * __restrict__ y, const int64_t m, const int64_t n,
                               const int64_t lo, const int64_t rows, const int64_t block_off) {
    const int64_t total = rows * n;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t) gridDim.x * blockDim.x) {
        const int64_t c = i / rows, r = i - c * rows;
        y[c * m + lo + r] = gathered[block_off + i];
    }
}

}  // namespace

struct b200_rowsplit {
    int         n_dev;
    int         dev[B200_MAX_PEERS];
    b200_ctx *  ctx[B200_MAX_PEERS];
    float       start[B200_MAX_PEERS + 1];   // cumulative row fractions
    cudaEvent_t ev[B200_MAX_PEERS];
    float *     stage[B200_MAX_PEERS];       // per device: its own shard result [n][rows_i]
    float *     gather[B200_MAX_PEERS];      // per device: concatenation of all shards' blocks (mode 0, n > 1)
    float *     xstage[B200_MAX_PEERS];      // per device: its copy of the main device's activations (b200_rowsplit_mul_mat_main)
    size_t      stage_size[B200_MAX_PEERS], gather_size[B200_MAX_PEERS], xstage_size[B200_MAX_PEERS];
    cudaEvent_t ev_main;                     // on the main context's device
    int         ev_main_dev;
    bool        nccl_tried, nccl_ok;
    nccl_api    nccl;
    nccl_comm_t comm[B200_MAX_PEERS];
};

struct b200_split_weights {
    int     type;
    int64_t k, m;
    void *  shard[B200_MAX_PEERS];
    int64_t lo[B200_MAX_PEERS], hi[B200_MAX_PEERS];
};

static int ensure(int device, float ** buf, size_t * have, size_t want) {
    if (want <= *have) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(device));
    B200_CUDA_TRY(cudaDeviceSynchronize());
    if (*buf) cudaFree(*buf);
    *buf = nullptr; *have = 0;
    cudaError_t err = cudaMalloc((void **) buf, want);
    if (err != cudaSuccess) { cudaGetLastError(); b200_set_error("rowsplit staging alloc: %s", cudaGetErrorString(err)); return B200_E_ALLOC; }
    *have = want;
    return B200_OK;
}

extern "C" {

int b200_rowsplit_create(const int * devices, int n_dev, const float * split, b200_rowsplit ** out) {
    device_guard guard;
    if (!out || !devices || n_dev < 1 || n_dev > B200_MAX_PEERS) { b200_set_error("rowsplit_create: bad arguments"); return B200_E_ARG; }
    *out = nullptr;
    b200_rowsplit * rs = (b200_rowsplit *) calloc(1, sizeof(b200_rowsplit));
    if (!rs) return B200_E_ALLOC;
    rs->n_dev = n_dev;
    float total = 0.0f;
    for (int i = 0; i < n_dev; ++i) total += split ? split[i] : 1.0f;
    float acc = 0.0f;
    for (int i = 0; i < n_dev; ++i) { rs->start[i] = total > 0 ? acc / total : 0.0f; acc += split ? split[i] : 1.0f; }
    rs->start[n_dev] = 1.0f;
    for (int i = 0; i < n_dev; ++i) {
        rs->dev[i] = devices[i];
        int rc = b200_ctx_create(devices[i], nullptr, &rs->ctx[i]);
        if (rc) { b200_rowsplit_destroy(rs); return rc; }
        // the automatic split-K of the tcgen05 GEMM depends on the tile count, i.e. on the shard height: shards never split K, so
        // that every split of a matrix runs one and the same K schedule (the unsplit matrix run with tc_split_k = 0 gives the same bits)
        rs->ctx[i]->tc_split_k = 0;
        cudaError_t err = cudaEventCreateWithFlags(&rs->ev[i], cudaEventDisableTiming);
        if (err != cudaSuccess) { b200_set_error("cudaEventCreate: %s", cudaGetErrorString(err)); b200_rowsplit_destroy(rs); return B200_E_CUDA; }
    }
    for (int i = 0; i < n_dev; ++i) {
        cudaSetDevice(devices[i]);
        for (int j = 0; j < n_dev; ++j) {
            if (i == j || devices[i] == devices[j]) continue;      // (one device listed twice: a split emulated on one GPU)
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[i], devices[j]);
            if (!can) { b200_set_error("devices %d and %d have no peer access", devices[i], devices[j]); b200_rowsplit_destroy(rs); return B200_E_UNSUPPORTED; }
            cudaError_t err = cudaDeviceEnablePeerAccess(devices[j], 0);
            if (err != cudaSuccess && err != cudaErrorPeerAccessAlreadyEnabled) {
                b200_set_error("cudaDeviceEnablePeerAccess(%d->%d): %s", devices[i], devices[j], cudaGetErrorString(err));
                b200_rowsplit_destroy(rs);
                return B200_E_CUDA;
            }
            cudaGetLastError();
        }
    }
    *out = rs;
    return B200_OK;
}

void b200_rowsplit_destroy(b200_rowsplit * rs) {
    device_guard guard;
    if (!rs) return;
    for (int i = 0; i < rs->n_dev; ++i) {
        if (rs->ctx[i]) { cudaSetDevice(rs->dev[i]); cudaDeviceSynchronize(); }
        if (rs->nccl_ok && rs->comm[i]) rs->nccl.CommDestroy(rs->comm[i]);
        if (rs->stage[i]) cudaFree(rs->stage[i]);
        if (rs->gather[i]) cudaFree(rs->gather[i]);
        if (rs->xstage[i]) cudaFree(rs->xstage[i]);
        if (rs->ev[i]) cudaEventDestroy(rs->ev[i]);
        if (rs->ctx[i]) b200_ctx_destroy(rs->ctx[i]);
    }
    if (rs->ev_main) { cudaSetDevice(rs->ev_main_dev); cudaEventDestroy(rs->ev_main); }
    free(rs);
}

int b200_rowsplit_n_dev(const b200_rowsplit * rs) { return rs ? rs->n_dev : 0; }

b200_ctx * b200_rowsplit_ctx(b200_rowsplit * rs, int idx) { return rs && idx >= 0 && idx < rs->n_dev ? rs->ctx[idx] : nullptr; }

int b200_rowsplit_rows(const b200_rowsplit * rs, int64_t m, int idx, int64_t * row_lo, int64_t * row_hi) {
    if (!rs || idx < 0 || idx >= rs->n_dev || m < 0) { b200_set_error("rowsplit_rows: bad arguments"); return B200_E_ARG; }
    auto cut = [&](int i) -> int64_t {
        if (i <= 0) return 0;
        if (i >= rs->n_dev) return m;
        int64_t r = (int64_t) ((double) m * (double) rs->start[i]);
        r -= r % B200_ROWSPLIT_ROUND;
        return r;
    };
    if (row_lo) *row_lo = cut(idx);
    if (row_hi) *row_hi = cut(idx + 1);
    return B200_OK;
}

int b200_rowsplit_upload(b200_rowsplit * rs, int type, const void * host_rows, int64_t k, int64_t m, b200_split_weights ** out) {
    device_guard guard;
    if (!rs || !host_rows || !out || b200_blck(type) == 0 || k % b200_blck(type) != 0) { b200_set_error("rowsplit_upload: bad arguments"); return B200_E_ARG; }
    b200_split_weights * w = (b200_split_weights *) calloc(1, sizeof(b200_split_weights));
    if (!w) return B200_E_ALLOC;
    w->type = type; w->k = k; w->m = m;
    const size_t rb = b200_row_size(type, k);
    for (int i = 0; i < rs->n_dev; ++i) {
        b200_rowsplit_rows(rs, m, i, &w->lo[i], &w->hi[i]);
        const int64_t rows = w->hi[i] - w->lo[i];
        if (rows <= 0) continue;
        int rc = b200_malloc(rs->dev[i], rb * rows + 512, &w->shard[i]);
        if (!rc) rc = b200_memcpy_h2d(rs->ctx[i], w->shard[i], (const char *) host_rows + rb * w->lo[i], rb * rows);
        if (rc) { b200_rowsplit_weights_free(rs, w); return rc; }
    }
    *out = w;
    return B200_OK;
}

void b200_rowsplit_weights_free(b200_rowsplit * rs, b200_split_weights * w) {
    device_guard guard;
    if (!rs || !w) return;
    for (int i = 0; i < rs->n_dev; ++i) if (w->shard[i]) b200_free(rs->dev[i], w->shard[i]);
    free(w);
}

int b200_rowsplit_synchronize(b200_rowsplit * rs) {
    device_guard guard;
    if (!rs) return B200_E_ARG;
    for (int i = 0; i < rs->n_dev; ++i) { int rc = b200_ctx_synchronize(rs->ctx[i]); if (rc) return rc; }
    return B200_OK;
}

int b200_rowsplit_mul_mat(b200_rowsplit * rs, const b200_split_weights * w, float * const * x_dev, int64_t n, float * const * y_dev, int mode) {
    device_guard guard;
    if (!rs || !w || !x_dev || !y_dev || n <= 0 || mode < 0 || mode > 2) { b200_set_error("rowsplit_mul_mat: bad arguments"); return B200_E_ARG; }
    const int G = rs->n_dev;
    if (mode == 0 && !rs->nccl_tried) {
        rs->nccl_tried = true;
        if (load_nccl(rs->nccl)) {
            const int rc = rs->nccl.CommInitAll(rs->comm, G, rs->dev);
            rs->nccl_ok = rc == 0;
            if (!rs->nccl_ok) b200_set_error("ncclCommInitAll: %s", rs->nccl.GetErrorString ? rs->nccl.GetErrorString(rc) : "?");
        } else {
            b200_set_error("libnccl.so.2 not found");
        }
    }
    if (mode == 0 && !rs->nccl_ok) return B200_E_NCCL;
    if (mode == 1 && n > 8) mode = 2;      // the fused store lives in the GEMV family

    const size_t rb = b200_row_size(w->type, w->k);
    size_t block_off[B200_MAX_PEERS + 1] = { 0 };
    for (int i = 0; i < G; ++i) block_off[i + 1] = block_off[i] + (size_t) (w->hi[i] - w->lo[i]) * n;

    // 1) every device computes its shard (into its staging block; mode 1 also stores into every GPU's y)
    for (int i = 0; i < G; ++i) {
        const int64_t rows = w->hi[i] - w->lo[i];
        if (rows <= 0) continue;
        int rc = ensure(rs->dev[i], &rs->stage[i], &rs->stage_size[i], (size_t) rows * n * 4);
        if (rc) return rc;
        b200_tensor a = {}, b = {}, d = {};
        a.data = w->shard[i]; a.type = w->type; a.ne[0] = w->k; a.ne[1] = rows; a.ne[2] = a.ne[3] = 1;
        a.nb[0] = b200_tsize(w->type); a.nb[1] = (int64_t) rb; a.nb[2] = a.nb[3] = (int64_t) rb * rows;
        b.data = x_dev[i]; b.type = B200_TYPE_F32; b.ne[0] = w->k; b.ne[1] = n; b.ne[2] = b.ne[3] = 1;
        b.nb[0] = 4; b.nb[1] = w->k * 4; b.nb[2] = b.nb[3] = w->k * 4 * n;
        d.data = rs->stage[i]; d.type = B200_TYPE_F32; d.ne[0] = rows; d.ne[1] = n; d.ne[2] = d.ne[3] = 1;
        d.nb[0] = 4; d.nb[1] = rows * 4; d.nb[2] = d.nb[3] = rows * 4 * n;
        rc = b200_mul_mat_ex(rs->ctx[i], &a, &b, &d, mode == 1 ? G : 0, y_dev, w->m, w->lo[i]);
        if (rc) return rc;
        rs->ctx[i]->n_pend = 16;       // (outputs written through peer mappings are not in the in-flight tracker: wait early next time)
    }

    // 2) exchange
    if (mode == 0) {
        const bool direct = n == 1;    // row slices are contiguous in y: broadcast in place
        for (int i = 0; i < G && !direct; ++i) {
            int rc = ensure(rs->dev[i], &rs->gather[i], &rs->gather_size[i], block_off[G] * 4);
            if (rc) return rc;
        }
        int nrc = rs->nccl.GroupStart();
        for (int r = 0; r < G && nrc == 0; ++r) {
            const size_t count = block_off[r + 1] - block_off[r];
            if (count == 0) continue;
            for (int i = 0; i < G && nrc == 0; ++i) {
                float * recv = direct ? y_dev[i] + w->lo[r] : rs->gather[i] + block_off[r];
                nrc = rs->nccl.Broadcast(rs->stage[i] /* only the root's matters */, recv, count, NCCL_FLOAT, r, rs->comm[i], rs->ctx[i]->stream);
            }
        }
        const int erc = rs->nccl.GroupEnd();
        if (nrc == 0) nrc = erc;
        if (nrc != 0) { b200_set_error("NCCL all-gather: %s", rs->nccl.GetErrorString ? rs->nccl.GetErrorString(nrc) : "?"); return B200_E_NCCL; }
        if (!direct) {
            for (int i = 0; i < G; ++i) {
                B200_CUDA_TRY(cudaSetDevice(rs->dev[i]));
                for (int r = 0; r < G; ++r) {
                    const int64_t rows = w->hi[r] - w->lo[r];
                    if (rows <= 0) continue;
                    const int64_t total = rows * n;
                    const int grid = (int) ((total + 255) / 256 < 1184 ? (total + 255) / 256 : 1184);
                    unstage_kernel<<<grid, 256, 0, rs->ctx[i]->stream>>>(rs->gather[i], y_dev[i], w->m, n, w->lo[r], rows, (int64_t) block_off[r]);
                    rs->ctx[i]->launches++;
                }
                B200_CUDA_TRY(cudaGetLastError());
            }
        }
        return B200_OK;     // NCCL ordered everything on the per-device streams
    }

    if (mode == 2) {
        for (int r = 0; r < G; ++r) {
            const int64_t rows = w->hi[r] - w->lo[r];
            if (rows <= 0) continue;
            B200_CUDA_TRY(cudaSetDevice(rs->dev[r]));
            for (int i = 0; i < G; ++i) {
                B200_CUDA_TRY(cudaMemcpy2DAsync(y_dev[i] + w->lo[r], (size_t) w->m * 4, rs->stage[r], (size_t) rows * 4, (size_t) rows * 4, (size_t) n,
                                                cudaMemcpyDefault, rs->ctx[r]->stream));
            }
        }
    }
    // modes 1 and 2: every stream waits until every device has finished writing
    for (int i = 0; i < G; ++i) {
        B200_CUDA_TRY(cudaSetDevice(rs->dev[i]));
        B200_CUDA_TRY(cudaEventRecord(rs->ev[i], rs->ctx[i]->stream));
    }
    for (int i = 0; i < G; ++i) {
        B200_CUDA_TRY(cudaSetDevice(rs->dev[i]));
        for (int j = 0; j < G; ++j) if (j != i) B200_CUDA_TRY(cudaStreamWaitEvent(rs->ctx[i]->stream, rs->ev[j], 0));
    }
    return B200_OK;
}

// MUL_MAT whose weights are row-split over the devices while src1 and dst live on ONE device (the contract of the reference's
// split buffer type: ggml_cuda_op_mul_mat broadcasts src1 from the main device and copies the dst row slices back to it,
// ggml-cuda.cu:1801-2119).  Ordered on main_ctx's stream: the other devices' streams wait for it, fetch the activations over
// NVLink, run the single-GPU kernel on their shard — whose epilogue stores the results straight into the main device's dst
// (peer stores, mm_store) — and main_ctx's stream waits for them.
int b200_rowsplit_mul_mat_main(b200_rowsplit * rs, b200_ctx * main_ctx, int type, int64_t k, int64_t m, void * const * shards,
                               const int64_t * row_lo, const int64_t * row_hi, const b200_tensor * src1, const b200_tensor * dst) {
    device_guard guard;
    if (!rs || !main_ctx || !shards || !row_lo || !row_hi || !src1 || !dst) { b200_set_error("rowsplit_mul_mat_main: bad arguments"); return B200_E_ARG; }
    const int64_t n = src1->ne[1];
    if (src1->type != B200_TYPE_F32 || dst->type != B200_TYPE_F32 || src1->ne[0] != k || src1->nb[0] != 4 || src1->ne[2] * src1->ne[3] != 1 ||
        dst->ne[0] != m || dst->ne[1] != n || dst->ne[2] * dst->ne[3] != 1 || dst->nb[0] != 4 || (n > 1 && dst->nb[1] != m * 4)) {
        b200_set_error("rowsplit_mul_mat_main: src1 [k, n] f32 and a contiguous dst [m, n] f32 without batch dims"); return B200_E_UNSUPPORTED;
    }
    if (n == 0 || m == 0) return B200_OK;
    const int G = rs->n_dev;
    const size_t rb = b200_row_size(type, k);
    if (!rs->ev_main || rs->ev_main_dev != main_ctx->device) {
        if (rs->ev_main) { cudaSetDevice(rs->ev_main_dev); cudaEventDestroy(rs->ev_main); rs->ev_main = nullptr; }
        B200_CUDA_TRY(cudaSetDevice(main_ctx->device));
        B200_CUDA_TRY(cudaEventCreateWithFlags(&rs->ev_main, cudaEventDisableTiming));
        rs->ev_main_dev = main_ctx->device;
    }
    B200_CUDA_TRY(cudaSetDevice(main_ctx->device));
    B200_CUDA_TRY(cudaEventRecord(rs->ev_main, main_ctx->stream));
    float * peer[1] = { (float *) dst->data };
    for (int i = 0; i < G; ++i) {
        const int64_t rows = row_hi[i] - row_lo[i];
        if (rows <= 0) continue;
        if (!shards[i]) { b200_set_error("rowsplit_mul_mat_main: shard %d is NULL", i); return B200_E_ARG; }
        b200_ctx * ci = rs->ctx[i];
        int rc = ensure(rs->dev[i], &rs->stage[i], &rs->stage_size[i], (size_t) rows * n * 4);
        if (rc) return rc;
        B200_CUDA_TRY(cudaSetDevice(rs->dev[i]));
        B200_CUDA_TRY(cudaStreamWaitEvent(ci->stream, rs->ev_main, 0));
        b200_tensor a = {}, b = *src1, d = {};
        if (rs->dev[i] != main_ctx->device) {
            rc = ensure(rs->dev[i], &rs->xstage[i], &rs->xstage_size[i], (size_t) k * n * 4);
            if (rc) return rc;
            B200_CUDA_TRY(cudaMemcpy2DAsync(rs->xstage[i], (size_t) k * 4, src1->data, (size_t) (n > 1 ? src1->nb[1] : k * 4), (size_t) k * 4, (size_t) n,
                                            cudaMemcpyDefault, ci->stream));
            b.data = rs->xstage[i]; b.nb[1] = k * 4; b.nb[2] = b.nb[3] = k * 4 * n;
        }
        a.data = shards[i]; a.type = type; a.ne[0] = k; a.ne[1] = rows; a.ne[2] = a.ne[3] = 1;
        a.nb[0] = b200_tsize(type); a.nb[1] = (int64_t) rb; a.nb[2] = a.nb[3] = (int64_t) rb * rows;
        d.data = rs->stage[i]; d.type = B200_TYPE_F32; d.ne[0] = rows; d.ne[1] = n; d.ne[2] = d.ne[3] = 1;
        d.nb[0] = 4; d.nb[1] = rows * 4; d.nb[2] = d.nb[3] = rows * 4 * n;
        ci->n_pend = 16;                                   // its activations arrive by a copy / another stream: wait before reading them
        rc = b200_mul_mat_ex(ci, &a, &b, &d, 1, peer, m, row_lo[i]);
        if (rc) return rc;
        B200_CUDA_TRY(cudaEventRecord(rs->ev[i], ci->stream));
    }
    B200_CUDA_TRY(cudaSetDevice(main_ctx->device));
    for (int i = 0; i < G; ++i) if (row_hi[i] > row_lo[i]) B200_CUDA_TRY(cudaStreamWaitEvent(main_ctx->stream, rs->ev[i], 0));
    main_ctx->n_pend = 16;
    return B200_OK;
}

}  // extern "C"
