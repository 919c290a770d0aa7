// comm.cu — SURVEY.md §8 row N3: the all-reduce behind ggml_backend_comm_{init, free, allreduce_tensor}
// (typedefs ggml/include/ggml-backend.h:205-208; consumer: the meta / tensor-parallel backend, ggml-backend-meta.cpp:1522-1553, 1949-2088;
// reference implementations: NCCL path ggml-cuda.cu:1184-1255, two-GPU "internal" pipeline :1260-1312).
//
// One process, one context (= device + stream) per participant; tensors[i] lives on participant i and every participant ends up with the
// element-wise F32 sum, in place.  No library collective: peer memory over NVLink / NVSwitch.
//   phase 1 (reduce-scatter): participant r sums chunk r of ALL participants' tensors, read through peer mappings, into its own chunk r
//   phase 2 (all-gather):     participant r copies every other chunk j from participant j (which reduced it) into its own tensor
// ordered by events only: "inputs complete" before phase 1, "chunks reduced" before phase 2, "peers done reading" before a participant's
// stream goes on (the caller may overwrite the tensor next).  Sums are formed in participant order 0..n-1 for every element, so all
// participants hold bit-identical results.  HBM / NVLink-bound: each participant reads n * ne / n + (n - 1) * ne / n elements.
#include "common.cuh"

namespace {

constexpr int COMM_MAX = 8;

struct comm_ptrs { float * p[COMM_MAX]; };

__global__ void __launch_bounds__(256) reduce_chunk_kernel(const comm_ptrs bufs, const int n, const int rank, const int64_t lo, const int64_t hi) {
    for (int64_t i = lo + (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (int64_t) gridDim.x * blockDim.x) {
        float s = bufs.p[0][i];
        for (int p = 1; p < n; ++p) s += bufs.p[p][i];
        bufs.p[rank][i] = s;
    }
}

__global__ void __launch_bounds__(256) gather_chunks_kernel(const comm_ptrs bufs, const int n, const int rank, const int64_t ne, const int64_t chunk) {
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < ne; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t j = i / chunk;
        if (j >= n) j = n - 1;
        if (j != rank) bufs.p[rank][i] = bufs.p[j][i];
    }
}

}  // namespace

struct b200_comm {
    int          n;
    b200_ctx *   ctx[COMM_MAX];
    cudaEvent_t  ev[3][COMM_MAX];          // per phase and participant
};

extern "C" {

int b200_comm_create(b200_ctx * const * ctxs, int n, b200_comm ** out) {
    if (!ctxs || !out || n < 1 || n > COMM_MAX) { b200_set_error("comm_create: 1..%d contexts", COMM_MAX); return B200_E_ARG; }
    for (int i = 0; i < n; ++i) if (!ctxs[i]) { b200_set_error("comm_create: context %d is NULL", i); return B200_E_ARG; }
    device_guard guard;
    // every participant reads every other participant's memory
    for (int i = 0; i < n; ++i) {
        B200_CUDA_TRY(cudaSetDevice(ctxs[i]->device));
        for (int j = 0; j < n; ++j) {
            if (ctxs[j]->device == ctxs[i]->device) continue;
            int can = 0;
            B200_CUDA_TRY(cudaDeviceCanAccessPeer(&can, ctxs[i]->device, ctxs[j]->device));
            if (!can) { b200_set_error("comm_create: devices %d and %d have no peer access", ctxs[i]->device, ctxs[j]->device); return B200_E_UNSUPPORTED; }
            const cudaError_t err = cudaDeviceEnablePeerAccess(ctxs[j]->device, 0);
            if (err != cudaSuccess && err != cudaErrorPeerAccessAlreadyEnabled) {
                b200_set_error("cudaDeviceEnablePeerAccess(%d->%d): %s", ctxs[i]->device, ctxs[j]->device, cudaGetErrorString(err));
                return B200_E_CUDA;
            }
            cudaGetLastError();
        }
    }
    b200_comm * c = (b200_comm *) calloc(1, sizeof(b200_comm));
    if (!c) { b200_set_error("comm_create: out of memory"); return B200_E_ALLOC; }
    c->n = n;
    for (int i = 0; i < n; ++i) {
        c->ctx[i] = ctxs[i];
        if (cudaSetDevice(ctxs[i]->device) != cudaSuccess) { free(c); b200_set_error("comm_create: cudaSetDevice failed"); return B200_E_CUDA; }
        for (int ph = 0; ph < 3; ++ph) {
            if (cudaEventCreateWithFlags(&c->ev[ph][i], cudaEventDisableTiming) != cudaSuccess) { free(c); b200_set_error("comm_create: event creation failed"); return B200_E_CUDA; }
        }
    }
    *out = c;
    return B200_OK;
}

void b200_comm_destroy(b200_comm * c) {
    if (!c) return;
    device_guard guard;
    for (int i = 0; i < c->n; ++i) {
        cudaSetDevice(c->ctx[i]->device);
        for (int ph = 0; ph < 3; ++ph) if (c->ev[ph][i]) cudaEventDestroy(c->ev[ph][i]);
    }
    free(c);
}

int b200_comm_size(const b200_comm * c) { return c ? c->n : 0; }

// data[i]: participant i's F32 tensor of ne elements (16-byte aligned, packed); zero[i] != 0: participant i contributes zeros (the reference
// clears tensors whose node was not computed on that backend, ggml-cuda.cu:1205-1211).  Asynchronous on every participant's stream.
int b200_comm_allreduce(b200_comm * c, void * const * data, const int * zero, int64_t ne) {
    if (!c || !data) { b200_set_error("comm_allreduce: NULL argument"); return B200_E_ARG; }
    if (ne == 0) return B200_OK;
    const int n = c->n;
    comm_ptrs P = {};
    for (int i = 0; i < n; ++i) {
        if (!data[i] || ((uintptr_t) data[i] & 15)) { b200_set_error("comm_allreduce: tensor %d is NULL or not 16-byte aligned", i); return B200_E_ARG; }
        P.p[i] = (float *) data[i];
    }
    device_guard guard;
    if (n == 1) {
        if (zero && zero[0]) { B200_CUDA_TRY(cudaSetDevice(c->ctx[0]->device)); B200_CUDA_TRY(cudaMemsetAsync(data[0], 0, (size_t) ne * 4, c->ctx[0]->stream)); }
        return B200_OK;
    }
    const int64_t chunk = (ne + n - 1) / n;
    auto all_wait = [&](int phase) -> int {            // every stream waits for every participant's event of this phase
        for (int i = 0; i < n; ++i) {
            B200_CUDA_TRY(cudaSetDevice(c->ctx[i]->device));
            for (int j = 0; j < n; ++j) if (j != i) B200_CUDA_TRY(cudaStreamWaitEvent(c->ctx[i]->stream, c->ev[phase][j], 0));
        }
        return B200_OK;
    };
    // inputs complete
    for (int i = 0; i < n; ++i) {
        B200_CUDA_TRY(cudaSetDevice(c->ctx[i]->device));
        if (zero && zero[i]) B200_CUDA_TRY(cudaMemsetAsync(data[i], 0, (size_t) ne * 4, c->ctx[i]->stream));
        B200_CUDA_TRY(cudaEventRecord(c->ev[0][i], c->ctx[i]->stream));
    }
    int rc = all_wait(0);
    if (rc) return rc;
    // phase 1: reduce my chunk
    for (int i = 0; i < n; ++i) {
        const int64_t lo = chunk * i < ne ? chunk * i : ne, hi = lo + chunk < ne ? lo + chunk : ne;
        B200_CUDA_TRY(cudaSetDevice(c->ctx[i]->device));
        if (hi > lo) {
            const int64_t blocks = (hi - lo + 255) / 256, cap = (int64_t) c->ctx[i]->sm_count * 8;
            reduce_chunk_kernel<<<(unsigned) (blocks < cap ? blocks : cap), 256, 0, c->ctx[i]->stream>>>(P, n, i, lo, hi);
            B200_CUDA_TRY(cudaGetLastError());
            c->ctx[i]->launches++;
        }
        B200_CUDA_TRY(cudaEventRecord(c->ev[1][i], c->ctx[i]->stream));
    }
    rc = all_wait(1);
    if (rc) return rc;
    // phase 2: fetch the other chunks
    for (int i = 0; i < n; ++i) {
        B200_CUDA_TRY(cudaSetDevice(c->ctx[i]->device));
        const int64_t blocks = (ne + 255) / 256, cap = (int64_t) c->ctx[i]->sm_count * 8;
        gather_chunks_kernel<<<(unsigned) (blocks < cap ? blocks : cap), 256, 0, c->ctx[i]->stream>>>(P, n, i, ne, chunk);
        B200_CUDA_TRY(cudaGetLastError());
        c->ctx[i]->launches++;
        c->ctx[i]->n_pend = 16;
        B200_CUDA_TRY(cudaEventRecord(c->ev[2][i], c->ctx[i]->stream));
    }
    // nobody's stream goes on (and overwrites its tensor) before its peers have read their chunks from it
    return all_wait(2);
}

}  // extern "C"
