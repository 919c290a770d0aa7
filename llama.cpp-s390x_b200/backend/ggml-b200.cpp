// ggml-b200.cpp — the ggml backend shell: a dynamically loadable ggml backend (GGML_BACKEND_DL plug-in) that makes the
// B200 quantized MUL_MAT / MUL_MAT_ID kernels available to unmodified llama.cpp, llama-bench and test-backend-ops.
//
// It implements the reference's four vtables + the backend (stream) vtable
//   ggml_backend_reg_i / ggml_backend_device_i / ggml_backend_buffer_type_i / ggml_backend_buffer_i / ggml_backend_i
//   (/root/reference/ggml/src/ggml-backend-impl.h:17-230)
// and is picked up through GGML_BACKEND_PATH by ggml_backend_load_all() (ggml/src/ggml-backend-reg.cpp:555-586) or
// through ggml_backend_load(path).  All compute and all device memory traffic goes through the C ABI of
// libggml_b200_kernels.so (include/ggml_b200.h); this file holds no arithmetic and no CUDA calls.  ggml-base symbols
// (ggml_backend_buffer_init, ggml_nbytes, ggml_abort, ...) are resolved from the host process at dlopen time, exactly
// like the reference's own dynamically loaded backends.
//
// Supported ops: MUL_MAT and MUL_MAT_ID with src0 in {Q4_0, Q8_0, Q4_K, Q6_K}, F32 src1/dst, plus the no-op view ops.
// Everything else is reported unsupported, so the scheduler places it on another backend (there is NO CPU fallback
// inside this backend).
#include "ggml-backend-impl.h"
#include "ggml-impl.h"
#include "ggml.h"

#include "ggml_b200.h"

#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#define GGML_B200_NAME     "B200"
#define GGML_B200_MAX_DEVS 16

extern "C" {
GGML_BACKEND_API ggml_backend_reg_t ggml_backend_b200_reg(void);
}

namespace {

// ---- per-device state --------------------------------------------------------------------------------------------------
struct b200_device_ctx {
    int         device = 0;
    std::string name;          // "B2000", "B2001", ... (what test-backend-ops -b / llama.cpp -dev match on)
    std::string description;
    b200_ctx *  util = nullptr; // context (stream) used by the buffers' synchronous copies
    std::mutex  mutex;
    ggml_backend_buffer_type buft;
};

struct graph_cache_entry {
    uint64_t key = 0; void * graph = nullptr; uint64_t alloc_gen = 0; int seen = 0; int capture_attempts = 0; bool failed = false; uint64_t stamp = 0;
    // eager launches or replay, whichever is faster for THIS graph on THIS host (GGML_B200_GRAPHS=1): two timed runs of each, then decided
    int   mode = 0;                         // 0 = still measuring, 1 = eager, 2 = replay
    int   eager_runs = 0, replay_runs = 0;  // timed runs started
    int   timing = 0;                       // a timed run is in flight: 1 = eager, 2 = replay
    float t_eager = 1e30f, t_replay = 1e30f;
};

struct b200_backend_ctx {
    int         device;
    std::string name;
    b200_ctx *  ctx;
    long        n_graphs = 0, n_mul_mat = 0, n_mul_mat_id = 0, n_fused_groups = 0;      // GGML_B200_STATS=1 prints them on free
    long        n_extra_ops = 0, n_fused_mm = 0;
    int         fuse = 1;                  // GGML_B200_FUSE: 0 = MUL_MAT + bias / gate + GLU run as separate nodes, 1 = fused when it pays, 2 = always
    bool        reorder = true;            // GGML_B200_REORDER=0: graph_optimize leaves the node order alone
    int         use_graphs = 1;            // GGML_B200_GRAPHS: 0 = every graph_compute launches its kernels one by one, 2 = a graph seen twice is captured
                                           // and replayed, 1 (default) = both are timed for every graph and the faster one is kept
    void *      t0 = nullptr, * t1 = nullptr;          // timed events of the run being measured
    long        n_graphs_eager_kept = 0, n_graphs_replay_kept = 0;
    graph_cache_entry graphs[8];           // captured cgraphs, least recently used replaced
    uint64_t    graph_clock = 0;
    long        n_graph_captures = 0, n_graph_replays = 0;
    void *      copy_event = nullptr;      // orders cpy_tensor_async after this backend's queued work
    b200_rowsplit * rs = nullptr;          // row-split driver of the split buffer type this backend last computed with
    const void *    rs_key = nullptr;
};

struct b200_buffer_ctx {
    int    device;
    void * base;
};

void check(int rc, const char * what) {
    if (rc != B200_OK) {
        GGML_ABORT("ggml-b200: %s failed (%d): %s", what, rc, b200_last_error());
    }
}

b200_ctx * device_util_ctx(b200_device_ctx * dctx) {
    std::lock_guard<std::mutex> lock(dctx->mutex);
    if (!dctx->util) {
        check(b200_ctx_create(dctx->device, nullptr, &dctx->util), "b200_ctx_create");
    }
    return dctx->util;
}

b200_tensor to_b200(const ggml_tensor * t) {
    b200_tensor r = {};
    r.data = t->data;
    r.type = (int32_t) t->type;        // enum ggml_type values are the ABI's type ids
    for (int i = 0; i < 4; ++i) { r.ne[i] = t->ne[i]; r.nb[i] = (int64_t) t->nb[i]; }
    return r;
}

// ---- buffer ----------------------------------------------------------------------------------------------------------------
ggml_backend_dev_t buffer_device(ggml_backend_buffer_t buffer) { return buffer->buft->device; }

void buffer_free(ggml_backend_buffer_t buffer) {
    b200_buffer_ctx * bctx = (b200_buffer_ctx *) buffer->context;
    b200_free(bctx->device, bctx->base);
    delete bctx;
}

void * buffer_get_base(ggml_backend_buffer_t buffer) { return ((b200_buffer_ctx *) buffer->context)->base; }

void buffer_memset_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor, uint8_t value, size_t offset, size_t size) {
    b200_ctx * u = device_util_ctx((b200_device_ctx *) buffer_device(buffer)->context);
    check(b200_memset(u, (char *) tensor->data + offset, value, size), "memset_tensor");
}

void buffer_set_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    b200_ctx * u = device_util_ctx((b200_device_ctx *) buffer_device(buffer)->context);
    check(b200_memcpy_h2d(u, (char *) tensor->data + offset, data, size), "set_tensor");
}

void buffer_get_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    b200_ctx * u = device_util_ctx((b200_device_ctx *) buffer_device(buffer)->context);
    check(b200_memcpy_d2h(u, data, (const char *) tensor->data + offset, size), "get_tensor");
}

bool buffer_is_b200(ggml_backend_buffer_t buffer);

bool buffer_cpy_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * src, ggml_tensor * dst) {
    ggml_backend_buffer_t src_buf = src->view_src ? src->view_src->buffer : src->buffer;
    if (!src_buf) return false;
    b200_ctx * u = device_util_ctx((b200_device_ctx *) buffer_device(buffer)->context);
    if (buffer_is_b200(src_buf)) {
        check(b200_memcpy_d2d(u, dst->data, src->data, ggml_nbytes(src)), "cpy_tensor (device)");
        return true;
    }
    if (ggml_backend_buffer_is_host(src_buf)) {
        check(b200_memcpy_h2d(u, dst->data, src->data, ggml_nbytes(src)), "cpy_tensor (host)");
        return true;
    }
    return false;
}

void buffer_clear(ggml_backend_buffer_t buffer, uint8_t value) {
    b200_buffer_ctx * bctx = (b200_buffer_ctx *) buffer->context;
    b200_ctx * u = device_util_ctx((b200_device_ctx *) buffer_device(buffer)->context);
    check(b200_memset(u, bctx->base, value, buffer->size), "clear");
}

const ggml_backend_buffer_i k_buffer_iface = {
    /* .free_buffer   = */ buffer_free,
    /* .get_base      = */ buffer_get_base,
    /* .init_tensor   = */ nullptr,
    /* .memset_tensor = */ buffer_memset_tensor,
    /* .set_tensor    = */ buffer_set_tensor,
    /* .get_tensor    = */ buffer_get_tensor,
    /* .set_tensor_2d = */ nullptr,
    /* .get_tensor_2d = */ nullptr,
    /* .cpy_tensor    = */ buffer_cpy_tensor,
    /* .clear         = */ buffer_clear,
    /* .reset         = */ nullptr,
};

// ---- buffer type -----------------------------------------------------------------------------------------------------------
const char * buft_get_name(ggml_backend_buffer_type_t buft) { return ((b200_device_ctx *) buft->device->context)->name.c_str(); }

ggml_backend_buffer_t buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    b200_device_ctx * dctx = (b200_device_ctx *) buft->device->context;
    void * base = nullptr;
    if (b200_malloc(dctx->device, size, &base) != B200_OK) {
        GGML_LOG_ERROR("ggml-b200: allocating %.2f MiB on device %d failed: %s\n", size / 1024.0 / 1024.0, dctx->device, b200_last_error());
        return nullptr;                    // callers check (ggml-backend.cpp:2106, test-backend-ops.cpp:1338)
    }
    return ggml_backend_buffer_init(buft, k_buffer_iface, new b200_buffer_ctx{dctx->device, base}, size);
}

size_t buft_get_alignment(ggml_backend_buffer_type_t) { return 128; }

// rows of quantized tensors are padded to a multiple of 512 elements so tiles may over-read in bounds
// (same contract as ggml-cuda.cu:802-816)
size_t buft_get_alloc_size(ggml_backend_buffer_type_t, const ggml_tensor * tensor) {
    size_t size = ggml_nbytes(tensor);
    const int64_t ne0 = tensor->ne[0];
    if (ggml_is_quantized(tensor->type) && ne0 % 512 != 0) {
        size += ggml_row_size(tensor->type, 512 - ne0 % 512);
    }
    return size;
}

bool buft_is_host(ggml_backend_buffer_type_t) { return false; }

const ggml_backend_buffer_type_i k_buft_iface = {
    /* .get_name       = */ buft_get_name,
    /* .alloc_buffer   = */ buft_alloc_buffer,
    /* .get_alignment  = */ buft_get_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ buft_get_alloc_size,
    /* .is_host        = */ buft_is_host,
};

// identify our buffers by their buffer type's vtable, never by context: ggml-base fabricates 0-byte buffers with a NULL
// context for weight placement probes (ggml-backend.cpp:38-45, llama-model-loader.cpp:919-931)
bool buft_is_b200(ggml_backend_buffer_type_t buft) { return buft && buft->iface.get_name == buft_get_name; }
bool buffer_is_b200(ggml_backend_buffer_t buffer) { return buffer && buft_is_b200(buffer->buft); }
bool buft_is_split(ggml_backend_buffer_type_t buft);

// ---- split buffer type ("ggml_backend_split_buffer_type": llama.cpp -sm row) --------------------------------------------------
// Replaces ggml_backend_cuda_split_buffer_type (ggml-cuda.cu:855-1144): a weight matrix allocated in this buffer type is cut by
// rows — cut points round_down(m * cumulative tensor_split, 128), get_row_split :868-881 — and shard i lives on device i.  The
// buffer's base address is a dummy (get_base :921-926); init_tensor allocates the shards (:928-976), set_tensor scatters host
// rows (:978-1015), get_tensor gathers them.  MUL_MAT with such a src0 runs on every device at once (b200_rowsplit_mul_mat_main);
// MUL_MAT_ID with split weights is not supported, like in the reference (:2628).
// GGML_B200_SPLIT_EMULATE=N (testing): N shards on ONE physical device, so a single-GPU box exercises the whole path.
struct b200_split_buft_ctx {
    int         main_device = 0;
    int         n_shards = 0;
    int         device[GGML_B200_MAX_DEVS];       // physical device of shard i
    float       cum[GGML_B200_MAX_DEVS + 1];      // cumulative row fractions
    std::string name;
    ggml_backend_buffer_type buft;
};

struct b200_split_extra {
    void *  shard[GGML_B200_MAX_DEVS] = {};
    int64_t lo[GGML_B200_MAX_DEVS] = {}, hi[GGML_B200_MAX_DEVS] = {};
};

struct b200_split_buffer_ctx {
    b200_split_buft_ctx * bt;
    std::vector<b200_split_extra *> extras;
};

const char * split_buft_get_name(ggml_backend_buffer_type_t buft) { return ((b200_split_buft_ctx *) buft->context)->name.c_str(); }
bool buft_is_split(ggml_backend_buffer_type_t buft) { return buft && buft->iface.get_name == split_buft_get_name; }

b200_device_ctx * reg_device_ctx(int device) {
    ggml_backend_reg_t reg = ggml_backend_b200_reg();
    return (b200_device_ctx *) ggml_backend_reg_dev_get(reg, (size_t) device)->context;
}

void split_rows(const b200_split_buft_ctx * bt, int64_t m, int i, int64_t * lo, int64_t * hi) {
    auto cut = [&](int j) -> int64_t {
        if (j <= 0) return 0;
        if (j >= bt->n_shards) return m;
        int64_t r = (int64_t) ((double) m * (double) bt->cum[j]);
        return r - r % B200_ROWSPLIT_ROUND;
    };
    *lo = cut(i); *hi = cut(i + 1);
}

size_t split_shard_bytes(const ggml_tensor * t, int64_t rows) {
    size_t size = ggml_row_size(t->type, t->ne[0]) * (size_t) rows;
    if (t->ne[0] % 512 != 0) size += ggml_row_size(t->type, 512 - t->ne[0] % 512);        // the kernels may over-read the last row in bounds
    return size;
}

void split_buffer_free(ggml_backend_buffer_t buffer) {
    b200_split_buffer_ctx * bc = (b200_split_buffer_ctx *) buffer->context;
    for (b200_split_extra * e : bc->extras) {
        for (int i = 0; i < bc->bt->n_shards; ++i) if (e->shard[i]) b200_free(bc->bt->device[i], e->shard[i]);
        delete e;
    }
    delete bc;
}

void * split_buffer_get_base(ggml_backend_buffer_t) { return (void *) 0x1000; }      // tensors in a split buffer have no single address

ggml_status split_buffer_init_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor) {
    GGML_ASSERT(tensor->view_src == nullptr);                       // views of split tensors are not supported (ggml-cuda.cu:929)
    GGML_ASSERT(ggml_is_contiguous(tensor) && tensor->ne[2] * tensor->ne[3] == 1 && "split tensors must be contiguous 2-D matrices");
    b200_split_buffer_ctx * bc = (b200_split_buffer_ctx *) buffer->context;
    b200_split_extra * e = new b200_split_extra;
    bc->extras.push_back(e);
    for (int i = 0; i < bc->bt->n_shards; ++i) {
        split_rows(bc->bt, tensor->ne[1], i, &e->lo[i], &e->hi[i]);
        const int64_t rows = e->hi[i] - e->lo[i];
        if (rows <= 0) continue;
        const size_t bytes = split_shard_bytes(tensor, rows);
        if (b200_malloc(bc->bt->device[i], bytes + 512, &e->shard[i]) != B200_OK) {
            GGML_LOG_ERROR("ggml-b200: split shard of %s on device %d: %s\n", tensor->name, bc->bt->device[i], b200_last_error());
            return GGML_STATUS_ALLOC_FAILED;
        }
        // the padding past the last row must not hold NaN patterns (it is multiplied by activations that are zero there)
        const size_t body = ggml_row_size(tensor->type, tensor->ne[0]) * (size_t) rows;
        check(b200_memset(device_util_ctx(reg_device_ctx(bc->bt->device[i])), (char *) e->shard[i] + body, 0, bytes + 512 - body), "split pad");
    }
    tensor->extra = e;
    return GGML_STATUS_SUCCESS;
}

void split_buffer_set_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    GGML_ASSERT(offset == 0 && size == ggml_nbytes(tensor) && "split tensors are set as a whole (ggml-cuda.cu:979-981)");
    b200_split_buffer_ctx * bc = (b200_split_buffer_ctx *) buffer->context;
    b200_split_extra * e = (b200_split_extra *) tensor->extra;
    const size_t rb = ggml_row_size(tensor->type, tensor->ne[0]);
    for (int i = 0; i < bc->bt->n_shards; ++i) {
        const int64_t rows = e->hi[i] - e->lo[i];
        if (rows <= 0) continue;
        check(b200_memcpy_h2d(device_util_ctx(reg_device_ctx(bc->bt->device[i])), e->shard[i], (const char *) data + rb * e->lo[i], rb * rows), "split set_tensor");
    }
}

void split_buffer_get_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    GGML_ASSERT(offset == 0 && size == ggml_nbytes(tensor));
    b200_split_buffer_ctx * bc = (b200_split_buffer_ctx *) buffer->context;
    const b200_split_extra * e = (const b200_split_extra *) tensor->extra;
    const size_t rb = ggml_row_size(tensor->type, tensor->ne[0]);
    for (int i = 0; i < bc->bt->n_shards; ++i) {
        const int64_t rows = e->hi[i] - e->lo[i];
        if (rows <= 0) continue;
        check(b200_memcpy_d2h(device_util_ctx(reg_device_ctx(bc->bt->device[i])), (char *) data + rb * e->lo[i], e->shard[i], rb * rows), "split get_tensor");
    }
}

void split_buffer_clear(ggml_backend_buffer_t, uint8_t) {}            // (no-op in the reference too, ggml-cuda.cu:1051-1054)

const ggml_backend_buffer_i k_split_buffer_iface = {
    /* .free_buffer   = */ split_buffer_free,
    /* .get_base      = */ split_buffer_get_base,
    /* .init_tensor   = */ split_buffer_init_tensor,
    /* .memset_tensor = */ nullptr,
    /* .set_tensor    = */ split_buffer_set_tensor,
    /* .get_tensor    = */ split_buffer_get_tensor,
    /* .set_tensor_2d = */ nullptr,
    /* .get_tensor_2d = */ nullptr,
    /* .cpy_tensor    = */ nullptr,
    /* .clear         = */ split_buffer_clear,
    /* .reset         = */ nullptr,
};

ggml_backend_buffer_t split_buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    // the shards are allocated tensor by tensor in init_tensor; `size` (the sum of get_alloc_size) is only accounted for
    return ggml_backend_buffer_init(buft, k_split_buffer_iface, new b200_split_buffer_ctx{(b200_split_buft_ctx *) buft->context, {}}, size);
}

size_t split_buft_get_alloc_size(ggml_backend_buffer_type_t buft, const ggml_tensor * tensor) {
    const b200_split_buft_ctx * bt = (const b200_split_buft_ctx *) buft->context;
    size_t total = 0;
    for (int i = 0; i < bt->n_shards; ++i) {
        int64_t lo, hi;
        split_rows(bt, tensor->ne[1], i, &lo, &hi);
        if (hi > lo) total += split_shard_bytes(tensor, hi - lo);
    }
    return total;
}

const ggml_backend_buffer_type_i k_split_buft_iface = {
    /* .get_name       = */ split_buft_get_name,
    /* .alloc_buffer   = */ split_buft_alloc_buffer,
    /* .get_alignment  = */ buft_get_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ split_buft_get_alloc_size,
    /* .is_host        = */ buft_is_host,
};

// the proc address llama.cpp asks for (src/llama-model.cpp:885-897): one buffer type object per (main device, split)
ggml_backend_buffer_type_t split_buffer_type(int main_device, const float * tensor_split) {
    static std::mutex mutex;
    static std::map<std::pair<int, std::vector<float>>, b200_split_buft_ctx *> cache;
    std::lock_guard<std::mutex> lock(mutex);
    const int n_phys = (int) ggml_backend_reg_dev_count(ggml_backend_b200_reg());
    if (n_phys < 1 || main_device < 0 || main_device >= n_phys) return nullptr;
    int emulate = 0;
    if (const char * e = getenv("GGML_B200_SPLIT_EMULATE")) emulate = atoi(e);
    const int n = emulate > 1 ? (emulate < 8 ? emulate : 8) : (n_phys < 8 ? n_phys : 8);
    std::vector<float> split((size_t) n, 0.0f);
    bool all_zero = true;
    for (int i = 0; i < n; ++i) { split[i] = tensor_split ? tensor_split[i] : 0.0f; all_zero = all_zero && split[i] == 0.0f; }
    if (all_zero) for (int i = 0; i < n; ++i) split[i] = 1.0f;          // (the reference defaults to the devices' memory sizes: equal here)
    auto key = std::make_pair(main_device, split);
    auto it = cache.find(key);
    if (it != cache.end()) return &it->second->buft;
    b200_split_buft_ctx * bt = new b200_split_buft_ctx;
    bt->main_device = main_device; bt->n_shards = n;
    float total = 0.0f, acc = 0.0f;
    for (int i = 0; i < n; ++i) total += split[i];
    for (int i = 0; i < n; ++i) { bt->device[i] = emulate > 1 ? main_device : i; bt->cum[i] = acc / total; acc += split[i]; }
    bt->cum[n] = 1.0f;
    bt->name = GGML_B200_NAME "_Split";
    bt->buft = ggml_backend_buffer_type{ k_split_buft_iface, ggml_backend_reg_dev_get(ggml_backend_b200_reg(), (size_t) main_device), bt };
    cache[key] = bt;
    return &bt->buft;
}

// src0 of a MUL_MAT node lives in a split buffer
bool split_src0(const ggml_tensor * node) {
    const ggml_tensor * a = node->src[0];
    return a && a->buffer && buft_is_split(a->buffer->buft);
}

int split_mul_mat(b200_backend_ctx * bctx, const ggml_tensor * node) {
    const ggml_tensor * a = node->src[0];
    const b200_split_buft_ctx * bt = (const b200_split_buft_ctx *) a->buffer->buft->context;
    const b200_split_extra * e = (const b200_split_extra *) a->extra;
    if (bctx->rs_key != bt) {
        if (bctx->rs) { b200_rowsplit_destroy(bctx->rs); bctx->rs = nullptr; }
        const int rc = b200_rowsplit_create(bt->device, bt->n_shards, nullptr, &bctx->rs);
        if (rc != B200_OK) return rc;
        bctx->rs_key = bt;
    }
    const b200_tensor b = to_b200(node->src[1]), d = to_b200(node);
    bctx->n_mul_mat++;
    return b200_rowsplit_mul_mat_main(bctx->rs, bctx->ctx, (int) a->type, a->ne[0], a->ne[1], e->shard, e->lo, e->hi, &b, &d);
}

// ---- backend (stream) -------------------------------------------------------------------------------------------------------
const char * backend_get_name(ggml_backend_t backend) { return ((b200_backend_ctx *) backend->context)->name.c_str(); }

void backend_free(ggml_backend_t backend) {
    b200_backend_ctx * bctx = (b200_backend_ctx *) backend->context;
    if (const char * e = getenv("GGML_B200_STATS"); e && *e == '1') {
        fprintf(stderr, "ggml-b200 stats: graphs=%ld mul_mat=%ld mul_mat_id=%ld fused_groups=%ld other_ops=%ld fused_mul_mat=%ld graph_captures=%ld graph_replays=%ld kept_eager=%ld kept_replay=%ld\n", bctx->n_graphs,
                bctx->n_mul_mat, bctx->n_mul_mat_id, bctx->n_fused_groups, bctx->n_extra_ops, bctx->n_fused_mm, bctx->n_graph_captures, bctx->n_graph_replays,
                bctx->n_graphs_eager_kept, bctx->n_graphs_replay_kept);
        for (const auto & c : bctx->graphs) if (c.mode) fprintf(stderr, "ggml-b200 stats: graph %016llx: eager %.3f ms, replay %.3f ms -> %s\n", (unsigned long long) c.key, c.t_eager, c.t_replay, c.mode == 1 ? "eager" : "replay");
    }
    if (bctx->t0) b200_event_destroy(bctx->device, bctx->t0);
    if (bctx->t1) b200_event_destroy(bctx->device, bctx->t1);
    for (auto & c : bctx->graphs) if (c.graph) b200_graph_free(c.graph);
    if (bctx->copy_event) b200_event_destroy(bctx->device, bctx->copy_event);
    if (bctx->rs) b200_rowsplit_destroy(bctx->rs);
    b200_ctx_destroy(bctx->ctx);
    delete bctx;
    delete backend;
}

void backend_synchronize(ggml_backend_t backend) {
    check(b200_ctx_synchronize(((b200_backend_ctx *) backend->context)->ctx), "synchronize");
}

void backend_set_tensor_async(ggml_backend_t backend, ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    check(b200_memcpy_async(((b200_backend_ctx *) backend->context)->ctx, (char *) tensor->data + offset, data, size, 0), "set_tensor_async");
}

void backend_get_tensor_async(ggml_backend_t backend, const ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    check(b200_memcpy_async(((b200_backend_ctx *) backend->context)->ctx, data, (const char *) tensor->data + offset, size, 1), "get_tensor_async");
}

bool is_view_op(ggml_op op) {
    return op == GGML_OP_NONE || op == GGML_OP_RESHAPE || op == GGML_OP_VIEW || op == GGML_OP_PERMUTE || op == GGML_OP_TRANSPOSE;
}

// ---- the other ops of a Llama graph: which forms the kernels take, and their dispatch ----------------
bool is_f32(const ggml_tensor * t) { return t && t->type == GGML_TYPE_F32; }
bool is_float_type(ggml_type t) { return t == GGML_TYPE_F32 || t == GGML_TYPE_F16 || t == GGML_TYPE_BF16; }
bool is_quant_weight(ggml_type t) {
    return t == GGML_TYPE_Q4_0 || t == GGML_TYPE_Q8_0 || t == GGML_TYPE_Q4_K || t == GGML_TYPE_Q6_K ||
           t == GGML_TYPE_Q4_1 || t == GGML_TYPE_Q5_0 || t == GGML_TYPE_Q5_1 || t == GGML_TYPE_Q5_K ||           // (N4: generic GEMV family; Q5_K also tcgen05)
           t == GGML_TYPE_Q2_K || t == GGML_TYPE_Q3_K || t == GGML_TYPE_IQ4_NL || t == GGML_TYPE_IQ4_XS || t == GGML_TYPE_MXFP4 || t == GGML_TYPE_NVFP4;
}

// GGML_B200_OPS=mul_mat: answer supports_op() for MUL_MAT / MUL_MAT_ID only (the round-1 behaviour: every other node of a graph runs
// on another backend, the scheduler copies activations across — kept so that split structure stays tested)
bool only_mul_mat() {
    static const bool v = [] { const char * e = getenv("GGML_B200_OPS"); return e && !strcmp(e, "mul_mat"); }();
    return v;
}

// GGML_B200_NO_OP=ADD,ROPE,...: decline single ops (bisecting a graph difference)
bool op_disabled(const ggml_tensor * op) {
    static const std::string list = [] { const char * e = getenv("GGML_B200_NO_OP"); return std::string(e ? e : ""); }();
    if (list.empty()) return false;
    const std::string name = std::string(",") + ggml_op_name(op->op) + ",";
    return (std::string(",") + list + ",").find(name) != std::string::npos;
}

bool extra_op_supported(const ggml_tensor * op) {
    if (only_mul_mat() || op_disabled(op)) return false;
    const ggml_tensor * a = op->src[0], * b = op->src[1], * c = op->src[2];
    switch (op->op) {
        case GGML_OP_ADD: case GGML_OP_SUB: case GGML_OP_MUL: case GGML_OP_DIV:
            return is_f32(a) && is_f32(b) && is_f32(op) && ggml_can_repeat(b, a) && ggml_are_same_shape(a, op);
        case GGML_OP_REPEAT:
            return is_f32(a) && is_f32(op) && ggml_can_repeat(a, op);
        case GGML_OP_ADD_ID:
            return is_f32(a) && is_f32(b) && is_f32(op) && c && c->type == GGML_TYPE_I32 && a->nb[0] == 4 && b->nb[0] == 4 && op->nb[0] == 4;
        case GGML_OP_GLU:
            return is_f32(a) && is_f32(op) && (!b || (is_f32(b) && ggml_is_contiguous_1(b))) && ggml_is_contiguous_1(a) && ggml_is_contiguous_1(op);
        case GGML_OP_UNARY:
            return is_f32(a) && is_f32(op) && ggml_is_contiguous(a) && ggml_is_contiguous(op) && b200_op_unary_supported((int) ggml_get_unary_op(op));
        case GGML_OP_SCALE:
            return is_f32(a) && is_f32(op) && ggml_is_contiguous(a) && ggml_is_contiguous(op);
        case GGML_OP_RMS_NORM:
            return is_f32(a) && is_f32(op) && a->nb[0] == 4 && op->nb[0] == 4;
        case GGML_OP_ROPE: {
            const int mode = ((const int32_t *) op->op_params)[2];
            if (mode != GGML_ROPE_TYPE_NORMAL && mode != GGML_ROPE_TYPE_NEOX) return false;
            if (!a || (a->type != GGML_TYPE_F32 && a->type != GGML_TYPE_F16) || op->type != a->type || !b || b->type != GGML_TYPE_I32) return false;
            if (c && c->type != GGML_TYPE_F32) return false;
            return a->nb[0] == ggml_type_size(a->type) && op->nb[0] == ggml_type_size(a->type) && (a->ne[0] % 2) == 0;
        }
        case GGML_OP_SOFT_MAX:
            return is_f32(a) && is_f32(op) && a->nb[0] == 4 && ggml_is_contiguous(op) &&
                   (!b || ((b->type == GGML_TYPE_F16 || b->type == GGML_TYPE_F32) && b->nb[0] == ggml_type_size(b->type))) && (!c || is_f32(c));
        case GGML_OP_CPY: case GGML_OP_CONT: case GGML_OP_DUP:
            return a && is_float_type(a->type) && is_float_type(op->type) && ggml_nelements(a) == ggml_nelements(op);
        case GGML_OP_GET_ROWS:
            return a && (is_float_type(a->type) || is_quant_weight(a->type)) && b && b->type == GGML_TYPE_I32 && is_f32(op) && op->nb[0] == 4;
        case GGML_OP_SET_ROWS:
            return is_f32(a) && a->nb[0] == 4 && b && (b->type == GGML_TYPE_I64 || b->type == GGML_TYPE_I32) && is_float_type(op->type);
        default:
            return false;
    }
}

int run_extra_op(b200_backend_ctx * bctx, const ggml_tensor * node) {
    b200_ctx * ctx = bctx->ctx;
    static const bool sync_ops = [] { const char * e = getenv("GGML_B200_SYNC_OPS"); return e && *e == '1'; }();      // debugging aid
    if (sync_ops) b200_ctx_synchronize(ctx);
    static const bool trace_ops = [] { const char * e = getenv("GGML_B200_TRACE_OPS"); return e && *e == '1'; }();    // debugging aid
    if (trace_ops) {
        fprintf(stderr, "ggml-b200 op %s %s dst %p ne [%ld %ld %ld %ld] nb [%zu %zu %zu %zu]", ggml_op_name(node->op), node->name, node->data,
                (long) node->ne[0], (long) node->ne[1], (long) node->ne[2], (long) node->ne[3], node->nb[0], node->nb[1], node->nb[2], node->nb[3]);
        for (int j = 0; j < 3 && node->src[j]; ++j) {
            const ggml_tensor * t = node->src[j];
            fprintf(stderr, " | src%d %s %p ne [%ld %ld %ld %ld] nb [%zu %zu %zu %zu]", j, t->name, t->data, (long) t->ne[0], (long) t->ne[1], (long) t->ne[2],
                    (long) t->ne[3], t->nb[0], t->nb[1], t->nb[2], t->nb[3]);
        }
        fprintf(stderr, "\n");
    }
    const b200_tensor d = to_b200(node);
    const b200_tensor a = node->src[0] ? to_b200(node->src[0]) : b200_tensor{};
    const b200_tensor b = node->src[1] ? to_b200(node->src[1]) : b200_tensor{};
    const b200_tensor c = node->src[2] ? to_b200(node->src[2]) : b200_tensor{};
    const float * fp = (const float *) node->op_params;
    switch (node->op) {
        case GGML_OP_ADD:      return b200_op_binary(ctx, B200_OP_ADD, &a, &b, &d);
        case GGML_OP_SUB:      return b200_op_binary(ctx, B200_OP_SUB, &a, &b, &d);
        case GGML_OP_MUL:      return b200_op_binary(ctx, B200_OP_MUL, &a, &b, &d);
        case GGML_OP_DIV:      return b200_op_binary(ctx, B200_OP_DIV, &a, &b, &d);
        case GGML_OP_REPEAT:   return b200_op_repeat(ctx, &a, &d);
        case GGML_OP_ADD_ID:   return b200_op_add_id(ctx, &a, &b, &c, &d);
        case GGML_OP_GLU:      return b200_op_glu(ctx, (int) ggml_get_glu_op(node), &a, node->src[1] ? &b : nullptr, &d, ggml_get_op_params_i32(node, 1),
                                                  ggml_get_op_params_f32(node, 2), ggml_get_op_params_f32(node, 3));
        case GGML_OP_UNARY:    return b200_op_unary(ctx, (int) ggml_get_unary_op(node), &a, &d);
        case GGML_OP_SCALE:    return b200_op_scale(ctx, &a, &d, fp[0], fp[1]);
        case GGML_OP_RMS_NORM: return b200_op_rms_norm(ctx, &a, &d, fp[0]);
        case GGML_OP_ROPE:     return b200_op_rope(ctx, &a, &b, node->src[2] ? &c : nullptr, &d, (const int32_t *) node->op_params);
        case GGML_OP_SOFT_MAX: return b200_op_soft_max(ctx, &a, node->src[1] ? &b : nullptr, node->src[2] ? &c : nullptr, &d, fp[0], fp[1]);
        case GGML_OP_CPY: case GGML_OP_CONT: case GGML_OP_DUP:
                               return b200_op_cpy(ctx, &a, &d);
        case GGML_OP_GET_ROWS: return b200_op_get_rows(ctx, &a, &b, &d);
        case GGML_OP_SET_ROWS: return b200_op_set_rows(ctx, &a, &b, &d);
        default:               return B200_E_UNSUPPORTED;
    }
}

// N2 (reference: ggml_cuda_should_fuse_mul_mat_vec_q, ggml-cuda.cu:2499-2535): the nodes starting at `i` are
//   up = MUL_MAT(Wu, x) [, ADD(up, bias)] , gate = MUL_MAT(Wg, x) [, ADD(gate, bias)] , GLU(gate, up)        -> one launch
//   y  = MUL_MAT(W, x) , ADD(y, bias)                                                                           -> one launch
// with one activation column per batch, quantized weights, and every intermediate used by the next node of the pattern only.
// Returns the number of nodes consumed (0: no fusion), the arguments in `f`.
struct fused_mm { const ggml_tensor * up, * gate, * x, * bias_up, * bias_gate, * out; int glu_op; bool add_only; };

bool bias_of(const ggml_tensor * add, const ggml_tensor * mm, const ggml_tensor ** bias) {
    if (add->op != GGML_OP_ADD) return false;
    const ggml_tensor * other = add->src[0] == mm ? add->src[1] : (add->src[1] == mm ? add->src[0] : nullptr);
    if (!other || add->src[0] != mm) return false;               // (the kernel adds into the product: keep src0 = product, as llama.cpp builds it)
    if (!is_f32(other) || !ggml_are_same_shape(other, mm) || other->nb[0] != 4) return false;
    *bias = other;
    return true;
}

int try_fuse_mul_mat(const ggml_cgraph * cgraph, int i, fused_mm & f, int mode) {
    if (only_mul_mat()) return 0;
    const ggml_tensor * n0 = cgraph->nodes[i];
    if (n0->op != GGML_OP_MUL_MAT || !is_quant_weight(n0->src[0]->type) || n0->ne[1] > 8 || split_src0(n0)) return 0;
    f = {};
    auto single_use = [&](int idx) { return ggml_node_has_n_uses(cgraph, idx, 1) && (cgraph->nodes[idx]->flags & GGML_TENSOR_FLAG_OUTPUT) == 0; };
    // first product (+ bias)
    int j = i + 1;
    const ggml_tensor * out1 = n0, * bias1 = nullptr;
    if (j < cgraph->n_nodes && single_use(i) && bias_of(cgraph->nodes[j], n0, &bias1)) { out1 = cgraph->nodes[j]; ++j; }
    const int after_first = j;
    // second product against the same activations (+ bias), then the GLU of the two (llama.cpp's build_ffn computes up first, other graphs
    // gate first: the GLU's src0 says which one is the gate)
    if (n0->ne[1] == 1 && j < cgraph->n_nodes && single_use(j - 1)) {
        const ggml_tensor * m1 = cgraph->nodes[j];
        if (m1->op == GGML_OP_MUL_MAT && m1->src[1] == n0->src[1] && m1->src[0]->type == n0->src[0]->type && ggml_are_same_shape(m1->src[0], n0->src[0]) &&
            ggml_are_same_stride(m1->src[0], n0->src[0]) && !split_src0(m1)) {
            const ggml_tensor * out2 = m1, * bias2 = nullptr;
            int k = j + 1;
            if (k < cgraph->n_nodes && single_use(j) && bias_of(cgraph->nodes[k], m1, &bias2)) { out2 = cgraph->nodes[k]; ++k; }
            if (k < cgraph->n_nodes && single_use(k - 1)) {
                const ggml_tensor * glu = cgraph->nodes[k];
                if (glu->op == GGML_OP_GLU && ggml_get_glu_op(glu) != GGML_GLU_OP_SWIGLU_OAI && glu->src[1] != nullptr) {
                    if (glu->src[0] == out2 && glu->src[1] == out1) {
                        f.up = n0->src[0]; f.bias_up = bias1; f.gate = m1->src[0]; f.bias_gate = bias2;
                    } else if (glu->src[0] == out1 && glu->src[1] == out2) {
                        f.up = m1->src[0]; f.bias_up = bias2; f.gate = n0->src[0]; f.bias_gate = bias1;
                    }
                    if (f.gate) { f.glu_op = (int) ggml_get_glu_op(glu); f.out = glu; j = k + 1; }
                }
            }
        }
    }
    if (!f.gate) {
        if (!bias1) return 0;
        f.up = n0->src[0]; f.bias_up = bias1; f.out = out1; j = after_first;
    }
    f.x = n0->src[1];
    const b200_tensor up = to_b200(f.up), x = to_b200(f.x), d = to_b200(f.out);
    const b200_tensor gt = f.gate ? to_b200(f.gate) : b200_tensor{}, bu = f.bias_up ? to_b200(f.bias_up) : b200_tensor{}, bg = f.bias_gate ? to_b200(f.bias_gate) : b200_tensor{};
    // MUL_MAT + ADD alone: the addend goes into the store of the fast GEMV kernels (any matrix size, up to 8 columns)
    if (!f.gate && b200_mul_mat_add_supported(&up, &x, &bu, &d)) { f.add_only = true; return j - i; }
    if (n0->ne[1] != 1) return 0;
    if (!b200_mul_mat_fused_supported(&up, f.gate ? &gt : nullptr, &x, f.bias_up ? &bu : nullptr, f.bias_gate ? &bg : nullptr, f.glu_op, &d)) return 0;
    // The fused launch walks the rows with the generic one-warp-per-row GEMV; the unfused path has the fast kernels (0.9 of the HBM peak on
    // Q4_K).  Measured on the Llama-3-8B decode graph (oracle/llama_step, 1 token, 512 past): everything fused 4.12 ms / step, nothing fused
    // 3.76 ms — so by default only matrices small enough for the saved launches to outweigh the slower rows are fused.
    if (mode == 1 && (ggml_nbytes(f.up) + (f.gate ? ggml_nbytes(f.gate) : 0)) > ((size_t) 8 << 20)) return 0;
    return j - i;
}

// graph_compute: every maximal run of MUL_MAT nodes (view ops in between need no work) goes to the kernels as ONE list,
// b200_mul_mat_batch — the list is executed as if in node order (dependencies between its nodes are found from the tensors'
// address ranges), runs of Q4_K decode matmuls become one persistent launch, nodes that share their activations (wq / wk / wv,
// ffn_gate / ffn_up of a llama graph) share one quantization of them.  The reference's CUDA backend reaches the same end by
// capturing the whole cgraph into a CUDA graph (ggml-cuda.cu:4457-4490).
ggml_status graph_compute_eager(b200_backend_ctx * bctx, ggml_cgraph * cgraph) {
    std::vector<b200_tensor> ta, tb, td;
    std::vector<const b200_tensor *> pa, pb, pd;
    auto flush = [&]() -> int {
        const int n = (int) ta.size();
        if (n == 0) return B200_OK;
        pa.resize(n); pb.resize(n); pd.resize(n);
        for (int j = 0; j < n; ++j) { pa[j] = &ta[j]; pb[j] = &tb[j]; pd[j] = &td[j]; }
        const int rc = n == 1 ? b200_mul_mat(bctx->ctx, pa[0], pb[0], pd[0]) : b200_mul_mat_batch(bctx->ctx, n, pa.data(), pb.data(), pd.data());
        bctx->n_mul_mat += n;
        bctx->n_fused_groups += n > 1;
        ta.clear(); tb.clear(); td.clear();
        return rc;
    };
    auto failed = [&](int rc, const ggml_tensor * node) -> ggml_status {
        if (rc == B200_E_ALLOC) return GGML_STATUS_ALLOC_FAILED;
        GGML_LOG_ERROR("ggml-b200: %s (%s) failed: %s\n", ggml_op_name(node->op), node->name, b200_last_error());
        return GGML_STATUS_FAILED;
    };
    const ggml_tensor * last_mm = nullptr;
    for (int i = 0; i < cgraph->n_nodes; ++i) {
        ggml_tensor * node = cgraph->nodes[i];
        if (ggml_is_empty(node) || is_view_op(node->op)) continue;
        if ((node->flags & GGML_TENSOR_FLAG_COMPUTE) == 0) continue;
        if (bctx->fuse) {
            fused_mm f;
            const int used = try_fuse_mul_mat(cgraph, i, f, bctx->fuse);
            if (used > 0) {
                int rc = flush();
                if (rc != B200_OK) return failed(rc, last_mm ? last_mm : node);
                const b200_tensor up = to_b200(f.up), x = to_b200(f.x), d = to_b200(f.out);
                const b200_tensor gt = f.gate ? to_b200(f.gate) : b200_tensor{}, bu = f.bias_up ? to_b200(f.bias_up) : b200_tensor{};
                const b200_tensor bg = f.bias_gate ? to_b200(f.bias_gate) : b200_tensor{};
                rc = f.add_only ? b200_mul_mat_add(bctx->ctx, &up, &x, &bu, &d)
                                : b200_mul_mat_fused(bctx->ctx, &up, f.gate ? &gt : nullptr, &x, f.bias_up ? &bu : nullptr, f.bias_gate ? &bg : nullptr, f.glu_op, &d);
                if (rc != B200_OK) return failed(rc, node);
                bctx->n_fused_mm++;
                i += used - 1;
                continue;
            }
        }
        // RMS_NORM followed by MUL with the norm weight (every norm of a Llama graph): one pass (reference: ggml_cuda_can_fuse of
        // {RMS_NORM, MUL}, ggml-cuda.cu)
        static const bool no_fuse_norm = [] { const char * e = getenv("GGML_B200_NO_FUSE_NORM"); return e && *e == '1'; }();
        if (bctx->fuse && !no_fuse_norm && node->op == GGML_OP_RMS_NORM && i + 1 < cgraph->n_nodes && extra_op_supported(node)) {
            const ggml_tensor * mul = cgraph->nodes[i + 1];
            if (mul->op == GGML_OP_MUL && ggml_node_has_n_uses(cgraph, i, 1) && (node->flags & GGML_TENSOR_FLAG_OUTPUT) == 0 && extra_op_supported(mul) &&
                (mul->src[0] == node || mul->src[1] == node)) {
                const ggml_tensor * wt = mul->src[0] == node ? mul->src[1] : mul->src[0];
                const char * xl = (const char *) node->src[0]->data, * dl = (const char *) mul->data;
                const bool clash = xl != dl && xl < dl + ggml_nbytes(mul) && dl < xl + ggml_nbytes(node->src[0]);     // (the same rows in place are fine)
                if (wt != node && !clash && ggml_are_same_shape(mul, node) && ggml_can_repeat(wt, node) && mul->nb[0] == 4) {
                    int rc = flush();
                    if (rc != B200_OK) return failed(rc, last_mm ? last_mm : node);
                    const b200_tensor a = to_b200(node->src[0]), w = to_b200(wt), d = to_b200(mul);
                    rc = b200_op_rms_norm_mul(bctx->ctx, &a, &w, &d, ggml_get_op_params_f32(node, 0));
                    if (rc != B200_OK) return failed(rc, node);
                    bctx->n_extra_ops++;
                    ++i;
                    continue;
                }
            }
        }
        switch (node->op) {
            case GGML_OP_MUL_MAT: {
                if (!is_quant_weight(node->src[0]->type)) {          // attention products: float src0
                    int rc = flush();
                    if (rc != B200_OK) return failed(rc, last_mm ? last_mm : node);
                    const b200_tensor a = to_b200(node->src[0]), b = to_b200(node->src[1]), d = to_b200(node);
                    rc = b200_mul_mat_f(bctx->ctx, &a, &b, &d);
                    if (rc != B200_OK) return failed(rc, node);
                    bctx->n_extra_ops++;
                    break;
                }
                if (split_src0(node)) {
                    int rc = flush();
                    if (rc == B200_OK) rc = split_mul_mat(bctx, node);
                    if (rc != B200_OK) return failed(rc, node);
                    break;
                }
                ta.push_back(to_b200(node->src[0])); tb.push_back(to_b200(node->src[1])); td.push_back(to_b200(node));
                last_mm = node;
            } break;
            case GGML_OP_MUL_MAT_ID: {
                int rc = flush();
                if (rc != B200_OK) return failed(rc, last_mm ? last_mm : node);
                const b200_tensor a = to_b200(node->src[0]), b = to_b200(node->src[1]), ids = to_b200(node->src[2]), d = to_b200(node);
                rc = b200_mul_mat_id(bctx->ctx, &a, &b, &ids, &d);
                bctx->n_mul_mat_id++;
                if (rc != B200_OK) return failed(rc, node);
            } break;
            default: {
                if (!extra_op_supported(node)) GGML_ABORT("ggml-b200: op %s reached graph_compute but supports_op() rejects it", ggml_op_name(node->op));
                int rc = flush();
                if (rc != B200_OK) return failed(rc, last_mm ? last_mm : node);
                rc = run_extra_op(bctx, node);
                if (rc != B200_OK) return failed(rc, node);
                bctx->n_extra_ops++;
            }
        }
    }
    const int rc = flush();
    if (rc != B200_OK) return failed(rc, last_mm ? last_mm : cgraph->nodes[0]);
    return GGML_STATUS_SUCCESS;
}

// CUDA-graph replay of a whole cgraph (reference: ggml-cuda.cu:3287 graph properties, :4457-4490 capture / update / launch).  A cgraph is
// identified by a hash over every node's op, type, shape, strides, parameters and data pointers (sources included): llama.cpp rebuilds the same
// graph for every decoded token until the KV-cache view grows past its padding, so a decode step is one cudaGraphLaunch instead of ~600
// kernel launches.  First sight of a key: eager (this also grows the scratch buffers); second sight: capture + launch; later: launch.


// (hashed 8 bytes at a time: a byte-wise FNV over the ~560 KB that describe a Llama-3-8B decode graph took 0.5 ms of every step — host time
//  in front of the first launch)
uint64_t graph_key(const ggml_cgraph * cgraph) {
    uint64_t h = 1469598103934665603ull;
    auto mix64 = [&](uint64_t w) { h = (h ^ w) * 0x9E3779B97F4A7C15ull; h ^= h >> 29; };
    auto mix_words = [&](const void * p, size_t n) {           // n is a multiple of 8 for every field hashed this way
        const char * b = (const char *) p;
        for (size_t i = 0; i < n; i += 8) { uint64_t w; memcpy(&w, b + i, 8); mix64(w); }
    };
    auto mix_tensor = [&](const ggml_tensor * t) {
        mix64((uint64_t) t->type); mix_words(t->ne, sizeof(t->ne)); mix_words(t->nb, sizeof(t->nb)); mix64((uint64_t) (uintptr_t) t->data);
    };
    static_assert(sizeof(((ggml_tensor *) nullptr)->op_params) % 8 == 0 && sizeof(((ggml_tensor *) nullptr)->ne) % 8 == 0, "hashed in 8-byte words");
    mix64((uint64_t) cgraph->n_nodes);
    for (int i = 0; i < cgraph->n_nodes; ++i) {
        const ggml_tensor * n = cgraph->nodes[i];
        mix64(((uint64_t) n->op << 32) | (uint32_t) n->flags); mix_words(n->op_params, sizeof(n->op_params));
        mix_tensor(n);
        for (int j = 0; j < GGML_MAX_SRC; ++j) {
            if (!n->src[j]) { mix64(~0ull); continue; }
            mix_tensor(n->src[j]);
        }
    }
    return h;
}

bool graph_capturable(const ggml_cgraph * cgraph) {
    int work = 0;
    for (int i = 0; i < cgraph->n_nodes; ++i) {
        const ggml_tensor * n = cgraph->nodes[i];
        if (is_view_op(n->op) || ggml_is_empty(n)) continue;
        if (n->op == GGML_OP_MUL_MAT && split_src0(n)) return false;        // several devices and their events: stays eager
        if (n->op == GGML_OP_MUL_MAT_ID) return false;                       // (its routing tables are sized on the host per call)
        ++work;
    }
    return work >= 8;                                                        // short graphs gain nothing
}

// Eager launches or CUDA-graph replay?  Measured on B200 (Llama-3-8B decode graph, ~520 launches): replay 2.92 ms, eager 2.72 ms — with
// programmatic dependent launch on every kernel the stream feeds the GPU faster than a graph's kernel-to-kernel dependencies resolve, as long
// as the host keeps up (here ~5 us per launch are available); a single MUL_MAT (test-backend-ops perf) is the other way round.  So neither
// is assumed: a graph is run eagerly twice and replayed twice under device timers, and the faster way is kept for that graph.
static void graph_collect_timing(b200_backend_ctx * bctx, graph_cache_entry * e) {
    if (!e->timing) return;
    float ms = 0.0f;
    if (b200_event_elapsed_ms(bctx->t0, bctx->t1, &ms) == B200_OK) {
        if (e->timing == 1) e->t_eager = std::min(e->t_eager, ms); else e->t_replay = std::min(e->t_replay, ms);
    }
    e->timing = 0;
}

static bool graph_timer_start(b200_backend_ctx * bctx) {
    if (!bctx->t0 && (b200_event_create_timed(bctx->device, &bctx->t0) != B200_OK || b200_event_create_timed(bctx->device, &bctx->t1) != B200_OK)) return false;
    return b200_event_record(bctx->ctx, bctx->t0) == B200_OK;
}

ggml_status backend_graph_compute(ggml_backend_t backend, ggml_cgraph * cgraph) {
    b200_backend_ctx * bctx = (b200_backend_ctx *) backend->context;
    bctx->n_graphs++;
    if (!bctx->use_graphs || !graph_capturable(cgraph)) return graph_compute_eager(bctx, cgraph);
    const uint64_t key = graph_key(cgraph);
    graph_cache_entry * e = nullptr, * oldest = &bctx->graphs[0];
    for (auto & c : bctx->graphs) if (c.timing) graph_collect_timing(bctx, &c);      // (one pair of timer events: a measurement is read before the next one starts)
    for (auto & c : bctx->graphs) {
        if (c.key == key && c.seen > 0) { e = &c; break; }
        if (c.stamp < oldest->stamp) oldest = &c;
    }
    if (!e) {
        e = oldest;
        if (e->graph) b200_graph_free(e->graph);
        *e = graph_cache_entry{};
        e->key = key;
    }
    e->stamp = ++bctx->graph_clock;
    e->seen++;
    const bool tune = bctx->use_graphs == 1;
    if (tune && e->mode == 0 && e->eager_runs >= 2 && e->replay_runs >= 2) {
        e->mode = e->t_replay <= e->t_eager ? 2 : 1;
        if (e->mode == 1) { bctx->n_graphs_eager_kept++; if (e->graph) { b200_graph_free(e->graph); e->graph = nullptr; } }
        else bctx->n_graphs_replay_kept++;
    }
    if (tune && (e->mode == 1 || e->failed)) return graph_compute_eager(bctx, cgraph);
    if (tune && e->mode == 0 && e->seen >= 2 && e->eager_runs < 2) {            // second and third sight: eager, timed
        const bool timed = graph_timer_start(bctx);
        const ggml_status st = graph_compute_eager(bctx, cgraph);
        if (timed && b200_event_record(bctx->ctx, bctx->t1) == B200_OK) e->timing = 1;
        e->eager_runs++;
        return st;
    }
    const uint64_t gen = b200_ctx_alloc_generation(bctx->ctx);
    if (e->graph && e->alloc_gen != gen) { b200_graph_free(e->graph); e->graph = nullptr; }       // a scratch buffer moved since the capture
    if (e->graph) {
        const bool timed = tune && e->mode == 0 && e->replay_runs < 2 && graph_timer_start(bctx);
        if (b200_graph_launch(bctx->ctx, e->graph) == B200_OK) {
            bctx->n_graph_replays++;
            if (tune && e->mode == 0 && e->replay_runs < 2) {
                if (timed && b200_event_record(bctx->ctx, bctx->t1) == B200_OK) e->timing = 2;
                e->replay_runs++;
            }
            return GGML_STATUS_SUCCESS;
        }
        GGML_LOG_ERROR("ggml-b200: graph launch failed (%s), running eagerly\n", b200_last_error());
        b200_graph_free(e->graph); e->graph = nullptr; e->failed = true;
    }
    if (e->seen < 2 || e->failed) return graph_compute_eager(bctx, cgraph);
    // seen before (and, when tuning, timed eagerly): record it
    const long s_mm = bctx->n_mul_mat, s_id = bctx->n_mul_mat_id, s_fg = bctx->n_fused_groups, s_ops = bctx->n_extra_ops, s_fm = bctx->n_fused_mm;
    if (b200_capture_begin(bctx->ctx) != B200_OK) { e->failed = true; return graph_compute_eager(bctx, cgraph); }
    const ggml_status st = graph_compute_eager(bctx, cgraph);
    void * g = nullptr;
    const int rc = b200_capture_end(bctx->ctx, &g);
    if (st != GGML_STATUS_SUCCESS || rc != B200_OK || !g) {
        // (typically a scratch buffer that had to grow: the eager run below grows it, the next sight of the graph captures)
        if (g) b200_graph_free(g);
        bctx->n_mul_mat = s_mm; bctx->n_mul_mat_id = s_id; bctx->n_fused_groups = s_fg; bctx->n_extra_ops = s_ops; bctx->n_fused_mm = s_fm;
        if (++e->capture_attempts >= 3) e->failed = true;
        if (!tune) e->seen = 1;
        return graph_compute_eager(bctx, cgraph);
    }
    e->graph = g; e->alloc_gen = b200_ctx_alloc_generation(bctx->ctx);
    bctx->n_graph_captures++;
    if (b200_graph_launch(bctx->ctx, e->graph) != B200_OK) {
        GGML_LOG_ERROR("ggml-b200: graph launch failed (%s)\n", b200_last_error());
        return GGML_STATUS_FAILED;
    }
    return GGML_STATUS_SUCCESS;
}

// ---- events, asynchronous copies (ggml_backend_cuda_event_record / _wait, ggml-cuda.cu:4498-4520; cpy_tensor_async :3177) ----
void backend_event_record(ggml_backend_t backend, ggml_backend_event_t event) {
    check(b200_event_record(((b200_backend_ctx *) backend->context)->ctx, event->context), "event_record");
}

void backend_event_wait(ggml_backend_t backend, ggml_backend_event_t event) {
    check(b200_event_wait(((b200_backend_ctx *) backend->context)->ctx, event->context), "event_wait");
}

bool backend_is_b200(ggml_backend_t backend);

// device-to-device copy between two of our backends (any two devices), ordered after the source backend's queued work and
// executed on the destination backend's stream
bool backend_cpy_tensor_async(ggml_backend_t backend_src, ggml_backend_t backend_dst, const ggml_tensor * src, ggml_tensor * dst) {
    if (!backend_is_b200(backend_src) || !backend_is_b200(backend_dst)) return false;
    ggml_backend_buffer_t sbuf = src->view_src ? src->view_src->buffer : src->buffer;
    ggml_backend_buffer_t dbuf = dst->view_src ? dst->view_src->buffer : dst->buffer;
    if (!buffer_is_b200(sbuf) || !buffer_is_b200(dbuf) || buft_is_split(sbuf->buft) || buft_is_split(dbuf->buft)) return false;
    b200_backend_ctx * cs = (b200_backend_ctx *) backend_src->context, * cd = (b200_backend_ctx *) backend_dst->context;
    if (cs != cd) {
        if (!cs->copy_event) check(b200_event_create(cs->device, &cs->copy_event), "event_create");
        check(b200_event_record(cs->ctx, cs->copy_event), "event_record");
        check(b200_event_wait(cd->ctx, cs->copy_event), "event_wait");
    }
    check(b200_memcpy_async(cd->ctx, dst->data, src->data, ggml_nbytes(src), 2), "cpy_tensor_async");
    return true;
}

// graph_optimize (called by ggml_backend_sched on every split before allocation): quantized MUL_MATs that share their activations are moved
// next to each other — llama.cpp emits wq, ROPE(q), wk, ROPE(k), wv …, and only adjacent MUL_MAT nodes reach the kernels as one list
// (b200_mul_mat_batch: one quantization of the activations, one launch for q / k / v).  A moved node depends on nothing but its weight and
// the shared activations, which are complete before the first MUL_MAT of the group, so the order stays a valid one.
void reorder_shared_activation_mul_mats(ggml_cgraph * cgraph);
void backend_graph_optimize(ggml_backend_t backend, ggml_cgraph * cgraph) {
    b200_backend_ctx * bctx = (b200_backend_ctx *) backend->context;
    if (!bctx->reorder || only_mul_mat()) return;
    reorder_shared_activation_mul_mats(cgraph);
}

void reorder_shared_activation_mul_mats(ggml_cgraph * cgraph) {
    auto groupable = [](const ggml_tensor * t) {
        return t->op == GGML_OP_MUL_MAT && t->src[0] && t->src[1] && is_quant_weight(t->src[0]->type) && !split_src0(t) && t->src[0]->op == GGML_OP_NONE;
    };
    for (int i = 0; i < cgraph->n_nodes; ++i) {
        ggml_tensor * first = cgraph->nodes[i];
        if (!groupable(first)) continue;
        int ins = i + 1, members = 1;
        for (int j = i + 1; j < cgraph->n_nodes && j <= i + 32 && members < 3; ++j) {
            ggml_tensor * t = cgraph->nodes[j];
            if (!groupable(t) || t->src[1] != first->src[1] || t->src[0]->type != first->src[0]->type) continue;
            // (an op between the two that works in place on the shared activations — a node that is a view of their buffer — pins the order)
            bool pinned = false;
            const ggml_tensor * xroot = first->src[1]->view_src ? first->src[1]->view_src : first->src[1];
            for (int k = ins; k < j && !pinned; ++k) {
                const ggml_tensor * m = cgraph->nodes[k];
                pinned = !is_view_op(m->op) && m->view_src && (m->view_src == xroot || m->view_src == first->src[1]);
            }
            if (pinned) break;
            if (j != ins) {                      // rotate nodes[ins .. j] right by one
                for (int k = j; k > ins; --k) cgraph->nodes[k] = cgraph->nodes[k - 1];
                cgraph->nodes[ins] = t;
            }
            ++ins; ++members;
        }
        i = ins - 1;
    }
}

const ggml_backend_i k_backend_iface = {
    /* .get_name            = */ backend_get_name,
    /* .free                = */ backend_free,
    /* .set_tensor_async    = */ backend_set_tensor_async,
    /* .get_tensor_async    = */ backend_get_tensor_async,
    /* .set_tensor_2d_async = */ nullptr,
    /* .get_tensor_2d_async = */ nullptr,
    /* .cpy_tensor_async    = */ backend_cpy_tensor_async,
    /* .synchronize         = */ backend_synchronize,
    /* .graph_plan_create   = */ nullptr,
    /* .graph_plan_free     = */ nullptr,
    /* .graph_plan_update   = */ nullptr,
    /* .graph_plan_compute  = */ nullptr,
    /* .graph_compute       = */ backend_graph_compute,
    /* .event_record        = */ backend_event_record,
    /* .event_wait          = */ backend_event_wait,
    /* .graph_optimize      = */ backend_graph_optimize,
};

bool backend_is_b200(ggml_backend_t backend) { return backend && backend->iface.get_name == backend_get_name; }

ggml_guid_t backend_guid() {
    static ggml_guid guid = { 0xb2, 0x00, 0x51, 0x10, 0x0a, 0x67, 0x67, 0x6d, 0x6c, 0x2d, 0x62, 0x32, 0x30, 0x30, 0x00, 0x01 };
    return &guid;
}

// ---- device ---------------------------------------------------------------------------------------------------------------
const char * device_get_name(ggml_backend_dev_t dev) { return ((b200_device_ctx *) dev->context)->name.c_str(); }
const char * device_get_description(ggml_backend_dev_t dev) { return ((b200_device_ctx *) dev->context)->description.c_str(); }

void device_get_memory(ggml_backend_dev_t dev, size_t * free, size_t * total) {
    b200_device_ctx * dctx = (b200_device_ctx *) dev->context;
    *free = *total = 0;
    b200_device_info(dctx->device, nullptr, 0, free, total, nullptr, nullptr, nullptr);
}

enum ggml_backend_dev_type device_get_type(ggml_backend_dev_t) { return GGML_BACKEND_DEVICE_TYPE_GPU; }

void device_get_props(ggml_backend_dev_t dev, ggml_backend_dev_props * props) {
    props->name        = device_get_name(dev);
    props->description = device_get_description(dev);
    props->type        = device_get_type(dev);
    props->device_id   = nullptr;
    device_get_memory(dev, &props->memory_free, &props->memory_total);
    props->caps = {
        /* .async                = */ true,       // graph_compute, set/get/cpy_tensor_async only queue work on the backend's stream
        /* .host_buffer          = */ true,       // pinned host buffer type below
        /* .buffer_from_host_ptr = */ false,
        /* .events               = */ true,
    };
}

ggml_backend_t device_init_backend(ggml_backend_dev_t dev, const char * /*params*/) {
    b200_device_ctx * dctx = (b200_device_ctx *) dev->context;
    b200_ctx * ctx = nullptr;
    if (b200_ctx_create(dctx->device, nullptr, &ctx) != B200_OK) {
        GGML_LOG_ERROR("ggml-b200: cannot create a context on device %d: %s\n", dctx->device, b200_last_error());
        return nullptr;
    }
    b200_backend_ctx * bctx = new b200_backend_ctx{dctx->device, dctx->name, ctx};
    if (const char * e = getenv("GGML_B200_FUSE")) bctx->fuse = atoi(e);
    if (const char * e = getenv("GGML_B200_GRAPHS"); e && *e >= '0' && *e <= '2') bctx->use_graphs = *e - '0';
    if (const char * e = getenv("GGML_B200_REORDER"); e && *e == '0') bctx->reorder = false;
    return new ggml_backend{
        /* .guid    = */ backend_guid(),
        /* .iface   = */ k_backend_iface,
        /* .device  = */ dev,
        /* .context = */ bctx,
    };
}

ggml_backend_buffer_type_t device_get_buffer_type(ggml_backend_dev_t dev) { return &((b200_device_ctx *) dev->context)->buft; }

bool device_supports_op(ggml_backend_dev_t, const ggml_tensor * op) {
    if (is_view_op(op->op)) return true;
    switch (op->op) {
        case GGML_OP_MUL_MAT: {
            if (!op->src[0] || !op->src[1]) return false;
            const b200_tensor a = to_b200(op->src[0]), b = to_b200(op->src[1]), d = to_b200(op);
            if (is_float_type(op->src[0]->type)) return !only_mul_mat() && !split_src0(op) && b200_mul_mat_f_supported(&a, &b, &d) != 0;
            if (split_src0(op) && (op->src[1]->ne[2] * op->src[1]->ne[3] != 1 || (op->ne[1] > 1 && op->nb[1] != (size_t) op->ne[0] * 4))) return false;
            return b200_mul_mat_supported(&a, &b, &d) != 0;
        }
        case GGML_OP_MUL_MAT_ID: {
            if (!op->src[0] || !op->src[1] || !op->src[2]) return false;
            if (split_src0(op)) return false;                      // like the reference (ggml-cuda.cu:2628)
            const b200_tensor a = to_b200(op->src[0]), b = to_b200(op->src[1]), ids = to_b200(op->src[2]), d = to_b200(op);
            return b200_mul_mat_id_supported(&a, &b, &ids, &d) != 0;
        }
        default:
            return extra_op_supported(op);
    }
}

bool device_supports_buft(ggml_backend_dev_t dev, ggml_backend_buffer_type_t buft) {
    if (buft_is_split(buft)) {
        // the main device's backend computes with split weights (the other devices take part through the row-split driver)
        return buft->device == dev;
    }
    return buft_is_b200(buft) && buft->device == dev;
}

// ---- pinned host buffer type (ggml_backend_cuda_host_buffer_type, ggml-cuda.cu:1493-1542): plain CPU buffers over page-locked
// memory, so the scheduler's input copies and llama.cpp's weight upload run asynchronously -------------------------------------
const char * host_buft_get_name(ggml_backend_buffer_type_t) { return GGML_B200_NAME "_Host"; }

void host_buffer_free(ggml_backend_buffer_t buffer) { b200_host_free(buffer->context); }

ggml_backend_buffer_t host_buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    void * ptr = nullptr;
    if (b200_host_malloc(size, &ptr) != B200_OK) {
        return ggml_backend_buft_alloc_buffer(ggml_backend_cpu_buffer_type(), size);      // pageable memory still works, synchronously
    }
    ggml_backend_buffer_t buffer = ggml_backend_cpu_buffer_from_ptr(ptr, size);
    buffer->buft = buft;
    buffer->iface.free_buffer = host_buffer_free;
    return buffer;
}

ggml_backend_buffer_type_t device_get_host_buffer_type(ggml_backend_dev_t dev) {
    static ggml_backend_buffer_type host_buft[GGML_B200_MAX_DEVS];
    static std::once_flag once;
    std::call_once(once, [] {
        ggml_backend_buffer_type_t cpu = ggml_backend_cpu_buffer_type();
        for (auto & b : host_buft) {
            b.iface = { host_buft_get_name, host_buft_alloc_buffer, cpu->iface.get_alignment, nullptr, cpu->iface.get_alloc_size, cpu->iface.is_host };
            b.device = nullptr; b.context = nullptr;
        }
    });
    b200_device_ctx * dctx = (b200_device_ctx *) dev->context;
    host_buft[dctx->device].device = dev;
    return &host_buft[dctx->device];
}

// ---- device events (ggml_backend_cuda_device_event_new / _free / _synchronize, ggml-cuda.cu:5430-5460) ------------------------
ggml_backend_event_t device_event_new(ggml_backend_dev_t dev) {
    void * ev = nullptr;
    if (b200_event_create(((b200_device_ctx *) dev->context)->device, &ev) != B200_OK) return nullptr;
    return new ggml_backend_event{ /* .device = */ dev, /* .context = */ ev };
}

void device_event_free(ggml_backend_dev_t dev, ggml_backend_event_t event) {
    b200_event_destroy(((b200_device_ctx *) dev->context)->device, event->context);
    delete event;
}

void device_event_synchronize(ggml_backend_dev_t, ggml_backend_event_t event) { check(b200_event_synchronize(event->context), "event_synchronize"); }

const ggml_backend_device_i k_device_iface = {
    /* .get_name             = */ device_get_name,
    /* .get_description      = */ device_get_description,
    /* .get_memory           = */ device_get_memory,
    /* .get_type             = */ device_get_type,
    /* .get_props            = */ device_get_props,
    /* .init_backend         = */ device_init_backend,
    /* .get_buffer_type      = */ device_get_buffer_type,
    /* .get_host_buffer_type = */ device_get_host_buffer_type,
    /* .buffer_from_host_ptr = */ nullptr,
    /* .supports_op          = */ device_supports_op,
    /* .supports_buft        = */ device_supports_buft,
    /* .offload_op           = */ nullptr,
    /* .event_new            = */ device_event_new,
    /* .event_free           = */ device_event_free,
    /* .event_synchronize    = */ device_event_synchronize,
};

// ---- registry --------------------------------------------------------------------------------------------------------------
struct b200_reg_ctx {
    std::vector<ggml_backend_device *> devices;
};

const char * reg_get_name(ggml_backend_reg_t) { return GGML_B200_NAME; }
size_t reg_get_device_count(ggml_backend_reg_t reg) { return ((b200_reg_ctx *) reg->context)->devices.size(); }

ggml_backend_dev_t reg_get_device(ggml_backend_reg_t reg, size_t index) {
    b200_reg_ctx * rctx = (b200_reg_ctx *) reg->context;
    GGML_ASSERT(index < rctx->devices.size());
    return rctx->devices[index];
}

ggml_backend_feature * get_features(ggml_backend_reg_t) {
    static ggml_backend_feature features[] = {
        { "ARCH",      "sm_100a" },
        { "TCGEN05",   "1" },
        { "Q8_1_EXACT","1" },
        { nullptr, nullptr },
    };
    return features;
}

// ---- ggml_backend_comm_{init, free, allreduce_tensor} (ggml-backend.h:205-208; reference ggml-cuda.cu:1184-1430): the all-reduce of the
// meta / tensor-parallel backend.  All backends must be B200 backends; a tensor that cannot be reduced here (type, layout) returns
// false and the meta backend uses its generic path for that call, like the reference's per-call fallback.
struct b200_comm_ctx { b200_comm * comm; std::vector<ggml_backend_t> backends; };

void * comm_init(ggml_backend_t * backends, size_t n_backends) {
    if (!backends || n_backends == 0 || n_backends > 8) return nullptr;
    std::vector<b200_ctx *> ctxs;
    for (size_t i = 0; i < n_backends; ++i) {
        if (!backend_is_b200(backends[i])) return nullptr;
        ctxs.push_back(((b200_backend_ctx *) backends[i]->context)->ctx);
    }
    b200_comm * comm = nullptr;
    if (b200_comm_create(ctxs.data(), (int) n_backends, &comm) != B200_OK) {
        GGML_LOG_ERROR("ggml-b200: comm_init failed: %s\n", b200_last_error());
        return nullptr;
    }
    return new b200_comm_ctx{comm, std::vector<ggml_backend_t>(backends, backends + n_backends)};
}

void comm_free(void * comm_ctx) {
    b200_comm_ctx * c = (b200_comm_ctx *) comm_ctx;
    if (!c) return;
    b200_comm_destroy(c->comm);
    delete c;
}

bool comm_allreduce_tensor(void * comm_ctx, ggml_tensor ** tensors) {
    b200_comm_ctx * c = (b200_comm_ctx *) comm_ctx;
    if (!c || !tensors || !tensors[0]) return false;
    const size_t n = c->backends.size();
    const int64_t ne = ggml_nelements(tensors[0]);
    if (ne == 0) return true;
    void * data[8]; int zero[8];
    for (size_t i = 0; i < n; ++i) {
        const ggml_tensor * t = tensors[i];
        if (!t || t->type != GGML_TYPE_F32 || ggml_nelements(t) != ne || !ggml_is_contiguously_allocated(t) || ((uintptr_t) t->data & 15)) return false;
        data[i] = t->data;
        zero[i] = (t->flags & GGML_TENSOR_FLAG_COMPUTE) == 0;         // not computed on this backend: contributes zeros (ggml-cuda.cu:1205-1211)
    }
    if (b200_comm_allreduce(c->comm, data, zero, ne) != B200_OK) {
        GGML_LOG_ERROR("ggml-b200: allreduce failed: %s\n", b200_last_error());
        return false;
    }
    return true;
}

void * reg_get_proc_address(ggml_backend_reg_t, const char * name) {
    if (strcmp(name, "ggml_backend_get_features") == 0) return (void *) get_features;
    if (strcmp(name, "ggml_backend_comm_init") == 0) return (void *) comm_init;                          // ggml_backend_comm_init_t
    if (strcmp(name, "ggml_backend_comm_free") == 0) return (void *) comm_free;                          // ggml_backend_comm_free_t
    if (strcmp(name, "ggml_backend_comm_allreduce_tensor") == 0) return (void *) comm_allreduce_tensor;  // ggml_backend_comm_allreduce_tensor_t
    if (strcmp(name, "ggml_backend_split_buffer_type") == 0) return (void *) split_buffer_type;      // ggml_backend_split_buffer_type_t
    return nullptr;
}

const ggml_backend_reg_i k_reg_iface = {
    /* .get_name         = */ reg_get_name,
    /* .get_device_count = */ reg_get_device_count,
    /* .get_device       = */ reg_get_device,
    /* .get_proc_address = */ reg_get_proc_address,
};

}  // namespace

ggml_backend_reg_t ggml_backend_b200_reg(void) {
    static ggml_backend_reg reg;
    static std::once_flag   once;
    std::call_once(once, [] {
        b200_reg_ctx * rctx = new b200_reg_ctx;
        reg = ggml_backend_reg{
            /* .api_version = */ GGML_BACKEND_API_VERSION,
            /* .iface       = */ k_reg_iface,
            /* .context     = */ rctx,
        };
        int n = b200_device_count();
        if (n > GGML_B200_MAX_DEVS) n = GGML_B200_MAX_DEVS;
        for (int i = 0; i < n; ++i) {
            b200_device_ctx * dctx = new b200_device_ctx;
            char devname[256] = "";
            b200_device_info(i, devname, sizeof(devname), nullptr, nullptr, nullptr, nullptr, nullptr);
            dctx->device      = i;
            dctx->name        = std::string(GGML_B200_NAME) + std::to_string(i);
            dctx->description = devname;
            ggml_backend_device * dev = new ggml_backend_device{
                /* .iface   = */ k_device_iface,
                /* .reg     = */ &reg,
                /* .context = */ dctx,
            };
            dctx->buft = ggml_backend_buffer_type{
                /* .iface   = */ k_buft_iface,
                /* .device  = */ dev,
                /* .context = */ nullptr,
            };
            rctx->devices.push_back(dev);
        }
    });
    return &reg;
}

static int ggml_backend_b200_score(void) { return b200_device_count() > 0 ? 100 : 0; }

// the node reordering of graph_optimize on its own (needs no device): oracle/reorder_check.cpp checks on the CPU that the new order is still
// a valid one and that q / k / v and gate / up end up adjacent
extern "C" {
GGML_BACKEND_API void ggml_backend_b200_reorder_graph(ggml_cgraph * cgraph);
void ggml_backend_b200_reorder_graph(ggml_cgraph * cgraph) { reorder_shared_activation_mul_mats(cgraph); }
}

GGML_BACKEND_DL_IMPL(ggml_backend_b200_reg)
GGML_BACKEND_DL_SCORE_IMPL(ggml_backend_b200_score)
