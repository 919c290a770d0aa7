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
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#define GGML_B200_NAME     "B200"
#define GGML_B200_MAX_DEVS 16

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

struct b200_backend_ctx {
    int         device;
    std::string name;
    b200_ctx *  ctx;
    long        n_graphs = 0, n_mul_mat = 0, n_mul_mat_id = 0, n_fused_groups = 0;      // GGML_B200_STATS=1 prints them on free
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

// ---- backend (stream) -------------------------------------------------------------------------------------------------------
const char * backend_get_name(ggml_backend_t backend) { return ((b200_backend_ctx *) backend->context)->name.c_str(); }

void backend_free(ggml_backend_t backend) {
    b200_backend_ctx * bctx = (b200_backend_ctx *) backend->context;
    if (const char * e = getenv("GGML_B200_STATS"); e && *e == '1') {
        fprintf(stderr, "ggml-b200 stats: graphs=%ld mul_mat=%ld mul_mat_id=%ld fused_groups=%ld\n", bctx->n_graphs, bctx->n_mul_mat,
                bctx->n_mul_mat_id, bctx->n_fused_groups);
    }
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

ggml_status backend_graph_compute(ggml_backend_t backend, ggml_cgraph * cgraph) {
    b200_backend_ctx * bctx = (b200_backend_ctx *) backend->context;
    bctx->n_graphs++;
    for (int i = 0; i < cgraph->n_nodes; ++i) {
        ggml_tensor * node = cgraph->nodes[i];
        if (ggml_is_empty(node) || is_view_op(node->op)) continue;
        if ((node->flags & GGML_TENSOR_FLAG_COMPUTE) == 0) continue;
        int rc;
        switch (node->op) {
            case GGML_OP_MUL_MAT: {
                // the MUL_MAT nodes that follow and read the same activations (wq / wk / wv, ffn_gate / ffn_up in a llama graph)
                // run in the same launch (b200_mul_mat_multi); they are independent: their weights are not results of the group
                ggml_tensor * group[3] = {node, nullptr, nullptr};
                int n_group = 1, last = i;
                for (int j = i + 1; j < cgraph->n_nodes && n_group < 3; ++j) {
                    ggml_tensor * nx = cgraph->nodes[j];
                    if (ggml_is_empty(nx) || is_view_op(nx->op)) continue;
                    if (nx->op != GGML_OP_MUL_MAT || nx->src[1] != node->src[1] || (nx->flags & GGML_TENSOR_FLAG_COMPUTE) == 0) break;
                    bool independent = true;
                    for (int g = 0; g < n_group; ++g) independent = independent && nx->src[0] != group[g] && nx->src[0]->view_src != group[g];
                    if (!independent) break;
                    group[n_group++] = nx;
                    last = j;
                }
                bctx->n_mul_mat += n_group;
                bctx->n_fused_groups += n_group > 1;
                if (n_group == 1) {
                    const b200_tensor a = to_b200(node->src[0]), b = to_b200(node->src[1]), d = to_b200(node);
                    rc = b200_mul_mat(bctx->ctx, &a, &b, &d);
                } else {
                    b200_tensor a[3], d[3];
                    const b200_tensor * ap[3], * dp[3];
                    for (int g = 0; g < n_group; ++g) { a[g] = to_b200(group[g]->src[0]); d[g] = to_b200(group[g]); ap[g] = &a[g]; dp[g] = &d[g]; }
                    const b200_tensor b = to_b200(node->src[1]);
                    rc = b200_mul_mat_multi(bctx->ctx, n_group, ap, &b, dp);
                    i = last;       // view nodes in between need no work
                }
            } break;
            case GGML_OP_MUL_MAT_ID: {
                const b200_tensor a = to_b200(node->src[0]), b = to_b200(node->src[1]), ids = to_b200(node->src[2]), d = to_b200(node);
                rc = b200_mul_mat_id(bctx->ctx, &a, &b, &ids, &d);
                bctx->n_mul_mat_id++;
            } break;
            default:
                GGML_ABORT("ggml-b200: op %s reached graph_compute but supports_op() rejects it", ggml_op_name(node->op));
        }
        if (rc == B200_E_ALLOC) return GGML_STATUS_ALLOC_FAILED;
        if (rc != B200_OK) {
            GGML_LOG_ERROR("ggml-b200: %s (%s) failed: %s\n", ggml_op_name(node->op), node->name, b200_last_error());
            return GGML_STATUS_FAILED;
        }
    }
    return GGML_STATUS_SUCCESS;
}

const ggml_backend_i k_backend_iface = {
    /* .get_name            = */ backend_get_name,
    /* .free                = */ backend_free,
    /* .set_tensor_async    = */ backend_set_tensor_async,
    /* .get_tensor_async    = */ backend_get_tensor_async,
    /* .set_tensor_2d_async = */ nullptr,
    /* .get_tensor_2d_async = */ nullptr,
    /* .cpy_tensor_async    = */ nullptr,
    /* .synchronize         = */ backend_synchronize,
    /* .graph_plan_create   = */ nullptr,
    /* .graph_plan_free     = */ nullptr,
    /* .graph_plan_update   = */ nullptr,
    /* .graph_plan_compute  = */ nullptr,
    /* .graph_compute       = */ backend_graph_compute,
    /* .event_record        = */ nullptr,
    /* .event_wait          = */ nullptr,
    /* .graph_optimize      = */ nullptr,
};

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
        /* .async                = */ false,
        /* .host_buffer          = */ false,
        /* .buffer_from_host_ptr = */ false,
        /* .events               = */ false,
    };
}

ggml_backend_t device_init_backend(ggml_backend_dev_t dev, const char * /*params*/) {
    b200_device_ctx * dctx = (b200_device_ctx *) dev->context;
    b200_ctx * ctx = nullptr;
    if (b200_ctx_create(dctx->device, nullptr, &ctx) != B200_OK) {
        GGML_LOG_ERROR("ggml-b200: cannot create a context on device %d: %s\n", dctx->device, b200_last_error());
        return nullptr;
    }
    return new ggml_backend{
        /* .guid    = */ backend_guid(),
        /* .iface   = */ k_backend_iface,
        /* .device  = */ dev,
        /* .context = */ new b200_backend_ctx{dctx->device, dctx->name, ctx},
    };
}

ggml_backend_buffer_type_t device_get_buffer_type(ggml_backend_dev_t dev) { return &((b200_device_ctx *) dev->context)->buft; }

bool device_supports_op(ggml_backend_dev_t, const ggml_tensor * op) {
    if (is_view_op(op->op)) return true;
    switch (op->op) {
        case GGML_OP_MUL_MAT: {
            if (!op->src[0] || !op->src[1]) return false;
            const b200_tensor a = to_b200(op->src[0]), b = to_b200(op->src[1]), d = to_b200(op);
            return b200_mul_mat_supported(&a, &b, &d) != 0;
        }
        case GGML_OP_MUL_MAT_ID: {
            if (!op->src[0] || !op->src[1] || !op->src[2]) return false;
            const b200_tensor a = to_b200(op->src[0]), b = to_b200(op->src[1]), ids = to_b200(op->src[2]), d = to_b200(op);
            return b200_mul_mat_id_supported(&a, &b, &ids, &d) != 0;
        }
        default:
            return false;
    }
}

bool device_supports_buft(ggml_backend_dev_t dev, ggml_backend_buffer_type_t buft) {
    return buft_is_b200(buft) && buft->device == dev;
}

const ggml_backend_device_i k_device_iface = {
    /* .get_name             = */ device_get_name,
    /* .get_description      = */ device_get_description,
    /* .get_memory           = */ device_get_memory,
    /* .get_type             = */ device_get_type,
    /* .get_props            = */ device_get_props,
    /* .init_backend         = */ device_init_backend,
    /* .get_buffer_type      = */ device_get_buffer_type,
    /* .get_host_buffer_type = */ nullptr,
    /* .buffer_from_host_ptr = */ nullptr,
    /* .supports_op          = */ device_supports_op,
    /* .supports_buft        = */ device_supports_buft,
    /* .offload_op           = */ nullptr,
    /* .event_new            = */ nullptr,
    /* .event_free           = */ nullptr,
    /* .event_synchronize    = */ nullptr,
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

void * reg_get_proc_address(ggml_backend_reg_t, const char * name) {
    if (strcmp(name, "ggml_backend_get_features") == 0) return (void *) get_features;
    return nullptr;
}

const ggml_backend_reg_i k_reg_iface = {
    /* .get_name         = */ reg_get_name,
    /* .get_device_count = */ reg_get_device_count,
    /* .get_device       = */ reg_get_device,
    /* .get_proc_address = */ reg_get_proc_address,
};

}  // namespace

extern "C" {
GGML_BACKEND_API ggml_backend_reg_t ggml_backend_b200_reg(void);
}

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

GGML_BACKEND_DL_IMPL(ggml_backend_b200_reg)
GGML_BACKEND_DL_SCORE_IMPL(ggml_backend_b200_score)
