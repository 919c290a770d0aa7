// oracle/e2e_plugin.cpp — MEASUREMENT HARNESS for bench.py's `e2e` key (links the reference's own libggml / libggml-base built by
// oracle/Makefile as the HOST framework; the thing measured is the plug-in libggml-b200.so, loaded through GGML_BACKEND_PATH like
// any other dynamic ggml backend).
//
// One decode step of the Llama-3 matmuls (the workload of bench.py) through the reference's public backend API, exactly as a
// reference caller drives a backend:
//     ggml_backend_tensor_set_async(activations, pinned host memory)      H2D, inside the timed region
//     ggml_backend_graph_compute_async(backend, graph of all MUL_MAT nodes)
//     ggml_backend_tensor_get_async(results -> pinned host memory)        D2H, inside the timed region
//     ggml_backend_synchronize(backend)
// The activation vectors of one step live in ONE device tensor (every MUL_MAT reads a view of it: q/k/v and gate/up share
// theirs, as in the model) and the result vectors in ONE device tensor (ggml_backend_tensor_alloc places every node's dst in it),
// so a step moves its inputs and its results with one copy each; pinned host memory comes from the plug-in's host buffer type.
// Weights are random VALID Q4_K blocks (4.22 GB for Llama-3-8B: far more than the L2), uploaded once, outside the timed region.
//
// usage: e2e_plugin <8b|70b|tiny> <steps> <warmup>        prints one JSON line
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "ggml.h"
#include "ggml-alloc.h"
#include "ggml-backend.h"
#include "ggml-cpu.h"

struct shape { const char * name; int64_t m, k; int count; int xgroup; };   // xgroup: matrices with the same id share activations

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static inline uint64_t rnd() { rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17; return rng_state; }

static inline uint16_t f32_to_f16(float f) { return ggml_fp32_to_fp16(f); }

// m rows of random valid Q4_K blocks (any byte pattern with sane fp16 scales is a legal tensor)
static void random_q4k(std::vector<uint8_t> & out, int64_t m, int64_t k) {
    const size_t nb = (size_t) m * (k / 256);
    out.resize(nb * 144);
    uint64_t * p = (uint64_t *) out.data();
    for (size_t i = 0; i < out.size() / 8; ++i) p[i] = rnd();
    for (size_t b = 0; b < nb; ++b) {
        const float d  = (0.25f + 0.75f * (float) (rnd() & 0xffff) / 65535.0f) * (0.02f / (15 * 63));
        const float dm = (0.25f + 0.75f * (float) (rnd() & 0xffff) / 65535.0f) * (0.02f / 63);
        uint16_t h[2] = { f32_to_f16(d), f32_to_f16(dm) };
        memcpy(out.data() + b * 144, h, 4);
    }
}

int main(int argc, char ** argv) {
    const std::string model = argc > 1 ? argv[1] : "8b";
    const int steps = argc > 2 ? atoi(argv[2]) : 20, warmup = argc > 3 ? atoi(argv[3]) : 3;
    std::vector<shape> layer;
    int n_layers; shape head;
    if (model == "70b") {
        layer = { {"wq", 8192, 8192, 1, 0}, {"wk", 1024, 8192, 1, 0}, {"wv", 1024, 8192, 1, 0}, {"wo", 8192, 8192, 1, 1},
                  {"ffn_gate", 28672, 8192, 1, 2}, {"ffn_up", 28672, 8192, 1, 2}, {"ffn_down", 8192, 28672, 1, 3} };
        n_layers = 80; head = {"output", 128256, 8192, 1, 4};
    } else if (model == "tiny") {
        layer = { {"wq", 512, 512, 1, 0}, {"wk", 256, 512, 1, 0}, {"wv", 256, 512, 1, 0}, {"wo", 512, 512, 1, 1},
                  {"ffn_gate", 1024, 512, 1, 2}, {"ffn_up", 1024, 512, 1, 2}, {"ffn_down", 512, 1024, 1, 3} };
        n_layers = 2; head = {"output", 2048, 512, 1, 4};
    } else {
        layer = { {"wq", 4096, 4096, 1, 0}, {"wk", 1024, 4096, 1, 0}, {"wv", 1024, 4096, 1, 0}, {"wo", 4096, 4096, 1, 1},
                  {"ffn_gate", 14336, 4096, 1, 2}, {"ffn_up", 14336, 4096, 1, 2}, {"ffn_down", 4096, 14336, 1, 3} };
        n_layers = 32; head = {"output", 128256, 4096, 1, 4};
    }

    ggml_backend_load_all();
    ggml_backend_dev_t dev = ggml_backend_dev_by_name("B2000");
    if (!dev) { fprintf(stderr, "e2e_plugin: device B2000 not registered (GGML_BACKEND_PATH?)\n"); return 3; }
    ggml_backend_t gpu = ggml_backend_dev_init(dev, nullptr);
    if (!gpu) { fprintf(stderr, "e2e_plugin: backend init failed\n"); return 3; }
    ggml_backend_dev_props props;
    ggml_backend_dev_get_props(dev, &props);

    // ---- the list of matmuls in model order ------------------------------------------------------------------------------------
    struct mm { int64_t m, k; int xslot; };
    std::vector<mm> mms;
    std::vector<int64_t> xslot_k;            // activation vectors of the step, in order
    for (int l = 0; l < n_layers; ++l) {
        int last_group = -1;
        for (const shape & s : layer) {
            if (s.xgroup != last_group) { xslot_k.push_back(s.k); last_group = s.xgroup; }
            mms.push_back({ s.m, s.k, (int) xslot_k.size() - 1 });
        }
    }
    xslot_k.push_back(head.k);
    mms.push_back({ head.m, head.k, (int) xslot_k.size() - 1 });
    const int n_mm = (int) mms.size();
    int64_t x_total = 0, y_total = 0;
    std::vector<int64_t> x_off, y_off;
    for (int64_t k : xslot_k) { x_off.push_back(x_total); x_total += k; }
    for (const mm & q : mms) { y_off.push_back(y_total); y_total += q.m; }

    // ---- weights (device), activations + results (one device tensor each) --------------------------------------------------------
    ggml_init_params ip = { ggml_tensor_overhead() * (size_t) (4 * n_mm + 16) + ggml_graph_overhead_custom(2 * n_mm + 16, false), nullptr, true };
    ggml_context * wctx = ggml_init(ip);
    std::vector<ggml_tensor *> W;
    for (const mm & q : mms) W.push_back(ggml_new_tensor_2d(wctx, GGML_TYPE_Q4_K, q.k, q.m));
    ggml_backend_buffer_t wbuf = ggml_backend_alloc_ctx_tensors(wctx, gpu);
    if (!wbuf) { fprintf(stderr, "e2e_plugin: weight allocation failed\n"); return 1; }
    ggml_backend_buffer_set_usage(wbuf, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
    double weight_bytes = 0;
    {
        std::vector<uint8_t> blob;
        for (int i = 0; i < n_mm; ++i) {
            random_q4k(blob, mms[i].m, mms[i].k);
            ggml_backend_tensor_set(W[i], blob.data(), 0, blob.size());
            weight_bytes += (double) blob.size();
        }
    }
    ggml_context * actx = ggml_init(ip);
    ggml_tensor * X = ggml_new_tensor_1d(actx, GGML_TYPE_F32, x_total);
    ggml_tensor * Y = ggml_new_tensor_1d(actx, GGML_TYPE_F32, y_total);
    ggml_backend_buffer_t abuf = ggml_backend_alloc_ctx_tensors(actx, gpu);
    if (!abuf) { fprintf(stderr, "e2e_plugin: activation allocation failed\n"); return 1; }

    // ---- the graph: every MUL_MAT reads a view of X; its dst is placed inside Y ---------------------------------------------------
    ggml_context * gctx = ggml_init(ip);
    ggml_cgraph * gf = ggml_new_graph_custom(gctx, 2 * n_mm + 16, false);
    for (int i = 0; i < n_mm; ++i) {
        ggml_tensor * xv = ggml_view_1d(gctx, X, mms[i].k, (size_t) x_off[mms[i].xslot] * sizeof(float));
        ggml_backend_view_init(xv);
        ggml_tensor * y = ggml_mul_mat(gctx, W[i], xv);
        ggml_backend_tensor_alloc(abuf, y, (char *) Y->data + (size_t) y_off[i] * sizeof(float));
        ggml_build_forward_expand(gf, y);
    }

    // ---- pinned host memory from the plug-in's host buffer type -------------------------------------------------------------------
    ggml_backend_buffer_type_t hbuft = ggml_backend_dev_host_buffer_type(dev);
    ggml_context * hctx = ggml_init(ip);
    ggml_tensor * Xh = ggml_new_tensor_1d(hctx, GGML_TYPE_F32, x_total);
    ggml_tensor * Yh = ggml_new_tensor_1d(hctx, GGML_TYPE_F32, y_total);
    ggml_backend_buffer_t hbuf = ggml_backend_alloc_ctx_tensors_from_buft(hctx, hbuft ? hbuft : ggml_backend_cpu_buffer_type());
    if (!hbuf) { fprintf(stderr, "e2e_plugin: host allocation failed\n"); return 1; }
    float * xh = (float *) Xh->data, * yh = (float *) Yh->data;
    for (int64_t i = 0; i < x_total; ++i) xh[i] = (float) ((double) (rnd() >> 11) / 9007199254740992.0 * 2.0 - 1.0);

    auto step = [&]() {
        ggml_backend_tensor_set_async(gpu, X, xh, 0, (size_t) x_total * sizeof(float));
        if (ggml_backend_graph_compute_async(gpu, gf) != GGML_STATUS_SUCCESS) { fprintf(stderr, "e2e_plugin: graph compute failed\n"); exit(1); }
        ggml_backend_tensor_get_async(gpu, Y, yh, 0, (size_t) y_total * sizeof(float));
        ggml_backend_synchronize(gpu);
    };
    for (int i = 0; i < warmup; ++i) step();
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < steps; ++i) step();
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() / steps;

    // ---- spot check against the reference's CPU backend: the second matmul of the list (wk: small), same bytes, same activations ---
    double err = -1.0;
    {
        ggml_backend_t cpu = ggml_backend_init_by_type(GGML_BACKEND_DEVICE_TYPE_CPU, nullptr);
        const int i = 1;
        std::vector<uint8_t> wb(ggml_nbytes(W[i]));
        ggml_backend_tensor_get(W[i], wb.data(), 0, wb.size());
        ggml_init_params cip = { ggml_tensor_overhead() * 8 + ggml_graph_overhead(), nullptr, true };
        ggml_context * c = ggml_init(cip);
        ggml_tensor * w = ggml_new_tensor_2d(c, GGML_TYPE_Q4_K, mms[i].k, mms[i].m);
        ggml_tensor * x = ggml_new_tensor_1d(c, GGML_TYPE_F32, mms[i].k);
        ggml_tensor * y = ggml_mul_mat(c, w, x);
        ggml_cgraph * g = ggml_new_graph(c);
        ggml_build_forward_expand(g, y);
        ggml_backend_buffer_t b = ggml_backend_alloc_ctx_tensors(c, cpu);
        ggml_backend_tensor_set(w, wb.data(), 0, wb.size());
        ggml_backend_tensor_set(x, xh + x_off[mms[i].xslot], 0, (size_t) mms[i].k * sizeof(float));
        ggml_backend_graph_compute(cpu, g);
        std::vector<float> ref((size_t) mms[i].m);
        ggml_backend_tensor_get(y, ref.data(), 0, ref.size() * sizeof(float));
        double num = 0, den = 0;
        for (size_t r = 0; r < ref.size(); ++r) { const double d = (double) yh[y_off[i] + r] - ref[r]; num += d * d; den += (double) ref[r] * ref[r]; }
        err = den > 0 ? num / den : (num > 0 ? INFINITY : 0.0);
        ggml_backend_buffer_free(b);
        ggml_free(c);
        ggml_backend_free(cpu);
    }

    double algo_bytes = 0;
    for (const mm & q : mms) algo_bytes += (double) q.m * q.k * 4.5 / 8 + (double) q.k * 4 + (double) q.m * 4;
    printf("{\"ms_per_step\": %.4f, \"gbps\": %.2f, \"n_matmuls\": %d, \"h2d_bytes_per_step\": %lld, \"d2h_bytes_per_step\": %lld, \"weights_gb\": %.3f, "
           "\"nmse_vs_cpu_backend\": %.3e, \"caps\": {\"async\": %d, \"host_buffer\": %d, \"events\": %d}, \"host_memory\": \"%s\", "
           "\"api\": \"ggml_backend_tensor_set_async + ggml_backend_graph_compute_async + ggml_backend_tensor_get_async + ggml_backend_synchronize "
           "on the plug-in libggml-b200.so (GGML_BACKEND_PATH), hosted by the reference's unmodified libggml\"}\n",
           ms, algo_bytes / ms / 1e6, n_mm, (long long) (x_total * 4), (long long) (y_total * 4), weight_bytes / 1e9, err,
           (int) props.caps.async, (int) props.caps.host_buffer, (int) props.caps.events, hbuft ? ggml_backend_buft_name(hbuft) : "pageable");
    const bool ok = err >= 0 && err <= 5e-4;
    ggml_backend_free(gpu);
    return ok ? 0 : 1;
}
