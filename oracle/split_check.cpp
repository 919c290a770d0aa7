// oracle/split_check.cpp — TEST INFRASTRUCTURE (never part of the product path).
//
// Exercises the plug-in's split buffer type the way llama.cpp's -sm row does (src/llama-model.cpp:879-902 asks the registry for
// "ggml_backend_split_buffer_type"; src/llama-model-loader.cpp:919-931 probes whether the device can run the op with a weight
// in that buffer type): the same quantized matrix is placed once in a split buffer and once in the device's plain buffer, the
// MUL_MAT of both is computed by the plug-in's backend and by the reference's CPU backend, and
//   * the split result must equal the unsplit result bit for bit (the shards run the same per-row arithmetic),
//   * both must be within test-backend-ops' NMSE of the CPU backend,
//   * get_tensor of the split weights must return the bytes that were set.
// On a box with one GPU run it with GGML_B200_SPLIT_EMULATE=N (N shards on the one device).
//
// usage: split_check <q4_0|q8_0|q4_K|q6_K> <m> <k> <n> [fraction of rows on shard 0, default 0.5]
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "ggml.h"
#include "ggml-alloc.h"
#include "ggml-backend.h"
#include "ggml-cpu.h"

static ggml_type parse_type(const char * s) {
    if (!strcmp(s, "q4_0")) return GGML_TYPE_Q4_0;
    if (!strcmp(s, "q8_0")) return GGML_TYPE_Q8_0;
    if (!strcmp(s, "q4_K")) return GGML_TYPE_Q4_K;
    if (!strcmp(s, "q6_K")) return GGML_TYPE_Q6_K;
    fprintf(stderr, "unknown type %s\n", s);
    exit(2);
}

static double nmse(const std::vector<float> & ref, const std::vector<float> & got) {
    double num = 0, den = 0;
    for (size_t i = 0; i < ref.size(); ++i) { const double d = (double) got[i] - ref[i]; num += d * d; den += (double) ref[i] * ref[i]; }
    return den > 0 ? num / den : (num > 0 ? INFINITY : 0.0);
}

// y = W x on `backend`, W in a buffer of `wbuft`
static std::vector<float> run(ggml_backend_t backend, ggml_backend_buffer_type_t wbuft, ggml_type type, int64_t m, int64_t k, int64_t n,
                              const std::vector<uint8_t> & wq, const std::vector<float> & x, bool check_readback) {
    ggml_init_params ip = { ggml_tensor_overhead() * 8 + ggml_graph_overhead(), nullptr, true };
    ggml_context * wctx = ggml_init(ip);
    ggml_tensor * W = ggml_new_tensor_2d(wctx, type, k, m);
    ggml_backend_buffer_t wbuf = ggml_backend_alloc_ctx_tensors_from_buft(wctx, wbuft);
    if (!wbuf) { fprintf(stderr, "weight allocation failed\n"); exit(1); }
    ggml_backend_buffer_set_usage(wbuf, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
    ggml_backend_tensor_set(W, wq.data(), 0, wq.size());
    if (check_readback) {
        std::vector<uint8_t> back(wq.size());
        ggml_backend_tensor_get(W, back.data(), 0, back.size());
        if (back != wq) { fprintf(stderr, "split_check: get_tensor does not return the bytes that were set\n"); exit(1); }
    }
    ggml_context * ctx = ggml_init(ip);
    ggml_tensor * X = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, k, n);
    ggml_tensor * Y = ggml_mul_mat(ctx, W, X);
    ggml_cgraph * gf = ggml_new_graph(ctx);
    ggml_build_forward_expand(gf, Y);
    ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors(ctx, backend);
    if (!buf) { fprintf(stderr, "activation allocation failed\n"); exit(1); }
    ggml_backend_tensor_set(X, x.data(), 0, x.size() * sizeof(float));
    std::vector<float> y((size_t) m * n);
    for (int rep = 0; rep < 2; ++rep) {
        if (ggml_backend_graph_compute(backend, gf) != GGML_STATUS_SUCCESS) { fprintf(stderr, "graph compute failed\n"); exit(1); }
    }
    ggml_backend_tensor_get(Y, y.data(), 0, y.size() * sizeof(float));
    ggml_backend_buffer_free(buf);
    ggml_backend_buffer_free(wbuf);
    ggml_free(ctx);
    ggml_free(wctx);
    return y;
}

int main(int argc, char ** argv) {
    if (argc < 5) { fprintf(stderr, "usage: %s <type> <m> <k> <n> [fraction on shard 0]\n", argv[0]); return 2; }
    const ggml_type type = parse_type(argv[1]);
    const int64_t m = atoll(argv[2]), k = atoll(argv[3]), n = atoll(argv[4]);
    const float f0 = argc > 5 ? (float) atof(argv[5]) : 0.5f;

    ggml_backend_load_all();
    ggml_backend_dev_t dev = ggml_backend_dev_by_name("B2000");
    if (!dev) { fprintf(stderr, "split_check: device B2000 not registered (GGML_BACKEND_PATH?)\n"); return 3; }
    auto split_fn = (ggml_backend_split_buffer_type_t) ggml_backend_reg_get_proc_address(ggml_backend_dev_backend_reg(dev), "ggml_backend_split_buffer_type");
    if (!split_fn) { fprintf(stderr, "split_check: the registry does not answer ggml_backend_split_buffer_type\n"); return 1; }
    float tensor_split[16] = { f0, 1.0f - f0 };
    const char * emu = getenv("GGML_B200_SPLIT_EMULATE");
    for (int i = 2; i < (emu ? atoi(emu) : 0); ++i) tensor_split[i] = (1.0f - f0);
    ggml_backend_buffer_type_t sbuft = split_fn(0, tensor_split);
    if (!sbuft) { fprintf(stderr, "split_check: no split buffer type\n"); return 1; }
    if (split_fn(0, tensor_split) != sbuft) { fprintf(stderr, "split_check: the same (device, split) must give the same buffer type object\n"); return 1; }
    if (!ggml_backend_dev_supports_buft(dev, sbuft)) { fprintf(stderr, "split_check: the main device rejects its own split buffer type\n"); return 1; }

    // the weight placement probe of llama.cpp (src/llama-model-loader.cpp:919-931): a zero-size buffer of the type, then supports_op
    {
        ggml_init_params ip = { ggml_tensor_overhead() * 8, nullptr, true };
        ggml_context * pctx = ggml_init(ip);
        ggml_tensor * w = ggml_new_tensor_2d(pctx, type, k, m);
        ggml_tensor * b = ggml_new_tensor_2d(pctx, GGML_TYPE_F32, k, n);
        ggml_tensor * op = ggml_mul_mat(pctx, w, b);
        w->buffer = ggml_backend_buft_alloc_buffer(sbuft, 0);
        const bool ok = ggml_backend_dev_supports_op(dev, op);
        ggml_tensor * as = ggml_new_tensor_3d(pctx, type, k, m, 4);
        ggml_tensor * ids = ggml_new_tensor_2d(pctx, GGML_TYPE_I32, 2, n);
        ggml_tensor * bb = ggml_new_tensor_3d(pctx, GGML_TYPE_F32, k, 1, n);
        ggml_tensor * opid = ggml_mul_mat_id(pctx, as, bb, ids);
        as->buffer = w->buffer;
        const bool ok_id = ggml_backend_dev_supports_op(dev, opid);
        ggml_backend_buffer_free(w->buffer);
        ggml_free(pctx);
        if (!ok)   { fprintf(stderr, "split_check: supports_op rejects MUL_MAT with a split weight (-sm row would not place weights here)\n"); return 1; }
        if (ok_id) { fprintf(stderr, "split_check: MUL_MAT_ID with split weights must be rejected (ggml-cuda.cu:2628)\n"); return 1; }
    }

    ggml_backend_t gpu = ggml_backend_dev_init(dev, nullptr);
    ggml_backend_t cpu = ggml_backend_init_by_type(GGML_BACKEND_DEVICE_TYPE_CPU, nullptr);
    if (!gpu || !cpu) { fprintf(stderr, "split_check: backend init failed\n"); return 3; }

    std::mt19937 rng(4321);
    std::vector<float> wf((size_t) k * m), x((size_t) k * n);
    std::normal_distribution<float> wd(0.0f, 1.0f / std::sqrt((float) k)), xd(0.0f, 1.0f);
    for (auto & v : wf) v = wd(rng);
    for (auto & v : x) v = xd(rng);
    std::vector<uint8_t> wq(ggml_row_size(type, k) * m);
    ggml_quantize_chunk(type, wf.data(), wq.data(), 0, m, k, nullptr);

    const std::vector<float> y_cpu   = run(cpu, ggml_backend_get_default_buffer_type(cpu), type, m, k, n, wq, x, false);
    const std::vector<float> y_plain = run(gpu, ggml_backend_dev_buffer_type(dev), type, m, k, n, wq, x, false);
    const std::vector<float> y_split = run(gpu, sbuft, type, m, k, n, wq, x, true);

    // Same bits: the shards run the per-row arithmetic of the unsplit kernel.  One documented exception: the unsplit tcgen05 GEMM may
    // split a long K in two when the matrix has few tiles (k >= 8192, n > 8), the shards never do — then the two K halves are
    // added in a different association and the results agree to fp32 rounding only.
    const bool may_split_k = n > 8 && k >= 8192;
    const bool same = may_split_k ? nmse(y_plain, y_split) <= 1e-10 : memcmp(y_plain.data(), y_split.data(), y_plain.size() * sizeof(float)) == 0;
    const double e_plain = nmse(y_cpu, y_plain), e_split = nmse(y_cpu, y_split);
    printf("split_check type=%s m=%lld k=%lld n=%lld buft=%s split==plain:%d nmse_plain=%.3e nmse_split=%.3e\n", argv[1], (long long) m, (long long) k,
           (long long) n, ggml_backend_buft_name(sbuft), (int) same, e_plain, e_split);
    ggml_backend_free(gpu);
    ggml_backend_free(cpu);
    const bool ok = same && e_plain <= 5e-4 && e_split <= 5e-4;
    printf(ok ? "OK\n" : "FAIL\n");
    return ok ? 0 : 1;
}
