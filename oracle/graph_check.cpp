// oracle/graph_check.cpp — TEST INFRASTRUCTURE (never part of the product path).
//
// Drives the ggml backend plug-in the way llama.cpp does, not op by op like test-backend-ops: a transformer-layer-shaped
// graph (rms_norm -> wq / wk / wv -> ... -> wo -> ffn_gate / ffn_up -> silu * up -> ffn_down, residuals) is scheduled by
// the reference's own ggml_backend_sched over {B200 plug-in, CPU}: the quantized weights live in a B200 buffer, so every
// MUL_MAT lands on the plug-in (ggml-backend.cpp ggml_backend_sched_backend_id_from_cur: the weight's buffer decides) and
// all other ops run on the CPU backend, with the scheduler copying activations across — exactly the split structure a
// llama.cpp user gets with this MUL_MAT-only backend.  The same graph is then computed by the CPU backend alone and the
// outputs are compared (NMSE).  Links the reference's libggml / libggml-base / libggml-cpu built by oracle/Makefile.
//
// usage: graph_check <weight type: q4_0|q8_0|q4_K|q6_K> <n_tokens> <n_layers>      (GGML_BACKEND_PATH = the plug-in)
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include "ggml.h"
#include "ggml-backend.h"
#include "ggml-cpu.h"

namespace {

constexpr int D = 1024, D_KV = 256, D_FF = 2816;

struct layer_w { ggml_tensor * wq, * wk, * wv, * wo, * wg, * wu, * wd; };

struct model {
    ggml_context *        ctx = nullptr;
    ggml_backend_buffer_t buf = nullptr;
    std::vector<layer_w>  layers;
};

ggml_type parse_type(const char * s) {
    if (!strcmp(s, "q4_0")) return GGML_TYPE_Q4_0;
    if (!strcmp(s, "q8_0")) return GGML_TYPE_Q8_0;
    if (!strcmp(s, "q4_K")) return GGML_TYPE_Q4_K;
    if (!strcmp(s, "q6_K")) return GGML_TYPE_Q6_K;
    fprintf(stderr, "unknown type %s\n", s);
    exit(2);
}

// quantized weights with the reference's own quantizer, identical bytes for both runs
std::vector<uint8_t> make_weight(ggml_type type, int64_t k, int64_t m, std::mt19937 & rng) {
    std::vector<float> f((size_t) k * m);
    std::normal_distribution<float> dist(0.0f, 1.0f / std::sqrt((float) k));
    for (auto & v : f) v = dist(rng);
    std::vector<uint8_t> q(ggml_row_size(type, k) * m);
    ggml_quantize_chunk(type, f.data(), q.data(), 0, m, k, nullptr);
    return q;
}

model load_model(ggml_backend_t backend, ggml_type type, int n_layers, const std::vector<std::vector<uint8_t>> & blobs) {
    model M;
    ggml_init_params ip = { ggml_tensor_overhead() * (size_t) (n_layers * 7 + 4), nullptr, true };
    M.ctx = ggml_init(ip);
    for (int l = 0; l < n_layers; ++l) {
        layer_w w;
        w.wq = ggml_new_tensor_2d(M.ctx, type, D, D);
        w.wk = ggml_new_tensor_2d(M.ctx, type, D, D_KV);
        w.wv = ggml_new_tensor_2d(M.ctx, type, D, D_KV);
        w.wo = ggml_new_tensor_2d(M.ctx, type, D, D);
        w.wg = ggml_new_tensor_2d(M.ctx, type, D, D_FF);
        w.wu = ggml_new_tensor_2d(M.ctx, type, D, D_FF);
        w.wd = ggml_new_tensor_2d(M.ctx, type, D_FF, D);
        M.layers.push_back(w);
    }
    M.buf = ggml_backend_alloc_ctx_tensors(M.ctx, backend);
    if (!M.buf) { fprintf(stderr, "weight allocation failed\n"); exit(1); }
    ggml_backend_buffer_set_usage(M.buf, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
    size_t b = 0;
    for (auto & w : M.layers) {
        for (ggml_tensor * t : { w.wq, w.wk, w.wv, w.wo, w.wg, w.wu, w.wd }) {
            ggml_backend_tensor_set(t, blobs[b].data(), 0, blobs[b].size());
            ++b;
        }
    }
    return M;
}

struct graph { ggml_context * ctx; ggml_cgraph * gf; ggml_tensor * inp, * out, * kv; };

graph build_graph(const model & M, int n_tokens) {
    graph G;
    ggml_init_params ip = { ggml_tensor_overhead() * 4096 + ggml_graph_overhead(), nullptr, true };
    G.ctx = ggml_init(ip);
    G.gf  = ggml_new_graph(G.ctx);
    G.inp = ggml_new_tensor_2d(G.ctx, GGML_TYPE_F32, D, n_tokens);
    ggml_set_input(G.inp);
    ggml_tensor * h  = G.inp;
    ggml_tensor * kv = nullptr;
    for (const layer_w & w : M.layers) {
        ggml_tensor * cur = ggml_rms_norm(G.ctx, h, 1e-5f);
        ggml_tensor * q   = ggml_mul_mat(G.ctx, w.wq, cur);
        ggml_tensor * k   = ggml_mul_mat(G.ctx, w.wk, cur);
        ggml_tensor * v   = ggml_mul_mat(G.ctx, w.wv, cur);
        // a stand-in for attention that keeps q, k and v all live: q scaled by tanh of the (k + v) mean pattern
        ggml_tensor * kvs = ggml_add(G.ctx, k, v);
        kv = kv ? ggml_add(G.ctx, kv, kvs) : kvs;
        ggml_tensor * gate = ggml_repeat(G.ctx, ggml_tanh(G.ctx, kvs), q);        // [256, n] tiled to [1024, n]
        ggml_tensor * att  = ggml_mul(G.ctx, q, gate);
        ggml_tensor * o    = ggml_mul_mat(G.ctx, w.wo, att);
        h = ggml_add(G.ctx, h, o);
        ggml_tensor * nrm = ggml_rms_norm(G.ctx, h, 1e-5f);
        ggml_tensor * g   = ggml_mul_mat(G.ctx, w.wg, nrm);
        ggml_tensor * u   = ggml_mul_mat(G.ctx, w.wu, nrm);
        ggml_tensor * act = ggml_swiglu_split(G.ctx, g, u);                       // as llm_graph_context::build_ffn does (src/llama-graph.cpp:1225)
        ggml_tensor * d   = ggml_mul_mat(G.ctx, w.wd, act);
        h = ggml_add(G.ctx, h, d);
    }
    G.out = h;  ggml_set_output(G.out);
    G.kv  = kv; ggml_set_output(G.kv);
    ggml_build_forward_expand(G.gf, G.out);
    ggml_build_forward_expand(G.gf, G.kv);
    return G;
}

struct result { std::vector<float> out, kv; int n_splits; };

result run(std::vector<ggml_backend_t> backends, ggml_backend_t weight_backend, ggml_type type, int n_tokens, int n_layers,
           const std::vector<std::vector<uint8_t>> & blobs, const std::vector<float> & x) {
    model M = load_model(weight_backend, type, n_layers, blobs);
    graph G = build_graph(M, n_tokens);
    ggml_backend_sched_t sched = ggml_backend_sched_new(backends.data(), nullptr, (int) backends.size(), 4096, false, true);
    if (!ggml_backend_sched_alloc_graph(sched, G.gf)) { fprintf(stderr, "sched alloc failed\n"); exit(1); }
    result R;
    for (int rep = 0; rep < 2; ++rep) {        // twice: the second pass re-uses every buffer (stale data would show)
        // the input is set before EVERY pass, as llama.cpp does: ggml-alloc may hand an input's memory to a later node of the graph
        // once the input has been consumed (seen with the REPEAT node of this graph on the CPU backend), so it does not survive a pass
        ggml_backend_tensor_set(G.inp, x.data(), 0, x.size() * sizeof(float));
        if (ggml_backend_sched_graph_compute(sched, G.gf) != GGML_STATUS_SUCCESS) { fprintf(stderr, "graph compute failed\n"); exit(1); }
    }
    R.n_splits = ggml_backend_sched_get_n_splits(sched);
    R.out.resize(ggml_nelements(G.out));
    R.kv.resize(ggml_nelements(G.kv));
    ggml_backend_tensor_get(G.out, R.out.data(), 0, R.out.size() * sizeof(float));
    ggml_backend_tensor_get(G.kv, R.kv.data(), 0, R.kv.size() * sizeof(float));
    ggml_backend_sched_free(sched);
    ggml_free(G.ctx);
    ggml_backend_buffer_free(M.buf);
    ggml_free(M.ctx);
    return R;
}

double nmse(const std::vector<float> & ref, const std::vector<float> & got) {
    double num = 0, den = 0;
    for (size_t i = 0; i < ref.size(); ++i) {
        const double d = (double) got[i] - ref[i];
        num += d * d; den += (double) ref[i] * ref[i];
    }
    return den > 0 ? num / den : (num > 0 ? INFINITY : 0.0);
}

}  // namespace

int main(int argc, char ** argv) {
    if (argc < 4) { fprintf(stderr, "usage: %s <q4_0|q8_0|q4_K|q6_K> <n_tokens> <n_layers>\n", argv[0]); return 2; }
    const ggml_type type = parse_type(argv[1]);
    const int n_tokens = atoi(argv[2]), n_layers = atoi(argv[3]);

    ggml_backend_load_all();
    ggml_backend_dev_t dev = ggml_backend_dev_by_name("B2000");
    if (!dev) { fprintf(stderr, "graph_check: device B2000 not registered (GGML_BACKEND_PATH?)\n"); return 3; }
    ggml_backend_t gpu = ggml_backend_dev_init(dev, nullptr);
    ggml_backend_t cpu = ggml_backend_init_by_type(GGML_BACKEND_DEVICE_TYPE_CPU, nullptr);
    if (!gpu || !cpu) { fprintf(stderr, "graph_check: backend init failed\n"); return 3; }

    std::mt19937 rng(1234);
    std::vector<std::vector<uint8_t>> blobs;
    for (int l = 0; l < n_layers; ++l) {
        const int64_t km[7][2] = { {D, D}, {D, D_KV}, {D, D_KV}, {D, D}, {D, D_FF}, {D, D_FF}, {D_FF, D} };
        for (auto & s : km) blobs.push_back(make_weight(type, s[0], s[1], rng));
    }
    std::vector<float> x((size_t) D * n_tokens);
    std::normal_distribution<float> dist(0.0f, 1.0f);
    for (auto & v : x) v = dist(rng);

    const result ref = run({ cpu }, cpu, type, n_tokens, n_layers, blobs, x);
    const result got = run({ gpu, cpu }, gpu, type, n_tokens, n_layers, blobs, x);

    const double e_out = nmse(ref.out, got.out), e_kv = nmse(ref.kv, got.kv);
    if (getenv("GRAPH_CHECK_DEBUG")) {
        for (size_t i = 0; i < 6 && i < ref.kv.size(); ++i) printf("kv[%zu] ref %g got %g | out ref %g got %g\n", i, ref.kv[i], got.kv[i], ref.out[i], got.out[i]);
    }
    bool finite = true;
    for (float v : got.out) finite = finite && std::isfinite(v);
    printf("graph_check type=%s n_tokens=%d n_layers=%d splits_cpu=%d splits_b200=%d nmse_out=%.3e nmse_kv=%.3e finite=%d\n",
           argv[1], n_tokens, n_layers, ref.n_splits, got.n_splits, e_out, e_kv, (int) finite);
    ggml_backend_free(gpu);
    ggml_backend_free(cpu);
    // every single MUL_MAT is within NMSE 5e-4 of the CPU backend (test-backend-ops); the CPU quantizes Q4_K / Q6_K activations
    // to Q8_K and this backend to Q8_1, so along a chain of 4 dependent matmuls per layer the differences add up
    const double tol = 5e-4 * 4 * n_layers;
    // GGML_B200_OPS=mul_mat: only the MUL_MATs run on the plug-in, so the scheduler must have cut the graph into many splits;
    // default: the plug-in takes every op of this graph and the whole graph is one split
    const char * ops_env = getenv("GGML_B200_OPS");
    const bool mm_only = ops_env && !strcmp(ops_env, "mul_mat");
    const bool ok = finite && e_out <= tol && e_kv <= tol && (mm_only ? got.n_splits > 1 : got.n_splits <= 2);
    printf(ok ? "OK\n" : "FAIL\n");
    return ok ? 0 : 1;
}
