// oracle/reorder_check.cpp — TEST HARNESS (CPU only, no device needed) for the node reordering of the plug-in's graph_optimize hook
// (backend/ggml-b200.cpp: reorder_shared_activation_mul_mats, exported for this harness as ggml_backend_b200_reorder_graph).
//
// Builds cgraphs with the reference's own ggml (no_alloc contexts: shapes and edges only), in the node order llama.cpp produces —
// wq, ROPE(q), wk, ROPE(k), wv ... gate, up — lets the plug-in reorder them and checks that
//   * the nodes are the same set and every node still comes after all of its sources (and after the tensor it is a view of);
//   * the quantized MUL_MATs that share their activations are adjacent afterwards (q / k / v, gate / up);
//   * MUL_MATs separated by an op that works in place on their shared activations are NOT moved across it;
//   * MUL_MATs with different activations, or float weights, stay where they were.
// usage: reorder_check <path to libggml-b200.so>          prints OK / a description of the first violation
#include <dlfcn.h>

#include <cstdio>
#include <cstdlib>
#include <set>
#include <string>
#include <vector>

#include "ggml.h"

typedef void (*reorder_fn)(ggml_cgraph *);

static int fail(const char * what, int i = -1) {
    if (i >= 0) printf("FAIL: %s (node %d)\n", what, i); else printf("FAIL: %s\n", what);
    return 1;
}

static int index_of(const ggml_cgraph * gf, const ggml_tensor * t) {
    for (int i = 0; i < ggml_graph_n_nodes((ggml_cgraph *) gf); ++i) if (ggml_graph_node((ggml_cgraph *) gf, i) == t) return i;
    return -1;
}

// every node after its sources and after its view source; same node set as before
static int check_valid(ggml_cgraph * gf, const std::vector<ggml_tensor *> & before) {
    const int n = ggml_graph_n_nodes(gf);
    if (n != (int) before.size()) return fail("node count changed");
    std::multiset<ggml_tensor *> a(before.begin(), before.end()), b;
    for (int i = 0; i < n; ++i) b.insert(ggml_graph_node(gf, i));
    if (a != b) return fail("node set changed");
    for (int i = 0; i < n; ++i) {
        const ggml_tensor * t = ggml_graph_node(gf, i);
        for (int j = 0; j < GGML_MAX_SRC; ++j) {
            if (!t->src[j]) continue;
            const int k = index_of(gf, t->src[j]);
            if (k >= i) return fail("a node precedes one of its sources", i);
        }
        if (t->view_src) {
            const int k = index_of(gf, t->view_src);
            if (k >= i) return fail("a view precedes its source", i);
        }
    }
    return 0;
}

static std::vector<ggml_tensor *> nodes_of(ggml_cgraph * gf) {
    std::vector<ggml_tensor *> v;
    for (int i = 0; i < ggml_graph_n_nodes(gf); ++i) v.push_back(ggml_graph_node(gf, i));
    return v;
}

int main(int argc, char ** argv) {
    const char * path = argc > 1 ? argv[1] : getenv("GGML_BACKEND_PATH");
    if (!path) { fprintf(stderr, "usage: %s <libggml-b200.so>\n", argv[0]); return 2; }
    void * h = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    if (!h) { printf("FAIL: dlopen: %s\n", dlerror()); return 1; }
    reorder_fn reorder = (reorder_fn) dlsym(h, "ggml_backend_b200_reorder_graph");
    if (!reorder) return fail("ggml_backend_b200_reorder_graph not exported");

    ggml_init_params ip = { (size_t) 64 << 20, nullptr, true };
    ggml_context * ctx = ggml_init(ip);
    const int64_t n_embd = 1024, n_head = 8, n_head_kv = 2, hd = 128, n_ff = 2048, n_tok = 3, n_layer = 3;

    // ---- case 1: decoder layers in llama.cpp's node order ---------------------------------------------------------------------
    {
        ggml_cgraph * gf = ggml_new_graph_custom(ctx, 4096, false);
        ggml_tensor * x   = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, n_embd, n_tok);
        ggml_tensor * pos = ggml_new_tensor_1d(ctx, GGML_TYPE_I32, n_tok);
        ggml_set_input(x); ggml_set_input(pos);
        std::vector<std::vector<ggml_tensor *>> groups;
        for (int l = 0; l < n_layer; ++l) {
            const ggml_type wt = l == 1 ? GGML_TYPE_Q8_0 : GGML_TYPE_Q4_K;
            ggml_tensor * wn = ggml_new_tensor_1d(ctx, GGML_TYPE_F32, n_embd);
            ggml_tensor * wq = ggml_new_tensor_2d(ctx, wt, n_embd, n_head * hd), * wk = ggml_new_tensor_2d(ctx, wt, n_embd, n_head_kv * hd);
            ggml_tensor * wv = ggml_new_tensor_2d(ctx, wt, n_embd, n_head_kv * hd), * wo = ggml_new_tensor_2d(ctx, wt, n_head * hd, n_embd);
            ggml_tensor * wg = ggml_new_tensor_2d(ctx, wt, n_embd, n_ff), * wu = ggml_new_tensor_2d(ctx, wt, n_embd, n_ff), * wd = ggml_new_tensor_2d(ctx, wt, n_ff, n_embd);
            ggml_tensor * cur = ggml_mul(ctx, ggml_rms_norm(ctx, x, 1e-5f), wn);
            ggml_tensor * q = ggml_mul_mat(ctx, wq, cur), * k = ggml_mul_mat(ctx, wk, cur), * v = ggml_mul_mat(ctx, wv, cur);
            ggml_tensor * qr = ggml_rope(ctx, ggml_reshape_3d(ctx, q, hd, n_head, n_tok), pos, (int) hd, 0);
            ggml_tensor * kr = ggml_rope(ctx, ggml_reshape_3d(ctx, k, hd, n_head_kv, n_tok), pos, (int) hd, 0);
            ggml_build_forward_expand(gf, qr);           // wq, reshape, ROPE ...
            ggml_build_forward_expand(gf, kr);           // ... wk, reshape, ROPE ...
            ggml_build_forward_expand(gf, v);            // ... wv
            // (attention itself does not matter here: the projection reads q; k and v only have to be in the graph)
            ggml_tensor * att = ggml_add(ctx, ggml_reshape_2d(ctx, qr, n_head * hd, n_tok),
                                         ggml_repeat(ctx, ggml_add(ctx, ggml_reshape_2d(ctx, kr, n_head_kv * hd, n_tok), v), q));
            ggml_tensor * x2  = ggml_add(ctx, x, ggml_mul_mat(ctx, wo, att));
            ggml_tensor * n2  = ggml_mul(ctx, ggml_rms_norm(ctx, x2, 1e-5f), wn);
            ggml_tensor * g = ggml_mul_mat(ctx, wg, n2);
            ggml_tensor * gs = ggml_silu(ctx, g);         // gate, SILU, up: the up product is separated from the gate product
            ggml_build_forward_expand(gf, gs);
            ggml_tensor * u = ggml_mul_mat(ctx, wu, n2);
            x = ggml_add(ctx, x2, ggml_mul_mat(ctx, wd, ggml_mul(ctx, gs, u)));
            ggml_build_forward_expand(gf, x);
            groups.push_back({q, k, v});
            groups.push_back({g, u});
        }
        const std::vector<ggml_tensor *> before = nodes_of(gf);
        for (const auto & grp : groups) {                 // the premise: they are NOT adjacent as built
            if (index_of(gf, grp[1]) == index_of(gf, grp[0]) + 1 && grp.size() == 2) return fail("test premise: gate / up already adjacent");
        }
        reorder(gf);
        if (int rc = check_valid(gf, before)) return rc;
        for (const auto & grp : groups) {
            const int i0 = index_of(gf, grp[0]);
            for (size_t j = 1; j < grp.size(); ++j) if (index_of(gf, grp[j]) != i0 + (int) j) return fail("MUL_MATs sharing their activations are not adjacent", index_of(gf, grp[j]));
        }
        printf("layers: %d nodes, %zu groups adjacent\n", ggml_graph_n_nodes(gf), groups.size());
    }

    // ---- case 2: an in-place op on the shared activations between two MUL_MATs pins the order ------------------------------------
    {
        ggml_cgraph * gf = ggml_new_graph_custom(ctx, 64, false);
        ggml_tensor * x0 = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, n_embd, 1);
        ggml_set_input(x0);
        ggml_tensor * x  = ggml_scale(ctx, x0, 0.5f);                              // the shared activations: a computed tensor
        ggml_tensor * w1 = ggml_new_tensor_2d(ctx, GGML_TYPE_Q4_K, n_embd, 256), * w2 = ggml_new_tensor_2d(ctx, GGML_TYPE_Q4_K, n_embd, 512);
        ggml_tensor * t1 = ggml_mul_mat(ctx, w1, x);
        ggml_tensor * xs = ggml_scale_inplace(ctx, x, 2.0f);                        // writes x's memory
        ggml_tensor * t2 = ggml_mul_mat(ctx, w2, x);                                // (reads x AFTER the scaling, by node order)
        ggml_build_forward_expand(gf, t1);
        ggml_build_forward_expand(gf, xs);
        ggml_build_forward_expand(gf, t2);
        const std::vector<ggml_tensor *> before = nodes_of(gf);
        reorder(gf);
        if (int rc = check_valid(gf, before)) return rc;
        if (nodes_of(gf) != before) return fail("a MUL_MAT was moved across an in-place op on its activations");
        printf("in-place pin: order kept\n");
    }

    // ---- case 3: different activations / float weights are left alone --------------------------------------------------------------
    {
        ggml_cgraph * gf = ggml_new_graph_custom(ctx, 64, false);
        ggml_tensor * xa = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, n_embd, 1), * xb = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, n_embd, 1);
        ggml_set_input(xa); ggml_set_input(xb);
        ggml_tensor * w1 = ggml_new_tensor_2d(ctx, GGML_TYPE_Q4_K, n_embd, 256), * w2 = ggml_new_tensor_2d(ctx, GGML_TYPE_Q4_K, n_embd, 256);
        ggml_tensor * wf = ggml_new_tensor_2d(ctx, GGML_TYPE_F16, n_embd, 256), * w3 = ggml_new_tensor_2d(ctx, GGML_TYPE_Q6_K, n_embd, 256);
        ggml_tensor * a = ggml_mul_mat(ctx, w1, xa);
        ggml_tensor * s = ggml_silu(ctx, a);
        ggml_tensor * b = ggml_mul_mat(ctx, w2, xb);          // other activations
        ggml_tensor * c = ggml_mul_mat(ctx, wf, xa);          // float weights
        ggml_tensor * d = ggml_mul_mat(ctx, w3, xa);          // another weight type: not one list
        for (ggml_tensor * t : {s, b, c, d}) ggml_build_forward_expand(gf, t);
        const std::vector<ggml_tensor *> before = nodes_of(gf);
        reorder(gf);
        if (int rc = check_valid(gf, before)) return rc;
        if (nodes_of(gf) != before) return fail("MUL_MATs that do not form a list were moved");
        printf("unrelated products: order kept\n");
    }
    ggml_free(ctx);
    printf("OK\n");
    return 0;
}
