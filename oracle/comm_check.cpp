// oracle/comm_check.cpp — TEST INFRASTRUCTURE (never part of the product path).
//
// The all-reduce the meta / tensor-parallel backend of the reference asks a backend registry for (ggml-backend.h:205-208; caller
// ggml-backend-meta.cpp:1522-1553): "ggml_backend_comm_init" over n backends of the plug-in, "ggml_backend_comm_allreduce_tensor" on one F32
// tensor per backend, "ggml_backend_comm_free".  Every backend's tensor must afterwards hold the element-wise sum in participant order
// (compared bit for bit with the same sum formed on the host); a tensor without GGML_TENSOR_FLAG_COMPUTE contributes zeros
// (ggml-cuda.cu:1205-1211).  With fewer GPUs than participants the backends share devices (peer access is then the same memory).
//
// usage: comm_check <n_backends> <n_elements> [index of a participant whose tensor is "not computed"]
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "ggml.h"
#include "ggml-alloc.h"
#include "ggml-backend.h"

int main(int argc, char ** argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s <n_backends> <n_elements> [not-computed index]\n", argv[0]); return 2; }
    const int n = atoi(argv[1]);
    const int64_t ne = atoll(argv[2]);
    const int skip = argc > 3 ? atoi(argv[3]) : -1;
    ggml_backend_load_all();
    ggml_backend_reg_t reg = ggml_backend_reg_by_name("B200");
    if (!reg) { fprintf(stderr, "comm_check: registry B200 not loaded (GGML_BACKEND_PATH?)\n"); return 3; }
    auto comm_init = (ggml_backend_comm_init_t) ggml_backend_reg_get_proc_address(reg, "ggml_backend_comm_init");
    auto comm_free = (ggml_backend_comm_free_t) ggml_backend_reg_get_proc_address(reg, "ggml_backend_comm_free");
    auto comm_allreduce = (ggml_backend_comm_allreduce_tensor_t) ggml_backend_reg_get_proc_address(reg, "ggml_backend_comm_allreduce_tensor");
    if (!comm_init || !comm_free || !comm_allreduce) { fprintf(stderr, "comm_check: the registry does not answer ggml_backend_comm_*\n"); return 1; }
    const size_t n_dev = ggml_backend_reg_dev_count(reg);
    std::vector<ggml_backend_t> backends;
    for (int i = 0; i < n; ++i) {
        ggml_backend_t b = ggml_backend_dev_init(ggml_backend_reg_dev_get(reg, (size_t) i % n_dev), nullptr);
        if (!b) { fprintf(stderr, "comm_check: backend init failed\n"); return 3; }
        backends.push_back(b);
    }
    void * comm = comm_init(backends.data(), backends.size());
    if (!comm) { fprintf(stderr, "comm_check: comm_init returned NULL\n"); return 1; }

    std::mt19937 rng(99);
    std::normal_distribution<float> dist(0.0f, 1.0f);
    std::vector<std::vector<float>> host(n, std::vector<float>((size_t) ne));
    std::vector<ggml_context *> ctxs;
    std::vector<ggml_backend_buffer_t> bufs;
    std::vector<ggml_tensor *> tensors;
    for (int i = 0; i < n; ++i) {
        for (auto & v : host[i]) v = dist(rng);
        ggml_init_params ip = { ggml_tensor_overhead() * 4, nullptr, true };
        ggml_context * c = ggml_init(ip);
        ggml_tensor * t = ggml_new_tensor_1d(c, GGML_TYPE_F32, ne);
        ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors(c, backends[i]);
        if (!buf) { fprintf(stderr, "comm_check: allocation failed\n"); return 1; }
        ggml_backend_tensor_set(t, host[i].data(), 0, (size_t) ne * 4);
        t->flags |= GGML_TENSOR_FLAG_COMPUTE;
        if (i == skip) { t->flags &= ~GGML_TENSOR_FLAG_COMPUTE; std::fill(host[i].begin(), host[i].end(), 0.0f); }
        ctxs.push_back(c); bufs.push_back(buf); tensors.push_back(t);
    }
    std::vector<float> want((size_t) ne);
    for (int64_t j = 0; j < ne; ++j) { float s = host[0][j]; for (int i = 1; i < n; ++i) s += host[i][j]; want[j] = s; }

    bool ok = true;
    for (int rep = 0; rep < 2 && ok; ++rep) {                       // twice: the second call reuses the events
        if (rep == 1) for (int i = 0; i < n; ++i) if (i != skip) ggml_backend_tensor_set(tensors[i], host[i].data(), 0, (size_t) ne * 4);
        if (!comm_allreduce(comm, tensors.data())) { fprintf(stderr, "comm_check: allreduce returned false\n"); ok = false; break; }
        for (int i = 0; i < n; ++i) ggml_backend_synchronize(backends[i]);
        for (int i = 0; i < n && ok; ++i) {
            std::vector<float> got((size_t) ne);
            ggml_backend_tensor_get(tensors[i], got.data(), 0, (size_t) ne * 4);
            if (memcmp(got.data(), want.data(), (size_t) ne * 4) != 0) {
                int64_t bad = 0; for (int64_t j = 0; j < ne; ++j) bad += got[j] != want[j];
                fprintf(stderr, "comm_check: participant %d differs in %lld of %lld elements (pass %d)\n", i, (long long) bad, (long long) ne, rep);
                ok = false;
            }
        }
    }
    // a type the all-reduce does not take must be declined, not mangled
    {
        ggml_tensor fake = *tensors[0];
        fake.type = GGML_TYPE_F16;
        std::vector<ggml_tensor *> ts = tensors; ts[0] = &fake;
        if (comm_allreduce(comm, ts.data())) { fprintf(stderr, "comm_check: an F16 tensor was accepted\n"); ok = false; }
    }
    printf("comm_check n=%d ne=%lld devices=%zu not_computed=%d\n", n, (long long) ne, n_dev, skip);
    comm_free(comm);
    for (int i = 0; i < n; ++i) { ggml_backend_buffer_free(bufs[i]); ggml_free(ctxs[i]); ggml_backend_free(backends[i]); }
    printf(ok ? "OK\n" : "FAIL\n");
    return ok ? 0 : 1;
}
