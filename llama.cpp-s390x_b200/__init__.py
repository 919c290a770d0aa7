"""llama.cpp-s390x_b200 — host-side mirror (Python) of the B200 quantized MUL_MAT / MUL_MAT_ID path.

The product is the native code next to this file:
  libggml_b200_kernels.so  C ABI (include/ggml_b200.h) over hand-written sm_100a CUDA kernels
  libggml-b200.so          ggml backend plug-in (GGML_BACKEND_PATH) for unmodified llama.cpp / test-backend-ops

This package only binds the C ABI with ctypes so tests and bench.py can drive it with torch tensors as device
memory.  There is no fallback of any kind: if the library or a B200 is missing, calls raise.

The directory name is not a Python identifier; load it with `load_package()` from tests/conftest.py / bench.py or
    importlib.util.spec_from_file_location("llama_cpp_s390x_b200", ".../llama.cpp-s390x_b200/__init__.py",
                                           submodule_search_locations=[...])
"""
from .binding import (  # noqa: F401
    B200Error, Context, F32, HostMM, I32, Q4_0, Q4_K, Q6_K, Q8_0, Q8_1, WEIGHT_TYPES, blck_size, build_native, check, lib, lib_path,
    row_size, type_size,
)
from .ops import MatmulBatch, MatmulBatchGather, stream_comm_init, gather_ll_init, mul_mat_batch, mm_ids, mul_mat, mul_mat_gather, mul_mat_gather_ll, mul_mat_id, mul_mat_multi, mul_mat_fused, mul_mat_add, graph_op, mul_mat_noalloc, quantize_row_q8_1, row_shard  # noqa: F401
