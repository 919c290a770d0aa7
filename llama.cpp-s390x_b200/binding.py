"""ctypes binding of include/ggml_b200.h (libggml_b200_kernels.so).  Fails loudly when the library is missing."""
from __future__ import annotations

import ctypes as C
import subprocess
from functools import lru_cache
from pathlib import Path

HERE = Path(__file__).resolve().parent

# enum ggml_type values (reference: ggml/include/ggml.h:389-432)
F32, Q4_0, Q8_0, Q8_1, Q4_K, Q6_K, I32 = 0, 2, 8, 9, 12, 14, 26
WEIGHT_TYPES = {"q4_0": Q4_0, "q8_0": Q8_0, "q4_K": Q4_K, "q6_K": Q6_K}

EXPORTS = [
    "b200_abi_version", "b200_last_error", "b200_device_count", "b200_device_info", "b200_blck_size", "b200_type_size",
    "b200_row_size", "b200_ctx_create", "b200_ctx_destroy", "b200_ctx_stream", "b200_ctx_synchronize",
    "b200_ctx_launch_count", "b200_ctx_set_option", "b200_quantize_row_q8_1", "b200_mul_mat", "b200_mul_mat_supported",
    "b200_mul_mat_noalloc",
    "b200_mul_mat_id", "b200_mul_mat_id_supported", "b200_mm_ids", "b200_weights_upload", "b200_weights_free",
    "b200_mul_mat_host", "b200_mul_mat_host_batch", "b200_mul_mat_gather", "b200_gather_sync_init", "b200_mul_mat_multi", "b200_mul_mat_batch", "b200_stream_comm_init", "b200_mul_mat_batch_gather", "b200_gather_ll_init", "b200_mul_mat_gather_ll", "b200_malloc", "b200_free", "b200_memcpy_h2d", "b200_memcpy_d2h", "b200_memcpy_d2d",
    "b200_memset", "b200_memcpy_async", "b200_host_malloc", "b200_host_free", "b200_event_create", "b200_event_destroy", "b200_event_create_timed", "b200_event_elapsed_ms",
    "b200_event_record", "b200_event_wait", "b200_event_synchronize", "b200_rowsplit_create", "b200_rowsplit_destroy", "b200_rowsplit_n_dev",
    "b200_rowsplit_rows", "b200_rowsplit_upload", "b200_rowsplit_weights_free", "b200_rowsplit_mul_mat", "b200_rowsplit_mul_mat_main",
    "b200_rowsplit_synchronize", "b200_rowsplit_ctx",
    "b200_comm_create", "b200_comm_destroy", "b200_comm_size", "b200_comm_allreduce", "b200_capture_begin", "b200_capture_end", "b200_graph_launch", "b200_graph_free", "b200_ctx_alloc_generation",
    "b200_op_binary", "b200_op_repeat", "b200_op_add_id", "b200_op_glu", "b200_op_unary", "b200_op_unary_supported", "b200_op_scale", "b200_op_rms_norm", "b200_op_rms_norm_mul", "b200_op_rope",
    "b200_op_soft_max", "b200_op_cpy", "b200_op_get_rows", "b200_op_set_rows", "b200_mul_mat_f", "b200_mul_mat_f_supported", "b200_mul_mat_add", "b200_mul_mat_add_supported", "b200_mul_mat_fused",
    "b200_mul_mat_fused_supported",
]


class B200Error(RuntimeError):
    pass


class HostMM(C.Structure):
    """struct b200_host_mm"""
    _fields_ = [("type", C.c_int32), ("_pad", C.c_int32), ("w_dev", C.c_void_p), ("m", C.c_int64), ("k", C.c_int64), ("n", C.c_int64),
                ("x_host", C.c_void_p), ("y_host", C.c_void_p)]


class Tensor(C.Structure):
    """struct b200_tensor"""
    _fields_ = [("data", C.c_void_p), ("type", C.c_int32), ("_pad", C.c_int32), ("ne", C.c_int64 * 4), ("nb", C.c_int64 * 4)]


def lib_path() -> Path:
    import os
    alt = os.environ.get("GGML_B200_KERNELS_SO")          # experiments: a differently configured build of the same library
    return Path(alt) if alt else HERE / "libggml_b200_kernels.so"


def build_native() -> None:
    """compile the kernels library (and, when /root/reference is present, the ggml backend shell) in-tree"""
    subprocess.run(["make", "-s", "-j8", "-C", str(HERE), "all"], check=True)


@lru_cache(maxsize=None)
def lib() -> C.CDLL:
    p = lib_path()
    if not p.exists():
        raise B200Error(f"{p} is missing: run __graft_entry__.build() (there is no fallback path)")
    L = C.CDLL(str(p))
    vp, i64, i32, sz = C.c_void_p, C.c_int64, C.c_int, C.c_size_t
    T = C.POINTER(Tensor)

    def sig(name, res, args):
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args

    sig("b200_abi_version", i32, [])
    sig("b200_last_error", C.c_char_p, [])
    sig("b200_device_count", i32, [])
    sig("b200_device_info", i32, [i32, C.c_char_p, sz, C.POINTER(sz), C.POINTER(sz), C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)])
    sig("b200_blck_size", i32, [i32]); sig("b200_type_size", sz, [i32]); sig("b200_row_size", sz, [i32, i64])
    sig("b200_ctx_create", i32, [i32, vp, C.POINTER(vp)])
    sig("b200_ctx_destroy", None, [vp]); sig("b200_ctx_stream", vp, [vp]); sig("b200_ctx_synchronize", i32, [vp])
    sig("b200_ctx_launch_count", i64, [vp]); sig("b200_ctx_set_option", i32, [vp, C.c_char_p, i32])
    sig("b200_quantize_row_q8_1", i32, [vp, vp, i64, vp, i64, i64])
    sig("b200_mul_mat", i32, [vp, T, T, T]); sig("b200_mul_mat_supported", i32, [T, T, T])
    sig("b200_mul_mat_noalloc", i32, [vp, T, T, T])
    f32 = C.c_float
    sig("b200_comm_create", i32, [C.POINTER(vp), i32, C.POINTER(vp)]); sig("b200_comm_destroy", None, [vp]); sig("b200_comm_size", i32, [vp])
    sig("b200_comm_allreduce", i32, [vp, C.POINTER(vp), C.POINTER(i32), i64])
    sig("b200_capture_begin", i32, [vp]); sig("b200_capture_end", i32, [vp, C.POINTER(vp)]); sig("b200_graph_launch", i32, [vp, vp])
    sig("b200_graph_free", None, [vp]); sig("b200_ctx_alloc_generation", C.c_uint64, [vp])
    sig("b200_op_binary", i32, [vp, i32, T, T, T]); sig("b200_op_repeat", i32, [vp, T, T]); sig("b200_op_add_id", i32, [vp, T, T, T, T])
    sig("b200_op_glu", i32, [vp, i32, T, T, T, i32, f32, f32]); sig("b200_op_unary", i32, [vp, i32, T, T]); sig("b200_op_unary_supported", i32, [i32])
    sig("b200_op_scale", i32, [vp, T, T, f32, f32]); sig("b200_op_rms_norm", i32, [vp, T, T, f32]); sig("b200_op_rms_norm_mul", i32, [vp, T, T, T, f32])
    sig("b200_op_rope", i32, [vp, T, T, T, T, C.POINTER(i32)]); sig("b200_op_soft_max", i32, [vp, T, T, T, T, f32, f32])
    sig("b200_op_cpy", i32, [vp, T, T]); sig("b200_op_get_rows", i32, [vp, T, T, T]); sig("b200_op_set_rows", i32, [vp, T, T, T])
    sig("b200_mul_mat_f", i32, [vp, T, T, T]); sig("b200_mul_mat_f_supported", i32, [T, T, T])
    sig("b200_mul_mat_add", i32, [vp, T, T, T, T]); sig("b200_mul_mat_add_supported", i32, [T, T, T, T])
    sig("b200_mul_mat_fused", i32, [vp, T, T, T, T, T, i32, T]); sig("b200_mul_mat_fused_supported", i32, [T, T, T, T, T, i32, T])
    sig("b200_mul_mat_id", i32, [vp, T, T, T, T]); sig("b200_mul_mat_id_supported", i32, [T, T, T, T])
    sig("b200_mm_ids", i32, [vp, vp, i64, i32, i32, i64, vp, vp])
    sig("b200_weights_upload", i32, [vp, i32, vp, i64, i64, C.POINTER(vp)]); sig("b200_weights_free", None, [vp, vp])
    sig("b200_mul_mat_host", i32, [vp, i32, vp, i64, i64, vp, i64, vp])
    sig("b200_mul_mat_host_batch", i32, [vp, C.POINTER(HostMM), i32])
    sig("b200_mul_mat_gather", i32, [vp, T, T, T, i32, C.POINTER(vp), i64, i64])
    sig("b200_gather_sync_init", i32, [vp, i32, i32, C.POINTER(vp)])
    sig("b200_mul_mat_multi", i32, [vp, i32, C.POINTER(T), T, C.POINTER(T)])
    sig("b200_mul_mat_batch", i32, [vp, i32, C.POINTER(T), C.POINTER(T), C.POINTER(T)])
    sig("b200_stream_comm_init", i32, [vp, i32, i32, C.POINTER(vp), i64])
    sig("b200_mul_mat_batch_gather", i32, [vp, i32, C.POINTER(T), C.POINTER(T), C.POINTER(T), C.POINTER(i64)])
    sig("b200_gather_ll_init", i32, [vp, i32, i32, C.POINTER(vp), i64])
    sig("b200_mul_mat_gather_ll", i32, [vp, i32, C.POINTER(T), T, C.POINTER(T), C.POINTER(i64)])
    sig("b200_malloc", i32, [i32, sz, C.POINTER(vp)]); sig("b200_free", i32, [i32, vp])
    sig("b200_memcpy_h2d", i32, [vp, vp, vp, sz]); sig("b200_memcpy_d2h", i32, [vp, vp, vp, sz]); sig("b200_memcpy_d2d", i32, [vp, vp, vp, sz])
    sig("b200_memset", i32, [vp, vp, i32, sz]); sig("b200_memcpy_async", i32, [vp, vp, vp, sz, i32])
    sig("b200_host_malloc", i32, [sz, C.POINTER(vp)]); sig("b200_host_free", i32, [vp])
    sig("b200_event_create", i32, [i32, C.POINTER(vp)]); sig("b200_event_destroy", i32, [i32, vp])
    sig("b200_event_create_timed", i32, [i32, C.POINTER(vp)]); sig("b200_event_elapsed_ms", i32, [vp, vp, C.POINTER(C.c_float)])
    sig("b200_event_record", i32, [vp, vp]); sig("b200_event_wait", i32, [vp, vp]); sig("b200_event_synchronize", i32, [vp])
    sig("b200_rowsplit_create", i32, [C.POINTER(i32), i32, C.POINTER(C.c_float), C.POINTER(vp)])
    sig("b200_rowsplit_destroy", None, [vp]); sig("b200_rowsplit_n_dev", i32, [vp])
    sig("b200_rowsplit_rows", i32, [vp, i64, i32, C.POINTER(i64), C.POINTER(i64)])
    sig("b200_rowsplit_upload", i32, [vp, i32, vp, i64, i64, C.POINTER(vp)]); sig("b200_rowsplit_weights_free", None, [vp, vp])
    sig("b200_rowsplit_mul_mat", i32, [vp, vp, C.POINTER(vp), i64, C.POINTER(vp), i32])
    sig("b200_rowsplit_mul_mat_main", i32, [vp, vp, i32, i64, i64, C.POINTER(vp), C.POINTER(i64), C.POINTER(i64), T, T])
    sig("b200_rowsplit_synchronize", i32, [vp]); sig("b200_rowsplit_ctx", vp, [vp, i32])
    return L


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise B200Error(f"{what} failed ({rc}): {lib().b200_last_error().decode(errors='replace')}")


def blck_size(t: int) -> int:
    return lib().b200_blck_size(t)


def type_size(t: int) -> int:
    return lib().b200_type_size(t)


def row_size(t: int, k: int) -> int:
    return lib().b200_row_size(t, k)


class Context:
    """b200_ctx: one device + one stream.  stream=None -> owns a stream; stream="torch" -> torch's current stream."""

    def __init__(self, device: int = 0, stream=None):
        if stream == "torch":
            import torch
            # torch's default stream is the legacy NULL stream (handle 0); hand it over as cudaStreamLegacy (0x1),
            # because a NULL handle asks b200_ctx_create for a private stream
            stream = torch.cuda.current_stream(device).cuda_stream or 0x1
        self.device = device
        h = C.c_void_p()
        check(lib().b200_ctx_create(device, C.c_void_p(stream) if stream else None, C.byref(h)), "b200_ctx_create")
        self.handle = h

    def synchronize(self) -> None:
        check(lib().b200_ctx_synchronize(self.handle), "b200_ctx_synchronize")

    @property
    def stream(self) -> int:
        return lib().b200_ctx_stream(self.handle) or 0

    @property
    def launches(self) -> int:
        return lib().b200_ctx_launch_count(self.handle)

    def set_option(self, name: str, value: int) -> int:
        rc = lib().b200_ctx_set_option(self.handle, name.encode(), value)
        if rc < 0:
            raise B200Error(f"unknown option {name}")
        return rc

    def close(self) -> None:
        if self.handle:
            lib().b200_ctx_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
