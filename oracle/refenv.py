"""ctypes loaders for the oracle.  TEST INFRASTRUCTURE ONLY (tests/, smoke(), bench.py cpu arm).

  load_oracle()  -> oracle/libggml_oracle.so   : our C restatement (ggml_oracle.c), always buildable
  load_ref()     -> oracle/_ref/libref_shim.so : the UNMODIFIED reference compiled from /root/reference,
                                                 or None when oracle/_ref was not built / did not travel

The reference's ggml-cpu comes in two builds (oracle/Makefile); `cpu_variant()` picks `native` only when
this host's /proc/cpuinfo has every ISA flag that build was compiled for, else `portable` (x86-64-v3).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from functools import lru_cache
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF_DIR = HERE / "_ref"

# enum ggml_type values (ggml/include/ggml.h:389-432)
F32, Q4_0, Q8_0, Q8_1, Q4_K, Q6_K, Q8_K, I32 = 0, 2, 8, 9, 12, 14, 15, 26
WEIGHT_TYPES = {"q4_0": Q4_0, "q8_0": Q8_0, "q4_K": Q4_K, "q6_K": Q6_K}
BLOCK = {Q4_0: (32, 18), Q8_0: (32, 34), Q8_1: (32, 36), Q4_K: (256, 144), Q6_K: (256, 210), Q8_K: (256, 292)}


def row_size(t: int, k: int) -> int:
    bs, sz = BLOCK[t]
    assert k % bs == 0
    return k // bs * sz


def build_oracle() -> None:
    """make the restatement (and, when /root/reference is present, oracle/_ref)."""
    subprocess.run(["make", "-s", "-j8", "-C", str(HERE), "all"], check=True)


@lru_cache(maxsize=None)
def cpu_variant() -> str:
    want = os.environ.get("ORACLE_CPU_VARIANT")
    if want:
        return want
    flags_file = REF_DIR / "native" / "cpu_flags.txt"
    try:
        need = set(flags_file.read_text().split())
        have: set[str] = set()
        for line in Path("/proc/cpuinfo").read_text().splitlines():
            if line.startswith("flags"):
                have = set(line.split(":", 1)[1].split())
                break
        if need and need <= have and (REF_DIR / "native" / "libggml-cpu.so").exists():
            return "native"
    except OSError:
        pass
    return "portable"


def ref_env(extra: dict | None = None) -> dict:
    """environment for running oracle/_ref binaries (test-backend-ops) as subprocesses."""
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = f"{REF_DIR / cpu_variant()}:{REF_DIR}:" + env.get("LD_LIBRARY_PATH", "")
    if extra:
        env.update(extra)
    return env


def _sig(fn, res, args):
    fn.restype, fn.argtypes = res, args
    return fn


@lru_cache(maxsize=None)
def load_oracle() -> C.CDLL:
    so = HERE / "libggml_oracle.so"
    if not so.exists():
        subprocess.run(["make", "-s", "-C", str(HERE), "restatement"], check=True)
    L = C.CDLL(str(so))
    vp, i64, i32, f32p = C.c_void_p, C.c_int64, C.c_int, C.c_void_p
    _sig(L.orc_blck_size, i32, [i32]); _sig(L.orc_type_size, C.c_size_t, [i32]); _sig(L.orc_row_size, C.c_size_t, [i32, i64])
    _sig(L.orc_fp16_to_fp32, C.c_float, [C.c_uint16]); _sig(L.orc_fp32_to_fp16, C.c_uint16, [C.c_float])
    for name in ("orc_quantize_row_q8_1", "orc_quantize_row_q8_0", "orc_quantize_row_q8_K", "orc_quantize_row_q4_0"):
        _sig(getattr(L, name), None, [f32p, vp, i64])
    _sig(L.orc_dequantize_row, i32, [i32, vp, f32p, i64])
    _sig(L.orc_vec_dot_type, i32, [i32])
    _sig(L.orc_vec_dot, C.c_float, [i32, i64, vp, vp])
    _sig(L.orc_mul_mat, i32, [i32, vp, f32p, f32p, i64, i64, i64, i64, i64, i64, i64, i32])
    _sig(L.orc_mul_mat_id, i32, [i32, vp, f32p, vp, f32p, i64, i64, i64, i64, i64, i64, i64, i32])
    _sig(L.orc_mul_mat_f64, i32, [i32, vp, f32p, vp, i64, i64, i64, i32])
    _sig(L.orc_nmse, C.c_double, [f32p, f32p, C.c_size_t])
    return L


@lru_cache(maxsize=None)
def load_ref():
    shim = REF_DIR / "libref_shim.so"
    if not shim.exists():
        return None
    mode = C.RTLD_GLOBAL
    C.CDLL(str(REF_DIR / "libggml-base.so"), mode=mode)
    C.CDLL(str(REF_DIR / cpu_variant() / "libggml-cpu.so"), mode=mode)  # satisfies soname libggml-cpu.so
    C.CDLL(str(REF_DIR / "libggml.so"), mode=mode)
    L = C.CDLL(str(shim), mode=mode)
    vp, i64, i32 = C.c_void_p, C.c_int64, C.c_int
    _sig(L.ref_blck_size, i32, [i32]); _sig(L.ref_type_size, C.c_size_t, [i32]); _sig(L.ref_row_size, C.c_size_t, [i32, i64])
    _sig(L.ref_fp16_to_fp32, C.c_float, [C.c_uint16]); _sig(L.ref_fp32_to_fp16, C.c_uint16, [C.c_float])
    _sig(L.ref_quantize_row_ref, i32, [i32, vp, vp, i64]); _sig(L.ref_quantize_row_cpu, i32, [i32, vp, vp, i64])
    _sig(L.ref_quantize_chunk, C.c_size_t, [i32, vp, vp, i64, i64])
    _sig(L.ref_dequantize_row, i32, [i32, vp, vp, i64]); _sig(L.ref_validate_row_data, i32, [i32, vp, C.c_size_t])
    _sig(L.ref_vec_dot_generic, C.c_float, [i32, i64, vp, vp])
    _sig(L.ref_mul_mat, i32, [i32, vp, vp, vp, i64, i64, i64, i64, i64, i64, i64, i32])
    _sig(L.ref_mul_mat_id, i32, [i32, vp, vp, vp, vp, i64, i64, i64, i64, i64, i64, i32])
    _sig(L.ref_time_mul_mat, C.c_double, [i32, vp, i64, i64, i64, i32, i32, i32])
    return L


def ptr(a) -> int:
    """address of a numpy array's data"""
    return a.ctypes.data
