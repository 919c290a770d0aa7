"""CPU tests of the boundary: the C-ABI library loads, exports every symbol include/ggml_b200.h declares, answers the
geometry queries, validates arguments without a device, and refuses to compute without a GPU (no fallback)."""
import ctypes as C
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def declared_symbols():
    text = (ROOT / "include" / "ggml_b200.h").read_text()
    return sorted(set(re.findall(r"B200_API[^;(]*?\b(b200_\w+)\s*\(", text)))


def test_header_symbols_exported(pkg):
    L = C.CDLL(str(pkg.lib_path()))
    syms = declared_symbols()
    assert len(syms) >= 35
    missing = [s for s in syms if not hasattr(L, s)]
    assert not missing, missing
    # and the Python binding knows every one of them
    from llama_cpp_s390x_b200 import binding
    assert sorted(binding.EXPORTS) == syms


def test_geometry(pkg):
    assert pkg.lib().b200_abi_version() == 1
    assert (pkg.blck_size(pkg.Q4_0), pkg.type_size(pkg.Q4_0)) == (32, 18)
    assert (pkg.blck_size(pkg.Q8_0), pkg.type_size(pkg.Q8_0)) == (32, 34)
    assert (pkg.blck_size(pkg.Q8_1), pkg.type_size(pkg.Q8_1)) == (32, 36)
    assert (pkg.blck_size(pkg.Q4_K), pkg.type_size(pkg.Q4_K)) == (256, 144)
    assert (pkg.blck_size(pkg.Q6_K), pkg.type_size(pkg.Q6_K)) == (256, 210)
    assert pkg.row_size(pkg.Q4_K, 4096) == 2304 and pkg.row_size(pkg.Q6_K, 4096) == 3360


def _t(pkg, type_, ne, nb, data=0x1000):
    from llama_cpp_s390x_b200.binding import Tensor
    t = Tensor()
    t.data, t.type = data, type_
    for i in range(4):
        t.ne[i], t.nb[i] = ne[i], nb[i]
    return t


def test_supports_op_validation_needs_no_device(pkg):
    L = pkg.lib()
    k, m, n = 4096, 128, 3
    rs = pkg.row_size(pkg.Q4_K, k)
    a = _t(pkg, pkg.Q4_K, (k, m, 1, 1), (144, rs, rs * m, rs * m))
    b = _t(pkg, pkg.F32, (k, n, 1, 1), (4, 4 * k, 4 * k * n, 4 * k * n))
    d = _t(pkg, pkg.F32, (m, n, 1, 1), (4, 4 * m, 4 * m * n, 4 * m * n))
    assert L.b200_mul_mat_supported(C.byref(a), C.byref(b), C.byref(d)) == 1
    bad_k = _t(pkg, pkg.F32, (k + 32, n, 1, 1), (4, 4 * k, 4 * k * n, 4 * k * n))
    assert L.b200_mul_mat_supported(C.byref(a), C.byref(bad_k), C.byref(d)) == 0
    f16 = _t(pkg, 1, (k, n, 1, 1), (2, 2 * k, 2 * k * n, 2 * k * n))          # F16 activations are not on this path
    assert L.b200_mul_mat_supported(C.byref(a), C.byref(f16), C.byref(d)) == 0
    q5 = _t(pkg, 13, (k, m, 1, 1), (176, 176 * 16, 176 * 16 * m, 176 * 16 * m))        # Q5_K weights: the generic GEMV family (round 2, N4)
    assert L.b200_mul_mat_supported(C.byref(q5), C.byref(b), C.byref(d)) == 1
    q2 = _t(pkg, 10, (k, m, 1, 1), (84, 84 * 16, 84 * 16 * m, 84 * 16 * m))    # Q2_K weights: generic GEMV family (N4)
    assert L.b200_mul_mat_supported(C.byref(q2), C.byref(b), C.byref(d)) == 1
    iq2 = _t(pkg, 16, (k, m, 1, 1), (66, 66 * 16, 66 * 16 * m, 66 * 16 * m))   # IQ2_XXS (grid code books) is not on this path
    assert L.b200_mul_mat_supported(C.byref(iq2), C.byref(b), C.byref(d)) == 0
    perm = _t(pkg, pkg.F32, (m, n, 1, 1), (4 * n, 4, 4 * m * n, 4 * m * n))    # transposed dst
    assert L.b200_mul_mat_supported(C.byref(a), C.byref(b), C.byref(perm)) == 0
    ids = _t(pkg, pkg.I32, (2, 5, 1, 1), (4, 32, 160, 160))                    # strided view rows
    as_ = _t(pkg, pkg.Q4_K, (k, m, 8, 1), (144, rs, rs * m, rs * m * 8))
    bb = _t(pkg, pkg.F32, (k, 1, 5, 1), (4, 4 * k, 4 * k, 4 * k * 5))
    cc = _t(pkg, pkg.F32, (m, 2, 5, 1), (4, 4 * m, 8 * m, 40 * m))
    assert L.b200_mul_mat_id_supported(C.byref(as_), C.byref(bb), C.byref(ids), C.byref(cc)) == 1


def test_no_cpu_fallback(pkg):
    """without a B200 the product path must fail loudly, not compute on the host"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    assert pkg.lib().b200_device_count() == 0
    with pytest.raises(pkg.B200Error):
        pkg.Context(0)


def test_backend_plugin_exports():
    so = ROOT / "llama.cpp-s390x_b200" / "libggml-b200.so"
    if not so.exists():
        pytest.skip("backend shell not built (needs the reference headers)")
    import subprocess
    out = subprocess.run(["nm", "-D", "--defined-only", str(so)], capture_output=True, text=True, check=True).stdout
    for sym in ("ggml_backend_init", "ggml_backend_score", "ggml_backend_b200_reg"):
        assert re.search(rf"\bT {sym}\b", out), sym


# enum ggml_type -> (block elements, block bytes) of every quantized weight type the path takes (ggml-common.h)
WEIGHT_FORMATS = {2: (32, 18), 3: (32, 20), 6: (32, 22), 7: (32, 24), 8: (32, 34), 10: (256, 84), 11: (256, 110), 12: (256, 144), 13: (256, 176),
                  14: (256, 210), 20: (32, 18), 23: (256, 136), 39: (32, 17), 40: (64, 36)}


@pytest.mark.parametrize("t", sorted(WEIGHT_FORMATS))
def test_weight_type_geometry_and_validation(pkg, t):
    """block geometry as the reference defines it, and b200_mul_mat_supported's stride rules per type (no device needed): whole blocks
    per row, 2-byte aligned rows — except MXFP4, whose 17-byte blocks make rows of any byte alignment legal"""
    L = pkg.lib()
    bs, sz = WEIGHT_FORMATS[t]
    assert (pkg.blck_size(t), pkg.type_size(t)) == (bs, sz)
    k, m, n = 3 * 256, 5, 2                                   # 3 super-blocks / 12 or 24 small blocks per row
    rs = pkg.row_size(t, k)
    assert rs == k // bs * sz
    a = _t(pkg, t, (k, m, 1, 1), (sz, rs, rs * m, rs * m))
    b = _t(pkg, pkg.F32, (k, n, 1, 1), (4, 4 * k, 4 * k * n, 4 * k * n))
    d = _t(pkg, pkg.F32, (m, n, 1, 1), (4, 4 * m, 4 * m * n, 4 * m * n))
    assert L.b200_mul_mat_supported(C.byref(a), C.byref(b), C.byref(d)) == 1
    short = _t(pkg, t, (k - bs // 2, m, 1, 1), (sz, rs, rs * m, rs * m))            # not a whole number of blocks
    assert L.b200_mul_mat_supported(C.byref(short), C.byref(b), C.byref(d)) == 0
    odd = _t(pkg, t, (k, m, 1, 1), (sz, rs + 1, (rs + 1) * m, (rs + 1) * m))       # rows one byte further apart
    assert L.b200_mul_mat_supported(C.byref(odd), C.byref(b), C.byref(d)) == (1 if t == 39 or (rs + 1) % 2 == 0 else 0)
