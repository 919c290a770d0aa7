"""the weight types beyond Q4_0 / Q8_0 / Q4_K / Q6_K — Q4_1, Q5_0, Q5_1, Q5_K, Q2_K, Q3_K, IQ4_NL, IQ4_XS, MXFP4, NVFP4 —
through the C ABI against the UNMODIFIED reference (oracle/_ref: ggml_quantize_chunk makes the blocks, ggml-cpu's MUL_MAT and
dequantize_row_* are the expected values).  The C restatement in oracle/ does not cover these types, so these tests need oracle/_ref."""
import ctypes as C
import zlib
from importlib import import_module

import numpy as np
import pytest

from oracle import refenv
from tests.util import NMSE_TOL, nmse

pytestmark = pytest.mark.gpu

# name -> (enum ggml_type, block elements, block bytes)   (ggml/include/ggml.h:389-432, ggml-common.h)
MORE = {
    "q4_1": (3, 32, 20), "q5_0": (6, 32, 22), "q5_1": (7, 32, 24), "q5_K": (13, 256, 176), "q2_K": (10, 256, 84), "q3_K": (11, 256, 110),
    "iq4_nl": (20, 32, 18), "iq4_xs": (23, 256, 136), "mxfp4": (39, 32, 17), "nvfp4": (40, 64, 36),
}


def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def quantize(ref, name, m, k, rng, scale=0.05):
    t, bs, sz = MORE[name]
    assert ref.ref_blck_size(t) == bs and ref.ref_type_size(t) == sz
    wf = (rng.standard_normal((m, k)) * scale).astype(np.float32)
    W = np.zeros((m, k // bs * sz), np.uint8)
    ref.ref_quantize_chunk(t, refenv.ptr(wf), refenv.ptr(W), m, k)
    return W


def ref_mul_mat(ref, t, W, X):
    m = W.shape[0]
    n, k = X.shape
    Y = np.zeros((n, m), np.float32)
    assert ref.ref_mul_mat(t, refenv.ptr(W), refenv.ptr(np.ascontiguousarray(X)), refenv.ptr(Y), m, n, k, 1, 1, 1, 1, 8) == 0
    return Y


@pytest.mark.parametrize("n", [1, 5, 8, 19])
@pytest.mark.parametrize("name", sorted(MORE))
def test_mul_mat_more_types_against_the_reference(pkg, gpu_ctx, ref, name, n):
    """n <= 8: the generic GEMV family; n = 19: three column tiles of it (Q5_K: the tcgen05 GEMM)"""
    if ref is None:
        pytest.skip("oracle/_ref did not travel")
    import torch
    t = MORE[name][0]
    m, k = 300, {256: 2304, 64: 2368, 32: 2336}[MORE[name][1]]      # (2336 / 32 = 73 blocks: MXFP4 rows of an odd number of bytes)
    rng = np.random.default_rng(zlib.crc32(repr((name, n)).encode()))
    W = quantize(ref, name, m, k, rng)
    X = rng.standard_normal((n, k)).astype(np.float32)
    want = ref_mul_mat(ref, t, W, X)
    got = pkg.mul_mat(gpu_ctx, t, dev(W), dev(X))
    torch.cuda.synchronize()
    assert nmse(want, got.cpu().numpy()) <= NMSE_TOL


@pytest.mark.parametrize("n", [64, 512])
def test_q5k_prefill_gemm_on_the_tensor_cores(pkg, gpu_ctx, ref, n):
    """Q5_K at n > 8 runs on the tcgen05 GEMM (weights converted to fp16 in tensor memory): against ggml-cpu, and bit-identical for
    single CTAs and CTA pairs"""
    if ref is None:
        pytest.skip("oracle/_ref did not travel")
    import torch
    t = MORE["q5_K"][0]
    m, k = 1000, 4096
    rng = np.random.default_rng(n)
    W = quantize(ref, "q5_K", m, k, rng)
    X = rng.standard_normal((n, k)).astype(np.float32)
    want = ref_mul_mat(ref, t, W, X)
    Wd, Xd = dev(W), dev(X)
    l0 = gpu_ctx.launches
    got = pkg.mul_mat(gpu_ctx, t, Wd, Xd)
    torch.cuda.synchronize()
    assert nmse(want, got.cpu().numpy()) <= NMSE_TOL
    assert gpu_ctx.launches - l0 <= 3                            # activation conversion + one GEMM (not n / 8 GEMV tiles)
    try:
        gpu_ctx.set_option("tc_pair", 2)
        pair = pkg.mul_mat(gpu_ctx, t, Wd, Xd)
        gpu_ctx.set_option("tc_pair", 0)
        single = pkg.mul_mat(gpu_ctx, t, Wd, Xd)
    finally:
        gpu_ctx.set_option("tc_pair", 1)
    assert torch.equal(pair, got) and torch.equal(single, got)


@pytest.mark.parametrize("n", [17, 130])
@pytest.mark.parametrize("name", sorted(set(MORE) - {"q5_K"}))
def test_prefill_of_types_without_a_converter(pkg, gpu_ctx, ref, name, n):
    """n >= 17: the matrix is dequantized into fp16 scratch and multiplied by the tcgen05 GEMM (three launches) instead of ceil(n / 8)
    column tiles of the generic GEMV; both against ggml-cpu"""
    if ref is None:
        pytest.skip("oracle/_ref did not travel")
    import torch
    t = MORE[name][0]
    m, k = 700, 2048
    rng = np.random.default_rng(zlib.crc32(repr((name, n, "tc")).encode()))
    W = quantize(ref, name, m, k, rng)
    X = rng.standard_normal((n, k)).astype(np.float32)
    want = ref_mul_mat(ref, t, W, X)
    Wd, Xd = dev(W), dev(X)
    l0 = gpu_ctx.launches
    got = pkg.mul_mat(gpu_ctx, t, Wd, Xd)
    torch.cuda.synchronize()
    assert gpu_ctx.launches - l0 == 3                            # dequantize, activations -> fp16, GEMM
    assert nmse(want, got.cpu().numpy()) <= NMSE_TOL
    try:
        gpu_ctx.set_option("tc_dequant", 0)
        tiles = pkg.mul_mat(gpu_ctx, t, Wd, Xd)
    finally:
        gpu_ctx.set_option("tc_dequant", 1)
    assert nmse(want, tiles.cpu().numpy()) <= NMSE_TOL


@pytest.mark.parametrize("name", sorted(MORE))
def test_get_rows_more_types(pkg, gpu_ctx, ref, name):
    if ref is None:
        pytest.skip("oracle/_ref did not travel")
    import torch
    t = MORE[name][0]
    k = 1024
    rng = np.random.default_rng(zlib.crc32(name.encode()))
    W = quantize(ref, name, 64, k, rng)
    rows = [3, 0, 63, 3, 17]
    want = np.zeros((len(rows), k), np.float32)
    for j, r in enumerate(rows):
        assert ref.ref_dequantize_row(t, refenv.ptr(np.ascontiguousarray(W[r])), refenv.ptr(want[j]), k) == 0
    ops = import_module(pkg.__name__ + ".ops")
    Wd = dev(W).reshape(1, 1, 64, -1)                  # (kept alive: the descriptor holds a raw pointer)
    idx = torch.tensor(rows, dtype=torch.int32, device="cuda").reshape(1, 1, 1, len(rows))
    out = torch.empty((1, 1, len(rows), k), device="cuda")
    ops.check(ops.lib().b200_op_get_rows(gpu_ctx.handle, C.byref(ops._desc(Wd, t, k_elems=k)), C.byref(ops._fdesc(idx)), C.byref(ops._fdesc(out))),
              "b200_op_get_rows")
    torch.cuda.synchronize()
    # (the reference's own build may contract d * q - m into an FMA: values agree to the last bit or two, not necessarily exactly)
    assert nmse(want, out[0, 0].cpu().numpy()) <= 1e-12


@pytest.mark.parametrize("name", ["mxfp4", "nvfp4", "q3_K", "iq4_xs"])
def test_mul_mat_id_more_types(pkg, gpu_ctx, ref, name):
    """MUL_MAT_ID (the MoE matmul: gpt-oss ships MXFP4 experts) for the added types, decode- and prompt-sized"""
    if ref is None:
        pytest.skip("oracle/_ref did not travel")
    import torch
    t = MORE[name][0]
    n_expert, n_used, m, k = 8, 2, 96, 512
    for n_tokens in (1, 37, 80):                                 # (80: 160 pairs, the experts' matrices go through fp16 scratch to the tcgen05 GEMM)
        rng = np.random.default_rng(zlib.crc32(repr((name, n_tokens)).encode()))
        AS = np.stack([quantize(ref, name, m, k, rng) for _ in range(n_expert)])
        B = rng.standard_normal((n_tokens, 1, k)).astype(np.float32)
        ids = np.stack([rng.permutation(n_expert)[:n_used] for _ in range(n_tokens)]).astype(np.int32)
        want = np.zeros((n_tokens, n_used, m), np.float32)
        assert ref.ref_mul_mat_id(t, refenv.ptr(AS), refenv.ptr(B), refenv.ptr(ids), refenv.ptr(want), m, k, n_expert, n_used, n_tokens, 1, 8) == 0
        got = pkg.mul_mat_id(gpu_ctx, t, dev(AS), dev(B), dev(ids))
        torch.cuda.synchronize()
        assert nmse(want, got.cpu().numpy()) <= NMSE_TOL
