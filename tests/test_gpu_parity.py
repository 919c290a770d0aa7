"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI, against the oracle on the same seeded inputs.

  * Q8_1 activation quantizer: BIT-EXACT against quantize_row_q8_1_ref (golden bytes + the C restatement on 1M blocks)
  * MUL_MAT / MUL_MAT_ID: NMSE <= 5e-4 against ggml-cpu semantics (tests/test-backend-ops.cpp:3946-3948, 4135-4137);
    in practice ~1e-6 (the GPU quantizes activations per 32 as Q8_1, ggml-cpu per 32 as Q8_0 or per 256 as Q8_K)
  * size-independent properties at BASELINE.json's full sizes (row-slice consistency, sampled rows vs oracle)
"""
import zlib

import numpy as np
import pytest

from oracle import refenv
from tests.util import (NMSE_TOL, TYPES, golden, nmse, oracle_mul_mat, oracle_mul_mat_f64, oracle_mul_mat_id, quantized_weights,
                        random_blocks)

pytestmark = pytest.mark.gpu


def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


# ---- (1) quantizer ------------------------------------------------------------------------------------------------------
def test_q8_1_golden_bit_exact(pkg, gpu_ctx):
    g = golden("q8_1")
    y = pkg.quantize_row_q8_1(gpu_ctx, dev(g["x"]))
    assert np.array_equal(y.cpu().numpy(), g["y"])


def test_q8_1_million_blocks_bit_exact(pkg, gpu_ctx, orc):
    rng = np.random.default_rng(1)
    rows, k = 2048, 16384                                  # 1,048,576 blocks
    x = rng.standard_normal((rows, k)).astype(np.float32)
    x *= rng.choice([1e-4, 1.0, 37.0, 1e5], (rows, 1)).astype(np.float32)
    x[5, :64] = 0.0
    x[6, 32:64] = np.round(x[6, 32:64] * 8) / 8            # plenty of exact ties after scaling
    want = np.zeros((rows, k // 32 * 36), np.uint8)
    for r in range(rows):
        orc.orc_quantize_row_q8_1(refenv.ptr(x[r]), refenv.ptr(want[r]), k)
    got = pkg.quantize_row_q8_1(gpu_ctx, dev(x)).cpu().numpy()
    assert np.array_equal(got, want)


def test_q8_1_strided_and_unaligned_rows(pkg, gpu_ctx, orc):
    import torch
    rng = np.random.default_rng(2)
    big = dev(rng.standard_normal((33, 96 + 5)).astype(np.float32))
    x = big[:, 5:5 + 96]                                   # row stride 101 floats, base offset 20 B: scalar-load variant
    want = np.zeros((33, 3 * 36), np.uint8)
    xh = np.ascontiguousarray(x.cpu().numpy())
    for r in range(33):
        orc.orc_quantize_row_q8_1(refenv.ptr(xh[r]), refenv.ptr(want[r]), 96)
    assert np.array_equal(pkg.quantize_row_q8_1(gpu_ctx, x).cpu().numpy(), want)
    with pytest.raises(ValueError):
        pkg.quantize_row_q8_1(gpu_ctx, torch.zeros((2, 48), device="cuda"))


# ---- (2) decode GEMV family -----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", list(TYPES))
@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 8])
@pytest.mark.parametrize("shape", [(96, 512), (257, 4096), (64, 14336)])
def test_gemv_vs_oracle(pkg, gpu_ctx, orc, ref, name, n, shape):
    t = TYPES[name]
    m, k = shape
    if k % refenv.BLOCK[t][0]:
        pytest.skip("k not a whole number of blocks")
    rng = np.random.default_rng(zlib.crc32(repr((name, n, shape)).encode()))
    W = quantized_weights(t, m, k, rng, ref if m * k <= 1 << 20 else None)
    X = rng.uniform(-1, 1, (n, k)).astype(np.float32)
    want = oracle_mul_mat(orc, t, W, X)
    got = pkg.mul_mat(gpu_ctx, t, dev(W), dev(X)).cpu().numpy()
    assert np.isfinite(got).all()
    e = nmse(want, got)
    assert e <= NMSE_TOL, e
    truth = oracle_mul_mat_f64(orc, t, W, X)
    assert nmse(truth, got) <= NMSE_TOL


@pytest.mark.parametrize("n", [1, 2, 4, 8])
def test_q4k_fast_kernel_matches_generic_kernel(pkg, gpu_ctx, n):
    """two independent CUDA kernels for the same op (fused-quantize fast path vs generic) agree to fp32 re-association"""
    rng = np.random.default_rng(5)
    m, k = 300, 2048
    W, X = dev(random_blocks(refenv.Q4_K, m, k, rng)), dev(rng.standard_normal((n, k)).astype(np.float32))
    fast = pkg.mul_mat(gpu_ctx, refenv.Q4_K, W, X).cpu().numpy()
    gpu_ctx.set_option("force_generic", 1)
    try:
        slow = pkg.mul_mat(gpu_ctx, refenv.Q4_K, W, X).cpu().numpy()
    finally:
        gpu_ctx.set_option("force_generic", 0)
    # n = 1: the fast path keeps the fp32 d of each activation block, the planes its fp16 rounding.  n >= 2 may run the
    # tensor-core GEMV, which quantizes activations per 256 values like the reference's Q8_K (not per 32 like Q8_1): the two
    # differ by activation-quantization noise (~1e-4), far below NMSE_TOL
    assert nmse(slow, fast) < (1e-6 if n == 1 else 3e-4)


@pytest.mark.parametrize("name", ["q4_K", "q6_K"])
@pytest.mark.parametrize("n", [1, 2, 3, 5, 8])
@pytest.mark.parametrize("shape", [(96, 512), (257, 4096), (64, 14336), (4099, 2048)])
def test_k_quant_tensor_core_gemv(pkg, gpu_ctx, orc, name, n, shape):
    """mmvq_mma.cu (forced): against the oracle and the f64 truth; the result is bit-identical for every K slicing, ring
    geometry and row sharding (fixed summation tree) — the property the row-split path relies on"""
    import torch
    m, k = shape
    t = TYPES[name]
    rng = np.random.default_rng(zlib.crc32(repr(("mma", name, n, shape)).encode()))
    W = random_blocks(t, m, k, rng)
    X = rng.standard_normal((n, k)).astype(np.float32)
    Wd, Xd = dev(W), dev(X)
    gpu_ctx.set_option("use_mma", 2)
    try:
        got = pkg.mul_mat(gpu_ctx, t, Wd, Xd)
        for ks in (1, 2, 4, 8, 16):
            gpu_ctx.set_option("mma_ks", ks)
            assert torch.equal(got, pkg.mul_mat(gpu_ctx, t, Wd, Xd)), ks
        gpu_ctx.set_option("mma_ks", 0)
        gpu_ctx.set_option("mma_ch", 1)
        assert torch.equal(got, pkg.mul_mat(gpu_ctx, t, Wd, Xd))
        gpu_ctx.set_option("mma_ch", 0)
        lo, hi = m // 3, m // 3 + max(1, m // 2)
        assert torch.equal(pkg.mul_mat(gpu_ctx, t, Wd[lo:hi].contiguous(), Xd), got[:, lo:hi])
    finally:
        gpu_ctx.set_option("use_mma", 1); gpu_ctx.set_option("mma_ks", 0); gpu_ctx.set_option("mma_ch", 0)
    got = got.cpu().numpy()
    assert np.isfinite(got).all()
    assert nmse(oracle_mul_mat(orc, t, W, X), got) <= NMSE_TOL
    assert nmse(oracle_mul_mat_f64(orc, t, W, X), got) <= NMSE_TOL


@pytest.mark.parametrize("name", ["q8_0", "q4_0"])
@pytest.mark.parametrize("n", [1, 3, 8])
@pytest.mark.parametrize("shape", [(96, 512), (257, 4096), (1030, 2048)])
def test_tensor_core_gemv_32_blocks(pkg, gpu_ctx, orc, name, n, shape):
    """mmvq_mma.cu for the 32-value block types (per-32 activation scales = the Q8_1 values): against the oracle, against the
    CUDA-core kernel that follows the same algebra (fp32 re-association only), and invariant under K slicing / row sharding"""
    import torch
    m, k = shape
    t = TYPES[name]
    rng = np.random.default_rng(zlib.crc32(repr(("mma32", name, n, shape)).encode()))
    W = random_blocks(t, m, k, rng)
    X = rng.standard_normal((n, k)).astype(np.float32)
    Wd, Xd = dev(W), dev(X)
    gpu_ctx.set_option("mma_min_n", 1)
    try:
        got = pkg.mul_mat(gpu_ctx, t, Wd, Xd)
        for ks in (1, 4, 16):
            gpu_ctx.set_option("mma_ks", ks)
            assert torch.equal(got, pkg.mul_mat(gpu_ctx, t, Wd, Xd)), ks
        gpu_ctx.set_option("mma_ks", 0)
        lo, hi = m // 3, m // 3 + max(1, m // 2)
        assert torch.equal(pkg.mul_mat(gpu_ctx, t, Wd[lo:hi].contiguous(), Xd), got[:, lo:hi])
        gpu_ctx.set_option("use_mma", 0)
        core = pkg.mul_mat(gpu_ctx, t, Wd, Xd).cpu().numpy()
    finally:
        gpu_ctx.set_option("use_mma", 1); gpu_ctx.set_option("mma_ks", 0); gpu_ctx.set_option("mma_min_n", 1)
    got = got.cpu().numpy()
    assert np.isfinite(got).all()
    assert nmse(core, got) < 1e-6           # same integers, fp32 d of the activation blocks vs their fp16 rounding in the planes
    assert nmse(oracle_mul_mat(orc, t, W, X), got) <= NMSE_TOL


def test_q4k_tensor_core_gemv_batches_and_zero_columns(pkg, gpu_ctx, orc):
    """batch dims with broadcast (several weight matrices behind one tensor map) and an all-zero activation column"""
    rng = np.random.default_rng(77)
    W = random_blocks(refenv.Q4_K, 2 * 160, 1024, rng).reshape(1, 2, 160, -1)
    X = rng.standard_normal((2, 4, 3, 1024)).astype(np.float32)
    X[1, 2, 1] = 0.0
    gpu_ctx.set_option("use_mma", 2)
    try:
        got = pkg.mul_mat(gpu_ctx, refenv.Q4_K, dev(W), dev(X)).cpu().numpy()
    finally:
        gpu_ctx.set_option("use_mma", 1)
    assert got.shape == (2, 4, 3, 160) and np.all(got[1, 2, 1] == 0.0)
    for i3 in range(2):
        for i2 in range(4):
            want = oracle_mul_mat(orc, refenv.Q4_K, np.ascontiguousarray(W[0, i2 // 2]), X[i3, i2])
            assert nmse(want, got[i3, i2]) <= NMSE_TOL


@pytest.mark.parametrize("name", list(TYPES))
def test_batch_broadcast_golden(pkg, gpu_ctx, name):
    t = TYPES[name]
    g = golden(name)
    got = pkg.mul_mat(gpu_ctx, t, dev(g["W"]), dev(g["X"])).cpu().numpy()
    assert nmse(g["Y"], got) <= NMSE_TOL
    got = pkg.mul_mat(gpu_ctx, t, dev(g["Wb"]), dev(g["Xb"])).cpu().numpy()      # W [k,m,2,1] x X [k,2,4,3]
    assert got.shape == g["Yb"].shape and nmse(g["Yb"], got) <= NMSE_TOL


@pytest.mark.parametrize("name", list(TYPES))
def test_noncontiguous_activation_rows(pkg, gpu_ctx, orc, name):
    """src1 as a view with a larger row stride (test-backend-ops k_v cases, tests/test-backend-ops.cpp:4018-4022)"""
    t = TYPES[name]
    rng = np.random.default_rng(9)
    m, k, n = 64, 512, 3
    W = random_blocks(t, m, k, rng)
    big = rng.uniform(-1, 1, (n, k + 256)).astype(np.float32)
    xv = dev(big)[:, :k]
    want = oracle_mul_mat(orc, t, W, np.ascontiguousarray(big[:, :k]))
    assert nmse(want, pkg.mul_mat(gpu_ctx, t, dev(W), xv).cpu().numpy()) <= NMSE_TOL


# ---- (3) prefill-sized n -------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", list(TYPES))
@pytest.mark.parametrize("n,m,k", [(9, 64, 512), (17, 130, 1024), (64, 256, 4096), (512, 384, 1024)])
def test_gemm_vs_oracle(pkg, gpu_ctx, orc, name, n, m, k):
    t = TYPES[name]
    rng = np.random.default_rng(n * 1000 + m)
    W = random_blocks(t, m, k, rng)
    X = rng.uniform(-1, 1, (n, k)).astype(np.float32)
    want = oracle_mul_mat(orc, t, W, X)
    got = pkg.mul_mat(gpu_ctx, t, dev(W), dev(X)).cpu().numpy()
    assert np.isfinite(got).all()
    assert nmse(want, got) <= NMSE_TOL


TC_TYPES = ["q4_K", "q4_0", "q8_0", "q6_K"]     # weight types the tcgen05 path covers (Q6_K: rows of a multiple of 2048 values, else CUDA cores)


@pytest.mark.parametrize("name", TC_TYPES)
@pytest.mark.parametrize("n,m,k", [(33, 200, 512), (130, 128, 256), (512, 1000, 2048), (77, 300, 4096)])
def test_tc_gemm_close_to_f64_truth_and_to_cuda_core_path(pkg, gpu_ctx, orc, name, n, m, k):
    """prefill path (tcgen05, fp16 operands / fp32 accumulate) vs (a) the exact product in f64 — fp16 rounding only, so far
    closer than the Q8_1-quantised reference gets — and (b) our CUDA-core path that follows ggml-cpu's Q8_1 arithmetic"""
    from tests.util import oracle_mul_mat_f64
    t = TYPES[name]
    if name == "q6_K" and k % 2048:
        pytest.skip("Q6_K rows of this length are not 16-byte aligned: CUDA-core path (covered by test_gemm_vs_oracle)")
    rng = np.random.default_rng(7 * n + m)
    W = random_blocks(t, m, k, rng)
    X = rng.standard_normal((n, k)).astype(np.float32)
    truth = oracle_mul_mat_f64(orc, t, W, X)
    Wd, Xd = dev(W), dev(X)
    gpu_ctx.set_option("use_tc", 1)
    got = pkg.mul_mat(gpu_ctx, t, Wd, Xd).cpu().numpy()
    gpu_ctx.set_option("use_tc", 0)
    core = pkg.mul_mat(gpu_ctx, t, Wd, Xd).cpu().numpy()
    gpu_ctx.set_option("use_tc", 1)
    assert nmse(truth, got) <= 1e-6
    assert nmse(core, got) <= 1e-4 and nmse(oracle_mul_mat(orc, t, W, X), got) <= NMSE_TOL


@pytest.mark.parametrize("name", TC_TYPES)
def test_tc_gemm_llama_shape_linearity_and_tiles(pkg, gpu_ctx, name):
    """full prefill size (m=4096, k=4096, n=512): column j of the batched result must equal the same column computed in a
    small batch (tiles / clusters / persistent scheduling cannot leak between columns), and W(2x) = 2 W(x) exactly
    (power-of-two scaling commutes with every rounding on the path)"""
    import torch
    t = TYPES[name]
    rng = np.random.default_rng(11)
    W = dev(np.tile(random_blocks(t, 256, 4096, rng), (16, 1)))
    X = torch.randn((512, 4096), device="cuda")
    Y = pkg.mul_mat(gpu_ctx, t, W, X)
    sel = [0, 1, 127, 128, 255, 256, 300, 511, 17, 64, 200, 444, 3, 5, 7, 9]
    Ys = pkg.mul_mat(gpu_ctx, t, W, X[sel].contiguous())
    assert torch.equal(Y[sel], Ys)
    assert torch.equal(pkg.mul_mat(gpu_ctx, t, W, 2.0 * X), 2.0 * Y)
    assert torch.equal(Y[:, :256], Y[:, 256:512]) and torch.equal(Y[:, :256], Y[:, 3840:])      # W repeats every 256 rows


@pytest.mark.parametrize("name", TC_TYPES)
@pytest.mark.parametrize("m,k,n", [(1000, 4096, 77), (4096, 4096, 512), (2048, 8192, 200)])
def test_tc_gemm_tilings_give_the_same_bits(pkg, gpu_ctx, orc, name, m, k, n):
    """the sum over K of an output element is formed in the same order whatever the tiling: every tile width (64 / 128 / 192 columns, the
    last tile of a row narrower), and CTA pairs on 256-row tiles (cta_group::2, option tc_pair = 2), must give the bits of the default
    plan; split K re-associates the two halves (fp32 rounding only).  Sampled rows against the oracle."""
    import torch
    t = TYPES[name]
    if t == TYPES["q6_K"] and k % 2048:
        pytest.skip("Q6_K takes the tensor-core path only for k % 2048 == 0")
    rng = np.random.default_rng(zlib.crc32(repr((name, m, k, n)).encode()))
    Wn = np.tile(random_blocks(t, 250, k, rng), ((m + 249) // 250, 1))[:m]
    W = dev(Wn)
    Xn = rng.standard_normal((n, k)).astype(np.float32)
    X = dev(Xn)
    gpu_ctx.set_option("tc_split_k", 0)
    try:
        base = pkg.mul_mat(gpu_ctx, t, W, X)
        rows = np.sort(rng.choice(m, 48, replace=False))
        assert nmse(oracle_mul_mat(orc, t, np.ascontiguousarray(Wn[rows]), Xn), base.cpu().numpy()[:, rows]) <= NMSE_TOL
        for nt in (64, 128, 192):
            for pair in (0, 2):
                gpu_ctx.set_option("tc_nt", nt); gpu_ctx.set_option("tc_pair", pair)
                assert torch.equal(pkg.mul_mat(gpu_ctx, t, W, X), base), (nt, pair)
        gpu_ctx.set_option("tc_nt", 0); gpu_ctx.set_option("tc_pair", 0); gpu_ctx.set_option("tc_split_k", 2)
        assert nmse(base.cpu().numpy(), pkg.mul_mat(gpu_ctx, t, W, X).cpu().numpy()) <= 1e-10
    finally:
        gpu_ctx.set_option("tc_nt", 0); gpu_ctx.set_option("tc_pair", 1); gpu_ctx.set_option("tc_split_k", 1)


# ---- (4) MUL_MAT_ID --------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", list(TYPES))
def test_mul_mat_id_golden_strided_ids(pkg, gpu_ctx, name):
    t = TYPES[name]
    g = golden("mmid")
    ids_view = dev(g[f"{name}_ids_full"])[:, :2]           # a view: row stride 4 ints, 2 used
    for b_ne1 in (2, 1):
        got = pkg.mul_mat_id(gpu_ctx, t, dev(g[f"{name}_AS"]), dev(g[f"{name}_B{b_ne1}"]), ids_view).cpu().numpy()
        assert nmse(g[f"{name}_C{b_ne1}"], got) <= NMSE_TOL


@pytest.mark.parametrize("name", list(TYPES))
@pytest.mark.parametrize("n_tokens,n_expert,n_used,b_ne1", [(1, 8, 2, 1), (4, 8, 2, 2), (5, 4, 4, 1), (17, 8, 2, 2), (129, 8, 4, 1), (32, 4, 1, 1)])
def test_mul_mat_id_vs_oracle(pkg, gpu_ctx, orc, name, n_tokens, n_expert, n_used, b_ne1):
    t = TYPES[name]
    rng = np.random.default_rng(n_tokens * 31 + n_used)
    m, k = 128, 512
    AS = random_blocks(t, n_expert * m, k, rng).reshape(n_expert, m, -1)
    B = rng.uniform(-1, 1, (n_tokens, b_ne1, k)).astype(np.float32)
    ids = np.stack([rng.permutation(n_expert)[:n_used] for _ in range(n_tokens)]).astype(np.int32)
    want = oracle_mul_mat_id(orc, t, AS, B, ids)
    got = pkg.mul_mat_id(gpu_ctx, t, dev(AS), dev(B), dev(ids)).cpu().numpy()
    assert np.isfinite(got).all()
    assert nmse(want, got) <= NMSE_TOL


def test_mm_ids_gather(pkg, gpu_ctx):
    rng = np.random.default_rng(3)
    for n_tokens, n_expert, n_used in [(1, 8, 2), (7, 4, 4), (300, 8, 2), (1000, 128, 8)]:
        ids = np.stack([rng.permutation(n_expert)[:n_used] for _ in range(n_tokens)]).astype(np.int32)
        src1, bounds = pkg.mm_ids(gpu_ctx, dev(ids), n_expert)
        src1, bounds = src1.cpu().numpy(), bounds.cpu().numpy()
        flat = ids.ravel()
        order = np.argsort(flat, kind="stable")            # stable = pair-index order inside an expert
        assert np.array_equal(src1, order.astype(np.int32))
        assert np.array_equal(bounds, np.concatenate([[0], np.cumsum(np.bincount(flat, minlength=n_expert))]).astype(np.int32))


# ---- full sizes: properties ----------------------------------------------------------------------------------------------
def test_config1_q4_0_4096x4096(pkg, gpu_ctx, orc, ref):
    """BASELINE.json configs[0]: MUL_MAT Q4_0 x F32, m=4096 n=1 k=4096"""
    rng = np.random.default_rng(4096)
    W = quantized_weights(refenv.Q4_0, 4096, 4096, rng, ref)
    X = rng.uniform(-1, 1, (1, 4096)).astype(np.float32)
    want = oracle_mul_mat(orc, refenv.Q4_0, W, X)
    got = pkg.mul_mat(gpu_ctx, refenv.Q4_0, dev(W), dev(X)).cpu().numpy()
    assert nmse(want, got) <= NMSE_TOL
    if ref is not None:                                      # and against the real reference backend, live
        Y = np.zeros((1, 4096), np.float32)
        assert ref.ref_mul_mat(refenv.Q4_0, refenv.ptr(W), refenv.ptr(X), refenv.ptr(Y), 4096, 1, 4096, 1, 1, 1, 1, 0) == 0
        assert nmse(Y, got) <= NMSE_TOL


@pytest.mark.parametrize("name,m,k", [("q4_K", 128256, 4096), ("q6_K", 128256, 4096), ("q4_K", 4096, 14336), ("q8_0", 14336, 4096)])
def test_llama3_shapes_sampled_rows_and_slices(pkg, gpu_ctx, orc, name, m, k):
    """full Llama-3-8B matrices: (a) 512 sampled rows against the oracle, (b) the result of a row slice computed on its own
    is bit-identical to the slice of the full result (the property the row-split path relies on), (c) n=1 equals column 0
    of n=4 computed with the same kernel family to fp32 re-association"""
    import torch
    t = TYPES[name]
    rng = np.random.default_rng(m + k)
    base = random_blocks(t, 2048, k, rng)
    reps = (m + 2047) // 2048
    W = np.concatenate([np.roll(base, i, axis=0) for i in range(reps)], 0)[:m]     # distinct row order per 2048-chunk
    X = rng.standard_normal((4, k)).astype(np.float32)
    Wd, Xd = dev(W), dev(X)
    y4 = pkg.mul_mat(gpu_ctx, t, Wd, Xd)
    y1 = pkg.mul_mat(gpu_ctx, t, Wd, Xd[:1])
    rows = np.sort(rng.choice(m, 512, replace=False))
    want = oracle_mul_mat(orc, t, np.ascontiguousarray(W[rows]), X)
    assert nmse(want, y4.cpu().numpy()[:, rows]) <= NMSE_TOL
    assert nmse(want[:1], y1.cpu().numpy()[:, rows]) <= NMSE_TOL
    lo, hi = 16000 if m > 20000 else 1024, 16000 + 4096 if m > 20000 else 2048
    ys = pkg.mul_mat(gpu_ctx, t, Wd[lo:hi], Xd[:1])
    assert torch.equal(ys, y1[:, lo:hi])


# ---- host-buffer entry (the e2e path) and back-to-back launches ---------------------------------------------------------------
def test_mul_mat_host_and_back_to_back_dependencies(pkg, gpu_ctx, orc):
    """(a) b200_mul_mat_host: H2D x, GEMV, D2H y per call with a different x every call; (b) a chain y1 = W1 x, y2 = W2 y1 launched
    back to back: the second GEMV's input IS the first one's output, so the deferred dependency wait must not be taken"""
    import ctypes as C
    import torch
    L = pkg.lib()
    rng = np.random.default_rng(77)
    m, k = 2048, 2048
    W1, W2 = random_blocks(refenv.Q4_K, m, k, rng), random_blocks(refenv.Q4_K, m, k, rng, scale=2.0)
    W1d, W2d = dev(W1), dev(W2)
    hctx = pkg.Context(0)
    xh = torch.empty((1, k)).pin_memory()
    yh = torch.empty((1, m)).pin_memory()
    for i in range(6):
        x = rng.standard_normal((1, k)).astype(np.float32)
        xh.copy_(torch.from_numpy(x))
        rc = L.b200_mul_mat_host(hctx.handle, refenv.Q4_K, W1d.data_ptr(), m, k, xh.data_ptr(), 1, yh.data_ptr())
        assert rc == 0, L.b200_last_error()
        assert nmse(oracle_mul_mat(orc, refenv.Q4_K, W1, x), yh.numpy()) <= NMSE_TOL
    hctx.close()
    for i in range(4):
        x = rng.standard_normal((1, k)).astype(np.float32)
        xd = dev(x)
        y1 = pkg.mul_mat(gpu_ctx, refenv.Q4_K, W1d, xd)
        y2 = pkg.mul_mat(gpu_ctx, refenv.Q4_K, W2d, y1)          # depends on the kernel right before it
        y3 = pkg.mul_mat(gpu_ctx, refenv.Q4_K, W1d, xd)          # independent: may overlap y2's tail
        w1 = oracle_mul_mat(orc, refenv.Q4_K, W1, x)
        assert nmse(w1, y1.cpu().numpy()) <= NMSE_TOL
        assert nmse(oracle_mul_mat(orc, refenv.Q4_K, W2, y1.cpu().numpy()), y2.cpu().numpy()) <= NMSE_TOL
        assert torch.equal(y1, y3)


def test_mul_mat_host_batch_matches_single_calls(pkg, gpu_ctx):
    """b200_mul_mat_host_batch: contiguous and scattered host vectors, mixed types and n; identical to b200_mul_mat_host"""
    import ctypes as C
    import torch
    L = pkg.lib()
    rng = np.random.default_rng(3)
    specs = [("q4_K", 512, 1024, 1), ("q4_K", 256, 1024, 1), ("q8_0", 128, 512, 2), ("q6_K", 300, 256, 1), ("q4_0", 64, 2048, 9)]
    xh = torch.randn((sum(k * n for _, _, k, n in specs) + 64,)).pin_memory()
    yh = torch.zeros((sum(m * n for _, m, _, n in specs) + 64,)).pin_memory()
    want = torch.zeros_like(yh)
    jobs = (pkg.HostMM * len(specs))()
    Ws, xo, yo = [], 0, 0
    hctx = pkg.Context(0)
    for j, (name, m, k, n) in enumerate(specs):
        if j == 3:
            xo += 16; yo += 16                               # break the contiguity once
        W = dev(random_blocks(TYPES[name], m, k, rng)); Ws.append(W)
        jobs[j].type, jobs[j].w_dev, jobs[j].m, jobs[j].k, jobs[j].n = TYPES[name], W.data_ptr(), m, k, n
        jobs[j].x_host, jobs[j].y_host = xh.data_ptr() + 4 * xo, yh.data_ptr() + 4 * yo
        pkg.check(L.b200_mul_mat_host(hctx.handle, TYPES[name], W.data_ptr(), m, k, xh.data_ptr() + 4 * xo, n, want.data_ptr() + 4 * yo), "single")
        xo += k * n; yo += m * n
    pkg.check(L.b200_mul_mat_host_batch(hctx.handle, jobs, len(specs)), "batch")
    assert torch.equal(yh, want) and float(yh.abs().max()) > 0


@pytest.mark.parametrize("name,ms,k,n", [
    ("q4_K", (512, 128, 128), 1024, 1),        # fused GEMV, q/k/v shaped
    ("q4_K", (1792, 1792), 512, 1),            # fused GEMV, gate/up shaped
    ("q4_K", (300, 77), 256, 1),               # ragged heights
    ("q4_K", (512, 128, 128), 1024, 40),       # tcgen05, three matrices
    ("q4_K", (300, 77), 512, 33),              # tcgen05, ragged heights (row tiles padded per matrix)
    ("q8_0", (256, 384), 512, 200),            # tcgen05, Q8_0
    ("q4_0", (128, 640, 64), 256, 129),        # tcgen05, Q4_0
    ("q6_K", (128, 64), 256, 1),               # not fusable: runs one by one
    ("q4_K", (128, 64), 256, 4),               # not fusable (n in 2..8): runs one by one
    ("q4_K", (64, 64, 64, 64), 256, 1),        # more than three matrices: runs one by one
])
def test_mul_mat_multi_equals_separate_calls(pkg, gpu_ctx, name, ms, k, n):
    """b200_mul_mat_multi: several matrices x the same activations == the same b200_mul_mat calls, bit for bit"""
    import torch
    rng = np.random.default_rng(11)
    t = TYPES[name]
    Ws = [dev(random_blocks(t, m, k, rng)) for m in ms]
    x = dev(rng.standard_normal((n, k)).astype(np.float32))
    want = [pkg.mul_mat(gpu_ctx, t, W, x) for W in Ws]
    outs = [torch.full((n, m), float("nan"), dtype=torch.float32, device="cuda") for m in ms]
    pkg.mul_mat_multi(gpu_ctx, t, Ws, x, outs)
    gpu_ctx.synchronize()
    for w, o in zip(want, outs):
        assert torch.equal(w, o)


def test_mul_mat_multi_chain_orders_dependent_launches(pkg, gpu_ctx):
    """fused launches in a dependency chain (x -> [a, b] -> a feeds the next pair ...) give the same as single calls"""
    import torch
    rng = np.random.default_rng(12)
    k = 512
    Ws = [dev(random_blocks(TYPES["q4_K"], k, k, rng)) for _ in range(6)]
    x0 = dev(rng.standard_normal((1, k)).astype(np.float32))

    def run(fused):
        x, outs = x0, []
        for i in range(0, 6, 2):
            if fused:
                a, b = pkg.mul_mat_multi(gpu_ctx, TYPES["q4_K"], Ws[i:i + 2], x)
            else:
                a, b = pkg.mul_mat(gpu_ctx, TYPES["q4_K"], Ws[i], x), pkg.mul_mat(gpu_ctx, TYPES["q4_K"], Ws[i + 1], x)
            outs += [a, b]
            x = a * 0.05
        gpu_ctx.synchronize()
        return outs

    for w, o in zip(run(False), run(True)):
        assert torch.isfinite(w).all() and torch.equal(w, o)


def test_gather_ll_world1_equals_mul_mat(pkg):
    """the LL exchange with a world of one rank (the GEMV sends {value, epoch} to its own slot, the receiver unpacks): equal to
    b200_mul_mat, over several launches (the two slots alternate) and 1..3 matrices per launch"""
    import torch
    rng = np.random.default_rng(21)
    ctx = pkg.Context(0, stream="torch")
    k, cap = 1024, 4096
    buf = torch.zeros((2 * cap, 2), dtype=torch.int32, device="cuda")
    pkg.gather_ll_init(ctx, 0, 1, [buf.data_ptr()], cap)
    x = dev(rng.standard_normal((1, k)).astype(np.float32))
    for ms in [(512,), (512, 128, 128), (1792, 1792), (300, 77), (4096,), (640, 64, 1)]:
        Ws = [dev(random_blocks(TYPES["q4_K"], m, k, rng)) for m in ms]
        want = [pkg.mul_mat(ctx, TYPES["q4_K"], W, x) for W in Ws]
        outs = [torch.full((1, m), float("nan"), dtype=torch.float32, device="cuda") for m in ms]
        pkg.mul_mat_gather_ll(ctx, TYPES["q4_K"], Ws, x, outs, [0] * len(ms))
        ctx.synchronize()
        for w, o in zip(want, outs):
            assert torch.equal(w, o)
        x = (outs[0][:, :1] * 0 + x)              # next round's activations depend on this round's result
    # too many results for a slot, overlapping output: refused
    with pytest.raises(pkg.B200Error):
        pkg.mul_mat_gather_ll(ctx, TYPES["q4_K"], [dev(random_blocks(TYPES["q4_K"], cap + 256, k, rng))], x,
                              [torch.empty((1, cap + 256), device="cuda")], [0])
    ctx.close()


def test_mul_mat_host_batch_shares_uploads_and_launches(pkg, gpu_ctx):
    """jobs that follow one another with the SAME host activations (q/k/v, gate/up) are uploaded once and fused into one launch;
    results identical to b200_mul_mat_host per job"""
    import torch
    L = pkg.lib()
    rng = np.random.default_rng(4)
    k = 1024
    # (type, m, which x) — x index repeats mark shared activations; a type change or a fourth job must break the group
    specs = [("q4_K", 512, 0), ("q4_K", 128, 0), ("q4_K", 128, 0), ("q4_K", 512, 1), ("q4_K", 1408, 2), ("q4_K", 1408, 2), ("q8_0", 256, 2),
             ("q4_K", 64, 3), ("q4_K", 64, 3), ("q4_K", 64, 3), ("q4_K", 64, 3), ("q6_K", 96, 4), ("q6_K", 96, 4)]
    xh = torch.randn((5, k)).pin_memory()
    yh = torch.zeros((sum(m for _, m, _ in specs),)).pin_memory()
    want = torch.zeros_like(yh)
    jobs = (pkg.HostMM * len(specs))()
    hctx = pkg.Context(0)
    Ws, yo = [], 0
    for j, (name, m, xi) in enumerate(specs):
        W = dev(random_blocks(TYPES[name], m, k, rng)); Ws.append(W)
        jobs[j].type, jobs[j].w_dev, jobs[j].m, jobs[j].k, jobs[j].n = TYPES[name], W.data_ptr(), m, k, 1
        jobs[j].x_host, jobs[j].y_host = xh[xi].data_ptr(), yh.data_ptr() + 4 * yo
        pkg.check(L.b200_mul_mat_host(hctx.handle, TYPES[name], W.data_ptr(), m, k, xh[xi].data_ptr(), 1, want.data_ptr() + 4 * yo), "single")
        yo += m
    n0 = hctx.launches
    pkg.check(L.b200_mul_mat_host_batch(hctx.handle, jobs, len(specs)), "batch")
    assert torch.equal(yh, want) and float(yh.abs().max()) > 0
    # Q4_K groups fuse: {0,1,2} {3} {4,5} {6: q8_0} {7,8,9} {10} -> the q8_0 / q6_K jobs go one by one (2 launches each: quantize + GEMV)
    assert hctx.launches - n0 < 2 * len(specs) - 4
    hctx.close()


@pytest.mark.parametrize("m_full,k,n,world", [(1024, 8192, 512, 8), (1024, 1024, 40, 2), (768, 2048, 300, 3)])
def test_mul_mat_gather_prefill_peer_stores_on_one_gpu(pkg, gpu_ctx, m_full, k, n, world):
    """row split at prefill size, all `world` ranks emulated on one GPU: every rank's tcgen05 epilogue stores its tiles into all
    `world` full-height outputs (b200_mul_mat_gather, n > 8) — each of them must equal the unsplit product (to fp32 re-association:
    the unsplit GEMM may split long K)"""
    import torch
    rng = np.random.default_rng(m_full + k + n)
    t = TYPES["q4_K"]
    W = dev(random_blocks(t, m_full, k, rng))
    X = dev(rng.standard_normal((n, k)).astype(np.float32))
    want = pkg.mul_mat(gpu_ctx, t, W, X)
    fulls = [torch.full((n, m_full), float("nan"), device="cuda") for _ in range(world)]
    rows = m_full // world
    for r in range(world):
        shard = W[r * rows:(r + 1) * rows].contiguous()
        ysh = torch.empty((n, rows), device="cuda")
        pkg.mul_mat_gather(gpu_ctx, t, shard, X, ysh, [f.data_ptr() for f in fulls], m_full, r * rows)
        gpu_ctx.synchronize()
        assert nmse(want[:, r * rows:(r + 1) * rows].cpu().numpy(), ysh.cpu().numpy()) < 1e-10
    for f in fulls:
        assert torch.isfinite(f).all()
        assert nmse(want.cpu().numpy(), f.cpu().numpy()) < 1e-10
