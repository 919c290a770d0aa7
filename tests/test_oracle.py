"""CPU tests: pin the oracle (oracle/ggml_oracle.c, our C restatement) against
  (a) the committed golden vectors generated from the unmodified reference (tests/golden/make_golden.py), and
  (b) the reference itself, live, when oracle/_ref is present (this container; it also travels to the GPU box).
Bit-exact for bytes and for floats computed in the same summation order."""
import numpy as np
import pytest

from oracle import refenv
from tests.util import TYPES, golden, nmse, oracle_mul_mat, oracle_mul_mat_f64, oracle_mul_mat_id, random_blocks


REASSOC_NMSE = 1e-11


def test_geometry(orc):
    for t, (bs, sz) in refenv.BLOCK.items():
        assert orc.orc_blck_size(t) == bs and orc.orc_type_size(t) == sz
        assert orc.orc_row_size(t, 4096) == 4096 // bs * sz


def test_fp16_roundtrip_all_halves(orc):
    hs = np.arange(0, 1 << 16, dtype=np.uint16)
    for h in hs[::7]:
        f = orc.orc_fp16_to_fp32(int(h))
        expect = np.array([h], np.uint16).view(np.float16)[0]
        if np.isnan(expect):
            assert np.isnan(f)
            continue
        assert f == float(expect)
        assert orc.orc_fp32_to_fp16(f) == int(h)


def test_q8_1_golden(orc):
    g = golden("q8_1")
    x, y = g["x"], g["y"]
    out = np.zeros_like(y)
    for r in range(x.shape[0]):
        orc.orc_quantize_row_q8_1(refenv.ptr(x[r]), refenv.ptr(out[r]), x.shape[1])
    assert np.array_equal(out, y)


@pytest.mark.parametrize("name", list(TYPES))
def test_dequant_vecdot_mulmat_golden(orc, name):
    t = TYPES[name]
    g = golden(name)
    W, deq, X, Y = g["W"], g["deq"], g["X"], g["Y"]
    m, k = deq.shape
    out = np.zeros((k,), np.float32)
    for r in range(m):
        assert orc.orc_dequantize_row(t, refenv.ptr(W[r]), refenv.ptr(out), k) == 0
        assert np.array_equal(out, deq[r])
    # activation format of the CPU path and the generic vec_dot
    at = orc.orc_vec_dot_type(t)
    act = np.zeros_like(g["act"])
    {refenv.Q8_0: orc.orc_quantize_row_q8_0, refenv.Q8_K: orc.orc_quantize_row_q8_K}[at](refenv.ptr(X[0]), refenv.ptr(act), k)
    assert np.array_equal(act, g["act"])
    vd = np.array([orc.orc_vec_dot(t, k, refenv.ptr(W[r]), refenv.ptr(act)) for r in range(m)], np.float32)
    if name == "q4_K":
        # gcc contracts `sumf -= dmin * sumi` (ggml-cpu/quants.c:713) into an FMA when the reference is built with
        # -march>=haswell; the restatement is compiled -ffp-contract=off.  20/48 values differ by 1 ulp.
        assert np.abs(vd - g["vec_dot"]).max() <= 2.4e-7 and nmse(g["vec_dot"], vd) < 1e-12
    else:
        assert np.array_equal(vd, g["vec_dot"])
    # whole op, plain and batched / broadcast.  The reference CPU backend dispatches its SIMD vec_dot (AVX2 here) even
    # with use_ref, so whole-op outputs equal the generic-order restatement only up to fp32 re-association:
    # measured NMSE 1e-14..7e-14 (max |diff| 4e-7); the bound below is 100x that, 8 orders under the op tolerance.
    assert nmse(Y, oracle_mul_mat(orc, t, W, X)) < REASSOC_NMSE
    assert nmse(g["Yb"], oracle_mul_mat(orc, t, g["Wb"], g["Xb"])) < REASSOC_NMSE
    # and the op is close to the exact-dequant f64 truth
    truth = oracle_mul_mat_f64(orc, t, W, X)
    assert nmse(truth, Y) < 5e-4


@pytest.mark.parametrize("name", list(TYPES))
def test_mul_mat_id_golden(orc, name):
    t = TYPES[name]
    g = golden("mmid")
    AS, full = g[f"{name}_AS"], g[f"{name}_ids_full"]
    ids = full[:, :2]
    for b_ne1 in (2, 1):
        out = oracle_mul_mat_id(orc, t, AS, g[f"{name}_B{b_ne1}"], ids)
        assert nmse(g[f"{name}_C{b_ne1}"], out) < REASSOC_NMSE


def test_q8_1_live_reference(orc, ref):
    if ref is None:
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(7)
    x = (rng.standard_normal((512, 1024)) * rng.choice([1e-3, 1.0, 50.0], (512, 1))).astype(np.float32)
    a = np.zeros((512, 1024 // 32 * 36), np.uint8)
    b = np.zeros_like(a)
    for r in range(x.shape[0]):
        orc.orc_quantize_row_q8_1(refenv.ptr(x[r]), refenv.ptr(a[r]), 1024)
        ref.ref_quantize_row_ref(refenv.Q8_1, refenv.ptr(x[r]), refenv.ptr(b[r]), 1024)
    assert np.array_equal(a, b)


@pytest.mark.parametrize("name", list(TYPES))
def test_mul_mat_live_reference_random_blocks(orc, ref, name):
    """random valid blocks (the generator the GPU tests use) through the reference CPU backend == the restatement"""
    if ref is None:
        pytest.skip("oracle/_ref not built")
    t = TYPES[name]
    rng = np.random.default_rng(11)
    m, n, k = 40, 3, 768
    W = random_blocks(t, m, k, rng)
    assert ref.ref_validate_row_data(t, refenv.ptr(W), W.nbytes) == 1
    X = rng.standard_normal((n, k)).astype(np.float32)
    Y = np.zeros((n, m), np.float32)
    assert ref.ref_mul_mat(t, refenv.ptr(W), refenv.ptr(X), refenv.ptr(Y), m, n, k, 1, 1, 1, 1, 4) == 0
    assert nmse(Y, oracle_mul_mat(orc, t, W, X)) < REASSOC_NMSE
