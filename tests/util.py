"""Shared helpers for the tests: synthetic weight blocks, NMSE, golden loading.  Test infrastructure only."""
from pathlib import Path

import numpy as np

from oracle import refenv

GOLDEN = Path(__file__).resolve().parent / "golden"
TYPES = refenv.WEIGHT_TYPES          # name -> enum ggml_type
NMSE_TOL = 5e-4                      # tests/test-backend-ops.cpp:3946-3948 (MUL_MAT), :4135-4137 (MUL_MAT_ID)


def golden(name):
    return np.load(GOLDEN / f"{name}.npz")


def nmse(a, b):
    """test-backend-ops.cpp:278-291: sum((a-b)^2) / sum(a^2), a = reference"""
    a = np.asarray(a, np.float64).ravel()
    b = np.asarray(b, np.float64).ravel()
    return float(((a - b) ** 2).sum() / max((a ** 2).sum(), 1e-300))


def random_blocks(t, m, k, rng, scale=0.02):
    """m rows of VALID random blocks of weight type t (no quantizer needed; any byte pattern with sane fp16 scales is a
    legal tensor).  Returns uint8 [m, row_size]."""
    bs, sz = refenv.BLOCK[t]
    nb = k // bs
    raw = rng.integers(0, 256, (m, nb, sz), dtype=np.uint8)

    def put_f16(col, vals):
        raw[:, :, col:col + 2] = vals.astype(np.float16).view(np.uint8).reshape(m, nb, 2)

    if t == refenv.Q4_0:
        put_f16(0, rng.uniform(0.25, 1.0, (m, nb)) * scale / 8 * rng.choice([-1.0, 1.0], (m, nb)))
    elif t == refenv.Q8_0:
        put_f16(0, rng.uniform(0.25, 1.0, (m, nb)) * scale / 127)
    elif t == refenv.Q4_K:
        put_f16(0, rng.uniform(0.25, 1.0, (m, nb)) * scale / (15 * 63))
        put_f16(2, rng.uniform(0.25, 1.0, (m, nb)) * scale / 63)
    elif t == refenv.Q6_K:
        put_f16(208, rng.uniform(0.25, 1.0, (m, nb)) * scale / (32 * 127))
    else:
        raise ValueError(t)
    return np.ascontiguousarray(raw.reshape(m, nb * sz))


def quantized_weights(t, m, k, rng, ref=None, scale=0.05):
    """gaussian weights quantized by the reference's ggml_quantize_chunk when oracle/_ref is present, else random blocks"""
    if ref is not None:
        wf = (rng.standard_normal((m, k)) * scale).astype(np.float32)
        W = np.zeros((m, refenv.row_size(t, k)), np.uint8)
        ref.ref_quantize_chunk(t, refenv.ptr(wf), refenv.ptr(W), m, k)
        return W
    return random_blocks(t, m, k, rng, scale)


def oracle_mul_mat(orc, t, W, X, threads=8):
    """ggml-cpu semantics through the C restatement.  W uint8 [(ne03, ne02,) m, rs], X f32 [(ne13, ne12,) n, k]"""
    W4 = W.reshape((1,) * (4 - W.ndim) + W.shape)
    X4 = np.ascontiguousarray(X.reshape((1,) * (4 - X.ndim) + X.shape))
    ne03, ne02, m, _ = W4.shape
    ne13, ne12, n, k = X4.shape
    Y = np.zeros((ne13, ne12, n, m), np.float32)
    rc = orc.orc_mul_mat(t, refenv.ptr(np.ascontiguousarray(W4)), refenv.ptr(X4), refenv.ptr(Y), m, n, k, ne02, ne03, ne12, ne13, threads)
    assert rc == 0, rc
    return Y.reshape(X.shape[:-1] + (m,))


def oracle_mul_mat_f64(orc, t, W, X, threads=8):
    m = W.shape[0]
    n, k = X.shape
    Y = np.zeros((n, m), np.float64)
    rc = orc.orc_mul_mat_f64(t, refenv.ptr(np.ascontiguousarray(W)), refenv.ptr(np.ascontiguousarray(X)), refenv.ptr(Y), m, n, k, threads)
    assert rc == 0, rc
    return Y


def oracle_mul_mat_id(orc, t, AS, B, ids, threads=8):
    """AS uint8 [n_expert, m, rs]; B f32 [n_tokens, b_ne1, k]; ids int32 [n_tokens, n_used] (contiguous)"""
    ne, m, _ = AS.shape
    nt, b_ne1, k = B.shape
    nu = ids.shape[1]
    ids = np.ascontiguousarray(ids, np.int32)
    Cc = np.zeros((nt, nu, m), np.float32)
    rc = orc.orc_mul_mat_id(t, refenv.ptr(np.ascontiguousarray(AS)), refenv.ptr(np.ascontiguousarray(B)), refenv.ptr(ids), refenv.ptr(Cc),
                            m, k, ne, nu, nt, b_ne1, nu, threads)
    assert rc == 0, rc
    return Cc
