"""GPU parity at BASELINE.json's OWN sizes (-m gpu), against the oracle on sampled rows (the oracle cannot run the whole problem
in seconds, but a sampled row of the full problem is an exact sub-problem: output rows are independent):

  * configs[2]  Llama-3-8B prefill, n = 512, Q8_0 (and Q4_K): 14336x4096, 4096x14336 (the split-K path), 128256x4096, and the fused
                q/k/v and gate/up launches of b200_mul_mat_multi at Llama heights;
  * configs[3]  Mixtral-8x7B MUL_MAT_ID, Q4_K, top-2 of 8 experts, as[4096 -> 14336] and as[14336 -> 4096] at 1, 8 and 512 tokens.

Tolerance: NMSE <= 5e-4 (tests/test-backend-ops.cpp:3946-3948, 4135-4137), per matmul."""
import zlib

import numpy as np
import pytest

from oracle import refenv
from tests.util import NMSE_TOL, TYPES, nmse, oracle_mul_mat, oracle_mul_mat_id, random_blocks

pytestmark = pytest.mark.gpu


def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def tall(t, m, k, rng):
    """m rows of valid random blocks: a 2048-row base repeated with a different row rotation per chunk (generation stays cheap,
    every row of a tile is still different from its neighbours)"""
    base = random_blocks(t, min(m, 2048), k, rng)
    reps = (m + 2047) // 2048
    return np.concatenate([np.roll(base, 7 * i, axis=0) for i in range(reps)], 0)[:m]


@pytest.mark.parametrize("name", ["q8_0", "q4_K"])
@pytest.mark.parametrize("m,k", [(14336, 4096), (4096, 14336), (128256, 4096)])
def test_prefill_n512_llama_shapes_sampled_rows_vs_oracle(pkg, gpu_ctx, orc, name, m, k):
    import torch
    t = TYPES[name]
    n = 512
    rng = np.random.default_rng(zlib.crc32(repr((name, m, k)).encode()))
    W = tall(t, m, k, rng)
    X = rng.standard_normal((n, k)).astype(np.float32)
    Y = pkg.mul_mat(gpu_ctx, t, dev(W), dev(X))
    assert bool(torch.isfinite(Y).all())
    got = Y.cpu().numpy()
    rows = np.sort(rng.choice(m, 192, replace=False))
    want = oracle_mul_mat(orc, t, np.ascontiguousarray(W[rows]), X)
    e = nmse(want, got[:, rows])
    assert e <= NMSE_TOL, e
    # the tail tile and the first tile separately (a wrong tile would drown in the average)
    for sl in (slice(0, 16), slice(-16, None)):
        r2 = np.arange(m)[sl]
        assert nmse(oracle_mul_mat(orc, t, np.ascontiguousarray(W[r2]), X[:64]), got[:64][:, r2]) <= NMSE_TOL


@pytest.mark.parametrize("name", ["q8_0", "q4_K"])
@pytest.mark.parametrize("heights", [(4096, 1024, 1024), (14336, 14336)])
def test_prefill_n512_fused_multi_vs_oracle(pkg, gpu_ctx, orc, name, heights):
    """wq / wk / wv and ffn_gate / ffn_up of Llama-3-8B through ONE b200_mul_mat_multi launch at n = 512"""
    import torch
    t = TYPES[name]
    k, n = 4096, 512
    rng = np.random.default_rng(zlib.crc32(repr((name, heights)).encode()))
    Ws = [tall(t, m, k, rng) for m in heights]
    X = rng.standard_normal((n, k)).astype(np.float32)
    Xd = dev(X)
    outs = pkg.mul_mat_multi(gpu_ctx, t, [dev(W) for W in Ws], Xd)
    for W, Y in zip(Ws, outs):
        m = W.shape[0]
        rows = np.sort(rng.choice(m, 128, replace=False))
        want = oracle_mul_mat(orc, t, np.ascontiguousarray(W[rows]), X)
        assert nmse(want, Y.cpu().numpy()[:, rows]) <= NMSE_TOL
        # and the fused launch == the matrix on its own, bit for bit
        assert torch.equal(Y, pkg.mul_mat(gpu_ctx, t, dev(W), Xd))


@pytest.mark.parametrize("m,k", [(14336, 4096), (4096, 14336)])
@pytest.mark.parametrize("n_tokens", [1, 8, 512])
def test_mixtral_mul_mat_id_sampled_rows_vs_oracle(pkg, gpu_ctx, orc, m, k, n_tokens):
    """BASELINE configs[3]: 8 experts of Mixtral's FFN shapes, top-2 routing, b broadcast over the slots (ne11 = 1) — the sizes
    bench.py's `moe` section times.  The oracle computes 160 sampled rows of every expert."""
    t = refenv.Q4_K
    NE, NU = 8, 2
    rng = np.random.default_rng(zlib.crc32(repr((m, k, n_tokens)).encode()))
    base = random_blocks(t, 2048, k, rng)
    AS = np.stack([np.concatenate([np.roll(base, 11 * e + 5 * i, axis=0) for i in range((m + 2047) // 2048)], 0)[:m] for e in range(NE)])
    B = rng.standard_normal((n_tokens, 1, k)).astype(np.float32)
    ids = np.stack([rng.permutation(NE)[:NU] for _ in range(n_tokens)]).astype(np.int32)
    got = pkg.mul_mat_id(gpu_ctx, t, dev(AS), dev(B), dev(ids)).cpu().numpy()          # [n_tokens, NU, m]
    assert np.isfinite(got).all()
    rows = np.sort(rng.choice(m, 160, replace=False))
    want = oracle_mul_mat_id(orc, t, np.ascontiguousarray(AS[:, rows]), B, ids)       # [n_tokens, NU, 160]
    e = nmse(want, got[:, :, rows])
    assert e <= NMSE_TOL, e
    # every (token, slot) pair on its own: a mis-routed pair would not show in the average
    worst = max(nmse(want[tk, s], got[tk, s, rows]) for tk in range(min(n_tokens, 24)) for s in range(NU))
    assert worst <= NMSE_TOL, worst
