"""Row split across GPUs (BASELINE configs[4]): single-process b200_rowsplit_* through the C ABI.
Needs >= 2 GPUs (`gpurun --gpus 2`); skipped otherwise.  Output rows are independent, so every exchange mode must give
BIT-IDENTICAL results to the single-GPU kernel on the unsplit matrix."""
import ctypes as C

import numpy as np
import pytest

from tests.util import TYPES, random_blocks

pytestmark = pytest.mark.gpu


def _n_gpus():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.fixture(scope="module")
def rs(pkg):
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    n = 2 if n < 4 else 4 if n < 8 else 8
    L = pkg.lib()
    devs = (C.c_int * n)(*range(n))
    h = C.c_void_p()
    pkg.check(L.b200_rowsplit_create(devs, n, None, C.byref(h)), "b200_rowsplit_create")
    yield L, h, n
    L.b200_rowsplit_destroy(h)


def _run(pkg, L, h, n_dev, t, W, X, mode):
    import torch
    m, k, n = W.shape[0], X.shape[1], X.shape[0]
    w = C.c_void_p()
    pkg.check(L.b200_rowsplit_upload(h, t, W.ctypes.data_as(C.c_void_p), k, m, C.byref(w)), "b200_rowsplit_upload")
    xs = [torch.from_numpy(X).to(f"cuda:{i}") for i in range(n_dev)]
    ys = [torch.full((n, m), float("nan"), device=f"cuda:{i}") for i in range(n_dev)]
    for i in range(n_dev):
        torch.cuda.synchronize(i)
    xp = (C.c_void_p * n_dev)(*[x.data_ptr() for x in xs])
    yp = (C.c_void_p * n_dev)(*[y.data_ptr() for y in ys])
    pkg.check(L.b200_rowsplit_mul_mat(h, w, xp, n, yp, mode), "b200_rowsplit_mul_mat")
    pkg.check(L.b200_rowsplit_synchronize(h), "b200_rowsplit_synchronize")
    out = [y.cpu().numpy() for y in ys]
    L.b200_rowsplit_weights_free(h, w)
    return out


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("name,m,k,n", [("q4_K", 8192, 8192, 1), ("q4_K", 8192, 4096, 4), ("q8_0", 1000, 1024, 2), ("q6_K", 4096, 2048, 1),
                                        ("q4_K", 4096, 4096, 64)])
def test_rowsplit_matches_single_gpu_bit_for_bit(pkg, gpu_ctx, rs, name, m, k, n, mode):
    import torch
    L, h, n_dev = rs
    t = TYPES[name]
    rng = np.random.default_rng(m + k + n)
    W = random_blocks(t, m, k, rng)
    X = rng.uniform(-1, 1, (n, k)).astype(np.float32)
    want = pkg.mul_mat(gpu_ctx, t, torch.from_numpy(W).cuda(), torch.from_numpy(X).cuda()).cpu().numpy()
    got = _run(pkg, L, h, n_dev, t, W, X, mode)
    for i, y in enumerate(got):
        assert np.array_equal(y, want), f"device {i}, mode {mode}"


def test_rowsplit_cut_points(pkg, rs):
    L, h, n_dev = rs
    lo, hi = C.c_int64(), C.c_int64()
    prev = 0
    for i in range(n_dev):
        pkg.check(L.b200_rowsplit_rows(h, 128256, i, C.byref(lo), C.byref(hi)), "b200_rowsplit_rows")
        assert lo.value == prev and (hi.value % 128 == 0 or i == n_dev - 1)
        prev = hi.value
    assert prev == 128256
