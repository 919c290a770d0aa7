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


def test_gather_ll_two_ranks_in_one_process(pkg):
    """b200_mul_mat_gather_ll between two GPUs (two contexts of one process, peer-mapped exchange buffers): each rank's
    full-height results == the unsplit single-GPU results, bit for bit, over several dependent launches"""
    import torch
    if _n_gpus() < 2:
        pytest.skip("needs >= 2 GPUs")
    if not (torch.cuda.can_device_access_peer(0, 1) and torch.cuda.can_device_access_peer(1, 0)):
        pytest.skip("no peer access between GPU 0 and 1")
    q4k = TYPES["q4_K"]
    rng = np.random.default_rng(5)
    k, cap, world = 2048, 8192, 2
    # make torch enable peer access in both directions (cudaMalloc'ed blocks become peer-mapped)
    torch.zeros(8, device="cuda:0").to("cuda:1"); torch.zeros(8, device="cuda:1").to("cuda:0")
    bufs = [torch.zeros((2 * cap, 2), dtype=torch.int32, device=f"cuda:{r}") for r in range(world)]
    ctxs = []
    for r in range(world):
        with torch.cuda.device(r):
            ctxs.append(pkg.Context(r, stream="torch"))
    for r in range(world):
        torch.cuda.synchronize(r)
        pkg.gather_ll_init(ctxs[r], r, world, [b.data_ptr() for b in bufs], cap)
    x_np = rng.standard_normal((1, k)).astype(np.float32)
    xs = [torch.from_numpy(x_np).to(f"cuda:{r}") for r in range(world)]
    for ms in [(2048,), (2048, 512, 512), (3584, 3584), (2048,), (256, 128)]:
        W_np = [random_blocks(q4k, m, k, rng) for m in ms]
        with torch.cuda.device(0):
            want = [pkg.mul_mat(ctxs[0], q4k, torch.from_numpy(W).to("cuda:0"), xs[0]) for W in W_np]
        # all allocations and uploads first: nothing but the launches may happen while a receiver is polling for its peer
        prep = []
        for r in range(world):
            shards, row0s = [], []
            for W, m in zip(W_np, ms):
                lo, rows = pkg.row_shard(m, r, world)
                shards.append(torch.from_numpy(W[lo:lo + rows]).to(f"cuda:{r}")); row0s.append(lo)
            o = [torch.full((1, m), float("nan"), dtype=torch.float32, device=f"cuda:{r}") for m in ms]
            prep.append((shards, row0s, o))
        for r in range(world):
            torch.cuda.synchronize(r)
        outs = []
        for r in range(world):
            shards, row0s, o = prep[r]
            pkg.mul_mat_gather_ll(ctxs[r], q4k, shards, xs[r], o, row0s)           # asynchronous: rank 0's receiver waits for rank 1's GEMV
            outs.append((o, shards))
        for r in range(world):
            ctxs[r].synchronize()
        for r in range(world):
            for w, o in zip(want, outs[r][0]):
                assert torch.equal(w.cpu(), o.cpu())
        # dependent activations for the next round: x += tiny * first result (same on both ranks)
        xs = [x + 1e-3 * outs[r][0][0][:, :k] if ms[0] >= k else x for r, x in enumerate(xs)]
    for c in ctxs:
        c.close()
