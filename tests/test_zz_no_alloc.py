"""b200_mul_mat_noalloc (-m gpu; collected last on purpose): the allocation-free MUL_MAT entry point, called through the C ABI,
against the existing entry point b200_mul_mat on the same inputs — on two calls in a row, because every call shares the one
static device scratch array and the second call sees the first call's leftovers.

  * vs b200_mul_mat's default route (fused / tensor-core GEMVs): NMSE <= 5e-4, the tolerance of every MUL_MAT test here
    (tests/test-backend-ops.cpp:3946-3948);
  * vs b200_mul_mat with force_generic = 1 (the same two kernels, planes in the context's cudaMalloc scratch): bit-identical;
  * sizes beyond B200_NOALLOC_SCRATCH_BYTES are rejected with a clear error, nothing is launched.
"""
import numpy as np
import pytest

from oracle import refenv
from tests.util import NMSE_TOL, nmse, random_blocks

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 1 << 20       # B200_NOALLOC_SCRATCH_BYTES (include/ggml_b200.h)


def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def run_pair(pkg, ctx, t, m, k, n, rng, batch=()):
    """one problem through both entry points; returns (noalloc, existing default route, existing generic route)"""
    import torch
    W = dev(random_blocks(t, m, k, rng))
    X = dev(rng.uniform(-1, 1, batch + (n, k)).astype(np.float32))
    out = torch.full(batch + (n, m), float("nan"), dtype=torch.float32, device="cuda")
    pkg.mul_mat_noalloc(ctx, t, W, X, out)
    want = pkg.mul_mat(ctx, t, W, X)
    prev = ctx.set_option("force_generic", 1)
    try:
        want_generic = pkg.mul_mat(ctx, t, W, X)
    finally:
        ctx.set_option("force_generic", prev)
    ctx.synchronize()
    return out.cpu().numpy(), want.cpu().numpy(), want_generic.cpu().numpy()


def test_noalloc_matches_mul_mat_two_calls_in_a_row(pkg, gpu_ctx):
    rng = np.random.default_rng(77)
    # first call: large planes (8 columns of k = 14336 -> 129 KB of the scratch); second call right after it: a different type,
    # fewer and shorter columns, so most of what the scratch holds is stale; third: batch dims + more than 8 columns per batch
    cases = [(refenv.Q4_K, 320, 14336, 8, ()), (refenv.Q8_0, 257, 1024, 3, ()), (refenv.Q6_K, 96, 2048, 11, (2,)), (refenv.Q4_0, 130, 4096, 1, ())]
    for t, m, k, n, batch in cases:
        got, want, want_generic = run_pair(pkg, gpu_ctx, t, m, k, n, rng, batch)
        assert np.isfinite(got).all(), (t, m, k, n)
        e = nmse(want, got)
        assert e <= NMSE_TOL, (t, m, k, n, e)
        assert np.array_equal(got, want_generic), (t, m, k, n)


def test_noalloc_rejects_sizes_beyond_the_static_scratch(pkg, gpu_ctx):
    import torch
    rng = np.random.default_rng(78)
    t, m, k = refenv.Q4_K, 64, 4096
    W = dev(random_blocks(t, m, k, rng))
    n = SCRATCH_BYTES // k + 1                       # int8 plane alone exceeds the array
    X = torch.zeros((n, k), dtype=torch.float32, device="cuda")
    out = torch.zeros((n, m), dtype=torch.float32, device="cuda")
    before = gpu_ctx.launches
    with pytest.raises(pkg.B200Error, match="static maximum"):
        pkg.mul_mat_noalloc(gpu_ctx, t, W, X, out)
    assert gpu_ctx.launches == before
    # the largest that fits still runs: ncols * k * (1 + 1/8) <= 1 MiB
    n_ok = SCRATCH_BYTES // (k + k // 8)
    X = dev(rng.uniform(-1, 1, (n_ok, k)).astype(np.float32))
    out = torch.empty((n_ok, m), dtype=torch.float32, device="cuda")
    pkg.mul_mat_noalloc(gpu_ctx, t, W, X, out)
    prev = gpu_ctx.set_option("force_generic", 1)
    try:
        want = pkg.mul_mat(gpu_ctx, t, W, X)
    finally:
        gpu_ctx.set_option("force_generic", prev)
    assert torch.equal(out, want)
