"""GPU parity (-m gpu) of the other graph ops of the C ABI at Llama-3-8B sizes, through the C ABI, against
plain PyTorch fp32 restatements of the reference's CPU forward functions (ggml-cpu/ops.cpp; lines cited in csrc/ops.cu) — the
reference's own harness runs the small grid (tests/test_gpu_backend_ops.py), these are the model-sized cases — and of the fused
MUL_MAT + bias + GLU launch against the separate calls and the oracle.  Float ops: relative tolerance written per test."""
import zlib

import numpy as np
import pytest

from tests.util import NMSE_TOL, TYPES, nmse, oracle_mul_mat, random_blocks

pytestmark = pytest.mark.gpu


def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def test_rms_norm_add_mul_scale(pkg, gpu_ctx):
    import torch
    g = torch.Generator(device="cuda"); g.manual_seed(3)
    x = torch.randn((1, 1, 7, 4096), device="cuda", generator=g) * 3.0
    w = torch.randn((1, 1, 1, 4096), device="cuda", generator=g)
    y = pkg.graph_op(gpu_ctx, "rms_norm", x, torch.empty_like(x), eps=1e-5)
    want = x / torch.sqrt((x.double() ** 2).mean(-1, keepdim=True).float() + 1e-5)
    assert float((y - want).abs().max()) <= 2e-6 * float(want.abs().max())
    z = pkg.graph_op(gpu_ctx, "mul", y, w, torch.empty_like(y))                       # broadcast over rows
    assert torch.equal(z, y * w)
    r = pkg.graph_op(gpu_ctx, "add", z, x, torch.empty_like(z))
    assert torch.equal(r, z + x)
    s = pkg.graph_op(gpu_ctx, "scale", r, torch.empty_like(r), scale=0.125, bias=1.0)
    assert torch.allclose(s, r * 0.125 + 1.0, rtol=1e-6, atol=1e-6)
    # strided source (a view of every second row), in place destination
    big = torch.randn((1, 1, 14, 4096), device="cuda", generator=g)
    view = big[:, :, ::2, :]
    out = pkg.graph_op(gpu_ctx, "add", view, x, torch.empty_like(x))
    assert torch.equal(out, view + x)


@pytest.mark.parametrize("mode", [0, 2])
def test_rope_llama3_heads(pkg, gpu_ctx, mode):
    """32 heads of 128, 5 positions, Llama-3 frequency factors, theta base 500000 (normal and NeoX pair layouts)"""
    import torch
    g = torch.Generator(device="cuda"); g.manual_seed(5)
    n_tok, n_head, hd = 5, 32, 128
    x = torch.randn((1, n_tok, n_head, hd), device="cuda", generator=g)
    pos = torch.tensor([0, 1, 17, 4095, 8191], dtype=torch.int32, device="cuda")
    ff = 1.0 + torch.rand((hd // 2,), device="cuda", generator=g) * 7.0
    y = pkg.graph_op(gpu_ctx, "rope", x, pos, ff, torch.empty_like(x), n_dims=hd, mode=mode, n_ctx_orig=8192, freq_base=500000.0)
    i = torch.arange(hd // 2, device="cuda", dtype=torch.float64)
    theta = pos.double()[:, None] * (500000.0 ** (-2.0 * i / hd))[None, :] / ff.double()[None, :]            # [tok, hd/2]
    c, s = torch.cos(theta)[None, :, None, :], torch.sin(theta)[None, :, None, :]
    xd = x.double()
    if mode == 0:
        x0, x1 = xd[..., 0::2], xd[..., 1::2]
        want = torch.stack([x0 * c - x1 * s, x0 * s + x1 * c], dim=-1).reshape(x.shape)
    else:
        x0, x1 = xd[..., : hd // 2], xd[..., hd // 2:]
        want = torch.cat([x0 * c - x1 * s, x0 * s + x1 * c], dim=-1)
    # angles up to 8191 rad in fp32: 1e-3 absolute on values of order 1 (the reference's own fp32 theta carries the same error)
    assert float((y.double() - want).abs().max()) <= 2e-3


def test_soft_max_with_f16_mask_and_scale(pkg, gpu_ctx):
    import torch
    g = torch.Generator(device="cuda"); g.manual_seed(7)
    n_kv, n_tok, n_head = 4096, 3, 32
    x = torch.randn((1, n_head, n_tok, n_kv), device="cuda", generator=g) * 4.0
    mask = torch.zeros((1, 1, n_tok, n_kv), device="cuda", dtype=torch.float16)
    for t in range(n_tok):
        mask[0, 0, t, 1000 + t:] = float("-inf")
    y = pkg.graph_op(gpu_ctx, "soft_max", x, mask, None, torch.empty_like(x), scale=0.088388)
    want = torch.softmax(x.double() * 0.088388 + mask.double(), dim=-1)
    assert float((y.double() - want).abs().max()) <= 1e-6
    assert float(y[..., 1002:].abs().max()) == 0.0 or bool((y[0, :, 0, 1000:] == 0).all())


def test_kv_store_and_attention_products(pkg, gpu_ctx):
    """SET_ROWS into an F16 cache by I64 index, then K.q and V.p over permuted views with grouped-query broadcast (8 KV heads, 32 heads)"""
    import torch
    g = torch.Generator(device="cuda"); g.manual_seed(9)
    n_kv, hd, n_head_kv, n_head = 256, 128, 8, 32
    cache = torch.zeros((1, 1, n_kv, n_head_kv * hd), device="cuda", dtype=torch.float16)
    k_new = torch.randn((1, 1, 3, n_head_kv * hd), device="cuda", generator=g)
    idx = torch.tensor([[[5, 200, 17]]], dtype=torch.int64, device="cuda").reshape(1, 1, 1, 3)
    pkg.graph_op(gpu_ctx, "set_rows", k_new, idx, cache)
    assert torch.equal(cache[0, 0, [5, 200, 17]], k_new[0, 0].half())
    assert float(cache[0, 0, 6].abs().max()) == 0.0
    cache.copy_(torch.randn(cache.shape, device="cuda", generator=g).half())
    # K as ggml sees it after view + permute: [hd, n_kv, n_head_kv] with nb1 = row of the cache, nb2 = hd elements
    k_view = cache.view(n_kv, n_head_kv, hd).permute(1, 0, 2).unsqueeze(0)          # torch (1, n_head_kv, n_kv, hd), strided
    q = torch.randn((1, n_head, 1, hd), device="cuda", generator=g)
    kq = pkg.graph_op(gpu_ctx, "mul_mat_f", k_view, q, torch.empty((1, n_head, 1, n_kv), device="cuda"))
    want = torch.einsum("hkd,hd->hk", k_view[0].float().repeat_interleave(n_head // n_head_kv, dim=0), q[0, :, 0])
    assert nmse(want.cpu().numpy(), kq[0, :, 0].cpu().numpy()) <= 1e-10


@pytest.mark.parametrize("n", [17, 512])
def test_f16_weights_prefill_on_the_tensor_cores(pkg, gpu_ctx, n):
    """MUL_MAT with a plain fp16 matrix and many columns: the tcgen05 GEMM reads the fp16 rows as they are (two launches), against the
    SIMT tile kernel (use_tc = 0) and the f64 product"""
    import torch
    g = torch.Generator(device="cuda"); g.manual_seed(n)
    m, k = 1000, 2048
    W = (torch.randn((1, 1, m, k), device="cuda", generator=g) * 0.05).half()
    X = torch.randn((1, 1, n, k), device="cuda", generator=g)
    l0 = gpu_ctx.launches
    got = pkg.graph_op(gpu_ctx, "mul_mat_f", W, X, torch.empty((1, 1, n, m), device="cuda"))
    torch.cuda.synchronize()
    assert gpu_ctx.launches - l0 == 2
    want = X[0, 0].double() @ W[0, 0].double().T
    assert nmse(want.cpu().numpy(), got[0, 0].cpu().numpy()) <= 1e-6
    try:
        gpu_ctx.set_option("use_tc", 0)
        simt = pkg.graph_op(gpu_ctx, "mul_mat_f", W, X, torch.empty((1, 1, n, m), device="cuda"))
    finally:
        gpu_ctx.set_option("use_tc", 1)
    assert nmse(want.cpu().numpy(), simt[0, 0].cpu().numpy()) <= 1e-10


def test_prompt_attention_products_on_the_tensor_cores(pkg, gpu_ctx):
    """K.q and V.p of a 512-token prompt (Llama-3-8B: 32 heads over 8 KV heads, head size 128): K and V^T are strided fp16 views of the cache,
    q is a permuted f32 view; the tcgen05 GEMM takes them as a batch (2 launches each) — against the SIMT tile kernel and torch"""
    import torch
    g = torch.Generator(device="cuda"); g.manual_seed(5)
    n_tok, n_kv, hd, n_head_kv, n_head, n_ctx = 512, 512, 128, 8, 32, 1024
    kc = (torch.randn((n_ctx, n_head_kv * hd), device="cuda", generator=g)).half()              # K cache: one row per position
    vc = (torch.randn((n_head_kv * hd, n_ctx), device="cuda", generator=g)).half()              # V cache, transposed: one row per channel
    k_view = kc[:n_kv].view(n_kv, n_head_kv, hd).permute(1, 0, 2).unsqueeze(0)                   # ggml [hd, n_kv, n_head_kv]
    v_view = vc.view(n_head_kv, hd, n_ctx)[:, :, :n_kv].unsqueeze(0)                             # ggml [n_kv, hd, n_head_kv]
    q = torch.randn((n_tok, n_head, hd), device="cuda", generator=g).permute(1, 0, 2).unsqueeze(0)      # ggml [hd, n_tok, n_head], strided
    p_ = torch.softmax(torch.randn((1, n_head, n_tok, n_kv), device="cuda", generator=g), dim=-1)
    rep = n_head // n_head_kv
    for a, b, m, want in ((k_view, q, n_kv, torch.einsum("hkd,htd->htk", k_view[0].double().repeat_interleave(rep, dim=0), q[0].double())),
                          (v_view, p_, hd, torch.einsum("hdk,htk->htd", v_view[0].double().repeat_interleave(rep, dim=0), p_[0].double()))):
        l0 = gpu_ctx.launches
        got = pkg.graph_op(gpu_ctx, "mul_mat_f", a, b, torch.empty((1, n_head, n_tok, m), device="cuda"))
        torch.cuda.synchronize()
        assert gpu_ctx.launches - l0 == 2
        assert nmse(want.cpu().numpy(), got[0].cpu().numpy()) <= 1e-6
        try:
            gpu_ctx.set_option("use_tc", 0)
            simt = pkg.graph_op(gpu_ctx, "mul_mat_f", a, b, torch.empty((1, n_head, n_tok, m), device="cuda"))
        finally:
            gpu_ctx.set_option("use_tc", 1)
        assert nmse(want.cpu().numpy(), simt[0].cpu().numpy()) <= 1e-10


@pytest.mark.parametrize("name", ["f32", "q4_K", "q6_K", "q8_0", "q4_0"])
def test_get_rows_token_embeddings(pkg, gpu_ctx, orc, name):
    import torch
    rng = np.random.default_rng(11)
    idx = torch.tensor([3, 0, 77, 3, 127], dtype=torch.int32, device="cuda").reshape(1, 1, 1, 5)
    if name == "f32":
        table = torch.randn((1, 1, 128, 4096), device="cuda")
        got = pkg.graph_op(gpu_ctx, "get_rows", table, idx, torch.empty((1, 1, 5, 4096), device="cuda"))
        assert torch.equal(got[0, 0], table[0, 0, idx.flatten().long()])
        return
    from oracle import refenv
    t = TYPES[name]
    W = np.ascontiguousarray(random_blocks(t, 128, 4096, rng))
    want = np.zeros((5, 4096), np.float32)
    for j, r in enumerate([3, 0, 77, 3, 127]):
        assert orc.orc_dequantize_row(t, refenv.ptr(W[r]), refenv.ptr(want[j]), 4096) == 0
    from importlib import import_module
    ops = import_module(pkg.__name__ + ".ops")
    Wd = dev(W).reshape(1, 1, 128, -1)                 # (kept alive: the descriptor holds a raw pointer)
    a = ops._desc(Wd, t, k_elems=4096)
    out = torch.empty((1, 1, 5, 4096), device="cuda")
    import ctypes as C
    ops.check(ops.lib().b200_op_get_rows(gpu_ctx.handle, C.byref(a), C.byref(ops._fdesc(idx)), C.byref(ops._fdesc(out))), "b200_op_get_rows")
    assert np.array_equal(out[0, 0].cpu().numpy(), want)          # dequantize_row_* bit for bit


@pytest.mark.parametrize("glu", [2, 1])          # SWIGLU, GEGLU
@pytest.mark.parametrize("name", ["q4_K", "q8_0", "q4_0", "q6_K"])
def test_fused_gate_up_glu_equals_separate_calls(pkg, gpu_ctx, orc, name, glu):
    """b200_mul_mat_fused (one launch) against b200_mul_mat x 2 + b200_op_binary x 2 + b200_op_glu, Llama-3-8B FFN shapes, and the up
    product against the oracle"""
    import torch
    t = TYPES[name]
    m, k = 14336, 4096
    rng = np.random.default_rng(zlib.crc32(repr((name, glu)).encode()))
    Wu = np.tile(random_blocks(t, 512, k, rng), (m // 512, 1)); Wg = np.roll(Wu, 131, axis=0).copy()
    X = rng.standard_normal((1, k)).astype(np.float32)
    bu = rng.standard_normal((1, m)).astype(np.float32); bg = rng.standard_normal((1, m)).astype(np.float32)
    Wud, Wgd, Xd, bud, bgd = dev(Wu), dev(Wg), dev(X), dev(bu), dev(bg)
    fused = pkg.mul_mat_fused(gpu_ctx, t, Wud, Xd, w_gate=Wgd, bias_up=bud, bias_gate=bgd, glu_op=glu)
    gpu_ctx.set_option("force_generic", 1)           # the fused launch walks the rows like the generic GEMV: same sums, same bits
    try:
        up = pkg.mul_mat(gpu_ctx, t, Wud, Xd); gate = pkg.mul_mat(gpu_ctx, t, Wgd, Xd)
    finally:
        gpu_ctx.set_option("force_generic", 0)
    up_b = pkg.graph_op(gpu_ctx, "add", up, bud, torch.empty_like(up)); gate_b = pkg.graph_op(gpu_ctx, "add", gate, bgd, torch.empty_like(gate))
    sep = pkg.graph_op(gpu_ctx, "glu", gate_b, up_b, torch.empty_like(up), op=glu)
    assert torch.equal(fused, sep)
    act = torch.nn.functional.silu(gate_b) if glu == 2 else torch.nn.functional.gelu(gate_b, approximate="tanh")
    assert nmse((act * up_b).cpu().numpy(), fused.cpu().numpy()) <= 1e-9
    rows = np.sort(rng.choice(m, 64, replace=False))
    assert nmse(oracle_mul_mat(orc, t, np.ascontiguousarray(Wu[rows]), X), up.cpu().numpy()[:, rows]) <= NMSE_TOL
    # bias only
    only = pkg.mul_mat_fused(gpu_ctx, t, Wud, Xd, bias_up=bud)
    assert torch.equal(only, up_b)


@pytest.mark.parametrize("n", [1, 3, 8])
@pytest.mark.parametrize("name,m,k", [("q4_K", 4096, 14336), ("q4_K", 1000, 4096), ("q8_0", 4096, 4096), ("q6_K", 2048, 4096), ("q4_0", 4096, 4096), ("q5_K", 512, 2048)])
def test_mul_mat_add_equals_mul_mat_then_add(pkg, gpu_ctx, name, m, k, n):
    """b200_mul_mat_add (the addend goes into the store of the fast GEMV kernels) == b200_mul_mat followed by the ADD, bit for bit, also
    when the result overwrites the addend (the residual stream in place)"""
    import torch
    from tests.util import TYPES as T2
    t = {**T2, "q5_K": 13}[name]
    rng = np.random.default_rng(zlib.crc32(repr((name, m, k, n)).encode()))
    if name == "q5_K":
        W = torch.randint(0, 256, (m, k // 256 * 176), dtype=torch.uint8, device="cuda")
        W.view(m, k // 256, 176)[:, :, 0:4] = torch.tensor([0x00, 0x1c, 0x00, 0x1c], dtype=torch.uint8, device="cuda")      # d = dmin = 2^-8
    else:
        W = dev(random_blocks(t, m, k, rng))
    X = torch.randn((n, k), device="cuda")
    R = torch.randn((n, m), device="cuda")
    want = pkg.mul_mat(gpu_ctx, t, W, X) + R
    got = pkg.mul_mat_add(gpu_ctx, t, W, X, R)
    assert torch.equal(got, want)
    R2 = R.clone()
    pkg.mul_mat_add(gpu_ctx, t, W, X, R2, out=R2)
    assert torch.equal(R2, want)
