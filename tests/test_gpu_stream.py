"""GPU tests (-m gpu) of b200_mul_mat_batch / the persistent job-table GEMV (mmvq_stream.cu), through the C ABI.

The kernel keeps gemv_q4k's per-row arithmetic, so every result must be BIT-IDENTICAL to b200_mul_mat on the same inputs — for
independent lists, for lists whose jobs feed each other (in-kernel completion counters), for lists mixed with jobs the kernel
does not take, under CUDA-graph replay, and against the oracle (NMSE <= 5e-4, tests/test-backend-ops.cpp:3946-3948).
"""
import zlib

import numpy as np
import pytest

from oracle import refenv
from tests.util import NMSE_TOL, nmse, oracle_mul_mat, random_blocks

pytestmark = pytest.mark.gpu
Q4_K = refenv.Q4_K


def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.fixture(autouse=True)
def every_list_takes_the_persistent_kernel(gpu_ctx):
    """b200_mul_mat_batch keeps lists shorter than stream_min_jobs (default 6) on the per-launch kernels; these tests are about the
    persistent kernel, so it takes every list here"""
    prev = gpu_ctx.set_option("stream_min_jobs", 1)
    prev_chain = gpu_ctx.set_option("stream_chain", 1)        # ... dependency chains included
    yield
    gpu_ctx.set_option("stream_min_jobs", prev)
    gpu_ctx.set_option("stream_chain", prev_chain)


def q4k(m, k, rng):
    return dev(random_blocks(Q4_K, m, k, rng))


def test_batch_independent_jobs_bit_identical_and_vs_oracle(pkg, gpu_ctx, orc):
    import torch
    rng = np.random.default_rng(11)
    # ragged row counts (fewer rows than CTAs, odd counts, one row), short and long rows, partial last chunks (k % 2048 != 0),
    # three matrices sharing their activations, the same x again later (a new group)
    shapes = [(1, 256), (7, 512), (147, 4096), (148, 4096), (149, 2048), (300, 14336), (1024, 4096), (1024, 4096), (333, 4096),
              (64, 768), (4096, 1280), (500, 4096)]
    xs_by_k = {}
    ws, xs = [], []
    for i, (m, k) in enumerate(shapes):
        ws.append(q4k(m, k, rng))
        if k not in xs_by_k or i in (3,):
            xs_by_k[k] = dev(rng.uniform(-1, 1, (1, k)).astype(np.float32))
        xs.append(xs_by_k[k])
    outs = pkg.mul_mat_batch(gpu_ctx, Q4_K, ws, xs)
    gpu_ctx.synchronize()
    for (m, k), w, x, y in zip(shapes, ws, xs, outs):
        want = pkg.mul_mat(gpu_ctx, Q4_K, w, x)
        assert torch.equal(y, want), (m, k)
        if m * k <= 1 << 22:
            e = nmse(oracle_mul_mat(orc, Q4_K, w.cpu().numpy(), x.cpu().numpy()), y.cpu().numpy())
            assert e <= NMSE_TOL, (m, k, e)


def chain(pkg, ctx, ws, x0, bufs):
    """y_i = W_i x_{i}, x_{i+1} = y_i, ping-pong over `bufs`; returns the list of output views"""
    xs, outs, x = [], [], x0
    for i, w in enumerate(ws):
        m = w.shape[0]
        y = bufs[i % len(bufs)].view(-1)[:m].view(1, m)
        xs.append(x)
        outs.append(y)
        x = y
    return xs, outs


@pytest.mark.parametrize("opts", [{}, {"stream_ll": 0}, {"stream_stages": 2}])
def test_batch_dependent_chain_matches_sequential_calls(pkg, gpu_ctx, opts):
    """every job reads the previous job's output (RAW) and overwrites the buffer the one before read (WAR); with the dependency
    carried in the data ({value, epoch} words, default), by completion counters only (stream_ll = 0) and
    with two ring slots per warp"""
    import torch
    prev = {k: gpu_ctx.set_option(k, v) for k, v in opts.items()}
    try:
        _chain_case(pkg, gpu_ctx)
    finally:
        for k, v in prev.items():
            gpu_ctx.set_option(k, v)


def _chain_case(pkg, gpu_ctx):
    import torch
    rng = np.random.default_rng(12)
    dims = [1024, 2048, 512, 4096, 256, 1024, 1024, 768, 256]           # k_0 = 1024, then m_i = k_{i+1}
    ws = [q4k(dims[i + 1], dims[i], rng) for i in range(len(dims) - 1)]
    x0 = dev((rng.standard_normal((1, dims[0])) * 40).astype(np.float32))
    A, B = (torch.zeros((1, 4096), device="cuda") for _ in range(2))
    xs, outs = chain(pkg, gpu_ctx, ws, x0, [A, B])
    # sequential reference
    want, x = [], x0
    for w in ws:
        x = pkg.mul_mat(gpu_ctx, Q4_K, w, x) * 1.0
        want.append(x)
    for rep in range(3):                                   # the completion counters must be back at zero after every launch
        A.fill_(float("nan")); B.fill_(float("nan"))
        batch = pkg.MatmulBatch(Q4_K, ws, xs, outs)
        batch.run(gpu_ctx)
        gpu_ctx.synchronize()
        assert torch.equal(outs[-1], want[-1]), rep
        assert torch.equal(outs[-2], want[-2]), rep


def test_batch_dependency_through_partial_overlap_and_tiny_jobs(pkg, gpu_ctx):
    """jobs with fewer rows than CTAs (most CTAs own no row of them) whose outputs are slices of a later job's activations"""
    import torch
    rng = np.random.default_rng(13)
    k0 = 512
    x0 = dev(rng.uniform(-1, 1, (1, k0)).astype(np.float32))
    buf = dev(rng.uniform(-1, 1, (1, 256)).astype(np.float32))
    w0, w1, w2 = q4k(64, k0, rng), q4k(3, k0, rng), q4k(200, 256, rng)
    y2 = torch.empty((1, 200), device="cuda")
    # reference, sequentially
    ref = buf.clone()
    ref[:, 0:64] = pkg.mul_mat(gpu_ctx, Q4_K, w0, x0)
    ref[:, 100:103] = pkg.mul_mat(gpu_ctx, Q4_K, w1, x0)
    want = pkg.mul_mat(gpu_ctx, Q4_K, w2, ref)
    for _ in range(2):
        cur = buf.clone()
        flat = cur.view(-1)
        pkg.mul_mat_batch(gpu_ctx, Q4_K, [w0, w1, w2], [x0, x0, cur], [flat[0:64].view(1, 64), flat[100:103].view(1, 3), y2])
        gpu_ctx.synchronize()
        assert torch.equal(cur, ref)
        assert torch.equal(y2, want)


def test_batch_mixed_with_jobs_the_stream_kernel_does_not_take(pkg, gpu_ctx):
    import torch
    rng = np.random.default_rng(14)
    k = 1024
    x1 = dev(rng.uniform(-1, 1, (1, k)).astype(np.float32))
    x3 = dev(rng.uniform(-1, 1, (3, k)).astype(np.float32))
    ws = [q4k(300, k, rng), q4k(260, k, rng), q4k(128, k, rng), q4k(512, k, rng), q4k(100, k, rng)]
    xs = [x1, x1, x3, x1, x1]                               # job 2 has three columns: a plain b200_mul_mat between two runs
    outs = pkg.mul_mat_batch(gpu_ctx, Q4_K, ws, xs)
    for w, x, y in zip(ws, xs, outs):
        assert torch.equal(y, pkg.mul_mat(gpu_ctx, Q4_K, w, x))
    # in-place is refused
    with pytest.raises(pkg.B200Error, match="overlaps"):
        sq = q4k(k, k, rng)
        pkg.mul_mat_batch(gpu_ctx, Q4_K, [sq, sq], [x1, x1], [x1, x1])


def test_batch_under_cuda_graph_replay(pkg, orc):
    """the job table is a kernel parameter: a captured launch replays with the same table, counters self-reset"""
    import torch
    rng = np.random.default_rng(15)
    stream = torch.cuda.Stream()
    ctx = pkg.Context(0, stream=stream.cuda_stream)
    ctx.set_option("stream_min_jobs", 1)
    ctx.set_option("stream_chain", 1)
    dims = [2048, 1024, 2048, 1024, 2048]
    ws = [q4k(dims[i + 1], dims[i], rng) for i in range(len(dims) - 1)]
    x0 = dev((rng.standard_normal((1, dims[0])) * 30).astype(np.float32))
    A, B = (torch.zeros((1, 2048), device="cuda") for _ in range(2))
    xs, outs = chain(pkg, ctx, ws, x0, [A, B])
    batch = pkg.MatmulBatch(Q4_K, ws, xs, outs)
    want, x = None, x0
    with torch.cuda.stream(stream):
        for w in ws:
            x = pkg.mul_mat(ctx, Q4_K, w, x) * 1.0
        want = x
        batch.run(ctx)
        stream.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            batch.run(ctx)
        for _ in range(5):
            A.zero_(); B.zero_()
            g.replay()
            stream.synchronize()
            assert torch.equal(outs[-1], want)
    ctx.close()


@pytest.mark.parametrize("m,k", [(4096, 4096), (14336, 4096), (4096, 14336)])
def test_batch_llama_shapes_sampled_rows_vs_oracle(pkg, gpu_ctx, orc, m, k):
    """BASELINE configs[1] shapes through the persistent kernel: q/k/v-style sharing, 512 sampled rows against the oracle"""
    import torch
    rng = np.random.default_rng(zlib.crc32(repr((m, k)).encode()))
    W = random_blocks(Q4_K, m, k, rng)
    X = rng.standard_normal((1, k)).astype(np.float32)
    Wd, Xd = dev(W), dev(X)
    outs = pkg.mul_mat_batch(gpu_ctx, Q4_K, [Wd, Wd[: m // 4], Wd], [Xd, Xd, Xd])
    rows = np.sort(rng.choice(m, 512, replace=False))
    want = oracle_mul_mat(orc, Q4_K, W[rows], X)
    got = outs[0].cpu().numpy()
    assert nmse(want, got[:, rows]) <= NMSE_TOL
    assert torch.equal(outs[0], outs[2]) and torch.equal(outs[1], outs[0][:, : m // 4])
    assert torch.equal(outs[0], pkg.mul_mat(gpu_ctx, Q4_K, Wd, Xd))


@pytest.mark.parametrize("world,slot", [(2, 16384), (4, 16384), (2, 8192)])
def test_row_split_list_all_ranks_in_one_launch(pkg, world, slot):
    """b200_mul_mat_batch_gather (the persistent kernel with the all-gather folded in) with `world` ranks emulated on ONE GPU inside
    ONE cooperative launch (gemv_q4k_stream<2>: CTAs [r G, (r + 1) G) are rank r; separate launches that wait for one another must
    not share a GPU).  Every rank has its own exchange buffer, activation images, flags, counters and vectors, works on its row
    shard and exchanges {value, epoch} words with the others exactly as over NVLink.  A dependent chain (q/k/v-style sharing, wo
    reads q, ...) must give, on every rank, bit for bit what sequential b200_mul_mat calls give on the unsplit matrices — three
    times in a row (both exchange slots, then slot reuse).  slot = 8192 entries do not hold the whole list: it is cut into two
    launches."""
    import ctypes as C

    import torch
    from llama_cpp_s390x_b200.binding import Tensor
    from llama_cpp_s390x_b200.ops import _desc
    rng = np.random.default_rng(21 + world)
    # (name, m, k): wq/wk/wv share x0; wo reads q; gate/up read wo's output; down reads gate's output; head reads down's output
    dims = [("wq", 1024, 512), ("wk", 256, 512), ("wv", 256, 512), ("wo", 512, 1024), ("gate", 1536, 512), ("up", 1536, 512), ("down", 512, 1536),
            ("head", 3072, 512)]
    Ws = {n: q4k(m, k, rng) for n, m, k in dims}
    x0 = dev((rng.standard_normal((1, 512)) * 30).astype(np.float32))
    ref_ctx = pkg.Context(0, stream="torch")
    want = {}
    want["wq"] = pkg.mul_mat(ref_ctx, Q4_K, Ws["wq"], x0); want["wk"] = pkg.mul_mat(ref_ctx, Q4_K, Ws["wk"], x0); want["wv"] = pkg.mul_mat(ref_ctx, Q4_K, Ws["wv"], x0)
    want["wo"] = pkg.mul_mat(ref_ctx, Q4_K, Ws["wo"], want["wq"])
    want["gate"] = pkg.mul_mat(ref_ctx, Q4_K, Ws["gate"], want["wo"]); want["up"] = pkg.mul_mat(ref_ctx, Q4_K, Ws["up"], want["wo"])
    want["down"] = pkg.mul_mat(ref_ctx, Q4_K, Ws["down"], want["gate"])
    want["head"] = pkg.mul_mat(ref_ctx, Q4_K, Ws["head"], want["down"])
    ref_ctx.synchronize()

    # one arena per rank: [x0 | y(wq) | y(wk) | ...], every vector 64-float aligned
    off, o = {}, 512
    for n, m, k in dims:
        off[n] = o
        o += (m + 63) // 64 * 64
    arena = torch.zeros((world, o), dtype=torch.float32, device="cuda")
    arena[:, :512] = x0

    def vec(name, length):
        start = 0 if name == "x0" else off[name]
        return arena[0, start:start + length].view(1, length)

    src = {"wq": "x0", "wk": "x0", "wv": "x0", "wo": "wq", "gate": "wo", "up": "wo", "down": "gate", "head": "down"}
    a = [_desc(Ws[n], Q4_K, k_elems=k) for n, m, k in dims]
    b = [_desc(vec(src[n], k), pkg.F32) for n, m, k in dims]
    d = [_desc(vec(n, m), pkg.F32) for n, m, k in dims]
    T = C.POINTER(Tensor)
    ap, bp, dp = ((T * len(dims))(*[C.pointer(t) for t in arr]) for arr in (a, b, d))
    ll = torch.zeros((world * 2 * slot,), dtype=torch.int64, device="cuda")
    stream = torch.cuda.Stream()
    ctx = pkg.Context(0, stream=stream.cuda_stream)
    torch.cuda.synchronize()
    L = pkg.lib()
    L.b200_dbg_batch_gather_emulated.restype = C.c_int
    L.b200_dbg_batch_gather_emulated.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p, C.c_int64, C.c_int, C.POINTER(T), C.POINTER(T), C.POINTER(T)]
    for rep in range(3):
        arena[:, 512:] = float("nan")
        torch.cuda.synchronize()
        pkg.check(L.b200_dbg_batch_gather_emulated(ctx.handle, world, arena.stride(0) * 4, ll.data_ptr(), slot, len(dims), ap, bp, dp), "b200_dbg_batch_gather_emulated")
        ctx.synchronize()
        for r in range(world):
            for n, m, k in dims:
                got = arena[r, off[n]:off[n] + m].view(1, m)
                assert torch.equal(got, want[n]), (rep, r, n)
    ctx.close()
    ref_ctx.close()
