"""GPU tool: tensor-core GEMV (mmvq_mma.cu) for Q4_K with 2..8 columns: agreement with the dp4a kernel / oracle, invariance of the
result under K slicing and row sharding, and bandwidth against the dp4a kernel (CUDA graph, rotating weight copies > L2)."""
import sys
from pathlib import Path
import numpy as np
import torch
ROOT = Path(__file__).resolve().parents[1]; sys.path.insert(0, str(ROOT))
from __graft_entry__ import load_package
from tests.util import nmse, random_blocks
pkg = load_package()
stream = torch.cuda.Stream()
ctx = pkg.Context(0, stream=stream.cuda_stream)
rng = np.random.default_rng(0)
Q4_K = 12


def run(W, X, **opts):
    for k_, v in opts.items():
        ctx.set_option(k_, v)
    with torch.cuda.stream(stream):
        y = pkg.mul_mat(ctx, Q4_K, W, X)
    stream.synchronize()
    for k_ in opts:
        ctx.set_option(k_, {"use_mma": 1, "mma_ks": 0, "mma_ch": 0}[k_])
    return y


for m, k, n in ((96, 512, 2), (257, 4096, 3), (64, 14336, 8), (1024, 4096, 8), (4099, 2048, 5)):
    W = torch.from_numpy(random_blocks(Q4_K, m, k, rng)).cuda()
    X = torch.from_numpy(rng.standard_normal((n, k)).astype(np.float32)).cuda()
    y0 = run(W, X, use_mma=0)
    y1 = run(W, X, use_mma=2)
    same = all(torch.equal(y1, run(W, X, use_mma=2, mma_ks=ks)) for ks in (1, 2, 4, 8, 16)) and torch.equal(y1, run(W, X, use_mma=2, mma_ch=1))
    lo, hi = m // 3, m // 3 + max(1, m // 2)
    shard = torch.equal(run(W[lo:hi].contiguous(), X, use_mma=2), y1[:, lo:hi])
    print(f"m={m} k={k} n={n}: nmse(mma, dp4a) = {nmse(y0.cpu().numpy(), y1.cpu().numpy()):.2e}  ks-invariant={same}  shard-invariant={shard}", flush=True)

BPW = 4.5
for n in (2, 4, 8):
    for mode in ("dp4a", "mma") + (("mma ch1", "mma ks=1", "mma ks=2", "mma ks=4", "mma ks=8", "mma ks=16") if "--sweep" in sys.argv else ()):
        ctx.set_option("use_mma", 0 if mode == "dp4a" else 2)
        ctx.set_option("mma_ks", int(mode.split("=")[1]) if "=" in mode else 0)
        ctx.set_option("mma_ch", 1 if "ch1" in mode else 0)
        row = []
        for m, k in ((1024, 4096), (4096, 4096), (14336, 4096), (4096, 14336), (128256, 4096)):
            base = torch.from_numpy(random_blocks(Q4_K, 256, k, rng)).cuda().repeat((m + 255) // 256, 1)[:m].contiguous()
            copies = max(2, min(24, int(400e6 / base.numel()))) if m < 100000 else 1
            Ws = [base.clone() for _ in range(copies)]
            X = torch.randn((n, k), device="cuda"); Y = torch.empty((n, m), device="cuda")
            iters = 48 if m < 100000 else 6
            with torch.cuda.stream(stream):
                pkg.mul_mat(ctx, Q4_K, Ws[0], X, out=Y)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=stream):
                    for i in range(iters):
                        pkg.mul_mat(ctx, Q4_K, Ws[i % copies], X, out=Y)
                g.replay(); stream.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream); g.replay(); g.replay(); e1.record(stream); stream.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / (2 * iters)
            gb = (m * k * BPW / 8 + n * k * 4 + m * n * 4) / us / 1e3
            row.append(f"{m}x{k}: {us:7.1f}us {gb:5.0f}GB/s")
            del Ws, base
        print(f"n={n} {mode:9s}: " + " | ".join(row), flush=True)
ctx.set_option("use_mma", 1); ctx.set_option("mma_ks", 0); ctx.set_option("mma_ch", 0)
