"""GPU tool: n = 1..8 GEMV bandwidth of every weight type (CUDA graph, rotating weight copies > L2)."""
import sys
from pathlib import Path
import numpy as np
import torch
ROOT = Path(__file__).resolve().parents[1]; sys.path.insert(0, str(ROOT))
from __graft_entry__ import load_package
from tests.util import random_blocks
pkg = load_package()
stream = torch.cuda.Stream()
ctx = pkg.Context(0, stream=stream.cuda_stream)
rng = np.random.default_rng(0)
BPW = {"q4_0": 4.5, "q8_0": 8.5, "q4_K": 4.5, "q6_K": 6.5625}
TYPES = {"q4_0": 2, "q8_0": 8, "q4_K": 12, "q6_K": 14}
ns = [int(v) for v in sys.argv[1:]] or [1]
for n in ns:
    for name, t in TYPES.items():
        row = []
        for m, k in ((4096, 4096), (14336, 4096), (4096, 14336), (128256, 4096)):
            base = torch.from_numpy(random_blocks(t, 256, k, rng)).cuda().repeat((m + 255) // 256, 1)[:m].contiguous()
            copies = max(2, min(24, int(400e6 / base.numel()))) if m < 100000 else 1
            Ws = [base.clone() for _ in range(copies)]
            X = torch.randn((n, k), device="cuda"); Y = torch.empty((n, m), device="cuda")
            iters = 48 if m < 100000 else 6
            with torch.cuda.stream(stream):
                pkg.mul_mat(ctx, t, Ws[0], X, out=Y)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=stream):
                    for i in range(iters):
                        pkg.mul_mat(ctx, t, Ws[i % copies], X, out=Y)
                g.replay(); stream.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream); g.replay(); g.replay(); e1.record(stream); stream.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / (2 * iters)
            gb = (m * k * BPW[name] / 8 + n * k * 4 + m * n * 4) / us / 1e3
            row.append(f"{m}x{k}: {us:7.1f}us {gb:5.0f}GB/s")
            del Ws, base
        print(f"n={n} {name}: " + " | ".join(row), flush=True)
