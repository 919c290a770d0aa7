"""GPU tool: where the time of small prefill GEMMs goes (eager loop vs CUDA graph, K sweep)."""
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
for opts in ({}, {"tc_pdl": 1}, {}, {"tc_pdl": 1}, {"tc_split_k": 2}, {"tc_nt": 256}):
    for k_, v in {"tc_split_k": 1, "tc_nt": 0, "tc_pdl": 0}.items():
        ctx.set_option(k_, v)
    for k_, v in opts.items():
        ctx.set_option(k_, v)
    row = []
    for m, k, n in ((4096, 256, 512), (1024, 4096, 512), (4096, 4096, 512), (14336, 4096, 512), (28672, 4096, 512), (128256, 4096, 512)):
        W = torch.from_numpy(random_blocks(12, 256, k, rng)).cuda().repeat((m + 255) // 256, 1)[:m].contiguous()
        X = torch.randn((n, k), device="cuda"); Y = torch.empty((n, m), device="cuda")
        with torch.cuda.stream(stream):
            for _ in range(3):
                pkg.mul_mat(ctx, 12, W, X, out=Y)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(20):
                pkg.mul_mat(ctx, 12, W, X, out=Y)
            e1.record(stream); stream.synchronize()
            eager = e0.elapsed_time(e1) / 20 * 1e3
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                for _ in range(20):
                    pkg.mul_mat(ctx, 12, W, X, out=Y)
            g.replay(); stream.synchronize()
            e0.record(stream); g.replay(); e1.record(stream); stream.synchronize()
            graph = e0.elapsed_time(e1) / 20 * 1e3
        row.append(f"{m}x{k}: eager {eager:6.1f} graph {graph:6.1f} us")
    print(f"{str(opts):34s} " + " | ".join(row), flush=True)
