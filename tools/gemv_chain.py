"""GPU tool: DEPENDENT chains of matmuls (y1 = W1 x, y2 = W2 y1, ...), the regime of a real decode step: per-matmul time with the
activations quantized in every CTA's prologue (mma_prequant 0) vs once per call by a small kernel (1 = when dependent, the default)."""
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
TYPES = {"q4_0": 2, "q8_0": 8, "q4_K": 12, "q6_K": 14}
m = k = 4096
L = 24
for name, t in TYPES.items():
    Ws = [torch.from_numpy(random_blocks(t, m, k, rng, scale=0.6)).cuda() for _ in range(L)]
    for n in (1, 4, 8):
        row = []
        for pre in (0, 1):
            ctx.set_option("mma_prequant", pre)
            bufs = [torch.randn((n, k), device="cuda")] + [torch.empty((n, m), device="cuda") for _ in range(L)]
            with torch.cuda.stream(stream):
                for i in range(L):
                    pkg.mul_mat(ctx, t, Ws[i], bufs[i], out=bufs[i + 1])
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=stream):
                    for i in range(L):
                        pkg.mul_mat(ctx, t, Ws[i], bufs[i], out=bufs[i + 1])
                g.replay(); stream.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream); g.replay(); g.replay(); e1.record(stream); stream.synchronize()
            row.append(e0.elapsed_time(e1) * 1e3 / (2 * L))
        print(f"{name} n={n} chain of {L} dependent {m}x{k}: prologue in every CTA {row[0]:6.2f} us / quantized once {row[1]:6.2f} us per matmul", flush=True)
ctx.set_option("mma_prequant", 1)
