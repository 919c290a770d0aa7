"""GPU tool: tcgen05 GEMM with single CTAs vs CTA pairs (cta_group::2): agreement and speed on the Llama-3-8B prefill shapes."""
import sys
from pathlib import Path
import numpy as np
import torch
ROOT = Path(__file__).resolve().parents[1]; sys.path.insert(0, str(ROOT))
from __graft_entry__ import load_package
from tests.util import nmse, random_blocks
pkg = load_package()
ctx = pkg.Context(0, stream="torch")
rng = np.random.default_rng(1)
TYPES = {"q4_K": 12, "q8_0": 8, "q4_0": 2}
for name, t in TYPES.items():
    for m, k, n in ((4096, 4096, 512), (14336, 4096, 512), (4096, 14336, 512), (128256, 4096, 512), (2048, 1024, 300)):
        W = torch.from_numpy(random_blocks(t, 256, k, rng)).cuda().repeat((m + 255) // 256, 1)[:m].contiguous()
        X = torch.randn((n, k), device="cuda")
        ys, row = [], []
        for pair in (0, 2):
            ctx.set_option("tc_pair", pair)
            Y = torch.empty((n, m), device="cuda")
            for _ in range(3):
                pkg.mul_mat(ctx, t, W, X, out=Y)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            iters = 20 if m < 100000 else 4
            e0.record()
            for _ in range(iters):
                pkg.mul_mat(ctx, t, W, X, out=Y)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / iters
            row.append(f"pair={pair}: {ms * 1e3:7.1f}us {2.0 * m * n * k / ms / 1e9:6.0f}TF")
            ys.append(Y.cpu().numpy())
        print(f"{name} {m}x{k}x{n}: " + " | ".join(row) + f" | nmse(pair, single) = {nmse(ys[0], ys[1]):.2e}", flush=True)
