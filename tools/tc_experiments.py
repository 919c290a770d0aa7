"""GPU tool: timing experiments on the tcgen05 GEMM with parts of the pipeline disabled (results are garbage by design)."""
import sys
from pathlib import Path
import numpy as np
import torch
ROOT = Path(__file__).resolve().parents[1]; sys.path.insert(0, str(ROOT))
from __graft_entry__ import load_package
from tests.util import random_blocks
pkg = load_package()
ctx = pkg.Context(0, stream="torch")
rng = np.random.default_rng(0)
for t, name in ((12, "q4_K"),):
    m, k, n = 128256, 4096, 512
    W = torch.from_numpy(random_blocks(t, 256, k, rng)).cuda().repeat((m + 255) // 256, 1)[:m].contiguous()
    X = torch.randn((n, k), device="cuda")
    Y = torch.empty((n, m), device="cuda")
    for dbg, label in ((0, "full"), (1, "no weight loads"), (3, "no weight loads, no convert/STS"), (4, "no TMA"), (8, "no MMA"),
                       (7, "MMA only"), (12, "producers only"), (11, "TMA only"), (15, "nothing"), (31, "nothing, no stores"), (16, "full, no stores"), (31 + 32, "nothing, no stores, no fence"), (31 + 64, "nothing, no stores, arrive instead of commit"), (31 + 96, "nothing, no stores, no fence, arrive")):
        ctx.set_option("tc_dbg", dbg)
        for _ in range(2):
            pkg.mul_mat(ctx, t, W, X, out=Y)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(4):
            pkg.mul_mat(ctx, t, W, X, out=Y)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 4
        print(f"{name} dbg={dbg:2d} {label:34s}: {ms * 1e3:8.1f} us  ({2.0 * m * n * k / ms / 1e9:6.0f} TF-equivalent)", flush=True)
