"""GPU tool: phase time stamps of the tensor-core GEMV (library built with -DMMA_TRACE_ENABLE)."""
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
t = int(sys.argv[1]) if len(sys.argv) > 1 else 8
for m, k, n in ((1024, 4096, 1), (4096, 4096, 1), (4096, 4096, 8), (14336, 4096, 8)):
    W = torch.from_numpy(random_blocks(t, 256, k, rng)).cuda().repeat(m // 256, 1).contiguous()
    X = torch.randn((n, k), device="cuda")
    print(f"type {t} {m}x{k} n={n}", flush=True)
    for _ in range(3):
        pkg.mul_mat(ctx, t, W, X); torch.cuda.synchronize()
