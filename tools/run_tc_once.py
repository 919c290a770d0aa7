"""GPU tool: a few launches of the tcgen05 prefill GEMM at one shape (for ncu)."""
import sys
from pathlib import Path
import numpy as np
import torch
ROOT = Path(__file__).resolve().parents[1]; sys.path.insert(0, str(ROOT))
from __graft_entry__ import load_package
from tests.util import random_blocks
pkg = load_package()
ctx = pkg.Context(0, stream="torch")
t = int(sys.argv[1]) if len(sys.argv) > 1 else 12
m, k, n = (int(v) for v in sys.argv[2:5]) if len(sys.argv) > 4 else (14336, 4096, 512)
rng = np.random.default_rng(0)
W = torch.from_numpy(random_blocks(t, 256, k, rng)).cuda().repeat((m + 255) // 256, 1)[:m].contiguous()
X = torch.randn((n, k), device="cuda")
Y = torch.empty((n, m), device="cuda")
for _ in range(4):
    pkg.mul_mat(ctx, t, W, X, out=Y)
torch.cuda.synchronize()
print("done", float(Y.abs().max()))
