import os, sys
from pathlib import Path
import numpy as np
import torch
ROOT = Path(__file__).resolve().parents[1]; sys.path.insert(0, str(ROOT))
os.environ.setdefault('GGML_B200_KERNELS_SO', str(ROOT / 'tools' / 'bin' / 'libggml_b200_kernels_trace.so'))
from __graft_entry__ import load_package
from tests.util import random_blocks
pkg = load_package()
ctx = pkg.Context(0, stream="torch")
rng = np.random.default_rng(0)
m, k, n = 14336, 4096, 512
W = torch.from_numpy(random_blocks(12, 256, k, rng)).cuda().repeat(m // 256, 1).contiguous()
X = torch.randn((n, k), device="cuda"); Y = torch.empty((n, m), device="cuda")
pkg.mul_mat(ctx, 12, W, X, out=Y); torch.cuda.synchronize()
pkg.mul_mat(ctx, 12, W, X, out=Y); torch.cuda.synchronize()
