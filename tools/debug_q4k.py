"""debug helper (GPU): fast Q4_K kernel vs generic kernel, per column / per k"""
import sys
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parents[1]; sys.path.insert(0, str(ROOT))
from __graft_entry__ import load_package
from oracle import refenv
from tests.util import random_blocks, nmse
pkg = load_package()
ctx = pkg.Context(0, stream="torch")
rng = np.random.default_rng(0)
for k in (512, 1024, 2048, 4096, 8192):
    for n in (5, 6, 7, 8):
        m = 64
        W = torch.from_numpy(random_blocks(12, m, k, rng)).cuda(); X = torch.from_numpy(rng.standard_normal((n, k)).astype(np.float32)).cuda()
        fast = pkg.mul_mat(ctx, 12, W, X).cpu().numpy()
        ctx.set_option("force_generic", 1); slow = pkg.mul_mat(ctx, 12, W, X).cpu().numpy(); ctx.set_option("force_generic", 0)
        percol = [nmse(slow[c], fast[c]) for c in range(n)]
        bad = np.argwhere(np.abs(slow - fast) > 1e-3 * np.abs(slow).max())
        print(k, n, ["%.1e" % e for e in percol], "bad:", len(bad), bad[:6].tolist())
