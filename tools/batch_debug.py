import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
from __graft_entry__ import load_package
from oracle import refenv
from tests.util import random_blocks, oracle_mul_mat, nmse
pkg = load_package(); orc = refenv.load_oracle()
ctx = pkg.Context(0, stream="torch")
rng = np.random.default_rng(1)
for t, n, k, bs in ((12, 1, 256, (3, 1)), (12, 1, 1024, (3, 2)), (8, 1, 256, (2, 3)), (12, 8, 256, (2, 3)), (12, 16, 256, (2, 3)), (12, 1, 256, (1, 1))):
    m = 16
    W = random_blocks(t, bs[0] * bs[1] * m, k, rng).reshape(bs[1], bs[0], m, -1)
    X = rng.uniform(-1, 1, (bs[1], bs[0], n, k)).astype(np.float32)
    want = oracle_mul_mat(orc, t, W, X)
    got = pkg.mul_mat(ctx, t, torch.from_numpy(W).cuda(), torch.from_numpy(X).cuda()).cpu().numpy()
    errs = [[round(nmse(want[i3, i2], got[i3, i2]), 8) for i2 in range(bs[0])] for i3 in range(bs[1])]
    print(t, n, k, bs, "nmse per batch:", errs, flush=True)
