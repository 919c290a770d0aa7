"""GPU tool: correctness (vs the CUDA-core generic kernel and, at small sizes, the oracle) and speed of the tcgen05 prefill GEMM."""
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
TYPES = {"q4_K": 12, "q8_0": 8, "q4_0": 2, "q6_K": 14}
if "--q6" in sys.argv:
    TYPES = {"q6_K": 14}
quick = "--quick" in sys.argv


def both(t, W, X):
    ctx.set_option("use_tc", 1)
    y1 = pkg.mul_mat(ctx, t, W, X)
    torch.cuda.synchronize()
    ctx.set_option("use_tc", 0)
    y0 = pkg.mul_mat(ctx, t, W, X)
    torch.cuda.synchronize()
    ctx.set_option("use_tc", 1)
    return y0.cpu().numpy(), y1.cpu().numpy()


cases = [(256, 1024, 64), (128, 256, 9), (200, 512, 130), (4096, 4096, 512), (300, 2048, 77)]
if quick:
    cases = cases[:2]
for name, t in TYPES.items():
    for m, k, n in cases:
        W = torch.from_numpy(random_blocks(t, m, k, rng)).cuda()
        X = torch.from_numpy(rng.uniform(-1, 1, (n, k)).astype(np.float32)).cuda()
        y0, y1 = both(t, W, X)
        print(f"{name} m={m} k={k} n={n}: nmse(tc, generic) = {nmse(y0, y1):.3e}  max|y|={np.abs(y0).max():.3f}", flush=True)

if not quick:
    # batch dims + broadcast, huge/small magnitudes
    W = torch.from_numpy(random_blocks(12, 2 * 256, 512, rng).reshape(2, 256, -1)).cuda()
    X = torch.from_numpy(rng.uniform(-1, 1, (2, 4, 33, 512)).astype(np.float32)).cuda()
    W4 = W.reshape(1, 2, 256, -1)
    y0, y1 = both(12, W4, X)
    print(f"batched q4_K: nmse = {nmse(y0, y1):.3e}", flush=True)
    for scale in (1e-30, 1e30):
        W = torch.from_numpy(random_blocks(8, 256, 512, rng)).cuda()
        X = torch.from_numpy((rng.uniform(-1, 1, (40, 512)) * scale).astype(np.float32)).cuda()
        y0, y1 = both(8, W, X)
        print(f"scale {scale:g}: nmse = {nmse(y0, y1):.3e} finite={np.isfinite(y1).all()}", flush=True)
    # MUL_MAT_ID, prefill-sized
    AS = torch.from_numpy(random_blocks(12, 8 * 512, 1024, rng).reshape(8, 512, -1)).cuda()
    B = torch.from_numpy(rng.uniform(-1, 1, (300, 1, 1024)).astype(np.float32)).cuda()
    ids = torch.from_numpy(np.stack([rng.permutation(8)[:2] for _ in range(300)]).astype(np.int32)).cuda()
    ctx.set_option("use_tc", 1); c1 = pkg.mul_mat_id(ctx, 12, AS, B, ids).cpu().numpy()
    ctx.set_option("use_tc", 0); c0 = pkg.mul_mat_id(ctx, 12, AS, B, ids).cpu().numpy()
    ctx.set_option("use_tc", 1)
    print(f"mul_mat_id q4_K 300 tokens: nmse = {nmse(c0, c1):.3e}", flush=True)

    # speed
    for cl in ((1, 0),):
        ctx.set_option("tc_split_k", cl[0]); ctx.set_option("tc_nt", cl[1])
        for name, t in TYPES.items():
            row = []
            for m, k, n in ((4096, 4096, 512), (14336, 4096, 512), (4096, 14336, 512), (128256, 4096, 512)):
                W = torch.from_numpy(random_blocks(t, 256, k, rng)).cuda().repeat((m + 255) // 256, 1)[:m].contiguous()
                X = torch.randn((n, k), device="cuda")
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
                row.append(f"{m}x{k}: {ms * 1e3:7.1f}us {2.0 * m * n * k / ms / 1e9:6.0f}TF")
            print(f"sched {cl} {name}: " + " | ".join(row), flush=True)
