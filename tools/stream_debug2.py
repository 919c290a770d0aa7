import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
from __graft_entry__ import load_package
import bench
pkg = load_package()
gen = torch.Generator(device="cuda"); gen.manual_seed(5)
ctx = pkg.Context(0, stream="torch")
world = 2
shapes = [("wq", 4096, 4096), ("wk", 1024, 4096), ("wv", 1024, 4096), ("wo", 4096, 4096), ("gate", 14336, 4096), ("up", 14336, 4096), ("down", 4096, 14336)]
xs = {4096: torch.randn((1, 4096), device="cuda"), 14336: torch.randn((1, 14336), device="cuda")}
ws, xl, names = [], [], []
for l in range(3):
    for n, m, k in shapes:
        ws.append(bench.device_random_q4k(torch, m // world, k, gen)); xl.append(xs[k]); names.append(f"{n}{l}")
outs = pkg.mul_mat_batch(ctx, 12, ws, xl)
ctx.synchronize()
bad = [(nm, int((y != pkg.mul_mat(ctx, 12, w, x)).sum())) for nm, w, x, y in zip(names, ws, xl, outs) if not torch.equal(y, pkg.mul_mat(ctx, 12, w, x))]
print("single-GPU batch mismatches:", bad)
# two ranks emulated
slot = 1 << 18
streams = [torch.cuda.Stream() for _ in range(world)]
ctxs = [pkg.Context(0, stream=s.cuda_stream) for s in streams]
bufs = [torch.zeros((2 * slot,), dtype=torch.int64, device="cuda") for _ in range(world)]
torch.cuda.synchronize()
full = [torch.cat([ws[i], ws[i]], 0) for i in range(len(ws))]     # both ranks hold the same shard bytes: full matrix = shard twice
batches, youts = [], []
for r in range(world):
    ctxs[r].set_option("stream_grid", 72)
    pkg.stream_comm_init(ctxs[r], r, world, [b.data_ptr() for b in bufs], slot)
    ys = [torch.full((1, 2 * w.shape[0]), float("nan"), device="cuda") for w in ws]
    batches.append(pkg.MatmulBatchGather(12, ws, xl, ys, [r * w.shape[0] for w in ws])); youts.append(ys)
torch.cuda.synchronize()
for rep in range(2):
    for r in range(world): batches[r].run(ctxs[r])
    for r in range(world): ctxs[r].synchronize()
for r in range(world):
    bad = []
    for i, (nm, w, x) in enumerate(zip(names, ws, xl)):
        want = pkg.mul_mat(ctx, 12, w, x); ctx.synchronize()
        y = youts[r][i]; rows = w.shape[0]
        if not (torch.equal(y[:, :rows], want) and torch.equal(y[:, rows:], want)):
            bad.append((nm, int((y[:, :rows] != want).sum()), int((y[:, rows:] != want).sum()), int(torch.isnan(y).sum())))
    print("rank", r, "mismatches:", bad)
