"""debug: row-split list emulated on one GPU (two contexts)"""
import sys, os
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
from __graft_entry__ import load_package
from tests.util import random_blocks
pkg = load_package()
world = int(sys.argv[1]) if len(sys.argv) > 1 else 2
own = len(sys.argv) > 2 and sys.argv[2] == "own"
rng = np.random.default_rng(3)
def q4k(m, k): return torch.from_numpy(random_blocks(12, m, k, rng)).cuda()
dims = [("a", 1024, 512), ("b", 512, 1024)]
Ws = {n: q4k(m, k) for n, m, k in dims}
x0 = torch.randn((1, 512), device="cuda")
ref = pkg.Context(0, stream="torch")
wa = pkg.mul_mat(ref, 12, Ws["a"], x0); wb = pkg.mul_mat(ref, 12, Ws["b"], wa); ref.synchronize()
slot = 8192
streams = [torch.cuda.Stream() for _ in range(world)]
ctxs = [pkg.Context(0, stream=None if own else s.cuda_stream) for s in streams]
bufs = [torch.zeros((2 * slot,), dtype=torch.int64, device="cuda") for _ in range(world)]
torch.cuda.synchronize()
batches, outs = [], []
for r in range(world):
    ctxs[r].set_option("stream_grid", 148 // world - 2)
    pkg.stream_comm_init(ctxs[r], r, world, [b.data_ptr() for b in bufs], slot)
    ya = torch.full((1, 1024), float("nan"), device="cuda"); yb = torch.full((1, 512), float("nan"), device="cuda")
    sh, r0 = [], []
    for n, m, k in dims:
        lo, rows = pkg.row_shard(m, r, world)
        sh.append(Ws[n][lo:lo + rows].contiguous()); r0.append(lo)
    nj = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    batches.append(pkg.MatmulBatchGather(12, sh[:nj], [x0, ya][:nj], [ya, yb][:nj], r0[:nj]))
    outs.append((ya, yb))
torch.cuda.synchronize()
print("launching", flush=True)
for r in range(world):
    batches[r].run(ctxs[r])
    print("launched", r, flush=True)
for r in range(world):
    ctxs[r].synchronize()
    print("synced", r, flush=True)
for r in range(world):
    print(r, torch.equal(outs[r][0], wa), torch.equal(outs[r][1], wb))
print("slot words rank0:", bufs[0][slot:slot + 4].tolist(), bufs[0][slot + 1024: slot + 1028].tolist())
