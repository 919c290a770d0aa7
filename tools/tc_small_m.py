"""GPU tool: tcgen05 GEMM on short shards (m = 128..1024, long K): split-K / no split / CUDA-core agreement."""
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
for m, k, n in ((128, 8192, 512), (1024, 8192, 512), (3584, 8192, 512), (1024, 28672, 512), (128, 4096, 512)):
    W = torch.from_numpy(random_blocks(12, m, k, rng)).cuda()
    X = torch.randn((n, k), device="cuda")
    out = {}
    for name, opts in (("auto", {}), ("nosplit", {"tc_split_k": 0}), ("split2", {"tc_split_k": 2}), ("core", {"use_tc": 0})):
        for k_, v in {"tc_split_k": 1, "use_tc": 1}.items():
            ctx.set_option(k_, v)
        for k_, v in opts.items():
            ctx.set_option(k_, v)
        out[name] = pkg.mul_mat(ctx, 12, W, X).cpu().numpy()
        torch.cuda.synchronize()
    ctx.set_option("tc_split_k", 1); ctx.set_option("use_tc", 1)
    print(f"{m}x{k}x{n}: " + "  ".join(f"nmse({a}, core) = {nmse(out['core'], out[a]):.2e}" for a in ("auto", "nosplit", "split2")), flush=True)
