"""where does a row-split step's time go?  The Llama-3-8B list with every matrix cut to 1/div of its rows, on ONE GPU through
b200_mul_mat_batch (no exchange at all): the persistent kernel's own floor for small shards"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from __graft_entry__ import load_package
import bench
pkg = load_package()
stream = torch.cuda.Stream()
ctx = pkg.Context(0, stream=stream.cuda_stream)
gen = torch.Generator(device="cuda"); gen.manual_seed(1)
for div in (1, 2, 4, 8, 16):
    xs = {4096: torch.randn((1, 4096), device="cuda"), 14336: torch.randn((1, 14336), device="cuda")}
    ws, xl, ys = [], [], []
    layer = [("wq", 4096, 4096), ("wk", 1024, 4096), ("wv", 1024, 4096), ("wo", 4096, 4096), ("g", 14336, 4096), ("u", 14336, 4096), ("d", 4096, 14336)]
    for l in range(32):
        for n, m, k in layer:
            ws.append(bench.device_random_q4k(torch, m // div, k, gen)); xl.append(xs[k]); ys.append(torch.empty((1, m // div), device="cuda"))
    ws.append(bench.device_random_q4k(torch, 128256 // div, 4096, gen)); xl.append(xs[4096]); ys.append(torch.empty((1, 128256 // div), device="cuda"))
    b = pkg.MatmulBatch(12, ws, xl, ys)
    with torch.cuda.stream(stream):
        b.run(ctx); stream.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            b.run(ctx)
        for _ in range(3): g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(10): g.replay()
        e1.record(stream); stream.synchronize()
    ms = e0.elapsed_time(e1) / 10
    nbytes = sum(w.numel() for w in ws)
    print(f"rows / {div:2d}: {ms:.4f} ms per step, {nbytes / 1e9:.3f} GB -> {nbytes / ms / 1e6:.0f} GB/s (HBM time {nbytes / 6.55e9:.3f} ms)", flush=True)
    del ws, ys, b, g
    torch.cuda.empty_cache()
