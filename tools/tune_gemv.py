"""GPU tool: per-shape time of the n=1 Q4_K GEMV under the ctx options (CTA shape, PDL, launch-floor probe)."""
import sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[1]; sys.path.insert(0, str(ROOT))
from __graft_entry__ import load_package
import bench
pkg = load_package()
stream = torch.cuda.Stream()
ctx = pkg.Context(0, stream=stream.cuda_stream)
gen = torch.Generator(device="cuda"); gen.manual_seed(0)
SHAPES = [("wk", 1024, 4096, 32), ("wq", 4096, 4096, 32), ("gate", 14336, 4096, 24), ("down", 4096, 14336, 24), ("output", 128256, 4096, 2)]
data = {}
for name, m, k, count in SHAPES:
    x = torch.randn((1, k), device="cuda")
    data[name] = [(bench.device_random_q4k(torch, m, k, gen), x, torch.empty((1, m), device="cuda")) for _ in range(count)]

def measure(name, iters):
    group = data[name]
    with torch.cuda.stream(stream):
        for W, x, y in group[:2]:
            pkg.mul_mat(ctx, 12, W, x, out=y)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            for i in range(iters):
                W, x, y = group[i % len(group)]
                pkg.mul_mat(ctx, 12, W, x, out=y)
        g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(5):
            g.replay()
        e1.record(stream)
        stream.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (5 * iters)

configs = [("cfg0 32w", {"gemv_cfg": 0}), ("cfg1 16wx2", {"gemv_cfg": 1}), ("cfg2 32w+l2pf", {"gemv_cfg": 2}),
           ("cfg0 nodefer", {"gemv_cfg": 0, "defer_wait": 0}), ("cfg0 nopdl", {"gemv_cfg": 0, "pdl": 0}), ("noop", {"noop": 1})]
for label, opts in configs:
    for k_, v in {"gemv_cfg": 0, "pdl": 1, "noop": 0, "defer_wait": 1}.items():
        ctx.set_option(k_, v)
    for k_, v in opts.items():
        ctx.set_option(k_, v)
    row = []
    for name, m, k, count in SHAPES:
        us = measure(name, 64 if m < 100000 else 8)
        row.append(f"{name} {us:7.2f}us {bench.algo_bytes(m, k) / us / 1e3:7.0f}GB/s")
    print(f"{label:12s} | " + " | ".join(row), flush=True)
