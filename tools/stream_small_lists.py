"""per-call time of short lists: b200_mul_mat_batch through the persistent kernel vs the per-launch kernels (CUDA graph of 32 calls)"""
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
cases = {"qkv": [(4096, 4096), (1024, 4096), (1024, 4096)], "wo": [(4096, 4096)], "gate_up": [(14336, 4096), (14336, 4096)], "down": [(4096, 14336)],
         "layer(7)": [(4096, 4096), (1024, 4096), (1024, 4096), (4096, 4096), (14336, 4096), (14336, 4096), (4096, 14336)]}
NL = 16
for name, shapes in cases.items():
    sets = []
    for l in range(NL):
        ws = [bench.device_random_q4k(torch, m, k, gen) for m, k in shapes]
        xs = {k: torch.randn((1, k), device="cuda") for _, k in shapes}
        sets.append(pkg.MatmulBatch(12, ws, [xs[k] for _, k in shapes], [torch.empty((1, m), device="cuda") for m, _ in shapes]))
    res = {}
    for mode, minjobs in (("stream", 1), ("per_launch", 10000)):
        ctx.set_option("stream_min_jobs", minjobs)
        with torch.cuda.stream(stream):
            for b in sets: b.run(ctx)
            stream.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                for b in sets: b.run(ctx)
            g.replay(); g.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(10): g.replay()
            e1.record(stream); stream.synchronize()
        res[mode] = e0.elapsed_time(e1) * 1e3 / (10 * NL)
    byt = sum(m * k * 4.5 / 8 for m, k in shapes)
    print(f"{name:10s} stream {res['stream']:7.2f} us  per-launch {res['per_launch']:7.2f} us   (HBM time {byt / 6.55e6:6.2f} us)")
