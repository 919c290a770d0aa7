"""prefill GEMM (n = 512 and MoE): per-shape time under the tile-width / split-K options (CUDA graph, 3 rotating weight copies)"""
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
n = 512
quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
shapes = ((1024, 4096), (4096, 4096), (14336, 4096), (4096, 14336), (128256, 4096))
for tname, t, make in (("q4_K", 12, bench.device_random_q4k), ("q8_0", 8, bench.device_random_q8_0)):
    for (m, k) in shapes:
        copies = 1 if m > 100000 else 3
        Ws = [make(torch, m, k, gen) for _ in range(copies)]
        X = torch.randn((n, k), device="cuda"); Y = torch.empty((n, m), device="cuda")
        line = f"{tname} {m}x{k}:"
        for nt in ((0,) if quick else (0, 64, 128, 192)):
            for sk in ((1,) if quick else (1, 2, 0)):
                ctx.set_option("tc_nt", nt); ctx.set_option("tc_split_k", sk)
                with torch.cuda.stream(stream):
                    for i in range(3): pkg.mul_mat(ctx, t, Ws[i % copies], X, out=Y)
                    stream.synchronize()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=stream):
                        for i in range(12): pkg.mul_mat(ctx, t, Ws[i % copies], X, out=Y)
                    g.replay()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(stream); g.replay(); e1.record(stream); stream.synchronize()
                us = e0.elapsed_time(e1) * 1e3 / 12
                line += f"  nt{nt}/sk{sk} {us:6.1f} ({2.0 * m * k * n / us * 1e-6:5.0f} TF)"
        print(line, flush=True)
        del Ws
ctx.set_option("tc_nt", 0); ctx.set_option("tc_split_k", 1)
