"""GPU tool: tcgen05 prefill GEMM, CTA pairs (cta_group::2, option tc_pair = 2) against single CTAs: same bits? time per shape (CUDA graph)."""
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
shapes = [(1024, 4096, 512), (4096, 4096, 512), (14336, 4096, 512), (4096, 14336, 512), (128256, 4096, 512), (4096, 4096, 100), (1000, 4096, 77)]
if len(sys.argv) > 1 and sys.argv[1] == "small":
    shapes = [(4096, 4096, 512), (1000, 4096, 77)]
for tname, t, make in (("q4_K", 12, bench.device_random_q4k), ("q8_0", 8, bench.device_random_q8_0)):
    for (m, k, n) in shapes:
        copies = 1 if m > 100000 else 3
        Ws = [make(torch, m, k, gen) for _ in range(copies)]
        X = torch.randn((n, k), device="cuda")
        out = {}
        line = f"{tname} {m}x{k}x{n}:"
        for pair in (0, 2):
            ctx.set_option("tc_pair", pair)
            Y = torch.empty((n, m), device="cuda")
            with torch.cuda.stream(stream):
                for i in range(3): pkg.mul_mat(ctx, t, Ws[0], X, out=Y)
                stream.synchronize()
                out[pair] = Y.clone()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=stream):
                    for i in range(12): pkg.mul_mat(ctx, t, Ws[i % copies], X, out=Y)
                g.replay()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream); g.replay(); e1.record(stream); stream.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / 12
            line += f"  pair={pair} {us:7.1f} us ({2.0 * m * k * n / us * 1e-6:5.0f} TF)"
        same = torch.equal(out[0], out[2])
        line += f"  same bits: {same}" + ("" if same else f"  max diff {float((out[0] - out[2]).abs().max()):.3e} of {float(out[0].abs().max()):.3e}")
        print(line, flush=True)
        del Ws
ctx.set_option("tc_pair", 1)
