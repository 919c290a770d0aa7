"""GPU tool: BASELINE config 4 — Mixtral-8x7B Q4_K MUL_MAT_ID (top-2 of 8 experts): ffn_gate/up [4096 -> 14336] and ffn_down
[14336 -> 4096] at 1..8 and 512 tokens.  Small batches: GB/s over the bytes of the experts actually touched; 512: TFLOP/s."""
import sys
from pathlib import Path
import numpy as np
import torch
ROOT = Path(__file__).resolve().parents[1]; sys.path.insert(0, str(ROOT))
from __graft_entry__ import load_package
from tests.util import random_blocks
pkg = load_package()
stream = torch.cuda.Stream()
ctx = pkg.Context(0, stream=stream.cuda_stream)
rng = np.random.default_rng(99)
Q4_K, NE, NU = 12, 8, 2
for m, k in ((14336, 4096), (4096, 14336)):
    base = torch.from_numpy(random_blocks(Q4_K, 256, k, rng)).cuda()
    AS = base.repeat(NE * m // 256, 1).reshape(NE, m, -1).contiguous()
    copies = [AS, AS.clone()]                                  # 2 x 264 MB > L2
    for n_tokens in (1, 2, 4, 8, 512):
        ids = torch.from_numpy(np.stack([rng.permutation(NE)[:NU] for _ in range(n_tokens)]).astype(np.int32)).cuda()
        B = torch.randn((n_tokens, 1, k), device="cuda")
        out = torch.empty((n_tokens, NU, m), device="cuda")
        res = {}
        for mode in ("default", "per-pair") if 1 < n_tokens <= 8 else ("default",):
            ctx.set_option("use_mma", 0 if mode == "per-pair" else 1)
            iters = 8
            with torch.cuda.stream(stream):
                pkg.mul_mat_id(ctx, Q4_K, copies[0], B, ids, out=out)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=stream):
                    for i in range(iters):
                        pkg.mul_mat_id(ctx, Q4_K, copies[i % 2], B, ids, out=out)
                g.replay(); stream.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream); g.replay(); g.replay(); e1.record(stream); stream.synchronize()
            res[mode] = e0.elapsed_time(e1) * 1e3 / (2 * iters)
        ctx.set_option("use_mma", 1)
        touched = len(np.unique(ids.cpu().numpy()))
        wbytes = touched * m * k * 4.5 / 8
        us = res["default"]
        extra = f" | per-pair GEMV: {res['per-pair']:.1f} us" if "per-pair" in res else ""
        print(f"as[{k},{m},8] tokens={n_tokens:3d}: {us:8.1f} us  experts touched {touched}  {wbytes / us / 1e3:6.0f} GB/s (touched expert bytes)  "
              f"{2.0 * m * k * n_tokens * NU / us / 1e6:8.1f} TFLOP/s{extra}", flush=True)
