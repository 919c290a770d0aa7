"""GPU tool: prefill MUL_MAT (n = 512) per weight type — the types with a tcgen05 converter of their own, and the others through the
fp16-scratch route (tc_dequant = 1) against the column tiles of the generic GEMV (tc_dequant = 0).  Weights quantized by the reference."""
import sys
from pathlib import Path
import numpy as np
import torch
ROOT = Path(__file__).resolve().parents[1]; sys.path.insert(0, str(ROOT))
from __graft_entry__ import load_package
from oracle import refenv
pkg = load_package()
ctx = pkg.Context(0, stream="torch")
ref = refenv.load_ref()
TYPES = {"q4_0": (2, 32, 18), "q4_1": (3, 32, 20), "q5_0": (6, 32, 22), "q5_1": (7, 32, 24), "q8_0": (8, 32, 34), "q2_K": (10, 256, 84), "q3_K": (11, 256, 110),
         "q4_K": (12, 256, 144), "q5_K": (13, 256, 176), "q6_K": (14, 256, 210), "iq4_nl": (20, 32, 18), "iq4_xs": (23, 256, 136), "mxfp4": (39, 32, 17),
         "nvfp4": (40, 64, 36)}
m, k, n = 4096, 4096, 512
rng = np.random.default_rng(0)
wf = (rng.standard_normal((256, k)) * 0.05).astype(np.float32)
X = torch.randn((n, k), device="cuda")
Y = torch.empty((n, m), device="cuda")


def timed(t, W, reps=10):
    for _ in range(2):
        pkg.mul_mat(ctx, t, W, X, out=Y)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        pkg.mul_mat(ctx, t, W, X, out=Y)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


for name, (t, bs, sz) in TYPES.items():
    Wb = np.zeros((256, k // bs * sz), np.uint8)
    ref.ref_quantize_chunk(t, refenv.ptr(wf), refenv.ptr(Wb), 256, k)
    W = torch.from_numpy(Wb).cuda().repeat(m // 256, 1).contiguous()
    ctx.set_option("tc_dequant", 1)
    a = timed(t, W)
    ctx.set_option("tc_dequant", 0)
    b = timed(t, W, reps=3)
    ctx.set_option("tc_dequant", 1)
    print(f"{name:7s} {m}x{k}x{n}: {a:8.1f} us = {2.0 * m * k * n / a / 1e6:7.1f} TF   (tc_dequant=0: {b:8.1f} us)", flush=True)
