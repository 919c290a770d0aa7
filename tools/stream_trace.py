"""Coordinator time stamps of the persistent job-table GEMV on the Llama-3-8B step wired as a dependent chain (debugging aid).
usage: python tools/stream_trace.py [independent|chain] [stream_ll=0|1]"""
import ctypes as C
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
from __graft_entry__ import load_package
import bench

pkg = load_package()
mode = sys.argv[1] if len(sys.argv) > 1 else "chain"
ll = int(sys.argv[2]) if len(sys.argv) > 2 else 1
div = int(sys.argv[3]) if len(sys.argv) > 3 else 1
stream = torch.cuda.Stream()
ctx = pkg.Context(0, stream=stream.cuda_stream)
ctx.set_option("stream_ll", ll)
ctx.set_option("stream_chain", 1); ctx.set_option("stream_min_jobs", 1)
gen = torch.Generator(device="cuda"); gen.manual_seed(1)
NL = 8
T = {n: (m, k) for n, m, k, _ in bench.LLAMA3_8B}
hid, ffn = 4096, 14336
h0 = torch.randn((1, hid), device="cuda"); q = torch.empty((1, 4096), device="cuda"); h1 = torch.empty((1, 4096), device="cuda")
kk = torch.empty((1, 1024), device="cuda"); v = torch.empty((1, 1024), device="cuda")
g = torch.empty((1, ffn), device="cuda"); u = torch.empty((1, ffn), device="cuda")
wire = {"wq": (h0, q), "wk": (h0, kk), "wv": (h0, v), "wo": (q, h1), "ffn_gate": (h1, g), "ffn_up": (h1, u), "ffn_down": (g, h0)}
fixed = {4096: torch.randn((1, 4096), device="cuda"), 14336: torch.randn((1, 14336), device="cuda")}
ws, xs, ys, names = [], [], [], []
for l in range(NL):
    for n in ("wq", "wk", "wv", "wo", "ffn_gate", "ffn_up", "ffn_down"):
        m, k = T[n]
        m //= div
        ws.append(bench.device_random_q4k(torch, m, k, gen))
        if mode == "chain":
            xs.append(wire[n][0]); ys.append(wire[n][1])
        else:
            xs.append(fixed[k]); ys.append(torch.empty((1, m), device="cuda"))
        names.append(n)
batch = pkg.MatmulBatch(12, ws, xs, ys)
with torch.cuda.stream(stream):
    for _ in range(3):
        batch.run(ctx)
    stream.synchronize()
    ctx.set_option("stream_trace", 1)
    batch = pkg.MatmulBatch(12, ws, xs, ys)
    batch.run(ctx)
    stream.synchronize()
n_groups = 4 * NL
MAXJ = 448
buf = np.zeros(2 * MAXJ * 4, np.uint64)
rc = pkg.lib().b200_dbg_stream_trace(ctx.handle, buf.ctypes.data_as(C.c_void_p), buf.size)
assert rc == 0, rc
tr = buf.reshape(2, MAXJ, 4).astype(np.int64)
t0 = tr[0, 0, 0]
gn = ["qkv", "wo", "gate_up", "down"]
print(f"mode={mode} stream_ll={ll}  (ns; CTA 0 | last CTA)  group: coordinator start, +flags seen, +copy issued, +image in shared memory   group-to-group")
for gi in range(n_groups):
    a, b = tr[0, gi], tr[1, gi]
    nxt = tr[0, gi + 1, 3] - a[3] if gi + 1 < n_groups else 0
    print(f"{gi:3d} {gn[gi % 4]:8s} {a[0]-t0:8d} flags+{a[1]-a[0]:6d} issue+{a[2]-a[1]:6d} landed+{a[3]-a[2]:6d} | last CTA flags at {b[1]-t0:8d} landed+{b[3]-b[2]:6d} | next group's x ready after {nxt:6d}")

