"""GPU tool: isolate differences between the tensor-core Q6_K GEMV and the CUDA-core kernel: unit weights, one-hot activations."""
import sys
from pathlib import Path
import numpy as np
import torch
ROOT = Path(__file__).resolve().parents[1]; sys.path.insert(0, str(ROOT))
from __graft_entry__ import load_package
pkg = load_package()
ctx = pkg.Context(0, stream="torch")
Q6_K = 14
m, k = 16, 2048
W = np.zeros((m, k // 256, 210), np.uint8)
W[:, :, 0:128] = 0x11; W[:, :, 128:192] = 0xAA; W[:, :, 192:208] = 1        # q = 33 -> q - 32 = 1, scales 1
W[:, :, 208] = 0x00; W[:, :, 209] = 0x3C                                      # d = 1.0
Wd = torch.from_numpy(W.reshape(m, -1)).cuda()
res = np.zeros(k, np.float32); ref = np.zeros(k, np.float32)
for base in range(0, k, 8):
    X = np.zeros((8, k), np.float32)
    for c in range(8):
        X[c, base + c] = 1.0
    Xd = torch.from_numpy(X).cuda()
    ctx.set_option("use_mma", 1); a = pkg.mul_mat(ctx, Q6_K, Wd, Xd).cpu().numpy()
    ctx.set_option("use_mma", 0); b = pkg.mul_mat(ctx, Q6_K, Wd, Xd).cpu().numpy()
    res[base:base + 8] = a[:, 0]; ref[base:base + 8] = b[:, 0]
np.set_printoptions(linewidth=220, precision=2, suppress=True)
print("reference (first 64):", ref[:64])
for sb in range(2):
    print(f"super-block {sb}:")
    print(res[256 * sb: 256 * sb + 256].reshape(8, 32))
print("rows agree:", np.abs(a - a[:, :1]).max())
