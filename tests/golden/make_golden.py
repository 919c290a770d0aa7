"""Generate tests/golden/*.npz from the UNMODIFIED reference compiled into oracle/_ref (oracle/Makefile).

Run here (the container that has /root/reference):   python tests/golden/make_golden.py
The fixtures are committed; tests never need /root/reference or oracle/_ref to check against them.

  q8_1.npz      x f32 [rows, k]  ->  quantize_row_q8_1_ref bytes (ggml-quants.c:260-293), incl. ties, zeros, tiny/huge values
  <type>.npz    W blocks from ggml_quantize_chunk (ggml.h:2764), dequantized rows (to_float), X f32,
                Y = MUL_MAT on the reference CPU backend with use_ref (ggml-cpu.c:1241-1439), batch/broadcast variant,
                generic vec_dot values (ggml-cpu/quants.c) on the CPU-side activation format
  mmid.npz      MUL_MAT_ID (ggml-cpu.c:1521-1694) for each type incl. a strided ids view and b broadcast
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import refenv  # noqa: E402

OUT = Path(__file__).resolve().parent


def q8_1_inputs(rng):
    k = 256
    rows = []
    rows.append(rng.standard_normal((24, k)).astype(np.float32))
    rows.append((rng.standard_normal((8, k)) * 1e-3).astype(np.float32))
    rows.append((rng.standard_normal((8, k)) * 1e4).astype(np.float32))
    rows.append(rng.uniform(-1, 1, (8, k)).astype(np.float32))
    z = np.zeros((4, k), np.float32)
    z[1, 5] = 1.0          # single spike
    z[2, :] = -3.25        # constant block
    z[3, ::2] = 1e-30      # tiny but normal
    rows.append(z)
    # exact ties: amax = 127 -> d = 1, id = 1; values n + 0.5 must round away from zero
    t = np.zeros((4, k), np.float32)
    for b in range(k // 32):
        blk = (rng.integers(-126, 126, 32) + 0.5).astype(np.float32)
        blk[0] = 127.0
        t[0, b * 32:(b + 1) * 32] = blk
        t[1, b * 32:(b + 1) * 32] = -blk
        blk2 = (rng.integers(-62, 62, 32) + 0.5).astype(np.float32) * 2.0   # amax 254 -> d = 2, exact halves again
        blk2[3] = -254.0
        t[2, b * 32:(b + 1) * 32] = blk2
        t[3, b * 32:(b + 1) * 32] = blk * 0.5                               # amax 63.5 -> d = 0.5
    rows.append(t)
    # cos pattern of tests/test-quantize-fns.cpp:35-39
    i = np.arange(k * 4, dtype=np.float32)
    rows.append((0.1 + 2 * np.cos(i)).astype(np.float32).reshape(4, k))
    return np.concatenate(rows, 0)


def main():
    R = refenv.load_ref()
    if R is None:
        raise SystemExit("oracle/_ref is not built: make -C oracle")
    rng = np.random.default_rng(20261104)

    x = q8_1_inputs(rng)
    nb = x.shape[1] // 32
    y = np.zeros((x.shape[0], nb * 36), np.uint8)
    for r in range(x.shape[0]):
        assert R.ref_quantize_row_ref(refenv.Q8_1, refenv.ptr(x[r]), refenv.ptr(y[r]), x.shape[1]) == 0
    np.savez_compressed(OUT / "q8_1.npz", x=x, y=y)

    mmid = {}
    for name, t in refenv.WEIGHT_TYPES.items():
        m, k, n = 48, 512, 5
        wf = (rng.standard_normal((m, k)) * 0.05).astype(np.float32)
        W = np.zeros((m, refenv.row_size(t, k)), np.uint8)
        R.ref_quantize_chunk(t, refenv.ptr(wf), refenv.ptr(W), m, k)
        deq = np.zeros((m, k), np.float32)
        for r in range(m):
            R.ref_dequantize_row(t, refenv.ptr(W[r]), refenv.ptr(deq[r]), k)
        X = rng.uniform(-1, 1, (n, k)).astype(np.float32)
        Y = np.zeros((n, m), np.float32)
        assert R.ref_mul_mat(t, refenv.ptr(W), refenv.ptr(X), refenv.ptr(Y), m, n, k, 1, 1, 1, 1, 4) == 0
        # batch + broadcast: W [k, m, 2, 1], X [k, n, 4, 3]  (r2 = 2, r3 = 3)
        Wb = np.zeros((1, 2, m, refenv.row_size(t, k)), np.uint8)
        wfb = (rng.standard_normal((2, m, k)) * 0.05).astype(np.float32)
        R.ref_quantize_chunk(t, refenv.ptr(wfb), refenv.ptr(Wb), 2 * m, k)
        Xb = rng.uniform(-1, 1, (3, 4, 2, k)).astype(np.float32)
        Yb = np.zeros((3, 4, 2, m), np.float32)
        assert R.ref_mul_mat(t, refenv.ptr(Wb), refenv.ptr(Xb), refenv.ptr(Yb), m, 2, k, 2, 1, 4, 3, 4) == 0
        # generic vec_dot on the CPU activation format (Q8_0 / Q8_K)
        at = {refenv.Q4_0: refenv.Q8_0, refenv.Q8_0: refenv.Q8_0, refenv.Q4_K: refenv.Q8_K, refenv.Q6_K: refenv.Q8_K}[t]
        act = np.zeros((refenv.row_size(at, k),), np.uint8)
        assert R.ref_quantize_row_ref(at, refenv.ptr(X[0]), refenv.ptr(act), k) == 0
        vd = np.array([R.ref_vec_dot_generic(t, k, refenv.ptr(W[r]), refenv.ptr(act)) for r in range(m)], np.float32)
        np.savez_compressed(OUT / f"{name}.npz", W=W, deq=deq, X=X, Y=Y, Wb=Wb, Xb=Xb, Yb=Yb, act=act, vec_dot=vd)

        # MUL_MAT_ID: 4 experts, top-2, 6 tokens; ids is a view into a wider [n_tokens, n_expert] tensor (test-backend-ops.cpp:4099-4118)
        ne, nu, nt, mm = 4, 2, 6, 32
        asf = (rng.standard_normal((ne, mm, k)) * 0.05).astype(np.float32)
        AS = np.zeros((ne, mm, refenv.row_size(t, k)), np.uint8)
        R.ref_quantize_chunk(t, refenv.ptr(asf), refenv.ptr(AS), ne * mm, k)
        full = np.stack([rng.permutation(ne) for _ in range(nt)]).astype(np.int32)
        ids = np.ascontiguousarray(full[:, :nu])
        for b_ne1 in (nu, 1):
            B = rng.uniform(-1, 1, (nt, b_ne1, k)).astype(np.float32)
            Cc = np.zeros((nt, nu, mm), np.float32)
            assert R.ref_mul_mat_id(t, refenv.ptr(AS), refenv.ptr(B), refenv.ptr(ids), refenv.ptr(Cc), mm, k, ne, nu, nt, b_ne1, 4) == 0
            mmid[f"{name}_B{b_ne1}"] = B
            mmid[f"{name}_C{b_ne1}"] = Cc
        mmid[f"{name}_AS"] = AS
        mmid[f"{name}_ids_full"] = full
    np.savez_compressed(OUT / "mmid.npz", **mmid)
    for f in sorted(OUT.glob("*.npz")):
        print(f.name, f.stat().st_size)


if __name__ == "__main__":
    main()
