"""GPU tool: unmodified test-backend-ops, MUL_MAT / MUL_MAT_ID / GET_ROWS / MUL_MAT_VEC_FUSION: cases run per weight type"""
import collections, re, subprocess, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]; sys.path.insert(0, str(ROOT))
from oracle import refenv
env = refenv.ref_env({"GGML_BACKEND_PATH": str(ROOT / "llama.cpp-s390x_b200" / "libggml-b200.so"), "GGML_B200_FUSE": "2"})
for op in ("MUL_MAT", "MUL_MAT_ID", "GET_ROWS", "MUL_MAT_VEC_FUSION"):
    p = subprocess.run([str(refenv.REF_DIR / "test-backend-ops"), "test", "-b", "B2000", "-o", op], env=env, capture_output=True, text=True, timeout=1500)
    out = re.sub(r"\x1b\[[0-9;]*m", "", p.stdout + p.stderr)
    ran = collections.Counter(); fails = []
    for l in out.splitlines():
        m = re.search(r"type(?:_a)?=(\w+)", l)
        if m and l.rstrip().endswith("OK"): ran[m.group(1)] += 1
        if "FAIL" in l: fails.append(l)
    print(op, "rc", p.returncode, dict(ran), "FAILS", len(fails))
    print("\n".join(fails[:10]))
