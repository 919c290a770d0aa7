import subprocess, sys
sys.path.insert(0, '/root/repo')
from oracle import refenv
env = refenv.ref_env({"GGML_BACKEND_PATH": "/root/repo/llama.cpp-s390x_b200/libggml-b200.so"})
p = subprocess.run([str(refenv.REF_DIR / "test-backend-ops"), "test", "-b", "B2000", "-o", "MUL_MAT"], env=env, capture_output=True, text=True, timeout=300)
out = p.stdout + p.stderr
lines = out.splitlines()
bad = [l for l in lines if "FAIL" in l or "NMSE" in l or "nan" in l.lower()]
print(len(lines), len(bad)); print("\n".join(bad[:40]))
