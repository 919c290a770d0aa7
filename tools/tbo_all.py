"""GPU tool: the reference's unmodified test-backend-ops against the plug-in — `test` over ALL ops (what supports_op() declines is skipped by the
harness), or the arguments given.  Prints the summary, every FAIL line, and per op how many cases ran."""
import collections, re, subprocess, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]; sys.path.insert(0, str(ROOT))
from oracle import refenv
args = sys.argv[1:] or ["test", "-b", "B2000"]
env = refenv.ref_env({"GGML_BACKEND_PATH": str(ROOT / "llama.cpp-s390x_b200" / "libggml-b200.so")})
p = subprocess.run([str(refenv.REF_DIR / "test-backend-ops"), *args], env=env, capture_output=True, text=True, timeout=1500)
out = re.sub(r"\x1b\[[0-9;]*m", "", p.stdout + p.stderr)
Path(ROOT / "gpurun_out").mkdir(exist_ok=True)
(ROOT / "gpurun_out" / "tbo_last.txt").write_text(out)
ran = collections.Counter(); bad = []
for l in out.splitlines():
    m = re.match(r"\s*([A-Z_0-9]+)\(", l)
    if m and l.rstrip().endswith("OK"): ran[m.group(1)] += 1
    if "FAIL" in l or "error" in l.lower(): bad.append(l)
print("rc", p.returncode)
print("\n".join(out.splitlines()[-4:]))
print("ran:", dict(ran))
print("bad lines:", len(bad))
print("\n".join(bad[:40]))
