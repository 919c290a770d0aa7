"""CPU tool: SASS opcode-class counts per kernel of the built kernels library -> profiles/r02_sass.txt (cuobjdump -sass, sm_100a)"""
import collections, re, subprocess, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
so = ROOT / "llama.cpp-s390x_b200" / "libggml_b200_kernels.so"
classes = ["UTCHMMA", "UTCIMMA", "UTCBAR", "STTM", "LDTM", "UTMALDG", "UBLKCP", "SYNCS", "IMMA", "HMMA", "IDP.4A", "LOP3", "HFMA2", "LDG", "STG", "LDS", "STS", "SHFL", "RED", "MEMBAR"]
fn = subprocess.run(["cuobjdump", "-sass", str(so)], capture_output=True, text=True).stdout
demangle = lambda s: subprocess.run(["c++filt", s], capture_output=True, text=True).stdout.strip()
out = ["SASS opcode-class counts per kernel of llama.cpp-s390x_b200/libggml_b200_kernels.so (cuobjdump -sass, sm_100a; round 2, final kernels)",
       "UTCHMMA = tcgen05.mma kind::f16, STTM / LDTM = tcgen05.st / tcgen05.ld, UTCBAR = tcgen05.commit, UTMALDG = TMA tensor load, UBLKCP = cp.async.bulk, "
       "SYNCS = mbarrier, IMMA = mma.sync s8/u8, IDP.4A = dp4a", ""]
cur, counts, n = None, collections.Counter(), 0
def flush():
    if cur:
        out.append(demangle(cur).replace("(anonymous namespace)::", ""))
        out.append("    instructions=%d " % n + " ".join(f"{k}={v}" for k, v in counts.items() if v))
for line in fn.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        flush(); cur, counts, n = m.group(1), collections.Counter(), 0
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        n += 1
        op = m.group(1)
        for c in classes:
            if op.startswith(c): counts[c] += 1
flush()
(ROOT / "profiles" / "r02_sass.txt").write_text("\n".join(out) + "\n")
print(len(out), "lines")
