#!/bin/bash
# GPU tool: default bench line + reference arm, summarised
OUT=gpurun_out/r2z; mkdir -p $OUT
S=$(date +%s); python bench.py "$@" > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$? $(( $(date +%s) - S )) s"
python - <<'PY'
import json
j=json.loads(open("gpurun_out/r2z/bench.json").read().strip().splitlines()[-1])
print({k:j.get(k) for k in ("metric","value","unit","ms_per_step","gpu_launches","dtype","scaling")})
print("roofline", j["roofline"]["frac"], j["roofline"]["traffic"])
if "e2e" in j: print("e2e", {k:j["e2e"].get(k) for k in ("value","ms_per_step","h2d_bytes_per_step","d2h_bytes_per_step")})
print("cpu", j.get("cpu_baseline"))
print("whole_graph", json.dumps(j.get("whole_graph"))[:1000])
print("dep", json.dumps(j.get("dependent_chain"))[:500])
if "prefill" in j: print("prefill", j["prefill"]["q4_K"]["tflops"], j["prefill"]["q8_0"]["tflops"])
if "decode_batch" in j: print("decode_batch", {k:v["ms_per_step"] for k,v in j["decode_batch"].items() if isinstance(v,dict)})
if "moe" in j: print("moe", {k:v["us"] for k,v in j["moe"].items() if isinstance(v,dict)})
print("clocks", j.get("clocks"))
PY
