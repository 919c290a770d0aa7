#!/bin/bash
# round-2 profiles, call F: launch list of ONE 512-token Llama-3-8B prompt graph through the plug-in (oracle/_ref/llama_step), eager launches
set -u
OUT=gpurun_out/r2v; mkdir -p $OUT
export LD_LIBRARY_PATH=$PWD/oracle/_ref/native:$PWD/oracle/_ref/portable:$PWD/oracle/_ref:${LD_LIBRARY_PATH:-} GGML_BACKEND_PATH=$PWD/llama.cpp-s390x_b200/libggml-b200.so GGML_B200_GRAPHS=0
CMD="oracle/_ref/llama_step 8b 512 0 1 0"
$CMD > $OUT/plain_f.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 1700 -c 900 --csv --log-file $OUT/r02_launches_llama_prompt.csv $CMD > $OUT/ncu_f.log 2>&1
echo "rc=$?"; tail -2 $OUT/plain_f.log | cut -c1-300
