#!/bin/bash
# round-2 profiles, call D: launch list of ONE Llama-3-8B decode step through the plug-in with every op on the device (oracle/_ref/llama_step)
set -u
OUT=gpurun_out/r2u; mkdir -p $OUT
export LD_LIBRARY_PATH=$PWD/oracle/_ref/native:$PWD/oracle/_ref/portable:$PWD/oracle/_ref:${LD_LIBRARY_PATH:-} GGML_BACKEND_PATH=$PWD/llama.cpp-s390x_b200/libggml-b200.so GGML_B200_GRAPHS=0
CMD="oracle/_ref/llama_step 8b 1 512 1 0"
$CMD > $OUT/plain_d.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 1100 -c 560 --csv --log-file $OUT/r02_launches_llama_step.csv $CMD > $OUT/ncu_d.log 2>&1
echo "rc=$?"; tail -2 $OUT/plain_d.log | cut -c1-300
