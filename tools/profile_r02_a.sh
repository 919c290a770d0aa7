#!/bin/bash
# round-2 profiles, call A: launch list of the bench command + the reference's own perf / support harness (run under gpurun)
set -u
OUT=gpurun_out/r2p; mkdir -p $OUT
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-extras --no-prefill --cpu-seconds 0"
$CMD > $OUT/plain_a.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/r02_launches_gemv.csv $CMD > $OUT/ncu_a.log 2>&1
echo "ncu launch list rc=$?"
export LD_LIBRARY_PATH=$PWD/oracle/_ref/portable:$PWD/oracle/_ref GGML_BACKEND_PATH=$PWD/llama.cpp-s390x_b200/libggml-b200.so
T=oracle/_ref/test-backend-ops
$T perf -b B2000 -o MUL_MAT -p "type_a=(q4_0|q8_0|q4_K|q6_K),type_b=f32.*k=14336" 2>&1 | sed "s/\x1b\[[0-9;]*m//g" > $OUT/r02_tbo_perf_b200.txt; echo "tbo perf B200 rc=$?"
$T perf -b CPU -o MUL_MAT -p "type_a=(q4_0|q8_0|q4_K|q6_K),type_b=f32.*k=14336" 2>&1 | sed "s/\x1b\[[0-9;]*m//g" > $OUT/r02_tbo_perf_cpu.txt; echo "tbo perf CPU rc=$?"
$T support -b B2000 -o MUL_MAT --output csv > $OUT/r02_tbo_support_mul_mat.csv 2>$OUT/support.err; echo "support MUL_MAT rc=$?"
$T support -b B2000 -o MUL_MAT_ID --output csv > $OUT/r02_tbo_support_mul_mat_id.csv 2>>$OUT/support.err; echo "support MUL_MAT_ID rc=$?"
tail -3 $OUT/r02_tbo_perf_b200.txt; wc -l $OUT/*.csv
