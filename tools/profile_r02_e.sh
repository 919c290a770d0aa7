#!/bin/bash
# round-2 profiles (final state of the kernels), call E: launch list of the bench command with DRAM bytes per launch, and the reference's own
# perf / support harness against the plug-in and the CPU backend (run under gpurun)
set -u
OUT=gpurun_out/r2p; mkdir -p $OUT
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-extras --no-prefill --cpu-seconds 0"
$CMD > $OUT/plain_e.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"gemv|mmq|quantize|tc_act|mm_ids|unstage|ll_unpack" -c 400 --csv --log-file $OUT/r02_launches_gemv.csv $CMD > $OUT/ncu_e.log 2>&1
echo "ncu launch list rc=$?"; wc -l $OUT/r02_launches_gemv.csv
export LD_LIBRARY_PATH=$PWD/oracle/_ref/native:$PWD/oracle/_ref/portable:$PWD/oracle/_ref GGML_BACKEND_PATH=$PWD/llama.cpp-s390x_b200/libggml-b200.so
T=oracle/_ref/test-backend-ops
$T perf -b B2000 -o MUL_MAT -p "type_a=(q4_0|q8_0|q4_K|q6_K),type_b=f32.*k=14336" 2>&1 | sed "s/\x1b\[[0-9;]*m//g" > $OUT/r02_tbo_perf_b200.txt; echo "tbo perf B200 rc=$?"
$T perf -b CPU -o MUL_MAT -p "type_a=(q4_0|q8_0|q4_K|q6_K),type_b=f32.*k=14336" 2>&1 | sed "s/\x1b\[[0-9;]*m//g" > $OUT/r02_tbo_perf_cpu.txt; echo "tbo perf CPU rc=$?"
$T support -b B2000 --output csv > $OUT/r02_tbo_support_all_ops.csv 2>$OUT/support.err; echo "support all ops rc=$?"
$T test -b B2000 2>&1 | sed "s/\x1b\[[0-9;]*m//g" | grep -v "not supported" | tail -3 > $OUT/r02_tbo_test_all_tail.txt
grep -c ",yes," $OUT/r02_tbo_support_all_ops.csv; tail -3 $OUT/r02_tbo_perf_b200.txt; cat $OUT/r02_tbo_test_all_tail.txt
