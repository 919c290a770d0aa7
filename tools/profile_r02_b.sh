#!/bin/bash
# round-2 profiles, call B: --set full capture of the persistent job-table GEMV in the bench step (run under gpurun)
set -u
OUT=gpurun_out/r2p; mkdir -p $OUT
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-extras --no-prefill --cpu-seconds 0"
$CMD > $OUT/plain_b.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gemv_q4k_stream -s 3 -c 1 -o $OUT/r02_gemv_q4k_stream $CMD > $OUT/ncu_b.log 2>&1
echo "ncu full rc=$?"; tail -3 $OUT/ncu_b.log; ls -la $OUT/*.ncu-rep
