#!/bin/bash
# round-2 profiles, call C: launch list (our kernels only) of the bench command (run under gpurun)
set -u
OUT=gpurun_out/r2p; mkdir -p $OUT
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-extras --no-prefill --cpu-seconds 0"
$CMD > $OUT/plain_c.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"gemv|mmq|quantize|tc_act|mm_ids|unstage|ll_unpack" -c 400 --csv --log-file $OUT/r02_launches_gemv.csv $CMD > $OUT/ncu_c.log 2>&1
echo "ncu launch list rc=$?"; wc -l $OUT/r02_launches_gemv.csv
