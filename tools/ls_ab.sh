#!/bin/bash
# GPU tool: the Llama-3-8B decode graph (and the 512-token prompt graph) on the plug-in under the shell's switches, same box, back to back
export LD_LIBRARY_PATH=$PWD/oracle/_ref/native:$PWD/oracle/_ref/portable:$PWD/oracle/_ref:$LD_LIBRARY_PATH GGML_BACKEND_PATH=$PWD/llama.cpp-s390x_b200/libggml-b200.so
for rep in 1; do
for cfg in "GGML_B200_GRAPHS=1" "GGML_B200_GRAPHS=0" "GGML_B200_GRAPHS=2" ; do
  echo "== $cfg: decode $(env $cfg oracle/_ref/llama_step 8b 1 512 40 0 2>&1 | grep -o '"ms_per_step_b200": [0-9.]*')  prompt $(env $cfg oracle/_ref/llama_step 8b 512 0 5 0 2>&1 | grep -o '"ms_per_step_b200": [0-9.]*')"
done; done
GGML_B200_STATS=1 oracle/_ref/llama_step 8b 1 512 20 0 2>&1 | grep "stats"
