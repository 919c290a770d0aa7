#!/bin/bash
# GPU tool: oracle/_ref/llama_step (a whole Llama decoder graph on the plug-in against the reference's CPU backend): tiny parity cases, then
# Llama-3-8B decode (1 token, 512 past) and prefill (512 tokens)
export LD_LIBRARY_PATH=$PWD/oracle/_ref/native:$PWD/oracle/_ref/portable:$PWD/oracle/_ref:$LD_LIBRARY_PATH GGML_BACKEND_PATH=$PWD/llama.cpp-s390x_b200/libggml-b200.so GGML_B200_STATS=1
for args in "tiny 1 40 2 1" "tiny 7 33 2 1" "tiny 64 0 2 1" "$@"; do echo "== $args"; timeout 600 oracle/_ref/llama_step $args 2>&1 | grep -v load_backend | tail -4; done
