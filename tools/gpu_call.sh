#!/bin/bash
mkdir -p gpurun_out
timeout 1700 python tools/llama7b_decode_e2e.py > gpurun_out/r02_llama7b_decode_e2e.log 2>&1; echo "rc=$?" >> gpurun_out/r02_llama7b_decode_e2e.log
grep -v "^  step" gpurun_out/r02_llama7b_decode_e2e.log
