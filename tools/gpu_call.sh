#!/bin/bash
# scratch: one GPU call
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_e.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_gpu_e.log
timeout 900 python bench.py > gpurun_out/r02_bench_n1_c.json 2> gpurun_out/r02_bench_n1_c.err; echo "bench rc=$?" >> gpurun_out/r02_bench_n1_c.err
tail -3 gpurun_out/r02_pytest_gpu_e.log; tail -3 gpurun_out/r02_bench_n1_c.err; head -c 3000 gpurun_out/r02_bench_n1_c.json
