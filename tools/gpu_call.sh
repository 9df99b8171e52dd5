#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/r02_pytest_multi_gpu_n2_b.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_multi_gpu_n2_b.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > gpurun_out/r02_bench_n2_d.json 2> gpurun_out/r02_bench_n2_d.err; echo "bench rc=$?" >> gpurun_out/r02_bench_n2_d.err
tail -3 gpurun_out/r02_pytest_multi_gpu_n2_b.log; tail -2 gpurun_out/r02_bench_n2_d.err; head -c 600 gpurun_out/r02_bench_n2_d.json
