#!/bin/bash
mkdir -p gpurun_out
N=8
timeout 400 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/r02_pytest_multi_gpu_n8_c.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_multi_gpu_n8_c.log
tail -3 gpurun_out/r02_pytest_multi_gpu_n8_c.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N > gpurun_out/r02_bench_n${N}_g.json 2> gpurun_out/r02_bench_n${N}_g.err; echo "bench rc=$?" >> gpurun_out/r02_bench_n${N}_g.err
tail -1 gpurun_out/r02_bench_n${N}_g.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_n${N}_g.json').read().strip().splitlines()[-1])
p=d['prefill']; print('value', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value']); print('prefill', p['value'], p['ms_per_pass'], p['kernels_only']['ms_per_pass'], p['gather']['ms_per_pass'])
PY
