#!/bin/bash
mkdir -p gpurun_out
N=4
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N > gpurun_out/r02_bench_n${N}_f.json 2> gpurun_out/r02_bench_n${N}_f.err; echo "bench rc=$?" >> gpurun_out/r02_bench_n${N}_f.err
tail -2 gpurun_out/r02_bench_n${N}_f.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_n${N}_f.json').read().strip().splitlines()[-1])
p=d['prefill']; print('value', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value']); print('prefill', p['value'], p['ms_per_pass'], p['kernels_only'], p['gather'])
PY
