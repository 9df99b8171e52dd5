#!/bin/bash
mkdir -p gpurun_out
N=${NGPU:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N > gpurun_out/r02_bench_n${N}_e.json 2> gpurun_out/r02_bench_n${N}_e.err; echo "bench rc=$?" >> gpurun_out/r02_bench_n${N}_e.err
tail -3 gpurun_out/r02_bench_n${N}_e.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_n${N}_e.json').read().strip().splitlines()[-1])
print('value', d['value'], d['ms_per_step']); e=d['e2e']; print('e2e', e['value'], e['ms_per_step'], e['per_matmul_host_call']['ms_per_step']); print('prefill', d['prefill']['value'])
PY
