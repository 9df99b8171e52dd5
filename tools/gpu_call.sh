#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r02_bench_n1_e.json 2> gpurun_out/r02_bench_n1_e.err; echo "bench rc=$?" >> gpurun_out/r02_bench_n1_e.err
tail -4 gpurun_out/r02_bench_n1_e.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_n1_e.json').read().strip().splitlines()[-1])
print('value', d['value'], d['ms_per_step']); e=d['e2e']; print(e['value'], e['ms_per_step']); print(json.dumps(e.get('list_api'))[:900])
PY
