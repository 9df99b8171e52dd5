#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r02_bench_n1_d.json 2> gpurun_out/r02_bench_n1_d.err; echo "bench rc=$?" >> gpurun_out/r02_bench_n1_d.err
tail -3 gpurun_out/r02_bench_n1_d.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_n1_d.json').read().strip().splitlines()[-1])
print(json.dumps(d['configs']['siblings']['q8_0'].get('prefill_n512'), indent=1))
print('e2e', d['e2e']['ms_per_step'], d['e2e']['value'], 'value', d['value'])
print('prefill', d['prefill']['value'], d['prefill']['roofline']['achieved'], d['prefill']['roofline']['frac'])
PY
