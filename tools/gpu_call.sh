set -x
test llama.cpp-q_4_0_b200/lib/libggml_b200.so -nt llama.cpp-q_4_0_b200/csrc/ggml-b200.cu || { echo STALE LIB; exit 9; }
mkdir -p gpurun_out
tools/ubench/fma_rate > gpurun_out/r02_ubench_fma_rate.log 2>&1; tail -4 gpurun_out/r02_ubench_fma_rate.log
timeout 300 tools/ubench/epi_rate > gpurun_out/r02_ubench_epi_rate.log 2>&1; cat gpurun_out/r02_ubench_epi_rate.log
timeout 900 python -m pytest tests/test_glue.py -x -q -m gpu > gpurun_out/r02_pytest_glue.log 2>&1; tail -15 gpurun_out/r02_pytest_glue.log
timeout 1200 python bench.py --steps 20 --warmup 3 --no-extra --no-prefill > gpurun_out/r02_bench_n1_b.json 2> gpurun_out/r02_bench_n1_b.err; tail -c 600 gpurun_out/r02_bench_n1_b.err; python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_bench_n1_b.json"))
print(d["value"], json.dumps(d["e2e"])[:1800])
PY
