test llama.cpp-q_4_0_b200/lib/libggml_b200.so -nt llama.cpp-q_4_0_b200/csrc/siblings.cu || { echo STALE LIB; exit 9; }
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_siblings.py -x -q -m gpu > gpurun_out/r02_pytest_siblings.log 2>&1; tail -2 gpurun_out/r02_pytest_siblings.log
for c in 1 2; do
GGML_B200_SIB_CTAS=$c timeout 600 python - > gpurun_out/r02_bench_siblings_c$c.log 2>&1 <<'PY'
import json, numpy as np, torch, bench
mod = bench.load_binding(); mod.init(0)
stream = torch.cuda.Stream(); mod.set_stream(stream.cuda_stream)
print(json.dumps(bench.siblings_leg(mod, stream, np.random.default_rng(1), 20, 3), indent=1))
PY
echo "== ctas $c"; grep "GBps\|us_per\|ms_per\|q8_0\|q4_1\|Error" gpurun_out/r02_bench_siblings_c$c.log
done
