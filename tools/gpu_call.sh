test llama.cpp-q_4_0_b200/lib/libggml_b200.so -nt llama.cpp-q_4_0_b200/csrc/siblings.cu || { echo STALE LIB; exit 9; }
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_siblings.py tests/test_glue.py -x -q -m gpu -k "sibling or siblings or block or edge or upload or reference_itself" > gpurun_out/r02_pytest_siblings_q5.log 2>&1; tail -3 gpurun_out/r02_pytest_siblings_q5.log
timeout 900 python - > gpurun_out/r02_bench_siblings_q5.log 2>&1 <<'PY'
import json, numpy as np, torch, bench
mod = bench.load_binding(); mod.init(0)
stream = torch.cuda.Stream(); mod.set_stream(stream.cuda_stream)
print(json.dumps(bench.siblings_leg(mod, stream, np.random.default_rng(1), 20, 3), indent=1))
PY
grep "GBps\|q8_0\|q4_1\|q5_0\|q5_1\|Error\|frac" gpurun_out/r02_bench_siblings_q5.log
