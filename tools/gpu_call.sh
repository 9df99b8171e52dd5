test llama.cpp-q_4_0_b200/lib/libggml_b200.so -nt llama.cpp-q_4_0_b200/csrc/gemm_q4_0_tc.cu || { echo STALE LIB; exit 9; }
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/r02_pytest_gpu_d.log 2>&1; tail -3 gpurun_out/r02_pytest_gpu_d.log
timeout 600 python tools/bench_gemm.py > gpurun_out/r02_bench_gemm_v3_16w.log 2>&1; cat gpurun_out/r02_bench_gemm_v3_16w.log
