test llama.cpp-q_4_0_b200/lib/libggml_b200.so -nt llama.cpp-q_4_0_b200/csrc/gemm_q4_0_tc3.cuh || { echo STALE LIB; exit 9; }
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -x -q -m gpu -k "gemm or prefill or auto_dispatch" > gpurun_out/t.log 2>&1; tail -2 gpurun_out/t.log
timeout 600 python tools/bench_gemm.py > gpurun_out/r02_bench_gemm_v3_16w_c.log 2>&1; cat gpurun_out/r02_bench_gemm_v3_16w_c.log
