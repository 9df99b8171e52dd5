test llama.cpp-q_4_0_b200/lib/libggml_b200.so -nt llama.cpp-q_4_0_b200/csrc/gemm_q4_0_tc3.cuh || { echo STALE LIB; exit 9; }
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q -m gpu -k "gemm or auto_dispatch or same_src1_batched or static" > gpurun_out/r02_pytest_gemm_v3.log 2>&1; tail -3 gpurun_out/r02_pytest_gemm_v3.log
timeout 600 python tools/bench_gemm.py > gpurun_out/r02_bench_gemm_v3.log 2>&1; cat gpurun_out/r02_bench_gemm_v3.log; cp gpurun_out/bench_gemm.json gpurun_out/r02_bench_gemm_v3.json
timeout 600 python tools/bench_gemm.py gemm_v3=0 > gpurun_out/r02_bench_gemm_v1.log 2>&1; cat gpurun_out/r02_bench_gemm_v1.log
timeout 120 python tools/trace_gemm.py > gpurun_out/r02_trace_gemm_v3.log 2>&1
