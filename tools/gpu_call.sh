test llama.cpp-q_4_0_b200/lib/libggml_b200.so -nt llama.cpp-q_4_0_b200/csrc/siblings.cu || { echo STALE LIB; exit 9; }
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_glue.py tests/test_gpu_siblings.py -x -q -m gpu > gpurun_out/r02_pytest_glue_siblings.log 2>&1; tail -8 gpurun_out/r02_pytest_glue_siblings.log
