test llama.cpp-q_4_0_b200/lib/libggml_b200.so -nt llama.cpp-q_4_0_b200/csrc/siblings.cu || { echo STALE LIB; exit 9; }
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_fullsize.py -x -q -m gpu --durations=5 > gpurun_out/r02_pytest_fullsize.log 2>&1; tail -14 gpurun_out/r02_pytest_fullsize.log
