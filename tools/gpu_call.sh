test llama.cpp-q_4_0_b200/lib/libggml_b200.so -nt llama.cpp-q_4_0_b200/csrc/siblings.cu || { echo STALE LIB; exit 9; }
mkdir -p gpurun_out
df -h /tmp | tail -1; free -g | head -2; nproc
timeout 1500 python -m pytest tests/test_glue.py -x -q -m gpu -k "falcon" --durations=5 > gpurun_out/r02_pytest_falcon.log 2>&1; tail -25 gpurun_out/r02_pytest_falcon.log
