mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_glue.py -x -q -m gpu -k "rwkv or gptneox or tiny_falcon" > gpurun_out/r02_pytest_arch.log 2>&1; tail -8 gpurun_out/r02_pytest_arch.log
