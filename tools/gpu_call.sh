set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r02_pytest_noalloc_alone.log 2>&1; tail -5 gpurun_out/r02_pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r02_pytest_gpu_b.log 2>&1; tail -5 gpurun_out/r02_pytest_gpu_b.log
