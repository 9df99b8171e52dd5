#!/bin/bash
# One GPU-box call as used at the end of a round (gpurun -- 'bash tools/gpu_call.sh'): the -m gpu tests, smoke(), the bench
# line and the reference arm; logs under gpurun_out/.  Edit freely: this is a scratch driver, not part of the product.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref_n1.json 2> gpurun_out/bench_ref_n1.err; echo "reference arm rc=$?"
