#!/bin/bash
# scratch: one GPU call
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_neighbours.py tests/test_glue.py -m gpu -x -q > gpurun_out/r02_pytest_neigh.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_neigh.log
timeout 300 python tools/trace_seam.py > gpurun_out/r02_trace_seam.log 2>&1; echo "trace rc=$?" >> gpurun_out/r02_trace_seam.log
GGML_B200_GLUE_PROFILE=1 timeout 300 python tools/probe_glue_e2e.py > gpurun_out/r02_probe_glue_e2e_b.log 2>&1; echo "probe rc=$?" >> gpurun_out/r02_probe_glue_e2e_b.log
tail -5 gpurun_out/r02_pytest_neigh.log; grep "device_acts" gpurun_out/r02_probe_glue_e2e_b.log
