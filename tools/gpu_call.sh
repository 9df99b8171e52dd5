#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_neighbours.py -m gpu -x -q > gpurun_out/r02_pytest_neigh_final.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_neigh_final.log
tail -6 gpurun_out/r02_pytest_neigh_final.log
