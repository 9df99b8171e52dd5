mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_glue.py -x -q -m gpu > gpurun_out/t.log 2>&1; tail -3 gpurun_out/t.log
GGML_B200_GLUE_PROFILE=1 timeout 600 python tools/probe_glue_e2e.py > gpurun_out/r02_probe_glue_e2e_b.log 2>&1; cat gpurun_out/r02_probe_glue_e2e_b.log | tail -8
