set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_glue.py -x -q -m gpu 2>&1 | tail -3
GGML_B200_GLUE_PROFILE=1 timeout 300 python tools/probe_glue_e2e.py 2>&1 | tail -12
GGML_B200_GLUE_PROFILE=1 GGML_B200_GLUE_GRAPHS=0 timeout 300 python tools/probe_glue_e2e.py 2>&1 | tail -12
