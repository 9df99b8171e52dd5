set -x
test llama.cpp-q_4_0_b200/lib/libggml_b200.so -nt llama.cpp-q_4_0_b200/csrc/set_q4_0.cu || { echo STALE LIB; exit 9; }
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 tools/check_sharded_set.py > gpurun_out/r02_check_sharded_set_n2.log 2>&1; grep "CHECK_SH\|FAIL" gpurun_out/r02_check_sharded_set_n2.log
timeout 1200 python bench.py --gpus 2 --steps 20 --warmup 3 --no-prefill > gpurun_out/r02_bench_n2_b.json 2> gpurun_out/r02_bench_n2_b.err; tail -c 800 gpurun_out/r02_bench_n2_b.err; python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_bench_n2_b.json"))
print(d["value"], d["ms_per_step"], d["config"]["one_launch_per_step"], d["roofline"]["frac"])
PY
