set -x
test llama.cpp-q_4_0_b200/lib/libggml_b200.so -nt llama.cpp-q_4_0_b200/csrc/set_q4_0.cu || { echo STALE LIB; exit 9; }
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 tools/check_sharded_set.py > gpurun_out/r02_check_sharded_set_n8.log 2>&1; grep "CHECK_SH\|FAIL" gpurun_out/r02_check_sharded_set_n8.log
timeout 1200 python bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r02_bench_n8_a.json 2> gpurun_out/r02_bench_n8_a.err; tail -c 800 gpurun_out/r02_bench_n8_a.err; head -c 4500 gpurun_out/r02_bench_n8_a.json
timeout 1200 python bench.py --gpus 4 --steps 20 --warmup 3 --no-prefill > gpurun_out/r02_bench_n4_a.json 2> gpurun_out/r02_bench_n4_a.err; head -c 300 gpurun_out/r02_bench_n4_a.json
