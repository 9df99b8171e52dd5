N=${1:-8}
test llama.cpp-q_4_0_b200/lib/libggml_b200.so -nt llama.cpp-q_4_0_b200/csrc/siblings.cu || { echo STALE LIB; exit 9; }
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 1800 python -m pytest tests/test_multi_gpu.py -q -m gpu > gpurun_out/r02_pytest_multi_gpu_n$N.log 2>&1; tail -4 gpurun_out/r02_pytest_multi_gpu_n$N.log; grep "CHECK_SHARDED\|oracle (" gpurun_out/r02_pytest_multi_gpu_n$N.log | head
for n in 8 4; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 100 --warmup 5 > gpurun_out/r02_bench_n${n}_c.json 2> gpurun_out/r02_bench_n${n}_c.err; echo "bench $n rc=$?"; head -c 400 gpurun_out/r02_bench_n${n}_c.json; echo
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29622 tools/check_sharded.py > gpurun_out/r02_check_sharded_n8.log 2>&1; grep "CHECK_SHARDED\|oracle (\|FAIL" gpurun_out/r02_check_sharded_n8.log | head
