set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_set.py -x -q -m gpu 2>&1 | tail -3
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29543 tools/check_sharded_set.py --quick > gpurun_out/r02_check_sharded_set_quick_n2.log 2>&1; tail -12 gpurun_out/r02_check_sharded_set_quick_n2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 tools/check_sharded_set.py > gpurun_out/r02_check_sharded_set_n2.log 2>&1; tail -12 gpurun_out/r02_check_sharded_set_n2.log
