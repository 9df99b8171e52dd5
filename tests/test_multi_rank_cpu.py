"""world_size = 2 and 4 on CPU (gloo): the host-side logic of the row-sharded path (comm.cu) without a GPU --
ggml_b200_shard_rows partitions, the slab all-gather layout [world][n][rows_max] and the scatter back to
dst[n][M].  Each rank computes ITS rows with the CPU oracle (the checker standing in for the kernels), the ranks
all-gather like ggml_b200_mul_mat_q4_0_sharded does with NCCL, and every rank must end with the full result.
The reference's own multi-device mechanism is the row split of GGML_BACKEND_GPU_SPLIT (ggml-cuda.cu:2780-2797,
gather :2514-2540)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port_no, M, K, n, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port_no)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle
    from conftest import load_binding
    mod = load_binding()
    P = oracle.port()
    W = P.quantize_q4_0(P.synthetic(M * K, 0x5EED0001, 0.02).reshape(M, K))      # identical on every rank
    X = P.synthetic(n * K, 0x5EED0002, 1.0, 100).reshape(n, K)
    lo, hi = mod.shard_rows(M, rank, world)
    rows_max = mod.shard_rows(M, 0, world)[1]
    slab = np.zeros((n, rows_max), np.float32)                                     # this rank's [n][rows_max] slab
    if hi > lo:
        slab[:, : hi - lo] = P.mul_mat(W[lo:hi], X, avx2_order=True)
    gathered = [torch.zeros(n, rows_max) for _ in range(world)]
    dist.all_gather(gathered, torch.from_numpy(slab))                              # ncclAllGather's stand-in
    g = torch.stack(gathered).numpy()                                              # [world][n][rows_max]
    # scatter_slabs_kernel restated: y[c][m] = g[owner(m)][c][m - owner(m) * per]
    per = rows_max
    Y = np.empty((n, M), np.float32)
    for m0 in range(0, M, per):
        r = min(m0 // per, world - 1)
        w = min(per, M - m0)
        Y[:, m0:m0 + w] = g[r][:, :w]
    Yfull = P.mul_mat(W, X, avx2_order=True)
    ok = np.array_equal(Y, Yfull)
    in_place_ok = (n == 1 and rows_max * world == M)                               # the zero-copy case of comm.cu
    if in_place_ok:
        ok = ok and np.array_equal(g.reshape(1, -1), Yfull)
    with open(os.path.join(out_dir, f"rank{rank}.txt"), "w") as f:
        f.write("ok" if ok else "mismatch")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,M,K,n", [(2, 512, 256, 1), (2, 300, 128, 3), (4, 1024, 128, 1), (4, 700, 64, 2)])
def test_row_shards_all_gather_and_scatter(tmp_path, world, M, K, n):
    port_no = _free_port()
    mp.spawn(_worker, args=(world, port_no, M, K, n, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert open(tmp_path / f"rank{r}.txt").read() == "ok"
