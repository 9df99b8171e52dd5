"""Multi-GPU check of ggml_b200_mul_mat_q4_0_sharded (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tools/check_sharded.py

For each shape every rank uploads (a) the full weight and (b) its row shard.  The sharded op -- peer-to-peer
exchange over NVLink (n = 1: fused into the GEMV, or as a stand-alone kernel with p2p_gather=5) and NCCL
all-gather (n > 1, or p2p_gather=0) -- must reproduce the single-GPU op of the
same library bit for bit on EVERY rank, also when 300 ops are queued back to back without a host sync (the
flag/parity protocol of p2p_gather_kernel) and when the sequence is replayed from a CUDA graph.  Prints one
"CHECK_SHARDED OK" line from rank 0, exits non-zero on any mismatch.  The kernels of the single-GPU op are
pinned against the oracle by tests/test_gpu_parity.py; this tool pins the sharding around them.
"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def random_q4_0(rng, M, K):
    nbk = K // 32
    blk = rng.integers(0, 256, size=(M, nbk, 18), dtype=np.uint8)
    d = (rng.standard_normal((M, nbk)).astype(np.float32) * 0.02).astype(np.float16)
    blk[:, :, :2] = d.view(np.uint8).reshape(M, nbk, 2)
    return blk


def main():
    import torch
    import torch.distributed as dist
    from conftest import load_binding

    os.environ.setdefault("NCCL_DEBUG", "WARN")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    mod = load_binding()
    torch.cuda.set_device(local)
    mod.init(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    uid = [mod.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    mod.comm_init(rank, world, uid[0])
    stream = torch.cuda.Stream()
    mod.set_stream(stream.cuda_stream)
    L = mod.lib()

    def err():
        return L.ggml_b200_last_error().decode()

    # (M, K): even shards, a ragged last shard, a shard-less last rank at world 8, vocabulary sized
    # + the LLaMA-65B matmuls of the scaling run (BASELINE configs[4])
    shapes = [(4096, 4096), (11008, 4096), (4544, 4544 // 64 * 64), (32000, 4096), (384, 256), (1000, 512),
              (8192, 8192), (22016, 8192), (8192, 22016)]
    import oracle
    try:
        R = oracle.ref()
    except Exception:   # noqa: BLE001
        R = None
    P = oracle.port() if R is None else None
    oracle_err = {}
    failures = []
    rng = np.random.default_rng(1234)                       # same stream on every rank: identical tensors
    weights = []
    for (M, K) in shapes:
        blk = random_q4_0(rng, M, K)
        lo, hi = mod.shard_rows(M, rank, world)
        weights.append((mod.Weight(blk), mod.Weight(blk, lo, hi), M, K, blk))

    def run_case(n, p2p, label):
        mod.set_tuning("p2p_gather", p2p)
        for (wf, ws, M, K, blk) in weights:
            x = torch.from_numpy(np.random.default_rng(M + K + n).standard_normal((n, K), dtype=np.float32)).cuda()
            y_ref = torch.zeros((n, M), dtype=torch.float32, device="cuda")
            y = torch.full((n, M), float("nan"), dtype=torch.float32, device="cuda")
            torch.cuda.synchronize()
            if L.ggml_b200_mul_mat_q4_0(wf.h, C.c_void_p(x.data_ptr()), n, K * 4, C.c_void_p(y_ref.data_ptr()), M * 4):
                raise RuntimeError(err())
            if L.ggml_b200_mul_mat_q4_0_sharded(ws.h, C.c_void_p(x.data_ptr()), n, K * 4, C.c_void_p(y.data_ptr()), M * 4):
                raise RuntimeError(err())
            mod.synchronize()
            if not torch.equal(y, y_ref):
                bad = int((y != y_ref).sum().item())
                failures.append(f"rank {rank} {label} M={M} K={K} n={n}: {bad} of {n * M} differ")
            if n == 1 and label == "p2p-fused":
                # every rank: its gathered result against the ORACLE (the reference itself, else the C port)
                xn = x.cpu().numpy()
                want = R.mul_mat(blk, xn, n_threads=4) if R is not None else P.mul_mat(blk, xn, avx2_order=True)
                e = float(np.max(np.abs(y.cpu().numpy() - want)) / max(float(np.max(np.abs(want))), 1e-30))
                oracle_err[f"{M}x{K}"] = e
                if not e <= 1e-4:
                    failures.append(f"rank {rank} {label} M={M} K={K}: differs from the oracle by {e:.3e}")

    run_case(1, 1, "p2p-fused")
    run_case(1, 5, "p2p-gather")
    run_case(1, 0, "nccl")
    run_case(7, 1, "nccl-n7")
    run_case(64, 1, "nccl-gemm")

    # ---- protocol stress: a chain y1 = W1 x, then many ops queued with no host sync, results kept per op
    mod.set_tuning("p2p_gather", 1)
    wf, ws, M, K, _ = weights[0]                             # 4096 x 4096: y feeds the next op's x
    reps = 300
    x0 = torch.from_numpy(np.random.default_rng(99).standard_normal((1, K), dtype=np.float32)).cuda()
    bufs_ref = [x0] + [torch.zeros((1, M), dtype=torch.float32, device="cuda") for _ in range(reps)]
    bufs = [x0] + [torch.zeros((1, M), dtype=torch.float32, device="cuda") for _ in range(reps)]
    torch.cuda.synchronize()
    for i in range(reps):                                    # values shrink/grow geometrically; finite for 300 steps at 0.02 scale? keep x fixed every 8
        src = bufs_ref[i] if i % 8 else x0
        L.ggml_b200_mul_mat_q4_0(wf.h, C.c_void_p(src.data_ptr()), 1, K * 4, C.c_void_p(bufs_ref[i + 1].data_ptr()), M * 4)
    mod.synchronize()
    dist.barrier()
    t0 = time.time()
    for i in range(reps):
        src = bufs[i] if i % 8 else x0
        if i % 5 == 3:
            mod.set_tuning("p2p_gather", 5)                  # mix in the stand-alone exchange kernel: one op numbering
        if L.ggml_b200_mul_mat_q4_0_sharded(ws.h, C.c_void_p(src.data_ptr()), 1, K * 4, C.c_void_p(bufs[i + 1].data_ptr()), M * 4):
            raise RuntimeError(err())
        mod.set_tuning("p2p_gather", 1)
    mod.synchronize()
    dt = (time.time() - t0) / reps * 1e6
    bad_ops = sum(0 if torch.equal(bufs[i + 1], bufs_ref[i + 1]) else 1 for i in range(reps))
    if bad_ops:
        failures.append(f"rank {rank} chained p2p: {bad_ops} of {reps} ops differ")

    # ---- the same chain from a CUDA graph, replayed 5 times (op numbers live on the device)
    graph_us = None
    try:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            mod.set_stream(torch.cuda.current_stream().cuda_stream)
            for i in range(reps // 3 * 3 - 1):               # an odd op count per replay: parity flips between replays
                src = bufs[i] if i % 8 else x0
                L.ggml_b200_mul_mat_q4_0_sharded(ws.h, C.c_void_p(src.data_ptr()), 1, K * 4, C.c_void_p(bufs[i + 1].data_ptr()), M * 4)
        mod.set_stream(stream.cuda_stream)
        nops = reps // 3 * 3 - 1
        for rep in range(5):
            for b in bufs[1:]:
                b.zero_()
            torch.cuda.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(stream):
                e0.record(stream)
                g.replay()
                e1.record(stream)
            stream.synchronize()
            graph_us = e0.elapsed_time(e1) * 1e3 / nops
            bad_ops = sum(0 if torch.equal(bufs[i + 1], bufs_ref[i + 1]) else 1 for i in range(nops))
            if bad_ops:
                failures.append(f"rank {rank} graph replay {rep}: {bad_ops} of {nops} ops differ")
        del g
    except Exception as e:   # noqa: BLE001
        failures.append(f"rank {rank} graph: {e}")
        mod.set_stream(stream.cuda_stream)

    # ---- same-src1 runs (overlap_same_src): wq/wk/wv-like triples and w3/w1-like pairs do not wait for each other nor
    # for each other's exchange; a consumer of their results must still see all of them.  60 rounds back to back,
    # eagerly and from a CUDA graph, against the same sequence on the unsharded weights.
    run_us = None
    (wf0, ws0, M0, K0, _b0), (wf1, ws1, M1, K1, _b1) = weights[0], weights[1]          # 4096x4096 and 11008x4096
    rounds = 60

    def make_bufs():
        return {k: torch.zeros((1, m), dtype=torch.float32, device="cuda") for k, m in
                (("a", M0), ("b", M0), ("c", M0), ("d", M0), ("e", M1), ("f", M1), ("g", M0))}

    def sequence(fn, w0, w1, B):
        for r in range(rounds):
            xin = x0 if r % 3 == 0 else B["g"]
            for k in ("a", "b", "c"):
                fn(w0, xin, K0, B[k], M0)                       # run of three on xin
            fn(w0, B["b"], K0, B["d"], M0)                      # consumes a member of the run
            for k in ("e", "f"):
                fn(w1, B["d"], K0, B[k], M1)                    # run of two
            fn(w0, B["c"], K0, B["g"], M0)                      # feeds the next round

    def op_full(w, x, K, y, M):
        assert L.ggml_b200_mul_mat_q4_0(w.h, C.c_void_p(x.data_ptr()), 1, K * 4, C.c_void_p(y.data_ptr()), M * 4) == 0, err()

    def op_shard(w, x, K, y, M):
        assert L.ggml_b200_mul_mat_q4_0_sharded(w.h, C.c_void_p(x.data_ptr()), 1, K * 4, C.c_void_p(y.data_ptr()), M * 4) == 0, err()

    Bref, Bsh = make_bufs(), make_bufs()
    torch.cuda.synchronize()
    sequence(op_full, wf0, wf1, Bref)
    mod.synchronize()
    mod.set_tuning("overlap_same_src", 1)
    try:
        for attempt in range(2):                                # second pass: run members are known, half-SM kernels
            for b_ in Bsh.values():
                b_.zero_()
            torch.cuda.synchronize()
            dist.barrier()
            sequence(op_shard, ws0, ws1, Bsh)
            mod.synchronize()
            for k in Bref:
                if not torch.equal(Bsh[k], Bref[k]):
                    failures.append(f"rank {rank} same-src runs, pass {attempt}: buffer {k} differs")
        g2 = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g2, stream=stream):
            mod.set_stream(torch.cuda.current_stream().cuda_stream)
            sequence(op_shard, ws0, ws1, Bsh)
        mod.set_stream(stream.cuda_stream)
        for rep in range(3):
            for b_ in Bsh.values():
                b_.zero_()
            torch.cuda.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(stream):
                e0.record(stream)
                g2.replay()
                e1.record(stream)
            stream.synchronize()
            run_us = e0.elapsed_time(e1) * 1e3 / (rounds * 7)
            for k in Bref:
                if not torch.equal(Bsh[k], Bref[k]):
                    failures.append(f"rank {rank} same-src runs, graph replay {rep}: buffer {k} differs")
        del g2
    except Exception as e:   # noqa: BLE001
        failures.append(f"rank {rank} same-src runs: {e}")
        mod.set_stream(stream.cuda_stream)
    mod.set_tuning("overlap_same_src", 0)

    # NCCL per-op time on the same chain for the log
    mod.set_tuning("p2p_gather", 0)
    dist.barrier()
    t0 = time.time()
    for i in range(reps):
        src = bufs[i] if i % 8 else x0
        L.ggml_b200_mul_mat_q4_0_sharded(ws.h, C.c_void_p(src.data_ptr()), 1, K * 4, C.c_void_p(bufs[i + 1].data_ptr()), M * 4)
    mod.synchronize()
    dt_nccl = (time.time() - t0) / reps * 1e6
    mod.set_tuning("p2p_gather", 1)

    allf = [None] * world
    dist.all_gather_object(allf, failures)
    flat = [f for fs in allf for f in fs]
    if rank == 0:
        for f in flat:
            print("FAIL", f)
        print(f"CHECK_SHARDED {'OK' if not flat else 'FAILED'} world={world} shapes={len(shapes)} "
              f"chain_us_per_op p2p={dt:.1f} (graph {graph_us if graph_us is None else round(graph_us, 2)}) nccl={dt_nccl:.1f} "
              f"same-src-runs graph us/op={run_us if run_us is None else round(run_us, 2)}")
        print("oracle (%s) max|dY|/max|Y| of rank 0's gathered n = 1 results: %s" % ("reference" if R is not None else "C port",
              " ".join("%s=%.2e" % kv for kv in oracle_err.items())))
    sys.stdout.flush()
    dist.barrier()
    torch.cuda.synchronize()
    os._exit(1 if flat else 0)


if __name__ == "__main__":
    main()
