"""Multi-GPU check of ggml_b200_set_create_sharded (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29543 \
        tools/check_sharded_set.py

A list of n = 1 mul_mats on row shards in ONE persistent launch per rank, the exchange inside it (tagged words into
every rank's arena over NVLink, gather warp per CTA).  On EVERY rank every dst must be bit-identical to the
single-GPU persistent launch over the unsharded weights (a row's arithmetic does not depend on who owns it), and on
rank 0 within 1e-4 of the ORACLE (the reference's own ggml_graph_compute) along the chain.  Lists:
  A  the LLaMA-65B shapes SCALE times (8192x8192 x4, 22016x8192 x2, 8192x22016, 32000x8192), external x, the dst buffers
     reused by three "layers" (write-after-write barriers), 20 launches back to back + CUDA-graph replays
  B  the same shapes chained like the model (wo reads wq's dst, w3/w1 read wo's, w2 reads w1's, the next layer w2's):
     every change of x is a gathered vector read out of the arena (linked runs)
  C  small ragged shapes with empty shards at 8 ranks, a dst read through a partial overlap (barrier that also waits for
     the gather warp) and an in-place chain
Prints "CHECK_SHARDED_SET OK" from rank 0, exits non-zero on any mismatch.
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def random_q4_0(rng, M, K, scale):
    nbk = K // 32
    blk = rng.integers(0, 256, size=(M, nbk, 18), dtype=np.uint8)
    d = (scale * (2.0 + np.abs(rng.standard_normal((M, nbk)).astype(np.float32))) / 8.0) * rng.choice(np.array([-1.0, 1.0], np.float32), size=(M, nbk))
    blk[:, :, :2] = d.astype(np.float16).view(np.uint8).reshape(M, nbk, 2)
    return blk


def main():
    import torch
    import torch.distributed as dist
    from conftest import load_binding

    os.environ.setdefault("NCCL_DEBUG", "WARN")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    quick = "--quick" in sys.argv
    mod = load_binding()
    torch.cuda.set_device(local)
    mod.init(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    uid = [mod.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    mod.comm_init(rank, world, uid[0])
    stream = torch.cuda.Stream()
    mod.set_stream(stream.cuda_stream)
    failures = []
    notes = []
    rng = np.random.default_rng(4321)                       # same stream on every rank: identical tensors

    def dev(a):
        return torch.from_numpy(np.ascontiguousarray(a)).cuda()

    def weights_for(shapes):
        out = []
        for (M, K) in shapes:
            blk = random_q4_0(rng, M, K, 0.6 / np.sqrt(K))      # gain ~1 per matmul: a chain stays in range
            lo, hi = mod.shard_rows(M, rank, world)
            out.append((blk, mod.Weight(blk), mod.Weight(blk, lo, hi), M, K))
        return out

    def compare(label, got, want, exact=True):
        for i, (g_, w_) in enumerate(zip(got, want)):
            if exact and not torch.equal(g_, w_):
                bad = int((g_ != w_).sum().item())
                failures.append(f"rank {rank} {label}: dst {i}: {bad} of {g_.numel()} values differ from the single-GPU launch")

    def oracle_check(label, W, x, y):
        import oracle
        R = oracle.ref()
        want = R.mul_mat(W, x, n_threads=8) if R is not None else oracle.port().mul_mat(W, x, avx2_order=True)
        err = float(np.max(np.abs(y - want)) / max(float(np.max(np.abs(want))), 1e-30))
        notes.append(f"{label} {W.shape[0]}x{W.shape[1] * 32}: {err:.2e}")
        if not err <= 1e-4:
            failures.append(f"rank {rank} {label}: max|dY|/max|Y| = {err:.3e} against the oracle")

    def run_both(label, ws, ops_idx, x_of, y_of, make, launches=3, graph=False, replays=10):
        """ops_idx: [(weight index, x key, y key)]; make() -> dict of fresh buffers (x inputs filled).  Runs the list on
        the full weights (single-GPU launch) and on the shards (collective launch), compares every buffer."""
        Bf, Bs = make(), make()
        torch.cuda.synchronize()
        full = mod.OpSet([(ws[w][1], Bf[xk].data_ptr(), Bf[yk].data_ptr()) for (w, xk, yk) in ops_idx])
        shard = mod.OpSet([(ws[w][2], Bs[xk].data_ptr(), Bs[yk].data_ptr()) for (w, xk, yk) in ops_idx], sharded=True)
        keys = sorted({yk for (_, _, yk) in ops_idx})
        inputs = {k: v.clone() for k, v in Bf.items() if k not in keys}
        ms = None
        for it in range(launches):
            for B in (Bf, Bs):
                for k in keys:
                    B[k].fill_(7.0)
                for k, v in inputs.items():
                    B[k].copy_(v)
            torch.cuda.synchronize()
            dist.barrier()
            with torch.cuda.stream(stream):
                full.run()
            stream.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(stream):
                e0.record(stream)
                shard.run()
                e1.record(stream)
            stream.synchronize()
            ms = e0.elapsed_time(e1)
            compare(f"{label} launch {it}", [Bs[k] for k in keys], [Bf[k] for k in keys])
        if graph:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                mod.set_stream(torch.cuda.current_stream().cuda_stream)
                shard.run()
            mod.set_stream(stream.cuda_stream)
            for k in keys:
                Bs[k].fill_(7.0)
            for k, v in inputs.items():
                Bs[k].copy_(v)
            torch.cuda.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(stream):
                e0.record(stream)
                for _ in range(replays):
                    g.replay()
                e1.record(stream)
            stream.synchronize()
            ms = e0.elapsed_time(e1) / replays
            compare(f"{label} graph x{replays}", [Bs[k] for k in keys], [Bf[k] for k in keys])
            del g
        info = (shard.barriers, shard.linked_runs, shard.n_ops, shard.bytes)
        dist.barrier()
        torch.cuda.synchronize()
        shard.free()
        full.free()
        return Bs, Bf, ms, info

    # ---- lists A and B: the LLaMA-65B shapes
    big = [(8192, 8192)] * 4 + [(22016, 8192)] * 2 + [(8192, 22016), (32000, 8192)]
    if quick:
        big = [(1024, 1024)] * 4 + [(2752, 1024)] * 2 + [(1024, 2752), (4000, 1024)]
    wb = weights_for(big)
    Kd = big[6][1]
    x_attn = np.random.default_rng(7).standard_normal((1, big[0][1]), dtype=np.float32)
    x_mid = np.random.default_rng(8).standard_normal((1, Kd), dtype=np.float32)

    def make_A():
        B = {"x_attn": dev(x_attn), "x_ao": dev(x_attn * 0.5), "x_ffn": dev(x_attn[:, ::-1].copy()), "x_mid": dev(x_mid), "x_out": dev(x_attn * 2)}
        for i, (M, K) in enumerate(big):
            B[f"y{i}"] = torch.zeros((1, M), dtype=torch.float32, device="cuda")
        return B
    xsrc = ["x_attn", "x_attn", "x_attn", "x_ao", "x_ffn", "x_ffn", "x_mid"]
    opsA = []
    for layer in range(3):
        opsA += [(i, xsrc[i], f"y{i}") for i in range(7)]
    opsA.append((7, "x_out", "y7"))
    Bs, Bf, msA, infoA = run_both("A (65B shapes, reused dst)", wb, opsA, None, None, make_A, launches=3 if quick else 5, graph=True)
    if rank == 0:
        for i, xk in ((0, "x_attn"), (4, "x_ffn"), (6, "x_mid"), (7, "x_out")):
            oracle_check("A", wb[i][0], Bs[xk].cpu().numpy(), Bs[f"y{i}"].cpu().numpy())

    def make_B():
        B = {"x0": dev(x_attn)}
        for layer in range(2):
            for i, (M, K) in enumerate(big[:7]):
                B[f"L{layer}y{i}"] = torch.zeros((1, M), dtype=torch.float32, device="cuda")
        B["out"] = torch.zeros((1, big[7][0]), dtype=torch.float32, device="cuda")
        return B
    opsB = []
    cur = "x0"
    for layer in range(2):
        p = f"L{layer}y"
        opsB += [(0, cur, p + "0"), (1, cur, p + "1"), (2, cur, p + "2"), (3, p + "0", p + "3"),
                 (4, p + "3", p + "4"), (5, p + "3", p + "5"), (6, p + "5", p + "6")]
        cur = p + "6"
    opsB.append((7, cur, "out"))
    Bs, Bf, msB, infoB = run_both("B (65B shapes, chained)", wb, opsB, None, None, make_B, launches=3, graph=True)
    if infoB[1] < 6:
        failures.append(f"rank {rank} B: expected linked runs, got {infoB}")
    if rank == 0:
        oracle_check("B", wb[3][0], Bs["L0y0"].cpu().numpy(), Bs["L0y3"].cpu().numpy())
        oracle_check("B", wb[6][0], Bs["L1y5"].cpu().numpy(), Bs["L1y6"].cpu().numpy())
        oracle_check("B", wb[7][0], Bs["L1y6"].cpu().numpy(), Bs["out"].cpu().numpy())
    for w_ in wb:
        w_[1].free(); w_[2].free()

    # ---- list C: ragged / empty shards, partial-overlap read (barrier + gather), in-place chain
    small = [(384, 256), (1000, 512), (512, 1024), (256, 384), (640, 640)]
    wc = weights_for(small)
    xs256 = np.random.default_rng(11).standard_normal((1, 256), dtype=np.float32)
    xs512 = np.random.default_rng(12).standard_normal((1, 512), dtype=np.float32)
    xs640 = np.random.default_rng(13).standard_normal((1, 640), dtype=np.float32)

    def make_C():
        return {"x256": dev(xs256), "x512": dev(xs512), "x640": dev(xs640),
                "a": torch.zeros((1, 384), dtype=torch.float32, device="cuda"),      # 384 rows: empty shards at 8 ranks
                "b": torch.zeros((1, 1024), dtype=torch.float32, device="cuda"),     # op1 writes [0,1000), op2 reads all 1024: partial overlap
                "c": torch.zeros((1, 512), dtype=torch.float32, device="cuda"),
                "d": torch.zeros((1, 256), dtype=torch.float32, device="cuda"),
                "p": dev(xs640), "q": torch.zeros((1, 640), dtype=torch.float32, device="cuda")}
    opsC = [(0, "x256", "a"),        # a = W0 x256
            (3, "a", "d"),           # d = W3 a           (linked: gathered vector, empty shards on high ranks)
            (1, "x512", "b"),        # b[0:1000] = W1 x512
            (2, "b", "c"),           # c = W2 b[0:1024]   (x overlaps op1's dst partially: barrier that waits for the gather)
            (4, "p", "q"),           # q = W4 p
            (4, "q", "p"),           # p = W4 q           (overwrites the x of the op before: WAR barrier)
            (4, "p", "q")]           # q = W4 p           (linked to the op before; WAW on q)
    Bs, Bf, msC, infoC = run_both("C (ragged, barriers)", wc, opsC, None, None, make_C, launches=6, graph=True, replays=1)      # (the list overwrites its own input p: one replay per reset)
    if rank == 0:
        oracle_check("C", wc[3][0], Bs["a"].cpu().numpy(), Bs["d"].cpu().numpy())
        oracle_check("C", wc[2][0], Bs["b"].cpu().numpy(), Bs["c"].cpu().numpy())

    allf = [None] * world
    dist.all_gather_object(allf, failures)
    flat = [f for fs in allf for f in fs]
    if rank == 0:
        for f in flat:
            print("FAIL", f)
        bytesA = infoA[3]
        print(f"CHECK_SHARDED_SET {'OK' if not flat else 'FAILED'} world={world} "
              f"A: {infoA[2]} ops/rank {infoA[0]} barriers {msA * 1e3:.0f} us per launch ({bytesA * world / msA / 1e6:.0f} GB/s aggregate) | "
              f"B: {infoB[1]} linked runs {msB * 1e3:.0f} us | C: {infoC[0]} barriers {infoC[1]} linked {msC * 1e3:.0f} us | oracle max rel err: {'; '.join(notes)}")
    sys.stdout.flush()
    dist.barrier()
    torch.cuda.synchronize()
    os._exit(1 if flat else 0)


if __name__ == "__main__":
    main()
