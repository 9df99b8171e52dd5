"""Small ncu target (GPU box): a few launches of each hot kernel through the C ABI on BASELINE
shapes -- the fused n = 1 GEMV on 4096x4096 and 11008x4096, and kernel 1 + the tcgen05 GEMM at
n = 512 on 4096x4096 and 11008x4096.  Kept short because `ncu --set full` replays each launch
~40 times.  Not part of the product."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    mod = bench.load_binding()
    mod.init(0)
    L_ = mod.lib()
    ns = (1, 512)
    if "--gemm-only" in sys.argv:
        ns = (512,)
    if "--gemv-only" in sys.argv:
        ns = (1,)
    for kv in sys.argv[1:]:
        if "=" in kv:
            k, v = kv.split("=")
            mod.set_tuning(k, int(v))
    rng = np.random.default_rng(3)
    if "--set-only" in sys.argv:
        # the persistent multi-op launch on 4 layers of the LLaMA-7B decode list (455 MB of weights: > 3x L2)
        shapes = [(4096, 4096)] * 4 + [(11008, 4096)] * 2 + [(4096, 11008)]
        groups = ["a", "a", "a", "b", "c", "c", "d"]
        blks = {s: bench.random_q4_0(rng, *s) for s in set(shapes)}
        xs = {g: mod.DevBuf.from_numpy(rng.standard_normal((1, K), dtype=np.float32)) for g, (M, K) in zip(groups, shapes)}
        ws, ops = [], []
        for layer in range(4):
            for g, (M, K) in zip(groups, shapes):
                w = mod.Weight(blks[(M, K)])
                ws.append(w)
                ops.append((w, xs[g].ptr, mod.DevBuf(M * 4)))
        keep = [o[2] for o in ops]
        s = mod.OpSet([(w, xp, y.ptr) for (w, xp, y) in ops])
        for _ in range(3):
            s.run()
        mod.synchronize()
        print("set bytes per launch:", s.bytes, "ops", s.n_ops, "barriers", s.barriers, "slots", s.slots)
        s.free()
        del keep
        mod.shutdown()
        print("ncu target done")
        return
    for (M, K) in [(4096, 4096), (11008, 4096)]:
        blk = bench.random_q4_0(rng, M, K)
        w = mod.Weight(blk)
        for n in ns:
            xh = rng.standard_normal((n, K), dtype=np.float32)
            x = mod.DevBuf.from_numpy(xh)
            ys = [mod.DevBuf(n * M * 4) for _ in range(3)]
            for _ in range(2):          # distinct dst per call: with overlap_same_src=1 the second pass runs as a same-src1 run
                for y in ys:
                    rc = L_.ggml_b200_mul_mat_q4_0(w.h, C.c_void_p(x.ptr), n, K * 4, C.c_void_p(y.ptr), M * 4)
                    assert rc == 0, L_.ggml_b200_last_error()
            mod.synchronize()
            if n == 1 and "--host" in sys.argv:      # the one-launch host entry point (gemv_q4_0_hostio_kernel)
                for _ in range(2):
                    mod.mul_mat_host(w, xh)
            x.free()
            for y in ys:
                y.free()
        w.free()
    mod.shutdown()
    print("ncu target done")


if __name__ == "__main__":
    main()
