"""Probe (GPU box): per-launch time of the n=1 kernels on L2-hot weights (same buffer every launch)
vs cold weights (distinct buffers > L2), long enough for clocks to settle.  Profiling aid."""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def timeit(mod, L_, stream, hs, x, y, M, K, nk, reps=200):
    def step():
        for i in range(nk):
            L_.ggml_b200_mul_mat_q4_0(hs[i % len(hs)], C.c_void_p(x.data_ptr()), 1, K * 4, C.c_void_p(y.data_ptr()), M * 4)
    with torch.cuda.stream(stream):
        step()
    stream.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        mod.set_stream(torch.cuda.current_stream().cuda_stream)
        step()
    mod.set_stream(stream.cuda_stream)
    with torch.cuda.stream(stream):
        for _ in range(20):
            g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            g.replay()
        e1.record(stream)
    stream.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * nk)


def main():
    mod = bench.load_binding()
    mod.init(0)
    L_ = mod.lib()
    stream = torch.cuda.Stream()
    mod.set_stream(stream.cuda_stream)
    rng = np.random.default_rng(1)
    for (M, K) in [(4096, 4096), (11008, 4096), (256, 4096), (148, 4096)]:
        blk = bench.random_q4_0(rng, M, K)
        ncopies = max(2, int(400e6 // (18 * M * K // 32)))
        hs = [L_.ggml_b200_upload_q4_0(blk.ctypes.data_as(C.c_void_p), M, K, (K // 32) * 18, 0, M) for _ in range(min(ncopies, 64))]
        x = torch.randn(1, K, device="cuda")
        y = torch.zeros(1, M, device="cuda")
        for label, t in [("fused1024 pdl", dict(gemv_fused=1, fused_threads=1024, fused_ctas_per_sm=1, fused_rif=1, pdl=1)),
                         ("fused1024 nopdl", dict(gemv_fused=1, fused_threads=1024, fused_ctas_per_sm=1, fused_rif=1, pdl=0)),
                         ("fused512x1rif2 pdl", dict(gemv_fused=1, fused_threads=512, fused_ctas_per_sm=1, fused_rif=2, pdl=1)),
                         ("unfused pdl", dict(gemv_fused=0, pdl=1))]:
            for k, v in t.items():
                mod.set_tuning(k, v)
            cold = timeit(mod, L_, stream, hs, x, y, M, K, len(hs))
            hot = timeit(mod, L_, stream, hs[:1], x, y, M, K, 64)
            print(f"{M:6d}x{K:<6d} {label:20s} cold {cold:6.2f} us   L2-hot {hot:6.2f} us   ideal-HBM {bench.algo_bytes(M, K) / 6547.8e3:5.2f} us", flush=True)
        for h in hs:
            L_.ggml_b200_free_weight(h)
    # empty-kernel floor: launch rate of a graph of trivial kernels = M=1 rows
    mod.shutdown()


if __name__ == "__main__":
    main()
