"""GEMV tuning sweep (GPU box): for each BASELINE n=1 shape and each (threads, ctas/SM, unroll)
times the GEMV kernel alone over enough distinct weight copies to exceed L2, via a CUDA graph.
Writes gpurun_out/sweep_gemv.json and prints a table.  Not part of the product."""
import ctypes as C
import itertools
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

SHAPES = [(4096, 4096), (11008, 4096), (4096, 11008), (32000, 4096), (4672, 4544), (18176, 4544), (4544, 18176),
          (8192, 8192), (22016, 8192)]


def main():
    quick = "--quick" in sys.argv
    repack = 0 if "--raw" in sys.argv else 1
    mod = bench.load_binding()
    mod.init(0)
    mod.set_repack(bool(repack))
    L_ = mod.lib()
    stream = torch.cuda.Stream()
    mod.set_stream(stream.cuda_stream)
    rng = np.random.default_rng(1)
    peak, _ = bench.peaks()
    base = dict(gemv_fused=1, pdl=1, gemv_unroll=4, fused_threads=1024, fused_ctas_per_sm=1, fused_rif=1)
    cfgs = []
    for pdl in (1, 0):
        cfgs.append(dict(base, pdl=pdl))
        cfgs.append(dict(base, pdl=pdl, fused_threads=512, fused_ctas_per_sm=2))
        cfgs.append(dict(base, pdl=pdl, fused_threads=512, fused_ctas_per_sm=1, fused_rif=2))
        cfgs.append(dict(base, pdl=pdl, fused_threads=512, fused_ctas_per_sm=1, fused_rif=2, gemv_unroll=2))
        cfgs.append(dict(base, pdl=pdl, fused_threads=512, fused_ctas_per_sm=1, fused_rif=4, gemv_unroll=2))
        cfgs.append(dict(base, pdl=pdl, fused_threads=512, fused_ctas_per_sm=2, fused_rif=2))
        cfgs.append(dict(base, pdl=pdl, fused_threads=256, fused_ctas_per_sm=2, fused_rif=2))
        cfgs.append(dict(base, pdl=pdl, fused_threads=256, fused_ctas_per_sm=2, fused_rif=4, gemv_unroll=2))
    cfgs.append(dict(base, gemv_fused=0))
    if quick:
        cfgs = cfgs[:3]
    results = []
    for (M, K) in (SHAPES[:4] if quick else SHAPES):
        wbytes = 18 * M * K // 32
        copies = max(4, int(600e6 // wbytes) + 1)
        blk = bench.random_q4_0(rng, M, K)
        hs = []
        for _ in range(copies):
            h = L_.ggml_b200_upload_q4_0(blk.ctypes.data_as(C.c_void_p), M, K, (K // 32) * 18, 0, M)
            assert h, L_.ggml_b200_last_error()
            hs.append(h)
        x = torch.randn(1, K, device="cuda")
        y = torch.zeros(1, M, device="cuda")
        best = None
        for cfg in cfgs:
            for k_, v_ in cfg.items():
                mod.set_tuning(k_, v_)

            def step():
                for h in hs:
                    L_.ggml_b200_mul_mat_q4_0(h, C.c_void_p(x.data_ptr()), 1, K * 4, C.c_void_p(y.data_ptr()), M * 4)
            with torch.cuda.stream(stream):
                step()
            stream.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                mod.set_stream(torch.cuda.current_stream().cuda_stream)
                step()
            mod.set_stream(stream.cuda_stream)
            reps = 5
            with torch.cuda.stream(stream):
                g.replay()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(reps):
                    g.replay()
                e1.record(stream)
            stream.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / (reps * copies)
            gbs = bench.algo_bytes(M, K) / us / 1e3
            r = dict(cfg, M=M, K=K, us=us, GBps=gbs, frac=gbs / peak)
            results.append(r)
            if best is None or gbs > best["GBps"]:
                best = r
            del g
        print(f"{M:6d}x{K:<6d} best {best['GBps']:7.0f} GB/s ({best['frac']*100:5.1f}% of measured) {best['us']:7.2f} us  "
              + " ".join(f"{k_}={v_}" for k_, v_ in best.items() if k_ in base and base[k_] != v_), flush=True)
        for r_ in sorted([x_ for x_ in results if x_["M"] == M and x_["K"] == K], key=lambda x_: x_["us"])[:6]:
            print("        %7.2f us  " % r_["us"] + " ".join(f"{k_}={v_}" for k_, v_ in r_.items() if k_ in base and base[k_] != v_), flush=True)
        for h in hs:
            L_.ggml_b200_free_weight(h)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    tag = "raw" if not repack else "planes"
    with open(os.path.join(ROOT, "gpurun_out", f"sweep_gemv_{tag}.json"), "w") as f:
        json.dump(results, f, indent=0)
    mod.shutdown()


if __name__ == "__main__":
    main()
