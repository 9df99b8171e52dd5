"""Timeline of back-to-back n=1 kernels (GPU box): per-CTA start / dependency-wait / prologue / end
timestamps from the kernels' optional trace, for one shape and a few configs.  Profiling aid."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def run(mod, L_, stream, hs, x, y, M, K, label, nk=12):
    ys = [torch.zeros_like(y) for _ in range(nk)]        # distinct dst per op: same-src1 ops may overlap

    def step():
        for i, h in enumerate(hs[:nk]):
            L_.ggml_b200_mul_mat_q4_0(h, C.c_void_p(x.data_ptr()), 1, K * 4, C.c_void_p(ys[i].data_ptr()), M * 4)
    with torch.cuda.stream(stream):
        step()
    stream.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        mod.set_stream(torch.cuda.current_stream().cuda_stream)
        step()
    mod.set_stream(stream.cuda_stream)
    with torch.cuda.stream(stream):
        g.replay()
    stream.synchronize()
    assert L_.ggml_b200_trace_begin(nk * 400) == 0
    with torch.cuda.stream(stream):
        g.replay()
    stream.synchronize()
    buf = np.zeros((nk * 400, 8), np.uint64)
    n = L_.ggml_b200_trace_end(buf.ctypes.data_as(C.c_void_p), nk * 400)
    rec = buf[:n]
    seqs = sorted(set(rec[:, 0].tolist()))
    t0 = rec[:, 2].min()
    print(f"--- {label}: {M}x{K}, {n} CTA records, {len(seqs)} kernels; times in us relative to first CTA start")
    print("  k  ctas  start[min..max]   wait_done[min..max]  prologue_done[med]  end[min..max]   span")
    prev_end = None
    for i, sq in enumerate(seqs):
        r = rec[rec[:, 0] == sq]
        f = lambda c: (r[:, c].astype(np.float64) - float(t0)) / 1e3
        st, wd, pd, en = f(2), f(3), f(4), f(5)
        gap = "" if prev_end is None else f"  end-to-end {en.max() - prev_end:5.2f}"
        print(f"  {i:2d} {len(r):4d}  {st.min():6.2f}..{st.max():6.2f}   {wd.min():6.2f}..{wd.max():6.2f}      {np.median(pd):6.2f}        {en.min():6.2f}..{en.max():6.2f}  {en.max() - st.min():5.2f}{gap}")
        prev_end = en.max()
    del g


def main():
    M, K = 4096, 4096
    if len(sys.argv) > 2 and sys.argv[1].isdigit():
        M, K = int(sys.argv[1]), int(sys.argv[2])
    mod = bench.load_binding()
    mod.init(0)
    L_ = mod.lib()
    stream = torch.cuda.Stream()
    mod.set_stream(stream.cuda_stream)
    rng = np.random.default_rng(1)
    blk = bench.random_q4_0(rng, M, K)
    hs = []
    for _ in range(16):
        h = L_.ggml_b200_upload_q4_0(blk.ctypes.data_as(C.c_void_p), M, K, (K // 32) * 18, 0, M)
        hs.append(h)
    x = torch.randn(1, K, device="cuda")
    y = torch.zeros(1, M, device="cuda")
    cfgs = [("fused 512x1 rif2 pdl", dict(fused_threads=512, fused_ctas_per_sm=1, fused_rif=2, pdl=1)),
            ("fused 1024x1 pdl", dict(fused_threads=1024, fused_ctas_per_sm=1, fused_rif=1, pdl=1)),
            ("fused 1024x1 pdl overlap_same_src", dict(fused_threads=1024, fused_ctas_per_sm=1, fused_rif=1, pdl=1, overlap_same_src=1))]
    cfgs.append(("fused 512x1 rif2 pdl overlap_same_src", dict(fused_threads=512, fused_ctas_per_sm=1, fused_rif=2, pdl=1, overlap_same_src=1)))
    if "--overlap-only" in sys.argv:
        cfgs = cfgs[-1:]
    for label, t in cfgs:
        for k, v in t.items():
            mod.set_tuning(k, v)
        run(mod, L_, stream, hs, x, y, M, K, label)
    mod.shutdown()


if __name__ == "__main__":
    main()
