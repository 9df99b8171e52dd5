"""GPU box: device timeline of ONE decode eval through the seam (tests/glue/layer_driver.c, LLaMA-7B shapes): per kernel
the first CTA's start, the end of the dependency wait and the last CTA's end (globaltimer stamps of the kernels' optional
trace), plus the idle time between kernels.  Shows where a segment's time goes.  Profiling aid, not part of the product."""
import ctypes as C
import os
import sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
mod = bench.load_binding()
mod.init(0)
L_ = mod.lib()
G = C.CDLL(os.path.join(bench.PKG, "ggml_glue", "_build", "libggml_b200_glue_ref.so"))
G.glue_model_create.restype = C.c_void_p
G.glue_model_create.argtypes = [C.c_int] * 5 + [C.c_uint]
G.glue_model_eval.restype = C.c_double
G.glue_model_eval.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
G.glue_model_free.argtypes = [C.c_void_p]
G.ggml_b200_glue_configure.argtypes = [C.c_int, C.c_int]
G.ggml_b200_glue_configure.restype = None
L_.ggml_b200_trace_begin.argtypes = [C.c_uint]
L_.ggml_b200_trace_end.argtypes = [C.c_void_p, C.c_uint]
n_layer = 8
m = G.glue_model_create(n_layer, 4096, 11008, 32000, 1, 11)
x = np.random.default_rng(9).standard_normal(4096).astype(np.float32)
lg = np.zeros(32000, np.float32)
for fuse in (0, 1):
    G.ggml_b200_glue_configure(-1, fuse)
    G.glue_model_eval(m, x.ctypes.data_as(C.c_void_p), lg.ctypes.data_as(C.c_void_p), 1, 4, 1)
    cap = 40000
    assert L_.ggml_b200_trace_begin(cap) == 0
    t = G.glue_model_eval(m, x.ctypes.data_as(C.c_void_p), lg.ctypes.data_as(C.c_void_p), 1, 1, 1)
    buf = np.zeros((cap, 8), np.uint64)
    n = L_.ggml_b200_trace_end(buf.ctypes.data_as(C.c_void_p), cap)
    rec = buf[:n]
    print(f"=== fuse_neighbours={fuse}: eval {t * 1e6:.0f} us, {n} CTA records")
    ks = []
    for sq in set(rec[:, 0].tolist()):
        r = rec[rec[:, 0] == sq]
        en = r[:, 5][r[:, 5] > 0]
        wd = r[:, 3][r[:, 3] > 0]
        ks.append((int(r[:, 2].min()), int(wd.min()) if len(wd) else 0, int(wd.max()) if len(wd) else 0, int(en.max()) if len(en) else 0, len(r), int(sq)))
    ks.sort()
    t0 = ks[0][0]
    prev_end = None
    busy = 0.0
    print("   start    wait_done(first..last)   end     dur   idle_before  ctas  kind")
    for i, (st, w0, w1, en, nc, sq) in enumerate(ks):
        idle = (st - prev_end) / 1e3 if prev_end else 0.0
        kind = "eltwise" if sq & 0x80000000 else "gemv"
        if i < 40:
            print(f"  {(st - t0) / 1e3:7.1f}  {(w0 - t0) / 1e3:7.1f}..{(w1 - t0) / 1e3:7.1f}  {(en - t0) / 1e3:7.1f}  {(en - st) / 1e3:5.1f}  {idle:7.1f}      {nc:4d}  {kind}")
        busy += (en - max(st, prev_end or st)) / 1e3
        prev_end = max(en, prev_end or en)
    print(f"  {len(ks)} kernels, span {(prev_end - t0) / 1e3:.0f} us, union of kernel time {busy:.0f} us")
G.glue_model_free(m)
