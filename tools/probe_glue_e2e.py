"""GPU box: the LLaMA-7B decode graph through the reference's seam (tests/glue/layer_driver.c), host time breakdown
(GGML_B200_GLUE_PROFILE=1), with and without segment replay (GGML_B200_GLUE_GRAPHS).  Not part of the product."""
import ctypes as C
import os
import sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
mod = bench.load_binding()
mod.init(0)
G = C.CDLL(os.path.join(bench.PKG, "ggml_glue", "_build", "libggml_b200_glue_ref.so"))
G.glue_model_create.restype = C.c_void_p
G.glue_model_create.argtypes = [C.c_int] * 5 + [C.c_uint]
G.glue_model_eval.restype = C.c_double
G.glue_model_eval.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
G.glue_model_free.argtypes = [C.c_void_p]
m = G.glue_model_create(32, 4096, 11008, 32000, 1, 11)
x = np.random.default_rng(9).standard_normal(4096).astype(np.float32)
lg = np.zeros(32000, np.float32)
G.ggml_b200_glue_configure.argtypes = [C.c_int, C.c_int]
G.ggml_b200_glue_configure.restype = None
G.ggml_b200_glue_set_speculate.argtypes = [C.c_int]
G.ggml_b200_glue_speculation_hits.restype = C.c_long
for acts, fuse, spec in ((1, 0, 0), (0, 1, 0), (1, 1, 0), (1, 1, 1), (1, 0, 0), (1, 1, 0), (1, 1, 1)):
    G.ggml_b200_glue_configure(-1, fuse)
    G.ggml_b200_glue_set_speculate(spec)
    G.glue_model_eval(m, x.ctypes.data_as(C.c_void_p), lg.ctypes.data_as(C.c_void_p), 1, 3, acts)
    G.ggml_b200_glue_profile_dump()
    t = G.glue_model_eval(m, x.ctypes.data_as(C.c_void_p), lg.ctypes.data_as(C.c_void_p), 1, 20, acts)
    print(f"device_acts={acts} fuse_neighbours={fuse} ahead={spec}: {t * 1e3:.3f} ms per eval, launches so far {mod.launch_count()}, segments run ahead so far {G.ggml_b200_glue_speculation_hits()}", flush=True)
    G.ggml_b200_glue_profile_dump()
G.glue_model_free(m)
