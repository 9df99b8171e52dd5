"""Runs tests/glue/falcon_driver.cpp's tiny_falcon_run (or gptneox_driver.cpp's tiny_gptneox_run: whichever the library
exports) from one of the two builds of the reference's arch/<arch>/<arch>.cpp (ggml_glue/_build/lib<arch>_ref_cpu.so = plain
CPU reference, lib<arch>_b200.so = the same source with the loader tag, -DGGML_USE_CUBLAS over ggml-b200-glue.c +
libggml_b200.so) in its own process and saves the prompt / decode logits.
usage: run_tiny_falcon.py <lib.so> <model.bin> <n_prompt> <out.npy> [n_threads]"""
import ctypes as C
import sys

import numpy as np

lib, model, n_prompt, out = sys.argv[1], sys.argv[2], int(sys.argv[3]), sys.argv[4]
threads = int(sys.argv[5]) if len(sys.argv) > 5 else 4
L = C.CDLL(lib)
run = L.tiny_gptneox_run if "gptneox" in lib else L.tiny_rwkv_run if "rwkv" in lib else L.tiny_falcon_run
run.restype = C.c_int
NV = 512
toks = [(7 * i + 3) % NV for i in range(n_prompt)]
arr = (C.c_int * n_prompt)(*toks)
lp, ld = np.zeros(NV, np.float32), np.zeros(NV, np.float32)
rc = run(model.encode(), arr, n_prompt, 11, threads, lp.ctypes.data_as(C.c_void_p), ld.ctypes.data_as(C.c_void_p), NV)
if rc != 0:
    sys.exit("tiny model run failed: %d" % rc)
np.save(out, np.stack([lp, ld]))
try:                                   # the drop-in build links libggml_b200.so: how many of its kernels ran in this process
    L.ggml_b200_launch_count.restype = C.c_uint64
    print("b200_kernel_launches=%d" % L.ggml_b200_launch_count(), file=sys.stderr)
except AttributeError:
    pass
