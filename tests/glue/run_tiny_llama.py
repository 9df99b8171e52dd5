"""Runs tests/glue/llama_driver.cpp's tiny_llama_run from one of the two builds of the reference's llama.cpp
(ggml_glue/_build/libllama_ref_cpu.so = plain CPU reference, libllama_b200.so = reference -DGGML_USE_CUBLAS over
ggml-b200-glue.c + libggml_b200.so) in its own process and saves the prompt / decode logits.
usage: run_tiny_llama.py <lib.so> <model.bin> <n_gpu_layers> <n_prompt> <out.npy>"""
import ctypes as C
import sys

import numpy as np

lib, model, ngl, n_prompt, out = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), sys.argv[5]
L = C.CDLL(lib)
L.tiny_llama_run.restype = C.c_int
NV = 512
toks = [1] + [(7 * i + 3) % NV for i in range(n_prompt - 1)]          # BOS first (llama.cpp:1336)
arr = (C.c_int * n_prompt)(*toks)
lp, ld = np.zeros(NV, np.float32), np.zeros(NV, np.float32)
rc = L.tiny_llama_run(model.encode(), ngl, arr, n_prompt, 11, 4, lp.ctypes.data_as(C.c_void_p), ld.ctypes.data_as(C.c_void_p), NV)
if rc != 0:
    sys.exit("tiny_llama_run failed: %d" % rc)
np.save(out, np.stack([lp, ld]))
