"""Runs tests/glue/llama_driver.cpp's tiny_llama_run from one of the two builds of the reference's llama.cpp
(ggml_glue/_build/libllama_ref_cpu.so = plain CPU reference, libllama_b200.so = reference -DGGML_USE_CUBLAS over
ggml-b200-glue.c + libggml_b200.so) in its own process and saves the prompt / decode logits.
usage: run_tiny_llama.py <lib.so> <model.bin> <n_gpu_layers> <n_prompt> <out.npy> [n_decode]
With n_decode: a decode loop of that many steps (tiny_llama_decode); saves [n_decode][n_vocab] logits and prints
"decode_ms_per_step=... replays=... ahead=..." (the glue's counters, -1 from the CPU build)."""
import ctypes as C
import sys

import numpy as np

lib, model, ngl, n_prompt, out = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), sys.argv[5]
L = C.CDLL(lib)
L.tiny_llama_run.restype = C.c_int
import os
NV = int(os.environ.get("TINY_LLAMA_N_VOCAB", "512"))
toks = [1] + [(7 * i + 3) % NV for i in range(n_prompt - 1)]          # BOS first (llama.cpp:1336)
arr = (C.c_int * n_prompt)(*toks)
if len(sys.argv) > 6:
    n_decode = int(sys.argv[6])
    L.tiny_llama_decode.restype = C.c_int
    steps = np.zeros((n_decode, NV), np.float32)
    ms = C.c_double(0.0)
    cnt = (C.c_long * 2)(0, 0)
    rc = L.tiny_llama_decode(model.encode(), ngl, arr, n_prompt, n_decode, int(os.environ.get("TINY_LLAMA_THREADS", "4")), steps.ctypes.data_as(C.c_void_p), C.byref(ms), cnt, NV)
    if rc != 0:
        sys.exit("tiny_llama_decode failed: %d" % rc)
    print(f"decode_ms_per_step={ms.value:.4f} replays={cnt[0]} ahead={cnt[1]}", flush=True)
    np.save(out, steps)
    sys.exit(0)
lp, ld = np.zeros(NV, np.float32), np.zeros(NV, np.float32)
rc = L.tiny_llama_run(model.encode(), ngl, arr, n_prompt, 11, 4, lp.ctypes.data_as(C.c_void_p), ld.ctypes.data_as(C.c_void_p), NV)
if rc != 0:
    sys.exit("tiny_llama_run failed: %d" % rc)
np.save(out, np.stack([lp, ld]))
