"""CPU tests of the drop-in boundary: the shared library loads, exports every symbol the header
declares, its host-side logic (shard arithmetic) is right, and compute entry points FAIL LOUDLY
without a GPU (no CPU fallback exists)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, load_binding


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "ggml-b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ggml_b200_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    mod = load_binding()
    L = C.CDLL(mod.LIB_PATH)
    names = _declared_symbols()
    assert len(names) >= 35
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_binding_covers_header():
    mod = load_binding()
    L = mod.lib()
    for n in _declared_symbols():
        assert getattr(L, n).argtypes is not None, f"{n} has no ctypes signature in ggml_b200.py"


def test_header_cites_reference_lines():
    text = open(os.path.join(ROOT, "include", "ggml-b200.h")).read()
    for cite in ("ggml-cuda.cu:2766-2809", "ggml.c:11373-11411", "ggml.c:1122-1280", "ggml-cuda.cu:2811-2828",
                 "ggml.c:2593-2600", "ggml-cuda.cu:1826-1861"):
        assert cite in text


def test_shard_rows_partition():
    """Equal shards rounded up to 128 rows (the GEMM M tile), covering [0, M) exactly once; the
    reference's own slicing is ggml-cuda.cu:2780-2797."""
    mod = load_binding()
    for M in (4096, 11008, 32000, 4672, 18176, 4544, 8192, 22016, 100, 128, 129):
        for world in (1, 2, 4, 8):
            prev = 0
            for r in range(world):
                lo, hi = mod.shard_rows(M, r, world)
                assert lo == prev and lo <= hi <= M
                if hi < M:
                    assert (hi - lo) % mod.SHARD_ALIGN == 0
                prev = hi
            assert prev == M
    assert mod.shard_rows(8192, 3, 8) == (3072, 4096)
    assert mod.shard_rows(22016, 7, 8) == (19712, 22016)      # 2816-row shards, short last one


def test_no_gpu_means_loud_failure():
    mod = load_binding()
    if mod.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(mod.B200Error) as e:
        mod.init(0)
    assert e.value.code == mod.ERR_NO_DEVICE
    assert "no CPU fallback" in str(e.value)
    # compute entry points refuse before init instead of computing anything on the host
    L = mod.lib()
    x = np.zeros((1, 64), np.float32)
    y = np.zeros((1, 4), np.float32)
    w = np.zeros((4, 2, 18), np.uint8)
    assert L.ggml_b200_upload_q4_0(w.ctypes.data_as(C.c_void_p), 4, 64, 36, 0, 4) is None
    assert L.ggml_b200_quantize_q8_0(x.ctypes.data_as(C.c_void_p), 1, 64, 256, y.ctypes.data_as(C.c_void_p)) == mod.ERR_STATE
    ops = (mod.SetOp * 1)()
    assert L.ggml_b200_set_create(C.cast(ops, C.c_void_p), 1) is None
    assert b"ggml_b200_init" in L.ggml_b200_last_error()
    assert L.ggml_b200_set_run(None) == mod.ERR_STATE


def test_product_never_imports_oracle():
    """The product path must not route through the oracle (or /root/reference) at run time."""
    pkg = os.path.join(ROOT, "llama.cpp-q_4_0_b200")
    for dp, _, fns in os.walk(pkg):
        if os.path.basename(dp) in ("build", "lib"):
            continue
        for fn in fns:
            if fn.endswith((".cu", ".cuh", ".h", ".c", ".cpp", ".py")):
                src = open(os.path.join(dp, fn), errors="replace").read()
                assert "import oracle" not in src and "oracle/" not in src.replace("oracle/_ref", "") or fn == "Makefile", fn
                assert "dlopen(\"/root/reference" not in src
