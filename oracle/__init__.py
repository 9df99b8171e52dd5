"""ctypes front door onto the CPU oracle.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package, and only as the checker (or the timed CPU arm).  The product
(llama.cpp-q_4_0_b200/) never does.

Two libraries:
  * ``port``  -- oracle/_build/libq4_0_oracle.so, our scalar C restatement (q4_0_oracle.c),
                 built on demand with gcc.
  * ``ref``   -- oracle/_ref/libggml_ref.so, the UNMODIFIED reference ggml.c + ref_shim.c, built
                 by oracle/Makefile where /root/reference exists; travels prebuilt to the GPU
                 box.  ``ref()`` returns None when it is not available.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PORT_SO = os.path.join(_HERE, "_build", "libq4_0_oracle.so")
_REF_SO = os.path.join(_HERE, "_ref", "libggml_ref.so")                 # -march=x86-64-v3
_REF_NATIVE_SO = os.path.join(_HERE, "_ref", "libggml_ref_native.so")   # -march=native of the build host
_REF_NATIVE_FLAGS = os.path.join(_HERE, "_ref", "native_flags.txt")


def _native_ok() -> bool:
    """The -march=native build may only run where every CPU flag of its build host exists."""
    try:
        need = set(open(_REF_NATIVE_FLAGS).read().split())
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return need <= set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return False

Q4_0_BYTES = 18
Q8_0_BYTES = 34
QK = 32

_f32p = C.POINTER(C.c_float)
_vp = C.c_void_p


def build(force: bool = False) -> None:
    """Compile the C restatement (always) and oracle/_ref (only where the reference is present)."""
    src = os.path.join(_HERE, "q4_0_oracle.c")
    need = force or not os.path.exists(_PORT_SO) or os.path.getmtime(_PORT_SO) < os.path.getmtime(src)
    ref_present = os.path.exists("/root/reference/ggml.c")
    need_ref = ref_present and (force or not os.path.exists(_REF_SO) or not os.path.exists(_REF_NATIVE_SO)
                                or os.path.getmtime(_REF_SO) < os.path.getmtime(os.path.join(_HERE, "ref_shim.c")))
    if need or need_ref:
        subprocess.run(["make", "-C", _HERE, "all"], check=True, stdout=subprocess.DEVNULL)


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(_vp)


class Port:
    """The scalar C restatement (q4_0_oracle.c)."""

    def __init__(self):
        build()
        L = C.CDLL(_PORT_SO)
        self.L = L
        L.oracle_fp16_to_fp32.restype = C.c_float
        L.oracle_fp16_to_fp32.argtypes = [C.c_uint16]
        L.oracle_fp32_to_fp16.restype = C.c_uint16
        L.oracle_fp32_to_fp16.argtypes = [C.c_float]
        for name in ("oracle_quantize_row_q4_0", "oracle_quantize_row_q8_0",
                     "oracle_quantize_row_q8_0_scalar_reference"):
            getattr(L, name).restype = None
            getattr(L, name).argtypes = [_vp, _vp, C.c_int64]
        L.oracle_dequantize_row_q4_0.restype = None
        L.oracle_dequantize_row_q4_0.argtypes = [_vp, _vp, C.c_int64]
        L.oracle_block_dots_q4_0_q8_0.restype = None
        L.oracle_block_dots_q4_0_q8_0.argtypes = [C.c_int64, _vp, _vp, _vp]
        for name in ("oracle_vec_dot_q4_0_q8_0_scalar", "oracle_vec_dot_q4_0_q8_0_avx2_order"):
            getattr(L, name).restype = C.c_float
            getattr(L, name).argtypes = [C.c_int64, _vp, _vp]
        L.oracle_mul_mat_q4_0_f32.restype = None
        L.oracle_mul_mat_q4_0_f32.argtypes = [C.c_int64, C.c_int64, C.c_int64, _vp, C.c_size_t,
                                              _vp, C.c_size_t, _vp, C.c_size_t, _vp, C.c_int]
        L.oracle_fill_synthetic.restype = None
        L.oracle_fill_synthetic.argtypes = [_vp, C.c_int64, C.c_uint32, C.c_float, C.c_int]
        L.oracle_fill_cos.restype = None
        L.oracle_fill_cos.argtypes = [_vp, C.c_int64, C.c_float]
        L.oracle_fnv1a64.restype = C.c_uint64
        L.oracle_fnv1a64.argtypes = [_vp, C.c_size_t]

    # -- scalars
    def fp16_to_fp32(self, h: int) -> float:
        return float(self.L.oracle_fp16_to_fp32(h))

    def fp32_to_fp16(self, f: float) -> int:
        return int(self.L.oracle_fp32_to_fp16(C.c_float(f)))

    # -- generators
    def synthetic(self, count: int, seed: int, scale: float = 1.0, outlier_every: int = 0) -> np.ndarray:
        out = np.empty(count, dtype=np.float32)
        self.L.oracle_fill_synthetic(_ptr(out), count, seed & 0xFFFFFFFF, scale, outlier_every)
        return out

    def cos_data(self, count: int, offset: float) -> np.ndarray:
        out = np.empty(count, dtype=np.float32)
        self.L.oracle_fill_cos(_ptr(out), count, offset)
        return out

    def fnv(self, a: np.ndarray) -> int:
        a = np.ascontiguousarray(a)
        return int(self.L.oracle_fnv1a64(_ptr(a), a.nbytes))

    # -- codecs: x is [..., K] f32 contiguous; returns uint8 [..., K/32, 18|34]
    def quantize_q4_0(self, x: np.ndarray) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        out = np.empty(x.shape[:-1] + (x.shape[-1] // QK, Q4_0_BYTES), dtype=np.uint8)
        self.L.oracle_quantize_row_q4_0(_ptr(x), _ptr(out), x.size)
        return out

    def dequantize_q4_0(self, w: np.ndarray) -> np.ndarray:
        w = np.ascontiguousarray(w, dtype=np.uint8)
        nblk = w.size // Q4_0_BYTES
        out = np.empty(w.shape[:-2] + (w.shape[-2] * QK,), dtype=np.float32)
        self.L.oracle_dequantize_row_q4_0(_ptr(w), _ptr(out), nblk * QK)
        return out

    def quantize_q8_0(self, x: np.ndarray, scalar_reference: bool = False) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        out = np.empty(x.shape[:-1] + (x.shape[-1] // QK, Q8_0_BYTES), dtype=np.uint8)
        fn = self.L.oracle_quantize_row_q8_0_scalar_reference if scalar_reference else self.L.oracle_quantize_row_q8_0
        fn(_ptr(x), _ptr(out), x.size)
        return out

    def block_dots(self, w_row: np.ndarray, q8_row: np.ndarray) -> np.ndarray:
        """int32 S_b for one weight row (nbk,18) against one q8 row (nbk,34)."""
        w_row = np.ascontiguousarray(w_row, dtype=np.uint8)
        q8_row = np.ascontiguousarray(q8_row, dtype=np.uint8)
        nbk = w_row.size // Q4_0_BYTES
        S = np.empty(nbk, dtype=np.int32)
        self.L.oracle_block_dots_q4_0_q8_0(nbk * QK, _ptr(w_row), _ptr(q8_row), _ptr(S))
        return S

    def block_dots_matrix(self, W: np.ndarray, Q8: np.ndarray) -> np.ndarray:
        """S[n][M][nbk] via numpy (vectorised restatement of ggml.c:2593-2600)."""
        W = np.ascontiguousarray(W, dtype=np.uint8)        # [M][nbk][18]
        Q8 = np.ascontiguousarray(Q8, dtype=np.uint8)      # [n][nbk][34]
        qs = W[:, :, 2:]
        lo = (qs & 0x0F).astype(np.int32) - 8               # elements 0..15
        hi = (qs >> 4).astype(np.int32) - 8                 # elements 16..31
        wq = np.concatenate([lo, hi], axis=-1)              # [M][nbk][32]
        yq = Q8[:, :, 2:].view(np.int8).astype(np.int32)    # [n][nbk][32]
        return np.einsum("mbk,nbk->nmb", wq, yq, optimize=True).astype(np.int32)

    def vec_dot(self, w_row: np.ndarray, q8_row: np.ndarray, avx2_order: bool = False) -> float:
        w_row = np.ascontiguousarray(w_row, dtype=np.uint8)
        q8_row = np.ascontiguousarray(q8_row, dtype=np.uint8)
        k = (w_row.size // Q4_0_BYTES) * QK
        fn = self.L.oracle_vec_dot_q4_0_q8_0_avx2_order if avx2_order else self.L.oracle_vec_dot_q4_0_q8_0_scalar
        return float(fn(k, _ptr(w_row), _ptr(q8_row)))

    def mul_mat(self, W: np.ndarray, X: np.ndarray, avx2_order: bool = True) -> np.ndarray:
        """W [M][nbk][18] u8, X [n][K] f32 -> Y [n][M] f32 (the op, ggml.c:11353-11411)."""
        W = np.ascontiguousarray(W, dtype=np.uint8)
        X = np.ascontiguousarray(X, dtype=np.float32)
        M, nbk = W.shape[0], W.shape[1]
        n, K = X.shape
        assert K == nbk * QK
        Y = np.empty((n, M), dtype=np.float32)
        wdata = np.empty(n * nbk * Q8_0_BYTES, dtype=np.uint8)
        self.L.oracle_mul_mat_q4_0_f32(M, K, n, _ptr(W), nbk * Q4_0_BYTES, _ptr(X), K * 4,
                                       _ptr(Y), M * 4, _ptr(wdata), 1 if avx2_order else 0)
        return Y


class Ref:
    """The reference itself (unmodified ggml.c) behind ref_shim.c."""

    def __init__(self):
        self.native = os.path.exists(_REF_NATIVE_SO) and _native_ok()
        L = C.CDLL(_REF_NATIVE_SO if self.native else _REF_SO)
        self.L = L
        L.ref_init.restype = None
        L.ref_fp16_to_fp32.restype = C.c_float
        L.ref_fp16_to_fp32.argtypes = [C.c_uint16]
        L.ref_fp32_to_fp16.restype = C.c_uint16
        L.ref_fp32_to_fp16.argtypes = [C.c_float]
        for name in ("ref_quantize_row_q4_0", "ref_quantize_row_q8_0", "ref_quantize_row_q8_0_reference"):
            getattr(L, name).restype = None
            getattr(L, name).argtypes = [_vp, _vp, C.c_int]
        L.ref_dequantize_row_q4_0.restype = None
        L.ref_dequantize_row_q4_0.argtypes = [_vp, _vp, C.c_int]
        L.ref_vec_dot_q4_0_q8_0.restype = C.c_float
        L.ref_vec_dot_q4_0_q8_0.argtypes = [C.c_int, _vp, _vp]
        if hasattr(L, "ref_time_vec_dot_q4_0_q8_0"):
            L.ref_time_vec_dot_q4_0_q8_0.restype = C.c_int64
            L.ref_time_vec_dot_q4_0_q8_0.argtypes = [C.c_int, _vp, C.c_int64, _vp, C.c_int, C.c_int, _vp]
        L.ref_quantize_q4_0.restype = C.c_size_t
        L.ref_quantize_q4_0.argtypes = [_vp, _vp, C.c_int, C.c_int]
        L.ref_mm_create.restype = _vp
        L.ref_mm_create.argtypes = [C.c_int64, C.c_int64, C.c_int64, _vp]
        L.ref_mm_run.restype = None
        L.ref_mm_run.argtypes = [_vp, _vp, _vp, C.c_int]
        L.ref_mm_compute.restype = None
        L.ref_mm_compute.argtypes = [_vp, C.c_int]
        L.ref_mm_x.restype = _vp
        L.ref_mm_x.argtypes = [_vp]
        L.ref_mm_y.restype = _vp
        L.ref_mm_y.argtypes = [_vp]
        L.ref_mm_free.restype = None
        L.ref_mm_free.argtypes = [_vp]
        if hasattr(L, "ref_type_vec_dot"):          # sibling formats (ref_shim.c, second half)
            L.ref_type_block_bytes.restype = C.c_int
            L.ref_type_block_bytes.argtypes = [C.c_int]
            L.ref_type_dot_block_bytes.restype = C.c_int
            L.ref_type_dot_block_bytes.argtypes = [C.c_int]
            for name in ("ref_type_quantize_row", "ref_type_dequantize_row", "ref_type_quantize_row_dot"):
                getattr(L, name).restype = None
                getattr(L, name).argtypes = [C.c_int, _vp, _vp, C.c_int]
            L.ref_type_vec_dot.restype = C.c_float
            L.ref_type_vec_dot.argtypes = [C.c_int, C.c_int, _vp, _vp]
            L.ref_mm_create_type.restype = _vp
            L.ref_mm_create_type.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_int64, _vp]
        if hasattr(L, "ref_neighbour_op"):          # element-wise nodes next to the op (ref_shim.c, last part)
            L.ref_neighbour_op.restype = C.c_int
            L.ref_neighbour_op.argtypes = [C.c_int, _vp, _vp, _vp, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int]
        L.ref_init()

    # -- the element-wise nodes next to the op, by the reference's own graph compute
    NEIGHBOUR_OPS = {"add": 0, "mul": 1, "silu": 2, "rms_norm": 3, "rope": 4}

    def neighbour(self, op: str, a: np.ndarray, b: np.ndarray | None = None, n_past: int = 0, n_threads: int = 1) -> np.ndarray:
        """a: f32 [ne2][ne1][ne0] (any leading dims are padded to three); rope: mode 0 over whole rows, positions n_past + i2."""
        a = np.ascontiguousarray(a, dtype=np.float32)
        sh = (1,) * (3 - a.ndim) + a.shape
        y = np.empty_like(a)
        bb = np.ascontiguousarray(b, dtype=np.float32) if b is not None else a
        rc = self.L.ref_neighbour_op(self.NEIGHBOUR_OPS[op], _ptr(a), _ptr(bb), _ptr(y), sh[2], sh[1], sh[0], n_past, n_threads)
        if rc != 0:
            raise RuntimeError("ref_neighbour_op failed: %d" % rc)
        return y

    # -- sibling formats: the reference's own functions by ggml_type (3 = Q4_1, 8 = Q8_0)
    def type_quantize(self, qtype: int, x: np.ndarray) -> np.ndarray:
        """quantize_row_q of the type (the weight-side codec): [..., K] f32 -> uint8 [..., K/32, block bytes]."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        bs = int(self.L.ref_type_block_bytes(qtype))
        out = np.empty(x.shape[:-1] + (x.shape[-1] // QK, bs), dtype=np.uint8)
        self.L.ref_type_quantize_row(qtype, _ptr(x), _ptr(out), x.size)
        return out

    def type_quantize_dot(self, qtype: int, x: np.ndarray) -> np.ndarray:
        """quantize_row_q_dot of the type (q8_0 for Q8_0, q8_1 for Q4_1), the build host's SIMD branch."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        bs = int(self.L.ref_type_dot_block_bytes(qtype))
        out = np.empty(x.shape[:-1] + (x.shape[-1] // QK, bs), dtype=np.uint8)
        self.L.ref_type_quantize_row_dot(qtype, _ptr(x), _ptr(out), x.size)
        return out

    def type_vec_dot(self, qtype: int, w_row: np.ndarray, q_row: np.ndarray) -> float:
        w_row = np.ascontiguousarray(w_row, dtype=np.uint8)
        q_row = np.ascontiguousarray(q_row, dtype=np.uint8)
        k = w_row.shape[0] * QK
        return float(self.L.ref_type_vec_dot(qtype, k, _ptr(w_row), _ptr(q_row)))

    def type_mul_mat(self, qtype: int, W: np.ndarray, X: np.ndarray, n_threads: int = 1) -> np.ndarray:
        """The whole operator through the reference's graph API for a weight of `qtype`."""
        W = np.ascontiguousarray(W, dtype=np.uint8)
        X = np.ascontiguousarray(X, dtype=np.float32)
        M, nbk = W.shape[0], W.shape[1]
        n, K = X.shape
        assert K == nbk * QK
        h = self.L.ref_mm_create_type(qtype, M, K, n, _ptr(W))
        assert h
        Y = np.empty((n, M), dtype=np.float32)
        self.L.ref_mm_run(h, _ptr(X), _ptr(Y), n_threads)
        self.L.ref_mm_free(h)
        return Y

    def fp16_to_fp32(self, h: int) -> float:
        return float(self.L.ref_fp16_to_fp32(h))

    def fp32_to_fp16(self, f: float) -> int:
        return int(self.L.ref_fp32_to_fp16(C.c_float(f)))

    def quantize_q4_0(self, x: np.ndarray) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        out = np.empty(x.shape[:-1] + (x.shape[-1] // QK, Q4_0_BYTES), dtype=np.uint8)
        # int k in the reference API: go row by row to stay below 2^31
        rows = x.reshape(-1, x.shape[-1])
        o2 = out.reshape(rows.shape[0], -1)
        for i in range(rows.shape[0]):
            self.L.ref_quantize_row_q4_0(_ptr(rows[i]), _ptr(o2[i]), rows.shape[1])
        return out

    def quantize_q4_0_bulk(self, x: np.ndarray) -> np.ndarray:
        """ggml_quantize_q4_0 (ggml.c:19157), the call benchmark-matmult.cpp:182 uses."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        out = np.empty(x.shape[:-1] + (x.shape[-1] // QK, Q4_0_BYTES), dtype=np.uint8)
        rows = x.reshape(-1, x.shape[-1])
        o2 = out.reshape(rows.shape[0], -1)
        step = max(1, (1 << 30) // rows.shape[1])
        for i in range(0, rows.shape[0], step):
            blk = rows[i:i + step]
            self.L.ref_quantize_q4_0(_ptr(blk), _ptr(o2[i:i + step]), blk.size, rows.shape[1])
        return out

    def dequantize_q4_0(self, w: np.ndarray) -> np.ndarray:
        w = np.ascontiguousarray(w, dtype=np.uint8)
        nblk = w.size // Q4_0_BYTES
        out = np.empty(w.shape[:-2] + (w.shape[-2] * QK,), dtype=np.float32)
        self.L.ref_dequantize_row_q4_0(_ptr(w), _ptr(out), nblk * QK)
        return out

    def quantize_q8_0(self, x: np.ndarray, scalar_reference: bool = False) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        out = np.empty(x.shape[:-1] + (x.shape[-1] // QK, Q8_0_BYTES), dtype=np.uint8)
        rows = x.reshape(-1, x.shape[-1])
        o2 = out.reshape(rows.shape[0], -1)
        fn = self.L.ref_quantize_row_q8_0_reference if scalar_reference else self.L.ref_quantize_row_q8_0
        for i in range(rows.shape[0]):
            fn(_ptr(rows[i]), _ptr(o2[i]), rows.shape[1])
        return out

    def vec_dot(self, w_row: np.ndarray, q8_row: np.ndarray) -> float:
        w_row = np.ascontiguousarray(w_row, dtype=np.uint8)
        q8_row = np.ascontiguousarray(q8_row, dtype=np.uint8)
        k = (w_row.size // Q4_0_BYTES) * QK
        return float(self.L.ref_vec_dot_q4_0_q8_0(k, _ptr(w_row), _ptr(q8_row)))

    def time_vec_dot(self, W: np.ndarray, q8_row: np.ndarray, warmup: int = 5, iters: int = 10) -> float:
        """best-of-`iters` seconds of one single-thread pass of vec_dot_q over all rows of W (test-quantize-perf method)"""
        W = np.ascontiguousarray(W, dtype=np.uint8)
        q8_row = np.ascontiguousarray(q8_row, dtype=np.uint8)
        rows = W.shape[0]
        k = (W.size // rows // Q4_0_BYTES) * QK
        sink = C.c_float(0)
        us = self.L.ref_time_vec_dot_q4_0_q8_0(k, _ptr(W), rows, _ptr(q8_row), warmup, iters, C.byref(sink))
        return max(us, 1) * 1e-6

    def mul_mat(self, W: np.ndarray, X: np.ndarray, n_threads: int = 1) -> np.ndarray:
        """The whole op through ggml_graph_compute (ref_shim.c)."""
        W = np.ascontiguousarray(W, dtype=np.uint8)
        X = np.ascontiguousarray(X, dtype=np.float32)
        M, nbk = W.shape[0], W.shape[1]
        n, K = X.shape
        assert K == nbk * QK
        h = self.L.ref_mm_create(M, K, n, _ptr(W))
        assert h, "ref_mm_create failed"
        try:
            Y = np.empty((n, M), dtype=np.float32)
            self.L.ref_mm_run(h, _ptr(X), _ptr(Y), n_threads)
        finally:
            self.L.ref_mm_free(h)
        return Y

    class Handle:
        """Persistent op instance for timing (bench.py cpu_baseline / --impl reference)."""

        def __init__(self, ref: "Ref", W: np.ndarray, n: int):
            W = np.ascontiguousarray(W, dtype=np.uint8)
            self.ref = ref
            self.M, self.nbk = W.shape[0], W.shape[1]
            self.K = self.nbk * QK
            self.n = n
            self.h = ref.L.ref_mm_create(self.M, self.K, n, _ptr(W))
            assert self.h

        def set_x(self, X: np.ndarray) -> None:
            X = np.ascontiguousarray(X, dtype=np.float32)
            C.memmove(self.ref.L.ref_mm_x(self.h), _ptr(X), X.nbytes)

        def compute(self, n_threads: int) -> None:
            self.ref.L.ref_mm_compute(self.h, n_threads)

        def y(self) -> np.ndarray:
            buf = (C.c_float * (self.n * self.M)).from_address(self.ref.L.ref_mm_y(self.h))
            return np.frombuffer(buf, dtype=np.float32).reshape(self.n, self.M).copy()

        def close(self) -> None:
            if self.h:
                self.ref.L.ref_mm_free(self.h)
                self.h = None


_port = None
_ref = None
_ref_tried = False


def port() -> Port:
    global _port
    if _port is None:
        _port = Port()
    return _port


def ref():
    """The compiled reference, or None where oracle/_ref was never built."""
    global _ref, _ref_tried
    if not _ref_tried:
        _ref_tried = True
        try:
            build()
        except Exception:
            pass
        if os.path.exists(_REF_SO):
            try:
                _ref = Ref()
            except OSError:
                _ref = None
    return _ref
