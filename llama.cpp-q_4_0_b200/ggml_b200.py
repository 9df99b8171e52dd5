"""ctypes binding of lib/libggml_b200.so (the C ABI in include/ggml-b200.h).

This is harness plumbing for tests/ and bench.py, not the product: the product is the shared
library, whose host side is C (the reference is C; see ggml_glue/ for the ggml_tensor glue).
Method names mirror the C entry points, which mirror the reference seams they replace
(ggml_cuda_transform_tensor -> upload_q4_0, ggml_cuda_mul_mat -> mul_mat_q4_0, ...).

There is no fallback of any kind: if the library is missing, or no sm_100 GPU is present,
loading / init raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libggml_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "ggml-b200.h")

OK = 0
ERR_NO_DEVICE, ERR_ARCH, ERR_CUDA, ERR_ARG, ERR_STATE = -1, -2, -3, -4, -5
LAYOUT_RAW, LAYOUT_PLANES = 0, 1
PATH_AUTO, PATH_GEMV, PATH_GEMM = 0, 1, 2
SHARD_ALIGN = 128
QK = 32

_vp = C.c_void_p
_i64 = C.c_int64
_sz = C.c_size_t


class B200Error(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"ggml-b200 error {code}: {msg}")
        self.code = code


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU or PyTorch fallback for this path)")
    L = C.CDLL(LIB_PATH)
    sig = {
        "ggml_b200_init": (C.c_int, [C.c_int]),
        "ggml_b200_shutdown": (None, []),
        "ggml_b200_device_count": (C.c_int, []),
        "ggml_b200_last_error": (C.c_char_p, []),
        "ggml_b200_sm_count": (C.c_int, []),
        "ggml_b200_set_repack": (None, [C.c_int]),
        "ggml_b200_get_repack": (C.c_int, []),
        "ggml_b200_set_gemm_min_n": (None, [C.c_int]),
        "ggml_b200_get_gemm_min_n": (C.c_int, []),
        "ggml_b200_set_path": (None, [C.c_int]),
        "ggml_b200_set_stream": (None, [_vp]),
        "ggml_b200_get_stream": (_vp, []),
        "ggml_b200_set_tuning": (C.c_int, [C.c_char_p, C.c_int]),
        "ggml_b200_host_malloc": (_vp, [_sz]),
        "ggml_b200_host_free": (None, [_vp]),
        "ggml_b200_upload_q4_0": (_vp, [_vp, _i64, _i64, _sz, _i64, _i64]),
        "ggml_b200_free_weight": (None, [_vp]),
        "ggml_b200_weight_rows": (_i64, [_vp]),
        "ggml_b200_weight_row_lo": (_i64, [_vp]),
        "ggml_b200_weight_M": (_i64, [_vp]),
        "ggml_b200_weight_K": (_i64, [_vp]),
        "ggml_b200_weight_layout": (C.c_int, [_vp]),
        "ggml_b200_weight_device_bytes": (_sz, [_vp]),
        "ggml_b200_download_q4_0": (C.c_int, [_vp, _vp]),
        "ggml_b200_quantize_q8_0": (C.c_int, [_vp, _i64, _i64, _sz, _vp]),
        "ggml_b200_quantize_q4_0": (C.c_int, [_vp, _i64, _i64, _sz, _vp]),
        "ggml_b200_dequantize_q4_0": (C.c_int, [_vp, _i64, _vp]),
        "ggml_b200_get_rows_q4_0": (C.c_int, [_vp, _vp, _i64, _vp, _sz]),
        "ggml_b200_mul_mat_q4_0": (C.c_int, [_vp, _vp, _i64, _sz, _vp, _sz]),
        "ggml_b200_mul_mat_q4_0_host": (C.c_int, [_vp, _vp, _i64, _sz, _vp, _sz]),
        "ggml_b200_mul_mat_q4_0_static": (C.c_int, [_vp, _vp, _i64, _sz, _vp, _sz]),
        "ggml_b200_upload_quant": (_vp, [C.c_int, _vp, _i64, _i64, _sz]),
        "ggml_b200_free_quant": (None, [_vp]),
        "ggml_b200_quant_type": (C.c_int, [_vp]),
        "ggml_b200_quant_bytes": (_sz, [_vp]),
        "ggml_b200_mul_mat_quant": (C.c_int, [_vp, _vp, _i64, _sz, _vp, _sz]),
        "ggml_b200_block_dots_quant": (C.c_int, [_vp, _vp, _i64, _sz, _vp, _vp, _sz]),
        "ggml_b200_quantize_dot_blocks": (C.c_int, [C.c_int, _vp, _i64, _i64, _sz, _vp]),
        "ggml_b200_static_max_elems": (_sz, []),
        "ggml_b200_block_dots_q4_0": (C.c_int, [_vp, _vp, _i64, _sz, _vp, C.c_int]),
        "ggml_b200_fused_checkpoint_q4_0": (C.c_int, [_vp, _vp, _vp, _vp]),
        "ggml_b200_comm_unique_id": (C.c_int, [_vp]),
        "ggml_b200_comm_init": (C.c_int, [C.c_int, C.c_int, _vp]),
        "ggml_b200_comm_shutdown": (None, []),
        "ggml_b200_comm_rank": (C.c_int, []),
        "ggml_b200_comm_world": (C.c_int, []),
        "ggml_b200_shard_rows": (None, [_i64, C.c_int, C.c_int, C.POINTER(_i64), C.POINTER(_i64)]),
        "ggml_b200_mul_mat_q4_0_sharded": (C.c_int, [_vp, _vp, _i64, _sz, _vp, _sz]),
        "ggml_b200_set_create": (_vp, [_vp, C.c_int]),
        "ggml_b200_set_create_sharded": (_vp, [_vp, C.c_int]),
        "ggml_b200_set_run": (C.c_int, [_vp]),
        "ggml_b200_set_free": (None, [_vp]),
        "ggml_b200_set_ops": (C.c_int, [_vp]),
        "ggml_b200_set_barriers": (C.c_int, [_vp]),
        "ggml_b200_set_slots": (C.c_int, [_vp]),
        "ggml_b200_set_warps": (C.c_int, [_vp]),
        "ggml_b200_set_plan": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
        "ggml_b200_set_linked_runs": (C.c_int, [_vp]),
        "ggml_b200_set_bytes": (C.c_uint64, [_vp]),
        "ggml_b200_set_checkpoint": (C.c_int, [_vp, _vp, C.c_int]),
        "ggml_b200_set_profile": (C.c_int, [_vp, _vp, C.c_int]),
        "ggml_b200_launch_count": (C.c_uint64, []),
        "ggml_b200_trace_begin": (C.c_int, [C.c_uint]),
        "ggml_b200_trace_end": (C.c_int, [_vp, C.c_uint]),
        "ggml_b200_gemm_trace_read": (C.c_int, [_vp, C.c_int]),
        "ggml_b200_dev_malloc": (_vp, [_sz]),
        "ggml_b200_dev_free": (None, [_vp]),
        "ggml_b200_memcpy_h2d": (C.c_int, [_vp, _vp, _sz]),
        "ggml_b200_memcpy_d2h": (C.c_int, [_vp, _vp, _sz]),
        "ggml_b200_synchronize": (C.c_int, []),
        "ggml_b200_memcpy_h2d_async": (C.c_int, [_vp, _vp, _sz]),
        "ggml_b200_memcpy_d2h_async": (C.c_int, [_vp, _vp, _sz]),
        "ggml_b200_rms_norm_f32": (C.c_int, [_vp, _vp, _i64, _i64, C.c_float]),
        "ggml_b200_mul_f32": (C.c_int, [_vp, _vp, _vp, _i64, _i64]),
        "ggml_b200_add_f32": (C.c_int, [_vp, _vp, _vp, _i64, _i64]),
        "ggml_b200_silu_f32": (C.c_int, [_vp, _vp, _i64]),
        "ggml_b200_silu_mul_f32": (C.c_int, [_vp, _vp, _vp, _i64, _vp]),
        "ggml_b200_rms_norm_mul_f32": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, C.c_float]),
        "ggml_b200_mark": (C.c_int, [C.c_int]),
        "ggml_b200_wait_mark": (C.c_int, [C.c_int]),
        "ggml_b200_rope_f32": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i64]),
        "ggml_b200_neigh_ok": (C.c_int, [_vp]),
        "ggml_b200_mul_mat_q4_0_add": (C.c_int, [_vp, _vp, _vp, _vp, _vp]),
        "ggml_b200_mul_mat_q4_0_swiglu": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp]),
        "ggml_b200_graph_begin": (C.c_int, []),
        "ggml_b200_graph_end": (C.c_int, [C.POINTER(_vp)]),
        "ggml_b200_graph_launch": (C.c_int, [_vp]),
        "ggml_b200_graph_free": (None, [_vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    return L


_L = None


def lib() -> C.CDLL:
    global _L
    if _L is None:
        _L = _load()
    return _L


def _check(rc: int) -> None:
    if rc != OK:
        raise B200Error(rc, lib().ggml_b200_last_error().decode(errors="replace"))


def shard_rows(M: int, rank: int, world: int) -> tuple[int, int]:
    """Row range rank owns (pure host arithmetic; usable without a GPU)."""
    lo, hi = _i64(), _i64()
    lib().ggml_b200_shard_rows(M, rank, world, C.byref(lo), C.byref(hi))
    return lo.value, hi.value


def device_count() -> int:
    return lib().ggml_b200_device_count()


def init(device: int = 0) -> None:
    _check(lib().ggml_b200_init(device))


def shutdown() -> None:
    lib().ggml_b200_shutdown()


def set_repack(on: bool) -> None:
    lib().ggml_b200_set_repack(1 if on else 0)


def set_path(path: int) -> None:
    lib().ggml_b200_set_path(path)


def set_gemm_min_n(n: int) -> None:
    lib().ggml_b200_set_gemm_min_n(n)


def set_tuning(key: str, value: int) -> None:
    _check(lib().ggml_b200_set_tuning(key.encode(), value))


def set_stream(handle: int | None) -> None:
    lib().ggml_b200_set_stream(_vp(handle) if handle else None)


def launch_count() -> int:
    return int(lib().ggml_b200_launch_count())


def sm_count() -> int:
    return lib().ggml_b200_sm_count()


def synchronize() -> None:
    _check(lib().ggml_b200_synchronize())


class DevBuf:
    """A raw device allocation owned by the library's allocator helpers."""

    def __init__(self, nbytes: int):
        self.nbytes = int(nbytes)
        self.ptr = lib().ggml_b200_dev_malloc(self.nbytes)
        if not self.ptr:
            raise B200Error(ERR_CUDA, lib().ggml_b200_last_error().decode(errors="replace"))

    @classmethod
    def from_numpy(cls, a: np.ndarray) -> "DevBuf":
        a = np.ascontiguousarray(a)
        b = cls(a.nbytes)
        _check(lib().ggml_b200_memcpy_h2d(b.ptr, a.ctypes.data_as(_vp), a.nbytes))
        return b

    def to_numpy(self, dtype, shape) -> np.ndarray:
        out = np.empty(shape, dtype=dtype)
        assert out.nbytes <= self.nbytes
        _check(lib().ggml_b200_memcpy_d2h(out.ctypes.data_as(_vp), self.ptr, out.nbytes))
        return out

    def free(self) -> None:
        if self.ptr:
            lib().ggml_b200_dev_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Weight:
    """A device-resident Q4_0 row shard (ggml_b200_upload_q4_0)."""

    def __init__(self, blocks: np.ndarray, row_lo: int = 0, row_hi: int | None = None):
        """blocks: uint8 [M][K/32][18] (the reference's block_q4_0 rows, contiguous)."""
        blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
        assert blocks.ndim == 3 and blocks.shape[2] == 18
        M, nbk = blocks.shape[0], blocks.shape[1]
        self.M, self.K = M, nbk * QK
        if row_hi is None:
            row_hi = M
        self.h = lib().ggml_b200_upload_q4_0(blocks.ctypes.data_as(_vp), M, self.K, nbk * 18, row_lo, row_hi)
        if not self.h:
            raise B200Error(ERR_ARG, lib().ggml_b200_last_error().decode(errors="replace"))
        self.row_lo, self.row_hi = row_lo, row_hi

    @property
    def rows(self) -> int:
        return int(lib().ggml_b200_weight_rows(self.h))

    @property
    def layout(self) -> int:
        return int(lib().ggml_b200_weight_layout(self.h))

    @property
    def device_bytes(self) -> int:
        return int(lib().ggml_b200_weight_device_bytes(self.h))

    def download(self) -> np.ndarray:
        out = np.empty((self.rows, self.K // QK, 18), dtype=np.uint8)
        _check(lib().ggml_b200_download_q4_0(self.h, out.ctypes.data_as(_vp)))
        return out

    def free(self) -> None:
        if self.h:
            lib().ggml_b200_free_weight(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


TYPE_Q4_1, TYPE_Q5_0, TYPE_Q5_1, TYPE_Q8_0 = 3, 6, 7, 8          # the reference's ggml_type values (ggml.h:230-242)
QUANT_BLOCK_BYTES = {TYPE_Q4_1: 20, TYPE_Q5_0: 22, TYPE_Q5_1: 24, TYPE_Q8_0: 34}
DOT_BLOCK_BYTES = {TYPE_Q4_1: 40, TYPE_Q5_0: 34, TYPE_Q5_1: 40, TYPE_Q8_0: 34}     # block_q8_1 (formats with a min term) / block_q8_0


class QuantWeight:
    """A device-resident Q8_0 / Q4_1 / Q5_0 / Q5_1 weight (ggml_b200_upload_quant): the op's sibling formats."""

    def __init__(self, qtype: int, blocks: np.ndarray):
        blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
        assert blocks.ndim == 3 and blocks.shape[2] == QUANT_BLOCK_BYTES[qtype]
        self.type, self.M, self.K = qtype, blocks.shape[0], blocks.shape[1] * QK
        self.h = lib().ggml_b200_upload_quant(qtype, blocks.ctypes.data_as(_vp), self.M, self.K, blocks.shape[1] * blocks.shape[2])
        if not self.h:
            raise B200Error(ERR_ARG, lib().ggml_b200_last_error().decode(errors="replace"))

    @property
    def device_bytes(self) -> int:
        return int(lib().ggml_b200_quant_bytes(self.h))

    def mul_mat(self, X: np.ndarray, checkpoint: bool = False):
        """X [n][K] f32 -> Y [n][M] f32 (and S [n][M][K/32] int32 with checkpoint=True)."""
        X = np.ascontiguousarray(X, dtype=np.float32)
        n = X.shape[0]
        assert X.shape[1] == self.K
        xd = DevBuf.from_numpy(X) if n else DevBuf(16)
        yd = DevBuf(max(16, n * self.M * 4))
        if checkpoint:
            sd = DevBuf(max(16, n * self.M * (self.K // QK) * 4))
            _check(lib().ggml_b200_block_dots_quant(self.h, xd.ptr, n, self.K * 4, sd.ptr, yd.ptr, self.M * 4))
            synchronize()
            return yd.to_numpy(np.float32, (n, self.M)), sd.to_numpy(np.int32, (n, self.M, self.K // QK))
        _check(lib().ggml_b200_mul_mat_quant(self.h, xd.ptr, n, self.K * 4, yd.ptr, self.M * 4))
        synchronize()
        return yd.to_numpy(np.float32, (n, self.M))

    def free(self) -> None:
        if self.h:
            lib().ggml_b200_free_quant(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def quantize_dot_blocks(qtype: int, x: np.ndarray) -> np.ndarray:
    """The activation blocks the sibling kernels' fused quantizer forms, reference format: uint8 [n][K/32][34 | 40]."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, K = x.shape
    xd = DevBuf.from_numpy(x)
    out = DevBuf(n * (K // QK) * DOT_BLOCK_BYTES[qtype])
    _check(lib().ggml_b200_quantize_dot_blocks(qtype, xd.ptr, n, K, K * 4, out.ptr))
    synchronize()
    return out.to_numpy(np.uint8, (n, K // QK, DOT_BLOCK_BYTES[qtype]))


def quantize_q8_0(x: np.ndarray) -> np.ndarray:
    """Kernel 1 on host data: f32 [n][K] -> reference-format blocks uint8 [n][K/32][34]."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, K = x.shape
    xd = DevBuf.from_numpy(x)
    out = DevBuf(max(1, n * (K // QK) * 34))
    _check(lib().ggml_b200_quantize_q8_0(xd.ptr, n, K, K * 4, out.ptr))
    synchronize()
    res = out.to_numpy(np.uint8, (n, K // QK, 34))
    xd.free()
    out.free()
    return res


def quantize_q4_0(x: np.ndarray) -> np.ndarray:
    """Weight quantizer on host data: f32 [M][K] -> reference-format blocks uint8 [M][K/32][18]."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    M, K = x.shape
    xd = DevBuf.from_numpy(x)
    out = DevBuf(max(1, M * (K // QK) * 18))
    _check(lib().ggml_b200_quantize_q4_0(xd.ptr, M, K, K * 4, out.ptr))
    synchronize()
    res = out.to_numpy(np.uint8, (M, K // QK, 18))
    xd.free()
    out.free()
    return res


def dequantize_q4_0(blocks: np.ndarray) -> np.ndarray:
    """uint8 [..., 18] blocks -> f32 [nblocks*32]."""
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
    nblk = blocks.size // 18
    bd = DevBuf.from_numpy(blocks)
    out = DevBuf(max(1, nblk * QK * 4))
    _check(lib().ggml_b200_dequantize_q4_0(bd.ptr, nblk, out.ptr))
    synchronize()
    res = out.to_numpy(np.float32, (nblk * QK,))
    bd.free()
    out.free()
    return res


def get_rows(w: Weight, ids: np.ndarray) -> np.ndarray:
    """Rows `ids` of an uploaded Q4_0 tensor, dequantized: f32 [len(ids)][K]."""
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    idd = DevBuf.from_numpy(ids) if ids.size else DevBuf(4)
    out = DevBuf(max(1, ids.size * w.K * 4))
    _check(lib().ggml_b200_get_rows_q4_0(w.h, idd.ptr, ids.size, out.ptr, w.K * 4))
    synchronize()
    res = out.to_numpy(np.float32, (ids.size, w.K))
    idd.free()
    out.free()
    return res


def mul_mat_host(w: Weight, x: np.ndarray, y: np.ndarray | None = None) -> np.ndarray:
    """The reference-facing call: host x [n][K] f32 -> host y [n][M] f32 (this shard's columns)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, K = x.shape
    assert K == w.K
    if y is None:
        y = np.zeros((n, w.M), dtype=np.float32)
    _check(lib().ggml_b200_mul_mat_q4_0_host(w.h, x.ctypes.data_as(_vp), n, K * 4,
                                             y.ctypes.data_as(_vp), w.M * 4))
    return y


def mul_mat_dev(w: Weight, x_dev_ptr: int, n: int, y_dev_ptr: int, nb11: int | None = None,
                nb1: int | None = None) -> None:
    _check(lib().ggml_b200_mul_mat_q4_0(w.h, _vp(x_dev_ptr), n, nb11 or w.K * 4, _vp(y_dev_ptr), nb1 or w.M * 4))


def mul_mat_add(w: Weight, x: np.ndarray, r: np.ndarray):
    """n = 1: (W x, W x + r) from ONE kernel (ggml_b200_mul_mat_q4_0_add; the residual add behind wo / w2)."""
    xd, rd = DevBuf.from_numpy(np.ascontiguousarray(x, np.float32)), DevBuf.from_numpy(np.ascontiguousarray(r, np.float32))
    yd, y2d = DevBuf(w.M * 4), DevBuf(w.M * 4)
    _check(lib().ggml_b200_mul_mat_q4_0_add(w.h, xd.ptr, yd.ptr, rd.ptr, y2d.ptr))
    synchronize()
    return yd.to_numpy(np.float32, (w.M,)), y2d.to_numpy(np.float32, (w.M,))


def mul_mat_swiglu(w1: Weight, w3: Weight, x: np.ndarray, want_silu: bool = True):
    """n = 1: (W1 x, W3 x, silu(W1 x), silu(W1 x) * W3 x) from ONE kernel (ggml_b200_mul_mat_q4_0_swiglu; the FFN gate)."""
    xd = DevBuf.from_numpy(np.ascontiguousarray(x, np.float32))
    y1, y3, ys, y2 = (DevBuf(w1.M * 4) for _ in range(4))
    _check(lib().ggml_b200_mul_mat_q4_0_swiglu(w1.h, w3.h, xd.ptr, y1.ptr, y3.ptr, ys.ptr if want_silu else None, y2.ptr))
    synchronize()
    out = (y1.to_numpy(np.float32, (w1.M,)), y3.to_numpy(np.float32, (w3.M,)), ys.to_numpy(np.float32, (w1.M,)) if want_silu else None,
           y2.to_numpy(np.float32, (w1.M,)))
    return out


def rms_norm(x: np.ndarray, w: np.ndarray | None = None, eps: float = 1e-6):
    """rows of x normalised (ggml_b200_rms_norm_f32); with w also the product with the norm weights (…_rms_norm_mul_f32).
    Returns y, or (y, y * w)."""
    x = np.ascontiguousarray(x, np.float32)
    ncols, nrows = x.shape[-1], x.size // x.shape[-1]
    xd, yd = DevBuf.from_numpy(x), DevBuf(x.nbytes)
    if w is None:
        _check(lib().ggml_b200_rms_norm_f32(xd.ptr, yd.ptr, ncols, nrows, C.c_float(eps)))
        synchronize()
        return yd.to_numpy(np.float32, x.shape)
    w = np.ascontiguousarray(w, np.float32)
    wd, y2d = DevBuf.from_numpy(w), DevBuf(x.nbytes)
    _check(lib().ggml_b200_rms_norm_mul_f32(xd.ptr, yd.ptr, wd.ptr, y2d.ptr, ncols, nrows, w.size, C.c_float(eps)))
    synchronize()
    return yd.to_numpy(np.float32, x.shape), y2d.to_numpy(np.float32, x.shape)


def eltwise(op: str, a: np.ndarray, b: np.ndarray | None = None):
    """The stand-alone element-wise kernels (add | mul | silu | silu_mul) on host arrays: the unfused twins in the tests."""
    a = np.ascontiguousarray(a, np.float32)
    ad, yd = DevBuf.from_numpy(a), DevBuf(a.size * 4)
    bd = DevBuf.from_numpy(np.ascontiguousarray(b, np.float32)) if b is not None else None
    if op == "add":
        _check(lib().ggml_b200_add_f32(ad.ptr, bd.ptr, yd.ptr, a.size, b.size))
    elif op == "mul":
        _check(lib().ggml_b200_mul_f32(ad.ptr, bd.ptr, yd.ptr, a.size, b.size))
    elif op == "silu":
        _check(lib().ggml_b200_silu_f32(ad.ptr, yd.ptr, a.size))
    elif op == "silu_mul":
        _check(lib().ggml_b200_silu_mul_f32(ad.ptr, bd.ptr, yd.ptr, a.size, None))
    else:
        raise ValueError(op)
    synchronize()
    return yd.to_numpy(np.float32, a.shape)


def mul_mat_sharded_dev(w: Weight, x_dev_ptr: int, n: int, y_dev_ptr: int) -> None:
    _check(lib().ggml_b200_mul_mat_q4_0_sharded(w.h, _vp(x_dev_ptr), n, w.K * 4, _vp(y_dev_ptr), w.M * 4))


class SetOp(C.Structure):
    """struct ggml_b200_set_op"""
    _fields_ = [("w", _vp), ("x_dev", _vp), ("y_dev", _vp)]


class SetPlanOp(C.Structure):
    """struct ggml_b200_set_plan_op"""
    _fields_ = [("x", C.c_uint64), ("y", C.c_uint64), ("K", C.c_int64), ("M", C.c_int64), ("row_lo", C.c_int64), ("row_hi", C.c_int64)]


SET_OP_BARRIER, SET_OP_QUANT, SET_OP_XEXT = 1, 2, 4


def set_plan(ops):
    """The hazard analysis of ggml_b200_set_create alone (host code, no device).
    ops: iterable of (x address, y address, M, K[, row_lo, row_hi]).
    Returns (barriers, linked_runs, [dict(barrier, new_run, ext, arrivals, link)] per op)."""
    ops = [tuple(o) + ((0, o[2]) if len(o) == 4 else ()) for o in ops]
    arr = (SetPlanOp * max(1, len(ops)))()
    for i, (x, y, M, K, lo, hi) in enumerate(ops):
        arr[i] = SetPlanOp(x, y, K, M, lo, hi)
    flags = (C.c_int * max(1, len(ops)))()
    link = (C.c_int * max(1, len(ops)))()
    linked = C.c_int(0)
    rc = lib().ggml_b200_set_plan(C.cast(arr, _vp), len(ops), flags, link, C.byref(linked))
    _check(min(rc, 0))
    out = [dict(barrier=bool(flags[i] & SET_OP_BARRIER), new_run=bool(flags[i] & SET_OP_QUANT), ext=bool(flags[i] & SET_OP_XEXT),
                arrivals=(flags[i] >> 8) & 0x7F, link=link[i]) for i in range(len(ops))]
    return rc, linked.value, out


class OpSet:
    """A list of n = 1 ops compiled for ONE persistent launch (ggml_b200_set_create).
    ops: iterable of (weight handle or Weight, x device pointer, y device pointer)."""

    def __init__(self, ops, sharded=False):
        """sharded: collective over the ranks of comm_init (ggml_b200_set_create_sharded): row shards, full dst on every rank"""
        ops = list(ops)
        arr = (SetOp * max(1, len(ops)))()
        for i, (w, xp, yp) in enumerate(ops):
            arr[i].w = w.h if isinstance(w, Weight) else w
            arr[i].x_dev = xp
            arr[i].y_dev = yp
        self.h = (lib().ggml_b200_set_create_sharded if sharded else lib().ggml_b200_set_create)(C.cast(arr, _vp), len(ops))
        if not self.h:
            raise B200Error(ERR_ARG, lib().ggml_b200_last_error().decode(errors="replace"))

    n_ops = property(lambda self: int(lib().ggml_b200_set_ops(self.h)))
    barriers = property(lambda self: int(lib().ggml_b200_set_barriers(self.h)))
    slots = property(lambda self: int(lib().ggml_b200_set_slots(self.h)))
    warps = property(lambda self: int(lib().ggml_b200_set_warps(self.h)))
    linked_runs = property(lambda self: int(lib().ggml_b200_set_linked_runs(self.h)))
    bytes = property(lambda self: int(lib().ggml_b200_set_bytes(self.h)))

    def run(self) -> None:
        _check(lib().ggml_b200_set_run(self.h))

    def checkpoint(self, S_ptrs) -> None:
        """S_ptrs: one device pointer (or None) per non-empty op: int32 [rows][K/32]."""
        arr = (_vp * len(S_ptrs))(*[_vp(p) if p else None for p in S_ptrs])
        _check(lib().ggml_b200_set_checkpoint(self.h, C.cast(arr, _vp), len(S_ptrs)))

    def profile(self) -> np.ndarray:
        """int64 [ctas][16] clock totals of one instrumented run (see ggml_b200_set_profile)."""
        out = np.zeros((sm_count(), 16), dtype=np.int64)
        n = lib().ggml_b200_set_profile(self.h, out.ctypes.data_as(_vp), out.shape[0])
        if n < 0:
            _check(n)
        return out[:n]

    def free(self) -> None:
        if self.h:
            lib().ggml_b200_set_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def block_dots(w: Weight, x: np.ndarray, path: int = PATH_GEMV) -> np.ndarray:
    """int32 S_b [n][rows][K/32] as the chosen kernel forms them."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, K = x.shape
    nbk = K // QK
    xd = DevBuf.from_numpy(x)
    S = DevBuf(max(4, n * w.rows * nbk * 4))
    _check(lib().ggml_b200_block_dots_q4_0(w.h, xd.ptr, n, K * 4, S.ptr, path))
    synchronize()
    out = S.to_numpy(np.int32, (n, w.rows, nbk))
    xd.free()
    S.free()
    return out


def fused_checkpoint(w: Weight, x: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    """(q8_0 blocks [K/32][34], S_b [rows][K/32]) as the fused n = 1 kernel forms them."""
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
    nbk = x.size // QK
    xd = DevBuf.from_numpy(x)
    S = DevBuf(max(4, w.rows * nbk * 4))
    B = DevBuf(nbk * 34)
    _check(lib().ggml_b200_fused_checkpoint_q4_0(w.h, xd.ptr, S.ptr, B.ptr))
    synchronize()
    out = (B.to_numpy(np.uint8, (nbk, 34)), S.to_numpy(np.int32, (w.rows, nbk)))
    for b in (xd, S, B):
        b.free()
    return out


def comm_unique_id() -> bytes:
    buf = C.create_string_buffer(128)
    _check(lib().ggml_b200_comm_unique_id(buf))
    return buf.raw


def comm_init(rank: int, world: int, uid: bytes) -> None:
    buf = C.create_string_buffer(uid, 128)
    _check(lib().ggml_b200_comm_init(rank, world, buf))


def comm_shutdown() -> None:
    lib().ggml_b200_comm_shutdown()
