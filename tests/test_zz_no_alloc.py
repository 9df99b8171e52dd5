"""The allocation-free entry point ggml_b200_mul_mat_q4_0_static (no counterpart in the
reference): same op as ggml_b200_mul_mat_q4_0 with kernel 1's output in a file-scope static __device__ array.
Collected last on purpose (a failure here must not hide the other tests)."""
import ctypes as C

import numpy as np
import pytest

from test_gpu_parity import TOL, assert_close, make_weights, make_x, oracle_Y

pytestmark = pytest.mark.gpu


def _call(b200, fn, w, xd, n, yd):
    rc = fn(w.h, C.c_void_p(xd.ptr), n, w.K * 4, C.c_void_p(yd.ptr), w.M * 4)
    assert rc == 0, b200.lib().ggml_b200_last_error()
    b200.synchronize()
    return yd.to_numpy(np.float32, (n, w.M)).copy()


@pytest.mark.parametrize("M,K,n", [(300, 1088, 64), (520, 2048, 5), (257, 1024, 1), (1024, 4096, 200)])
def test_static_scratch_entry_matches_the_allocating_one(b200, port, ref, M, K, n):
    """n = 64 / 200: kernel 1 + tensor-core GEMM; n = 5: kernel 1 + GEMV; n = 1: fused kernel (no scratch at all).
    Two calls in a row with different activations (the array keeps the previous call's contents), each compared with
    the existing entry point (tolerance of the parity tests; the bits are in fact identical) and with the oracle."""
    L_ = b200.lib()
    W = make_weights(port, ref, M, K, 2100 + n)
    w = b200.Weight(W)
    yd = b200.DevBuf(n * M * 4)
    for call in range(2):
        X = make_x(port, n, K, 7000 + 13 * call + n)
        xd = b200.DevBuf.from_numpy(X)
        want = _call(b200, L_.ggml_b200_mul_mat_q4_0, w, xd, n, yd)
        L_.ggml_b200_memcpy_h2d(yd.ptr, np.zeros(n * M, np.float32).ctypes.data_as(C.c_void_p), n * M * 4)
        got = _call(b200, L_.ggml_b200_mul_mat_q4_0_static, w, xd, n, yd)
        assert_close(got, want)
        assert float(np.max(np.abs(got - want))) <= TOL * float(np.max(np.abs(want)))
        assert np.array_equal(got, want)              # same kernels on the same bits
        assert_close(got, oracle_Y(port, ref, W, X))
        # and the allocating entry point still works after it (its workspace was put back)
        again = _call(b200, L_.ggml_b200_mul_mat_q4_0, w, xd, n, yd)
        assert np.array_equal(again, want)
    w.free()


def test_static_scratch_entry_rejects_what_does_not_fit(b200, port, ref):
    L_ = b200.lib()
    cap = L_.ggml_b200_static_max_elems()
    assert cap == 4 << 20
    K = 1024
    w = b200.Weight(make_weights(port, ref, 64, K, 2200))
    n = cap // K + 1
    xd = b200.DevBuf(n * K * 4)
    yd = b200.DevBuf(n * 64 * 4)
    rc = L_.ggml_b200_mul_mat_q4_0_static(w.h, C.c_void_p(xd.ptr), n, K * 4, C.c_void_p(yd.ptr), 64 * 4)
    assert rc == b200.ERR_ARG
    assert b"static scratch" in L_.ggml_b200_last_error()
    w.free()
