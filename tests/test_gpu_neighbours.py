"""The decode op with its neighbour folded into the store (gemv_q4_0_neigh.cuh) through the C ABI:
the matmul results bit-identical to ggml_b200_mul_mat_q4_0, the fused result bit-identical to the stand-alone element-wise
kernel applied to it, and both within the bars against the oracle."""
import numpy as np
import pytest

from oracle import neighbours as nb
from test_gpu_parity import make_weights, make_x, oracle_Y, assert_close

pytestmark = pytest.mark.gpu

SHAPES = [(256, 64), (4096, 4096), (1000, 4544), (11008, 4096), (4096, 11008)]


@pytest.mark.parametrize("M,K", SHAPES)
def test_mul_mat_add(b200, port, ref, M, K):
    W = make_weights(port, ref, M, K)
    x = make_x(port, 1, K)
    r = port.synthetic(M, 0xADD + M, 1.0).astype(np.float32)
    w = b200.Weight(W)
    y_plain = b200.mul_mat_host(w, x)[0]
    y, y2 = b200.mul_mat_add(w, x[0], r)
    assert np.array_equal(y.view(np.uint32), y_plain.view(np.uint32))
    assert np.array_equal(y2.view(np.uint32), b200.eltwise("add", y_plain, r).view(np.uint32))
    assert np.array_equal(y2.view(np.uint32), nb.add(y_plain, r).view(np.uint32))
    assert_close(y[None], oracle_Y(port, ref, W, x))
    # in place over the residual row, as ggml_add_inplace would ask for
    rd = b200.DevBuf.from_numpy(r)
    xd = b200.DevBuf.from_numpy(x[0])
    yd = b200.DevBuf(M * 4)
    b200._check(b200.lib().ggml_b200_mul_mat_q4_0_add(w.h, xd.ptr, yd.ptr, rd.ptr, rd.ptr))
    b200.synchronize()
    assert np.array_equal(rd.to_numpy(np.float32, (M,)).view(np.uint32), y2.view(np.uint32))
    w.free()


@pytest.mark.parametrize("M,K", SHAPES)
def test_mul_mat_swiglu(b200, port, ref, M, K):
    W1 = make_weights(port, ref, M, K, seed=0x5EED0011)
    W3 = make_weights(port, ref, M, K, seed=0x5EED0033)
    x = make_x(port, 1, K)
    w1, w3 = b200.Weight(W1), b200.Weight(W3)
    g_plain, u_plain = b200.mul_mat_host(w1, x)[0], b200.mul_mat_host(w3, x)[0]
    g, u, s, y2 = b200.mul_mat_swiglu(w1, w3, x[0])
    assert np.array_equal(g.view(np.uint32), g_plain.view(np.uint32))
    assert np.array_equal(u.view(np.uint32), u_plain.view(np.uint32))
    assert np.array_equal(s.view(np.uint32), b200.eltwise("silu", g_plain).view(np.uint32))
    assert np.array_equal(y2.view(np.uint32), b200.eltwise("silu_mul", g_plain, u_plain).view(np.uint32))
    # the oracle's restatement of the table look-up: at most one fp16 ulp apart (expf), mostly equal
    s_ref = nb.silu(g_plain)
    ulp = np.maximum(np.abs(s_ref), 2.0 ** -14) * 2.0 ** -10
    assert np.all(np.abs(s - s_ref) <= ulp)
    assert np.mean(s == s_ref) > 0.99
    assert_close(g[None], oracle_Y(port, ref, W1, x))
    assert_close(u[None], oracle_Y(port, ref, W3, x))
    _, _, s_none, y2b = b200.mul_mat_swiglu(w1, w3, x[0], want_silu=False)
    assert s_none is None and np.array_equal(y2b.view(np.uint32), y2.view(np.uint32))
    w1.free(); w3.free()


def test_contract(b200, port, ref):
    W = make_weights(port, ref, 128, 64)
    W2 = make_weights(port, ref, 256, 64)
    w, w2 = b200.Weight(W), b200.Weight(W2)
    assert b200.lib().ggml_b200_neigh_ok(w.h) == 1
    d = b200.DevBuf(4096)
    with pytest.raises(b200.B200Error):
        b200._check(b200.lib().ggml_b200_mul_mat_q4_0_add(w.h, d.ptr, d.ptr, None, d.ptr))
    with pytest.raises(b200.B200Error):
        b200._check(b200.lib().ggml_b200_mul_mat_q4_0_swiglu(w.h, w2.h, d.ptr, d.ptr, d.ptr, None, d.ptr))
    w.free(); w2.free()


@pytest.mark.parametrize("npos,rows,ne0", [(1, 32, 128), (7, 32, 128), (1, 71, 64), (3, 5, 2)])
def test_rope_mode0(b200, npos, rows, ne0):
    """ggml_b200_rope_f32 against the oracle's restatement on the same angle table: bit-exact (the same four products and
    two sums per pair), in place and out of place."""
    rng = np.random.default_rng(npos * 1000 + ne0)
    x = rng.standard_normal((npos, rows, ne0), dtype=np.float32)
    tab = nb.rope_table(13, ne0, ne0, npos)
    want = nb.rope(x, tab)
    xd, td = b200.DevBuf.from_numpy(x), b200.DevBuf.from_numpy(tab)
    yd = b200.DevBuf(x.nbytes)
    b200._check(b200.lib().ggml_b200_rope_f32(xd.ptr, yd.ptr, td.ptr, ne0, rows, npos))
    b200.synchronize()
    assert np.array_equal(yd.to_numpy(np.float32, x.shape).view(np.uint32), want.view(np.uint32))
    b200._check(b200.lib().ggml_b200_rope_f32(xd.ptr, xd.ptr, td.ptr, ne0, rows, npos))
    b200.synchronize()
    assert np.array_equal(xd.to_numpy(np.float32, x.shape).view(np.uint32), want.view(np.uint32))
    assert np.allclose(np.linalg.norm(want, axis=-1), np.linalg.norm(x, axis=-1), rtol=1e-5)      # a rotation


@pytest.mark.parametrize("nrows,ncols", [(1, 4096), (1, 4544), (3, 8192), (2, 11008), (1, 64), (5, 100), (1, 20000)])
def test_rms_norm_kernels(b200, nrows, ncols):
    """Both rms_norm kernels (rows held in registers up to 16384 aligned floats, the strided one otherwise) against the oracle's
    restatement of ggml_compute_forward_rms_norm_f32: the double-precision sum is a tree here and a serial loop there, which
    moves the float mean by at most its last bit -- one ulp on the result; with the norm weights the product is exact on top."""
    rng = np.random.default_rng(nrows * 7 + ncols)
    x = rng.standard_normal((nrows, ncols), dtype=np.float32) * np.float32(2.5)
    w = (1.0 + 0.1 * rng.standard_normal(ncols)).astype(np.float32)
    want = nb.rms_norm(x)
    y = b200.rms_norm(x)
    assert np.all(np.abs(y - want) <= np.spacing(np.abs(want)))
    assert np.mean(y == want) > 0.5
    y1, y2 = b200.rms_norm(x, w)
    assert np.array_equal(y1.view(np.uint32), y.view(np.uint32))
    assert np.array_equal(y2.view(np.uint32), (y1 * w[None, :]).astype(np.float32).view(np.uint32))
