"""GPU parity tests (run with -m gpu on a B200).  Everything goes through the C ABI
(lib/libggml_b200.so via ctypes); the oracle is only the checker.

Bars (BASELINE.json north_star):
  * q8_0 activation blocks: bit-exact (34-byte compare)
  * int32 per-block dots S_b: bit-exact
  * fp32 outputs: max|dY| / max|Y| per column <= 1e-4, and max|dY| / sum_b|S_b*dw*dx| <= 1e-4
    (per-element relative error is meaningless on near-zero outputs; only the
    accumulation order differs from the CPU)
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-4


def make_weights(port, ref, M, K, seed=0x5EED0001):
    wf = port.synthetic(M * K, seed + M * 31 + K, 0.02).reshape(M, K)
    q = ref.quantize_q4_0_bulk(wf) if ref is not None else port.quantize_q4_0(wf)
    return q


def make_x(port, n, K, seed=0x5EED0002):
    return port.synthetic(n * K, seed + n, 1.0, 100).reshape(n, K)


def oracle_Y(port, ref, W, X, threads=8):
    """CPU truth: the compiled reference when present (fast, multi-threaded), else the C port."""
    if ref is not None:
        return ref.mul_mat(W, X, n_threads=threads)
    return port.mul_mat(W, X, avx2_order=True)


def scale_sum(port, W, X):
    """sum_b |S_b * dw * dx| per output, the magnitude actually summed (metric ii)."""
    q8 = port.quantize_q8_0(X)
    S = port.block_dots_matrix(W, q8).astype(np.float64)                     # [n][M][nbk]
    dw = W[:, :, :2].copy().view(np.float16).astype(np.float64)[..., 0]      # [M][nbk]
    dx = q8[:, :, :2].copy().view(np.float16).astype(np.float64)[..., 0]     # [n][nbk]
    return np.sum(np.abs(S * dw[None] * dx[:, None, :]), axis=-1)            # [n][M]


def assert_close(Y, Yref, denom2=None):
    assert Y.shape == Yref.shape
    assert np.isfinite(Y).all()
    d = np.abs(Y.astype(np.float64) - Yref.astype(np.float64))
    col_max = np.max(np.abs(Yref), axis=1, keepdims=True)
    col_max[col_max == 0] = 1.0
    m1 = float(np.max(d / col_max))
    assert m1 <= TOL, f"max|dY|/max|Y| = {m1:.3e}"
    if denom2 is not None:
        dd = np.where(denom2 > 0, denom2, 1.0)
        m2 = float(np.max(d / dd))
        assert m2 <= TOL, f"max|dY|/sum|terms| = {m2:.3e}"


# ---------------------------------------------------------------------------- kernel 1
def test_quantize_q8_0_golden_bit_exact(b200, golden):
    assert np.array_equal(b200.quantize_q8_0(golden["cos_b"][None, :]), golden["cos_b_q8_0"][None])
    assert np.array_equal(b200.quantize_q8_0(golden["op_X"]), golden["op_q8"])


def test_quantize_q8_0_edge_cases_bit_exact(b200, golden):
    """zero blocks, single non-zero, exact .5 ties (even rounding), amax in last lane, fp16
    denormal / zero / inf scale, inf input."""
    got = b200.quantize_q8_0(golden["edge_X"])
    want = golden["edge_q8"]
    for i in range(want.shape[0]):
        assert np.array_equal(got[i], want[i]), f"edge row {i}"


@pytest.mark.parametrize("n,K", [(1, 64), (1, 4096), (3, 4544), (7, 11008), (33, 4096), (512, 4096), (2, 18176)])
def test_quantize_q8_0_random_bit_exact(b200, port, n, K):
    rng = np.random.default_rng(n * 1000 + K)
    x = (rng.standard_normal((n, K)) * rng.choice([1e-4, 1.0, 30.0], size=(n, 1))).astype(np.float32)
    assert np.array_equal(b200.quantize_q8_0(x), port.quantize_q8_0(x))


def test_quantize_q8_0_strided_and_unaligned_rows(b200, port):
    """src1 rows with a byte stride nb11 > K*4 (ggml views) and a base that is only 4-byte aligned."""
    import ctypes as C
    n, K, ld = 5, 256, 300
    big = port.synthetic(n * ld + 1, 99, 1.0)
    for off in (0, 1):
        x = np.lib.stride_tricks.as_strided(big[off:], shape=(n, K), strides=(ld * 4, 4))
        xd = b200.DevBuf.from_numpy(big)
        out = b200.DevBuf(n * (K // 32) * 34)
        rc = b200.lib().ggml_b200_quantize_q8_0(C.c_void_p(xd.ptr + 4 * off), n, K, ld * 4, out.ptr)
        assert rc == 0, b200.lib().ggml_b200_last_error()
        b200.synchronize()
        got = out.to_numpy(np.uint8, (n, K // 32, 34))
        assert np.array_equal(got, port.quantize_q8_0(np.ascontiguousarray(x)))


# ---------------------------------------------------------------------------- upload / repack
@pytest.mark.parametrize("repack", [True, False])
def test_upload_download_round_trip(b200, port, ref, repack):
    M, K = 200, 4544
    W = make_weights(port, ref, M, K)
    b200.set_repack(repack)
    try:
        w = b200.Weight(W)
        assert w.layout == (b200.LAYOUT_PLANES if repack else b200.LAYOUT_RAW)
        assert w.device_bytes == M * (K // 32) * 18          # algorithmic bytes unchanged by the repack
        assert np.array_equal(w.download(), W)
        ws = b200.Weight(W, 64, 130)                          # a row shard
        assert ws.rows == 66 and np.array_equal(ws.download(), W[64:130])
        w.free(); ws.free()
    finally:
        b200.set_repack(True)


# ---------------------------------------------------------------------------- kernel 2: S_b and Y
@pytest.mark.parametrize("repack", [True, False])
@pytest.mark.parametrize("n", [1, 2, 3, 5, 8, 11])
def test_gemv_block_dots_bit_exact(b200, port, ref, repack, n):
    M, K = 70, 1088            # ragged: M not a multiple of anything, nbk = 34 (partial lane sweep)
    W = make_weights(port, ref, M, K)
    X = make_x(port, n, K)
    b200.set_repack(repack)
    try:
        w = b200.Weight(W)
        S = b200.block_dots(w, X, b200.PATH_GEMV)
        w.free()
    finally:
        b200.set_repack(True)
    want = port.block_dots_matrix(W, port.quantize_q8_0(X))
    assert np.array_equal(S, want)


def test_gemv_golden_op(b200, golden, port):
    b200.set_path(b200.PATH_GEMV)
    try:
        w = b200.Weight(golden["op_W"])
        Y = b200.mul_mat_host(w, golden["op_X"])
        w.free()
    finally:
        b200.set_path(b200.PATH_AUTO)
    assert_close(Y, golden["op_Y"], scale_sum(port, golden["op_W"], golden["op_X"]))


def test_gemv_cos_vectors_of_reference_test(b200, golden):
    """tests/test-quantize-fns.cpp's own vectors: one row, K = 4096 -> the 4470.44141 dot."""
    w = b200.Weight(golden["cos_a_q4_0"][None])
    S = b200.block_dots(w, golden["cos_b"][None], b200.PATH_GEMV)
    Y = b200.mul_mat_host(w, golden["cos_b"][None])
    w.free()
    assert np.array_equal(S[0, 0], golden["cos_S"]) and int(S.sum()) == -1034162
    assert abs(float(Y[0, 0]) - float(golden["cos_dot"][0])) <= 1e-6 * 4470.44141 * 4


@pytest.mark.parametrize("repack", [True, False])
@pytest.mark.parametrize("M,K,n", [(1, 64, 1), (33, 64, 1), (4096, 4096, 1), (4672, 4544, 1), (1000, 11008, 1),
                                   (777, 4096, 2), (512, 4544, 4), (300, 18176, 3), (256, 4096, 9)])
def test_gemv_outputs_vs_oracle(b200, port, ref, repack, M, K, n):
    W = make_weights(port, ref, M, K)
    X = make_x(port, n, K)
    b200.set_repack(repack)
    b200.set_path(b200.PATH_GEMV)
    try:
        w = b200.Weight(W)
        Y = b200.mul_mat_host(w, X)
        w.free()
    finally:
        b200.set_repack(True)
        b200.set_path(b200.PATH_AUTO)
    assert_close(Y, oracle_Y(port, ref, W, X), scale_sum(port, W, X) if M * K * n <= 1 << 24 else None)


def test_extreme_scales_and_zero_inputs(b200, port, ref):
    """all-zero activations -> exact zeros; zero weight blocks; tiny / huge activation scales."""
    M, K = 64, 512
    W = make_weights(port, ref, M, K)
    W[5, 3, :] = 0                                # an all-zero weight block (d = 0, q = 0)
    w = b200.Weight(W)
    Y0 = b200.mul_mat_host(w, np.zeros((2, K), np.float32))
    assert np.array_equal(Y0, np.zeros((2, M), np.float32))
    for scale in (1e-6, 1e3):
        X = make_x(port, 2, K) * np.float32(scale)
        assert_close(b200.mul_mat_host(w, X), oracle_Y(port, ref, W, X), scale_sum(port, W, X))
    w.free()


def test_empty_and_sharded_rows(b200, port, ref):
    M, K, n = 300, 1024, 2
    W = make_weights(port, ref, M, K)
    X = make_x(port, n, K)
    Yref = oracle_Y(port, ref, W, X)
    # n = 0 is a no-op
    w = b200.Weight(W)
    assert b200.mul_mat_host(w, np.zeros((0, K), np.float32)).shape == (0, M)
    w.free()
    # two row shards write disjoint column ranges of the same dst (ggml-cuda.cu:2780-2797 semantics)
    Y = np.full((n, M), np.nan, np.float32)
    for lo, hi in ((0, 128), (128, 300)):
        ws = b200.Weight(W, lo, hi)
        b200.mul_mat_host(ws, X, Y)
        ws.free()
    assert_close(Y, Yref, scale_sum(port, W, X))
    # an empty shard leaves dst untouched
    Y2 = np.full((n, M), 7.0, np.float32)
    we = b200.Weight(W, 300, 300)
    b200.mul_mat_host(we, X, Y2)
    we.free()
    assert np.all(Y2 == 7.0)


def test_device_pointer_entry_with_strides(b200, port, ref):
    """ggml_b200_mul_mat_q4_0 with device pointers and padded src1 / dst rows."""
    M, K, n = 130, 256, 3
    ldx, ldy = K + 8, M + 4
    W = make_weights(port, ref, M, K)
    X = make_x(port, n, K)
    Xp = np.zeros((n, ldx), np.float32); Xp[:, :K] = X
    w = b200.Weight(W)
    xd = b200.DevBuf.from_numpy(Xp)
    yd = b200.DevBuf.from_numpy(np.full((n, ldy), -5.0, np.float32))
    b200.mul_mat_dev(w, xd.ptr, n, yd.ptr, nb11=ldx * 4, nb1=ldy * 4)
    b200.synchronize()
    Yp = yd.to_numpy(np.float32, (n, ldy))
    w.free()
    assert np.all(Yp[:, M:] == -5.0)
    assert_close(Yp[:, :M], oracle_Y(port, ref, W, X), scale_sum(port, W, X))


def test_linearity_property_full_size(b200, port, ref):
    """Size-independent property at a BASELINE shape (11008 x 4096): with activations that are
    exactly representable after quantization (integers * power of two, amax = 127 per block), the
    op is exact in the q8 domain, so Y(2x) == 2 * Y(x) bit-for-bit."""
    M, K = 11008, 4096
    W = make_weights(port, ref, M, K)
    rng = np.random.default_rng(5)
    xi = rng.integers(-127, 128, size=(1, K)).astype(np.float32)
    xi[:, ::32] = 127.0
    w = b200.Weight(W)
    Y1 = b200.mul_mat_host(w, xi * np.float32(0.25))
    Y2 = b200.mul_mat_host(w, xi * np.float32(0.5))
    w.free()
    assert np.array_equal((Y1 * np.float32(2)).view(np.uint32), Y2.view(np.uint32))
    assert_close(Y1, oracle_Y(port, ref, W, xi * np.float32(0.25)))


def test_launch_counter_counts_kernels(b200, port, ref):
    W = make_weights(port, ref, 64, 256)
    w = b200.Weight(W)
    c0 = b200.launch_count()
    b200.mul_mat_host(w, make_x(port, 1, 256))
    assert b200.launch_count() - c0 == 1          # n = 1: kernel 1 is fused into kernel 2
    c0 = b200.launch_count()
    b200.mul_mat_host(w, make_x(port, 2, 256))
    assert b200.launch_count() - c0 == 2          # kernel 1 + kernel 2
    w.free()


@pytest.mark.parametrize("M,K", [(70, 1088), (4096, 4096), (300, 18176), (1, 64), (149, 4544), (1000, 11008),
                                 (500, 22016), (297, 8192), (2000, 256), (40, 24576 + 64)])
def test_fused_kernel_checkpoints_bit_exact(b200, port, ref, golden, M, K):
    """The n = 1 kernel quantizes inside the CTA: its q8_0 blocks and S_b must be the reference's,
    and its output must equal the unfused two-kernel path bit for bit (same accumulation order)."""
    W = make_weights(port, ref, M, K)
    x = make_x(port, 1, K)
    w = b200.Weight(W)
    q8, S = b200.fused_checkpoint(w, x)
    want_q8 = port.quantize_q8_0(x)
    assert np.array_equal(q8, want_q8[0])
    assert np.array_equal(S, port.block_dots_matrix(W, want_q8)[0])
    Ys = b200.mul_mat_host(w, x)                      # the one-launch host entry point (gemv_q4_0_hostio.cuh)
    try:
        q8r, Sr = b200.fused_checkpoint(w, x)
        b200.set_tuning("host_fast", 0)
        Yf = b200.mul_mat_host(w, x)                  # fused kernel behind staged copies
        b200.set_tuning("gemv_fused", 0)
        Yu = b200.mul_mat_host(w, x)                  # kernel 1 + kernel 2 as two launches
    finally:
        b200.set_tuning("gemv_fused", 1)
        b200.set_tuning("host_fast", 1)
    w.free()
    assert np.array_equal(q8r, want_q8[0]) and np.array_equal(Sr, S)
    assert np.array_equal(Yf.view(np.uint32), Yu.view(np.uint32))   # same accumulation order
    Yo = oracle_Y(port, ref, W, x)
    assert_close(Ys, Yo, scale_sum(port, W, x) if M * K <= 1 << 24 else None)
    assert_close(Yf, Yo)


def test_fused_kernel_edge_rows(b200, golden, port, ref):
    """edge activations (ties, zeros, denormal / inf scales) through the fused quantizer."""
    W = make_weights(port, ref, 8, 64)
    w = b200.Weight(W)
    for i, row in enumerate(golden["edge_X"]):
        q8, _ = b200.fused_checkpoint(w, row)
        assert np.array_equal(q8, golden["edge_q8"][i]), f"edge row {i}"
    w.free()


@pytest.mark.parametrize("pdl", [0, 1])
def test_back_to_back_ops_with_and_without_pdl(b200, port, ref, pdl):
    """Consecutive ops share the activation workspace and overlap their prologues under
    programmatic dependent launch; results must not depend on it."""
    M, K = 600, 2048
    W1, W2 = make_weights(port, ref, M, K, 1), make_weights(port, ref, M, K, 2)
    w1, w2 = b200.Weight(W1), b200.Weight(W2)
    b200.set_tuning("pdl", pdl)
    try:
        for n in (1, 3):
            xs = [make_x(port, n, K, 10 + i) for i in range(4)]
            xd = [b200.DevBuf.from_numpy(x) for x in xs]
            yd = [b200.DevBuf(n * M * 4) for _ in xs]
            for rep in range(3):
                for i in range(4):
                    b200.mul_mat_dev(w1 if i % 2 == 0 else w2, xd[i].ptr, n, yd[i].ptr)
            b200.synchronize()
            for i in range(4):
                Y = yd[i].to_numpy(np.float32, (n, M))
                assert_close(Y, oracle_Y(port, ref, W1 if i % 2 == 0 else W2, xs[i]))
    finally:
        b200.set_tuning("pdl", 1)
    w1.free(); w2.free()


def test_same_src1_ops_overlap_without_changing_results(b200, port, ref):
    """overlap_same_src: ops that read the same x (wq/wk/wv, w1/w3) start without waiting for each other; an op
    that consumes one of their results, reuses a dst, or writes x must still see stream order.  A decode-like chain
    (3 independent ops -> op on the first result -> 2 independent ops -> op that overwrites x) is run 40 times back
    to back with and without the key; every buffer must be bit-identical, and correct against the oracle."""
    K = M = 1024
    Ws = [make_weights(port, ref, M, K, 30 + i) for i in range(7)]
    ws = [b200.Weight(W) for W in Ws]
    x0 = make_x(port, 1, K, 77)

    def run(overlap):
        b200.set_tuning("overlap_same_src", overlap)
        xd = b200.DevBuf.from_numpy(x0)
        yd = [b200.DevBuf(M * 4) for _ in range(7)]
        for rep in range(40):
            for i in (0, 1, 2):
                b200.mul_mat_dev(ws[i], xd.ptr, 1, yd[i].ptr)         # same src1, distinct dst: may overlap
            b200.mul_mat_dev(ws[3], yd[0].ptr, 1, yd[3].ptr)          # reads a result of the run: must wait
            for i in (4, 5):
                b200.mul_mat_dev(ws[i], yd[3].ptr, 1, yd[i].ptr)      # a new run on another x
            b200.mul_mat_dev(ws[5], yd[3].ptr, 1, yd[5].ptr)          # same dst again: must not overlap its twin
            b200.mul_mat_dev(ws[6], yd[4].ptr, 1, yd[6].ptr)
        b200.synchronize()
        out = [y.to_numpy(np.float32, (1, M)) for y in yd]
        return out

    try:
        serial = run(0)
        overlapped = run(1)
    finally:
        b200.set_tuning("overlap_same_src", 0)
    for i in range(7):
        assert np.array_equal(serial[i], overlapped[i]), f"buffer {i} differs with overlap_same_src"
    for i in (0, 1, 2):
        assert_close(serial[i], oracle_Y(port, ref, Ws[i], x0))
    assert_close(serial[3], oracle_Y(port, ref, Ws[3], serial[0]))
    assert_close(serial[6], oracle_Y(port, ref, Ws[6], serial[4]))
    for w in ws:
        w.free()


@pytest.mark.parametrize("n", [5, 48, 200])
def test_same_src1_batched_ops_reuse_quantized_activations(b200, port, ref, n):
    """n > 1 with overlap_same_src: ops that share src1 skip kernel 1 (the activations of the run's first op are
    reused) and, on the tensor-core path, start without waiting for their predecessor.  Same chain as the n = 1
    test, run back to back with and without the key: bit-identical buffers, correct against the oracle."""
    K = M = 1024
    Ws = [make_weights(port, ref, M, K, 60 + i) for i in range(5)]
    ws = [b200.Weight(W) for W in Ws]
    x0 = make_x(port, n, K, 91)

    def run(overlap):
        b200.set_tuning("overlap_same_src", overlap)
        xd = b200.DevBuf.from_numpy(x0)
        yd = [b200.DevBuf(n * M * 4) for _ in range(5)]
        for rep in range(10):
            for i in (0, 1, 2):
                b200.mul_mat_dev(ws[i], xd.ptr, n, yd[i].ptr)         # same src1, distinct dst
            b200.mul_mat_dev(ws[3], yd[1].ptr, n, yd[3].ptr)          # consumes a member: new x, must re-quantize
            b200.mul_mat_dev(ws[4], yd[1].ptr, n, yd[4].ptr)          # same src1 as the previous op
            b200.mul_mat_dev(ws[4], yd[1].ptr, n, yd[4].ptr)          # same dst again: no overlap with its twin
        b200.synchronize()
        return [y.to_numpy(np.float32, (n, M)) for y in yd]

    try:
        serial = run(0)
        overlapped = run(1)
    finally:
        b200.set_tuning("overlap_same_src", 0)
    for i in range(5):
        assert np.array_equal(serial[i], overlapped[i]), f"buffer {i} differs with overlap_same_src"
    for i in (0, 1, 2):
        assert_close(serial[i], oracle_Y(port, ref, Ws[i], x0))
    assert_close(serial[3], oracle_Y(port, ref, Ws[3], serial[1]))
    assert_close(serial[4], oracle_Y(port, ref, Ws[4], serial[1]))
    for w in ws:
        w.free()


# ---------------------------------------------------------------------------- weight-side codec
def test_quantize_q4_0_golden_and_edges_bit_exact(b200, golden):
    """quantize_row_q4_0 (ggml.c:918-957) on the device: the reference test's cos vector and the edge rows (sign of d
    decided by the FIRST element of largest |x|, -0.0 scale of a zero block, clamp to 15) byte for byte."""
    q = b200.quantize_q4_0(golden["cos_a"].reshape(1, -1))
    assert np.array_equal(q.reshape(-1, 18), golden["cos_a_q4_0"].reshape(-1, 18))
    assert q[0, 0, 0] == 0x33 and q[0, 0, 1] == 0xB4 and list(q[0, 0, 2:6]) == [0xF0, 0xA4, 0x3B, 0x0F]
    e = b200.quantize_q4_0(golden["edge_W"])
    assert np.array_equal(e, golden["edge_W_q4_0"])


@pytest.mark.parametrize("M,K", [(1, 64), (7, 4096), (300, 4544), (33, 11008)])
def test_quantize_q4_0_random_bit_exact(b200, port, ref, M, K):
    wf = port.synthetic(M * K, 0xABC + M + K, 0.02, 50).reshape(M, K)
    want = ref.quantize_q4_0_bulk(wf) if ref is not None else port.quantize_q4_0(wf)
    assert np.array_equal(b200.quantize_q4_0(wf), want)


def test_quantize_q4_0_strided_rows_and_device_upload(b200, port, ref):
    """Row stride > 4*K, and the on-device path f32 -> blocks -> weight handle without a host round trip."""
    M, K = 70, 1088
    wide = port.synthetic(M * (K + 24), 0x77, 0.02).reshape(M, K + 24)
    xd = b200.DevBuf.from_numpy(wide)
    out = b200.DevBuf(M * (K // 32) * 18)
    b200._check(b200.lib().ggml_b200_quantize_q4_0(xd.ptr + 16, M, K, (K + 24) * 4, out.ptr))   # rows start 4 floats in
    b200.synchronize()
    got = out.to_numpy(np.uint8, (M, K // 32, 18))
    want = port.quantize_q4_0(np.ascontiguousarray(wide[:, 4:4 + K]))
    assert np.array_equal(got, want)
    h = b200.lib().ggml_b200_upload_q4_0(out.ptr, M, K, (K // 32) * 18, 0, M)                     # device pointer in
    assert h
    back = np.empty((M, K // 32, 18), np.uint8)
    b200._check(b200.lib().ggml_b200_download_q4_0(h, back.ctypes.data))
    assert np.array_equal(back, want)
    x = make_x(port, 1, K, 5)
    y = np.zeros((1, M), np.float32)
    b200._check(b200.lib().ggml_b200_mul_mat_q4_0_host(h, x.ctypes.data, 1, K * 4, y.ctypes.data, M * 4))
    assert_close(y, oracle_Y(port, ref, want, x))
    b200.lib().ggml_b200_free_weight(h)


def assert_same_floats(a, b):
    """Bit-equal, except that any NaN matches any NaN (0 * inf: x86 yields 0xFFC00000, the GPU 0x7FFFFFFF)."""
    a, b = np.asarray(a, np.float32).reshape(-1), np.asarray(b, np.float32).reshape(-1)
    na, nb = np.isnan(a), np.isnan(b)
    assert np.array_equal(na, nb)
    assert np.array_equal(a.view(np.uint32)[~na], b.view(np.uint32)[~nb])


def test_dequantize_q4_0_bit_exact(b200, golden, port):
    """dequantize_row_q4_0 (ggml.c:1500-1518): exact, including the quantize -> dequantize round trip error bound of
    the reference's own test (tests/test-quantize-fns.cpp:16,129: rmse <= 0.002 on the cos vector)."""
    got = b200.dequantize_q4_0(golden["cos_a_q4_0"])
    assert np.array_equal(got.view(np.uint32), golden["cos_a_dequant"].reshape(-1).view(np.uint32))
    assert_same_floats(b200.dequantize_q4_0(golden["edge_W_q4_0"]), golden["edge_W_dequant"])       # row 3: d = inf, (8-8)*inf = NaN
    n = got.size    # array_rmse of the reference test is sqrt(sum d^2) / n (tests/test-quantize-fns.cpp:33-41)
    err = float(np.sqrt(np.sum((got.astype(np.float64) - golden["cos_a"].astype(np.float64)) ** 2))) / n
    assert err <= 0.002 and abs(err - 0.001121) < 2e-6      # "q4_0 abs err 0.001121"
    W = port.quantize_q4_0(port.synthetic(50 * 2048, 0x99, 0.02).reshape(50, 2048))
    assert np.array_equal(b200.dequantize_q4_0(W).view(np.uint32), port.dequantize_q4_0(W).reshape(-1).view(np.uint32))


@pytest.mark.parametrize("repack", [True, False])
def test_get_rows_q4_0(b200, port, repack):
    """get_rows on a Q4_0 tensor (ggml.c:11868-11890; the token-embedding lookup) from both HBM layouts, full tensor
    and a row shard; an id outside the shard is refused."""
    M, K = 400, 1088
    W = port.quantize_q4_0(port.synthetic(M * K, 0x55, 0.02).reshape(M, K))
    full = port.dequantize_q4_0(W).reshape(M, K)
    b200.set_repack(repack)
    try:
        w = b200.Weight(W)
        ids = np.array([0, 399, 17, 17, 256, 3], np.int32)
        assert np.array_equal(b200.get_rows(w, ids).view(np.uint32), full[ids].view(np.uint32))
        assert b200.get_rows(w, np.zeros(0, np.int32)).shape == (0, K)
        ws = b200.Weight(W, 128, 256)
        ids2 = np.array([128, 255, 200], np.int32)
        assert np.array_equal(b200.get_rows(ws, ids2).view(np.uint32), full[ids2].view(np.uint32))
        with pytest.raises(b200.B200Error):
            b200.get_rows(ws, np.array([100], np.int32))
        w.free(); ws.free()
    finally:
        b200.set_repack(True)


# ---------------------------------------------------------------------------- kernel 3: tcgen05 GEMM
@pytest.mark.parametrize("M,K,n", [(128, 128, 128), (128, 256, 128), (256, 1024, 64), (300, 1088, 200), (70, 64, 5),
                                   (129, 4544, 130), (512, 4096, 512)])
def test_gemm_block_dots_bit_exact(b200, port, ref, M, K, n):
    """int32 S_b out of TMEM (tcgen05.mma kind::i8, one MMA per Q4_0 block) vs the scalar definition
    (ggml.c:2593-2600): ragged M (TMA zero-fill rows), ragged n, K tail (K/32 not a multiple of 4)."""
    W = make_weights(port, ref, M, K)
    X = make_x(port, n, K)
    w = b200.Weight(W)
    S = b200.block_dots(w, X, b200.PATH_GEMM)
    w.free()
    want = port.block_dots_matrix(W, port.quantize_q8_0(X))
    assert np.array_equal(S, want)


@pytest.mark.parametrize("M,K,n", [(128, 128, 128), (300, 1088, 200), (4096, 4096, 512), (4672, 4544, 130),
                                   (1000, 11008, 64), (256, 18176, 33), (2048, 8192, 512)])
def test_gemm_outputs_vs_oracle(b200, port, ref, M, K, n):
    W = make_weights(port, ref, M, K)
    X = make_x(port, n, K)
    b200.set_path(b200.PATH_GEMM)
    try:
        w = b200.Weight(W)
        Y = b200.mul_mat_host(w, X)
        w.free()
    finally:
        b200.set_path(b200.PATH_AUTO)
    assert_close(Y, oracle_Y(port, ref, W, X, threads=16), scale_sum(port, W, X) if M * K * n <= 1 << 26 else None)


def test_gemm_golden_op_and_gemv_agree(b200, golden, port):
    """The golden op (reference ggml_graph_compute output) through the tensor-core path, and the two
    device paths against each other on the same inputs."""
    w = b200.Weight(golden["op_W"])
    b200.set_path(b200.PATH_GEMM)
    try:
        Yg = b200.mul_mat_host(w, golden["op_X"])
        b200.set_path(b200.PATH_GEMV)
        Yv = b200.mul_mat_host(w, golden["op_X"])
    finally:
        b200.set_path(b200.PATH_AUTO)
    w.free()
    den = scale_sum(port, golden["op_W"], golden["op_X"])
    assert_close(Yg, golden["op_Y"], den)
    assert_close(Yg, Yv, den)


def test_auto_dispatch_by_n(b200, port, ref):
    """n >= gemm_min_n takes kernel 3 (1 quantize + 1 GEMM launch), below it the GEMV path."""
    M, K = 256, 512
    W = make_weights(port, ref, M, K)
    w = b200.Weight(W)
    nmin = b200.lib().ggml_b200_get_gemm_min_n()
    for n in (nmin - 1, nmin, 3 * nmin + 1):
        X = make_x(port, n, K)
        Y = b200.mul_mat_host(w, X)
        assert_close(Y, oracle_Y(port, ref, W, X), scale_sum(port, W, X))
    w.free()


def test_gemm_sharded_rows_disjoint(b200, port, ref):
    M, K, n = 512, 1024, 96
    W = make_weights(port, ref, M, K)
    X = make_x(port, n, K)
    Y = np.full((n, M), np.nan, np.float32)
    b200.set_path(b200.PATH_GEMM)
    try:
        for lo, hi in ((0, 128), (128, 384), (384, 512)):
            ws = b200.Weight(W, lo, hi)
            b200.mul_mat_host(ws, X, Y)
            ws.free()
    finally:
        b200.set_path(b200.PATH_AUTO)
    assert_close(Y, oracle_Y(port, ref, W, X), scale_sum(port, W, X))
