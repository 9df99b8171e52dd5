"""Sibling formats Q8_0, Q4_1, Q5_0 and Q5_1 on the GPU against the oracle: the activation blocks the kernel's
fused quantizer forms and the int32 block dots bit-exact, the fp32 result within 1e-4 (north_star's tolerance,
relative to the largest |y| of the column)."""
import numpy as np
import pytest

from oracle import siblings as sib

pytestmark = pytest.mark.gpu
TYPES = [sib.TYPE_Q8_0, sib.TYPE_Q4_1, sib.TYPE_Q5_0, sib.TYPE_Q5_1]
TOL = 1e-4


def _weights(qtype, seed, M, K):
    """Random reference-format blocks: sane fp16 scales (and mins), every quant byte value."""
    rng = np.random.default_rng(seed)
    nbk = K // 32
    W = np.zeros((M, nbk, sib.BLOCK_BYTES[qtype]), dtype=np.uint8)
    d = (rng.uniform(0.002, 0.02, size=(M, nbk)) * rng.choice([-1.0, 1.0], size=(M, nbk))).astype(np.float16)
    W[..., 0:2] = d.view(np.uint8).reshape(M, nbk, 2)
    off = 2
    if qtype in (sib.TYPE_Q4_1, sib.TYPE_Q5_1):
        m = rng.uniform(-0.2, 0.2, size=(M, nbk)).astype(np.float16)
        W[..., 2:4] = m.view(np.uint8).reshape(M, nbk, 2)
        off = 4
    W[..., off:] = rng.integers(0, 256, size=W[..., off:].shape, dtype=np.uint8)
    return W


def _x(seed, n, K, scale=1.0):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, K), dtype=np.float32) * np.float32(scale)
    if n:
        x[0, :32] = 0.0
        if K > 72:
            x[n - 1, 64:72] = 0.5
            x[n - 1, 72] = 63.5
    return x


@pytest.mark.parametrize("qtype", TYPES)
def test_activation_blocks_bit_exact(b200, qtype):
    for seed, scale, K in ((1, 1.0, 4096), (2, 1e-3, 4544), (3, 250.0, 11008)):
        x = _x(seed, 3, K, scale)
        x[1, 7] = 1e30
        x[2, 33] = np.inf
        assert np.array_equal(b200.quantize_dot_blocks(qtype, x), sib.quantize_dot(qtype, x))


@pytest.mark.parametrize("qtype", TYPES)
@pytest.mark.parametrize("M,K,n", [(64, 4096, 1), (130, 4544, 1), (257, 4096, 3), (96, 11008, 4), (33, 4544, 9), (40, 256, 5), (8, 32, 2)])
def test_block_dots_bit_exact_and_result_within_tolerance(b200, qtype, M, K, n):
    W = _weights(qtype, 10 + M, M, K)
    X = _x(20 + n, n, K)
    w = b200.QuantWeight(qtype, W)
    Y, S = w.mul_mat(X, checkpoint=True)
    Q = sib.quantize_dot(qtype, X)
    assert np.array_equal(S, sib.block_dots(qtype, W, Q))
    Yo = sib.mul_mat(qtype, W, X)
    err = np.max(np.abs(Y - Yo), axis=1) / np.maximum(np.max(np.abs(Yo), axis=1), 1e-30)
    assert np.all(err <= TOL), err
    # the product path (no checkpoint store; n = 1 takes the one-column instantiation) gives the same floats
    Y2 = w.mul_mat(X)
    assert np.array_equal(Y2, Y)
    w.free()


@pytest.mark.parametrize("M,K,n", [(128, 128, 32), (200, 4096, 33), (257, 4544, 130), (130, 11008, 64), (512, 4096, 160)])
def test_q8_0_prefill_on_the_tensor_cores(b200, M, K, n):
    """n >= gemm_min_n: Q8_0's int8 plane is kernel 3's A operand as it lies (tcgen05 kind::i8, WFMT = 8).  Block dots
    bit-exact, result within 1e-4, the column-pass kernel (forced with set_path) agrees; ragged rows / columns / K tail."""
    qtype = sib.TYPE_Q8_0
    W = _weights(qtype, 40 + M, M, K)
    X = _x(50 + n, n, K)
    w = b200.QuantWeight(qtype, W)
    c0 = b200.launch_count()
    Y, S = w.mul_mat(X, checkpoint=True)
    assert b200.launch_count() - c0 == 2                     # kernel 1 + kernel 3, not n / 4 column passes
    Q = sib.quantize_dot(qtype, X)
    assert np.array_equal(S, sib.block_dots(qtype, W, Q))
    Yo = sib.mul_mat(qtype, W, X)
    err = np.max(np.abs(Y - Yo), axis=1) / np.maximum(np.max(np.abs(Yo), axis=1), 1e-30)
    assert np.all(err <= TOL), err
    assert np.array_equal(w.mul_mat(X), Y)
    b200.set_path(b200.PATH_GEMV)
    try:
        Yc = w.mul_mat(X)
    finally:
        b200.set_path(b200.PATH_AUTO)
    errc = np.max(np.abs(Y - Yc), axis=1) / np.maximum(np.max(np.abs(Yo), axis=1), 1e-30)
    assert np.all(errc <= 1e-5), errc
    w.free()


@pytest.mark.parametrize("qtype", TYPES)
def test_against_the_reference_itself(b200, ref, qtype):
    """Where oracle/_ref is present: weights quantized by the reference's own quantize_row_q, the op through its graph."""
    M, K, n = 192, 4544, 2
    rng = np.random.default_rng(77)
    W = ref.type_quantize(qtype, rng.standard_normal((M, K), dtype=np.float32) * np.float32(0.05))
    X = _x(78, n, K)
    w = b200.QuantWeight(qtype, W)
    Y = w.mul_mat(X)
    Yr = ref.type_mul_mat(qtype, W, X, n_threads=2)
    err = np.max(np.abs(Y - Yr), axis=1) / np.max(np.abs(Yr), axis=1)
    assert np.all(err <= TOL), err
    w.free()


@pytest.mark.parametrize("qtype", TYPES)
def test_edge_cases(b200, qtype):
    W = _weights(qtype, 5, 16, 4096)
    w = b200.QuantWeight(qtype, W)
    assert w.mul_mat(np.zeros((0, 4096), dtype=np.float32)).shape == (0, 16)          # n = 0
    Y = w.mul_mat(np.zeros((2, 4096), dtype=np.float32))                              # all-zero activations: d = 0 everywhere
    assert np.array_equal(Y, np.zeros((2, 16), dtype=np.float32))
    # a row stride smaller than a row is an argument error, not a crash
    import ctypes as C
    xd = b200.DevBuf(4096 * 4)
    yd = b200.DevBuf(16 * 4)
    rc = b200.lib().ggml_b200_mul_mat_quant(w.h, xd.ptr, 1, 4096, yd.ptr, 16 * 4)
    assert rc == b200.ERR_ARG
    w.free()


def test_upload_rejects_other_types(b200):
    import ctypes as C
    buf = np.zeros(18 * 4, dtype=np.uint8)
    h = b200.lib().ggml_b200_upload_quant(2, buf.ctypes.data_as(C.c_void_p), 1, 128, 0)       # Q4_0 has its own entry point
    assert not h
    assert b"type 2" in b200.lib().ggml_b200_last_error()
