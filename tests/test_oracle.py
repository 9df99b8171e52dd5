"""CPU tests (no GPU): the oracle's C restatement against (a) the committed golden vectors that
were produced by running the reference, (b) the known answers SURVEY 8(c) recorded, and (c) the
reference itself where oracle/_ref exists.  Thresholds of the reference's own test
(tests/test-quantize-fns.cpp:16-20) are re-asserted on its own generator."""
import numpy as np
import pytest


def test_fp16_to_fp32_table(golden, port):
    got = np.array([np.float32(port.fp16_to_fp32(h)).view(np.uint32) for h in range(65536)], np.uint32)
    want = golden["fp16_to_fp32_bits"]
    nan = (want & 0x7FFFFFFF) > 0x7F800000
    assert np.array_equal(got[~nan], want[~nan])
    assert np.all((got[nan] & 0x7FFFFFFF) > 0x7F800000)


def test_fp32_to_fp16_samples(golden, port):
    fs, want = golden["fp32_samples"], golden["fp32_to_fp16"]
    got = np.array([port.fp32_to_fp16(float(f)) for f in fs], np.uint16)
    assert np.array_equal(got, want)


def test_known_answers_survey_8c(golden, port):
    """First-block bytes, the 4470.44141 dot and sum S_b recorded from the reference."""
    qa = port.quantize_q4_0(golden["cos_a"])
    qb = port.quantize_q8_0(golden["cos_b"])
    assert (int(qa[0, 0]) | int(qa[0, 1]) << 8) == 0xB433
    assert list(qa[0, 2:6]) == [0xF0, 0xA4, 0x3B, 0x0F]
    assert (int(qb[0, 0]) | int(qb[0, 1]) << 8) == 0x2433
    assert list(qb[0, 2:6].view(np.int8)) == [72, -45, -115, -74]
    d = np.float32(port.vec_dot(qa, qb, avx2_order=True))
    assert d.view(np.uint32) == 0x458BB388          # 4470.44141
    assert int(port.block_dots(qa, qb).sum()) == -1034162


def test_cos_generator_matches_fixture(golden, port):
    # libm cosf could differ on another C library; then the fixtures (not a re-run) are the inputs
    a = port.cos_data(4096, 0.0)
    if not np.array_equal(a, golden["cos_a"]):
        pytest.skip("this libm's cosf differs from the container that made the fixtures")
    assert np.array_equal(port.cos_data(4096, 1.0), golden["cos_b"])


def test_quantizers_bit_exact_vs_golden(golden, port):
    assert np.array_equal(port.quantize_q4_0(golden["cos_a"]), golden["cos_a_q4_0"])
    assert np.array_equal(port.quantize_q8_0(golden["cos_b"]), golden["cos_b_q8_0"])
    assert np.array_equal(port.quantize_q8_0(golden["op_X"]), golden["op_q8"])
    assert np.array_equal(port.dequantize_q4_0(golden["cos_a_q4_0"]).view(np.uint32),
                          golden["cos_a_dequant"].view(np.uint32))


def test_q8_edge_cases_bit_exact(golden, port):
    """zeros, single non-zero, .5 ties, amax in last lane, denormal / zero / inf d, inf input."""
    got = port.quantize_q8_0(golden["edge_X"])
    assert np.array_equal(got, golden["edge_q8"])
    # and the op's quantizer is NOT the scalar _reference (ties-away, id = 1/d): SURVEY 7-2
    sc = port.quantize_q8_0(golden["edge_X"], scalar_reference=True)
    finite = np.isfinite(golden["edge_X"]).all(axis=1)
    assert np.array_equal(sc[finite], golden["edge_q8_scalar_reference"][finite])
    assert not np.array_equal(sc[2], got[2]), "tie row must differ between AVX2 and scalar semantics"


def test_block_dots_and_dot_orders(golden, port):
    qa, qb = golden["cos_a_q4_0"], golden["cos_b_q8_0"]
    S = port.block_dots(qa, qb)
    assert np.array_equal(S, golden["cos_S"])
    assert np.array_equal(port.block_dots_matrix(qa[None], qb[None])[0, 0], S)
    assert np.float32(port.vec_dot(qa, qb, True)).view(np.uint32) == golden["cos_dot"].view(np.uint32)[0]
    # scalar order differs only by accumulation order
    assert abs(port.vec_dot(qa, qb, False) - float(golden["cos_dot"][0])) < 1e-6 * abs(float(golden["cos_dot"][0])) * 8


def test_op_bit_exact_vs_reference_graph(golden, port):
    """The restated operator (AVX2 lane order) reproduces ggml_graph_compute bit for bit."""
    Y = port.mul_mat(golden["op_W"], golden["op_X"], avx2_order=True)
    assert np.array_equal(Y.view(np.uint32), golden["op_Y"].view(np.uint32))
    # thread count does not change values (row partition only, ggml.c:11376-11383)
    assert np.array_equal(golden["op_Y"].view(np.uint32), golden["op_Y_t4"].view(np.uint32))
    Ys = port.mul_mat(golden["op_W"], golden["op_X"], avx2_order=False)
    assert np.max(np.abs(Ys - Y)) / np.max(np.abs(Y)) < 1e-5


def test_reference_test_thresholds(golden, port):
    """tests/test-quantize-fns.cpp:129-151 on its own generator, thresholds :16-20."""
    a, b = golden["cos_a"], golden["cos_b"]
    qa = port.quantize_q4_0(a)
    # array_rmse (test-quantize-fns.cpp:33-40) is sqrtf(sum of squared diffs) / n
    rmse = float(np.sqrt(np.sum((port.dequantize_q4_0(qa).astype(np.float64) - a) ** 2))) / a.size
    assert rmse < 0.002
    assert abs(rmse - 0.001121) < 2e-6      # the value test-quantize-fns -v prints here
    dot_ref = float(np.sum(a.astype(np.float64) * b.astype(np.float64)))
    err = abs(port.vec_dot(qa, port.quantize_q8_0(b), True) - dot_ref) / a.size
    assert err < 0.02


# ---- against the compiled reference itself (present in the build container, prebuilt on the box)
def test_port_vs_reference_random(port, ref):
    if ref is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    rng = np.random.default_rng(7)
    for scale in (1.0, 1e-3, 50.0):
        x = (rng.standard_normal((64, 1024)) * scale).astype(np.float32)
        assert np.array_equal(port.quantize_q8_0(x), ref.quantize_q8_0(x))
        assert np.array_equal(port.quantize_q4_0(x), ref.quantize_q4_0(x))
    # 100k N(0,1) blocks: SIMD semantics reproduced in all of them (SURVEY 7-2 probe)
    x = rng.standard_normal((100000 // 32, 32 * 32)).astype(np.float32)
    assert np.array_equal(port.quantize_q8_0(x), ref.quantize_q8_0(x))
    assert not np.array_equal(port.quantize_q8_0(x, scalar_reference=True), ref.quantize_q8_0(x))


def test_port_op_vs_reference_op(port, ref):
    if ref is None:
        pytest.skip("oracle/_ref not built")
    M, K, n = 256, 4096, 3
    W = ref.quantize_q4_0_bulk(port.synthetic(M * K, 0x5EED0001, 0.02).reshape(M, K))
    X = port.synthetic(n * K, 0x5EED0002, 1.0, 100).reshape(n, K)
    Yr = ref.mul_mat(W, X, n_threads=3)
    Yp = port.mul_mat(W, X, avx2_order=True)
    assert np.array_equal(Yr.view(np.uint32), Yp.view(np.uint32))
    for i in range(0, M, 37):
        assert np.float32(ref.vec_dot(W[i], port.quantize_q8_0(X[1]))).view(np.uint32) == Yp[1, i].view(np.uint32)


def test_fp32_to_fp16_vs_reference_dense(port, ref):
    if ref is None:
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(3)
    bits = rng.integers(0, 2 ** 32, size=20000, dtype=np.uint64).astype(np.uint32)
    fs = bits.view(np.float32)
    for f in fs[np.isfinite(fs)][:20000]:
        assert port.fp32_to_fp16(float(f)) == ref.fp32_to_fp16(float(f)), float(f)
