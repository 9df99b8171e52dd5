"""The numpy restatement of the sibling formats (oracle/siblings.py) pinned against the reference itself (oracle/_ref):
bit-exact for blocks and integer block dots, and for the fp32 results wherever the emulated fma does not double-round."""
import numpy as np
import pytest

from oracle import siblings as sib

TYPES = [sib.TYPE_Q8_0, sib.TYPE_Q4_1, sib.TYPE_Q5_0, sib.TYPE_Q5_1]


def _rows(seed, n, K, scale=1.0, huge=1e30):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, K), dtype=np.float32) * np.float32(scale)
    x[0, :32] = 0.0                      # an all-zero block: d = 0, id = 0
    if n > 1:
        x[1, 5] = huge                   # an outlier (1e30 overflows the fp16 scale of q8_0: blocks only)
        x[1, 40:48] = 0.5                # exact ties after scaling are possible
    return x


@pytest.mark.parametrize("qtype", TYPES)
def test_activation_blocks_bit_exact(ref, qtype):
    for seed, scale in ((1, 1.0), (2, 1e-3), (3, 300.0)):
        x = _rows(seed, 5, 4544, scale)
        assert np.array_equal(sib.quantize_dot(qtype, x), ref.type_quantize_dot(qtype, x))


def test_weight_codecs_bit_exact(ref):
    x = _rows(7, 6, 4096)
    assert np.array_equal(sib.quantize_q4_1_reference(x), ref.type_quantize(sib.TYPE_Q4_1, x))
    assert np.array_equal(sib.quantize_q5_0_reference(x), ref.type_quantize(sib.TYPE_Q5_0, x))
    assert np.array_equal(sib.quantize_q5_1_reference(x), ref.type_quantize(sib.TYPE_Q5_1, x))
    # quantize_row_q of Q8_0 is the SIMD quantize_row_q8_0 (ggml.c:1660-1666): the activation contract
    assert np.array_equal(sib.quantize_dot(sib.TYPE_Q8_0, x), ref.type_quantize(sib.TYPE_Q8_0, x))


def _weights(ref, qtype, seed, M, K):
    rng = np.random.default_rng(seed)
    return ref.type_quantize(qtype, rng.standard_normal((M, K), dtype=np.float32) * np.float32(0.05))


@pytest.mark.parametrize("qtype", TYPES)
def test_vec_dot_and_op_against_the_reference(ref, qtype):
    M, K, n = 48, 4544, 3
    W = _weights(ref, qtype, 11, M, K)
    X = _rows(12, n, K, huge=3e3)
    Q = ref.type_quantize_dot(qtype, X)
    Y = sib.mul_mat(qtype, W, X)
    Yref = ref.type_mul_mat(qtype, W, X, n_threads=2)
    # the graph op == its own vec_dot per (row, column)
    for c in range(n):
        for m in (0, 7, M - 1):
            assert np.float32(ref.type_vec_dot(qtype, W[m], Q[c])) == Yref[c, m]
    scale = np.abs(Yref).max()
    assert np.max(np.abs(Y - Yref)) <= 2e-6 * scale
    assert np.mean(Y == Yref) > 0.9          # the emulated fma double-rounds only in rare ties


@pytest.mark.parametrize("qtype", TYPES)
def test_block_dots_are_the_integer_part_of_vec_dot(ref, qtype):
    # one block, scales 1: vec_dot == S_b (+ m * s for Q4_1 with m = 0)
    rng = np.random.default_rng(5)
    W = np.zeros((4, 2, sib.BLOCK_BYTES[qtype]), dtype=np.uint8)
    one = np.array([1.0], dtype=np.float16).view(np.uint8)
    W[..., 0:2] = one
    off = {sib.TYPE_Q8_0: 2, sib.TYPE_Q4_1: 4, sib.TYPE_Q5_0: 2, sib.TYPE_Q5_1: 4}[qtype]     # after d (and m = 0): quants, or qh + quants
    W[..., off:] = rng.integers(0, 256, size=W[..., off:].shape, dtype=np.uint8)
    X = np.rint(rng.standard_normal((2, 64)) * 40).astype(np.float32)
    X[:, 0] = 127.0                                            # amax = 127 -> d = 1, the quants are the values
    X[:, 32] = -127.0
    Q = ref.type_quantize_dot(qtype, X)
    S = sib.block_dots(qtype, W, Q)
    for c in range(2):
        for m in range(4):
            assert float(S[c, m].sum()) == ref.type_vec_dot(qtype, W[m], Q[c])
