"""BASELINE.json's shapes at FULL size against the oracle (parity hole 3c): every distinct matmul of LLaMA-7B,
Falcon-7B and LLaMA-65B at n = 1 through the product's decode path (the fused quantize + GEMV launch and the persistent
list kernel), and the prefill shapes at n = 512 through the tcgen05 GEMM.  The oracle is the reference itself
(oracle/_ref, multi-threaded) where it is present, else the C port for the n = 1 cases.  Tolerance: north_star's 1e-4 of
the column's largest |y|."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-4

DECODE_SHAPES = [(32000, 4096), (11008, 4096), (4096, 11008), (4096, 4096),          # LLaMA-7B
                 (4672, 4544), (4544, 4544), (18176, 4544), (4544, 18176),           # Falcon-7B
                 (8192, 8192), (22016, 8192), (8192, 22016)]                         # LLaMA-65B
PREFILL_SHAPES = [(4672, 4544), (11008, 4096), (18176, 4544), (4544, 18176), (32000, 4096)]


def _blocks(seed, M, K):
    """Random Q4_0 blocks written directly (every nibble value, fp16 scales of either sign around 0.01)."""
    rng = np.random.default_rng(seed)
    W = rng.integers(0, 256, size=(M, K // 32, 18), dtype=np.uint8)
    d = (rng.uniform(0.004, 0.02, size=(M, K // 32)) * rng.choice([-1.0, 1.0], size=(M, K // 32))).astype(np.float16)
    W[..., 0:2] = d.view(np.uint8).reshape(M, K // 32, 2)
    return W


def _x(seed, n, K):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, K), dtype=np.float32)
    x[:, ::97] *= 30.0                      # outliers: blocks with very different scales
    x[0, :64] = 0.0
    return x


def _oracle(port, ref, W, X, threads=16):
    if ref is not None:
        return ref.mul_mat(W, X, n_threads=threads)
    return port.mul_mat(W, X, avx2_order=True)


def _rel(Y, Yo):
    cm = np.max(np.abs(Yo), axis=1, keepdims=True)
    cm[cm == 0] = 1.0
    return float(np.max(np.abs(Y.astype(np.float64) - Yo) / cm))


@pytest.mark.parametrize("M,K", DECODE_SHAPES)
def test_full_size_decode_against_the_oracle(b200, port, ref, M, K):
    W = _blocks(1000 + M + K, M, K)
    X = _x(7 + K, 1, K)
    Yo = _oracle(port, ref, W, X)
    w = b200.Weight(W)
    Y = b200.mul_mat_host(w, X)                                  # n = 1: the one-launch fused path
    e1 = _rel(Y, Yo)
    assert e1 <= TOL, f"fused n = 1: {e1:.3e}"
    # the same op as a one-entry persistent list (the kernel the headline number is measured on)
    xd = b200.DevBuf.from_numpy(X)
    yd = b200.DevBuf(M * 4)
    s = b200.OpSet([(w, xd.ptr, yd.ptr)])
    s.run()
    b200.synchronize()
    Ys = yd.to_numpy(np.float32, (1, M))
    s.free()
    e2 = _rel(Ys, Yo)
    assert e2 <= TOL, f"persistent list: {e2:.3e}"
    w.free()


@pytest.mark.parametrize("M,K", PREFILL_SHAPES)
def test_full_size_prefill_n512_against_the_oracle(b200, port, ref, M, K):
    if ref is None:
        pytest.skip("n = 512 at full size needs the compiled reference (oracle/_ref) as the oracle")
    W = _blocks(2000 + M + K, M, K)
    X = _x(9 + K, 512, K)
    Yo = _oracle(port, ref, W, X)
    w = b200.Weight(W)
    b200.set_path(b200.PATH_GEMM)
    try:
        Y = b200.mul_mat_host(w, X)
    finally:
        b200.set_path(b200.PATH_AUTO)
    e = _rel(Y, Yo)
    assert e <= TOL, f"tcgen05 GEMM n = 512: {e:.3e}"
    # size-independent property: the GEMV path on the same inputs gives the same S_b, hence floats within the same bound
    b200.set_path(b200.PATH_GEMV)
    try:
        Yv = b200.mul_mat_host(w, X[:8])
    finally:
        b200.set_path(b200.PATH_AUTO)
    assert _rel(Yv, Yo[:8]) <= TOL
    w.free()
