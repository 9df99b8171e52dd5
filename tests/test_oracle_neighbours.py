"""oracle/neighbours.py (the checker of the neighbour-fused kernels) pinned against the reference itself:
add, mul, silu (fp16 table), rms_norm and rope (mode 0) run by the unmodified ggml.c through oracle/ref_shim.c.  CPU only."""
import numpy as np
import pytest

from oracle import neighbours as nb


@pytest.fixture(scope="module")
def R(ref):
    if ref is None or not hasattr(ref.L, "ref_neighbour_op"):
        pytest.skip("oracle/_ref not built with the neighbour ops (needs /root/reference at build time)")
    return ref


def test_add_and_mul_are_one_fp32_operation(R):
    rng = np.random.default_rng(1)
    a = rng.standard_normal((3, 5, 64), dtype=np.float32) * np.float32(7.0)
    b = rng.standard_normal((3, 5, 64), dtype=np.float32)
    assert np.array_equal(R.neighbour("add", a, b).view(np.uint32), nb.add(a, b).view(np.uint32))
    assert np.array_equal(R.neighbour("mul", a, b).view(np.uint32), (a * b).astype(np.float32).view(np.uint32))


def test_silu_is_the_fp16_table(R):
    rng = np.random.default_rng(2)
    x = np.concatenate([rng.standard_normal(4096, dtype=np.float32) * np.float32(4.0),
                        np.array([0.0, -0.0, 1e-8, -1e-8, 20.0, -20.0, 70000.0, -70000.0, 6e-5, -6e-5], np.float32)])
    got = R.neighbour("silu", x[None, None, :])[0, 0]
    want = nb.silu(x)
    # numpy's float32 exp vs the host libm's expf: equal or one fp16 ulp apart on a rounding boundary
    ok = np.isfinite(want)
    ulp = np.maximum(np.abs(want[ok]), 2.0 ** -14) * 2.0 ** -10
    assert np.all(np.abs(got[ok] - want[ok]) <= ulp)
    assert np.mean(got[ok] == want[ok]) > 0.995
    # the result IS an fp16 value: silu goes through the table
    assert np.array_equal(got[ok], got[ok].astype(np.float16).astype(np.float32))


def test_rms_norm(R):
    rng = np.random.default_rng(3)
    x = rng.standard_normal((2, 3, 4096), dtype=np.float32) * np.float32(3.0)
    got = R.neighbour("rms_norm", x)
    want = nb.rms_norm(x)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("npos,rows,ne0,n_past", [(1, 32, 128, 0), (1, 32, 128, 57), (5, 4, 64, 9), (3, 2, 128, 200)])
def test_rope_mode0(R, npos, rows, ne0, n_past):
    rng = np.random.default_rng(4 + n_past)
    x = rng.standard_normal((npos, rows, ne0), dtype=np.float32)
    got = R.neighbour("rope", x, n_past=n_past)
    want = nb.rope(x, nb.rope_table(n_past, ne0, ne0, npos))
    # the table comes from the host's libm like the reference's angles, the rotation is four products and two sums: the same bits
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.allclose(np.linalg.norm(got, axis=-1), np.linalg.norm(x, axis=-1), rtol=1e-5)
