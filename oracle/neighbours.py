"""TEST INFRASTRUCTURE (oracle): numpy restatement of the element-wise nodes that sit right behind a Q4_0 mul_mat in a
LLaMA decode graph -- the checker for the neighbour-fused kernels.  Never imported by the product.

  add   ggml_compute_forward_add_f32 (ggml.c:8480-8560): one fp32 addition per element
  silu  ggml_vec_silu_f32 with GGML_SILU_FP16 (ggml.c:122, :3519-3538, table built at :4245-4252):
        y = fp16_to_fp32(table[fp32_to_fp16(x)]),  table[h] = fp32_to_fp16(silu_f32(fp16_to_fp32(h))),
        silu_f32(v) = v / (1.0f + expf(-v))
  mul   ggml_compute_forward_mul_f32 (ggml.c:9160-9240): one fp32 multiplication per element

Parity: PINNED against the reference itself (tests/test_oracle_neighbours.py, CPU only): the unmodified ggml.c runs add,
mul, silu, rms_norm and rope through oracle/ref_shim.c (ref_neighbour_op) -- add / mul / rms_norm / rope agree bit for bit
(rope's angles come from the host's libm here as in the reference), silu to the fp16 table entry except where numpy's
float32 exp and the host's expf straddle a rounding boundary (callers allow one fp16 ulp).
"""
import numpy as np


def add(a, b):
    return (np.asarray(a, np.float32) + np.asarray(b, np.float32)).astype(np.float32)


def silu(x):
    h = np.asarray(x, np.float32).astype(np.float16).astype(np.float32)
    with np.errstate(over="ignore"):
        s = (h / (np.float32(1.0) + np.exp(-h, dtype=np.float32))).astype(np.float32)
    return s.astype(np.float16).astype(np.float32)


def silu_mul(a, b):
    return (silu(a) * np.asarray(b, np.float32)).astype(np.float32)


def _libm():
    """the host's libm (powf / cosf / sinf as the reference calls them); None where it cannot be loaded"""
    import ctypes
    import ctypes.util
    try:
        m = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
        for f in (m.cosf, m.sinf):
            f.restype = ctypes.c_float
            f.argtypes = [ctypes.c_float]
        m.powf.restype = ctypes.c_float
        m.powf.argtypes = [ctypes.c_float, ctypes.c_float]
        return m
    except OSError:
        return None


def rope_table(n_past, n_dims, ne0, npos):
    """cos / sin of the angles ggml_compute_forward_rope_f32 walks through (ggml.c:12771-12813, mode 0):
    theta_scale = powf(10000.0, -2.0f / n_dims); per row theta = p, then theta *= theta_scale per pair, all in float32.
    Through the host's libm like the reference (an ulp of difference in theta_scale would grow with every pair);
    numpy's float32 functions stand in where libm cannot be loaded.  -> [npos][ne0 / 2][2]"""
    m = _libm()
    if m is not None:
        scale = np.float32(m.powf(10000.0, float(np.float32(-2.0) / np.float32(n_dims))))
        cos, sin = (lambda t: np.float32(m.cosf(float(t)))), (lambda t: np.float32(m.sinf(float(t))))
    else:
        scale = np.float32(np.power(np.float32(10000.0), np.float32(-2.0) / np.float32(n_dims), dtype=np.float32))
        cos, sin = (lambda t: np.cos(t, dtype=np.float32)), (lambda t: np.sin(t, dtype=np.float32))
    tab = np.zeros((npos, ne0 // 2, 2), np.float32)
    for i2 in range(npos):
        theta = np.float32(n_past + i2)
        for k in range(ne0 // 2):
            tab[i2, k, 0] = cos(theta)
            tab[i2, k, 1] = sin(theta)
            theta = np.float32(theta * scale)
    return tab


def rope(x, tab):
    """x [npos][rows][ne0]: adjacent pairs rotated, four float32 products and two sums per pair, no fused multiply-add."""
    x = np.asarray(x, np.float32)
    x0, x1 = x[..., 0::2], x[..., 1::2]
    c, s = tab[:, None, :, 0], tab[:, None, :, 1]
    y = np.empty_like(x)
    y[..., 0::2] = (x0 * c).astype(np.float32) - (x1 * s).astype(np.float32)
    y[..., 1::2] = (x0 * s).astype(np.float32) + (x1 * c).astype(np.float32)
    return y


def rms_norm(x, eps=1e-6):
    """ggml_compute_forward_rms_norm_f32 (ggml.c:10400-10450): per row, sum of squares accumulated in double in index order,
    mean = float(sum / ne0), scale = 1 / sqrtf(mean + eps), y = x * scale."""
    x = np.asarray(x, np.float32)
    sq = (x * x).astype(np.float32).astype(np.float64)            # the product is formed in float, then widened
    s = np.zeros(x.shape[:-1], np.float64)
    for i in range(x.shape[-1]):
        s = s + sq[..., i]
    mean = (s / x.shape[-1]).astype(np.float32)
    scale = (np.float32(1.0) / np.sqrt(mean + np.float32(eps), dtype=np.float32)).astype(np.float32)
    return (x * scale[..., None]).astype(np.float32)
