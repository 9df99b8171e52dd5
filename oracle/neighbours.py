"""TEST INFRASTRUCTURE (oracle): numpy restatement of the element-wise nodes that sit right behind a Q4_0 mul_mat in a
LLaMA decode graph -- the checker for the neighbour-fused kernels.  Never imported by the product.

  add   ggml_compute_forward_add_f32 (ggml.c:8480-8560): one fp32 addition per element
  silu  ggml_vec_silu_f32 with GGML_SILU_FP16 (ggml.c:122, :3519-3538, table built at :4245-4252):
        y = fp16_to_fp32(table[fp32_to_fp16(x)]),  table[h] = fp32_to_fp16(silu_f32(fp16_to_fp32(h))),
        silu_f32(v) = v / (1.0f + expf(-v))
  mul   ggml_compute_forward_mul_f32 (ggml.c:9160-9240): one fp32 multiplication per element

Parity: pinned through the compiled reference (tests/test_glue.py compares whole layer graphs with these nodes against
the reference's CPU build); numpy's float32 exp may differ from glibc's expf in the last place, which the fp16 rounding
of the table entry hides except on a rounding boundary -- callers allow one fp16 ulp on silu.
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


def rope_table(n_past, n_dims, ne0, npos):
    """cos / sin of the angles ggml_compute_forward_rope_f32 walks through (ggml.c:12771-12813, mode 0): theta = p, then
    theta *= theta_scale per pair, all in float32.  (numpy's float32 cos / sin stand in for the host's cosf / sinf: equal
    or one ulp apart; the product gets its table from the host's libm through the glue.)  -> [npos][ne0 / 2][2]"""
    scale = np.float32(np.power(np.float32(10000.0), np.float32(-2.0) / np.float32(n_dims), dtype=np.float32))
    tab = np.zeros((npos, ne0 // 2, 2), np.float32)
    for i2 in range(npos):
        theta = np.float32(n_past + i2)
        for k in range(ne0 // 2):
            tab[i2, k, 0] = np.cos(theta, dtype=np.float32)
            tab[i2, k, 1] = np.sin(theta, dtype=np.float32)
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
