"""CPU restatement (numpy) of the reference's Q8_0, Q4_1, Q5_0 and Q5_1 paths through ggml_compute_forward_mul_mat_q_f32.

TEST INFRASTRUCTURE ONLY: imported by tests/ (and never by the product).  Pinned against the reference itself
(oracle/_ref, `Ref.type_*`) in tests/test_oracle_siblings.py.

What is restated, with the reference lines followed:
  * block layouts: block_q4_1 {fp16 d, fp16 m, u8 qs[16]} ggml.c:877-883; block_q8_0 {fp16 d, i8 qs[32]} ggml.c:902-907;
    block_q8_1 {f32 d, f32 s, i8 qs[32]} ggml.c:909-915
  * block_q5_0 {fp16 d, u8 qh[4], u8 qs[16]} ggml.c:885-891; block_q5_1 {fp16 d, fp16 m, u8 qh[4], u8 qs[16]} ggml.c:893-900:
    element j = nibble | bit j of qh << 4; Q5_0 value = that - 16 (ggml_vec_dot_q5_0_q8_0 ggml.c:2894-2917, against q8_0),
    Q5_1 value = that, plus m * s (ggml_vec_dot_q5_1_q8_1 ggml.c:3144-3170, against q8_1); quantize_row_q5_0/1_reference
    ggml.c:1000-1092
  * quantize_row_q8_0 / quantize_row_q8_1, AVX2 branch: d = amax / 127, id = 127 / amax (two IEEE divides),
    q = round-half-even(x * id), s = d * sum(q)                                            ggml.c:1192-1250, :1402-1480
  * quantize_row_q4_1_reference (the weight-side codec of the tests' inputs)              ggml.c:959-995
  * quantize_row_q8_0_reference used for Q8_0 WEIGHTS by quantize_row_q (= quantize_row_q8_0, SIMD) -- the tests
    take weight blocks from the reference when it is present, else from these scalar restatements
  * ggml_vec_dot_q8_0_q8_0 and ggml_vec_dot_q4_1_q8_1, AVX2 branch: per block d = d_w * d_x (fp32 product), eight fp32
    lanes acc_l = fma(d, float(sum of 4 adjacent products), acc_l), result = hsum_float_8(acc) (+ summs, summs += m * s
    per block)                                                                              ggml.c:3296-3313, :2680-2708
"""
import numpy as np

QK = 32
TYPE_Q4_1, TYPE_Q5_0, TYPE_Q5_1, TYPE_Q8_0 = 3, 6, 7, 8
BLOCK_BYTES = {TYPE_Q4_1: 20, TYPE_Q5_0: 22, TYPE_Q5_1: 24, TYPE_Q8_0: 34}
DOT_BYTES = {TYPE_Q4_1: 40, TYPE_Q5_0: 34, TYPE_Q5_1: 40, TYPE_Q8_0: 34}       # block_q8_1 where the format has a min term
Q8_1_ACT = (TYPE_Q4_1, TYPE_Q5_1)

f32 = np.float32


def _f16_to_f32(u16: np.ndarray) -> np.ndarray:
    return u16.view(np.float16).astype(np.float32)


def quantize_dot(qtype: int, x: np.ndarray) -> np.ndarray:
    """Activation rows [n][K] f32 -> blocks uint8 [n][K/32][34 (q8_0) | 40 (q8_1)], AVX2 contract."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, K = x.shape
    xb = x.reshape(n, K // QK, QK)
    amax = np.max(np.abs(xb), axis=-1).astype(f32)
    with np.errstate(divide="ignore", invalid="ignore"):
        d = (amax / f32(127.0)).astype(f32)
        idv = np.where(amax != 0, (f32(127.0) / amax).astype(f32), f32(0.0)).astype(f32)
    v = (xb * idv[..., None]).astype(f32)
    q = np.rint(v)                                               # round half to even == _MM_ROUND_NEAREST
    q = np.where(np.isnan(v), -128.0, q)                         # cvtps_epi32(NaN) = INT_MIN -> saturating packs = -128
    q = np.clip(q, -128, 127).astype(np.int8)
    out = np.zeros((n, K // QK, DOT_BYTES[qtype]), dtype=np.uint8)
    if qtype not in Q8_1_ACT:
        out[..., 0:2] = d.astype(np.float16).view(np.uint16)[..., None].view(np.uint8).reshape(n, K // QK, 2)
        out[..., 2:] = q.view(np.uint8)
    else:
        s = (d * q.astype(np.int32).sum(axis=-1).astype(f32)).astype(f32)
        out[..., 0:4] = d.view(np.uint8).reshape(n, K // QK, 4)
        out[..., 4:8] = s.view(np.uint8).reshape(n, K // QK, 4)
        out[..., 8:] = q.view(np.uint8)
    return out


def quantize_q4_1_reference(x: np.ndarray) -> np.ndarray:
    """quantize_row_q4_1_reference (ggml.c:959-995): rows [..., K] f32 -> uint8 [..., K/32, 20]."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    xb = x.reshape(x.shape[:-1] + (x.shape[-1] // QK, QK))
    mn = xb.min(axis=-1).astype(f32)
    mx = xb.max(axis=-1).astype(f32)
    d = ((mx - mn).astype(f32) / f32(15.0)).astype(f32)
    with np.errstate(divide="ignore"):
        idv = np.where(d != 0, (f32(1.0) / d).astype(f32), f32(0.0)).astype(f32)
    t = ((xb - mn[..., None]).astype(f32) * idv[..., None]).astype(f32)
    qi = np.minimum(15, (t + f32(0.5)).astype(f32).astype(np.int32).astype(np.int8).astype(np.int32)).astype(np.uint8)   # (int8_t)(x + 0.5f), MIN(15, .)
    out = np.zeros(xb.shape[:-1] + (20,), dtype=np.uint8)
    out[..., 0:2] = d.astype(np.float16).view(np.uint16)[..., None].view(np.uint8).reshape(xb.shape[:-1] + (2,))
    out[..., 2:4] = mn.astype(np.float16).view(np.uint16)[..., None].view(np.uint8).reshape(xb.shape[:-1] + (2,))
    out[..., 4:] = qi[..., :16] | (qi[..., 16:] << 4)
    return out


def quantize_q8_0_reference(x: np.ndarray) -> np.ndarray:
    """quantize_row_q8_0_reference (ggml.c:1097-1120): d = amax / 127, id = 1 / d, q = roundf(x * id)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    xb = x.reshape(x.shape[:-1] + (x.shape[-1] // QK, QK))
    amax = np.max(np.abs(xb), axis=-1).astype(f32)
    d = (amax / f32(127.0)).astype(f32)
    with np.errstate(divide="ignore"):
        idv = np.where(d != 0, (f32(1.0) / d).astype(f32), f32(0.0)).astype(f32)
    t = (xb * idv[..., None]).astype(f32)
    q = (np.sign(t) * np.floor(np.abs(t).astype(np.float64) + 0.5)).astype(np.int32).astype(np.int8)     # roundf: half away from zero
    out = np.zeros(xb.shape[:-1] + (34,), dtype=np.uint8)
    out[..., 0:2] = d.astype(np.float16).view(np.uint16)[..., None].view(np.uint8).reshape(xb.shape[:-1] + (2,))
    out[..., 2:] = q.view(np.uint8)
    return out


def _put_f16(out, lo, v):
    out[..., lo:lo + 2] = v.astype(np.float16).view(np.uint16)[..., None].view(np.uint8).reshape(v.shape + (2,))


def _q5_pack(out, off, qi):
    """qi uint8 [..., 32] in 0..31 -> qh (32 fifth bits, element j = bit j) at off, nibbles at off + 4 (ggml.c:1030-1041)."""
    qh = np.zeros(qi.shape[:-1], dtype=np.uint32)
    for j in range(32):
        qh |= ((qi[..., j].astype(np.uint32) >> 4) & 1) << np.uint32(j)
    out[..., off:off + 4] = qh[..., None].view(np.uint8).reshape(qi.shape[:-1] + (4,))
    out[..., off + 4:] = (qi[..., :16] & 0x0F) | ((qi[..., 16:] & 0x0F) << 4)


def quantize_q5_0_reference(x: np.ndarray) -> np.ndarray:
    """quantize_row_q5_0_reference (ggml.c:1000-1042): d = max / -16 (max = the signed value of largest magnitude, first one
    wins), q = MIN(31, (int8_t)(x * id + 16.5f))."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    xb = x.reshape(x.shape[:-1] + (x.shape[-1] // QK, QK))
    idx = np.argmax(np.abs(xb), axis=-1)                       # first maximum, as the strict '<' of the loop
    mx = np.take_along_axis(xb, idx[..., None], axis=-1)[..., 0]
    d = (mx / f32(-16.0)).astype(f32)
    with np.errstate(divide="ignore"):
        idv = np.where(d != 0, (f32(1.0) / d).astype(f32), f32(0.0)).astype(f32)
    t = ((xb * idv[..., None]).astype(f32) + f32(16.5)).astype(f32)
    qi = np.minimum(31, t.astype(np.int32).astype(np.int8).astype(np.int32)).astype(np.uint8)
    out = np.zeros(xb.shape[:-1] + (22,), dtype=np.uint8)
    _put_f16(out, 0, d)
    _q5_pack(out, 2, qi)
    return out


def quantize_q5_1_reference(x: np.ndarray) -> np.ndarray:
    """quantize_row_q5_1_reference (ggml.c:1048-1088): d = (max - min) / 31, q = (uint8_t)((x - min) * id + 0.5f)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    xb = x.reshape(x.shape[:-1] + (x.shape[-1] // QK, QK))
    mn = xb.min(axis=-1).astype(f32)
    mx = xb.max(axis=-1).astype(f32)
    d = ((mx - mn).astype(f32) / f32(31.0)).astype(f32)
    with np.errstate(divide="ignore"):
        idv = np.where(d != 0, (f32(1.0) / d).astype(f32), f32(0.0)).astype(f32)
    t = (((xb - mn[..., None]).astype(f32) * idv[..., None]).astype(f32) + f32(0.5)).astype(f32)
    qi = t.astype(np.int32).astype(np.uint8)
    out = np.zeros(xb.shape[:-1] + (24,), dtype=np.uint8)
    _put_f16(out, 0, d)
    _put_f16(out, 2, mn)
    _q5_pack(out, 4, qi)
    return out


def unpack_weights(qtype: int, W: np.ndarray):
    """W uint8 [M][nbk][bytes] -> (quants int32 [M][nbk][32] in ELEMENT order, d f32 [M][nbk], m f32 [M][nbk])."""
    W = np.ascontiguousarray(W, dtype=np.uint8)
    if qtype == TYPE_Q8_0:
        d = _f16_to_f32(W[..., 0:2].copy().view(np.uint16)[..., 0])
        return W[..., 2:].view(np.int8).astype(np.int32), d, np.zeros_like(d)
    d = _f16_to_f32(W[..., 0:2].copy().view(np.uint16)[..., 0])
    if qtype in (TYPE_Q5_0, TYPE_Q5_1):
        off = 2 if qtype == TYPE_Q5_0 else 4
        m = _f16_to_f32(W[..., 2:4].copy().view(np.uint16)[..., 0]) if qtype == TYPE_Q5_1 else np.zeros_like(d)
        qh = W[..., off:off + 4].copy().view(np.uint32)[..., 0]
        bits = ((qh[..., None] >> np.arange(32, dtype=np.uint32)) & 1).astype(np.int32)          # element j = bit j
        qs = W[..., off + 4:]
        q = np.concatenate([(qs & 0x0F).astype(np.int32), (qs >> 4).astype(np.int32)], axis=-1) | (bits << 4)
        return (q - 16 if qtype == TYPE_Q5_0 else q), d, m
    m = _f16_to_f32(W[..., 2:4].copy().view(np.uint16)[..., 0])
    qs = W[..., 4:]
    return np.concatenate([(qs & 0x0F).astype(np.int32), (qs >> 4).astype(np.int32)], axis=-1), d, m


def unpack_dot(qtype: int, Q: np.ndarray):
    """Q uint8 [n][nbk][34|40] -> (quants int32 [n][nbk][32], d f32 [n][nbk], s f32 [n][nbk])."""
    Q = np.ascontiguousarray(Q, dtype=np.uint8)
    if qtype not in Q8_1_ACT:
        d = _f16_to_f32(Q[..., 0:2].copy().view(np.uint16)[..., 0])
        return Q[..., 2:].view(np.int8).astype(np.int32), d, np.zeros_like(d)
    d = Q[..., 0:4].copy().view(np.float32)[..., 0]
    s = Q[..., 4:8].copy().view(np.float32)[..., 0]
    return Q[..., 8:].view(np.int8).astype(np.int32), d, s


def block_dots(qtype: int, W: np.ndarray, Q: np.ndarray) -> np.ndarray:
    """int32 S[n][M][nbk] = sum_j w_j * x_j per block (exact)."""
    wq, _, _ = unpack_weights(qtype, W)
    xq, _, _ = unpack_dot(qtype, Q)
    return np.einsum("mbk,nbk->nmb", wq, xq, optimize=True).astype(np.int32)


def _fma32(a: np.ndarray, b: np.ndarray, c: np.ndarray) -> np.ndarray:
    # a * b is exact in float64 for fp32 operands; one rounding to 53 bits on the add, one to 24: the double rounding
    # differs from a true fp32 fma only in rare ties, far inside the parity tolerance (the bit-exact truth is oracle/_ref)
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(f32)


def mul_mat(qtype: int, W: np.ndarray, X: np.ndarray) -> np.ndarray:
    """The op for a Q8_0 / Q4_1 weight: W [M][nbk][bytes] u8, X [n][K] f32 -> Y [n][M] f32, AVX2 operation order."""
    Q = quantize_dot(qtype, X)
    wq, dw, mw = unpack_weights(qtype, W)
    xq, dx, sx = unpack_dot(qtype, Q)
    M, nbk = dw.shape
    n = dx.shape[0]
    # eight lanes, lane l = sum of the 4 adjacent products 4l..4l+3 (maddubs + madd, ggml.c:641-670)
    lanes = np.einsum("mblk,nblk->nmbl", wq.reshape(M, nbk, 8, 4), xq.reshape(n, nbk, 8, 4), optimize=True).astype(f32)   # exact ints
    acc = np.zeros((n, M, 8), dtype=f32)
    summs = np.zeros((n, M), dtype=f32)
    for b in range(nbk):
        d = (dw[None, :, b] * dx[:, None, b]).astype(f32)                  # d0 * d1, one fp32 product
        acc = _fma32(np.broadcast_to(d[..., None], acc.shape), lanes[:, :, b, :], acc)
        if qtype in Q8_1_ACT:
            summs = (summs + (mw[None, :, b] * sx[:, None, b]).astype(f32)).astype(f32)      # summs += m * s (no contraction: -std=c11)
    # hsum_float_8 (ggml.c:591-597): (lo4 + hi4), then movehl add, then the last pair
    t = (acc[..., 0:4] + acc[..., 4:8]).astype(f32)
    u = (t[..., 0:2] + t[..., 2:4]).astype(f32)
    r = (u[..., 0] + u[..., 1]).astype(f32)
    return (r + summs).astype(f32) if qtype in Q8_1_ACT else r
