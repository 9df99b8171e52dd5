"""Generates tests/golden/q4_0_golden.npz by RUNNING THE REFERENCE (oracle/_ref, the unmodified
/root/reference/ggml.c) in the build container.  The fixture travels to the GPU box, where
/root/reference does not exist.  Re-run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402


def edge_rows(P):
    """Activation rows covering the edge set of SURVEY 8(d); K = 64 per case, one case per row."""
    rows = []
    z = np.zeros(64, np.float32)
    rows.append(z.copy())                                              # all-zero blocks
    a = z.copy(); a[5] = 3.25; rows.append(a)                          # single non-zero
    t = np.zeros(64, np.float32)                                       # exact .5 ties with amax = 127
    t[:32] = [127.0, 0.5, 1.5, 2.5, -0.5, -1.5, -2.5, 3.5] + [0.0] * 24
    t[32:] = [-127.0, 62.5, -62.5, 126.5, -126.5, 0.5, -0.5, 1.5] + [0.0] * 24
    rows.append(t)
    l = P.synthetic(64, 0x11, 1.0); l[31] = 40.0; l[63] = -40.0; rows.append(l)   # amax in the last lane
    rows.append(P.synthetic(64, 0x12, 1e-6))                           # d is an fp16 denormal
    rows.append(P.synthetic(64, 0x13, 1e-9))                           # d underflows to 0 in fp16
    rows.append(P.synthetic(64, 0x14, 3e6))                            # d = amax/127 overflows fp16 -> inf
    h = P.synthetic(64, 0x15, 1.0); h[3] = np.float32(8.3e6); rows.append(h)
    rows.append(-np.abs(P.synthetic(64, 0x16, 2.0)))                   # all negative
    i = P.synthetic(64, 0x17, 1.0); i[7] = np.inf; rows.append(i)      # inf input: id = 0, inf*0 = NaN -> -128
    rows.append(P.cos_data(64, 0.0)); rows.append(P.cos_data(64, 1.0))
    return np.stack(rows).astype(np.float32)


def main():
    P, R = oracle.port(), oracle.ref()
    assert R is not None, "needs oracle/_ref (the compiled reference)"
    out = {}
    # 1. the reference's own test generator, tests/test-quantize-fns.cpp:26-30,110-111
    a, b = P.cos_data(4096, 0.0), P.cos_data(4096, 1.0)
    qa, qb = R.quantize_q4_0(a), R.quantize_q8_0(b)
    out["cos_a"], out["cos_b"] = a, b
    out["cos_a_q4_0"], out["cos_b_q8_0"] = qa, qb
    out["cos_a_dequant"] = R.dequantize_q4_0(qa)
    out["cos_dot"] = np.array([R.vec_dot(qa, qb)], np.float32)
    out["cos_S"] = P.block_dots(qa, qb)          # scalar definition; checked == -1034162 below
    assert int(out["cos_S"].sum()) == -1034162 and out["cos_dot"].view(np.uint32)[0] == 0x458BB388
    # 2. a small random op: M = 96, K = 512, n = 5 (weights incl. a negative-max block)
    M, K, n = 96, 512, 5
    wf = P.synthetic(M * K, 0x5EED0001, 0.02).reshape(M, K)
    wf[3, 40] = -0.5
    x = P.synthetic(n * K, 0x5EED0002, 1.0, 97).reshape(n, K)
    W = R.quantize_q4_0_bulk(wf)
    out["op_W"], out["op_X"] = W, x
    out["op_q8"] = R.quantize_q8_0(x)
    out["op_Y"] = R.mul_mat(W, x, n_threads=1)
    out["op_Y_t4"] = R.mul_mat(W, x, n_threads=4)
    # 3. edge rows for the activation quantizer
    e = edge_rows(P)
    out["edge_X"] = e
    out["edge_q8"] = R.quantize_q8_0(e)
    out["edge_q8_scalar_reference"] = R.quantize_q8_0(e, scalar_reference=True)
    # 3b. edge rows for the WEIGHT quantizer (ggml.c:918-957): ties between +a and -a in both orders decide the sign
    # of d, an all-zero block gives d = -0.0, the max element lands on q = 0 and its mirror on 16 -> clamped to 15
    w = np.zeros((8, 64), np.float32)
    w[0, :32] = [1.5, -1.5] + [0.25] * 30;  w[0, 32:] = [-1.5, 1.5] + [-0.25] * 30        # +a first / -a first
    w[1, 5] = 2.0; w[1, 40] = -2.0                                                          # mirrors across the nibble halves
    w[2] = P.synthetic(64, 0x31, 1e-7)                                                      # d below fp16 denormals
    w[3] = P.synthetic(64, 0x32, 4e5)                                                       # d overflows fp16
    w[4] = P.synthetic(64, 0x33, 0.02); w[4, 31] = 0.3; w[4, 63] = -0.3                     # max in the last lane
    w[5] = np.linspace(-1.0, 1.0, 64, dtype=np.float32)                                     # many .5 boundaries
    w[6, :32] = np.arange(32, dtype=np.float32) / 31.0 * 8.0 - 4.0; w[6, 32:] = -w[6, :32]
    w[7] = P.cos_data(64, 2.0)
    out["edge_W"] = w
    out["edge_W_q4_0"] = R.quantize_q4_0_bulk(w)
    assert np.array_equal(out["edge_W_q4_0"], P.quantize_q4_0(w))
    out["edge_W_dequant"] = R.dequantize_q4_0(out["edge_W_q4_0"])
    # 4. fp16 conversion table (65536 halves -> f32 bits) and a sample of f32 -> fp16
    out["fp16_to_fp32_bits"] = np.array([np.float32(R.fp16_to_fp32(h)).view(np.uint32) for h in range(65536)], np.uint32)
    fs = np.concatenate([P.synthetic(4096, 0x21, 1.0), P.synthetic(2048, 0x22, 1e-5), P.synthetic(2048, 0x23, 3e4),
                         np.array([0.0, -0.0, 65504.0, 65519.99, 65520.0, 1e9, 5.9604645e-8, 2.9802322e-8, 2.98023224e-8 * 1.0000001,
                                   6.1035156e-5, 6.0975552e-5, np.inf, -np.inf], np.float32)]).astype(np.float32)
    out["fp32_samples"] = fs
    out["fp32_to_fp16"] = np.array([R.fp32_to_fp16(float(f)) for f in fs], np.uint16)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "q4_0_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
