"""Kernel-3 pipeline timeline (GPU box): arms the clock64 trace of CTA (0,0), runs one 4096x4096 n=512
GEMM and prints, per k-stage / block, when each role passed its hand-off.  Not part of the product."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    mod = bench.load_binding()
    mod.init(0)
    L_ = mod.lib()
    for kv in sys.argv[1:]:
        if "=" in kv:
            k, v = kv.split("=")
            mod.set_tuning(k, int(v))
    mod.set_path(mod.PATH_GEMM)
    M, K, n = 4096, 4096, 512
    rng = np.random.default_rng(3)
    w = mod.Weight(bench.random_q4_0(rng, M, K))
    x = mod.DevBuf.from_numpy(rng.standard_normal((n, K), dtype=np.float32))
    y = mod.DevBuf(n * M * 4)
    for it in range(3):
        if it == 2:
            mod.set_tuning("gemm_trace", 1)
        rc = L_.ggml_b200_mul_mat_q4_0(w.h, C.c_void_p(x.ptr), n, K * 4, C.c_void_p(y.ptr), M * 4)
        assert rc == 0, L_.ggml_b200_last_error()
        mod.synchronize()
    t = np.zeros(4 * 2048, dtype=np.int64)
    rc = L_.ggml_b200_gemm_trace_read(t.ctypes.data_as(C.c_void_p), t.size)
    assert rc == 0
    t = t.reshape(4, 2048)
    t0 = min(v for v in t.ravel() if v > 0)
    rel = lambda v: (v - t0) if v > 0 else -1
    nbk = K // 32
    print("MMA issuer: cycle at which block b was issued (after its d_empty wait); delta to previous")
    prev = 0
    for b in range(0, nbk):
        v = rel(t[1, b])
        if b < 24 or b % 16 == 0 or b > nbk - 6:
            e = [rel(t[2, 8 * b + k]) for k in range(6)]
            print(f"  blk {b:3d}  t={v:8d}  d={v - prev:6d}   epi0: wait_start={e[0]:8d} +d_full={e[1]-e[0]:5d} +ld0={e[2]-e[1]:5d} +fma0={e[3]-e[2]:5d} +ld1={e[4]-e[3]:5d} +fma1={e[5]-e[4]:5d} | warp3 start {rel(t[2, 8*b+6]):8d} warp5 start {rel(t[2, 8*b+7]):8d}")
        prev = v
    print("per stage: MMA got op_full | TMA after raw_empty, op_empty, x_empty | unpack(t0): start, raw_full, waits done, stored")
    for s in range(0, nbk // 4):
        if s < 16 or s % 8 == 0:
            print(f"  st {s:3d}  mma wait_op_full from {rel(t[1, 1024+s]):8d} to {rel(t[1, 512+s]):8d} | tma {rel(t[0, 3*s]):8d} {rel(t[0, 3*s+1]):8d} {rel(t[0, 3*s+2]):8d} | "
                  f"unp {rel(t[3, 4*s]):8d} {rel(t[3, 4*s+1]):8d} {rel(t[3, 4*s+2]):8d} {rel(t[3, 4*s+3]):8d}")
    tot = rel(max(t[1, :nbk]))
    print(f"MMA issue span {tot} cycles = {tot / nbk:.0f} per block")
    mod.shutdown()


if __name__ == "__main__":
    main()
