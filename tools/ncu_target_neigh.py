"""Small ncu target (GPU box): the neighbour-fused decode kernels through the C ABI on LLaMA-7B shapes -- w1/w3 + SiLU*mul
(11008x4096 twice, one kernel), wo + residual (4096x4096), w2 + residual (4096x11008) -- with distinct weight copies per
launch so that every launch streams from HBM.  Not part of the product."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    mod = bench.load_binding()
    mod.init(0)
    rng = np.random.default_rng(5)
    gate = [mod.Weight(bench.random_q4_0(rng, 11008, 4096)) for _ in range(3)]
    up = [mod.Weight(bench.random_q4_0(rng, 11008, 4096)) for _ in range(3)]
    wo = [mod.Weight(bench.random_q4_0(rng, 4096, 4096)) for _ in range(8)]
    w2 = [mod.Weight(bench.random_q4_0(rng, 4096, 11008)) for _ in range(4)]
    x = rng.standard_normal(4096, dtype=np.float32)
    xf = rng.standard_normal(11008, dtype=np.float32)
    r = rng.standard_normal(4096, dtype=np.float32)
    for rep in range(2):
        for i in range(3):
            mod.mul_mat_swiglu(gate[i], up[i], x)
        for w in wo:
            mod.mul_mat_add(w, x, r)
        for w in w2:
            mod.mul_mat_add(w, xf, r)
    mod.shutdown()


if __name__ == "__main__":
    main()
