"""Prefill GEMM timing (GPU box): for each BASELINE n=512 shape, times the whole op (kernel 1 +
kernel 3) and kernel 3 alone (activations pre-quantized) through the C ABI with CUDA events on the
launching stream, cycling over distinct weight copies.  Prints int8 TOPS = 2*M*K*n / t and the
fraction of the dense int8 peaks of record (4500 nominal; 2 x MEASURED_PEAKS bf16).
Writes gpurun_out/bench_gemm.json.  Not part of the product."""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

SHAPES = [(4096, 4096), (11008, 4096), (4096, 11008), (32000, 4096), (4672, 4544), (18176, 4544), (4544, 18176),
          (8192, 8192), (22016, 8192)]


def main():
    n = 512
    tunes = []
    shapes = SHAPES
    for a in sys.argv[1:]:
        if a.startswith("--n="):
            n = int(a[4:])
        elif a == "--quick":
            shapes = SHAPES[:2]
        elif "=" in a:
            tunes.append(a)
    mod = bench.load_binding()
    mod.init(0)
    L_ = mod.lib()
    for kv in tunes:
        k, v = kv.split("=")
        mod.set_tuning(k, int(v))
    stream = torch.cuda.Stream()
    mod.set_stream(stream.cuda_stream)
    mod.set_path(mod.PATH_GEMM)
    rng = np.random.default_rng(1)
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            bf16 = float(json.load(f)["bf16_tflops"])
    except Exception:
        bf16 = 1700.0
    results = []
    for (M, K) in shapes:
        wbytes = 18 * M * K // 32
        copies = max(2, min(8, int(300e6 // wbytes) + 1))
        blk = bench.random_q4_0(rng, M, K)
        hs = []
        for _ in range(copies):
            h = L_.ggml_b200_upload_q4_0(blk.ctypes.data_as(C.c_void_p), M, K, (K // 32) * 18, 0, M)
            assert h, L_.ggml_b200_last_error()
            hs.append(h)
        x = torch.randn(n, K, device="cuda")
        y = torch.zeros(n, M, device="cuda")

        def step():
            for h in hs:
                rc = L_.ggml_b200_mul_mat_q4_0(h, C.c_void_p(x.data_ptr()), n, K * 4, C.c_void_p(y.data_ptr()), M * 4)
                assert rc == 0, L_.ggml_b200_last_error()

        def timed(reps=10):
            with torch.cuda.stream(stream):
                step()
                step()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(reps):
                    step()
                e1.record(stream)
            stream.synchronize()
            return e0.elapsed_time(e1) * 1e3 / (reps * copies)

        us_op = timed()
        mod.set_tuning("skip_quantize", 1)
        us_k3 = timed()
        mod.set_tuning("skip_quantize", 0)
        ops = 2.0 * M * K * n
        r = dict(M=M, K=K, n=n, us_op=us_op, us_gemm=us_k3, tops_op=ops / us_op / 1e6, tops_gemm=ops / us_k3 / 1e6)
        r["frac_nominal_4500"] = r["tops_gemm"] / 4500.0
        r["frac_2x_measured_bf16"] = r["tops_gemm"] / (2 * bf16)
        results.append(r)
        print(f"{M:6d}x{K:<6d} n={n}  op {us_op:8.1f} us {r['tops_op']:7.1f} TOPS | gemm alone {us_k3:8.1f} us {r['tops_gemm']:7.1f} TOPS "
              f"= {r['frac_nominal_4500']*100:5.1f}% of 4500, {r['frac_2x_measured_bf16']*100:5.1f}% of 2x measured bf16", flush=True)
        for h in hs:
            L_.ggml_b200_free_weight(h)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "bench_gemm.json"), "w") as f:
        json.dump(results, f, indent=0)
    mod.shutdown()


if __name__ == "__main__":
    main()
