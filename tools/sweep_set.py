"""Sweep of the persistent multi-op GEMV (ggml_b200_set_*) on the LLaMA-7B decode list (GPU box): one process,
weights uploaded once, an OpSet per (consumer warps, rows per warp, ring slots) configuration, CUDA-event timing
of back-to-back launches, and the instrumented run's per-CTA clock breakdown.  Not part of the product."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--workload", default="llama7b_decode_n1")
    ap.add_argument("--configs", default="16:0,12:0,8:0,16:0:bar,16:0:nb",
                    help="warps:slots[:nb]; nb = without dependency barriers (NOT the op semantics: streaming ceiling of the kernel only)")
    ap.add_argument("--bench-list", action="store_true",
                    help="buffers as bench.py has them (x from outside the list, one dst per position in the layer, reused by every layer) instead of the chained list")
    ap.add_argument("--own-dst", action="store_true", help="with --bench-list: a dst per op (no write-after-write barriers)")
    ap.add_argument("--one-x", action="store_true", help="with --bench-list: one x per K (runs change only where K changes)")
    args = ap.parse_args()
    import torch
    mod = bench.load_binding()
    mod.init(0)
    stream = torch.cuda.Stream()
    mod.set_stream(stream.cuda_stream)
    L, per_layer, once = bench.WORKLOADS[args.workload]
    L = min(L, args.layers)
    rng = np.random.default_rng(7)
    shapes = []
    for (M, K, c) in per_layer:
        shapes += [(M, K)] * c
    blks = {}
    handles = []
    for layer in range(L):
        for (M, K) in shapes:
            if (M, K) not in blks:
                blks[(M, K)] = bench.random_q4_0(rng, M, K, 0.6 / np.sqrt(K))
            handles.append((mod.Weight(blks[(M, K)]), M, K))
    for (M, K) in once:
        handles.append((mod.Weight(bench.random_q4_0(rng, M, K, 0.6 / np.sqrt(K))), M, K))
    total = sum(bench.algo_bytes(M, K) for (_, M, K) in handles)
    # the list as the model chains it (llama.cpp:1394-1504 minus the element-wise ops): wq/wk/wv read the layer input,
    # wo reads wq's result, w3/w1 read wo's, w2 reads w1's, the next layer reads w2's; every matmul has its own dst
    if args.bench_list:
        grp_of = [0, 0, 0, 1, 2, 2, 3]
        xg, ygb, ops = {}, {}, []
        for idx, (h, M, K) in enumerate(handles):
            pos = idx % 7 if idx < L * 7 else 7 + idx - L * 7
            g_ = grp_of[pos] if pos < 7 else 4
            if args.one_x:
                g_ = K
            if args.own_dst:
                pos = idx
            if g_ not in xg:
                xg[g_] = torch.from_numpy(rng.standard_normal((1, K), dtype=np.float32)).cuda()
            if pos not in ygb:
                ygb[pos] = torch.zeros((1, M), dtype=torch.float32, device="cuda")
            ops.append((h, xg[g_].data_ptr(), ygb[pos].data_ptr(), ygb[pos]))
    else:
        ops, h_in = bench.chain_ops(torch, handles, shapes, L)
    torch.cuda.synchronize()
    print(f"list: {len(ops)} ops, {total / 1e9:.3f} GB algorithmic per run", flush=True)

    def timed(fn, steps):
        with torch.cuda.stream(stream):
            for _ in range(3):
                fn()
        stream.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(steps):
                fn()
            e1.record(stream)
        stream.synchronize()
        return e0.elapsed_time(e1) / steps

    # baseline in the same session: one launch per mul_mat, CUDA graph, same-src1 overlap on
    mod.set_tuning("overlap_same_src", 1)

    def per_op():
        for (w, xp, yp, *_) in ops:
            mod.mul_mat_dev(w, xp, 1, yp)
    with torch.cuda.stream(stream):
        per_op(); per_op()
    stream.synchronize()
    yg = {yp: t for (w, xp, yp, t) in ops}
    ops = [(w, xp, yp) for (w, xp, yp, t) in ops]
    want = {p: y.cpu().numpy().copy() for p, y in yg.items()}
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        mod.set_stream(torch.cuda.current_stream().cuda_stream)
        per_op()
    mod.set_stream(stream.cuda_stream)
    ms = timed(g.replay, args.steps)
    print(f"per-op launches (graph): {ms * 1e3:8.1f} us/run  {total / ms / 1e6:7.1f} GB/s", flush=True)

    for cfg in args.configs.split(","):
        parts_ = cfg.split(":")
        warps, slots = int(parts_[0]), int(parts_[1])
        fly = [int(v[3:]) for v in parts_[2:] if v.startswith("fly")]
        parts_ = [v for v in parts_ if not v.startswith("fly")]
        mod.set_tuning("set_split_barrier", 0 if "ns" in parts_ else 1)
        slack = [int(v[1:]) for v in parts_[2:] if v[0] == "s" and v[1:].isdigit()]
        mod.set_tuning("set_barrier_slack", slack[0] if slack else (1 << 20))
        parts_ = [v for v in parts_ if v != "ns" and not (v[0] == "s" and v[1:].isdigit())]
        mod.set_tuning("set_fly", fly[0] if fly else 0)
        mod.set_tuning("set_link", 0 if parts_[2:] in (["bar"], ["nb"]) else 1)
        mod.set_tuning("set_warps", warps)
        mod.set_tuning("set_slots", slots)
        s = mod.OpSet(ops)
        for y in yg.values():
            y.zero_()
        torch.cuda.synchronize()
        ms = timed(s.run, args.steps)
        same = sum(np.array_equal(y.cpu().numpy(), want[p]) for p, y in yg.items())
        err = max(float(np.max(np.abs(y.cpu().numpy() - want[p])) / max(float(np.max(np.abs(want[p]))), 1e-30)) for p, y in yg.items())
        print(f"set{' NO-BARRIERS' if parts_[2:] == ['nb'] else ' BARRIERS-ONLY' if parts_[2:] == ['bar'] else ''} warps={warps:2d} fly={fly[0] if fly else 0} slots={s.slots:2d} barriers={s.barriers} linked={s.linked_runs}: {ms * 1e3:8.1f} us/run  "
              f"{total / ms / 1e6:7.1f} GB/s  dst buffers bit-identical to per-op: {same}/{len(yg)}, max|dY|/max|Y| = {err:.2e}", flush=True)
        s.profile()
        prof = s.profile().astype(np.float64)
        tot = prof[:, 0]
        names = ["kernel clk", "consumer waits for weights", "dependency barriers", "waiting for x (x warps; tagged x: poll + quantize)", "producer waits for a slot"]
        parts = [f"{names[0]} mean {tot.mean():.0f} max {tot.max():.0f}"]
        for i in range(1, 5):
            parts.append(f"{names[i]} {100 * (prof[:, i] / tot).mean():.1f}% (max {100 * (prof[:, i] / tot).max():.1f}%)")
        parts.append(f"warp 0: {prof[:, 6].mean():.0f} tiles, {(prof[:, 5] / np.maximum(prof[:, 6], 1)).mean():.0f} clk from data ready to y stored per tile "
                     f"({100 * (prof[:, 5] / tot).mean():.1f}% of the kernel)")
        parts.append(f"in the tile loops {100 * (prof[:, 7] / tot).mean():.1f}%")
        parts.append(f"loop: top-of-iteration gaps {100 * (prof[:, 12] / tot).mean():.1f}% over {prof[:, 13].mean():.0f} iterations, after last tile of an op {100 * (prof[:, 14] / tot).mean():.1f}%, next-op descriptor {100 * (prof[:, 15] / tot).mean():.1f}%")
        nb_ = max(s.barriers, 1)
        print("    profile: " + "; ".join(parts), flush=True)
        s.free()
    mod.shutdown()


if __name__ == "__main__":
    main()
