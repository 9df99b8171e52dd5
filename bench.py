#!/usr/bin/env python
"""bench.py -- Q4_0 x F32 mul_mat throughput on B200 (BASELINE.json metric), one JSON line.

A "step" is one pass of the hot path over one synthetic decode matmul set:
  N = 1 : LLaMA-7B Q4_0 decode matmul set, n = 1 (BASELINE configs[1]): 32 layers x
          (4x 4096x4096 attn, 2x 11008x4096 FFN up/gate, 1x 4096x11008 FFN down) + 32000x4096
          output = 225 matmuls, 3.7 GB of Q4_0 weights streamed once per step (>> 126 MB L2).
  N > 1 : LLaMA-65B decode matmul set (BASELINE configs[4]: 8192x8192, 22016x8192, 8192x22016,
          + 32000x8192 output), every weight row-sharded over the N ranks
          (ggml_b200_shard_rows) and every dst re-assembled by an NCCL all-gather over NVLink.

value   = algorithmic bytes of the set (18*M*K/32 + 4*K + 4*M per matmul) / step time,
          activations already resident in HBM, CUDA-graph replay, CUDA events, max over ranks.
e2e     = the same metric through the reference-facing call ggml_b200_mul_mat_q4_0_host with
          pinned HOST buffers: H2D of x and D2H of y inside the timed region for every matmul.
roofline= the GEMV kernel alone (activations pre-quantized), algorithmic bytes / CUDA-event time,
          against MEASURED_PEAKS.json hbm_gbs.
cpu_baseline = the reference's own CPU code (oracle/_ref, ggml_graph_compute with all host
          threads) on a bounded sample of the same workload.
`--impl reference` times only that CPU arm and prints it with "impl": "reference".
"""
import argparse
import ctypes as C
import importlib.util
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "llama.cpp-q_4_0_b200")
sys.path.insert(0, ROOT)

METRIC = "q4_0_mul_mat_n1_GBps"
UNIT = "GB/s"

WORKLOADS = {
    # name: (n_layers, per-layer [(M, K, count)], [(M, K) once])
    "llama7b_decode_n1": (32, [(4096, 4096, 4), (11008, 4096, 2), (4096, 11008, 1)], [(32000, 4096)]),
    "llama65b_decode_n1_rowsharded": (80, [(8192, 8192, 4), (22016, 8192, 2), (8192, 22016, 1)], [(32000, 8192)]),
}


def gemm_traffic():
    """DRAM bytes of one 4096x4096 n=512 launch of kernel 3 from the committed ncu capture (profiles/traffic.json)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            d = json.load(f)
        return {"bytes_per_launch_4096x4096_n512": d["gemm_dram_bytes_4096x4096_n512"], "algorithmic_read_bytes": 18 * 4096 * 4096 // 32 + 512 * 4096 + 512 * 128 * 4}
    except Exception:   # noqa: BLE001
        return None


def algo_bytes(M, K, n=1):
    return 18 * M * K // 32 + 4 * K * n + 4 * M * n


def matmul_list(workload, layers=None):
    L, per_layer, once = WORKLOADS[workload]
    if layers is not None:
        L = min(L, layers)
    out = []
    for _ in range(L):
        for (M, K, c) in per_layer:
            out += [(M, K)] * c
    out += list(once)
    return out


def load_binding():
    name = "ggml_b200"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, os.path.join(PKG, "ggml_b200.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def random_q4_0(rng, M, K, scale=0.02):
    """random-init Q4_0 blocks: uniform nibbles, fp16 scale ~ |N(0, 0.02)| * 3 / 8 with random sign
    (the scale a block of N(0, 0.02) weights gets from d = max / -8, ggml.c:932).  scale = 0.6 / sqrt(K) gives a
    matmul of gain ~1, so that a chain of 129 of them (chain_ops) stays in fp32 range."""
    nbk = K // 32
    blk = np.empty((M, nbk, 18), dtype=np.uint8)
    blk[:, :, 2:] = rng.integers(0, 256, size=(M, nbk, 16), dtype=np.uint8)
    d = (scale * (2.0 + np.abs(rng.standard_normal((M, nbk), dtype=np.float32))) / 8.0)
    d *= rng.choice(np.array([-1.0, 1.0], np.float32), size=(M, nbk))
    blk[:, :, :2] = d.astype(np.float16).view(np.uint8).reshape(M, nbk, 2)
    return blk


def chain_ops(torch, handles, layer_shapes, L, x0=None):
    """The decode list chained as the model chains it (llama.cpp:1394-1504 without the element-wise ops between the
    matmuls): wq, wk, wv read the layer input; wo reads wq's result; w3, w1 read wo's; w2 reads w1's; the next layer reads
    w2's; the output matmul reads the last layer's.  Every matmul has its own dst.  Works for any per-layer shape list
    [(M, K)] * 7 laid out (attn x4, up/gate x2, down x1).  Returns ([(handle, x_ptr, y_ptr, y_tensor)], input tensor -- keep it alive)."""
    assert len(layer_shapes) == 7
    ops = []
    first = handles[0]
    h_in = x0 if x0 is not None else torch.randn((1, first[2]), dtype=torch.float32, device="cuda")
    cur = h_in
    n_layer_ops = L * 7
    for idx, (h, M, K) in enumerate(handles):
        y = torch.zeros((1, M), dtype=torch.float32, device="cuda")
        if idx < n_layer_ops:
            pos = idx % 7
            if pos == 0:
                layer_in = cur
            src = {0: layer_in, 1: layer_in, 2: layer_in}.get(pos)
            if pos == 3:
                src = ops[idx - 3][3]          # wq's result
            elif pos in (4, 5):
                src = ops[idx - (pos - 3)][3]  # wo's result
            elif pos == 6:
                src = ops[idx - 1][3]          # w1's result (the second of the up/gate pair)
                cur = y
        else:
            src = cur
        assert src.shape[1] == K, (idx, src.shape, K)
        ops.append((h, src.data_ptr(), y.data_ptr(), y))
    return ops, h_in


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy bandwidth)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi sampled DURING the timed regions (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons, pw = [], None, set(), []
        for (t, line) in self.rows:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                if t0 <= t <= t1 + 0.1:
                    sm.append(float(parts[0]))
                    pw.append(float(parts[2]))
                mx = float(parts[1])
            except ValueError:
                continue
            if t0 <= t <= t1 + 0.1:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------- CPU arm
def cpu_reference_arm(workload, steps, warmup, budget_s=20.0):
    """The reference's own CPU path (ggml_graph_compute -> ggml_compute_forward_mul_mat_q_f32 ->
    AVX2 quantize_row_q8_0 + ggml_vec_dot_q4_0_q8_0) from oracle/_ref, all host threads, on a
    bounded sample: 2 layers of the set + the output matmul, distinct weights per matmul
    (so the sample, ~0.3-1 GB, does not sit in the CPU's caches).  Falls back to the scalar C port
    (kind "port", 1 thread, smaller sample) where oracle/_ref is absent."""
    import oracle
    R = oracle.ref()
    cores = os.cpu_count() or 1
    rng = np.random.default_rng(1234)
    L, per_layer, once = WORKLOADS[workload]
    if R is not None:
        shapes = matmul_list(workload, layers=2)
        kind = "reference"
        handles = []
        for (M, K) in shapes:
            h = R.Handle(R, random_q4_0(rng, M, K), 1)
            h.set_x(rng.standard_normal((1, K), dtype=np.float32))
            handles.append(h)
        sample_bytes = sum(algo_bytes(M, K) for (M, K) in shapes)

        def one_step():
            for h in handles:
                h.compute(cores)
        sample = (f"2 of {L} layers + output matmul of {workload} ({len(shapes)} matmuls, {sample_bytes / 1e6:.0f} MB), "
                  f"ggml_graph_compute n_threads={cores}, {'-march=native' if R.native else '-march=x86-64-v3'} build of the unmodified reference")
    else:
        P = oracle.port()
        kind, cores = "port", 1
        shapes = [(M, K) for (M, K, _) in per_layer]
        Ws = [random_q4_0(rng, M, K) for (M, K) in shapes]
        Xs = [rng.standard_normal((1, K), dtype=np.float32) for (_, K) in shapes]
        sample_bytes = sum(algo_bytes(M, K) for (M, K) in shapes)

        def one_step():
            for W, X in zip(Ws, Xs):
                P.mul_mat(W, X, avx2_order=True)
        sample = f"one matmul per distinct shape of {workload} ({sample_bytes / 1e6:.0f} MB), scalar C port, 1 thread"
    for _ in range(max(1, min(warmup, 3))):
        one_step()
    t_start = time.perf_counter()
    done = 0
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        one_step()
        times.append(time.perf_counter() - t0)
        done += 1
        if time.perf_counter() - t_start > budget_s:
            break
    mean = float(np.mean(times))
    if R is not None:
        for h in handles:
            h.close()
    return {"value": sample_bytes / mean / 1e9, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": sample, "steps_timed": done, "ms_per_sample_step": mean * 1e3,
            "best_value": sample_bytes / float(np.min(times)) / 1e9}


def run_reference_impl(args, workload, rank):
    if rank != 0:
        return
    cb = cpu_reference_arm(workload, args.steps, args.warmup, budget_s=60.0)
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": cb["steps_timed"], "warmup": min(args.warmup, 3), "ms_per_step": cb["ms_per_sample_step"],
        "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None,
        "dtype": "int8 dot (q4_0 x q8_0), fp32 scale+accumulate", "data": "synthetic",
        "config": {"workload": workload, "n": 1, "note": "reference CPU path; each step is a bounded sample of the workload"},
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------- GPU arm
class PinnedArray:
    """numpy view over pinned host memory from ggml_b200_host_malloc (ggml_cuda_host_malloc's mirror)."""

    def __init__(self, mod, shape, dtype=np.float32):
        self.mod = mod
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self.ptr = mod.lib().ggml_b200_host_malloc(max(nbytes, 16))
        if not self.ptr:
            raise RuntimeError("pinned allocation failed")
        buf = (C.c_char * max(nbytes, 16)).from_address(self.ptr)
        self.a = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def free(self):
        if self.ptr:
            self.mod.lib().ggml_b200_host_free(self.ptr)
            self.ptr = None


PREFILL_SHAPES = [(4096, 4096, 4), (11008, 4096, 2), (4096, 11008, 1), (32000, 4096, 1)]   # LLaMA-7B layer + output, n = 512


def prefill_leg(mod, stream, rng, steps, warmup, n=512):
    """BASELINE configs[2]: LLaMA-7B Q4_0 prefill, n = 512 tokens, tensor-core GEMM path (kernel 1 + kernel 3).
    One pass = the 7 matmuls of a layer + the output matmul; int8 TOPS = 2*M*K*n / time, CUDA events on the
    launching stream.  Also the same pass through the host-buffer call (H2D of x, D2H of y inside the timing)."""
    import torch
    L_ = mod.lib()
    handles, ops = [], 0.0
    for (M, K, c) in PREFILL_SHAPES:
        blk = random_q4_0(rng, M, K)
        for _ in range(c):
            h = L_.ggml_b200_upload_q4_0(blk.ctypes.data_as(C.c_void_p), M, K, (K // 32) * 18, 0, M)
            if not h:
                raise RuntimeError(L_.ggml_b200_last_error().decode())
            handles.append((h, M, K))
            ops += 2.0 * M * K * n
    Ks = sorted({K for (_, _, K) in handles})
    Ms = sorted({M for (_, M, _) in handles})
    xs = {K: torch.randn(n, K, device="cuda") for K in Ks}
    ys = {M: torch.zeros(n, M, device="cuda") for M in Ms}
    # device-resident leg: buffers as the model graph has them (wq, wk, wv share the normed rows; w3, w1 too; every
    # matmul has its own dst), so same-src1 ops may reuse the quantized activations and overlap (overlap_same_src)
    groups = ["attn_in", "attn_in", "attn_in", "attn_out", "ffn_in", "ffn_in", "ffn_mid", "out_in"]
    assert len(handles) == len(groups)
    xg, dev_ops = {}, []
    for (h, M, K), grp in zip(handles, groups):
        if grp not in xg:
            xg[grp] = xs[K].clone()
        dev_ops.append((h, M, K, xg[grp], torch.zeros(n, M, device="cuda")))

    def one_pass():
        for (h, M, K, xb, yb) in dev_ops:
            rc = L_.ggml_b200_mul_mat_q4_0(h, C.c_void_p(xb.data_ptr()), n, K * 4, C.c_void_p(yb.data_ptr()), M * 4)
            if rc:
                raise RuntimeError(L_.ggml_b200_last_error().decode())
    c0 = mod.launch_count()
    with torch.cuda.stream(stream):
        for _ in range(max(3, warmup)):
            one_pass()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            one_pass()
        e1.record(stream)
    stream.synchronize()
    ms = e0.elapsed_time(e1) / steps
    launches = (mod.launch_count() - c0) // (max(3, warmup) + steps)
    # kernel 3 alone (activations already quantized) for the roofline line
    mod.set_tuning("skip_quantize", 1)
    with torch.cuda.stream(stream):
        one_pass()
        e0.record(stream)
        for _ in range(steps):
            one_pass()
        e1.record(stream)
    stream.synchronize()
    mod.set_tuning("skip_quantize", 0)
    ms_gemm = e0.elapsed_time(e1) / steps
    # host buffers
    xh = {K: PinnedArray(mod, (n, K)) for K in Ks}
    yh = {M: PinnedArray(mod, (n, M)) for M in Ms}
    for K in Ks:
        xh[K].a[:] = xs[K].cpu().numpy()

    def one_pass_host():
        for (h, M, K) in handles:
            rc = L_.ggml_b200_mul_mat_q4_0_host(h, C.c_void_p(xh[K].ptr), n, K * 4, C.c_void_p(yh[M].ptr), M * 4)
            if rc:
                raise RuntimeError(L_.ggml_b200_last_error().decode())
    one_pass_host()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    hsteps = max(2, steps // 4)
    for _ in range(hsteps):
        one_pass_host()
    torch.cuda.synchronize()
    ms_host = (time.perf_counter() - t0) * 1e3 / hsteps
    for p in list(xh.values()) + list(yh.values()):
        p.free()
    for (h, _, _) in handles:
        L_.ggml_b200_free_weight(h)
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            pk = json.load(f)
        bf16, src = float(pk["bf16_tflops"]), "2 x MEASURED_PEAKS.json bf16_tflops (cuBLAS bf16 burst; dense int8 runs at twice the bf16 rate)"
    except Exception:
        bf16, src = 1590.0, "2 x fallback 1.59 PFLOP/s bf16 (B200_PROFILING.md)"
    tops, tops_gemm = ops / (ms * 1e-3) / 1e12, ops / (ms_gemm * 1e-3) / 1e12
    return {
        "metric": "q4_0_mul_mat_n512_int8_TOPS", "value": tops, "unit": "TOPS", "ms_per_pass": ms,
        "config": {"workload": "llama7b_prefill_n512", "n": n, "matmuls_per_pass": len(handles),
                   "shapes": [[M, K, c] for (M, K, c) in PREFILL_SHAPES], "ops_per_pass": ops,
                   "timed": "kernel 1 (quantize_row_q8_0) + kernel 3 (tcgen05 kind::i8 GEMM) per matmul, x/y resident in HBM"},
        "gpu_launches_per_pass": int(launches),
        "e2e": {"value": ops / (ms_host * 1e-3) / 1e12, "unit": "TOPS", "ms_per_pass": ms_host,
                "h2d_bytes_per_pass": sum(4 * K * n for (_, _, K) in handles), "d2h_bytes_per_pass": sum(4 * M * n for (_, M, _) in handles),
                "api": "ggml_b200_mul_mat_q4_0_host (pinned host x/y)"},
        "roofline": {"kernel": "gemm_q4_0_tc_kernel (TMA + nibble unpack + tcgen05.mma kind::i8 + TMEM + scale epilogue)",
                     "bound": "tensor", "achieved": tops_gemm, "peak": 2 * bf16, "unit": "TOP/s", "frac": tops_gemm / (2 * bf16),
                     "frac_of_nominal_4500": tops_gemm / 4500.0, "peak_source": src, "traffic": gemm_traffic(),
                     "ms_per_pass_kernel3": ms_gemm,
                     "note": "the per-block fp16 x fp16 scale epilogue (2 fp32 FMA per int32 S_b on the CUDA cores) bounds this "
                             "formulation at 64/256 = 25% of the tensor rate with 128x128 tiles; see DESIGN.md kernel 3"},
    }



def run_gpu_arm(args, workload, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"      # keep stdout to the one JSON line (the version banner goes to stdout)
    mod = load_binding()
    torch.cuda.set_device(local_rank)
    mod.init(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        uid = [mod.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        mod.comm_init(rank, world, uid[0])

    stream = torch.cuda.Stream()
    mod.set_stream(stream.cuda_stream)
    if args.unfused:
        mod.set_tuning("gemv_fused", 0)
    if args.no_pdl:
        mod.set_tuning("pdl", 0)
    for kv in args.tune:
        k, v = kv.split("=")
        mod.set_tuning(k, int(v))

    L, per_layer, once = WORKLOADS[workload]
    if args.layers:
        L = min(L, args.layers)
    rng = np.random.default_rng(20261104 + 7919 * rank)

    # ---- weights: one random-init layer generated on the host, uploaded once per layer so every
    # layer has its own HBM copy (timing needs distinct addresses, not distinct values)
    t_setup = time.time()
    layer_shapes = []
    for (M, K, c) in per_layer:
        layer_shapes += [(M, K)] * c
    host_blocks = {}
    weights = []        # (Weight, M, K)

    # the C ABI takes the FULL tensor pointer + [row_lo,row_hi); to avoid generating rows this rank
    # never uploads, hand it a pointer displaced by -row_lo rows (only rows [lo,hi) are read).
    def upload_shard(M, K, key):
        lo, hi = mod.shard_rows(M, rank, world)
        if key not in host_blocks:
            host_blocks[key] = random_q4_0(rng, max(hi - lo, 1), K)
        blk = host_blocks[key]
        nb1 = (K // 32) * 18
        base = blk.ctypes.data - lo * nb1
        h = mod.lib().ggml_b200_upload_q4_0(C.c_void_p(base), M, K, nb1, lo, hi)
        if not h:
            raise RuntimeError(mod.lib().ggml_b200_last_error().decode())
        return h

    handles = []
    for layer in range(L):
        for i, (M, K) in enumerate(layer_shapes):
            handles.append((upload_shard(M, K, ("L", i)), M, K))
    for i, (M, K) in enumerate(once):
        handles.append((upload_shard(M, K, ("O", i)), M, K))
    setup_s = time.time() - t_setup
    total_bytes = sum(algo_bytes(M, K) for (_, M, K) in handles)
    weight_bytes = sum(18 * M * K // 32 for (_, M, K) in handles)

    # ---- activations: one x per distinct K, one y per distinct M (device resident)
    Ks = sorted({K for (_, _, K) in handles})
    Ms = sorted({M for (_, M, _) in handles})
    xs = {K: torch.from_numpy(rng.standard_normal((1, K), dtype=np.float32)).cuda() for K in Ks}
    if world > 1:
        for K in Ks:
            dist.broadcast(xs[K], src=0)       # x is replicated on every rank
    ys = {M: torch.zeros((1, M), dtype=torch.float32, device="cuda") for M in Ms}
    torch.cuda.synchronize()
    L_ = mod.lib()
    op = L_.ggml_b200_mul_mat_q4_0_sharded if world > 1 else L_.ggml_b200_mul_mat_q4_0

    # Buffers as the model graph has them (llama.cpp:1394-1504): wq, wk, wv read the same normed row; wo reads the
    # attention output; w3, w1 read the same normed row; w2 reads the gated product; every matmul has its own dst.
    src_group = ["attn_in", "attn_in", "attn_in", "attn_out", "ffn_in", "ffn_in", "ffn_mid"]
    assert len(layer_shapes) == len(src_group)
    xg = {}
    yg = {}
    ops = []
    for idx, (h, M, K) in enumerate(handles):
        pos = idx % len(layer_shapes) if idx < L * len(layer_shapes) else len(layer_shapes) + (idx - L * len(layer_shapes))
        grp = src_group[pos] if pos < len(src_group) else "out_in"
        if grp not in xg:
            xg[grp] = xs[K].clone()
        if pos not in yg:
            yg[pos] = torch.zeros((1, M), dtype=torch.float32, device="cuda")
        assert xg[grp].shape[1] == K and yg[pos].shape[1] == M
        ops.append((h, M, K, xg[grp].data_ptr(), yg[pos].data_ptr()))
    torch.cuda.synchronize()
    if not args.no_overlap:
        # this process owns the stream: nothing but library ops sits between two matmuls that share src1
        mod.set_tuning("overlap_same_src", 1)

    def one_step_per_op():
        for (h, M, K, xp, yp) in ops:
            rc = op(h, C.c_void_p(xp), 1, K * 4, C.c_void_p(yp), M * 4)
            if rc:
                raise RuntimeError(L_.ggml_b200_last_error().decode())

    # ---- single GPU: the whole list as ONE persistent launch (ggml_b200_set_*).  Before it is timed
    # every dst buffer must be bit-identical to the per-op path's; otherwise the per-op kernels are timed instead.
    opset = None
    set_note = None
    if world == 1 and not args.no_set and not args.unfused:
        try:
            with torch.cuda.stream(stream):
                one_step_per_op()
            stream.synchronize()
            want = {pos: y.cpu().numpy().copy() for pos, y in yg.items()}
            for y in yg.values():
                y.zero_()
            torch.cuda.synchronize()
            opset = mod.OpSet([(h, xp, yp) for (h, M, K, xp, yp) in ops])
            with torch.cuda.stream(stream):
                opset.run()
            stream.synchronize()
            def set_check():
                """K <= 4096: bits of the per-op kernel; K > 4096: four warp partials per row in a fixed order."""
                exact = 0
                for pos, y in yg.items():
                    got = y.cpu().numpy()
                    if np.array_equal(got, want[pos]):
                        exact += 1
                        continue
                    K_pos = next(K for (h, M, K, xp, yp) in ops if yp == y.data_ptr())
                    err = float(np.max(np.abs(got - want[pos])) / max(float(np.max(np.abs(want[pos]))), 1e-30))
                    if K_pos <= 4096 or not err <= 1e-5:
                        raise RuntimeError(f"dst buffer {pos} (K = {K_pos}) differs from the per-op path: max|dY|/max|Y| = {err:.3e}")
                return exact
            n_exact = set_check()
            set_note = {"ops": opset.n_ops, "dependency_barriers": opset.barriers, "ring_slots": opset.slots,
                        "checked": f"{n_exact}/{len(yg)} dst buffers bit-identical to ggml_b200_mul_mat_q4_0 per op (all with K <= 4096), "
                                   "the others within 1e-5 of max|y| (fixed-order four-warp sums)"}
        except Exception as e:   # noqa: BLE001
            print(f"[bench] persistent multi-op launch unavailable ({e}); timing one launch per mul_mat", file=sys.stderr)
            opset = None
            torch.cuda.synchronize()      # a trapped kernel leaves a sticky error: fail here, loudly
    if args.probe_set:
        # child process of main(): is the persistent launch healthy on this box (parity above, then graph replays)?
        ok = opset is not None
        if ok:
            gp = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gp, stream=stream):
                mod.set_stream(torch.cuda.current_stream().cuda_stream)
                opset.run()
            mod.set_stream(stream.cuda_stream)
            for y in yg.values():
                y.zero_()
            torch.cuda.synchronize()
            with torch.cuda.stream(stream):
                for _ in range(20):
                    gp.replay()
            stream.synchronize()
            try:
                set_check()
            except RuntimeError as e:
                print(e, flush=True)
                ok = False
        print("probe-set:", "ok" if ok else "FAILED", flush=True)
        os._exit(0 if ok else 3)

    def one_step():
        if opset is not None:
            opset.run()
        else:
            one_step_per_op()

    # ---- sanity (outside any timed region): the last matmul of the set against the oracle
    parity = None
    if rank == 0 and not args.no_check:
        import oracle
        P = oracle.port()
        hW, M, K = handles[0]
        with torch.cuda.stream(stream):
            rc = L_.ggml_b200_mul_mat_q4_0(hW, C.c_void_p(xs[K].data_ptr()), 1, K * 4, C.c_void_p(ys[M].data_ptr()), M * 4)
        stream.synchronize()
        lo, hi = mod.shard_rows(M, rank, world)
        Wh = host_blocks[("L", 0)][: hi - lo]
        Yo = P.mul_mat(Wh, xs[K].cpu().numpy(), avx2_order=True)
        Yg = ys[M].cpu().numpy()[:, lo:hi]
        parity = float(np.max(np.abs(Yg - Yo)) / max(np.max(np.abs(Yo)), 1e-30))
        if not parity <= 1e-4:
            raise RuntimeError(f"bench sanity check failed: max|dY|/max|Y| = {parity:.3e}")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    barrier()      # rank 0 spent time in the CPU check; peers must not sit in a gather kernel waiting for it

    # ---- the persistent launch is timed only where it is the faster of the two paths on this list; both numbers are
    # reported (config.one_launch_per_step).  NOTE on what the list is: every matmul reads an activation buffer that no
    # matmul of the list writes (the model's element-wise ops sit between them and are not part of this set), so the
    # only data hazards are the dst buffers the layers reuse.  One launch per mul_mat serialises on stream order
    # (except ops that share src1); the persistent launch orders by those hazards.  On a list CHAINED like the model
    # (x of an op = dst of an earlier one) it is the slower path: tools/sweep_set.py, profiles/r01_set_sweep_v3*.log, r01_set_sweep_v5_slots.log.
    if opset is not None:
        def quick(fn, n=30):
            with torch.cuda.stream(stream):
                for _ in range(3):
                    fn()
            stream.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(stream):
                e0.record(stream)
                for _ in range(n):
                    fn()
                e1.record(stream)
            stream.synchronize()
            return e0.elapsed_time(e1) / n
        g_po = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g_po, stream=stream):
            mod.set_stream(torch.cuda.current_stream().cuda_stream)
            one_step_per_op()
        mod.set_stream(stream.cuda_stream)
        ms_po, ms_set = quick(g_po.replay), quick(opset.run)
        del g_po
        set_note["ms_per_step_one_launch_per_step"] = ms_set
        set_note["ms_per_step_one_launch_per_mul_mat"] = ms_po
        set_note["GBps_one_launch_per_mul_mat"] = total_bytes / (ms_po * 1e-3) / 1e9
        set_note["timed_path"] = "one launch per step" if ms_set < ms_po else "one launch per mul_mat (faster here)"
        if not ms_set < ms_po:
            opset.free()
            opset = None

    # ---- capture one step in a CUDA graph (launch-bound otherwise: 450 short kernels per step)
    with torch.cuda.stream(stream):
        for _ in range(2):
            one_step()
    stream.synchronize()
    graph = None
    launches_per_step = None
    if not args.no_graph:
        try:
            graph = torch.cuda.CUDAGraph()
            c0 = mod.launch_count()
            with torch.cuda.graph(graph, stream=stream):
                mod.set_stream(torch.cuda.current_stream().cuda_stream)
                one_step()
            launches_per_step = mod.launch_count() - c0
            mod.set_stream(stream.cuda_stream)
        except Exception as e:   # noqa: BLE001
            if rank == 0:
                print(f"[bench] CUDA graph capture failed ({e}); timing direct launches", file=sys.stderr)
            graph = None
            mod.set_stream(stream.cuda_stream)
    if launches_per_step is None:
        c0 = mod.launch_count()
        with torch.cuda.stream(stream):
            one_step()
        stream.synchronize()
        launches_per_step = mod.launch_count() - c0

    def timed(fn, steps, warmup):
        with torch.cuda.stream(stream):
            for _ in range(warmup):
                fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(steps):
                fn()
            e1.record(stream)
        stream.synchronize()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    step_fn = (lambda: graph.replay()) if graph is not None else one_step
    sampler = ClockSampler(local_rank) if rank == 0 else None
    t_clock0 = time.time()
    total_ms = timed(step_fn, args.steps, args.warmup)
    ms_per_step = total_ms / args.steps
    value = total_bytes / (ms_per_step * 1e-3) / 1e9

    # ---- roofline leg: the dominant kernel.  With the fused n = 1 kernel the whole op IS one
    # kernel launch, so its average launch duration is the step time / launches; with --unfused
    # (kernel 1 + kernel 2) the GEMV is re-timed alone on pre-quantized activations.
    roof = None
    if world == 1:
        peak, peak_src = peaks()
        n_launch = len(handles)
        if args.unfused:
            mod.set_tuning("skip_quantize", 1)
            g2 = None
            try:
                if graph is not None:
                    g2 = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g2, stream=stream):
                        mod.set_stream(torch.cuda.current_stream().cuda_stream)
                        one_step()
                    mod.set_stream(stream.cuda_stream)
                gemv_ms = timed((lambda: g2.replay()) if g2 is not None else one_step, args.steps, args.warmup) / args.steps
            finally:
                mod.set_tuning("skip_quantize", 0)
                mod.set_stream(stream.cuda_stream)
            kname = "gemv_q4_0_planes_kernel<1,4> (activations pre-quantized by quantize_q8_0_kernel)"
        else:
            gemv_ms = ms_per_step
            kname = ("gemv_q4_0_set_kernel (persistent: all mul_mats of the step in one launch, quantize_row_q8_0 fused in)" if opset is not None
                     else "gemv_q4_0_fused_kernel (quantize_row_q8_0 fused in; one launch per mul_mat)")
            if opset is not None:
                n_launch = 1
        achieved = total_bytes / (gemv_ms * 1e-3) / 1e9
        roof = {"kernel": kname, "bound": "hbm", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "frac_of_nominal_8000": achieved / 8000.0,
                "peak_source": peak_src, "traffic": None,
                "launches_per_step": n_launch, "avg_launch_us": gemv_ms * 1e3 / n_launch,
                "algorithmic_bytes_per_step": total_bytes,
                "note": ("one persistent launch per step: launch duration = step time" if opset is not None else
                         "launches overlap under programmatic dependent launch; avg = step time / launches")}
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                roof["traffic"] = json.load(f).get("set_dram_bytes_per_launch" if opset is not None else "gemv_dram_bytes_per_launch")
        except Exception:
            pass

    # ---- e2e leg: host buffers through the reference-facing call, H2D + D2H inside the timed region
    e2e = None
    if world == 1:
        xh = {K: PinnedArray(mod, (1, K)) for K in Ks}
        yh = {M: PinnedArray(mod, (1, M)) for M in Ms}
        for K in Ks:
            xh[K].a[:] = xs[K].cpu().numpy()

        def one_step_host():
            for (h, M, K) in handles:
                rc = L_.ggml_b200_mul_mat_q4_0_host(h, C.c_void_p(xh[K].ptr), 1, K * 4, C.c_void_p(yh[M].ptr), M * 4)
                if rc:
                    raise RuntimeError(L_.ggml_b200_last_error().decode())
        e2e_steps = args.steps if args.e2e_steps is None else args.e2e_steps
        for _ in range(max(1, min(args.warmup, 3))):
            one_step_host()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            one_step_host()
        torch.cuda.synchronize()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
        e2e = {"value": total_bytes / (e2e_ms * 1e-3) / 1e9, "unit": UNIT,
               "h2d_bytes_per_step": sum(4 * K for (_, _, K) in handles),
               "d2h_bytes_per_step": sum(4 * M for (_, M, _) in handles),
               "ms_per_step": e2e_ms, "steps": e2e_steps,
               "api": "ggml_b200_mul_mat_q4_0_host (pinned host x/y, synchronous per matmul like the ggml hook)"}
        # the D2H result of the last matmul must equal the resident path's
        if not args.no_check:
            hW, M, K = handles[-1]
            with torch.cuda.stream(stream):
                L_.ggml_b200_mul_mat_q4_0(hW, C.c_void_p(xs[K].data_ptr()), 1, K * 4, C.c_void_p(ys[M].data_ptr()), M * 4)
            stream.synchronize()
            assert np.array_equal(ys[M].cpu().numpy(), yh[M].a), "e2e result differs from resident result"
        for p in list(xh.values()) + list(yh.values()):
            p.free()
    else:
        # multi-GPU e2e: x from pinned host -> every rank's HBM, sharded op + all-gather, y back to host
        xh = {K: PinnedArray(mod, (1, K)) for K in Ks}
        yh = {M: PinnedArray(mod, (1, M)) for M in Ms}
        xt = {K: torch.from_numpy(xh[K].a) for K in Ks}
        yt = {M: torch.from_numpy(yh[M].a) for M in Ms}
        for K in Ks:
            xh[K].a[:] = xs[K].cpu().numpy()

        def one_step_host():
            for (h, M, K) in handles:
                xs[K].copy_(xt[K], non_blocking=True)
                rc = op(h, C.c_void_p(xs[K].data_ptr()), 1, K * 4, C.c_void_p(ys[M].data_ptr()), M * 4)
                if rc:
                    raise RuntimeError(L_.ggml_b200_last_error().decode())
                yt[M].copy_(ys[M], non_blocking=True)
                stream.synchronize()
        e2e_steps = args.steps if args.e2e_steps is None else args.e2e_steps
        with torch.cuda.stream(stream):
            one_step_host()
            barrier()
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                one_step_host()
        barrier()
        e2e_ms_t = torch.tensor([(time.perf_counter() - t0) * 1e3 / e2e_steps], dtype=torch.float64, device="cuda")
        dist.all_reduce(e2e_ms_t, op=dist.ReduceOp.MAX)
        e2e_ms = float(e2e_ms_t.item())
        e2e = {"value": total_bytes / (e2e_ms * 1e-3) / 1e9, "unit": UNIT,
               "h2d_bytes_per_step": sum(4 * K for (_, _, K) in handles) * world,
               "d2h_bytes_per_step": sum(4 * M for (_, M, _) in handles) * world,
               "ms_per_step": e2e_ms, "steps": e2e_steps,
               "api": "pinned host x -> HBM on every rank, ggml_b200_mul_mat_q4_0_sharded, y -> pinned host, per matmul"}
    prefill = None
    if world == 1 and not args.no_prefill:
        prefill = prefill_leg(mod, stream, rng, max(5, min(args.steps // 10, 20)), args.warmup)
    t_clock1 = time.time()
    clocks = sampler.stop(t_clock0, t_clock1) if sampler else None

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_reference_arm(workload, 10, 1, budget_s=20.0)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if world > 1 else "weak",
            "vs_baseline": None, "dtype": "int8 dot (q4_0 x q8_0, dp4a), fp32 scale+accumulate", "data": "synthetic",
            "config": {"workload": workload, "n": 1, "layers": L, "matmuls_per_step": len(handles),
                       "weight_bytes_per_step": weight_bytes, "algorithmic_bytes_per_step": total_bytes,
                       "l2": "inputs larger than L2: the set's weights (%.2f GB per rank) are streamed once per step" % (weight_bytes / world / 1e9),
                       "weights": "random-init Q4_0 blocks (one layer generated on the host, a separate HBM copy per layer)",
                       "layout": "planes (one-time repack at upload)", "cuda_graph": graph is not None,
                       "fused_quantize": not args.unfused, "pdl": not args.no_pdl,
                       "same_src1_overlap": not args.no_overlap, "tune": args.tune,
                       "one_launch_per_step": set_note,
                       "parallelism": ("row-sharded x%d + peer-to-peer gather kernel per matmul (NVLink stores + flags; NCCL only for n > 1)" % world) if world > 1 else "single GPU",
                       "setup_s": round(setup_s, 2), "sanity_max_rel_err": parity},
            "e2e": e2e, "gpu_launches": int(launches_per_step * args.steps), "gpu_launches_per_step": int(launches_per_step),
            "clocks": clocks,
        }
        if roof:
            line["roofline"] = roof
        if prefill:
            line["prefill"] = prefill
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)

    if world > 1:
        # the captured graph holds NCCL kernels of the library's communicator: tearing the communicator down under
        # it can block forever (seen on 2 GPUs), so multi-rank runs leave by the short way once every rank is done
        sys.stdout.flush()
        dist.barrier()
        torch.cuda.synchronize()
        os._exit(0)
    del graph
    if opset is not None:
        opset.free()
    for (h, _, _) in handles:
        L_.ggml_b200_free_weight(h)
    mod.shutdown()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=list(WORKLOADS))
    ap.add_argument("--layers", type=int, default=None, help="use fewer layers of the set (debug)")
    ap.add_argument("--e2e-steps", type=int, default=None)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--unfused", action="store_true", help="kernel 1 + kernel 2 as separate launches")
    ap.add_argument("--no-pdl", action="store_true")
    ap.add_argument("--no-set", action="store_true",
                    help="never time the persistent multi-op launch (ggml_b200_set_*); by default it is probed in a child process, "
                         "checked against the per-op path and timed when it is the faster of the two on this list")
    ap.add_argument("--set", action="store_true", help=argparse.SUPPRESS)      # (the default now; kept for old command lines)
    ap.add_argument("--probe-set", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--no-probe", action="store_true", help="skip the child-process health probe of the persistent launch (profiler runs)")
    ap.add_argument("--no-overlap", action="store_true", help="every matmul waits for the previous one even when they only share src1")
    ap.add_argument("--tune", action="append", default=[], help="key=value for ggml_b200_set_tuning")
    ap.add_argument("--no-check", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-prefill", action="store_true", help="skip the n = 512 tensor-core leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under it
        port = 29500 + (os.getpid() % 2000)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(port)] + sys.argv
        sys.exit(subprocess.call(cmd))
    workload = args.workload or ("llama7b_decode_n1" if args.gpus == 1 else "llama65b_decode_n1_rowsharded")
    if args.impl == "reference":
        run_reference_impl(args, workload, rank)
        return
    if world == 1 and not args.no_set and not args.unfused and not args.probe_set and not args.no_probe:
        # The persistent launch spins on device-side barriers: prove it healthy in a CHILD process first (a trapped
        # kernel poisons its CUDA context), and time one launch per mul_mat if it is not.
        cmd = [sys.executable, os.path.abspath(__file__), "--probe-set", "--set", "--layers", "2", "--workload", workload,
               "--no-cpu", "--no-prefill", "--no-check"] + [a for kv in args.tune for a in ("--tune", kv)]
        try:
            pr = subprocess.run(cmd, timeout=600, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            healthy = pr.returncode == 0
            tail = pr.stdout[-400:]
        except subprocess.TimeoutExpired:
            healthy, tail = False, "probe timed out"
        if not healthy:
            print(f"[bench] persistent multi-op launch failed its probe ({tail!r}); timing one launch per mul_mat", file=sys.stderr)
            args.no_set = True
    run_gpu_arm(args, workload, rank, world, local_rank)


if __name__ == "__main__":
    main()
