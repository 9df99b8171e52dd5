"""GPU box: the reference's OWN llama.cpp (loader, llama_eval, graph, CPU attention) on a random-init model with
LLaMA-7B's dimensions (3.9 GB of Q4_0 blocks, the reference's file format), decode loop: milliseconds per token with every
Q4_0 matmul offloaded through the seam (-ngl 99, ggml_glue/_build/libllama_b200.so) against the plain CPU build of the same
sources (libllama_ref_cpu.so, all host threads).  Logits of the first steps compared.  Not part of the product."""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "glue"))
import make_tiny_llama  # noqa: E402

BUILD = os.path.join(ROOT, "llama.cpp-q_4_0_b200", "ggml_glue", "_build")


def run(lib, model, ngl, n_prompt, n_decode, out, env):
    e = dict(os.environ)
    e.update(env)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "glue", "run_tiny_llama.py"), os.path.join(BUILD, lib), model, str(ngl), str(n_prompt), out, str(n_decode)],
                       capture_output=True, text=True, timeout=1500, env=e)
    if r.returncode != 0:
        sys.exit(r.stdout[-2000:] + r.stderr[-3000:])
    for ln in r.stderr.splitlines():
        if ln.startswith("ggml-b200 glue:"):
            print("    " + ln)
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("decode_ms_per_step=")][-1]
    return np.load(out), line


def main():
    d = tempfile.mkdtemp(prefix="llama7b_")
    model = os.path.join(d, "llama7b_q4_0.bin")
    make_tiny_llama.write_model(model, hp=make_tiny_llama.HP_7B, direct_blocks=True)
    print("model", os.path.getsize(model) / 1e9, "GB", flush=True)
    base = {"TINY_LLAMA_N_VOCAB": "32000", "GGML_B200_GLUE_PROFILE": "1"}
    threads = str(min(16, os.cpu_count() or 1))
    gpu, line = run("libllama_b200.so", model, 99, 8, 48, os.path.join(d, "gpu.npy"), dict(base, TINY_LLAMA_THREADS="1"))
    print("B200 seam, 1 host thread:        ", line, flush=True)
    gpu4, line = run("libllama_b200.so", model, 99, 8, 48, os.path.join(d, "gpu4.npy"), dict(base, TINY_LLAMA_THREADS="4"))
    print("B200 seam, 4 host threads:       ", line, flush=True)
    ns, line = run("libllama_b200.so", model, 99, 8, 48, os.path.join(d, "ns.npy"), dict(base, TINY_LLAMA_THREADS="1", GGML_B200_GLUE_SPECULATE="0"))
    print("B200 seam, no run-ahead:         ", line, flush=True)
    ng, line = run("libllama_b200.so", model, 99, 8, 48, os.path.join(d, "ng.npy"), dict(base, TINY_LLAMA_THREADS="1", GGML_B200_GLUE_GRAPHS="0"))
    print("B200 seam, no recorded segments: ", line, flush=True)
    cpu, line = run("libllama_ref_cpu.so", model, 0, 8, 6, os.path.join(d, "cpu.npy"), dict(base, TINY_LLAMA_THREADS=threads))
    print(f"CPU reference build, {threads} threads:", line, flush=True)
    print("bit-identical across the seam's modes:", bool(np.array_equal(gpu, ns) and np.array_equal(gpu, ng)), " 1 vs 4 host threads:", bool(np.array_equal(gpu, gpu4)))
    for i in range(cpu.shape[0]):
        err = float(np.abs(cpu[i] - gpu[i]).max() / np.abs(cpu[i]).max())
        cos = float(np.dot(cpu[i], gpu[i]) / (np.linalg.norm(cpu[i]) * np.linalg.norm(gpu[i])))
        print(f"  step {i}: max|dlogit|/max|logit| = {err:.3e}  cosine = {cos:.6f}  arg-max {int(cpu[i].argmax())} / {int(gpu[i].argmax())}")
    os.remove(model)


if __name__ == "__main__":
    main()
