"""Row-sharded op on real GPUs (skipped on a one-GPU box): ggml_b200_mul_mat_q4_0_sharded -- the peer-to-peer gather
kernel (n = 1) and the NCCL all-gather (n > 1) -- must equal the single-GPU op bit for bit on every rank, under
back-to-back queuing and CUDA-graph replay.  The work is done by tools/check_sharded.py under torchrun, one
process per GPU, as the library is meant to be driven (reference mechanism: GGML_BACKEND_GPU_SPLIT,
ggml-cuda.cu:2780-2797)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:   # noqa: BLE001
        return 0


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_equals_single_gpu(world):
    if _gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29600 + world), os.path.join(ROOT, "tools", "check_sharded.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "CHECK_SHARDED OK" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 8])
def test_sharded_persistent_list_equals_single_gpu(world):
    """ggml_b200_set_create_sharded: one persistent launch per rank with the exchange inside it (tools/check_sharded_set.py):
    the LLaMA-65B shapes, bit-identical to the single-GPU launch on every rank, within 1e-4 of the oracle on rank 0."""
    if _gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29700 + world), os.path.join(ROOT, "tools", "check_sharded_set.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "CHECK_SHARDED_SET OK" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2])
def test_seam_with_row_sharded_weights(world):
    """The reference's graph compute over the seam with a communicator up (tools/check_glue_multi_rank.py): the layer graph's
    matrices tagged GGML_BACKEND_GPU_SPLIT are row-sharded by ggml_cuda_transform_tensor, every matmul node is a sharded op
    (host-resident and device-resident activations, recorded segments included); every rank's logits against the CPU path."""
    if _gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    if not os.path.exists(os.path.join(ROOT, "llama.cpp-q_4_0_b200", "ggml_glue", "_build", "libggml_b200_glue_ref.so")):
        pytest.skip("ggml_glue/_build/libggml_b200_glue_ref.so not built (needs /root/reference at build time)")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29800 + world), os.path.join(ROOT, "tools", "check_glue_multi_rank.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("ALL OK") == world and "MISMATCH" not in r.stdout
