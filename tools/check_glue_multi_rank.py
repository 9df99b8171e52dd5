"""GPU box, under torchrun: the seam with a communicator up -- tests/glue/layer_driver.c's layer graph with its matrices tagged
GGML_BACKEND_GPU_SPLIT (row shards per rank, ggml_cuda_transform_tensor), device-resident and host-resident activations,
every rank's logits against the reference CPU path of the same graph.  Not part of the product."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    from conftest import load_binding
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ["GGML_B200_DEVICE"] = str(local)
    os.environ["GLUE_DRIVER_SPLIT"] = "1"
    mod = load_binding()
    torch.cuda.set_device(local)
    mod.init(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    uid = [mod.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    mod.comm_init(rank, world, uid[0])
    import bench
    G = C.CDLL(os.path.join(bench.PKG, "ggml_glue", "_build", "libggml_b200_glue_ref.so"))
    G.glue_model_create.restype = C.c_void_p
    G.glue_model_create.argtypes = [C.c_int] * 5 + [C.c_uint]
    G.glue_model_eval.restype = C.c_double
    G.glue_model_eval.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
    G.glue_model_free.argtypes = [C.c_void_p]

    def logits(dims, offload, acts, x, steps=3):
        m = G.glue_model_create(*dims, offload, 7)
        out = np.zeros(dims[3], np.float32)
        t = G.glue_model_eval(m, x.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), 1, steps, acts)
        G.glue_model_free(m)
        return out, t

    ok = True
    for dims in ((3, 256, 704, 512), (4, 1024, 2752, 4000)):
        x = np.random.default_rng(4).standard_normal(dims[1]).astype(np.float32)
        cpu, _ = logits(dims, 0, 0, x)
        for acts in (0, 1):
            dist.barrier()
            got, t = logits(dims, 1, acts, x, steps=4)
            err = float(np.abs(got - cpu).max() / np.abs(cpu).max())
            good = err <= 2e-3 and int(got.argmax()) == int(cpu.argmax())
            ok = ok and good
            print(f"rank {rank}/{world} dims {dims} device_acts={acts}: max|dlogit|/max|logit| = {err:.3e} {'ok' if good else 'MISMATCH'}  {t * 1e3:.3f} ms per eval, launches {mod.launch_count()}", flush=True)
    dist.barrier()
    dims = (2, 4096, 11008, 32000)
    x = np.random.default_rng(5).standard_normal(dims[1]).astype(np.float32)
    m = G.glue_model_create(*dims, 1, 9)
    out = np.zeros(dims[3], np.float32)
    for acts in (1, 0):
        dist.barrier()
        t = G.glue_model_eval(m, x.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), 1, 12, acts)
        print(f"rank {rank}/{world} 2 layers of LLaMA-7B shapes + output, device_acts={acts}: {t * 1e3:.3f} ms per eval", flush=True)
    G.glue_model_free(m)
    dist.barrier()
    print(f"rank {rank}: {'ALL OK' if ok else 'FAILED'}", flush=True)
    mod.comm_shutdown() if hasattr(mod, "comm_shutdown") else None
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
