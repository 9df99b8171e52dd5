"""Writes a random-init RWKV model in the format arch/rwkv/rwkv.cpp reads (rwkv_file_loader, rwkv.cpp:495-605: magic 'ggjt' v1,
hparams n_vocab / n_ctx / n_embd / n_layer / rescale_every / ftype, vocabulary WITHOUT scores (rwkv.cpp:539-560), tensors as in the other ggjt
files), the 7 matrices of every block and the head as Q4_0 blocks, layer norms and the time-mix / decay vectors F32.
Names and shapes: rwkv.cpp:1157-1217.  n_layer must be 12, 24, 32 or 40 (rwkv.cpp:1062-1073); 12 takes the smallest table."""
import os
import struct
import sys

import numpy as np

GGML_TYPE_F32, GGML_TYPE_Q4_0 = 0, 2
HP = dict(n_vocab=512, n_ctx=128, n_embd=256, n_layer=12, rescale_every=0, ftype=2)


def write_model(path, seed=555, hp=HP):
    rng = np.random.default_rng(seed)
    E, V = hp["n_embd"], hp["n_vocab"]

    def tensor(f, name, data, ttype):
        shape = data.shape if ttype == GGML_TYPE_F32 else (data.shape[0], data.shape[1] * 32)
        ne = shape[::-1]
        nb = name.encode()
        f.write(struct.pack("<III", len(ne), len(nb), ttype))
        f.write(struct.pack("<%dI" % len(ne), *ne))
        f.write(nb)
        f.write(b"\0" * (-f.tell() & 31))
        f.write(np.ascontiguousarray(data).tobytes())

    def q4(rows, cols, std):
        blk = rng.integers(0, 256, size=(rows, cols // 32, 18), dtype=np.uint8)
        d = (std / 4.6 * rng.uniform(0.7, 1.3, size=(rows, cols // 32))).astype(np.float16)
        blk[..., 0:2] = d.view(np.uint8).reshape(rows, cols // 32, 2)
        return blk

    def vec(mean, std, lo=None, hi=None):
        v = mean + std * rng.standard_normal(E)
        if lo is not None:
            v = np.clip(v, lo, hi)
        return v.astype(np.float32)

    s = 1.0 / np.sqrt(E)
    with open(path, "wb") as f:
        f.write(struct.pack("<II", 0x67676A74, 1))
        f.write(struct.pack("<6I", V, hp["n_ctx"], E, hp["n_layer"], hp["rescale_every"], hp["ftype"]))
        for i in range(V):
            tok = ("<%d>" % i).encode()
            f.write(struct.pack("<I", len(tok)) + tok)
        tensor(f, "rwkv.embeddings.weight", q4(V, E, 1.0), GGML_TYPE_Q4_0)
        tensor(f, "rwkv.blocks.0.pre_ln.weight", vec(1.0, 0.1), GGML_TYPE_F32)
        tensor(f, "rwkv.blocks.0.pre_ln.bias", vec(0.0, 0.02), GGML_TYPE_F32)
        tensor(f, "rwkv.ln_out.weight", vec(1.0, 0.1), GGML_TYPE_F32)
        tensor(f, "rwkv.ln_out.bias", vec(0.0, 0.02), GGML_TYPE_F32)
        tensor(f, "head.weight", q4(V, E, 5.0 * s), GGML_TYPE_Q4_0)     # rwkv_get_logits returns soft-max probabilities: keep them unsaturated
        for l in range(hp["n_layer"]):
            p = "rwkv.blocks.%d" % l
            tensor(f, p + ".ln1.weight", vec(1.0, 0.1), GGML_TYPE_F32)
            tensor(f, p + ".ln1.bias", vec(0.0, 0.02), GGML_TYPE_F32)
            for nm in ("time_mix_key", "time_mix_value", "time_mix_receptance"):
                tensor(f, p + ".attention." + nm, vec(0.5, 0.2, 0.05, 0.95), GGML_TYPE_F32)
            tensor(f, p + ".attention.time_first", vec(0.3, 0.2), GGML_TYPE_F32)
            tensor(f, p + ".attention.time_decay", vec(-1.0, 0.5, -3.0, 0.5), GGML_TYPE_F32)
            tensor(f, p + ".attention.key.weight", q4(E, E, 0.8 * s), GGML_TYPE_Q4_0)
            tensor(f, p + ".attention.value.weight", q4(E, E, 1.0 * s), GGML_TYPE_Q4_0)
            tensor(f, p + ".attention.receptance.weight", q4(E, E, 1.0 * s), GGML_TYPE_Q4_0)
            tensor(f, p + ".attention.output.weight", q4(E, E, 0.6 * s), GGML_TYPE_Q4_0)
            tensor(f, p + ".ln2.weight", vec(1.0, 0.1), GGML_TYPE_F32)
            tensor(f, p + ".ln2.bias", vec(0.0, 0.02), GGML_TYPE_F32)
            tensor(f, p + ".feed_forward.time_mix_key", vec(0.5, 0.2, 0.05, 0.95), GGML_TYPE_F32)
            tensor(f, p + ".feed_forward.time_mix_receptance", vec(0.5, 0.2, 0.05, 0.95), GGML_TYPE_F32)
            tensor(f, p + ".feed_forward.key.weight", q4(4 * E, E, 1.0 * s), GGML_TYPE_Q4_0)
            tensor(f, p + ".feed_forward.receptance.weight", q4(E, E, 1.0 * s), GGML_TYPE_Q4_0)
            tensor(f, p + ".feed_forward.value.weight", q4(E, 4 * E, 0.3 * s), GGML_TYPE_Q4_0)
    return path


if __name__ == "__main__":
    out = sys.argv[1] if len(sys.argv) > 1 else "/tmp/tiny_rwkv_q4_0.bin"
    write_model(out)
    print("wrote", out, os.path.getsize(out), "bytes")
