"""Writes a tiny random-init LLaMA model in the reference's own file format (ggjt v3, llama.cpp:405-500: magic,
version, hparams, vocabulary, then per tensor n_dims / name_len / type / ne[] / name / pad-to-32 / data), with every
2-D weight stored as Q4_0 blocks (ftype 2) and the norm weights as F32 -- what `quantize` produces from a real
checkpoint.  The reference's unmodified loader (llama_model_load_internal, llama.cpp:897-1170) reads it, tags the
offloaded layers' matrices GGML_BACKEND_GPU_SPLIT and hands them to ggml_cuda_transform_tensor, i.e. to the glue.

n_layer = 26 makes the loader pick its smallest buffer table (MODEL_3B, llama.cpp:919)."""
import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

GGML_TYPE_F32, GGML_TYPE_Q4_0 = 0, 2
HP = dict(n_vocab=512, n_embd=256, n_mult=64, n_head=4, n_layer=26, n_rot=64, ftype=2)


def n_ff(hp=HP):
    return ((2 * (4 * hp["n_embd"]) // 3 + hp["n_mult"] - 1) // hp["n_mult"]) * hp["n_mult"]    # llama.cpp:930


HP_7B = dict(n_vocab=32000, n_embd=4096, n_mult=256, n_head=32, n_layer=32, n_rot=128, ftype=2)     # LLaMA-7B's own dimensions


def write_model(path, seed=1234, hp=HP, direct_blocks=False):
    """direct_blocks: the Q4_0 blocks are drawn directly (random nibbles, fp16 scale ~ scale / 4.6) instead of quantizing
    random floats -- what makes a 7B-sized file (3.9 GB) a matter of seconds."""
    import oracle
    P = oracle.port()
    rng = np.random.default_rng(seed)
    E, V, F = hp["n_embd"], hp["n_vocab"], n_ff(hp)

    def tensor(f, name, data, ttype):
        shape = data.shape if ttype == GGML_TYPE_F32 else (data.shape[0], data.shape[1] * 32)   # rows, K
        ne = shape[::-1]                                                                         # ne[0] = inner dim
        nb = name.encode()
        f.write(struct.pack("<III", len(ne), len(nb), ttype))
        f.write(struct.pack("<%dI" % len(ne), *ne))
        f.write(nb)
        f.write(b"\0" * (-f.tell() & 31))
        f.write(np.ascontiguousarray(data).tobytes())

    def q4(rows, cols, scale):
        if direct_blocks:
            blk = rng.integers(0, 256, size=(rows, cols // 32, 18), dtype=np.uint8)
            d = (scale / 4.6 * rng.uniform(0.7, 1.3, size=(rows, cols // 32))).astype(np.float16)
            blk[..., 0:2] = d.view(np.uint8).reshape(rows, cols // 32, 2)
            return blk
        w = (rng.standard_normal((rows, cols)) * scale).astype(np.float32)
        return P.quantize_q4_0(w)                                                                # [rows][cols/32][18]

    with open(path, "wb") as f:
        f.write(struct.pack("<II", 0x67676A74, 3))
        f.write(struct.pack("<7I", hp["n_vocab"], hp["n_embd"], hp["n_mult"], hp["n_head"], hp["n_layer"], hp["n_rot"], hp["ftype"]))
        for i in range(V):
            tok = ("<%d>" % i).encode()
            f.write(struct.pack("<I", len(tok)) + tok + struct.pack("<f", -float(i)))
        tensor(f, "tok_embeddings.weight", q4(V, E, 1.0), GGML_TYPE_Q4_0)
        tensor(f, "norm.weight", (1.0 + 0.1 * rng.standard_normal(E)).astype(np.float32), GGML_TYPE_F32)
        tensor(f, "output.weight", q4(V, E, 0.08), GGML_TYPE_Q4_0)
        for l in range(hp["n_layer"]):
            p = "layers.%d." % l
            tensor(f, p + "attention_norm.weight", (1.0 + 0.1 * rng.standard_normal(E)).astype(np.float32), GGML_TYPE_F32)
            for nm in ("wq", "wk", "wv", "wo"):
                tensor(f, p + "attention.%s.weight" % nm, q4(E, E, 0.06), GGML_TYPE_Q4_0)
            tensor(f, p + "ffn_norm.weight", (1.0 + 0.1 * rng.standard_normal(E)).astype(np.float32), GGML_TYPE_F32)
            tensor(f, p + "feed_forward.w1.weight", q4(F, E, 0.06), GGML_TYPE_Q4_0)
            tensor(f, p + "feed_forward.w2.weight", q4(E, F, 0.04), GGML_TYPE_Q4_0)
            tensor(f, p + "feed_forward.w3.weight", q4(F, E, 0.06), GGML_TYPE_Q4_0)
    return path


if __name__ == "__main__":
    out = sys.argv[1] if len(sys.argv) > 1 else "/tmp/tiny_llama_q4_0.bin"
    write_model(out)
    print("wrote", out, os.path.getsize(out), "bytes")
