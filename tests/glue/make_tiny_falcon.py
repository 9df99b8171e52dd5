"""Writes a random-init Falcon model in the format arch/falcon/falcon.cpp reads (falcon_file_loader, falcon.cpp:423-533:
magic 'ggjt' v1, hparams n_vocab / n_embd / n_head / n_layer / parallel_attn / ftype, vocabulary with scores, then per
tensor n_dims / name_len / type / ne[] / name / pad-to-32 / data), every 2-D weight as Q4_0 blocks (ftype 2), layer-norm
weights and biases F32.  Tensor names and shapes: falcon.cpp:1000-1026 (multi-query: qkv rows = n_embd + 2 * head_dim).

n_layer must be 32 (falcon.cpp:907-918 sizes its buffers by layer count; n_embd < 4544 -> its smallest table).
write_model(..., n_embd=4544, n_head=71) gives Falcon-7B's own shapes (4672x4544, 4544x4544, 18176x4544, 4544x18176)."""
import os
import struct
import sys

import numpy as np

GGML_TYPE_F32, GGML_TYPE_Q4_0 = 0, 2
HP = dict(n_vocab=512, n_embd=256, n_head=4, n_layer=32, parallel_attn=1, ftype=2)


def write_model(path, seed=4321, hp=HP):
    rng = np.random.default_rng(seed)
    E, V, H = hp["n_embd"], hp["n_vocab"], hp["n_head"]
    hd = E // H
    qkv = E + 2 * hd

    def tensor(f, name, data, ttype):
        shape = data.shape if ttype == GGML_TYPE_F32 else (data.shape[0], data.shape[1] * 32)   # rows, K
        ne = shape[::-1]
        nb = name.encode()
        f.write(struct.pack("<III", len(ne), len(nb), ttype))
        f.write(struct.pack("<%dI" % len(ne), *ne))
        f.write(nb)
        f.write(b"\0" * (-f.tell() & 31))
        f.write(np.ascontiguousarray(data).tobytes())

    def q4(rows, cols, std):
        """Q4_0 blocks written directly: weights uniform in [-8 d, 7 d] with d ~ std / 4.6 (no float pass: a 7B-sized file
        is generated in seconds)."""
        blk = rng.integers(0, 256, size=(rows, cols // 32, 18), dtype=np.uint8)
        d = (std / 4.6 * rng.uniform(0.7, 1.3, size=(rows, cols // 32))).astype(np.float16)
        blk[..., 0:2] = d.view(np.uint8).reshape(rows, cols // 32, 2)
        return blk

    def ln(f, prefix):
        tensor(f, prefix + ".weight", (1.0 + 0.1 * rng.standard_normal(E)).astype(np.float32), GGML_TYPE_F32)
        tensor(f, prefix + ".bias", (0.02 * rng.standard_normal(E)).astype(np.float32), GGML_TYPE_F32)

    s = 1.0 / np.sqrt(E)
    with open(path, "wb") as f:
        f.write(struct.pack("<II", 0x67676A74, 1))
        f.write(struct.pack("<6I", V, E, H, hp["n_layer"], hp["parallel_attn"], hp["ftype"]))
        for i in range(V):
            tok = ("<%d>" % i).encode()
            f.write(struct.pack("<I", len(tok)) + tok + struct.pack("<f", -float(i)))
        tensor(f, "transformer.word_embeddings.weight", q4(V, E, 1.0), GGML_TYPE_Q4_0)
        ln(f, "transformer.ln_f")
        tensor(f, "lm_head.weight", q4(V, E, 1.5 * s), GGML_TYPE_Q4_0)
        for l in range(hp["n_layer"]):
            p = "transformer.h.%d" % l
            ln(f, p + ".input_layernorm")
            tensor(f, p + ".self_attention.query_key_value.weight", q4(qkv, E, 1.2 * s), GGML_TYPE_Q4_0)
            tensor(f, p + ".self_attention.dense.weight", q4(E, E, 0.6 * s), GGML_TYPE_Q4_0)
            tensor(f, p + ".mlp.dense_h_to_4h.weight", q4(4 * E, E, 1.2 * s), GGML_TYPE_Q4_0)
            tensor(f, p + ".mlp.dense_4h_to_h.weight", q4(E, 4 * E, 0.3 * s), GGML_TYPE_Q4_0)
    return path


if __name__ == "__main__":
    out = sys.argv[1] if len(sys.argv) > 1 else "/tmp/tiny_falcon_q4_0.bin"
    hp = dict(HP)
    if len(sys.argv) > 2 and sys.argv[2] == "7b":
        hp.update(n_embd=4544, n_head=71)
    write_model(out, hp=hp)
    print("wrote", out, os.path.getsize(out), "bytes")
