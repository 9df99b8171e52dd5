"""Writes a random-init GPT-NeoX model in the format arch/gptneox/gptneox.cpp reads (gptneox_file_loader,
gptneox.cpp:426-536: magic 'ggjt' v1, hparams n_vocab / n_ctx / n_embd / n_head / n_layer / n_rot / use_parallel_residual /
ftype, vocabulary with scores, tensors as in the other ggjt files), 2-D weights as Q4_0 blocks, layer norms and the
biases of every linear layer F32.  Names and shapes: gptneox.cpp:1000-1026 (fused QKV: 3 * n_embd rows).
n_layer must be 16, 36 or 44 (gptneox.cpp:910-921); 16 with n_embd < 6144 takes the smallest buffer table."""
import os
import struct
import sys

import numpy as np

GGML_TYPE_F32, GGML_TYPE_Q4_0 = 0, 2
HP = dict(n_vocab=512, n_ctx=128, n_embd=256, n_head=4, n_layer=16, n_rot=16, use_parallel_residual=1, ftype=2)


def write_model(path, seed=987, hp=HP):
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

    def vec(n, mean, std):
        return (mean + std * rng.standard_normal(n)).astype(np.float32)

    s = 1.0 / np.sqrt(E)
    with open(path, "wb") as f:
        f.write(struct.pack("<II", 0x67676A74, 1))
        f.write(struct.pack("<8I", V, hp["n_ctx"], E, hp["n_head"], hp["n_layer"], hp["n_rot"], hp["use_parallel_residual"], hp["ftype"]))
        for i in range(V):
            tok = ("<%d>" % i).encode()
            f.write(struct.pack("<I", len(tok)) + tok + struct.pack("<f", -float(i)))
        tensor(f, "gpt_neox.embed_in.weight", q4(V, E, 1.0), GGML_TYPE_Q4_0)
        tensor(f, "gpt_neox.final_layer_norm.weight", vec(E, 1.0, 0.1), GGML_TYPE_F32)
        tensor(f, "gpt_neox.final_layer_norm.bias", vec(E, 0.0, 0.02), GGML_TYPE_F32)
        tensor(f, "embed_out.weight", q4(V, E, 1.5 * s), GGML_TYPE_Q4_0)
        for l in range(hp["n_layer"]):
            p = "gpt_neox.layers.%d" % l
            tensor(f, p + ".input_layernorm.weight", vec(E, 1.0, 0.1), GGML_TYPE_F32)
            tensor(f, p + ".input_layernorm.bias", vec(E, 0.0, 0.02), GGML_TYPE_F32)
            tensor(f, p + ".attention.query_key_value.weight", q4(3 * E, E, 1.2 * s), GGML_TYPE_Q4_0)
            tensor(f, p + ".attention.query_key_value.bias", vec(3 * E, 0.0, 0.02), GGML_TYPE_F32)
            tensor(f, p + ".attention.dense.weight", q4(E, E, 0.6 * s), GGML_TYPE_Q4_0)
            tensor(f, p + ".attention.dense.bias", vec(E, 0.0, 0.02), GGML_TYPE_F32)
            tensor(f, p + ".post_attention_layernorm.weight", vec(E, 1.0, 0.1), GGML_TYPE_F32)
            tensor(f, p + ".post_attention_layernorm.bias", vec(E, 0.0, 0.02), GGML_TYPE_F32)
            tensor(f, p + ".mlp.dense_h_to_4h.weight", q4(4 * E, E, 1.2 * s), GGML_TYPE_Q4_0)
            tensor(f, p + ".mlp.dense_h_to_4h.bias", vec(4 * E, 0.0, 0.02), GGML_TYPE_F32)
            tensor(f, p + ".mlp.dense_4h_to_h.weight", q4(E, 4 * E, 0.3 * s), GGML_TYPE_Q4_0)
            tensor(f, p + ".mlp.dense_4h_to_h.bias", vec(E, 0.0, 0.02), GGML_TYPE_F32)
    return path


if __name__ == "__main__":
    out = sys.argv[1] if len(sys.argv) > 1 else "/tmp/tiny_gptneox_q4_0.bin"
    write_model(out)
    print("wrote", out, os.path.getsize(out), "bytes")
