"""Build step of ggml_glue/Makefile: writes _build/<arch>_seam.cpp = the reference's arch/<arch>/<arch>.cpp, read where it
lies, plus the loader tag INTEGRATION.md section 1 documents (the arch/* loaders create tensors without a backend tag, so
nothing would ever reach the offload seam): after load_all_data, every 2-D Q4_0 matrix that only feeds ggml_mul_mat is
tagged GGML_BACKEND_GPU and handed to ggml_cuda_transform_tensor, exactly what llama.cpp:678-685 does for its own
tensors.  The token-embedding table stays on the host (ggml_get_rows reads it on the CPU).  The output is a build
artefact under _build/ (git-ignored); no reference source enters the repository.
usage: patch_arch.py falcon|gptneox|rwkv <reference source.cpp> <out.cpp>"""
import sys

arch, src, out = sys.argv[1], sys.argv[2], sys.argv[3]
TAGGED = {
    # falcon.cpp:1000-1026
    "falcon": ("for (auto & layer : model.layers) { offload(layer.attn_qkv_w); offload(layer.attn_proj_w); offload(layer.mlp_fc_w); offload(layer.mlp_proj_w); }\n"
               "        offload(model.lmh_w);"),
    # gptneox.cpp:1000-1026
    "gptneox": ("for (auto & layer : model.layers) { offload(layer.c_attn_attn_w); offload(layer.c_attn_proj_w); offload(layer.c_mlp_fc_w); offload(layer.c_mlp_proj_w); }\n"
                "        offload(model.lmh_g);"),
    # rwkv.cpp:1157-1217
    "rwkv": ("for (auto & layer : model.layers) { offload(layer.attn_k_w); offload(layer.attn_v_w); offload(layer.attn_r_w); offload(layer.attn_out_w);\n"
             "                                            offload(layer.ff_k_w); offload(layer.ff_r_w); offload(layer.ff_v_w); }\n"
             "        offload(model.lmh_w);"),
}
text = open(src).read()
anchor = "    ml->load_all_data(progress_callback, progress_callback_user_data, use_mlock ? &lctx.model.mlock_mmap : NULL);\n"
assert text.count(anchor) == 1, src + ": load_all_data call site not found"
tag = anchor + """#ifdef GGML_USE_CUBLAS
    {   // the offload tag (see tests/glue/patch_arch.py)
        auto offload = [](struct ggml_tensor * t) {
            if (t->type == GGML_TYPE_Q4_0 && t->n_dims == 2) { t->backend = GGML_BACKEND_GPU; ggml_cuda_transform_tensor(t->data, t); }
        };
        %s
    }
#endif
""" % TAGGED[arch]
text = text.replace(anchor, tag)
inc = '#include "../../ggml.h"\n' if arch == "rwkv" else '#include "../ggml.h"\n'
assert text.count(inc) == 1
text = text.replace(inc, inc + '#ifdef GGML_USE_CUBLAS\n#include "ggml-cuda.h"\n#endif\n')
open(out, "w").write(text)
