// falcon_driver.cpp -- test driver over the reference's PUBLIC arch/falcon/falcon.h API (falcon_init_from_file,
// falcon_eval, falcon_get_logits; falcon.cpp:2189-2260, :2848-2870): loads a model file with the reference's loader,
// evaluates a prompt (n > 1) and then one more token (n = 1), returns both logit vectors.  Built twice by
// ggml_glue/Makefile: against the plain CPU reference (the truth) and against the reference compiled -DGGML_USE_CUBLAS
// with the loader tag (tests/glue/patch_arch.py) + ggml-b200-glue.c + libggml_b200.so.  Test infrastructure only.
#include <cstdio>
#include <cstring>
#include <vector>

#include "falcon.h"

extern "C" int tiny_falcon_run(const char * path, const int * tokens, int n_prompt, int next_token, int n_threads,
                               float * logits_prompt, float * logits_decode, int n_vocab_expected) {
    falcon_context_params p = falcon_context_default_params();
    p.n_ctx = 128;
    p.n_batch = 64;
    p.seed = 1;
    p.use_mmap = false;
    p.f16_kv = true;
    falcon_context * ctx = falcon_init_from_file(path, p);
    if (!ctx) return -1;
    const int nv = falcon_n_vocab(ctx);
    if (nv != n_vocab_expected) { falcon_free(ctx); return -2; }
    std::vector<falcon_token> toks(tokens, tokens + n_prompt);
    if (falcon_eval(ctx, toks.data(), n_prompt, 0, n_threads) != 0) { falcon_free(ctx); return -3; }
    memcpy(logits_prompt, falcon_get_logits(ctx), sizeof(float) * nv);
    falcon_token t = next_token;
    if (falcon_eval(ctx, &t, 1, n_prompt, n_threads) != 0) { falcon_free(ctx); return -4; }
    memcpy(logits_decode, falcon_get_logits(ctx), sizeof(float) * nv);
    falcon_free(ctx);
    return 0;
}
