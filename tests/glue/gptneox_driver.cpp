// gptneox_driver.cpp -- test driver over the reference's PUBLIC arch/gptneox/gptneox.h API (gptneox_init_from_file,
// gptneox_eval, gptneox_get_logits): the same as falcon_driver.cpp for the GPT-NeoX loader and graph (fused QKV with
// biases, parallel residual; gptneox.cpp:1100-1400).  Built twice by ggml_glue/Makefile (CPU truth / drop-in with the
// loader tag of tests/glue/patch_arch.py).  Test infrastructure only.
#include <cstdio>
#include <cstring>
#include <vector>

#include "gptneox.h"

extern "C" int tiny_gptneox_run(const char * path, const int * tokens, int n_prompt, int next_token, int n_threads,
                                float * logits_prompt, float * logits_decode, int n_vocab_expected) {
    gptneox_context_params p = gptneox_context_default_params();
    p.n_ctx = 128;
    p.n_batch = 64;
    p.seed = 1;
    p.use_mmap = false;
    p.f16_kv = true;
    gptneox_context * ctx = gptneox_init_from_file(path, p);
    if (!ctx) return -1;
    const int nv = gptneox_n_vocab(ctx);
    if (nv != n_vocab_expected) { gptneox_free(ctx); return -2; }
    std::vector<gptneox_token> toks(tokens, tokens + n_prompt);
    if (gptneox_eval(ctx, toks.data(), n_prompt, 0, n_threads) != 0) { gptneox_free(ctx); return -3; }
    memcpy(logits_prompt, gptneox_get_logits(ctx), sizeof(float) * nv);
    gptneox_token t = next_token;
    if (gptneox_eval(ctx, &t, 1, n_prompt, n_threads) != 0) { gptneox_free(ctx); return -4; }
    memcpy(logits_decode, gptneox_get_logits(ctx), sizeof(float) * nv);
    gptneox_free(ctx);
    return 0;
}
