// rwkv_driver.cpp -- test driver over the reference's PUBLIC arch/rwkv/rwkv.h API (rwkv_init_from_file, rwkv_eval,
// rwkv_get_logits): RWKV evaluates one token per call over a graph it builds once (rwkv.cpp:1682-1760), so the
// "prompt" is n_prompt calls and the "decode" one more; both logit vectors are returned.  Built twice by
// ggml_glue/Makefile (CPU truth / drop-in with the loader tag of tests/glue/patch_arch.py).  Test infrastructure only.
#include <cstdio>
#include <cstring>

#include "rwkv.h"

extern "C" int tiny_rwkv_run(const char * path, const int * tokens, int n_prompt, int next_token, int n_threads,
                             float * logits_prompt, float * logits_decode, int n_vocab_expected) {
    (void)n_threads;
    rwkv_context_params p = rwkv_context_default_params();
    p.n_ctx = 128;
    p.seed = 1;
    p.use_mmap = false;
    rwkv_context * ctx = rwkv_init_from_file(path, p);
    if (!ctx) return -1;
    const int nv = rwkv_n_vocab(ctx);
    if (nv != n_vocab_expected) { rwkv_free(ctx); return -2; }
    for (int i = 0; i < n_prompt; i++)
        if (rwkv_eval(ctx, tokens[i], NULL) != 0) { rwkv_free(ctx); return -3; }
    memcpy(logits_prompt, rwkv_get_logits(ctx), sizeof(float) * nv);
    if (rwkv_eval(ctx, next_token, NULL) != 0) { rwkv_free(ctx); return -4; }
    memcpy(logits_decode, rwkv_get_logits(ctx), sizeof(float) * nv);
    rwkv_free(ctx);
    return 0;
}
