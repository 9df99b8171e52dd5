// llama_driver.cpp -- test driver over the reference's PUBLIC llama.h API (llama_init_from_file, llama_eval,
// llama_get_logits): loads a model file with the reference's unmodified loader, evaluates a prompt (n > 1) and then
// one more token (n = 1), returns both logit vectors.  Built twice by ggml_glue/Makefile: against the plain CPU
// reference (the truth) and against the reference compiled -DGGML_USE_CUBLAS + ggml-b200-glue.c + libggml_b200.so
// (the drop-in).  Test infrastructure only.
#include <cstdio>
#include <cstring>
#include <vector>

#include "llama.h"

extern "C" int tiny_llama_run(const char * path, int n_gpu_layers, const int * tokens, int n_prompt, int next_token,
                              int n_threads, float * logits_prompt, float * logits_decode, int n_vocab_expected) {
    static bool backend = false;
    if (!backend) { llama_init_backend(false); backend = true; }
    llama_context_params p = llama_context_default_params();
    p.n_ctx = 128;
    p.n_batch = 64;
    p.n_gpu_layers = n_gpu_layers;
    p.seed = 1;
    p.use_mmap = false;
    p.f16_kv = true;
    llama_context * ctx = llama_init_from_file(path, p);
    if (!ctx) return -1;
    const int nv = llama_n_vocab(ctx);
    if (nv != n_vocab_expected) { llama_free(ctx); return -2; }
    std::vector<llama_token> toks(tokens, tokens + n_prompt);
    if (llama_eval(ctx, toks.data(), n_prompt, 0, n_threads) != 0) { llama_free(ctx); return -3; }
    memcpy(logits_prompt, llama_get_logits(ctx), sizeof(float) * nv);
    llama_token t = next_token;
    if (llama_eval(ctx, &t, 1, n_prompt, n_threads) != 0) { llama_free(ctx); return -4; }
    memcpy(logits_decode, llama_get_logits(ctx), sizeof(float) * nv);
    llama_free(ctx);
    return 0;
}

// a prompt, then n_decode single-token evals over a fixed token sequence (the decode loop of examples/main): every
// step's logits, the mean wall-clock milliseconds per decode step, and -- from the B200 build -- how many recorded
// segments the glue replayed and how many it had submitted ahead of the CPU node in front of them.
extern "C" long ggml_b200_glue_replays(void) __attribute__((weak));
extern "C" long ggml_b200_glue_speculation_hits(void) __attribute__((weak));
extern "C" int tiny_llama_decode(const char * path, int n_gpu_layers, const int * tokens, int n_prompt, int n_decode, int n_threads,
                                 float * logits_steps, double * ms_per_step, long * counters, int n_vocab_expected) {
    static bool backend = false;
    if (!backend) { llama_init_backend(false); backend = true; }
    llama_context_params p = llama_context_default_params();
    p.n_ctx = 256;
    p.n_batch = 64;
    p.n_gpu_layers = n_gpu_layers;
    p.seed = 1;
    p.use_mmap = false;
    p.f16_kv = true;
    llama_context * ctx = llama_init_from_file(path, p);
    if (!ctx) return -1;
    const int nv = llama_n_vocab(ctx);
    if (nv != n_vocab_expected || n_prompt + n_decode > 256) { llama_free(ctx); return -2; }
    std::vector<llama_token> toks(tokens, tokens + n_prompt);
    if (llama_eval(ctx, toks.data(), n_prompt, 0, n_threads) != 0) { llama_free(ctx); return -3; }
    const int64_t t0 = llama_time_us();
    for (int i = 0; i < n_decode; i++) {
        llama_token t = (13 * i + 5) % nv;
        if (llama_eval(ctx, &t, 1, n_prompt + i, n_threads) != 0) { llama_free(ctx); return -4; }
        memcpy(logits_steps + (size_t)i * nv, llama_get_logits(ctx), sizeof(float) * nv);
    }
    *ms_per_step = (double)(llama_time_us() - t0) * 1e-3 / (n_decode > 0 ? n_decode : 1);
    counters[0] = ggml_b200_glue_replays ? ggml_b200_glue_replays() : -1;
    counters[1] = ggml_b200_glue_speculation_hits ? ggml_b200_glue_speculation_hits() : -1;
    llama_free(ctx);
    return 0;
}
