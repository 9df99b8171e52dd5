/*
 * layer_driver.c -- TEST / BENCH INFRASTRUCTURE: a LLaMA decode graph (n = 1) built with the UNMODIFIED reference
 * graph API exactly as llama.cpp builds it (llama.cpp:1290-1540: rms_norm, mul by the norm weights, wq/wk/wv, wo, add,
 * rms_norm, mul, w3, w1, silu, mul, w2, add per layer; final norm and output matmul), with the offload calls llama.cpp
 * makes for -ngl n_layer + 1 (ggml_cuda_transform_tensor on the weights at "load", ggml_cuda_assign_buffers on the
 * results of the offloaded nodes) and the reference's own ggml_graph_compute.  The attention core between wq/wk/wv and
 * wo (rope, KV cache, KQ soft-max: F16 matmuls and CPU nodes, not on this back-end's path) is replaced by ONE CPU node
 * (q + k + v), so the residency pattern is the real one: q, k, v come back to the host, wo's input goes up, everything
 * else stays where the seam put it.  Weights are random-init Q4_0 (one host copy per distinct shape, a separate
 * upload per layer).  Linked into ggml_glue/_build/libggml_b200_glue_ref.so.
 */
#include "ggml.h"
#include "ggml-cuda.h"
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

struct glue_layer { struct ggml_tensor * attn_norm, * wq, * wk, * wv, * wo, * ffn_norm, * w1, * w2, * w3; };
struct glue_model {
    int n_layer, n_embd, n_ff, n_vocab, offload;
    struct ggml_context * ctx;       /* tensor structs only (no_alloc) */
    struct glue_layer * layers;
    struct ggml_tensor * norm, * output;
    void ** host;                    /* host buffers we own */
    int n_host;
    void * cbuf;                     /* compute buffer for the per-eval graph */
    size_t cbuf_size;
    int cbuf_pinned;
};

static uint64_t g_rng;
static uint32_t rnd(void) { g_rng ^= g_rng << 13; g_rng ^= g_rng >> 7; g_rng ^= g_rng << 17; return (uint32_t)(g_rng >> 16); }

/* random-init Q4_0 [M][K]: uniform nibbles, d = +-(0.6/sqrt(K)) * (2 + u) / 8: a matmul of gain ~1, so a deep chain stays in range */
static void * random_q4_0(int64_t M, int64_t K) {
    const int64_t nb = M * (K / 32);
    unsigned char * p = (unsigned char *)malloc((size_t)nb * 18);
    const float s = 0.6f / sqrtf((float)K);
    for (int64_t i = 0; i < nb; i++) {
        const uint32_t r = rnd();
        float d = s * (2.0f + (float)(r & 0xFFFF) / 65536.0f) / 8.0f;
        if (r & 0x10000) d = -d;
        const ggml_fp16_t h = ggml_fp32_to_fp16(d);
        memcpy(p + 18 * i, &h, 2);
        for (int j = 0; j < 16; j += 4) { const uint32_t q = rnd() ^ (rnd() << 16); memcpy(p + 18 * i + 2 + j, &q, 4); }
    }
    return p;
}

static struct ggml_tensor * make_weight(struct glue_model * m, enum ggml_type type, int64_t ne0, int64_t ne1, void * data, int own) {
    struct ggml_tensor * t = ne1 > 0 ? ggml_new_tensor_2d(m->ctx, type, ne0, ne1) : ggml_new_tensor_1d(m->ctx, type, ne0);
    t->data = data;
    if (own) m->host[m->n_host++] = data;
    if (m->offload) {
        /* llama.cpp:678-685: the loader tags the tensor and hands the bytes over; the callee copies */
        /* GLUE_DRIVER_SPLIT=1: matrices tagged like llama.cpp's LLAMA_BACKEND_OFFLOAD_SPLIT -- row-sharded when a communicator is up */
        const char * split = getenv("GLUE_DRIVER_SPLIT");
        t->backend = (split && atoi(split) && ne1 > 0 && type != GGML_TYPE_F32) ? GGML_BACKEND_GPU_SPLIT : GGML_BACKEND_GPU;
        ggml_cuda_transform_tensor(data, t);
    }
    return t;
}

void glue_model_free(struct glue_model * m);

/* offload: 0 = everything on the reference CPU path, 1 = weights uploaded through the seam */
struct glue_model * glue_model_create(int n_layer, int n_embd, int n_ff, int n_vocab, int offload, unsigned seed) {
    struct glue_model * m = (struct glue_model *)calloc(1, sizeof(*m));
    m->n_layer = n_layer; m->n_embd = n_embd; m->n_ff = n_ff; m->n_vocab = n_vocab; m->offload = offload;
    g_rng = 0x9E3779B97F4A7C15ull ^ ((uint64_t)seed << 20) ^ seed;
    struct ggml_init_params ip = { (size_t)(n_layer * 9 + 8) * (ggml_tensor_overhead() + 64) + (1u << 20), NULL, true };
    m->ctx = ggml_init(ip);
    if (!m->ctx) { free(m); return NULL; }
    if (offload) ggml_cuda_set_scratch_size((size_t)64 << 20);
    m->layers = (struct glue_layer *)calloc((size_t)n_layer, sizeof(struct glue_layer));
    m->host = (void **)calloc((size_t)n_layer * 9 + 16, sizeof(void *));
    /* one host copy per distinct shape when the weights go to the device (timing needs distinct HBM addresses, not
     * distinct values); distinct values per layer on the CPU path and for the small parity models */
    const int share = offload && (int64_t)n_layer * n_embd * n_ff > ((int64_t)1 << 28);
    void * s_attn = NULL, * s_up = NULL, * s_down = NULL;
    for (int il = 0; il < n_layer; il++) {
        struct glue_layer * L = &m->layers[il];
        float * an = (float *)malloc(sizeof(float) * n_embd), * fn = (float *)malloc(sizeof(float) * n_embd);
        for (int i = 0; i < n_embd; i++) { an[i] = 0.5f + (float)(rnd() & 0xFFFF) / 65536.0f; fn[i] = 0.5f + (float)(rnd() & 0xFFFF) / 65536.0f; }
        L->attn_norm = make_weight(m, GGML_TYPE_F32, n_embd, 0, an, 1);
        L->ffn_norm = make_weight(m, GGML_TYPE_F32, n_embd, 0, fn, 1);
        const int fresh = !share || il == 0;
        if (fresh) { s_attn = random_q4_0(n_embd, n_embd); s_up = random_q4_0(n_ff, n_embd); s_down = random_q4_0(n_embd, n_ff); }
        L->wq = make_weight(m, GGML_TYPE_Q4_0, n_embd, n_embd, s_attn, fresh);
        L->wk = make_weight(m, GGML_TYPE_Q4_0, n_embd, n_embd, fresh && !share ? random_q4_0(n_embd, n_embd) : s_attn, fresh && !share);
        L->wv = make_weight(m, GGML_TYPE_Q4_0, n_embd, n_embd, fresh && !share ? random_q4_0(n_embd, n_embd) : s_attn, fresh && !share);
        L->wo = make_weight(m, GGML_TYPE_Q4_0, n_embd, n_embd, fresh && !share ? random_q4_0(n_embd, n_embd) : s_attn, fresh && !share);
        L->w1 = make_weight(m, GGML_TYPE_Q4_0, n_embd, n_ff, s_up, fresh);
        L->w3 = make_weight(m, GGML_TYPE_Q4_0, n_embd, n_ff, fresh && !share ? random_q4_0(n_ff, n_embd) : s_up, fresh && !share);
        L->w2 = make_weight(m, GGML_TYPE_Q4_0, n_ff, n_embd, s_down, fresh);
    }
    float * nw = (float *)malloc(sizeof(float) * n_embd);
    for (int i = 0; i < n_embd; i++) nw[i] = 0.5f + (float)(rnd() & 0xFFFF) / 65536.0f;
    m->norm = make_weight(m, GGML_TYPE_F32, n_embd, 0, nw, 1);
    m->output = make_weight(m, GGML_TYPE_Q4_0, n_embd, n_vocab, random_q4_0(n_vocab, n_embd), 1);
    m->cbuf_size = (size_t)(n_layer * 24 + 16) * (ggml_tensor_overhead() + 256 + sizeof(float) * (size_t)(n_ff > n_embd ? n_ff : n_embd)) +
                   sizeof(float) * (size_t)n_vocab * 4 + ((size_t)16 << 20);
    m->cbuf = offload ? ggml_cuda_host_malloc(m->cbuf_size) : NULL;      /* pinned like llama.cpp's compute buffers (llama-util.h) */
    m->cbuf_pinned = m->cbuf != NULL;
    if (!m->cbuf) m->cbuf = malloc(m->cbuf_size);
    return m;
}

static void off(const struct glue_model * m, int acts, struct ggml_tensor * t) { if (m->offload && acts) ggml_cuda_assign_buffers(t); }

/* One decode step per `steps`: x_in (n_embd floats, host) -> logits (n_vocab floats, host).  device_acts = 1: the
 * graph builder asks for device-resident results like llama.cpp does (ggml_cuda_assign_buffers); 0: every offloaded
 * matmul takes host activations.  Returns the mean wall-clock seconds per step (graph build + ggml_graph_compute). */
double glue_model_eval(struct glue_model * m, const float * x_in, float * logits, int n_threads, int steps, int device_acts) {
    double total = 0.0;
    for (int s = 0; s < steps; s++) {
        const int64_t t0 = ggml_time_us();
        struct ggml_init_params ip = { m->cbuf_size, m->cbuf, false };
        struct ggml_context * c = ggml_init(ip);
        if (!c) return -1.0;
        struct ggml_cgraph gf;
        memset(&gf, 0, sizeof(gf));
        gf.n_threads = n_threads;
        struct ggml_tensor * inpL = ggml_new_tensor_1d(c, GGML_TYPE_F32, m->n_embd);
        memcpy(inpL->data, x_in, sizeof(float) * m->n_embd);
        struct ggml_tensor * cur;
        for (int il = 0; il < m->n_layer; il++) {
            const struct glue_layer * L = &m->layers[il];
            struct ggml_tensor * inpSA = inpL;
            cur = ggml_rms_norm(c, inpL);                 off(m, device_acts, cur);
            cur = ggml_mul(c, cur, L->attn_norm);         off(m, device_acts, cur);
            struct ggml_tensor * q = ggml_mul_mat(c, L->wq, cur);      /* results on the host: rope / KV cache / soft-max are CPU nodes */
            struct ggml_tensor * k = ggml_mul_mat(c, L->wk, cur);
            struct ggml_tensor * v = ggml_mul_mat(c, L->wv, cur);
            struct ggml_tensor * att = ggml_add(c, ggml_add(c, q, k), v);      /* stand-in for the attention core */
            cur = ggml_mul_mat(c, L->wo, att);            off(m, device_acts, cur);
            struct ggml_tensor * inpFF = ggml_add(c, cur, inpSA);      off(m, device_acts, inpFF);
            cur = ggml_rms_norm(c, inpFF);                off(m, device_acts, cur);
            cur = ggml_mul(c, cur, L->ffn_norm);          off(m, device_acts, cur);
            struct ggml_tensor * tmp = ggml_mul_mat(c, L->w3, cur);    off(m, device_acts, tmp);
            cur = ggml_mul_mat(c, L->w1, cur);            off(m, device_acts, cur);
            cur = ggml_silu(c, cur);                      off(m, device_acts, cur);
            cur = ggml_mul(c, cur, tmp);                  off(m, device_acts, cur);
            cur = ggml_mul_mat(c, L->w2, cur);            off(m, device_acts, cur);
            cur = ggml_add(c, cur, inpFF);                off(m, device_acts, cur);
            inpL = cur;
        }
        cur = ggml_rms_norm(c, inpL);                     off(m, device_acts, cur);      /* offload_func_nr, -ngl > n_layer */
        cur = ggml_mul(c, cur, m->norm);                  /* not offloaded by llama.cpp (llama.cpp:1530) */
        cur = ggml_mul_mat(c, m->output, cur);
        ggml_build_forward_expand(&gf, cur);
        ggml_graph_compute(c, &gf);
        memcpy(logits, cur->data, sizeof(float) * m->n_vocab);
        ggml_free(c);
        total += (double)(ggml_time_us() - t0) * 1e-6;
    }
    return total / (steps > 0 ? steps : 1);
}

void glue_model_free(struct glue_model * m) {
    if (!m) return;
    if (m->offload) {
        for (int il = 0; il < m->n_layer; il++) {
            struct glue_layer * L = &m->layers[il];
            struct ggml_tensor * ts[9] = { L->attn_norm, L->wq, L->wk, L->wv, L->wo, L->ffn_norm, L->w1, L->w2, L->w3 };
            for (int i = 0; i < 9; i++) ggml_cuda_free_data(ts[i]);
        }
        ggml_cuda_free_data(m->norm);
        ggml_cuda_free_data(m->output);
    }
    for (int i = 0; i < m->n_host; i++) free(m->host[i]);
    if (m->cbuf_pinned) ggml_cuda_host_free(m->cbuf); else free(m->cbuf);
    free(m->host);
    free(m->layers);
    ggml_free(m->ctx);
    free(m);
}
