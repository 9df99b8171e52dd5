/*
 * ref_shim.c -- thin C front door onto the UNMODIFIED reference ggml.c (compiled from
 * /root/reference where it lies; see oracle/Makefile).  TEST INFRASTRUCTURE ONLY.
 *
 * Exposes the reference's own static functions through its public test hook
 * ggml_internal_get_quantize_fn (ggml.h:1533, ggml.c:1712-1715) and the whole operator through
 * the public graph API (ggml_mul_mat + ggml_build_forward + ggml_graph_compute), the same way
 * examples/benchmark/benchmark-matmult.cpp:181-190,224 drives it.
 */
#include "ggml.h"
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

static int g_inited = 0;

/* ggml_init fills the fp16 table on first call (ggml.c:4245-4252) */
void ref_init(void) {
    if (g_inited) return;
    struct ggml_init_params p = { 1024, NULL, true };
    struct ggml_context * ctx = ggml_init(p);
    ggml_free(ctx);
    g_inited = 1;
}

float    ref_fp16_to_fp32(uint16_t h) { ref_init(); return ggml_fp16_to_fp32(h); }
uint16_t ref_fp32_to_fp16(float f)    { ref_init(); return ggml_fp32_to_fp16(f); }

void ref_quantize_row_q4_0(const float * x, void * y, int k) {
    ref_init();
    ggml_internal_get_quantize_fn(GGML_TYPE_Q4_0).quantize_row_q(x, y, k);
}

void ref_dequantize_row_q4_0(const void * x, float * y, int k) {
    ref_init();
    ggml_internal_get_quantize_fn(GGML_TYPE_Q4_0).dequantize_row_q(x, y, k);
}

/* quantize_row_q_dot of Q4_0 == quantize_row_q8_0 (ggml.c:1623), SIMD branch of the build host */
void ref_quantize_row_q8_0(const float * x, void * y, int k) {
    ref_init();
    ggml_internal_get_quantize_fn(GGML_TYPE_Q4_0).quantize_row_q_dot(x, y, k);
}

void ref_quantize_row_q8_0_reference(const float * x, void * y, int k) {
    ref_init();
    ggml_internal_get_quantize_fn(GGML_TYPE_Q8_0).quantize_row_q_reference(x, y, k);
}

float ref_vec_dot_q4_0_q8_0(int k, const void * x, const void * y) {
    ref_init();
    float s = 0.0f;
    ggml_internal_get_quantize_fn(GGML_TYPE_Q4_0).vec_dot_q(k, &s, x, y);
    return s;
}

/* single-thread vec_dot_q throughput, the method of tests/test-quantize-perf.cpp:85-108 (warm-up passes, then the
 * best of `iters` timed passes): one pass = vec_dot of every row of W [rows][k/32 blocks] against one q8_0 row.
 * Returns the best pass in microseconds; *sink keeps the sums alive. */
int64_t ref_time_vec_dot_q4_0_q8_0(int k, const void * W, int64_t rows, const void * y, int warmup, int iters, float * sink) {
    ref_init();
    quantize_fns_t f = ggml_internal_get_quantize_fn(GGML_TYPE_Q4_0);
    const size_t row_bytes = (size_t)(k / 32) * 18;
    float acc = 0.0f;
    int64_t best = -1;
    for (int it = 0; it < warmup + iters; it++) {
        const int64_t t0 = ggml_time_us();
        for (int64_t r = 0; r < rows; r++) {
            float s = 0.0f;
            f.vec_dot_q(k, &s, (const char *)W + r * row_bytes, y);
            acc += s;
        }
        const int64_t dt = ggml_time_us() - t0;
        if (it >= warmup && (best < 0 || dt < best)) best = dt;
    }
    if (sink) *sink = acc;
    return best;
}

/* bulk weight quantizer, ggml.c:19157-19178 */
size_t ref_quantize_q4_0(const float * src, void * dst, int n, int k) {
    ref_init();
    int64_t hist[16] = {0};
    return ggml_quantize_q4_0(src, dst, n, k, hist);
}

/* ---- whole operator through the graph scheduler ------------------------------------------ */
struct ref_mm {
    struct ggml_context * ctx;
    struct ggml_tensor  * w;
    struct ggml_tensor  * x;
    struct ggml_tensor  * y;
    struct ggml_cgraph    gf;
    int64_t M, K, n;
};

/* W: M*K/32 q4_0 blocks, contiguous.  Returns an opaque handle. */
struct ref_mm * ref_mm_create(int64_t M, int64_t K, int64_t n, const void * W) {
    ref_init();
    struct ref_mm * h = (struct ref_mm *)calloc(1, sizeof(*h));
    const size_t wbytes = (size_t)M * (K / 32) * 18;
    size_t ctx_size = wbytes + (size_t)n * K * 4 + (size_t)n * M * 4
                    + (size_t)n * (K / 32) * 34 + (16u << 20);
    struct ggml_init_params p = { ctx_size, NULL, false };
    h->ctx = ggml_init(p);
    if (!h->ctx) { free(h); return NULL; }
    h->M = M; h->K = K; h->n = n;
    h->w = ggml_new_tensor_2d(h->ctx, GGML_TYPE_Q4_0, K, M);
    memcpy(h->w->data, W, wbytes);
    h->x = ggml_new_tensor_2d(h->ctx, GGML_TYPE_F32, K, n);
    h->y = ggml_mul_mat(h->ctx, h->w, h->x);
    h->gf = ggml_build_forward(h->y);
    return h;
}

/* X: [n][K] f32 contiguous, Y: [n][M] f32 contiguous */
void ref_mm_run(struct ref_mm * h, const float * X, float * Y, int n_threads) {
    memcpy(h->x->data, X, (size_t)h->n * h->K * 4);
    h->gf.n_threads = n_threads;
    ggml_graph_compute(h->ctx, &h->gf);
    memcpy(Y, h->y->data, (size_t)h->n * h->M * 4);
}

/* compute only (inputs already in place): what the CPU baseline times */
void ref_mm_compute(struct ref_mm * h, int n_threads) {
    h->gf.n_threads = n_threads;
    ggml_graph_compute(h->ctx, &h->gf);
}

float * ref_mm_x(struct ref_mm * h) { return (float *)h->x->data; }
float * ref_mm_y(struct ref_mm * h) { return (float *)h->y->data; }

void ref_mm_free(struct ref_mm * h) {
    if (!h) return;
    ggml_free(h->ctx);
    free(h);
}

int64_t ref_time_us(void) { ref_init(); return ggml_time_us(); }

/* ---- sibling formats: the same hooks by type ------------------------------------------------
 * quantize_fns[] (ggml.c:1619-1670): Q4_1 -> quantize_row_q_dot = quantize_row_q8_1, vec_dot_q = ggml_vec_dot_q4_1_q8_1;
 * Q8_0 -> quantize_row_q8_0, ggml_vec_dot_q8_0_q8_0. */
int ref_type_block_bytes(int type) { ref_init(); return (int)ggml_type_size((enum ggml_type)type); }
int ref_type_dot_block_bytes(int type) {
    ref_init();
    return (int)ggml_type_size(ggml_internal_get_quantize_fn((size_t)type).vec_dot_type);
}
void ref_type_quantize_row(int type, const float * x, void * y, int k) {
    ref_init();
    ggml_internal_get_quantize_fn((size_t)type).quantize_row_q(x, y, k);
}
void ref_type_dequantize_row(int type, const void * x, float * y, int k) {
    ref_init();
    ggml_internal_get_quantize_fn((size_t)type).dequantize_row_q(x, y, k);
}
void ref_type_quantize_row_dot(int type, const float * x, void * y, int k) {
    ref_init();
    ggml_internal_get_quantize_fn((size_t)type).quantize_row_q_dot(x, y, k);
}
float ref_type_vec_dot(int type, int k, const void * x, const void * y) {
    ref_init();
    float s = 0.0f;
    ggml_internal_get_quantize_fn((size_t)type).vec_dot_q(k, &s, x, y);
    return s;
}
/* the whole operator for a weight of `type` (W: M * K/32 blocks of that type, contiguous) */
struct ref_mm * ref_mm_create_type(int type, int64_t M, int64_t K, int64_t n, const void * W) {
    ref_init();
    struct ref_mm * h = (struct ref_mm *)calloc(1, sizeof(*h));
    const size_t wbytes = (size_t)M * (K / 32) * ggml_type_size((enum ggml_type)type);
    size_t ctx_size = wbytes + (size_t)n * K * 4 + (size_t)n * M * 4 + (size_t)n * (K / 32) * 40 + (16u << 20);
    struct ggml_init_params p = { ctx_size, NULL, false };
    h->ctx = ggml_init(p);
    if (!h->ctx) { free(h); return NULL; }
    h->M = M; h->K = K; h->n = n;
    h->w = ggml_new_tensor_2d(h->ctx, (enum ggml_type)type, K, M);
    memcpy(h->w->data, W, wbytes);
    h->x = ggml_new_tensor_2d(h->ctx, GGML_TYPE_F32, K, n);
    h->y = ggml_mul_mat(h->ctx, h->w, h->x);
    h->gf = ggml_build_forward(h->y);
    return h;
}

/* ---- the element-wise nodes next to the op, run by the reference's own ggml_graph_compute: what
 * oracle/neighbours.py is pinned against.  op: 0 add(a, b)  1 mul(a, b)  2 silu(a)  3 rms_norm(a) (rows of ne0)
 * 4 rope(a, n_past, n_dims = ne0, mode 0) with a = [ne0][ne1][ne2].  Result into y (same element count as a). */
int ref_neighbour_op(int op, const float * a, const float * b, float * y, int64_t ne0, int64_t ne1, int64_t ne2, int n_past, int n_threads) {
    ref_init();
    const size_t n = (size_t)ne0 * ne1 * ne2;
    struct ggml_init_params p = { n * 4 * 4 + (8u << 20), NULL, false };
    struct ggml_context * ctx = ggml_init(p);
    if (!ctx) return -1;
    struct ggml_tensor * ta = ggml_new_tensor_3d(ctx, GGML_TYPE_F32, ne0, ne1, ne2);
    memcpy(ta->data, a, n * 4);
    struct ggml_tensor * tb = NULL, * r = NULL;
    if (op == 0 || op == 1) { tb = ggml_new_tensor_3d(ctx, GGML_TYPE_F32, ne0, ne1, ne2); memcpy(tb->data, b, n * 4); }
    switch (op) {
        case 0: r = ggml_add(ctx, ta, tb); break;
        case 1: r = ggml_mul(ctx, ta, tb); break;
        case 2: r = ggml_silu(ctx, ta); break;
        case 3: r = ggml_rms_norm(ctx, ta); break;
        case 4: r = ggml_rope(ctx, ta, n_past, (int)ne0, 0, 0); break;
        default: ggml_free(ctx); return -2;
    }
    struct ggml_cgraph gf = ggml_build_forward(r);
    gf.n_threads = n_threads;
    ggml_graph_compute(ctx, &gf);
    memcpy(y, r->data, n * 4);
    ggml_free(ctx);
    return 0;
}
