/*
 * graph_driver.c -- TEST INFRASTRUCTURE: drives the UNMODIFIED reference graph API (ggml_mul_mat +
 * ggml_build_forward + ggml_graph_compute, as examples/benchmark/benchmark-matmult.cpp:181-224 does) with the
 * weight tagged for offload, so the op runs through the reference's own seam (ggml_cuda_transform_tensor at load,
 * ggml_cuda_compute_forward inside ggml_compute_forward) into ggml-b200-glue.c.  Linked with the reference
 * ggml.c built -DGGML_USE_CUBLAS; see llama.cpp-q_4_0_b200/ggml_glue/Makefile.
 */
#include "ggml.h"
#include "ggml-cuda.h"
#include <stdlib.h>
#include <string.h>

int ggml_b200_glue_ready(void);

/* wtype: 0 = Q4_0 blocks [M][K/32][18], 1 = F32 [M][K], 3 / 6 / 7 / 8 = Q4_1 / Q5_0 / Q5_1 / Q8_0 blocks (the reference's type values).  offload: 0 = CPU, 1 = GGML_BACKEND_GPU, 2 = GPU_SPLIT.
 * Returns 0, or -1 when the context could not be created. */
int glue_graph_mul_mat(const void * wdata, int wtype, int64_t M, int64_t K, const float * x, int64_t n, float * y,
                       int n_threads, int offload) {
    const enum ggml_type t = wtype == 0 ? GGML_TYPE_Q4_0 : wtype == 1 ? GGML_TYPE_F32 : (enum ggml_type)wtype;
    const size_t wbytes = wtype == 1 ? (size_t)M * K * 4 : (size_t)M * (K / 32) * ggml_type_size(t);
    size_t mem = wbytes + (size_t)n * K * 4 + (size_t)n * M * 4 + (size_t)n * (K / 32) * 40 * 2 + (64u << 20);
    struct ggml_init_params ip = { mem, NULL, false };
    struct ggml_context * ctx = ggml_init(ip);
    if (!ctx) return -1;
    struct ggml_tensor * W = ggml_new_tensor_2d(ctx, t, K, M);
    memcpy(W->data, wdata, wbytes);
    if (offload) {
        /* what llama.cpp's loader does for an offloaded layer (llama.cpp:678-685) */
        W->backend = offload == 2 ? GGML_BACKEND_GPU_SPLIT : GGML_BACKEND_GPU;
        ggml_cuda_transform_tensor(W->data, W);
    }
    struct ggml_tensor * X = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, K, n);
    memcpy(X->data, x, (size_t)n * K * 4);
    struct ggml_tensor * Y = ggml_mul_mat(ctx, W, X);
    struct ggml_cgraph gf = ggml_build_forward(Y);
    gf.n_threads = n_threads;
    ggml_graph_compute(ctx, &gf);
    memcpy(y, Y->data, (size_t)n * M * 4);
    if (offload) ggml_cuda_free_data(W);
    ggml_free(ctx);
    return 0;
}

int glue_ready(void) { return ggml_b200_glue_ready(); }
int glue_cpu_has_cublas(void) { return ggml_cpu_has_cublas(); }
