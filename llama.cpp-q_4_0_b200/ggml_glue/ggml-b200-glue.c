/*
 * ggml-b200-glue.c -- the reference's own GPU offload seam, re-implemented on top of the C ABI of
 * libggml_b200.so (include/ggml-b200.h).
 *
 * The reference reaches its GPU back-end through the functions declared in ggml-cuda.h and, inside
 * ggml.c, through three call sites guarded by GGML_USE_CUBLAS: ggml_init_cublas() at the first
 * ggml_init (ggml.c:4282-4286), the pre-hook ggml_cuda_compute_forward() at the top of
 * ggml_compute_forward (ggml.c:15645-15652) and ggml_cuda_can_mul_mat() in the task planner
 * (ggml.c:17283-17287); the model loaders call ggml_cuda_transform_tensor / ggml_cuda_free_data
 * (llama.cpp:678-685, :241-245).  This file defines EVERY symbol of ggml-cuda.h, so an UNMODIFIED
 * ggml.c / llama.cpp built with -DGGML_USE_CUBLAS links against this object + libggml_b200.so
 * instead of ggml-cuda.o: zero source changes on the reference side.
 *
 * Scope: only GGML_OP_MUL_MAT with a GGML_TYPE_Q4_0 src0 that was uploaded through
 * ggml_cuda_transform_tensor and an F32 src1 is taken; every other node returns false from the hook
 * and runs on the reference's CPU code unchanged.  Tensors of any other type that a loader tags
 * GGML_BACKEND_GPU are kept on the host (copied, re-tagged GGML_BACKEND_CPU) so real model graphs
 * keep working.  There is no CPU fallback for the taken op: a failing device call prints and aborts
 * like the reference's CUDA_CHECK (ggml-cuda.cu:22-30).
 *
 * Built only where the reference headers exist (ggml_glue/Makefile, -I$(REF)); plain C.
 */
#include "ggml-cuda.h"   /* the reference's header: struct ggml_tensor_extra_gpu + prototypes */
#include "ggml.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/ggml-b200.h"

#define B200_MAGIC_WEIGHT ((void *)(uintptr_t)0xB200B200u)   /* extra->data_device[1]: data_device[0] is a ggml_b200_weight */
#define B200_MAGIC_HOST   ((void *)(uintptr_t)0xB200C0DEu)   /* extra->data_device[1]: data_device[0] is our host copy */

static int    g_ready = 0;          /* ggml_b200_init succeeded */
static int    g_main_device = 0;
static float *g_x_dev = NULL, *g_y_dev = NULL;   /* staging for the multi-rank path */
static size_t g_x_cap = 0, g_y_cap = 0;

#define B200_CHECK(call)                                                                                   \
    do {                                                                                                   \
        int rc_ = (call);                                                                                  \
        if (rc_ != GGML_B200_OK) {                                                                         \
            fprintf(stderr, "ggml-b200 error %d at %s:%d: %s\n  %s\n", rc_, __FILE__, __LINE__, #call,     \
                    ggml_b200_last_error());                                                               \
            exit(1);                                                                                       \
        }                                                                                                  \
    } while (0)

/* ggml_init_cublas (ggml-cuda.cu:1826-1861): called once from the first ggml_init.  On a host without a
 * usable B200 this leaves the back-end off: nothing can be uploaded, so nothing is ever offloaded. */
void ggml_init_cublas(void) {
    if (g_ready) return;
    const char * dev = getenv("GGML_B200_DEVICE");
    if (dev) g_main_device = atoi(dev);
    if (ggml_b200_device_count() <= 0) return;
    if (ggml_b200_init(g_main_device) == GGML_B200_OK) g_ready = 1;
    else fprintf(stderr, "ggml-b200: init failed, Q4_0 offload disabled: %s\n", ggml_b200_last_error());
}

/* ggml_cuda_set_main_device (ggml-cuda.cu:2943): one process drives one GPU here */
void ggml_cuda_set_main_device(int main_device) {
    if (g_ready && main_device != g_main_device)
        fprintf(stderr, "ggml-b200: already initialised on device %d; ignoring main_device=%d (set GGML_B200_DEVICE)\n", g_main_device, main_device);
    else g_main_device = main_device;
}

/* ggml_cuda_set_tensor_split (ggml-cuda.cu:1863-1882): the row split is fixed by (rank, world) of the NCCL
 * communicator (ggml_b200_shard_rows); fractions from the command line are not used */
void ggml_cuda_set_tensor_split(const float * tensor_split) { (void)tensor_split; }
void ggml_cuda_set_scratch_size(size_t scratch_size) { (void)scratch_size; }   /* activations stay on the host */
void ggml_cuda_free_scratch(void) {}

/* activations are not kept device-resident by this back-end: leave the tensor on the CPU (ggml-cuda.cu:2830-2935) */
void ggml_cuda_assign_buffers(struct ggml_tensor * tensor) { (void)tensor; }
void ggml_cuda_assign_buffers_no_scratch(struct ggml_tensor * tensor) { (void)tensor; }
void ggml_cuda_assign_buffers_force_inplace(struct ggml_tensor * tensor) { (void)tensor; }

/* pinned host memory (ggml-cuda.cu:1884-1905); NULL makes the caller fall back to malloc */
void * ggml_cuda_host_malloc(size_t size) { return g_ready ? ggml_b200_host_malloc(size) : NULL; }
void   ggml_cuda_host_free(void * ptr) { if (g_ready) ggml_b200_host_free(ptr); }

static ggml_b200_weight * weight_of(const struct ggml_tensor * t) {
    if (t == NULL || t->extra == NULL) return NULL;
    if (t->backend != GGML_BACKEND_GPU && t->backend != GGML_BACKEND_GPU_SPLIT) return NULL;
    const struct ggml_tensor_extra_gpu * e = (const struct ggml_tensor_extra_gpu *)t->extra;
    return e->data_device[1] == B200_MAGIC_WEIGHT ? (ggml_b200_weight *)e->data_device[0] : NULL;
}

/* ggml_cuda_transform_tensor (ggml-cuda.cu:2766-2809): the loader hands over the host bytes of a tensor it
 * tagged GPU / GPU_SPLIT.  The callee must copy: the caller frees `data` right after (llama.cpp:682-684). */
void ggml_cuda_transform_tensor(void * data, struct ggml_tensor * tensor) {
    struct ggml_tensor_extra_gpu * extra = (struct ggml_tensor_extra_gpu *)calloc(1, sizeof(*extra));
    GGML_ASSERT(extra != NULL);
    const int q4_2d = tensor->type == GGML_TYPE_Q4_0 && tensor->ne[2] == 1 && tensor->ne[3] == 1 && (tensor->ne[0] % 64) == 0;
    if (g_ready && q4_2d) {
        const int64_t K = tensor->ne[0], M = tensor->ne[1];
        int64_t lo = 0, hi = M;
        if (tensor->backend == GGML_BACKEND_GPU_SPLIT && ggml_b200_comm_world() > 1)
            ggml_b200_shard_rows(M, ggml_b200_comm_rank(), ggml_b200_comm_world(), &lo, &hi);
        ggml_b200_weight * w = ggml_b200_upload_q4_0(data, M, K, tensor->nb[1], lo, hi);
        if (w == NULL) {
            fprintf(stderr, "ggml-b200: upload of %s failed: %s\n", tensor->name, ggml_b200_last_error());
            exit(1);
        }
        extra->data_device[0] = w;
        extra->data_device[1] = B200_MAGIC_WEIGHT;
        tensor->extra = extra;
        return;
    }
    /* not on this back-end's path: keep it for the CPU code */
    const size_t nbytes = ggml_nbytes(tensor);
    void * copy = malloc(nbytes ? nbytes : 1);
    GGML_ASSERT(copy != NULL);
    memcpy(copy, data, nbytes);
    extra->data_device[0] = copy;
    extra->data_device[1] = B200_MAGIC_HOST;
    tensor->extra = extra;
    tensor->data = copy;
    tensor->backend = GGML_BACKEND_CPU;
}

/* ggml_cuda_free_data (ggml-cuda.cu:2811-2828) */
void ggml_cuda_free_data(struct ggml_tensor * tensor) {
    if (tensor == NULL || tensor->extra == NULL) return;
    struct ggml_tensor_extra_gpu * e = (struct ggml_tensor_extra_gpu *)tensor->extra;
    if (e->data_device[1] == B200_MAGIC_WEIGHT) ggml_b200_free_weight((ggml_b200_weight *)e->data_device[0]);
    else if (e->data_device[1] == B200_MAGIC_HOST) { free(e->data_device[0]); if (tensor->data == e->data_device[0]) tensor->data = NULL; }
    else return;   /* not ours */
    free(e);
    tensor->extra = NULL;
}

/* ggml_cuda_can_mul_mat (ggml-cuda.cu:2594-2610) as the planner uses it (ggml.c:17283-17287: n_tasks = 1).
 * The reference also grabs large CPU-resident matmuls and uploads them per call; this back-end only takes
 * weights that were uploaded at load time. */
bool ggml_cuda_can_mul_mat(const struct ggml_tensor * src0, const struct ggml_tensor * src1, struct ggml_tensor * dst) {
    /* 2-D only (every Q4_0 weight of the models is; a batched src1 falls through to the CPU code instead of aborting),
     * K a multiple of 64 (ggml.c:2343-2344) */
    return g_ready && weight_of(src0) != NULL && src0->type == GGML_TYPE_Q4_0 && src1->type == GGML_TYPE_F32 &&
           dst->type == GGML_TYPE_F32 && src1->backend == GGML_BACKEND_CPU && dst->backend == GGML_BACKEND_CPU &&
           src0->ne[2] == 1 && src0->ne[3] == 1 && src1->ne[2] == 1 && src1->ne[3] == 1 && src0->ne[0] % 64 == 0;
}

size_t ggml_cuda_mul_mat_get_wsize(const struct ggml_tensor * src0, const struct ggml_tensor * src1, struct ggml_tensor * dst) {
    (void)src0; (void)src1; (void)dst;
    return 0;   /* activations are quantized on the device: no host work buffer (ggml.c:17325-17326 is not needed) */
}

static void ensure_dev(float ** p, size_t * cap, size_t need) {
    if (need <= *cap) return;
    if (*p) ggml_b200_dev_free(*p);
    *p = (float *)ggml_b200_dev_malloc(need);
    if (*p == NULL) { fprintf(stderr, "ggml-b200: %s\n", ggml_b200_last_error()); exit(1); }
    *cap = need;
}

/* ggml_cuda_mul_mat (ggml-cuda.h:21; the COMPUTE phase of ggml_compute_forward_mul_mat_q_f32, ggml.c:11373-11411).
 * dst is complete in host memory on return, as ggml_cuda_op guarantees (ggml-cuda.cu:2514-2552). */
void ggml_cuda_mul_mat(const struct ggml_tensor * src0, const struct ggml_tensor * src1, struct ggml_tensor * dst, void * wdata, size_t wsize) {
    (void)wdata; (void)wsize;
    ggml_b200_weight * w = weight_of(src0);
    GGML_ASSERT(w != NULL);
    /* the asserts of the CPU op (ggml.c:11267-11290) that matter here */
    GGML_ASSERT(src1->nb[0] == sizeof(float) && dst->nb[0] == sizeof(float));
    GGML_ASSERT(src1->ne[2] == 1 && src1->ne[3] == 1 && dst->ne[2] == 1 && dst->ne[3] == 1);
    GGML_ASSERT(src1->ne[0] == src0->ne[0] && dst->ne[0] == src0->ne[1] && dst->ne[1] == src1->ne[1]);
    const int64_t n = src1->ne[1], K = src0->ne[0], M = src0->ne[1];
    if (ggml_b200_comm_world() > 1 && ggml_b200_weight_rows(w) != M) {
        /* row shard: replicate x on the device, sharded op + NCCL all-gather, full dst back to the host */
        ensure_dev(&g_x_dev, &g_x_cap, (size_t)n * K * sizeof(float));
        ensure_dev(&g_y_dev, &g_y_cap, (size_t)n * M * sizeof(float));
        for (int64_t c = 0; c < n; c++)
            B200_CHECK(ggml_b200_memcpy_h2d(g_x_dev + c * K, (const char *)src1->data + c * src1->nb[1], (size_t)K * sizeof(float)));
        B200_CHECK(ggml_b200_mul_mat_q4_0_sharded(w, g_x_dev, n, (size_t)K * sizeof(float), g_y_dev, (size_t)M * sizeof(float)));
        B200_CHECK(ggml_b200_synchronize());
        for (int64_t c = 0; c < n; c++)
            B200_CHECK(ggml_b200_memcpy_d2h((char *)dst->data + c * dst->nb[1], g_y_dev + c * M, (size_t)M * sizeof(float)));
        return;
    }
    B200_CHECK(ggml_b200_mul_mat_q4_0_host(w, (const float *)src1->data, n, src1->nb[1], (float *)dst->data, dst->nb[1]));
}

/* ggml_cuda_mul (ggml-cuda.cu:2580): element-wise ops are outside this back-end; never routed here */
void ggml_cuda_mul(const struct ggml_tensor * src0, const struct ggml_tensor * src1, struct ggml_tensor * dst) {
    (void)src0; (void)src1; (void)dst;
    fprintf(stderr, "ggml-b200: ggml_cuda_mul is not offloaded by this back-end\n");
    abort();
}

/* ggml_cuda_compute_forward (ggml-cuda.cu:2966-3082): the pre-hook of ggml_compute_forward (ggml.c:15645-15649).
 * Called by every worker thread for every phase; exactly one thread launches, in COMPUTE
 * (ggml-cuda.cu:3013-3020).  Returns true iff the CPU code must be skipped. */
bool ggml_cuda_compute_forward(struct ggml_compute_params * params, struct ggml_tensor * tensor) {
    if (tensor->op != GGML_OP_MUL_MAT || !ggml_cuda_can_mul_mat(tensor->src0, tensor->src1, tensor)) return false;
    if (params->ith != 0) return true;
    if (params->type == GGML_TASK_INIT || params->type == GGML_TASK_FINALIZE) return true;
    ggml_cuda_mul_mat(tensor->src0, tensor->src1, tensor, params->wdata, params->wsize);
    return true;
}

/* spelled-out aliases under this back-end's own name */
bool ggml_b200_compute_forward(struct ggml_compute_params * params, struct ggml_tensor * tensor) { return ggml_cuda_compute_forward(params, tensor); }
bool ggml_b200_can_mul_mat(const struct ggml_tensor * src0, const struct ggml_tensor * src1, struct ggml_tensor * dst) { return ggml_cuda_can_mul_mat(src0, src1, dst); }
void ggml_b200_transform_tensor(void * data, struct ggml_tensor * tensor) { ggml_cuda_transform_tensor(data, tensor); }
void ggml_b200_free_data(struct ggml_tensor * tensor) { ggml_cuda_free_data(tensor); }
int  ggml_b200_glue_ready(void) { return g_ready; }
