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
 * Scope: GGML_OP_MUL_MAT with a GGML_TYPE_Q4_0 src0 that was uploaded through
 * ggml_cuda_transform_tensor and an F32 src1, plus the element-wise nodes that sit between consecutive Q4_0 matmuls of
 * a LLaMA layer (RMS_NORM, MUL by the norm weights, ADD of the residual, SILU) when the graph builder asked for their
 * results to live on the device (ggml_cuda_assign_buffers, llama.cpp:1307-1313): activations then stay in HBM from
 * one matmul to the next, launches are asynchronous and the stream is synchronised once per CPU node instead of once
 * per offloaded node.  Every other node returns false from the hook and runs on the reference's CPU code unchanged;
 * a device-resident tensor it reads (directly or through reshape/view/permute/transpose) is copied to its host
 * buffer first.  Q4_0 matrices and F32 vectors (norm weights) tagged GGML_BACKEND_GPU are uploaded; tensors of other
 * types are kept on the host (copied, re-tagged GGML_BACKEND_CPU) so real model graphs keep working.  There is no CPU
 * fallback for a taken op: a failing device call prints and aborts like the reference's CUDA_CHECK (ggml-cuda.cu:22-30).
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
#define B200_MAGIC_QWEIGHT ((void *)(uintptr_t)0xB200B208u)  /* extra->data_device[1]: data_device[0] is a ggml_b200_qweight (Q8_0 / Q4_1) */
#define B200_MAGIC_HOST   ((void *)(uintptr_t)0xB200C0DEu)   /* extra->data_device[1]: data_device[0] is our host copy */
#define B200_MAGIC_DEVF32 ((void *)(uintptr_t)0xB200F320u)   /* ... data_device[0] = device copy of an F32 weight, [2] = our host copy (tensor->data) */
#define B200_MAGIC_ACT    ((void *)(uintptr_t)0xB200AC70u)   /* ... data_device[0] = device buffer of an activation, [2] != NULL: tensor->data holds the same values */

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
void ggml_cuda_set_tensor_split(const float * tensor_split) {
    static int warned = 0;
    if (tensor_split != NULL && !warned) {
        for (int i = 0; i < GGML_CUDA_MAX_DEVICES; i++) if (tensor_split[i] != 0.0f && !warned) {
            fprintf(stderr, "ggml-b200: tensor_split is ignored: rows are split evenly over the ranks of the communicator (ggml_b200_shard_rows)\n");
            warned = 1;
        }
    }
}

/* ---- device-resident activations (ggml-cuda.cu:2830-2935) -------------------------------------------------------
 * ggml_cuda_set_scratch_size gives the size of a VRAM ring (llama.cpp:1115-1116: n_batch MB); ggml_cuda_assign_buffers,
 * called by the graph builder on the RESULT tensor of a node, takes the next slice of the ring and tags the tensor
 * GGML_BACKEND_GPU.  Only nodes this back-end computes are taken (the others stay on the CPU and read device-resident
 * sources through ensure_host below): RMS_NORM, MUL, ADD, SILU on contiguous F32 and MUL_MAT with an uploaded Q4_0
 * weight.  The extra structs come from a ring too (the reference leaks one per call). */
static size_t g_scratch_size = 0, g_scratch_off = 0;
static char * g_scratch = NULL;
#define B200_EXTRA_POOL 16384
static struct ggml_tensor_extra_gpu g_extra_pool[B200_EXTRA_POOL];
static size_t g_extra_next = 0;
static int    g_pending = 0;        /* asynchronous copies in flight that a CPU node must wait for */

void ggml_cuda_set_scratch_size(size_t scratch_size) {
    if (scratch_size != g_scratch_size && g_scratch != NULL) { ggml_b200_synchronize(); ggml_b200_dev_free(g_scratch); g_scratch = NULL; }
    g_scratch_size = scratch_size;
    g_scratch_off = 0;
}
void ggml_b200_glue_profile_dump(void);
void ggml_cuda_free_scratch(void) {
    ggml_b200_glue_profile_dump();
    if (g_scratch == NULL) return;
    ggml_b200_synchronize();
    ggml_b200_dev_free(g_scratch);
    g_scratch = NULL;
}

static ggml_b200_weight * weight_of(const struct ggml_tensor * t);
static ggml_b200_qweight * qweight_of(const struct ggml_tensor * t);
static int is_act(const struct ggml_tensor * t) {     /* (by the extra alone: the backend tag flips to CPU once the host buffer holds the values too) */
    return t != NULL && t->extra != NULL && ((const struct ggml_tensor_extra_gpu *)t->extra)->data_device[1] == B200_MAGIC_ACT;
}
static int is_devf32(const struct ggml_tensor * t) {
    return t != NULL && t->extra != NULL && ((const struct ggml_tensor_extra_gpu *)t->extra)->data_device[1] == B200_MAGIC_DEVF32;
}
static float * dev_ptr(const struct ggml_tensor * t) { return (float *)((struct ggml_tensor_extra_gpu *)t->extra)->data_device[0]; }
static int f32_contig(const struct ggml_tensor * t) { return t != NULL && t->type == GGML_TYPE_F32 && ggml_is_contiguous(t); }

#define ROPE_TAB_BYTES ((size_t)256 << 10)
/* can this node run on the device?  (shape rules of the kernels in csrc/elementwise.cuh and of the op) */
static int node_supported(const struct ggml_tensor * t) {
    if (!g_ready || t == NULL || !f32_contig(t)) return 0;
    switch (t->op) {
        case GGML_OP_RMS_NORM:
        case GGML_OP_SILU:
            return f32_contig(t->src0);
        case GGML_OP_MUL:
        case GGML_OP_ADD:
            /* src1 = src0's shape, or one row repeated over src0's rows (norm weights: ggml.c:9199-9210) */
            return f32_contig(t->src0) && f32_contig(t->src1) && t->src0->ne[0] == t->src1->ne[0] &&
                   (ggml_nelements(t->src1) == ggml_nelements(t->src0) || ggml_nelements(t->src1) == t->src1->ne[0]);
        case GGML_OP_RESHAPE:
            /* a contiguous reshape of a device-resident tensor is the same buffer under another shape */
            return f32_contig(t->src0) && is_act(t->src0);
        case GGML_OP_ROPE: {
            /* mode 0 over whole rows (llama.cpp:1342-1347); tables of up to 256 KB (N * ne0 floats) */
            if (!f32_contig(t->src0) || t->src1 == NULL || t->src1->type != GGML_TYPE_I32 || ggml_nelements(t->src1) != 4 || t->src1->data == NULL) return 0;
            const int32_t * pr = (const int32_t *)t->src1->data;
            return pr[2] == 0 && pr[1] == t->ne[0] && (t->ne[0] & 1) == 0 && t->ne[3] == 1 && pr[0] >= 0 &&
                   (size_t)t->ne[2] * (size_t)t->ne[0] * sizeof(float) <= ROPE_TAB_BYTES;
        }
        case GGML_OP_MUL_MAT:
            return ((weight_of(t->src0) != NULL && t->src0->type == GGML_TYPE_Q4_0) || qweight_of(t->src0) != NULL) && f32_contig(t->src1) &&
                   t->src0->ne[2] == 1 && t->src0->ne[3] == 1 && t->src1->ne[2] == 1 && t->src1->ne[3] == 1 && t->src0->ne[0] % 64 == 0;
        default:
            return 0;
    }
}

static int g_stage_out_next = 0;
static int g_computed_since_assign = 0;   /* a node has run since the last ggml_cuda_assign_buffers: the next call belongs to a NEW graph */
/* fingerprint of the graph being built: what the builder asked this back-end to keep on the device, in order (op, shape,
 * where it landed).  Two graphs in a row with the same fingerprint run the same segments in the same order (see
 * speculation in flush_segment); host addresses are left out, they move with n_past. */
static uint64_t g_fp_build = 0, g_fp_running = 0, g_fp_previous = 0;
static int g_ring_wrapped = 0;            /* the scratch ring came round within this graph: two of its tensors share a buffer */
static void fp_mix(uint64_t v) { g_fp_build = (g_fp_build ^ v) * 0x100000001b3ull + 0x9e3779b97f4a7c15ull; }
static void new_graph_reset(void);
static void assign_buffers_impl(struct ggml_tensor * tensor, int scratch, int force_inplace) {
    if (!g_ready || tensor == NULL) return;
    if (g_computed_since_assign) new_graph_reset();
    if (g_computed_since_assign) {
        /* a graph builder assigns all its buffers before it computes: this is the first call of the next graph.  Start the
         * ring over, so that a graph that repeats (every decode step) gets the same device addresses -- what lets the
         * segments below be replayed as recorded CUDA graphs.  Tensors of the previous graph are dead. */
        g_computed_since_assign = 0;
        g_scratch_off = 0;
        g_stage_out_next = 0;
    }
    if (scratch && g_scratch_size == 0) return;                 /* llama.cpp low_vram: ggml_cuda_set_scratch_size(0) */
    /* like the reference (ggml-cuda.cu:2835-2841): a reshape between this node and a device-resident tensor comes along */
    if (tensor->op == GGML_OP_ROPE && tensor->src0 != NULL && tensor->src0->op == GGML_OP_RESHAPE && !is_act(tensor->src0))
        assign_buffers_impl(tensor->src0, scratch, force_inplace);
    if (!node_supported(tensor)) return;                        /* stays on the CPU */
    const size_t size = (ggml_nbytes(tensor) + 255) & ~(size_t)255;
    struct ggml_tensor_extra_gpu * extra = &g_extra_pool[g_extra_next++ % B200_EXTRA_POOL];
    memset(extra, 0, sizeof(*extra));
    const int inplace = force_inplace || tensor->op == GGML_OP_RESHAPE || (tensor->src0 != NULL && tensor->src0->data == tensor->data);
    if (inplace && is_act(tensor->src0)) {
        extra->data_device[0] = dev_ptr(tensor->src0);          /* an _inplace node writes its src0's buffer */
    } else if (scratch) {
        GGML_ASSERT(size <= g_scratch_size);
        if (g_scratch == NULL) {
            g_scratch = (char *)ggml_b200_dev_malloc(g_scratch_size);
            if (g_scratch == NULL) { fprintf(stderr, "ggml-b200: %s\n", ggml_b200_last_error()); exit(1); }
            g_scratch_off = 0;
        }
        if (g_scratch_off + size > g_scratch_size) { g_scratch_off = 0; g_ring_wrapped = 1; }     /* ring: older tensors are dead by now (as in the reference) */
        extra->data_device[0] = g_scratch + g_scratch_off;
        g_scratch_off += size;
    } else {
        void * p = ggml_b200_dev_malloc(size);                  /* (never freed by the reference either: ggml-cuda.cu:2878-2882) */
        if (p == NULL) { fprintf(stderr, "ggml-b200: %s\n", ggml_b200_last_error()); exit(1); }
        extra->data_device[0] = p;
    }
    fp_mix((uint64_t)tensor->op); fp_mix((uint64_t)tensor->ne[0]); fp_mix((uint64_t)tensor->ne[1]); fp_mix((uint64_t)tensor->ne[2] * 65537u + (uint64_t)tensor->ne[3]);
    fp_mix((uint64_t)(uintptr_t)extra->data_device[0]);
    extra->data_device[1] = B200_MAGIC_ACT;
    extra->data_device[2] = NULL;                               /* the host buffer does not hold the values */
    tensor->backend = GGML_BACKEND_GPU;
    tensor->extra = extra;
}
void ggml_cuda_assign_buffers(struct ggml_tensor * tensor) { assign_buffers_impl(tensor, 1, 0); }
void ggml_cuda_assign_buffers_no_scratch(struct ggml_tensor * tensor) { assign_buffers_impl(tensor, 0, 0); }
void ggml_cuda_assign_buffers_force_inplace(struct ggml_tensor * tensor) { assign_buffers_impl(tensor, 0, 1); }

/* pinned host memory (ggml-cuda.cu:1884-1905); NULL makes the caller fall back to malloc */
void * ggml_cuda_host_malloc(size_t size) { return g_ready ? ggml_b200_host_malloc(size) : NULL; }
void   ggml_cuda_host_free(void * ptr) { if (g_ready) ggml_b200_host_free(ptr); }

static ggml_b200_weight * weight_of(const struct ggml_tensor * t) {
    if (t == NULL || t->extra == NULL) return NULL;
    if (t->backend != GGML_BACKEND_GPU && t->backend != GGML_BACKEND_GPU_SPLIT) return NULL;
    const struct ggml_tensor_extra_gpu * e = (const struct ggml_tensor_extra_gpu *)t->extra;
    return e->data_device[1] == B200_MAGIC_WEIGHT ? (ggml_b200_weight *)e->data_device[0] : NULL;
}

/* the sibling formats: Q8_0, Q4_1, Q5_0 and Q5_1 weights, single GPU */
static ggml_b200_qweight * qweight_of(const struct ggml_tensor * t) {
    if (t == NULL || t->extra == NULL) return NULL;
    if (t->backend != GGML_BACKEND_GPU && t->backend != GGML_BACKEND_GPU_SPLIT) return NULL;
    const struct ggml_tensor_extra_gpu * e = (const struct ggml_tensor_extra_gpu *)t->extra;
    return e->data_device[1] == B200_MAGIC_QWEIGHT ? (ggml_b200_qweight *)e->data_device[0] : NULL;
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
    const int sib_2d = (tensor->type == GGML_TYPE_Q8_0 || tensor->type == GGML_TYPE_Q4_1 || tensor->type == GGML_TYPE_Q5_0 || tensor->type == GGML_TYPE_Q5_1) && tensor->ne[2] == 1 && tensor->ne[3] == 1 &&
                       (tensor->ne[0] % 64) == 0 && ggml_b200_comm_world() <= 1;
    if (g_ready && sib_2d) {
        ggml_b200_qweight * w = ggml_b200_upload_quant((int)tensor->type, data, tensor->ne[1], tensor->ne[0], tensor->nb[1]);
        if (w == NULL) {
            fprintf(stderr, "ggml-b200: upload of %s failed: %s\n", tensor->name, ggml_b200_last_error());
            exit(1);
        }
        extra->data_device[0] = w;
        extra->data_device[1] = B200_MAGIC_QWEIGHT;
        tensor->extra = extra;
        return;
    }
    const size_t nbytes = ggml_nbytes(tensor);
    void * copy = malloc(nbytes ? nbytes : 1);
    GGML_ASSERT(copy != NULL);
    memcpy(copy, data, nbytes);
    if (g_ready && tensor->type == GGML_TYPE_F32 && nbytes > 0 && nbytes <= ((size_t)1 << 24)) {
        /* norm weights (llama.cpp:1048-1070 tags them with the layer's backend): a device copy for MUL nodes that run on
         * the device, and the host copy stays readable for nodes that run on the CPU */
        void * d = ggml_b200_dev_malloc(nbytes);
        if (d == NULL) { fprintf(stderr, "ggml-b200: %s\n", ggml_b200_last_error()); exit(1); }
        B200_CHECK(ggml_b200_memcpy_h2d(d, copy, nbytes));
        extra->data_device[0] = d;
        extra->data_device[1] = B200_MAGIC_DEVF32;
        extra->data_device[2] = copy;
        tensor->extra = extra;
        tensor->data = copy;
        tensor->backend = GGML_BACKEND_CPU;      /* nodes that run on the CPU may read it (ggml.c:15650-15651 asserts CPU sources) */
        return;
    }
    /* not on this back-end's path: keep it for the CPU code */
    extra->data_device[0] = copy;
    extra->data_device[1] = B200_MAGIC_HOST;
    tensor->extra = extra;
    tensor->data = copy;
    tensor->backend = GGML_BACKEND_CPU;
}

/* ggml_cuda_free_data (ggml-cuda.cu:2811-2828) */
static void forget_recorded_segments(void);
void ggml_cuda_free_data(struct ggml_tensor * tensor) {
    if (tensor == NULL || tensor->extra == NULL) return;
    struct ggml_tensor_extra_gpu * e = (struct ggml_tensor_extra_gpu *)tensor->extra;
    /* recorded segments hold the device addresses of the weights they use: none may outlive one of them (a later
     * upload can get the same handle address back) */
    if (e->data_device[1] == B200_MAGIC_WEIGHT || e->data_device[1] == B200_MAGIC_QWEIGHT || e->data_device[1] == B200_MAGIC_DEVF32) forget_recorded_segments();
    if (e->data_device[1] == B200_MAGIC_WEIGHT) ggml_b200_free_weight((ggml_b200_weight *)e->data_device[0]);
    else if (e->data_device[1] == B200_MAGIC_QWEIGHT) ggml_b200_free_quant((ggml_b200_qweight *)e->data_device[0]);
    else if (e->data_device[1] == B200_MAGIC_HOST) { free(e->data_device[0]); if (tensor->data == e->data_device[0]) tensor->data = NULL; }
    else if (e->data_device[1] == B200_MAGIC_DEVF32) { ggml_b200_dev_free(e->data_device[0]); free(e->data_device[2]); if (tensor->data == e->data_device[2]) tensor->data = NULL; }
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
    return g_ready && ((weight_of(src0) != NULL && src0->type == GGML_TYPE_Q4_0) || qweight_of(src0) != NULL) && src1->type == GGML_TYPE_F32 &&
           dst->type == GGML_TYPE_F32 && (src1->backend == GGML_BACKEND_CPU || is_act(src1)) && (dst->backend == GGML_BACKEND_CPU || is_act(dst)) &&
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
    ggml_b200_qweight * qw = qweight_of(src0);
    GGML_ASSERT(w != NULL || qw != NULL);
    /* the asserts of the CPU op (ggml.c:11267-11290) that matter here */
    GGML_ASSERT(src1->nb[0] == sizeof(float) && dst->nb[0] == sizeof(float));
    GGML_ASSERT(src1->ne[2] == 1 && src1->ne[3] == 1 && dst->ne[2] == 1 && dst->ne[3] == 1);
    GGML_ASSERT(src1->ne[0] == src0->ne[0] && dst->ne[0] == src0->ne[1] && dst->ne[1] == src1->ne[1]);
    const int64_t n = src1->ne[1], K = src0->ne[0], M = src0->ne[1];
    if (qw != NULL) {
        /* Q8_0 / Q4_1: activations to the device, the op, dst back (no host-pointer entry point for the siblings) */
        ensure_dev(&g_x_dev, &g_x_cap, (size_t)n * K * sizeof(float));
        ensure_dev(&g_y_dev, &g_y_cap, (size_t)n * M * sizeof(float));
        for (int64_t c = 0; c < n; c++)
            B200_CHECK(ggml_b200_memcpy_h2d(g_x_dev + c * K, (const char *)src1->data + c * src1->nb[1], (size_t)K * sizeof(float)));
        B200_CHECK(ggml_b200_mul_mat_quant(qw, g_x_dev, n, (size_t)K * sizeof(float), g_y_dev, (size_t)M * sizeof(float)));
        B200_CHECK(ggml_b200_synchronize());
        for (int64_t c = 0; c < n; c++)
            B200_CHECK(ggml_b200_memcpy_d2h((char *)dst->data + c * dst->nb[1], g_y_dev + c * M, (size_t)M * sizeof(float)));
        return;
    }
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

/* ---- nodes on the device -----------------------------------------------------------------------------------------
 * A node that runs on the device is first turned into a PLAN (which kernel, which pointers, which copies) and appended
 * to the current SEGMENT: the device nodes between two CPU nodes of the graph.  A segment is submitted when a CPU node
 * needs its results, or as soon as a node with a host-resident result has been recorded.  Submitting = the plans' calls
 * in order; a segment that was seen before with the very same plans (llama.cpp's graphs and buffers repeat every
 * eval: same scratch addresses, same compute buffer) is replayed as ONE CUDA graph launch instead, so the host pays one
 * call per segment instead of one per node and copy.  Two peephole fusions at record time: RMS_NORM followed by the MUL
 * that reads it, and SILU followed by the MUL that reads it, become one kernel each (both results are still written). */
enum { PL_MUL_MAT, PL_RMS_NORM, PL_RMS_NORM_MUL, PL_MUL, PL_ADD, PL_SILU, PL_SILU_MUL, PL_MUL_MAT_ADD, PL_SWIGLU, PL_ROPE };
struct plan {
    int kind;
    ggml_b200_weight * w, * w3;      /* w3: the second weight of PL_SWIGLU */
    float * y3, * ys;                /* PL_SWIGLU: W3 x and silu(W1 x) */
    ggml_b200_qweight * qw;          /* a Q8_0 / Q4_1 weight instead of w */
    const float * a, * b;           /* device sources */
    float * y, * y2;                /* device results (y2: the second result of a fused pair) */
    int64_t n, nb, ncols, nrows, K, M;
    const void * h2d_src[2]; void * h2d_dst[2]; size_t h2d_bytes[2];     /* staged uploads of host-resident sources */
    void * d2h_dst; const void * d2h_src; size_t d2h_bytes;                /* result back to a host-resident tensor */
    int sharded;
};
#define SEG_MAX 64
#define SEG_CACHE 512
static struct plan g_seg[SEG_MAX];
static int g_seg_n = 0;
/* beside the plans (not part of the replay key): the tensors a plan computes and whether a later plan of the segment
 * reads them.  Results nothing in the segment reads (wq/wk/wv of a layer whose attention runs on the CPU) are what the
 * next CPU nodes will ask for: their copies to the host are queued right behind the segment instead of one by one, each
 * with its own wait (flush_segment). */
static struct ggml_tensor * g_seg_out[SEG_MAX][2];
static unsigned char g_seg_read[SEG_MAX][2];
static void seg_mark_read(const struct ggml_tensor * t) {
    for (int i = 0; i < g_seg_n; i++) for (int o = 0; o < 2; o++) if (g_seg_out[i][o] == t) g_seg_read[i][o] = 1;
}
static struct ggml_tensor * g_seg_last = NULL;       /* the tensor the last plan computes (for the peephole) */
static struct { struct plan * plans; int n; void * exec; int seen; int next; } g_cache[SEG_CACHE];
static int g_cache_n = 0, g_cache_cursor = 0;
static int g_use_graphs = -1;
static int g_fuse_neigh = 1;           /* GGML_B200_GLUE_FUSE=0: a matmul's ADD / SiLU*MUL neighbours stay kernels of their own */
/* tests and benches: override the two environment switches in-process (-1 leaves a switch alone).  Takes effect at the
 * next recorded node; cached segments of the other setting simply stop matching. */
void ggml_b200_glue_configure(int graphs, int fuse);
static void glue_env(void);
void ggml_b200_glue_configure(int graphs, int fuse) {
    glue_env();
    if (graphs >= 0) g_use_graphs = graphs;
    if (fuse >= 0) g_fuse_neigh = fuse;
}
static void glue_env(void) {
    if (g_use_graphs >= 0) return;
    const char * e = getenv("GGML_B200_GLUE_GRAPHS"); g_use_graphs = e ? atoi(e) : 1;
    e = getenv("GGML_B200_GLUE_FUSE"); g_fuse_neigh = e ? atoi(e) : 1;
}

static float *g_stage_in[2] = {NULL, NULL}, *g_stage_out[4] = {NULL, NULL, NULL, NULL};
static size_t g_stage_in_cap[2] = {0, 0}, g_stage_out_cap[4] = {0, 0, 0, 0};

/* Small rows cross PCIe through two pinned arenas of the glue's own instead of the tensors' host buffers:
 *   up:   the CPU copies the row into the up arena when the node is recorded (its producer has run), the segment's copy
 *         kernel reads it from there;
 *   down: the segment's copy kernel writes the row into the down arena, the CPU copies it into the tensor's buffer right
 *         after the wait (g_post).
 * The addresses a recorded segment holds are then the arenas' -- the same at every step -- although the graph's own
 * host buffers move with n_past (llama.cpp's scratch buffers hold KQ, whose size grows by a row per token, ahead of the
 * tensors this back-end reads and writes): segments keep replaying.  It also makes pageable host buffers (arch/...,
 * plain malloc) as fast as pinned ones.  Arena offsets start over with every new graph (assign_buffers_impl), so that
 * every row of one evaluation has a place of its own: a segment submitted ahead of time (see speculation below) cannot
 * overwrite a row the CPU has not collected yet. */
#define ARENA_BYTES ((size_t)32 << 20)
#define ARENA_ROW_MAX ((size_t)256 << 10)         /* larger rows (prompt batches) go straight between the tensor and HBM */
static char * g_up = NULL, * g_down = NULL;
static size_t g_up_off = 0, g_down_off = 0;
#define POST_MAX 256
static struct { void * dst; const void * src; size_t bytes; } g_post[POST_MAX];
static int g_post_n = 0;
static int g_wait_mark = -1;            /* the mark that covers everything the CPU must wait for (speculation below), -1: the whole stream */
static void wait_pending(void);
static void flush_segment(void);
static int arenas_ready(void) {
    if (g_up == NULL) {
        g_up = (char *)ggml_b200_host_malloc(ARENA_BYTES);
        g_down = (char *)ggml_b200_host_malloc(ARENA_BYTES);
        if (g_up == NULL || g_down == NULL) { g_up = g_down = NULL; return 0; }
    }
    return 1;
}

/* GGML_B200_GLUE_PROFILE=1: where the host's time goes (printed by ggml_b200_glue_profile_dump / at scratch free) */
#include <time.h>
static struct { double flush_s, wait_s, record_s; long flushes, replays, captures, direct, waits, nodes; double wait_by_n[SEG_MAX + 1]; long waits_by_n[SEG_MAX + 1]; } g_prof;
static int g_last_flush_n = 0;         /* plans in the segment submitted last (profile: waits by segment length) */
static int g_prof_on = -1;
static double g_prof_cpu_by_op[GGML_OP_COUNT], g_prof_last_t = 0.0;
static int g_prof_last_op = -1;
static long g_spec_hits = 0, g_spec_misses = 0;      /* segments submitted ahead that came / did not come (flush_segment) */
static double now_s(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec; }
void ggml_b200_glue_profile_dump(void) {
    if (g_prof_on > 0 && g_prof.nodes)
        fprintf(stderr, "ggml-b200 glue: %ld device nodes, %ld segments (%ld replayed, %ld recorded, %ld direct) %.3f ms submitting, %ld waits %.3f ms, %.3f ms recording, %ld run ahead\n",
                g_prof.nodes, g_prof.flushes, g_prof.replays, g_prof.captures, g_prof.direct, g_prof.flush_s * 1e3, g_prof.waits, g_prof.wait_s * 1e3, g_prof.record_s * 1e3, g_spec_hits);
    if (g_prof_on > 0 && g_prof.nodes) {
        fprintf(stderr, "ggml-b200 glue: waits by kernels in the segment before them (kernels: count x mean us):");
        for (int i = 0; i <= SEG_MAX; i++) if (g_prof.waits_by_n[i]) fprintf(stderr, "  %d: %ld x %.1f", i, g_prof.waits_by_n[i], g_prof.wait_by_n[i] * 1e6 / (double)g_prof.waits_by_n[i]);
        fprintf(stderr, "\n");
    }
    if (g_prof_on > 0 && g_prof.nodes) {
        fprintf(stderr, "ggml-b200 glue: host time between hook calls by the op the CPU ran (ms):");
        for (int i = 0; i < GGML_OP_COUNT; i++) if (g_prof_cpu_by_op[i] > 1e-4) fprintf(stderr, "  op%d %.2f", i, g_prof_cpu_by_op[i] * 1e3);
        fprintf(stderr, "\n");
        memset(g_prof_cpu_by_op, 0, sizeof(g_prof_cpu_by_op));
    }
    memset(&g_prof, 0, sizeof(g_prof));
}
static void wait_pending(void) {
    if (g_pending) {
        const double t0 = g_prof_on > 0 ? now_s() : 0.0;
        if (g_wait_mark >= 0) B200_CHECK(ggml_b200_wait_mark(g_wait_mark)); else B200_CHECK(ggml_b200_synchronize());
        g_wait_mark = -1;
        g_pending = 0;
        if (g_prof_on > 0) { const double dt = now_s() - t0; g_prof.wait_s += dt; g_prof.waits++; g_prof.wait_by_n[g_last_flush_n] += dt; g_prof.waits_by_n[g_last_flush_n]++; }
        for (int i = 0; i < g_post_n; i++) memcpy(g_post[i].dst, g_post[i].src, g_post[i].bytes);
        g_post_n = 0;
    }
}

static void issue_d2h(const struct plan * p) {
    if (p->d2h_bytes) B200_CHECK(ggml_b200_memcpy_d2h_async(p->d2h_dst, p->d2h_src, p->d2h_bytes));
}
/* The copies of host-resident results (wq / wk / wv of a layer whose attention runs on the CPU) can wait for the end of
 * the segment -- the CPU only looks after the flush -- so that the matmuls ahead of them run back to back (and, sharing
 * src1, side by side).  Not when a staging block would come round again first (4 in rotation), nor when a later plan of
 * the segment uploads from a host buffer such a copy fills. */
static int d2h_deferrable(void) {
    int copies = 0;
    for (int i = 0; i < g_seg_n; i++) if (g_seg[i].d2h_bytes) copies++;
    if (copies == 0 || copies > 4) return 0;
    for (int i = 0; i < g_seg_n; i++) {
        if (!g_seg[i].d2h_bytes) continue;
        const char * lo = (const char *)g_seg[i].d2h_dst, * hi = lo + g_seg[i].d2h_bytes;
        for (int k = i + 1; k < g_seg_n; k++) for (int j = 0; j < 2; j++) {
            const char * a = (const char *)g_seg[k].h2d_src[j];
            if (g_seg[k].h2d_bytes[j] && a < hi && a + g_seg[k].h2d_bytes[j] > lo) return 0;
        }
    }
    return 1;
}
static void issue(const struct plan * p) {
    for (int i = 0; i < 2; i++) if (p->h2d_bytes[i]) B200_CHECK(ggml_b200_memcpy_h2d_async(p->h2d_dst[i], p->h2d_src[i], p->h2d_bytes[i]));
    switch (p->kind) {
        case PL_MUL_MAT:
            if (p->qw)      B200_CHECK(ggml_b200_mul_mat_quant(p->qw, p->a, p->n, (size_t)p->K * sizeof(float), p->y, (size_t)p->M * sizeof(float)));
            else if (p->sharded) B200_CHECK(ggml_b200_mul_mat_q4_0_sharded(p->w, p->a, p->n, (size_t)p->K * sizeof(float), p->y, (size_t)p->M * sizeof(float)));
            else            B200_CHECK(ggml_b200_mul_mat_q4_0(p->w, p->a, p->n, (size_t)p->K * sizeof(float), p->y, (size_t)p->M * sizeof(float)));
            break;
        case PL_RMS_NORM:     B200_CHECK(ggml_b200_rms_norm_f32(p->a, p->y, p->ncols, p->nrows, 1e-6f)); break;      /* eps: ggml.c:10417 */
        case PL_RMS_NORM_MUL: B200_CHECK(ggml_b200_rms_norm_mul_f32(p->a, p->y, p->b, p->y2, p->ncols, p->nrows, p->nb, 1e-6f)); break;
        case PL_MUL:          B200_CHECK(ggml_b200_mul_f32(p->a, p->b, p->y, p->n, p->nb)); break;
        case PL_ADD:          B200_CHECK(ggml_b200_add_f32(p->a, p->b, p->y, p->n, p->nb)); break;
        case PL_SILU:         B200_CHECK(ggml_b200_silu_f32(p->a, p->y, p->n)); break;
        case PL_SILU_MUL:     B200_CHECK(ggml_b200_silu_mul_f32(p->a, p->b, p->y2, p->n, p->y)); break;
        case PL_MUL_MAT_ADD:  B200_CHECK(ggml_b200_mul_mat_q4_0_add(p->w, p->a, p->y, p->b, p->y2)); break;
        case PL_ROPE:         B200_CHECK(ggml_b200_rope_f32(p->a, p->y, p->b, p->ncols, p->nrows, p->n)); break;
        case PL_SWIGLU:       B200_CHECK(ggml_b200_mul_mat_q4_0_swiglu(p->w, p->w3, p->a, p->y, p->y3, p->ys, p->y2)); break;
    }
}
static void issue_segment(void) {
    const int defer = d2h_deferrable();
    for (int i = 0; i < g_seg_n; i++) { issue(&g_seg[i]); if (!defer) issue_d2h(&g_seg[i]); }
    if (defer) for (int i = 0; i < g_seg_n; i++) issue_d2h(&g_seg[i]);
}

/* Speculation.  The CPU nodes of a decode graph cut it into segments, and some segments need nothing those CPU nodes
 * make: wv behind "rope(wk x)", wq behind the KV-cache copies (llama.cpp:1394-1440) -- their src1 is the normed row the
 * segment BEFORE the CPU node left in HBM.  Submitted only when the hook reaches them, each costs a full host round trip
 * (12-15 us against 3 us of work, profiles/r02_trace_seam_after.log).  So when a recorded segment is submitted, the
 * segment that followed it last time is submitted right behind it if
 *   - this graph has the same fingerprint as the previous one (same device nodes in the same order: a decode loop),
 *   - it is recorded (replayable) and consists of n = 1 matmuls (and the rope behind one) only, every source in HBM, every result going either
 *     down to the host through a staging block and the down arena, or into its tensor's own device buffer while the
 *     scratch ring has not come round within this graph (no other tensor of the graph shares that buffer): it writes
 *     nothing a CPU node in between could read.
 * When the hook arrives there, the recorded plans are compared as always; equal (they are, or the fingerprint would
 * differ) means the work is done or in flight, and the CPU just collects the rows.  The wait in between is for a mark
 * behind the FIRST segment, not for the stream.  GGML_B200_GLUE_SPECULATE=0 turns it off. */
static int g_speculate = -1;
static int g_spec_ok = 0;               /* fingerprint of this graph == fingerprint of the previous one */
static int g_spec_inflight = -1;        /* cache index of the segment submitted ahead, -1: none */
static int g_last_hit = -1;             /* cache index of the segment submitted last in this graph (for the successor links) */
static int g_mark_next = 0, g_spec_mark = -1;
static void new_graph_reset(void) {
    /* (called at the first ggml_cuda_assign_buffers of a graph: the previous graph has been computed) */
    if (g_spec_inflight >= 0) { B200_CHECK(ggml_b200_synchronize()); g_spec_inflight = -1; }
    if (g_pending || g_post_n) { g_pending = 1; wait_pending(); }
    g_up_off = g_down_off = 0;
    g_fp_build = 0;
    g_ring_wrapped = 0;
    g_last_hit = -1;
}
static int speculatable(int c) {
    if (c < 0 || g_cache[c].exec == NULL) return 0;
    for (int i = 0; i < g_cache[c].n; i++) {
        const struct plan * p = &g_cache[c].plans[i];
        if (p->h2d_bytes[0] || p->h2d_bytes[1]) return 0;
        if (p->kind == PL_ROPE && p->n == 1 && !p->d2h_bytes) continue;      /* rope(wq x) in place on the matmul's own result; its table is in HBM before any segment of the eval */
        if (p->kind != PL_MUL_MAT || p->n != 1 || p->sharded) return 0;
        if (p->d2h_bytes && ((const char *)p->d2h_dst < g_down || (const char *)p->d2h_dst >= g_down + ARENA_BYTES)) return 0;
    }
    return 1;
}

static void forget_recorded_segments(void) {
    if (g_cache_n == 0) return;
    if (g_ready) B200_CHECK(ggml_b200_synchronize());
    for (int c = 0; c < g_cache_n; c++) { free(g_cache[c].plans); ggml_b200_graph_free(g_cache[c].exec); }
    g_cache_n = 0; g_cache_cursor = 0;
    g_last_hit = -1; g_spec_inflight = -1;
}
void ggml_b200_glue_set_speculate(int on);
void ggml_b200_glue_set_speculate(int on) { g_speculate = on != 0; }
long ggml_b200_glue_speculation_hits(void);
long ggml_b200_glue_speculation_hits(void) { return g_spec_hits; }
static long g_replays_total = 0;
long ggml_b200_glue_replays(void);
long ggml_b200_glue_replays(void) { return g_replays_total; }

/* submit the current segment (asynchronous; copies in it make g_pending) */
static void flush_segment(void) {
    if (g_seg_n == 0) return;
    if (g_prof_on < 0) g_prof_on = getenv("GGML_B200_GLUE_PROFILE") != NULL;
    const double t0 = g_prof_on > 0 ? now_s() : 0.0;
    if (g_prof_on > 0) { g_prof.flushes++; g_prof.nodes += g_seg_n; }
    g_last_flush_n = g_seg_n;
    glue_env();
    int hit = -1;
    /* only decode segments are replayed: an n > 1 matmul goes through the library's activation workspace, which may
     * move when a larger batch comes by (a recorded launch would keep the old address); there the launches are not
     * the bottleneck anyway */
    int graphable = g_use_graphs;
    for (int i = 0; i < g_seg_n; i++) if (g_seg[i].kind == PL_MUL_MAT && g_seg[i].n != 1) graphable = 0;
    if (graphable) {
        for (int k = 0; k < g_cache_n && hit < 0; k++) {
            const int c = (g_cache_cursor + k) % g_cache_n;
            if (g_cache[c].n == g_seg_n && memcmp(g_cache[c].plans, g_seg, sizeof(struct plan) * (size_t)g_seg_n) == 0) hit = c;
        }
    }
    if (g_spec_inflight >= 0 && g_spec_inflight != hit) {
        /* cannot happen while the fingerprint holds; the segment submitted ahead has only written staging blocks and
         * arena rows of its own, but stop guessing for good */
        fprintf(stderr, "ggml-b200 glue: a segment submitted ahead did not come (fingerprint %016llx): speculation off\n", (unsigned long long)g_fp_running);
        g_speculate = 0; g_spec_misses++;
        g_spec_inflight = -1;
        B200_CHECK(ggml_b200_synchronize());
    }
    g_wait_mark = -1;
    int launched_ahead = 0, queued_now = 0;
    if (hit >= 0 && g_cache[hit].exec != NULL) {
        if (g_spec_inflight == hit) { launched_ahead = 1; g_spec_inflight = -1; g_spec_hits++; }
        else B200_CHECK(ggml_b200_graph_launch(g_cache[hit].exec));
        g_cache_cursor = (hit + 1) % g_cache_n;
        g_prof.replays++; g_replays_total++;
    } else if (hit >= 0) {
        g_prof.captures++;
        /* second time this very segment comes by: record it while submitting it */
        void * exec = NULL;
        B200_CHECK(ggml_b200_graph_begin());
        issue_segment();
        B200_CHECK(ggml_b200_graph_end(&exec));
        g_cache[hit].exec = exec;
        B200_CHECK(ggml_b200_graph_launch(exec));
        g_cache_cursor = (hit + 1) % g_cache_n;
    } else {
        g_prof.direct++;
        issue_segment();
        if (graphable) {
            if (g_cache_n == SEG_CACHE) forget_recorded_segments();       /* the graphs changed (another model / batch size): start over */
            g_cache[g_cache_n].plans = (struct plan *)malloc(sizeof(struct plan) * (size_t)g_seg_n);
            memcpy(g_cache[g_cache_n].plans, g_seg, sizeof(struct plan) * (size_t)g_seg_n);
            g_cache[g_cache_n].n = g_seg_n;
            g_cache[g_cache_n].exec = NULL;
            g_cache[g_cache_n].next = -1;
            g_cache_n++;
        }
    }
    for (int i = 0; i < g_seg_n; i++) if (g_seg[i].d2h_bytes || g_seg[i].h2d_bytes[0] || g_seg[i].h2d_bytes[1]) g_pending = 1;
    for (int i = 0; i < g_seg_n; i++) for (int o = 0; o < 2; o++) {
        struct ggml_tensor * t = g_seg_out[i][o];
        if (t == NULL || g_seg_read[i][o] || !is_act(t) || t->data == NULL) continue;
        struct ggml_tensor_extra_gpu * e = (struct ggml_tensor_extra_gpu *)t->extra;
        if (e->data_device[2] != NULL) continue;
        B200_CHECK(ggml_b200_memcpy_d2h_async(t->data, e->data_device[0], ggml_nbytes(t)));     /* in graph order: t's node is behind us */
        e->data_device[2] = t->data;
        g_pending = 1;
        queued_now = 1;
    }
    /* successor links, the mark the CPU will wait for, and the next segment ahead of time */
    const int replayed = hit >= 0 && g_cache[hit].exec != NULL;
    if (g_last_hit >= 0) g_cache[g_last_hit].next = replayed ? hit : -1;
    g_last_hit = replayed ? hit : -1;
    if (g_speculate < 0) { const char * e = getenv("GGML_B200_GLUE_SPECULATE"); g_speculate = e ? atoi(e) : 1; }
    int have_mark = 0;
    if (launched_ahead && !queued_now) { g_wait_mark = g_spec_mark; have_mark = 1; }     /* nothing new behind it */
    if (replayed && g_speculate > 0 && g_spec_ok) {
        const int nx = g_cache[hit].next;
        if (speculatable(nx)) {
            if (!have_mark) {
                const int m = g_mark_next++ % GGML_B200_MARKS;
                B200_CHECK(ggml_b200_mark(m));                /* behind this segment and its copies ... */
                g_wait_mark = m;
            }
            B200_CHECK(ggml_b200_graph_launch(g_cache[nx].exec));   /* ... and ahead of the CPU node in front of the next one */
            g_spec_mark = g_mark_next++ % GGML_B200_MARKS;
            B200_CHECK(ggml_b200_mark(g_spec_mark));
            g_spec_inflight = nx;
        }
    }
    g_seg_n = 0;
    g_seg_last = NULL;
    if (g_prof_on > 0) g_prof.flush_s += now_s() - t0;
}

/* growing a staging block: everything queued may still use the old one */
static void stage_grow(float ** p, size_t * cap, size_t need) {
    if (need <= *cap) return;
    flush_segment();
    g_pending = 1;
    wait_pending();
    ensure_dev(p, cap, (need + 4095) & ~(size_t)4095);
}
/* the values of `t` on the device: its own buffer, the device copy of a norm weight, or a staged upload of its host data */
static const float * plan_src(struct plan * p, const struct ggml_tensor * t, int slot) {
    if (is_act(t) || is_devf32(t)) return dev_ptr(t);
    const size_t nbytes = ggml_nbytes(t);
    stage_grow(&g_stage_in[slot], &g_stage_in_cap[slot], nbytes);
    const void * src = t->data;
    /* (a row still on its way down into t->data -- a host-resident result a device node reads -- must land first) */
    for (int i = 0; i < g_post_n; i++) if ((const char *)g_post[i].dst < (const char *)t->data + nbytes && (const char *)t->data < (const char *)g_post[i].dst + g_post[i].bytes) {
        flush_segment(); g_pending = 1; wait_pending(); break;
    }
    if (nbytes <= ARENA_ROW_MAX && arenas_ready()) {
        const size_t need = (nbytes + 255) & ~(size_t)255;
        if (g_up_off + need > ARENA_BYTES) { flush_segment(); g_pending = 1; wait_pending(); g_up_off = 0; }
        memcpy(g_up + g_up_off, t->data, nbytes);
        src = g_up + g_up_off;
        g_up_off += need;
    }
    p->h2d_src[slot] = src; p->h2d_dst[slot] = g_stage_in[slot]; p->h2d_bytes[slot] = nbytes;
    return g_stage_in[slot];
}
static float * plan_dst(struct plan * p, struct ggml_tensor * t) {
    if (is_act(t)) {
        struct ggml_tensor_extra_gpu * e = (struct ggml_tensor_extra_gpu *)t->extra;
        e->data_device[2] = NULL;                                    /* rewritten: the host buffer is stale */
        return (float *)e->data_device[0];
    }
    /* result of a host-resident tensor: a staging block (a few in rotation, so that wq/wk/wv do not wait for each
     * other's copy), then back to the tensor's host buffer */
    const int k = g_stage_out_next++ % 4;
    const size_t nbytes = ggml_nbytes(t);
    stage_grow(&g_stage_out[k], &g_stage_out_cap[k], nbytes);
    void * dst = t->data;
    if (nbytes <= ARENA_ROW_MAX && arenas_ready()) {
        const size_t need = (nbytes + 255) & ~(size_t)255;
        if (g_down_off + need > ARENA_BYTES || g_post_n == POST_MAX) { flush_segment(); g_pending = 1; wait_pending(); if (g_down_off + need > ARENA_BYTES) g_down_off = 0; }
        dst = g_down + g_down_off;
        g_down_off += need;
        g_post[g_post_n].dst = t->data; g_post[g_post_n].src = dst; g_post[g_post_n].bytes = nbytes; g_post_n++;
    }
    p->d2h_dst = dst; p->d2h_src = g_stage_out[k]; p->d2h_bytes = nbytes;
    return g_stage_out[k];
}

/* a node that runs on the CPU reads `t`: if its values live on the device only (t itself, or the tensor a chain of
 * reshape / view / permute / transpose nodes looks into), bring them to the host buffer and tag the tensor CPU */
static void ensure_host(struct ggml_tensor * t) {
    for (int depth = 0; t != NULL && depth < 8; depth++) {
        if (is_act(t)) {
            struct ggml_tensor_extra_gpu * e = (struct ggml_tensor_extra_gpu *)t->extra;
            if (e->data_device[2] == NULL) {
                flush_segment();
                B200_CHECK(ggml_b200_memcpy_d2h_async(t->data, e->data_device[0], ggml_nbytes(t)));
                g_pending = 1;
                g_wait_mark = -1;                                /* queued behind whatever went ahead: wait for the stream */
                e->data_device[2] = t->data;
            }
            t->backend = GGML_BACKEND_CPU;
            return;
        }
        if (t->op != GGML_OP_RESHAPE && t->op != GGML_OP_VIEW && t->op != GGML_OP_PERMUTE && t->op != GGML_OP_TRANSPOSE) return;
        t = t->src0;
    }
}

/* The angles of a rope node, computed as the reference computes them (ggml.c:12771-12813: theta = p, then theta *= theta_scale
 * per pair, cosf / sinf of the host's libm) and kept in ONE device block whose address the recorded segments hold: it is
 * refilled when the position changes (once per token: q and k of every layer share it), by a copy queued on the stream
 * ahead of the segments that read it. */
#include <math.h>
static float * g_rope_dev = NULL;
static float * g_rope_pin[2] = {NULL, NULL};
static int g_rope_flip = 0;
static struct { int n_past, n_dims, ne0, npos, valid; } g_rope_key;
static const float * rope_table(int n_past, int n_dims, int ne0, int npos) {
    if (g_rope_dev == NULL) {
        g_rope_dev = (float *)ggml_b200_dev_malloc(ROPE_TAB_BYTES);
        g_rope_pin[0] = (float *)ggml_b200_host_malloc(ROPE_TAB_BYTES);
        g_rope_pin[1] = (float *)ggml_b200_host_malloc(ROPE_TAB_BYTES);
        if (g_rope_dev == NULL || g_rope_pin[0] == NULL || g_rope_pin[1] == NULL) { fprintf(stderr, "ggml-b200: %s\n", ggml_b200_last_error()); exit(1); }
    }
    if (g_rope_key.valid && g_rope_key.n_past == n_past && g_rope_key.n_dims == n_dims && g_rope_key.ne0 == ne0 && g_rope_key.npos == npos) return g_rope_dev;
    float * tab = g_rope_pin[g_rope_flip ^= 1];
    const float theta_scale = powf(10000.0, -2.0f / n_dims);
    for (int i2 = 0; i2 < npos; i2++) {
        float theta = (float)(n_past + i2);
        for (int k = 0; k < ne0 / 2; k++) {
            tab[((size_t)i2 * (ne0 / 2) + k) * 2] = cosf(theta);
            tab[((size_t)i2 * (ne0 / 2) + k) * 2 + 1] = sinf(theta);
            theta *= theta_scale;
        }
    }
    B200_CHECK(ggml_b200_memcpy_h2d_async(g_rope_dev, tab, (size_t)npos * (size_t)ne0 * sizeof(float)));
    g_rope_key.n_past = n_past; g_rope_key.n_dims = n_dims; g_rope_key.ne0 = ne0; g_rope_key.npos = npos; g_rope_key.valid = 1;
    return g_rope_dev;
}

static void device_node(struct ggml_tensor * t) {
    glue_env();
    if (g_seg_n == SEG_MAX) flush_segment();
    struct plan * p = &g_seg[g_seg_n];
    memset(p, 0, sizeof(*p));
    p->n = ggml_nelements(t);
    const int fuse = g_use_graphs != 0 && g_seg_n > 0 && g_seg_last != NULL && t->op == GGML_OP_MUL && t->src0 == g_seg_last && is_act(t) &&
                     !g_seg[g_seg_n - 1].d2h_bytes && (is_act(t->src1) || is_devf32(t->src1));
    if (fuse && g_seg[g_seg_n - 1].kind == PL_RMS_NORM) {
        struct plan * q = &g_seg[g_seg_n - 1];                   /* y = rms_norm(a); y2 = y * w: one kernel */
        q->kind = PL_RMS_NORM_MUL; q->b = dev_ptr(t->src1); q->nb = ggml_nelements(t->src1); q->y2 = plan_dst(q, t);
        seg_mark_read(t->src0); seg_mark_read(t->src1);
        g_seg_out[g_seg_n - 1][1] = t; g_seg_read[g_seg_n - 1][1] = 0;
        g_seg_last = t;
        return;
    }
    if (fuse && g_seg[g_seg_n - 1].kind == PL_SILU && ggml_nelements(t->src1) == p->n) {
        struct plan * q = &g_seg[g_seg_n - 1];                   /* y = silu(a); y2 = y * b */
        q->kind = PL_SILU_MUL; q->b = dev_ptr(t->src1); q->y2 = plan_dst(q, t);
        seg_mark_read(t->src0); seg_mark_read(t->src1);
        g_seg_out[g_seg_n - 1][1] = t; g_seg_read[g_seg_n - 1][1] = 0;
        g_seg_last = t;
        /* ... and when a and b are the results of the two decode matmuls right before it, over the same x (the LLaMA FFN
         * gate, llama.cpp:1479-1497): all three become ONE kernel that streams both weights */
        if (g_fuse_neigh && g_seg_n >= 3) {
            struct plan * m1 = &g_seg[g_seg_n - 3], * m2 = &g_seg[g_seg_n - 2];
            const int pair = m1->kind == PL_MUL_MAT && m2->kind == PL_MUL_MAT && m1->w && m2->w && m1->n == 1 && m2->n == 1 &&
                             !m1->sharded && !m2->sharded && m1->a == m2->a && m1->K == m2->K && m1->M == m2->M &&
                             !m1->d2h_bytes && !m2->d2h_bytes && !m1->h2d_bytes[0] && !m2->h2d_bytes[0] && !q->h2d_bytes[0] && !q->h2d_bytes[1] && !q->d2h_bytes &&
                             ggml_b200_neigh_ok(m1->w) && ggml_b200_neigh_ok(m2->w);
            struct plan * gate = NULL, * up = NULL;
            if (pair && m1->y == q->a && m2->y == q->b) { gate = m1; up = m2; }
            if (pair && m2->y == q->a && m1->y == q->b) { gate = m2; up = m1; }
            if (gate != NULL) {
                struct plan f;
                memset(&f, 0, sizeof(f));
                f.kind = PL_SWIGLU; f.w = gate->w; f.w3 = up->w; f.a = gate->a; f.y = gate->y; f.y3 = up->y; f.ys = q->y; f.y2 = q->y2;
                f.n = 1; f.K = gate->K; f.M = gate->M;
                *m1 = f;
                g_seg_out[g_seg_n - 3][0] = t; g_seg_out[g_seg_n - 3][1] = NULL; g_seg_read[g_seg_n - 3][0] = 0; g_seg_read[g_seg_n - 3][1] = 0;
                g_seg_n -= 2;
            }
        }
        return;
    }
    /* the same gate in the order ggml_build_forward_expand gives llama.cpp's graph (src0 first): w1, silu, w3, mul */
    if (g_fuse_neigh && g_use_graphs != 0 && t->op == GGML_OP_MUL && is_act(t) && g_seg_n >= 3 && is_act(t->src0) && is_act(t->src1) &&
        ggml_nelements(t->src0) == p->n && ggml_nelements(t->src1) == p->n) {
        struct plan * gate = &g_seg[g_seg_n - 3], * sl = &g_seg[g_seg_n - 2], * up = &g_seg[g_seg_n - 1];
        const float * s0 = dev_ptr(t->src0), * s1 = dev_ptr(t->src1);
        if (gate->kind == PL_MUL_MAT && sl->kind == PL_SILU && up->kind == PL_MUL_MAT && gate->w && up->w && gate->n == 1 && up->n == 1 &&
            !gate->sharded && !up->sharded && sl->a == gate->y && gate->a == up->a && gate->K == up->K && gate->M == up->M && gate->M == p->n &&
            ((s0 == sl->y && s1 == up->y) || (s1 == sl->y && s0 == up->y)) &&
            !gate->d2h_bytes && !up->d2h_bytes && !sl->d2h_bytes && !gate->h2d_bytes[0] && !up->h2d_bytes[0] && !sl->h2d_bytes[0] &&
            ggml_b200_neigh_ok(gate->w) && ggml_b200_neigh_ok(up->w)) {
            struct plan f;
            memset(&f, 0, sizeof(f));
            f.kind = PL_SWIGLU; f.w = gate->w; f.w3 = up->w; f.a = gate->a; f.y = gate->y; f.y3 = up->y; f.ys = sl->y; f.y2 = plan_dst(&f, t);
            f.n = 1; f.K = gate->K; f.M = gate->M;
            *gate = f;
            seg_mark_read(t->src0); seg_mark_read(t->src1);
            g_seg_out[g_seg_n - 3][0] = t; g_seg_out[g_seg_n - 3][1] = NULL; g_seg_read[g_seg_n - 3][0] = 0; g_seg_read[g_seg_n - 3][1] = 0;
            g_seg_n -= 2;
            g_seg_last = t;
            return;
        }
    }
    /* a decode matmul and the ADD that reads its result (residual behind wo / w2, llama.cpp:1462, :1504): one kernel */
    if (g_fuse_neigh && g_use_graphs != 0 && g_seg_n > 0 && g_seg_last != NULL && t->op == GGML_OP_ADD && is_act(t) &&
        (t->src0 == g_seg_last || t->src1 == g_seg_last) && t->src0 != t->src1) {
        struct plan * q = &g_seg[g_seg_n - 1];
        const struct ggml_tensor * other = t->src0 == g_seg_last ? t->src1 : t->src0;
        if (q->kind == PL_MUL_MAT && q->w != NULL && q->n == 1 && !q->sharded && !q->d2h_bytes && (is_act(other) || is_devf32(other)) &&
            ggml_nelements(other) == q->M && ggml_nelements(t) == q->M && ggml_b200_neigh_ok(q->w)) {
            q->kind = PL_MUL_MAT_ADD; q->b = dev_ptr(other); q->y2 = plan_dst(q, t);
            seg_mark_read(t->src0); seg_mark_read(t->src1);
            g_seg_out[g_seg_n - 1][1] = t; g_seg_read[g_seg_n - 1][1] = 0;
            g_seg_last = t;
            return;
        }
    }
    switch (t->op) {
        case GGML_OP_MUL_MAT: {
            p->kind = PL_MUL_MAT;
            p->w = weight_of(t->src0);
            p->qw = qweight_of(t->src0);
            p->K = t->src0->ne[0]; p->M = t->src0->ne[1]; p->n = t->src1->ne[1];
            GGML_ASSERT(t->src1->ne[0] == p->K && t->ne[0] == p->M && t->ne[1] == p->n);
            p->a = plan_src(p, t->src1, 0);
            p->y = plan_dst(p, t);
            p->sharded = p->w != NULL && ggml_b200_comm_world() > 1 && ggml_b200_weight_rows(p->w) != p->M;
        } break;
        case GGML_OP_RMS_NORM:
            p->kind = PL_RMS_NORM; p->ncols = t->src0->ne[0]; p->nrows = ggml_nrows(t->src0);
            p->a = plan_src(p, t->src0, 0); p->y = plan_dst(p, t);
            break;
        case GGML_OP_MUL:
        case GGML_OP_ADD:
            p->kind = t->op == GGML_OP_MUL ? PL_MUL : PL_ADD; p->nb = ggml_nelements(t->src1);
            p->a = plan_src(p, t->src0, 0); p->b = plan_src(p, t->src1, 1); p->y = plan_dst(p, t);
            break;
        case GGML_OP_SILU:
            p->kind = PL_SILU; p->a = plan_src(p, t->src0, 0); p->y = plan_dst(p, t);
            break;
        case GGML_OP_RESHAPE:
            ((struct ggml_tensor_extra_gpu *)t->extra)->data_device[2] = NULL;      /* same buffer as src0: nothing to run */
            return;
        case GGML_OP_ROPE: {
            const int32_t * pr = (const int32_t *)t->src1->data;
            p->kind = PL_ROPE; p->ncols = t->ne[0]; p->nrows = t->ne[1]; p->n = t->ne[2];
            p->b = rope_table(pr[0], pr[1], (int)t->ne[0], (int)t->ne[2]);
            p->a = plan_src(p, t->src0, 0); p->y = plan_dst(p, t);
        } break;
        default:
            GGML_ASSERT(!"device_node: unsupported op");
    }
    /* (plan_src / plan_dst may have flushed the segment to grow a staging block: append at the current end) */
    if (p != &g_seg[g_seg_n]) g_seg[g_seg_n] = *p;
    seg_mark_read(t->src0); seg_mark_read(t->src1);
    g_seg_out[g_seg_n][0] = t; g_seg_out[g_seg_n][1] = NULL; g_seg_read[g_seg_n][0] = 0; g_seg_read[g_seg_n][1] = 0;
    g_seg_n++;
    g_seg_last = t;
    /* a host-resident result is copied back inside the segment (stream order); the segment is submitted when a CPU node
     * needs something (the hook below).  Submitting here, per result, cost two extra graph launches and waits per layer. */
}

/* ggml_cuda_compute_forward (ggml-cuda.cu:2966-3082): the pre-hook of ggml_compute_forward (ggml.c:15645-15649).
 * Called by every worker thread for every phase; the INIT phase always runs first and on one thread
 * (ggml.c:17114-17117), exactly one thread launches, in COMPUTE (ggml-cuda.cu:3013-3020).  Returns true iff the CPU
 * code must be skipped.  A node runs on the device when this back-end has a kernel for it AND its result or one of its
 * sources lives there; a Q4_0 mul_mat with everything on the host takes the one-launch host path.
 * A graph must not END on a device node whose result lives on the host (nothing would wait for its copy); llama.cpp's
 * graphs end on the output matmul, whose sources and result are on the host: the synchronous host path. */
bool ggml_cuda_compute_forward(struct ggml_compute_params * params, struct ggml_tensor * tensor) {
    if (!g_ready) return false;
    if (g_prof_on > 0 && params->ith == 0) {                  /* profile: wall time between hook calls, by the op the CPU ran */
        const double t = now_s();
        if (g_prof_last_op >= 0 && g_prof_last_op < GGML_OP_COUNT) g_prof_cpu_by_op[g_prof_last_op] += t - g_prof_last_t;
        g_prof_last_t = t;
        g_prof_last_op = -1;
    }
    if (!g_computed_since_assign && params->ith == 0) {       /* first node of a graph: its fingerprint is complete */
        g_fp_previous = g_fp_running; g_fp_running = g_fp_build;
        g_spec_ok = g_fp_running != 0 && g_fp_running == g_fp_previous && !g_ring_wrapped;
    }
    g_computed_since_assign = 1;
    const int on_device = is_act(tensor) || is_act(tensor->src0) || is_act(tensor->src1);
    const int host_mul_mat = tensor->op == GGML_OP_MUL_MAT && !on_device && ggml_cuda_can_mul_mat(tensor->src0, tensor->src1, tensor);
    /* (a reshape is a device node only where assign_buffers made it an alias of its source's buffer) */
    if ((on_device && node_supported(tensor) && (tensor->op != GGML_OP_RESHAPE || is_act(tensor))) || host_mul_mat) {
        if (params->ith != 0) return true;
        if (params->type == GGML_TASK_INIT || params->type == GGML_TASK_FINALIZE) return true;
        if (host_mul_mat) { flush_segment(); wait_pending(); ggml_cuda_mul_mat(tensor->src0, tensor->src1, tensor, params->wdata, params->wsize); }
        else if (g_prof_on > 0) { const double t0 = now_s(); device_node(tensor); g_prof.record_s += now_s() - t0; }
        else device_node(tensor);
        return true;
    }
    /* the reference's CPU code runs this node: before any thread starts on it (INIT), submit what is recorded, bring
     * device-resident sources home and wait for the copies in flight */
    if (g_prof_on > 0 && params->ith == 0) g_prof_last_op = (int)tensor->op;
    if (params->ith == 0 && params->type == GGML_TASK_INIT) {
        flush_segment();
        ensure_host(tensor->src0);
        ensure_host(tensor->src1);
        for (int i = 0; i < GGML_MAX_OPT; i++) ensure_host(tensor->opt[i]);
        wait_pending();
    }
    return false;
}

/* spelled-out aliases under this back-end's own name */
bool ggml_b200_compute_forward(struct ggml_compute_params * params, struct ggml_tensor * tensor) { return ggml_cuda_compute_forward(params, tensor); }
bool ggml_b200_can_mul_mat(const struct ggml_tensor * src0, const struct ggml_tensor * src1, struct ggml_tensor * dst) { return ggml_cuda_can_mul_mat(src0, src1, dst); }
void ggml_b200_transform_tensor(void * data, struct ggml_tensor * tensor) { ggml_cuda_transform_tensor(data, tensor); }
void ggml_b200_free_data(struct ggml_tensor * tensor) { ggml_cuda_free_data(tensor); }
int  ggml_b200_glue_ready(void) { return g_ready; }
