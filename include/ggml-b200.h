/*
 * ggml-b200.h -- C ABI of the B200-native (sm_100a) Q4_0 x F32 mul_mat.
 *
 * Drop-in boundary for ONE path of ggml: GGML_OP_MUL_MAT with a GGML_TYPE_Q4_0 src0 and an F32
 * src1 (ggml_compute_forward_mul_mat -> ggml_compute_forward_mul_mat_q_f32, reference
 * ggml.c:11426 / ggml.c:11226).  Plain C linkage, plain pointers and sizes, no C++ or torch
 * types.  One process drives one GPU (one rank per GPU); a multi-GPU job is N processes that
 * each upload their row shard and meet in ggml_b200_mul_mat_q4_0_sharded (NCCL all-gather over NVLink).
 *
 * Every entry point cites the reference interface it replaces (paths under the reference
 * tree).  The ggml_tensor-level glue that binds these to the reference's offload seam
 * (ggml_cuda_compute_forward etc.) is llama.cpp-q_4_0_b200/ggml_glue/ggml-b200-glue.c; see
 * INTEGRATION.md.
 *
 * Error convention: the reference aborts (GGML_ASSERT, ggml.h:204-210; CUDA_CHECK -> exit(1),
 * ggml-cuda.cu:22-30).  Here every call returns an int status (0 = ok, negative = error, text
 * via ggml_b200_last_error()) so a harness can skip on a GPU-less host; the glue turns a
 * non-zero status into the reference's print + abort.  There is NO CPU fallback anywhere.
 */
#ifndef GGML_B200_H
#define GGML_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GGML_B200_OK              0
#define GGML_B200_ERR_NO_DEVICE  (-1)   /* no CUDA device / driver */
#define GGML_B200_ERR_ARCH       (-2)   /* device is not sm_100 (the only code in the fatbin) */
#define GGML_B200_ERR_CUDA       (-3)   /* a CUDA / NCCL call failed, see ggml_b200_last_error */
#define GGML_B200_ERR_ARG        (-4)   /* shape / alignment / state violates the op contract */
#define GGML_B200_ERR_STATE      (-5)   /* ggml_b200_init not called, or comm not initialised */

/* weight layouts in HBM */
#define GGML_B200_LAYOUT_RAW     0      /* the reference's 18-byte block stream, as on disk */
#define GGML_B200_LAYOUT_PLANES  1      /* split planes: fp16 d[M][K/32] + 16-byte nibbles[M][K/32] */

/* kernel selection for ggml_b200_set_path (0 = choose by n, the default) */
#define GGML_B200_PATH_AUTO      0
#define GGML_B200_PATH_GEMV      1      /* dp4a + warp-shuffle kernel (any n, tiled over columns) */
#define GGML_B200_PATH_GEMM      2      /* TMA + tcgen05 kind::i8 + TMEM kernel */

typedef struct ggml_b200_weight ggml_b200_weight;   /* opaque: one device-resident Q4_0 row shard */

/* ---- life cycle ---------------------------------------------------------------------------
 * replaces ggml_init_cublas (ggml-cuda.cu:1826-1861, called from ggml_init ggml.c:4282-4286)
 * and ggml_cuda_set_main_device (ggml-cuda.h:33).  Selects `device`, checks it is sm_100,
 * creates the stream and the workspace.  Idempotent for the same device. */
int          ggml_b200_init(int device);
void         ggml_b200_shutdown(void);
int          ggml_b200_device_count(void);                 /* <= 0 when no usable GPU */
const char * ggml_b200_last_error(void);
int          ggml_b200_sm_count(void);

/* run-time knobs (the reference's equivalents are build-time GGML_CUDA_DMMV_* macros,
 * ggml-cuda.cu:203-215).  Also read once from the environment at init:
 * GGML_B200_REPACK=0|1, GGML_B200_GEMM_MIN_N=<n>, GGML_B200_PATH=auto|gemv|gemm */
void         ggml_b200_set_repack(int on);                 /* layout of FUTURE uploads; default 1 */
int          ggml_b200_get_repack(void);
void         ggml_b200_set_gemm_min_n(int n);              /* n >= this takes the tensor-core path */
int          ggml_b200_get_gemm_min_n(void);
void         ggml_b200_set_path(int path);
/* kernel tuning knobs for sweeps ("gemv_threads", "gemv_ctas_per_sm", "gemv_unroll", "gemv_nc",
 * "gemm_*", "set_*"); returns GGML_B200_ERR_ARG for an unknown key or a value out of range */
int          ggml_b200_set_tuning(const char * key, int value);
void         ggml_b200_set_stream(void * cuda_stream);     /* NULL = the library's own stream */
void *       ggml_b200_get_stream(void);

/* pinned host memory: mirrors ggml_cuda_host_malloc / ggml_cuda_host_free (ggml-cuda.cu:1884-1905) */
void *       ggml_b200_host_malloc(size_t size);
void         ggml_b200_host_free(void * ptr);

/* ---- load-time upload ---------------------------------------------------------------------
 * replaces ggml_cuda_transform_tensor (ggml-cuda.cu:2766-2809; caller llama.cpp:678-685).
 * `blocks` = host (or device) pointer to the tensor data: M rows of K/32 block_q4_0 (ggml.c:870-875),
 * row stride nb1 bytes (== 18*K/32 for every tensor ggml creates).  Rows [row_lo, row_hi) are
 * COPIED to the device (the caller may free `blocks` afterwards, llama.cpp:682-684) and, when
 * repack is on, re-laid once into the plane layout.  Returns NULL on error. */
ggml_b200_weight * ggml_b200_upload_q4_0(const void * blocks, int64_t M, int64_t K, size_t nb1,
                                         int64_t row_lo, int64_t row_hi);
/* replaces ggml_cuda_free_data (ggml-cuda.cu:2811-2828; caller llama.cpp:241-245) */
void         ggml_b200_free_weight(ggml_b200_weight * w);
int64_t      ggml_b200_weight_rows(const ggml_b200_weight * w);   /* row_hi - row_lo */
int64_t      ggml_b200_weight_row_lo(const ggml_b200_weight * w);
int64_t      ggml_b200_weight_M(const ggml_b200_weight * w);      /* rows of the whole tensor */
int64_t      ggml_b200_weight_K(const ggml_b200_weight * w);
int          ggml_b200_weight_layout(const ggml_b200_weight * w);
size_t       ggml_b200_weight_device_bytes(const ggml_b200_weight * w);
/* inverse of the repack: writes this shard's rows back as 18-byte blocks (host pointer,
 * rows*K/32*18 bytes).  Used by the parity tests to prove the repack is lossless. */
int          ggml_b200_download_q4_0(const ggml_b200_weight * w, void * blocks_out);

/* ---- weight-side Q4_0 codec on the device ---------------------------------------
 * replaces quantize_row_q4_0 / quantize_row_q4_0_reference (ggml.c:918-957; bulk form ggml_quantize_q4_0
 * :19157-19178, threaded caller llama.cpp:2452-2505): x_dev = nrows rows of K floats, row stride nb1 bytes;
 * blocks_dev receives nrows*K/32 block_q4_0 (18 bytes each, rows dense).  Bit-identical to the reference, ties
 * and sign of d included.  The blocks may be handed to ggml_b200_upload_q4_0, which accepts device pointers too. */
int          ggml_b200_quantize_q4_0(const float * x_dev, int64_t nrows, int64_t K, size_t nb1, void * blocks_dev);
/* replaces dequantize_row_q4_0 (ggml.c:1500-1518): nblocks 18-byte blocks -> 32*nblocks floats */
int          ggml_b200_dequantize_q4_0(const void * blocks_dev, int64_t nblocks, float * y_dev);
/* replaces ggml_compute_forward_get_rows_q for a Q4_0 src0 (ggml.c:11868-11890; token embeddings,
 * llama.cpp:1358): y_dev row i (row stride nb1 bytes, >= 4*K) = dequantized row ids_dev[i] of the tensor.
 * ids are int32 rows of the WHOLE tensor; GGML_B200_ERR_ARG if one is outside this shard. */
int          ggml_b200_get_rows_q4_0(const ggml_b200_weight * w, const int32_t * ids_dev, int64_t n,
                                     float * y_dev, size_t nb1);

/* ---- kernel 1: activation quantizer -------------------------------------------------------
 * replaces quantize_row_q8_0 (ggml.c:1122-1280, AVX2 branch :1192-1250 is the contract) as the
 * op's INIT phase calls it (ggml.c:11353-11367).  x_dev: n rows of K floats, row stride nb11
 * bytes.  blocks34_dev: n*K/32 reference-format block_q8_0 (34 bytes each, ggml.c:902-907),
 * byte-identical to the reference's work buffer.  Device pointers. */
int          ggml_b200_quantize_q8_0(const float * x_dev, int64_t n, int64_t K, size_t nb11,
                                     void * blocks34_dev);

/* ---- the op -------------------------------------------------------------------------------
 * replaces ggml_cuda_mul_mat (ggml-cuda.cu:2671-2690) / the COMPUTE phase of
 * ggml_compute_forward_mul_mat_q_f32 (ggml.c:11373-11411) for this shard's rows:
 *   y[c*nb1/4 + row_lo + r] = sum_b fp32(dw[r,b]) * fp32(dx[c,b]) * S_b(r,c)      0 <= c < n
 * x: n rows of K floats (stride nb11 bytes, nb10 == 4 as ggml.c:11279 requires); y: n rows of
 * M floats (stride nb1 bytes, nb0 == 4, ggml.c:11282) -- the FULL dst; a shard writes only its
 * own column range [row_lo, row_hi).  Quantizes x on the device (kernel 1), then runs the GEMV
 * (kernel 2) or the tensor-core GEMM (kernel 3).
 *   _dev : x and y are device pointers; asynchronous on the current stream.
 *   _host: x and y are host pointers (what a ggml graph with CPU-resident activations hands
 *          over); H2D + kernels + D2H + stream sync, complete on return like
 *          ggml_cuda_op (ggml-cuda.cu:2514-2552). */
int          ggml_b200_mul_mat_q4_0(const ggml_b200_weight * w, const float * x_dev, int64_t n,
                                    size_t nb11, float * y_dev, size_t nb1);
int          ggml_b200_mul_mat_q4_0_host(const ggml_b200_weight * w, const float * x_host, int64_t n,
                                         size_t nb11, float * y_host, size_t nb1);
/* The _dev op WITHOUT ANY ALLOCATION.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, dispatch and result bits as ggml_b200_mul_mat_q4_0; the quantized activations (kernel 1's output)
 * go to a file-scope `static __device__` array instead of the growing workspace, and no cudaMalloc / cudaFree /
 * allocating library routine is called.  Limit: n * K <= ggml_b200_static_max_elems() = 4 Mi activations per call
 * (e.g. n = 512 at K = 8192, n = 1024 at K = 4096; the array is 6 MiB and every GPU that loads the library holds it
 * until exit); larger calls return GGML_B200_ERR_ARG.  All calls share the array and it keeps the last call's
 * contents; calls that overlap on the GPU (different streams) race on it -- serialise them. */
int          ggml_b200_mul_mat_q4_0_static(const ggml_b200_weight * w, const float * x_dev, int64_t n,
                                           size_t nb11, float * y_dev, size_t nb1);
size_t       ggml_b200_static_max_elems(void);

/* ---- the op's element-wise neighbours in a LLaMA layer ----------------------------
 * so that activations stay in HBM between consecutive Q4_0 mul_mats (what ggml_cuda_assign_buffers,
 * ggml-cuda.cu:2830-2892, achieves in the reference).  Device pointers, asynchronous on the current stream, results
 * follow the reference's CPU code:
 *   rms_norm : ggml_compute_forward_rms_norm_f32 (ggml.c:10400-10450; sum of squares in double, eps 1e-6 by the caller)
 *   mul, add : ggml_compute_forward_mul_f32 / _add_f32 (ggml.c:9160-9240, :8480-8560); src1 has nb elements and is
 *              repeated along src0 (nb == n, or one row broadcast over the rows: the norm weights)
 *   silu     : ggml_vec_silu_f32 with GGML_SILU_FP16 (ggml.c:3530-3538): y = fp16(silu(fp16(x)))
 *   silu_mul : silu(x) * b in one kernel (llama.cpp:1492-1498, the FFN gate) */
int          ggml_b200_rms_norm_f32(const float * x_dev, float * y_dev, int64_t ncols, int64_t nrows, float eps);
int          ggml_b200_mul_f32(const float * a_dev, const float * b_dev, float * y_dev, int64_t n, int64_t nb);
int          ggml_b200_add_f32(const float * a_dev, const float * b_dev, float * y_dev, int64_t n, int64_t nb);
int          ggml_b200_silu_f32(const float * x_dev, float * y_dev, int64_t n);
int          ggml_b200_silu_mul_f32(const float * x_dev, const float * b_dev, float * y_dev, int64_t n,
                                    float * silu_dev /* NULL, or also receives silu(x) */);
/* rope, mode 0 (ggml_compute_forward_rope_f32, ggml.c:12714-12850; behind wq / wk in llama.cpp:1342-1347): x_dev is
 * [npos][rows_per_pos][ne0], adjacent pairs (2k, 2k+1) of every row are rotated by cos_sin_dev[pos][k] = {cos, sin}; the
 * caller supplies the angles (the glue runs the reference's float recurrence theta_k = p * scale^k on the host, so they are
 * the reference's bit for bit).  y_dev may be x_dev (rope_inplace). */
int          ggml_b200_rope_f32(const float * x_dev, float * y_dev, const float * cos_sin_dev, int64_t ne0,
                                int64_t rows_per_pos, int64_t npos);
/* rms_norm and the MUL by the norm weights that follows it in every layer, one kernel; both results are written */
int          ggml_b200_rms_norm_mul_f32(const float * x_dev, float * y_norm_dev, const float * w_dev, float * y_dev,
                                        int64_t ncols, int64_t nrows, int64_t nw, float eps);
/* The decode (n = 1) op with the element-wise node BEHIND it folded into the kernel's store: no launch and no trip of
 * the row through L2 for that node (the old back-end ran them as kernels of their own,
 * ggml-cuda.cu:2830-2892).  Every result the graph names is still written; y / y1 / y3 are bit-identical to
 * ggml_b200_mul_mat_q4_0, the fused result applies the same single fp32 operation as ggml_b200_add_f32 / _silu_mul_f32.
 *   _add    : y = W x;  y2 = y + r                       residual add behind wo / w2 (llama.cpp:1462, :1504)
 *   _swiglu : y1 = W1 x; y3 = W3 x; silu = silu(y1); y2 = silu * y3     the FFN gate (llama.cpp:1479-1497): x is read
 *             once, row i of both weights streams through the same warp
 * Device pointers, asynchronous; n = 1 only, the whole weight on this GPU in the plane layout (ggml_b200_neigh_ok
 * returns 1), GGML_B200_ERR_ARG otherwise. */
int          ggml_b200_neigh_ok(const ggml_b200_weight * w);
int          ggml_b200_mul_mat_q4_0_add(const ggml_b200_weight * w, const float * x_dev, float * y_dev,
                                        const float * r_dev, float * y2_dev);
int          ggml_b200_mul_mat_q4_0_swiglu(const ggml_b200_weight * w1, const ggml_b200_weight * w3, const float * x_dev,
                                           float * y1_dev, float * y3_dev, float * silu_dev /* may be NULL */,
                                           float * y2_dev);
/* Marks on the current stream: ggml_b200_wait_mark(slot) returns when everything queued before ggml_b200_mark(slot) has
 * finished, whatever was queued behind it (the glue submits the next segment of a decode graph ahead of the CPU node in
 * front of it and waits for the one before; the reference waits for whole streams, ggml-cuda.cu:2514-2552). */
#define GGML_B200_MARKS 8
int          ggml_b200_mark(int slot);
int          ggml_b200_wait_mark(int slot);
/* A sequence of this library's asynchronous calls (ops, element-wise kernels, async copies) recorded once and submitted
 * as one CUDA graph launch afterwards: begin ... calls ... end -> handle; launch runs it on the current stream.  The
 * glue uses it for the device nodes between two CPU nodes of a ggml graph (same pointers every eval).  No allocation
 * may happen between begin and end. */
int          ggml_b200_graph_begin(void);
int          ggml_b200_graph_end(void ** exec_out);
int          ggml_b200_graph_launch(void * exec);
void         ggml_b200_graph_free(void * exec);

/* ---- a list of n = 1 ops in ONE launch ----------------------------------------
 * replaces, for the run of GGML_OP_MUL_MAT nodes of a decode step, the per-node loop of
 * ggml_graph_compute (ggml.c:17541-17720) entering ggml_compute_forward_mul_mat_q_f32
 * (ggml.c:11226-11424) once per node: one persistent kernel streams the weights of all ops back to
 * back (gemv_q4_0_set.cuh).  Semantics and result BITS are those of
 *     for each op in order: ggml_b200_mul_mat_q4_0(op.w, op.x_dev, 1, 4*K, op.y_dev, 4*M)
 * Ordering follows the DATA hazards between the ops: ops that share x_dev run back to back
 * (wq/wk/wv, w3/w1; llama.cpp:1394-1504); an op whose x_dev is exactly the dst of one earlier op
 * waits for that op's rows only (they travel as tagged words, no barrier); any other overlap of an
 * op's x or dst with an earlier op's x or dst is a grid-wide dependency barrier, split in two: every
 * CTA arrives when it has finished the last op the run conflicts with, and the wait sits at the head
 * of the run (ops later, normally long satisfied).  x_dev 16-byte aligned (ggml tensors are), weights
 * in the plane layout.  Tuning keys (ggml_b200_set_tuning): "set_warps" consumer warps per CTA
 * (8/12/16/20, default 12), "set_slots" ring depth (0 = consumer groups + 1), "set_link" 0 = barriers
 * instead of tagged words, "set_split_barrier" 0 = arrive right before the wait, "set_barrier_slack".  create copies the list; the buffers must stay valid.  run is
 * asynchronous on the current stream and may be captured in a CUDA graph. */
typedef struct ggml_b200_set ggml_b200_set;
typedef struct ggml_b200_set_op {
    const ggml_b200_weight * w;
    const float *            x_dev;   /* K floats */
    float *                  y_dev;   /* the FULL dst row (M floats); a shard writes [row_lo,row_hi) */
} ggml_b200_set_op;
ggml_b200_set * ggml_b200_set_create(const ggml_b200_set_op * ops, int n_ops);
/* The same list on row shards, one rank per GPU (replaces, for a decode step, GGML_BACKEND_GPU_SPLIT's per-op
 * row slices + per-column cudaMemcpyAsync gather, ggml-cuda.cu:2353-2552): COLLECTIVE over the communicator of
 * ggml_b200_comm_init -- every rank passes the same list with its own shards (ggml_b200_shard_rows) and its own
 * replicated x / full-size dst buffers.  One persistent launch per rank and step; the exchange is INSIDE it: a
 * finished row goes to this rank's dst and, as a tagged 8-byte word, into every rank's arena over NVLink; runs whose
 * x is another op's dst read the gathered vector from the arena; a gather warp per CTA completes every dst on every
 * rank (the contract of ggml_b200_mul_mat_q4_0_sharded).  No NCCL call and no launch per mul_mat.  With a world of 1
 * it is ggml_b200_set_create.  Every rank must call run the same number of times. */
ggml_b200_set * ggml_b200_set_create_sharded(const ggml_b200_set_op * ops, int n_ops);
int          ggml_b200_set_run(ggml_b200_set * s);
void         ggml_b200_set_free(ggml_b200_set * s);
int          ggml_b200_set_ops(const ggml_b200_set * s);       /* ops with a non-empty shard */
int          ggml_b200_set_barriers(const ggml_b200_set * s);  /* grid-wide dependency barriers per launch */
int          ggml_b200_set_linked_runs(const ggml_b200_set * s); /* runs whose x is read from its producer op's tagged rows */
int          ggml_b200_set_slots(const ggml_b200_set * s);     /* depth of the shared-memory ring */
int          ggml_b200_set_warps(const ggml_b200_set * s);     /* consumer warps per CTA ("set_warps", fewer when the ring is shallower than the groups) */
uint64_t     ggml_b200_set_bytes(const ggml_b200_set * s);     /* algorithmic bytes per run */
/* The hazard analysis of create, alone (pure host code: no device, nothing dereferenced; x and y are addresses as
 * integers): flags_out[i] = what the kernel is told about op i -- bit 0 a dependency barrier is waited for at its head,
 * bit 1 it starts a new run (x is (re)quantized), bit 2 no op of the list writes its x, bits 8..14 how many barriers
 * every CTA arrives at when it has finished op i; link_out[i] = the op whose result rows op i's run reads as x (tagged
 * words), or -1.  Ops with an empty shard get 0 / -1.  Returns the number of barriers, or an error (< 0). */
typedef struct ggml_b200_set_plan_op {
    uint64_t x, y;                    /* addresses of src1 and of the FULL dst row */
    int64_t  K, M, row_lo, row_hi;    /* the weight shard: rows [row_lo, row_hi) of M, K columns */
} ggml_b200_set_plan_op;
int          ggml_b200_set_plan(const ggml_b200_set_plan_op * ops, int n_ops, int * flags_out, int * link_out,
                                int * linked_runs_out);
/* parity checkpoint of the same device code: S_dev[i] = [rows_i][K_i/32] int32 block dots of the
 * i-th non-empty op (device pointers, or NULL to skip that op); also writes y like run */
int          ggml_b200_set_checkpoint(ggml_b200_set * s, int32_t * const * S_dev, int n_ops);
/* profiling aid: one synchronous run of an instrumented build of the same kernel; out_host[cta*16 + i] = clock64
 * totals {whole kernel, a consumer warp waiting for weights, in dependency barriers, quantizing x, the producer
 * waiting for a free ring slot, that warp from data ready to y stored, its tiles, in the tile loops, 0 x 4 (the
 * barrier's pieces moved to the sentinel thread), loop bookkeeping x 4}.  Returns the number of CTAs written (<= max_ctas) or an error. */
int          ggml_b200_set_profile(ggml_b200_set * s, long long * out_host, int max_ctas);

/* ---- parity checkpoint --------------------------------------------------------------------
 * int32 per-block dots S_b = sum_j (q4_j - 8) * q8_j exactly as the kernels form them
 * (definition: scalar loop ggml.c:2593-2600).  S_dev: [n][rows][K/32] int32, device pointer.
 * path: GGML_B200_PATH_GEMV or GGML_B200_PATH_GEMM (the same device code as the op, with the
 * dump store enabled). */
int          ggml_b200_block_dots_q4_0(const ggml_b200_weight * w, const float * x_dev, int64_t n,
                                       size_t nb11, int32_t * S_dev, int path);

/* the same checkpoint for the n = 1 kernel that fuses kernel 1 into kernel 2: the q8_0 blocks its
 * in-CTA quantizer formed (blocks34_dev: K/32 reference-format blocks, may be NULL) and its S_b
 * (S_dev: [rows][K/32] int32, may be NULL).  x_dev: one row of K floats. */
int          ggml_b200_fused_checkpoint_q4_0(const ggml_b200_weight * w, const float * x_dev,
                                             int32_t * S_dev, void * blocks34_dev);

/* ---- sibling weight formats: Q8_0, Q4_1, Q5_0 and Q5_1 through the same op ----------------------------
 * The reference's operator is format blind (ggml_compute_forward_mul_mat_q_f32, ggml.c:11268-11411); quantize_fns[]
 * (ggml.c:1619-1670) binds per type the activation quantizer and the block dot:
 *   GGML_TYPE_Q8_0 (8): block_q8_0 weights (ggml.c:902-907) x quantize_row_q8_0 activations, ggml_vec_dot_q8_0_q8_0 (ggml.c:3231-3330)
 *   GGML_TYPE_Q4_1 (3): block_q4_1 weights (ggml.c:877-883) x quantize_row_q8_1 activations (ggml.c:1318-1480),
 *                       ggml_vec_dot_q4_1_q8_1 (ggml.c:2609-2734)
 *   GGML_TYPE_Q5_0 (6): block_q5_0 (ggml.c:885-891) x q8_0 activations, ggml_vec_dot_q5_0_q8_0 (ggml.c:2735-2973)
 *   GGML_TYPE_Q5_1 (7): block_q5_1 (ggml.c:893-900) x q8_1 activations, ggml_vec_dot_q5_1_q8_1 (ggml.c:2975-3229)
 * `type` takes the reference's enum values (ggml.h:230-242).  Decode path (dp4a GEMV, the activation quantizer fused into
 * the kernel, passes of 4 columns); no tensor-core path for these formats. */
#define GGML_B200_TYPE_Q4_1 3
#define GGML_B200_TYPE_Q5_0 6
#define GGML_B200_TYPE_Q5_1 7
#define GGML_B200_TYPE_Q8_0 8
typedef struct ggml_b200_qweight ggml_b200_qweight;
/* blocks: M rows of K/32 reference-format blocks (34 / 20 / 22 / 24 B), row stride nb1 bytes (0 = contiguous); host memory */
ggml_b200_qweight * ggml_b200_upload_quant(int type, const void * blocks, int64_t M, int64_t K, size_t nb1);
void         ggml_b200_free_quant(ggml_b200_qweight * w);
int          ggml_b200_quant_type(const ggml_b200_qweight * w);
size_t       ggml_b200_quant_bytes(const ggml_b200_qweight * w);      /* device bytes of the planes */
/* dst[c][m] = row m of w . column c of x: same argument meaning as ggml_b200_mul_mat_q4_0 (device pointers, byte strides) */
int          ggml_b200_mul_mat_quant(const ggml_b200_qweight * w, const float * x_dev, int64_t n, size_t nb11,
                                     float * y_dev, size_t nb1);
/* parity checkpoints: the same kernel with the S_b store enabled (S_dev: [n][M][K/32] int32), and the activation blocks
 * the fused quantizer forms, in the reference's format (block_q8_0 34 B for Q8_0, block_q8_1 40 B for Q4_1) */
int          ggml_b200_block_dots_quant(const ggml_b200_qweight * w, const float * x_dev, int64_t n, size_t nb11,
                                        int32_t * S_dev, float * y_dev, size_t nb1);
int          ggml_b200_quantize_dot_blocks(int type, const float * x_dev, int64_t n, int64_t K, size_t nb11, void * blocks_dev);

/* ---- multi-GPU: row shards + all-gather ---------------------------------------------------
 * replaces the row-split placement and gather of GGML_BACKEND_GPU_SPLIT (ggml_cuda_set_tensor_split
 * ggml-cuda.cu:1863-1882, slices :2780-2797, gather by cudaMemcpyAsync :2514-2540) with one
 * rank per GPU and one NCCL all-gather over NVLink.
 * unique_id: the 128-byte ncclUniqueId, produced on rank 0 by ggml_b200_comm_unique_id and
 * broadcast by the launcher (torch.distributed, MPI, a file ...). */
int          ggml_b200_comm_unique_id(void * out128);
int          ggml_b200_comm_init(int rank, int world, const void * unique_id128);
void         ggml_b200_comm_shutdown(void);
int          ggml_b200_comm_rank(void);
int          ggml_b200_comm_world(void);
/* the row range rank r of `world` owns for a tensor of M rows (equal shards rounded to
 * `GGML_B200_SHARD_ALIGN` rows; the reference rounds to its own granularity at
 * ggml-cuda.cu:2784-2797) */
#define GGML_B200_SHARD_ALIGN 128
void         ggml_b200_shard_rows(int64_t M, int rank, int world, int64_t * row_lo, int64_t * row_hi);
/* mul_mat on this rank's shard + all-gather so that EVERY rank ends with the full y[n][M]
 * (device pointers; asynchronous on the current stream).  x must already be replicated. */
int          ggml_b200_mul_mat_q4_0_sharded(const ggml_b200_weight * w, const float * x_dev, int64_t n,
                                            size_t nb11, float * y_dev, size_t nb1);

/* ---- introspection for bench.py / tests ---------------------------------------------------*/
/* number of kernels THIS library launched since init (the bench's gpu_launches claim) */
uint64_t     ggml_b200_launch_count(void);
/* per-CTA timeline trace of the n = 1 kernels (profiling aid): records of 8 x u64
 * {kernel seq, blockIdx | smid << 32, t_start, t_after_dependency_wait, t_after_prologue, t_end, 0, 0}
 * in globaltimer ns.  trace_end returns the number of records copied to `out`. */
int          ggml_b200_trace_begin(unsigned int capacity_records);
int          ggml_b200_trace_end(unsigned long long * out, unsigned int max_records);
/* kernel 3 timeline (profiling aid): after ggml_b200_set_tuning("gemm_trace", 1), CTA (0,0) of every GEMM
 * launch stamps clock64 at its pipeline hand-offs; out = 4 roles (TMA, MMA, epilogue warp 0, unpack) x 2048 */
int          ggml_b200_gemm_trace_read(long long * out, int count);
/* plain device memory helpers so a ctypes harness needs no other CUDA binding */
void *       ggml_b200_dev_malloc(size_t size);
void         ggml_b200_dev_free(void * ptr);
int          ggml_b200_memcpy_h2d(void * dst_dev, const void * src_host, size_t size);
int          ggml_b200_memcpy_d2h(void * dst_host, const void * src_dev, size_t size);
int          ggml_b200_synchronize(void);
/* the same copies without the stream sync (the glue keeps activations on the device and syncs once per CPU node) */
int          ggml_b200_memcpy_h2d_async(void * dst_dev, const void * src_host, size_t size);
int          ggml_b200_memcpy_d2h_async(void * dst_host, const void * src_dev, size_t size);

#ifdef __cplusplus
}
#endif
#endif /* GGML_B200_H */
