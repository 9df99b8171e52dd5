// comm.cu -- multi-GPU: row shards + NCCL all-gather over NVLink (one rank per GPU).
//
// Replaces the reference's only multi-device mechanism, GGML_BACKEND_GPU_SPLIT: row slices per
// device (ggml-cuda.cu:2780-2797) gathered with per-column cudaMemcpyAsync (:2514-2540) from a
// single host thread.  Here every rank owns rows [row_lo,row_hi) of each weight, computes its
// slab of dst, and one ncclAllGather reassembles dst on every rank.
//
// NCCL is bound at run time (dlopen "libnccl.so.2") so the library has no link-time dependency
// and shares the copy torch already loaded when driven from a torch.distributed launcher.
#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "p2p.cuh"
#include "state.h"

namespace b200 {

typedef struct ncclComm * ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclFloat32 = 7 };

// ---- n = 1 gather without a collective library call (protocol: p2p.cuh) -----------------------------------
// Stand-alone form, used after a GEMV that could not take the fused kernel (gemv_q4_0_p2p.cuh): this rank's rows
// are already in y; a few CTAs send them to every peer and collect the peers' rows.
// mode (tuning key p2p_gather): 1 = the protocol; 3 = no poll, 4 = no store, no poll (timing experiments only)
constexpr int P2P_GATHER_CTAS = 8;

__global__ void __launch_bounds__(1024)
p2p_gather_kernel(const __grid_constant__ P2PExchange ex, float * __restrict__ y, int mode) {
    asm volatile("griddepcontrol.wait;" ::: "memory");     // this rank's rows come from the GEMV just before
    __shared__ unsigned s_seq;
    if (threadIdx.x == 0) { s_seq = p2p_counter(ex) + P2P_MAX_RUN; p2p_head_publish_and_release(ex, s_seq); }   // always a dependent op
    __syncthreads();
    const unsigned seq = s_seq;
    int lo_me, n_me, j0, j1;
    p2p_rows_of(ex, ex.me, lo_me, n_me);
    p2p_cta_slice(n_me, j0, j1);
    if (mode != 4) p2p_send(ex, seq, y + lo_me + j0, lo_me + j0, j1 - j0);
    if (mode == 1) p2p_receive(ex, seq, y);
}

struct Nccl {
    void * lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char * (*GetErrorString)(ncclResult_t) = nullptr;
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    float * gather_buf = nullptr; size_t gather_cap = 0;   // [world][n][rows_max]
    float * slab_buf = nullptr;   size_t slab_cap = 0;     // [n][rows_max] (this rank)
    // n = 1 fast path: peer-to-peer gather over NVLink (see p2p_gather_kernel)
    unsigned char * arena = nullptr;                       // this rank's IPC-exported arena
    P2PPeers peers;                                        // every rank's arena as mapped in this process
    bool p2p_ok = false;
    // prefill exchange (n >= gemm_min_n): kernel 3 stores its tiles straight into the peers' landing buffers
    float * pf_local = nullptr;                            // [2 regions][pf_cap floats] + PF_SLOTS arrival counters, IPC-exported
    void * pf_mapped[P2P_MAX_WORLD] = {};                  // the same block of every rank as mapped here
    size_t pf_cap = 0;
    unsigned long long pf_seq = 0;                         // sharded prefill ops so far (identical on every rank)
    unsigned pf_target[64] = {};                           // arrivals a counter slot must have seen once its current op is complete
};
constexpr int PF_SLOTS = 64;
static Nccl nc;
static int g_p2p_enabled = 1;   // tuning key "p2p_gather"

static int nccl_load() {
    if (nc.lib) return GGML_B200_OK;
    const char * names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char * nm : names) { nc.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL); if (nc.lib) break; }
    if (!nc.lib) { set_error("cannot dlopen libnccl.so.2: %s", dlerror()); return GGML_B200_ERR_CUDA; }
#define SYM(field, name) *(void **)(&nc.field) = dlsym(nc.lib, name); if (!nc.field) { set_error("libnccl lacks %s", name); return GGML_B200_ERR_CUDA; }
    SYM(GetUniqueId, "ncclGetUniqueId")
    SYM(CommInitRank, "ncclCommInitRank")
    SYM(CommDestroy, "ncclCommDestroy")
    SYM(AllGather, "ncclAllGather")
    SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
    return GGML_B200_OK;
}

#define NK(call)                                                                              \
    do {                                                                                      \
        ncclResult_t r_ = (call);                                                             \
        if (r_ != 0) { set_error("NCCL error %d (%s): %s", r_, nc.GetErrorString(r_), #call); return GGML_B200_ERR_CUDA; } \
    } while (0)

// gather_buf[r][c][j] (rank r's slab, rows_r(r) wide but stored with stride rows_max) -> y[c][row_lo(r) + j]
__global__ void scatter_slabs_kernel(const float * __restrict__ gbuf, float * __restrict__ y, long long ldy,
                                     int n, long long M, int world, long long rows_max) {
    const long long total = (long long)n * M;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long c = i / M, m = i - c * M;
        // owner of row m under ggml_b200_shard_rows
        long long per = ((M + world - 1) / world + GGML_B200_SHARD_ALIGN - 1) / GGML_B200_SHARD_ALIGN * GGML_B200_SHARD_ALIGN;
        long long r = m / per;
        if (r >= world) r = world - 1;
        const long long j = m - r * per;
        y[c * ldy + m] = gbuf[(r * n + c) * rows_max + j];
    }
}

// Prefill exchange, receiving side: wait until every peer's tiles of this op have landed in this rank's region (one
// release-add per tile on the op's counter slot, gemm_q4_0_tc3.cuh), then move the peers' rows into dst.  The peers are
// other GPUs driven by other processes: the poll cannot starve what it waits for.
__global__ void __launch_bounds__(512) pf_collect_kernel(const float * __restrict__ region, const unsigned * __restrict__ ctr, unsigned target,
                                                          float * __restrict__ y, long long ldy, int n, int M, int world, int me, int per) {
    pdl_wait();
    pdl_launch_dependents();
    if (threadIdx.x == 0) {
        const unsigned long long t0 = p2p_now();
        unsigned v;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
            if (p2p_now() - t0 > P2P_TIMEOUT_NS) __trap();
        } while ((int)(v - target) < 0);
    }
    __syncthreads();
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nthr = (long long)gridDim.x * blockDim.x;
    for (int r = 0; r < world; r++) {
        if (r == me) continue;
        const int lo = min(r * per, M), width = min(lo + per, M) - lo;
        if (width <= 0) continue;
        const bool vec = (width & 3) == 0 && (lo & 3) == 0 && (M & 3) == 0 && (ldy & 3) == 0 &&
                         ((reinterpret_cast<uintptr_t>(region) | reinterpret_cast<uintptr_t>(y)) & 15) == 0;
        if (vec) {
            const int w4 = width >> 2;
            for (long long i = tid; i < (long long)n * w4; i += nthr) {
                const long long c = i / w4; const int j = (int)(i - c * w4) * 4;
                *reinterpret_cast<float4 *>(y + c * ldy + lo + j) = __ldcg(reinterpret_cast<const float4 *>(region + c * M + lo + j));
            }
        } else {
            for (long long i = tid; i < (long long)n * width; i += nthr) {
                const long long c = i / width; const int j = (int)(i - c * width);
                y[c * ldy + lo + j] = __ldcg(region + c * M + lo + j);
            }
        }
    }
}

void comm_set_p2p(int on) { g_p2p_enabled = on; }

// Export this rank's arena, learn everybody's handle through the communicator, map the peers.
// Any failure leaves p2p_ok false and the NCCL gather in charge.
static int p2p_setup() {
    nc.p2p_ok = false;
    if (nc.world > P2P_MAX_WORLD) return GGML_B200_OK;
    const char * env = getenv("GGML_B200_P2P");
    if (env && atoi(env) == 0) return GGML_B200_OK;
    CK(cudaMalloc(&nc.arena, P2P_ARENA_BYTES));
    CK(cudaMemset(nc.arena, 0, P2P_ARENA_BYTES));
    cudaIpcMemHandle_t mine;
    cudaIpcMemHandle_t all[P2P_MAX_WORLD];
    int ok = cudaIpcGetMemHandle(&mine, nc.arena) == cudaSuccess;
    if (!ok) { cudaGetLastError(); memset(&mine, 0, sizeof(mine)); }
    // handles (+ an "I could export" byte) travel through one all-gather
    constexpr size_t REC = sizeof(cudaIpcMemHandle_t) + 16;
    unsigned char * dev = nullptr;
    CK(cudaMalloc(&dev, REC * (nc.world + 1)));
    unsigned char rec[REC] = {0};
    memcpy(rec, &mine, sizeof(mine));
    rec[sizeof(mine)] = (unsigned char)ok;
    CK(cudaMemcpyAsync(dev + REC * nc.world, rec, REC, cudaMemcpyHostToDevice, g.stream));
    NK(nc.AllGather(dev + REC * nc.world, dev, REC / 4, ncclFloat32, nc.comm, g.stream));
    unsigned char host[REC * P2P_MAX_WORLD];
    CK(cudaMemcpyAsync(host, dev, REC * nc.world, cudaMemcpyDeviceToHost, g.stream));
    CK(cudaStreamSynchronize(g.stream));
    CK(cudaFree(dev));
    int all_ok = 1;
    for (int r = 0; r < nc.world; r++) { memcpy(&all[r], host + REC * r, sizeof(all[r])); all_ok &= host[REC * r + sizeof(all[r])]; }
    int mapped = all_ok;
    for (int r = 0; r < P2P_MAX_WORLD; r++) nc.peers.arena[r] = nullptr;
    nc.peers.arena[nc.rank] = nc.arena;
    for (int r = 0; r < nc.world && mapped; r++) {
        if (r == nc.rank) continue;
        void * p = nullptr;
        if (cudaIpcOpenMemHandle(&p, all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); mapped = 0; break; }
        nc.peers.arena[r] = (unsigned char *)p;
    }
    // every rank must agree, or one would wait for flags nobody raises
    int * flag_dev = nullptr;
    CK(cudaMalloc(&flag_dev, 4 * (nc.world + 1)));
    float f = mapped ? 1.0f : 0.0f;
    CK(cudaMemcpyAsync(flag_dev + nc.world, &f, 4, cudaMemcpyHostToDevice, g.stream));
    NK(nc.AllGather(flag_dev + nc.world, flag_dev, 1, ncclFloat32, nc.comm, g.stream));
    float fl[P2P_MAX_WORLD];
    CK(cudaMemcpyAsync(fl, flag_dev, 4 * nc.world, cudaMemcpyDeviceToHost, g.stream));
    CK(cudaStreamSynchronize(g.stream));
    CK(cudaFree(flag_dev));
    bool agree = true;
    for (int r = 0; r < nc.world; r++) agree = agree && fl[r] == 1.0f;
    nc.p2p_ok = agree;
    if (getenv("GGML_B200_VERBOSE")) fprintf(stderr, "ggml_b200: rank %d peer-to-peer gather %s\n", nc.rank, agree ? "on" : "off (NCCL gather)");
    return GGML_B200_OK;
}

bool comm_p2p_ready() { return nc.comm != nullptr && nc.p2p_ok; }

int comm_share(void * local, void ** mapped) {
    for (int r = 0; r < P2P_MAX_WORLD; r++) mapped[r] = nullptr;
    mapped[nc.rank] = local;
    if (nc.world == 1) return GGML_B200_OK;
    if (!comm_p2p_ready()) { set_error("peer access is not set up"); return GGML_B200_ERR_STATE; }
    cudaIpcMemHandle_t mine;
    CK(cudaIpcGetMemHandle(&mine, local));
    constexpr size_t REC = sizeof(cudaIpcMemHandle_t);
    static_assert(REC % 4 == 0, "handles travel as floats");
    unsigned char * dev = nullptr;
    CK(cudaMalloc(&dev, REC * (nc.world + 1)));
    CK(cudaMemcpyAsync(dev + REC * nc.world, &mine, REC, cudaMemcpyHostToDevice, g.stream));
    NK(nc.AllGather(dev + REC * nc.world, dev, REC / 4, ncclFloat32, nc.comm, g.stream));
    cudaIpcMemHandle_t all[P2P_MAX_WORLD];
    CK(cudaMemcpyAsync(all, dev, REC * nc.world, cudaMemcpyDeviceToHost, g.stream));
    CK(cudaStreamSynchronize(g.stream));
    CK(cudaFree(dev));
    for (int r = 0; r < nc.world; r++) {
        if (r == nc.rank) continue;
        void * p = nullptr;
        CK(cudaIpcOpenMemHandle(&p, all[r], cudaIpcMemLazyEnablePeerAccess));
        mapped[r] = p;
    }
    return GGML_B200_OK;
}

void comm_unshare(void ** mapped) {
    for (int r = 0; r < P2P_MAX_WORLD; r++) {
        if (mapped[r] && r != nc.rank) cudaIpcCloseMemHandle(mapped[r]);
        mapped[r] = nullptr;
    }
}

// the landing block of the prefill exchange: grown collectively (every rank sees the same shapes in the same order)
static int pf_ensure(size_t need_floats) {
    if (need_floats <= nc.pf_cap) return GGML_B200_OK;
    CK(cudaStreamSynchronize(g.stream));                       // this rank's tiles for the old block are out
    if (nc.pf_local) { comm_unshare(nc.pf_mapped); CK(cudaFree(nc.pf_local)); nc.pf_local = nullptr; nc.pf_cap = 0; }
    size_t cap = (size_t)8 << 20;                              // floats per region
    while (cap < need_floats) cap *= 2;
    const size_t bytes = 2 * cap * sizeof(float) + PF_SLOTS * sizeof(unsigned);
    CK(cudaMalloc(&nc.pf_local, bytes));
    CK(cudaMemsetAsync(reinterpret_cast<unsigned char *>(nc.pf_local) + 2 * cap * sizeof(float), 0, PF_SLOTS * sizeof(unsigned), g.stream));
    CK(cudaStreamSynchronize(g.stream));
    int rc = comm_share(nc.pf_local, nc.pf_mapped);            // (all-gather of the handles: also the barrier behind the old block)
    if (rc) return rc;
    nc.pf_cap = cap;
    nc.pf_seq = 0;
    memset(nc.pf_target, 0, sizeof(nc.pf_target));
    return GGML_B200_OK;
}

static void p2p_teardown() {
    if (nc.pf_local) { comm_unshare(nc.pf_mapped); cudaFree(nc.pf_local); nc.pf_local = nullptr; nc.pf_cap = 0; nc.pf_seq = 0; }
    for (int r = 0; r < P2P_MAX_WORLD; r++) {
        if (nc.peers.arena[r] && r != nc.rank) cudaIpcCloseMemHandle(nc.peers.arena[r]);
        nc.peers.arena[r] = nullptr;
    }
    if (nc.arena) { cudaFree(nc.arena); nc.arena = nullptr; }
    nc.p2p_ok = false;
}

}  // namespace b200

using namespace b200;

extern "C" {

void ggml_b200_shard_rows(int64_t M, int rank, int world, int64_t * row_lo, int64_t * row_hi) {
    if (world < 1) world = 1;
    int64_t per = ((M + world - 1) / world + GGML_B200_SHARD_ALIGN - 1) / GGML_B200_SHARD_ALIGN * GGML_B200_SHARD_ALIGN;
    int64_t lo = per * rank, hi = per * (rank + 1);
    if (lo > M) lo = M;
    if (hi > M) hi = M;
    if (row_lo) *row_lo = lo;
    if (row_hi) *row_hi = hi;
}

int ggml_b200_comm_unique_id(void * out128) {
    int rc = nccl_load();
    if (rc) return rc;
    ncclUniqueId id;
    NK(nc.GetUniqueId(&id));
    memcpy(out128, &id, sizeof(id));
    return GGML_B200_OK;
}

int ggml_b200_comm_init(int rank, int world, const void * unique_id128) {
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (world < 1 || rank < 0 || rank >= world) { set_error("bad rank %d / world %d", rank, world); return GGML_B200_ERR_ARG; }
    if (nc.comm) { set_error("communicator already initialised"); return GGML_B200_ERR_STATE; }
    nc.rank = rank; nc.world = world;
    if (world == 1) return GGML_B200_OK;
    int rc = nccl_load();
    if (rc) return rc;
    ncclUniqueId id;
    memcpy(&id, unique_id128, sizeof(id));
    NK(nc.CommInitRank(&nc.comm, world, id, rank));
    return p2p_setup();
}

void ggml_b200_comm_shutdown(void) {
    if (nc.comm) cudaDeviceSynchronize();   // peers may still be storing into this arena
    p2p_teardown();
    if (nc.comm) { nc.CommDestroy(nc.comm); nc.comm = nullptr; }
    if (nc.gather_buf) { cudaFree(nc.gather_buf); nc.gather_buf = nullptr; nc.gather_cap = 0; }
    if (nc.slab_buf) { cudaFree(nc.slab_buf); nc.slab_buf = nullptr; nc.slab_cap = 0; }
    nc.rank = 0; nc.world = 1;
}

int ggml_b200_comm_rank(void) { return nc.rank; }
int ggml_b200_comm_world(void) { return nc.world; }

int ggml_b200_mul_mat_q4_0_sharded(const ggml_b200_weight * w, const float * x_dev, int64_t n, size_t nb11, float * y_dev, size_t nb1) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!w || !x_dev || !y_dev) { set_error("null argument"); return GGML_B200_ERR_ARG; }
    if (nc.world == 1) return ggml_b200_mul_mat_q4_0(w, x_dev, n, nb11, y_dev, nb1);
    // the op contract of the single-GPU entry (ggml.c:11279-11285): n, src1 and dst row strides
    if (n < 0 || n > INT32_MAX) { set_error("bad n"); return GGML_B200_ERR_ARG; }
    if (nb11 < (size_t)w->K * 4 || (nb11 & 3)) { set_error("src1 row stride %zu < K*4 or not a multiple of 4", nb11); return GGML_B200_ERR_ARG; }
    if (nb1 < (size_t)w->M * 4 || (nb1 & 3)) { set_error("dst row stride %zu < M*4 or not a multiple of 4", nb1); return GGML_B200_ERR_ARG; }
    if (!nc.comm) { set_error("ggml_b200_comm_init was not called"); return GGML_B200_ERR_STATE; }
    int64_t lo, hi;
    ggml_b200_shard_rows(w->M, nc.rank, nc.world, &lo, &hi);
    if (lo != w->row_lo || hi != w->row_hi) { set_error("weight shard [%lld,%lld) is not rank %d's shard [%lld,%lld)", (long long)w->row_lo, (long long)w->row_hi, nc.rank, (long long)lo, (long long)hi); return GGML_B200_ERR_ARG; }
    if (n == 0) return GGML_B200_OK;
    int64_t l0, h0;
    ggml_b200_shard_rows(w->M, 0, nc.world, &l0, &h0);
    const int64_t rows_max = h0 - l0;   // rank 0 always holds a full-width shard
    const size_t slab = (size_t)n * rows_max;

    if (n == 1 && nc.p2p_ok && g_p2p_enabled && w->M <= P2P_MAX_M) {
        P2PExchange ex;
        ex.peers = nc.peers; ex.me = nc.rank; ex.world = nc.world; ex.per = (int)rows_max; ex.M = (int)w->M;
        // one kernel: GEMV of this rank's rows + delivery to every peer + collection of theirs
        if (g_p2p_enabled == 1 && fused_p2p_ok(w)) return launch_fused_p2p(w, x_dev, nb11, y_dev, ex);
        // otherwise GEMV straight into dst[lo,hi), then the stand-alone exchange kernel
        int rc = launch_slab(w, x_dev, 1, nb11, y_dev + lo, rows_max);
        if (rc) return rc;
        if (g_p2p_enabled == 2) return GGML_B200_OK;    // timing experiment: this rank's rows only
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = P2P_GATHER_CTAS; cfg.blockDim = 1024; cfg.stream = g.stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr; cfg.numAttrs = g.tune.pdl ? 1 : 0;
        CK(cudaLaunchKernelEx(&cfg, p2p_gather_kernel, ex, y_dev, g_p2p_enabled == 5 ? 1 : g_p2p_enabled));
        g.launches++;
        return GGML_B200_OK;
    }

    // n >= gemm_min_n: kernel 3 delivers its tiles to every peer itself and a collect kernel picks up theirs -- no
    // collective call, the exchange rides under the tiles still being computed (tuning key p2p_gather = 0: NCCL below)
    cudaStreamCaptureStatus cap_st = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(g.stream, &cap_st);
    if (n > 1 && nc.p2p_ok && g_p2p_enabled == 1 && cap_st == cudaStreamCaptureStatusNone && g.path != GGML_B200_PATH_GEMV &&
        (g.path == GGML_B200_PATH_GEMM || n >= g.gemm_min_n) && (w->rows() == 0 || gemm_available(w)) && !g.tune.skip_quantize) {
        int rc = pf_ensure((size_t)n * (size_t)w->M);
        if (rc) return rc;
        const unsigned long long seq = nc.pf_seq++;
        const int region = (int)(seq & 1), slot = (int)(seq % PF_SLOTS);
        const int ctiles = (int)((n + 127) / 128);
        unsigned expected = 0;
        float * dst[P2P_MAX_WORLD]; unsigned * ctr[P2P_MAX_WORLD];
        int nd = 0;
        for (int r = 0; r < nc.world; r++) {
            if (r == nc.rank) continue;
            int64_t rl, rh;
            ggml_b200_shard_rows(w->M, r, nc.world, &rl, &rh);
            expected += (unsigned)(((rh - rl) + 127) / 128) * (unsigned)ctiles;
            float * blk = (float *)nc.pf_mapped[r];
            dst[nd] = blk + (size_t)region * nc.pf_cap + lo;       // my rows inside r's landing region ([n][M])
            ctr[nd] = reinterpret_cast<unsigned *>(blk + 2 * nc.pf_cap) + slot;
            nd++;
        }
        const unsigned target = (nc.pf_target[slot] += expected);
        if (w->rows() > 0) {
            rc = launch_quantize(x_dev, n, w->K, nb11, nullptr);
            if (rc) return rc;
            rc = launch_gemm_peers(w, y_dev + lo, (long long)(nb1 / 4), (int)n, dst, ctr, nd, (long long)w->M);
            if (rc) return rc;
        }
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = (unsigned)(2 * g.sm_count); cfg.blockDim = 512; cfg.stream = g.stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr; cfg.numAttrs = g.tune.pdl ? 1 : 0;
        CK(cudaLaunchKernelEx(&cfg, pf_collect_kernel, (const float *)(nc.pf_local + (size_t)region * nc.pf_cap),
                              (const unsigned *)(reinterpret_cast<unsigned *>(nc.pf_local + 2 * nc.pf_cap) + slot), target, y_dev, (long long)(nb1 / 4), (int)n, (int)w->M,
                              nc.world, nc.rank, (int)rows_max));
        g.launches++;
        return GGML_B200_OK;
    }

    if (slab * nc.world * 4 > nc.gather_cap) {
        if (nc.gather_buf) CK(cudaFree(nc.gather_buf));
        nc.gather_buf = nullptr; nc.gather_cap = 0;
        CK(cudaMalloc(&nc.gather_buf, slab * nc.world * 4));
        nc.gather_cap = slab * nc.world * 4;
    }
    // n == 1 and every shard full-width: y itself is the gather buffer ([world][rows_max] == [M])
    const bool in_place = (n == 1) && (rows_max * nc.world == w->M);
    float * gbuf = in_place ? y_dev : nc.gather_buf;
    float * mine = gbuf + (size_t)nc.rank * slab;
    // this rank's slab, dense [n][rows_max]: reuse the op with a fake dst whose column 0 is `mine`
    int rc = launch_slab(w, x_dev, n, nb11, mine, rows_max);
    if (rc) return rc;
    NK(nc.AllGather(mine, gbuf, slab, ncclFloat32, nc.comm, g.stream));
    if (!in_place) {
        const long long total = (long long)n * w->M;
        const int threads = 256;
        long long blocks = (total + threads - 1) / threads;
        if (blocks > g.sm_count * 8) blocks = g.sm_count * 8;
        scatter_slabs_kernel<<<(unsigned)blocks, threads, 0, g.stream>>>(gbuf, y_dev, (long long)(nb1 / 4), (int)n, w->M, nc.world, rows_max);
        g.launches++;
        CK(cudaGetLastError());
    }
    return GGML_B200_OK;
}

}  // extern "C"
