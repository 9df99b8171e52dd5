// state.h -- process-wide state shared by the translation units of libggml_b200.so.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#include "../../include/ggml-b200.h"
#include "p2p.cuh"

// one device-resident Q4_0 row shard (opaque in the C ABI)
struct ggml_b200_weight {
    int64_t M = 0, K = 0;            // shape of the WHOLE tensor (ne01, ne00)
    int64_t row_lo = 0, row_hi = 0;  // rows this shard holds
    int     layout = GGML_B200_LAYOUT_PLANES;
    void *  raw = nullptr;           // LAYOUT_RAW: 18-byte blocks, rows contiguous
    uint4 * wq = nullptr;            // LAYOUT_PLANES: nibbles [rows][nbk]
    __half * wd = nullptr;           // LAYOUT_PLANES: scales  [rows][nbk]
    size_t  dev_bytes = 0;
    void *  gemm = nullptr;          // tensor-core path's per-weight state (TMA descriptors)
    mutable bool run_member = false; // seen in a run of n = 1 ops sharing src1 (overlap_same_src): gets the half-SM kernel
    int64_t rows() const { return row_hi - row_lo; }
};

namespace b200 {

constexpr int MAX_DYN_SMEM = 227 * 1024;

struct Tune {
    int gemv_threads = 512;      // threads per CTA of kernel 2
    int gemv_ctas_per_sm = 2;    // grid = SMs * this
    int gemv_unroll = 4;         // 16-byte weight loads in flight per lane
    int gemv_nc = 0;             // cap on columns per pass (0 = by n)
    int fused_rif = 1;           // rows in flight per warp of the fused kernel
    int gemv_fused = 1;          // n == 1: quantize inside the GEMV kernel (one launch per op)
    int fused_threads = 1024;    // threads per CTA of the fused kernel
    int fused_ctas_per_sm = 1;
    int small_copy = 1;          // rows <= 256 KB to / from mapped pinned host memory move with a kernel, not the copy engine
    int pdl = 1;                 // programmatic dependent launch between consecutive kernels
    int overlap_same_src = 0;    // n == 1: an op that reads the same x as the previous one does not wait for it (caller owns the stream)
    int host_fast = 1;           // n == 1 host entry: single-launch path with mapped staging rows and a completion word
    int set_coop = 1;            // persistent multi-op kernel: cooperative launch (the driver checks co-residency of the grid)
    int set_warps = 12;          // its consumer warps per CTA (8, 12, 16 or 20)
    int set_link = 1;            // x that is exactly the dst of one earlier op is read from that op's tagged rows (0: grid barrier instead)
    int set_fly = 0;             // bulk copies in flight per CTA (0 = as many as the ring has free slots)
    int set_split_barrier = 1;   // dependency barriers: arrive after the last conflicting op, wait at the run that needs it (0: arrive right before the wait)
    int set_barrier_slack = 1 << 20;   // a dependency barrier also waits for everything up to this many ops before its run (default: only for what the run conflicts with, i.e. arrive as early as possible)
    int set_xbufs = 2;           // quantized-x buffers in shared memory (2: the x warps may be a whole run ahead)
    int set_xext = 1;            // x that nothing in the list writes may be read before the run's barrier and through L1
    int set_table = 1;           // stage the op descriptors in shared memory when the list fits
    int set_slots = 0;           // cap on its ring depth (0 = as many as fit)
    int skip_quantize = 0;       // bench only: reuse the activations quantized by the previous call
};

struct Ctx {
    bool inited = false;
    int device = -1;
    int sm_count = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    int repack = 1;
    int gemm_min_n = 32;
    int path = GGML_B200_PATH_AUTO;
    Tune tune;
    // activation work buffer (device replacement of params->wdata, ggml.c:11354)
    int8_t * xq = nullptr;  size_t xq_cap = 0;
    int2 *   xds = nullptr; size_t xds_cap = 0;
    float *  xdT = nullptr; size_t xdT_cap = 0; long long xdT_n_pad = 0;   // scales block-major, GEMM epilogue
    // staging for the host-pointer entry point
    float * xstage = nullptr; size_t xstage_cap = 0;
    float * ystage = nullptr; size_t ystage_cap = 0;
    void *  sstage = nullptr; size_t sstage_cap = 0;
    // n = 1 host entry point without driver copies or stream syncs (gemv_q4_0_hostio.cuh)
    struct { float * hx = nullptr; size_t hx_cap = 0; float * hy = nullptr; size_t hy_cap = 0;
             float * xdev = nullptr; size_t xdev_cap = 0; float * ydev = nullptr; size_t ydev_cap = 0; unsigned * counters = nullptr;
             unsigned seq = 0; } hio;
    uint64_t launches = 0;
    // the run of n = 1 fused launches that read the same x (overlap_same_src): the ops still possibly in flight
    struct { const float * x = nullptr; int64_t K = 0; cudaStream_t stream = nullptr; uint64_t launches_after = 0;
             int count = 0; const float * y_lo[8]; const float * y_hi[8]; const ggml_b200_weight * head = nullptr; } run;
    // the same for n > 1: ops that share src1 reuse the quantized activations of the run's first op
    struct { const float * x = nullptr; int64_t K = 0, n = 0; size_t nb11 = 0; cudaStream_t stream = nullptr; uint64_t launches_after = 0;
             int count = 0; const float * y_lo[8]; const float * y_hi[8]; } runN;
    char err[512] = {0};
};

extern Ctx g;
// One process = one device = one Ctx: the public entry points that touch it (workspaces, the same-src1 run records, the
// staging rows) are serialised by this lock, so two host threads calling the C ABI cannot corrupt each other.
extern std::recursive_mutex g_api_mu;
#define B200_API_LOCK std::lock_guard<std::recursive_mutex> api_lock_(::b200::g_api_mu)
// kernels' dynamic-shared-memory attributes are per device: set again after ggml_b200_shutdown() + init on another one
extern unsigned g_init_epoch;
#define B200_SET_SMEM_ATTR(kern, bytes)                                                                    \
    do {                                                                                                   \
        static unsigned seen_ = 0;                                                                         \
        if (seen_ != ::b200::g_init_epoch) {                                                               \
            CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)));    \
            seen_ = ::b200::g_init_epoch;                                                                  \
        }                                                                                                  \
    } while (0)

void set_error(const char * fmt, ...);
int  cuda_fail(cudaError_t e, const char * what, const char * file, int line);

#define CK(call)                                                                   \
    do {                                                                           \
        cudaError_t e_ = (call);                                                   \
        if (e_ != cudaSuccess) return ::b200::cuda_fail(e_, #call, __FILE__, __LINE__); \
    } while (0)

int ensure_workspace(int64_t n, int64_t K);
int launch_quantize(const float * x_dev, int64_t n, int64_t K, size_t nb11, void * blocks34_dev);
int launch_gemv(const ggml_b200_weight * w, float * y, long long ldy, int n);
// kernel 2 or 3 by n, into y (first element of this shard's column 0), activations already quantized
int launch_shard(const ggml_b200_weight * w, float * y, long long ldy, int n);
int launch_slab(const ggml_b200_weight * w, const float * x_dev, int64_t n, size_t nb11, float * y, long long ldy);
void comm_set_p2p(int on);
// collective over the communicator: export `local` (a cudaMalloc'd block) over CUDA IPC and map every peer's block;
// mapped[me] = local.  comm_unshare closes the peers' mappings.
bool comm_p2p_ready();
int  comm_share(void * local, void ** mapped /* [P2P_MAX_WORLD] */);
void comm_unshare(void ** mapped);
// n = 1 row shard with the multi-GPU exchange fused in (gemv_q4_0_p2p.cuh); y_dev = the whole dst column
bool fused_p2p_ok(const ggml_b200_weight * w);
int launch_fused_p2p(const ggml_b200_weight * w, const float * x_dev, size_t nb11, float * y_dev, const P2PExchange & ex);
int mul_mat_dev(const ggml_b200_weight * w, const float * x_dev, int64_t n, size_t nb11, float * y_dev, size_t nb1);

// kernel 3 (gemm_q4_0_tc.cu)
int  gemm_init();
void gemm_shutdown();
int  gemm_set_tuning(const char * key, int value);
int  gemm_xd_perm();   // 1: kernel 1 writes the block-major activation scales in kernel 3's (v3) column order
bool gemm_available(const ggml_b200_weight * w);
int  gemm_prepare_weight(ggml_b200_weight * w);
void gemm_release_weight(ggml_b200_weight * w);
// y: first output element of this shard (column 0), ldy in floats; S != nullptr: also dump S_b
int  launch_gemm(const ggml_b200_weight * w, float * y, long long ldy, int n, int * S, int independent = 0);
// multi-GPU prefill: kernel 3 also stores every finished tile into the ndst peers' landing buffers (column stride ld) and
// release-adds 1 to ctr[i] per tile (comm.cu: prefill exchange)
int  launch_gemm_peers(const ggml_b200_weight * w, float * y, long long ldy, int n, float * const * dst, unsigned * const * ctr, int ndst, long long ld);
// the same kernel over a Q8_0 weight's int8 plane (siblings.cu); handle = its TMA descriptor
void * gemm_prepare_q8(const void * wq, int64_t M, int64_t K);
void gemm_release_q8(void * handle);
int  launch_gemm_q8(const void * handle, const void * wd, int64_t M, int64_t K, float * y, long long ldy, int n, int * S);

}  // namespace b200

extern "C" int ggml_b200_set_tuning(const char * key, int value);
