// ggml-b200.cu -- the C ABI declared in include/ggml-b200.h: init, load-time upload (+ repack),
// op dispatch, host-buffer path, parity dumps.  One process drives one GPU.
//
// Reference seams replaced (paths under the reference tree): ggml_init_cublas
// (ggml-cuda.cu:1826-1861), ggml_cuda_transform_tensor (:2766-2809), ggml_cuda_free_data
// (:2811-2828), ggml_cuda_mul_mat / ggml_cuda_op (:2671-2690, :2286-2567) and, behind them,
// ggml_compute_forward_mul_mat_q_f32 (ggml.c:11226-11424).  Nothing here is a port of those:
// the reference dequantizes to f32 and calls cuBLAS / a float DMMV kernel; this file quantizes
// the activations to q8_0 on the device and multiplies in the integer domain like the CPU path.
//
// There is NO CPU fallback: without a sm_100 device every compute entry point fails loudly.
#include "../../include/ggml-b200.h"

#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <mutex>

#include "common.cuh"
#include "gemv_q4_0.cuh"
#include "gemv_q4_0_fused.cuh"
#include "gemv_q4_0_p2p.cuh"
#include "gemv_q4_0_hostio.cuh"
#include "gemv_q4_0_neigh.cuh"
#include "quantize_q8_0.cuh"
#include "repack.cuh"
#include "codec_q4_0.cuh"
#include "elementwise.cuh"
#include "state.h"

namespace b200 {

Ctx g;
std::recursive_mutex g_api_mu;
unsigned g_init_epoch = 0;
static std::mutex g_mu;
static unsigned long long * h_trace_dev = nullptr;
static unsigned int h_trace_cap = 0;

void set_error(const char * fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g.err, sizeof(g.err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char * what, const char * file, int line) {
    set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    if (getenv("GGML_B200_VERBOSE")) fprintf(stderr, "ggml-b200: %s\n", g.err);
    return GGML_B200_ERR_CUDA;
}

int ensure_workspace(int64_t n, int64_t K) {
    const size_t need_q = (size_t)n * K;
    const size_t need_ds = (size_t)n * (K / QK) * sizeof(int2);
    if (need_q > g.xq_cap) {
        if (g.xq) CK(cudaFree(g.xq));
        g.xq = nullptr; g.xq_cap = 0;
        CK(cudaMalloc(&g.xq, need_q));
        g.xq_cap = need_q;
    }
    if (need_ds > g.xds_cap) {
        if (g.xds) CK(cudaFree(g.xds));
        g.xds = nullptr; g.xds_cap = 0;
        CK(cudaMalloc(&g.xds, need_ds));
        g.xds_cap = need_ds;
    }
    return GGML_B200_OK;
}

static int ensure_stage(void ** p, size_t * cap, size_t need) {
    if (need > *cap) {
        if (*p) CK(cudaFree(*p));
        *p = nullptr; *cap = 0;
        CK(cudaMalloc(p, need));
        *cap = need;
    }
    return GGML_B200_OK;
}

// every kernel goes through here: optional programmatic-dependent-launch edge to the previous
// kernel in the stream (the kernels call griddepcontrol.wait before touching activations)
template <typename... KArgs, typename... Args>
static int launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = g.stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = g.tune.pdl ? 1 : 0;
    CK(cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...));
    g.launches++;
    return GGML_B200_OK;
}

// ---- kernel 1 launch ------------------------------------------------------------------------
int launch_quantize(const float * x_dev, int64_t n, int64_t K, size_t nb11, void * blocks34_dev) {
    if (n <= 0) return GGML_B200_OK;
    int rc = ensure_workspace(n, K);
    if (rc) return rc;
    const long long total = (long long)n * (K / 8);
    const int threads = 256;
    const unsigned blocks = (unsigned)((total + threads - 1) / threads);
    const bool aligned = (((uintptr_t)x_dev | nb11) & 15) == 0;
    unsigned char * b34 = (unsigned char *)blocks34_dev;
    const long long nn = n;
    // the GEMM path also wants the scales block-major ([K/32][n_pad] floats)
    float * xdT = nullptr;
    const int perm = gemm_xd_perm();
    const long long n_pad = perm ? (n + 63) & ~63ll : (n + 3) & ~3ll;
    if (n >= 2 && g.path != GGML_B200_PATH_GEMV) {
        const size_t need = (size_t)(K / QK) * n_pad * sizeof(float);
        if (need > g.xdT_cap) {
            if (g.xdT) CK(cudaFree(g.xdT));
            g.xdT = nullptr; g.xdT_cap = 0;
            CK(cudaMalloc(&g.xdT, need));
            g.xdT_cap = need;
        }
        xdT = g.xdT;
    }
    g.xdT_n_pad = n_pad;
    if (aligned) {
        if (b34) return launch_k(quantize_q8_0_kernel<true, true>, blocks, threads, 0, x_dev, nn, (int)K, nb11, g.xq, g.xds, b34, xdT, n_pad, perm);
        return launch_k(quantize_q8_0_kernel<true, false>, blocks, threads, 0, x_dev, nn, (int)K, nb11, g.xq, g.xds, b34, xdT, n_pad, perm);
    }
    if (b34) return launch_k(quantize_q8_0_kernel<false, true>, blocks, threads, 0, x_dev, nn, (int)K, nb11, g.xq, g.xds, b34, xdT, n_pad, perm);
    return launch_k(quantize_q8_0_kernel<false, false>, blocks, threads, 0, x_dev, nn, (int)K, nb11, g.xq, g.xds, b34, xdT, n_pad, perm);
}

// ---- kernel 2 launch ------------------------------------------------------------------------
template <int NC, int UNROLL, bool DUMP>
static int launch_gemv_planes_t(const ggml_b200_weight * w, float * y, long long ldy, int n, int * S) {
    const int nbk = (int)(w->K / QK);
    const size_t smem = gemv_smem_bytes(nbk, NC);
    auto kern = gemv_q4_0_planes_kernel<NC, UNROLL, DUMP>;
    B200_SET_SMEM_ATTR(kern, MAX_DYN_SMEM);
    const int grid = g.sm_count * g.tune.gemv_ctas_per_sm;
    return launch_k(kern, grid, g.tune.gemv_threads, smem, (const uint4 *)w->wq, (const __half *)w->wd, (const int8_t *)g.xq, (const int2 *)g.xds, y, (int)w->rows(), nbk, ldy, n, S);
}

template <int NC, bool DUMP>
static int launch_gemv_raw_t(const ggml_b200_weight * w, float * y, long long ldy, int n, int * S) {
    const int nbk = (int)(w->K / QK);
    const size_t smem = gemv_smem_bytes(nbk, NC);
    auto kern = gemv_q4_0_raw_kernel<NC, DUMP>;
    B200_SET_SMEM_ATTR(kern, MAX_DYN_SMEM);
    const int grid = g.sm_count * g.tune.gemv_ctas_per_sm;
    return launch_k(kern, grid, g.tune.gemv_threads, smem, (const uint32_t *)w->raw, (const int8_t *)g.xq, (const int2 *)g.xds, y, (int)w->rows(), nbk, ldy, n, S);
}

template <bool DUMP>
static int launch_gemv_any(const ggml_b200_weight * w, float * y, long long ldy, int n, int * S) {
    const int nbk = (int)(w->K / QK);
    // columns per pass: as many as n needs (1, 2, 4, 8), capped by the knob and by shared memory
    int nc = n >= 8 ? 8 : n >= 3 ? 4 : n;
    if (g.tune.gemv_nc > 0 && nc > g.tune.gemv_nc) nc = g.tune.gemv_nc;
    while (nc > 1 && gemv_smem_bytes(nbk, nc) > MAX_DYN_SMEM) nc >>= 1;
    if (gemv_smem_bytes(nbk, nc) > MAX_DYN_SMEM) {
        set_error("K = %lld needs %zu bytes of shared memory for the staged activations (max %d)",
                  (long long)w->K, gemv_smem_bytes(nbk, 1), MAX_DYN_SMEM);
        return GGML_B200_ERR_ARG;
    }
    if (w->layout == GGML_B200_LAYOUT_RAW) {
        switch (nc) {
            case 1: return launch_gemv_raw_t<1, DUMP>(w, y, ldy, n, S);
            case 2: return launch_gemv_raw_t<2, DUMP>(w, y, ldy, n, S);
            case 4: return launch_gemv_raw_t<4, DUMP>(w, y, ldy, n, S);
            default: return launch_gemv_raw_t<8, DUMP>(w, y, ldy, n, S);
        }
    }
    const int u = g.tune.gemv_unroll;
    switch (nc) {
        case 1: return u >= 4 ? launch_gemv_planes_t<1, 4, DUMP>(w, y, ldy, n, S) : u == 2 ? launch_gemv_planes_t<1, 2, DUMP>(w, y, ldy, n, S) : launch_gemv_planes_t<1, 1, DUMP>(w, y, ldy, n, S);
        case 2: return u >= 4 ? launch_gemv_planes_t<2, 4, DUMP>(w, y, ldy, n, S) : launch_gemv_planes_t<2, 2, DUMP>(w, y, ldy, n, S);
        case 4: return u >= 4 ? launch_gemv_planes_t<4, 4, DUMP>(w, y, ldy, n, S) : launch_gemv_planes_t<4, 2, DUMP>(w, y, ldy, n, S);
        default: return launch_gemv_planes_t<8, 2, DUMP>(w, y, ldy, n, S);
    }
}

// ---- independent consecutive ops (tuning key overlap_same_src) -------------------------------------------
// A decode graph multiplies the same activation row by several weights in a row (wq/wk/wv, w1/w3).  Stream order
// makes each wait for the one before although they share nothing but the read-only x.  With the key set the CALLER
// promises that nothing but this library's ops is enqueued on the stream between two ops with the same src1
// pointer; the library then lets such an op start without waiting for its predecessor (gemv_q4_0_fused.cuh) when
// its dst overlaps neither x nor the dst of any op of the current run.
static bool ranges_overlap(const float * a_lo, const float * a_hi, const float * b_lo, const float * b_hi) { return a_lo < b_hi && b_lo < a_hi; }

static int independent_of_run(const ggml_b200_weight * w, const float * x, const float * y, const float * y_hi, int max_run = 8) {
    auto & r = g.run;
    if (!g.tune.overlap_same_src || !g.tune.pdl || r.count == 0 || r.count >= max_run) return 0;
    if (r.x != x || r.K != w->K || r.stream != g.stream || r.launches_after != g.launches) return 0;
    if (ranges_overlap(y, y_hi, x, x + w->K)) return 0;
    for (int i = 0; i < r.count; i++) if (ranges_overlap(y, y_hi, r.y_lo[i], r.y_hi[i])) return 0;
    return 1;
}

static void run_append(const ggml_b200_weight * w, const float * x, const float * y, const float * y_hi, int independent) {
    auto & r = g.run;
    if (!independent) { r.count = 0; r.x = x; r.K = w->K; r.stream = g.stream; r.head = w; }
    else { w->run_member = true; if (r.head) r.head->run_member = true; }   // next pass: both take the half-SM kernel
    // an op whose dst overlaps x ends the run: the next reader of x must wait for it
    if (ranges_overlap(y, y_hi, x, x + w->K)) { r.count = 0; r.x = nullptr; return; }
    r.y_lo[r.count] = y; r.y_hi[r.count] = y_hi; r.count++;
    r.launches_after = g.launches;
}

// n == 1, plane layout: kernel 1 fused into kernel 2 (one launch per op)
template <int UNROLL, int RIF, int MAXT, bool DUMP>
static int launch_fused_t(const ggml_b200_weight * w, const float * x, float * y, int * S, unsigned char * b34) {
    const int nbk = (int)(w->K / QK);
    auto kern = gemv_q4_0_fused_kernel<UNROLL, RIF, MAXT, DUMP>;
    B200_SET_SMEM_ATTR(kern, MAX_DYN_SMEM);
    int threads = g.tune.fused_threads > MAXT ? MAXT : g.tune.fused_threads;
    const int grid = g.sm_count * g.tune.fused_ctas_per_sm;
    const int aligned = (((uintptr_t)x) & 15) == 0;
    const int independent = (S || b34) ? 0 : independent_of_run(w, x, y, y + w->rows());
    int rc = launch_k(kern, grid, threads, gemv_smem_bytes(nbk, 1), (const uint4 *)w->wq, (const __half *)w->wd, x, y, (int)w->rows(), nbk, aligned, S, b34, (unsigned)g.launches,
                      independent);
    if (rc == GGML_B200_OK && !(S || b34)) run_append(w, x, y, y + w->rows(), independent);
    return rc;
}

bool fused_ok(const ggml_b200_weight * w, int64_t n) {
    return n == 1 && g.tune.gemv_fused && !g.tune.skip_quantize && w->layout == GGML_B200_LAYOUT_PLANES &&
           g.path != GGML_B200_PATH_GEMM && gemv_smem_bytes((int)(w->K / QK), 1) <= (size_t)MAX_DYN_SMEM;
}

int launch_fused(const ggml_b200_weight * w, const float * x, float * y, int * S, unsigned char * b34) {
    // 1024 x 1 CTA/SM is best up to ~8k rows; beyond, 512 x 2 CTAs/SM (finer row interleave) wins
    // (profiles/r01_sweep_gemv_v2.log); an explicit fused_threads knob overrides
    // members of a same-src1 run take 512 threads (half an SM) with two rows in flight per warp, so that the run's
    // kernels, which do not wait for each other, are resident together and overlap end to end
    const bool member = g.tune.overlap_same_src && g.tune.pdl && w->run_member && !(S || b34);
    const bool big = !member && g.tune.fused_threads > 512;
    const int rif = member ? 2 : g.tune.fused_rif;
    if (S || b34) return big ? launch_fused_t<4, 1, 1024, true>(w, x, y, S, b34) : rif >= 2 ? launch_fused_t<4, 2, 512, true>(w, x, y, S, b34) : launch_fused_t<4, 1, 512, true>(w, x, y, S, b34);
    if (big) return g.tune.gemv_unroll >= 4 ? launch_fused_t<4, 1, 1024, false>(w, x, y, S, b34) : launch_fused_t<2, 1, 1024, false>(w, x, y, S, b34);
    if (rif >= 4) return launch_fused_t<2, 4, 512, false>(w, x, y, S, b34);
    if (rif >= 2) return g.tune.gemv_unroll >= 4 ? launch_fused_t<4, 2, 512, false>(w, x, y, S, b34) : launch_fused_t<2, 2, 512, false>(w, x, y, S, b34);
    return g.tune.gemv_unroll >= 4 ? launch_fused_t<4, 1, 512, false>(w, x, y, S, b34) : launch_fused_t<2, 1, 512, false>(w, x, y, S, b34);
}

int launch_gemv(const ggml_b200_weight * w, float * y, long long ldy, int n) {
    return launch_gemv_any<false>(w, y, ldy, n, nullptr);
}

int launch_gemv_dump(const ggml_b200_weight * w, float * y, long long ldy, int n, int * S) {
    return launch_gemv_any<true>(w, y, ldy, n, S);
}

static bool use_gemm(const ggml_b200_weight * w, int64_t n) {
    if (g.path == GGML_B200_PATH_GEMV) return false;
    if (!gemm_available(w)) return false;
    if (g.path == GGML_B200_PATH_GEMM) return true;
    return n >= g.gemm_min_n;
}

static int check_op_args(const ggml_b200_weight * w, const void * x, int64_t n, size_t nb11, const void * y, size_t nb1) {
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!w || !x || !y) { set_error("null argument"); return GGML_B200_ERR_ARG; }
    if (n < 0 || n > INT32_MAX) { set_error("bad n"); return GGML_B200_ERR_ARG; }
    // ggml.c:11279 (nb10 == sizeof(float)) and :11282-11285 (dense dst rows)
    if (nb11 < (size_t)w->K * 4 || (nb11 & 3)) { set_error("src1 row stride %zu < K*4 or not a multiple of 4", nb11); return GGML_B200_ERR_ARG; }
    if (nb1 < (size_t)w->M * 4 || (nb1 & 3)) { set_error("dst row stride %zu < M*4 or not a multiple of 4", nb1); return GGML_B200_ERR_ARG; }
    return GGML_B200_OK;
}

int launch_shard(const ggml_b200_weight * w, float * y, long long ldy, int n) {
    if (use_gemm(w, n)) return launch_gemm(w, y, ldy, n, nullptr);
    return launch_gemv(w, y, ldy, n);
}

// n == 1 row shard, gather fused in (multi-GPU; comm.cu decides when)
bool fused_p2p_ok(const ggml_b200_weight * w) {
    return g.tune.gemv_fused && !g.tune.skip_quantize && (w->layout == GGML_B200_LAYOUT_PLANES || w->rows() == 0) && g.path != GGML_B200_PATH_GEMM &&
           gemv_p2p_smem_bytes((int)(w->K / QK), (int)w->rows(), g.sm_count) <= (size_t)MAX_DYN_SMEM - 1024;
}

template <int MAXT>
static int launch_fused_p2p_t(const ggml_b200_weight * w, const float * x_dev, float * y_dev, const P2PExchange & ex, int run_index) {
    const int nbk = (int)(w->K / QK);
    auto kern = gemv_q4_0_fused_p2p_kernel<4, MAXT>;
    B200_SET_SMEM_ATTR(kern, MAX_DYN_SMEM - 1024);
    const int aligned = (((uintptr_t)x_dev) & 15) == 0;
    return launch_k(kern, g.sm_count, MAXT, gemv_p2p_smem_bytes(nbk, (int)w->rows(), g.sm_count), (const uint4 *)w->wq, (const __half *)w->wd, x_dev, y_dev,
                    (int)w->rows(), nbk, aligned, (unsigned)g.launches, ex, run_index);
}

int launch_fused_p2p(const ggml_b200_weight * w, const float * x_dev, size_t nb11, float * y_dev, const P2PExchange & ex) {
    (void)nb11;   // n == 1: a single activation row
    // ops that only share src1 with their predecessors (wq/wk/wv, w3/w1) neither wait for them nor for their
    // exchange: position in the run = op-number offset (p2p.cuh); members of a run take the half-SM variant so
    // that the next op computes while this one waits for its peers
    const int independent = independent_of_run(w, x_dev, y_dev, y_dev + w->M, P2P_MAX_RUN);
    const int run_index = independent ? g.run.count : 0;
    const bool member = g.tune.overlap_same_src && g.tune.pdl && w->run_member;
    const int rc = member ? launch_fused_p2p_t<512>(w, x_dev, y_dev, ex, run_index) : launch_fused_p2p_t<1024>(w, x_dev, y_dev, ex, run_index);
    if (rc == GGML_B200_OK) run_append(w, x_dev, y_dev, y_dev + w->M, independent);
    return rc;
}

// this shard's slab of dst, dense: y = first element of column 0, ldy floats between columns
int launch_slab(const ggml_b200_weight * w, const float * x_dev, int64_t n, size_t nb11, float * y, long long ldy) {
    if (n == 0 || w->rows() == 0) return GGML_B200_OK;
    if (fused_ok(w, n)) return launch_fused(w, x_dev, y, nullptr, nullptr);
    if (!g.tune.skip_quantize) {
        int rc = launch_quantize(x_dev, n, w->K, nb11, nullptr);
        if (rc) return rc;
    }
    return launch_shard(w, y, ldy, (int)n);
}

// x_dev/y_dev device pointers; y_dev addresses the full dst, this shard writes [row_lo,row_hi)
int mul_mat_dev(const ggml_b200_weight * w, const float * x_dev, int64_t n, size_t nb11, float * y_dev, size_t nb1) {
    if (n == 0 || w->rows() == 0) return GGML_B200_OK;
    if (fused_ok(w, n)) return launch_fused(w, x_dev, y_dev + w->row_lo, nullptr, nullptr);
    float * y = y_dev + w->row_lo;
    const long long ldy = (long long)(nb1 / 4);
    // n > 1, same src1 as the previous op of this library on this stream, dst disjoint (overlap_same_src contract,
    // see independent_of_run): kernel 1 already ran for this x -- reuse its output, and let the GEMM start without
    // waiting for its predecessor
    const float * y_hi = y + (n - 1) * ldy + w->rows();
    auto & r = g.runN;
    bool same = g.tune.overlap_same_src && g.tune.pdl && r.count > 0 && r.count < 8 && r.x == x_dev && r.n == n && r.K == w->K && r.nb11 == nb11 &&
                r.stream == g.stream && r.launches_after == g.launches && !ranges_overlap(y, y_hi, x_dev, (const float *)((const char *)x_dev + n * nb11));
    for (int i = 0; same && i < r.count; i++) same = !ranges_overlap(y, y_hi, r.y_lo[i], r.y_hi[i]);
    if (!same && !g.tune.skip_quantize) {
        int rc = launch_quantize(x_dev, n, w->K, nb11, nullptr);
        if (rc) return rc;
    }
    const int rc = use_gemm(w, n) ? launch_gemm(w, y, ldy, (int)n, nullptr, same ? 1 : 0) : launch_gemv(w, y, ldy, (int)n);
    if (rc) return rc;
    if (!same) { r.count = 0; r.x = x_dev; r.n = n; r.K = w->K; r.nb11 = nb11; r.stream = g.stream; }
    if (ranges_overlap(y, y_hi, x_dev, (const float *)((const char *)x_dev + n * nb11))) { r.count = 0; r.x = nullptr; return GGML_B200_OK; }
    r.y_lo[r.count] = y; r.y_hi[r.count] = y_hi; r.count++;
    r.launches_after = g.launches;
    return GGML_B200_OK;
}

}  // namespace b200

using namespace b200;

// =============================================================================================
extern "C" {

int ggml_b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

const char * ggml_b200_last_error(void) { return g.err; }
int ggml_b200_sm_count(void) { return g.sm_count; }

int ggml_b200_init(int device) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (g.inited && g.device == device) return GGML_B200_OK;
    if (g.inited) { set_error("already initialised on device %d", g.device); return GGML_B200_ERR_STATE; }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0) {
        cudaGetLastError();
        set_error("no CUDA device (%s); this library has no CPU fallback", e == cudaSuccess ? "count 0" : cudaGetErrorString(e));
        return GGML_B200_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= ndev) { set_error("device %d out of range (0..%d)", device, ndev - 1); return GGML_B200_ERR_ARG; }
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        set_error("device %d is sm_%d%d; this library contains sm_100a code only", device, prop.major, prop.minor);
        return GGML_B200_ERR_ARCH;
    }
    CK(cudaSetDevice(device));
    g.device = device;
    g.sm_count = prop.multiProcessorCount;
    CK(cudaStreamCreateWithFlags(&g.own_stream, cudaStreamNonBlocking));
    g.stream = g.own_stream;
    const char * s;
    if ((s = getenv("GGML_B200_REPACK")))     g.repack = atoi(s) != 0;
    if ((s = getenv("GGML_B200_GEMM_MIN_N"))) g.gemm_min_n = atoi(s);
    if ((s = getenv("GGML_B200_PATH")))       g.path = !strcmp(s, "gemv") ? GGML_B200_PATH_GEMV : !strcmp(s, "gemm") ? GGML_B200_PATH_GEMM : GGML_B200_PATH_AUTO;
    int rc = gemm_init();
    if (rc) return rc;
    g_init_epoch++;
    g.inited = true;
    return GGML_B200_OK;
}

// GGML_B200_VERBOSE: where the host path's time goes (ns totals), printed at shutdown
static struct { uint64_t ops, copy_in, launch, spin, copy_out; } g_hio_stats = {};
static void hostio_release();

static cudaEvent_t g_marks[GGML_B200_MARKS] = {};      // ggml_b200_mark / _wait_mark
void ggml_b200_shutdown(void) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g.inited) return;
    ggml_b200_comm_shutdown();
    cudaStreamSynchronize(g.stream);
    if (g.xq) cudaFree(g.xq);
    if (g.xds) cudaFree(g.xds);
    if (g.xdT) cudaFree(g.xdT);
    if (g.xstage) cudaFree(g.xstage);
    if (g.ystage) cudaFree(g.ystage);
    if (g.sstage) cudaFree(g.sstage);
    if (getenv("GGML_B200_VERBOSE") && g_hio_stats.ops) {
        const double n = (double)g_hio_stats.ops * 1e3;
        unsigned long long prof[4] = {0, 0, 0, 0};
        if (g.hio.counters) cudaMemcpy(prof, (const char *)g.hio.counters + 16, sizeof(prof), cudaMemcpyDeviceToHost);
        fprintf(stderr, "ggml_b200: host path kernel: start -> x published %.2f us (mean), -> y stores issued %.2f us (max)\n", prof[1] / n, prof[2] / 1e3);
        fprintf(stderr, "ggml_b200: host path, %llu ops: copy-in %.2f us, launch %.2f us, wait for the first 4 rows %.2f us, rest of the copy-out %.2f us per op\n",
                (unsigned long long)g_hio_stats.ops, g_hio_stats.copy_in / n, g_hio_stats.launch / n, g_hio_stats.spin / n, g_hio_stats.copy_out / n);
    }
    for (int i = 0; i < GGML_B200_MARKS; i++) if (g_marks[i]) { cudaEventDestroy(g_marks[i]); g_marks[i] = nullptr; }
    hostio_release();
    gemm_shutdown();
    if (g.own_stream) cudaStreamDestroy(g.own_stream);
    const Tune t = g.tune;
    g = Ctx();
    g.tune = t;
}

void   ggml_b200_set_repack(int on) { g.repack = on != 0; }
int    ggml_b200_get_repack(void) { return g.repack; }
void   ggml_b200_set_gemm_min_n(int n) { g.gemm_min_n = n; }
int    ggml_b200_get_gemm_min_n(void) { return g.gemm_min_n; }
void   ggml_b200_set_path(int path) { g.path = path; }
void   ggml_b200_set_stream(void * s) { g.stream = s ? (cudaStream_t)s : g.own_stream; }
void * ggml_b200_get_stream(void) { return (void *)g.stream; }
uint64_t ggml_b200_launch_count(void) { return g.launches; }

int ggml_b200_set_tuning(const char * key, int value) {
    B200_API_LOCK;
    if (!strcmp(key, "gemv_threads"))      { if (value < 32 || value > 512 || (value & 31)) return GGML_B200_ERR_ARG; g.tune.gemv_threads = value; }
    else if (!strcmp(key, "gemv_ctas_per_sm")) { if (value < 1 || value > 16) return GGML_B200_ERR_ARG; g.tune.gemv_ctas_per_sm = value; }
    else if (!strcmp(key, "gemv_unroll"))  { g.tune.gemv_unroll = value; }
    else if (!strcmp(key, "gemv_nc"))      { g.tune.gemv_nc = value; }
    else if (!strcmp(key, "fused_rif"))    { g.tune.fused_rif = value; }
    else if (!strcmp(key, "gemv_fused"))   { g.tune.gemv_fused = value != 0; }
    else if (!strcmp(key, "pdl"))          { g.tune.pdl = value != 0; }
    else if (!strcmp(key, "fused_threads")) { if (value < 32 || value > 1024 || (value & 31)) return GGML_B200_ERR_ARG; g.tune.fused_threads = value; }
    else if (!strcmp(key, "small_copy"))   { g.tune.small_copy = value != 0; }
    else if (!strcmp(key, "fused_ctas_per_sm")) { if (value < 1 || value > 16) return GGML_B200_ERR_ARG; g.tune.fused_ctas_per_sm = value; }
    else if (!strcmp(key, "skip_quantize")) { g.tune.skip_quantize = value != 0; }
    else if (!strcmp(key, "host_fast"))    { g.tune.host_fast = value != 0; }
    else if (!strcmp(key, "overlap_same_src")) { g.tune.overlap_same_src = value != 0; g.run.count = 0; g.runN.count = 0; }
    else if (!strcmp(key, "set_coop"))     { g.tune.set_coop = value != 0; }
    else if (!strcmp(key, "set_warps"))    { if (value != 8 && value != 12 && value != 16 && value != 20) return GGML_B200_ERR_ARG; g.tune.set_warps = value; }
    else if (!strcmp(key, "set_link"))     { g.tune.set_link = value != 0; }
    else if (!strcmp(key, "set_fly"))      { g.tune.set_fly = value; }
    else if (!strcmp(key, "set_xext"))     { g.tune.set_xext = value != 0; }
    else if (!strcmp(key, "set_barrier_slack")) { g.tune.set_barrier_slack = value; }
    else if (!strcmp(key, "set_xbufs"))    { if (value < 1 || value > 2) return GGML_B200_ERR_ARG; g.tune.set_xbufs = value; }
    else if (!strcmp(key, "set_split_barrier")) { g.tune.set_split_barrier = value != 0; }
    else if (!strcmp(key, "set_table"))    { g.tune.set_table = value != 0; }
    else if (!strcmp(key, "set_slots"))    { g.tune.set_slots = value; }
    else if (!strcmp(key, "p2p_gather"))   { comm_set_p2p(value); }
    else return gemm_set_tuning(key, value);
    return GGML_B200_OK;
}

// pinned blocks handed out here are mapped into the device's address space (unified addressing): the asynchronous
// copies below move small rows to / from them with a kernel instead of a copy-engine transfer (see small_copy)
struct PinnedRange { const char * base; size_t size; };
static PinnedRange g_pinned[64];
static int g_pinned_n = 0;
static std::mutex g_pinned_mu;
void * ggml_b200_host_malloc(size_t size) {
    void * p = nullptr;
    if (cudaMallocHost(&p, size) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    std::lock_guard<std::mutex> lk(g_pinned_mu);
    if (g_pinned_n < 64) g_pinned[g_pinned_n++] = { (const char *)p, size };
    return p;
}
void ggml_b200_host_free(void * p) {
    if (!p) return;
    {
        std::lock_guard<std::mutex> lk(g_pinned_mu);
        for (int i = 0; i < g_pinned_n; i++) if (g_pinned[i].base == (const char *)p) { g_pinned[i] = g_pinned[--g_pinned_n]; break; }
    }
    cudaFreeHost(p);
}
static bool in_pinned_block(const void * p, size_t n) {
    std::lock_guard<std::mutex> lk(g_pinned_mu);
    for (int i = 0; i < g_pinned_n; i++)
        if ((const char *)p >= g_pinned[i].base && (const char *)p + n <= g_pinned[i].base + g_pinned[i].size) return true;
    return false;
}

void * ggml_b200_dev_malloc(size_t size) {
    void * p = nullptr;
    if (cudaMalloc(&p, size ? size : 1) != cudaSuccess) { cuda_fail(cudaGetLastError(), "cudaMalloc", __FILE__, __LINE__); return nullptr; }
    return p;
}
void ggml_b200_dev_free(void * p) { if (p) cudaFree(p); }
int ggml_b200_memcpy_h2d(void * d, const void * s, size_t n) { CK(cudaMemcpyAsync(d, s, n, cudaMemcpyHostToDevice, g.stream)); CK(cudaStreamSynchronize(g.stream)); return GGML_B200_OK; }
int ggml_b200_memcpy_d2h(void * d, const void * s, size_t n) { CK(cudaMemcpyAsync(d, s, n, cudaMemcpyDeviceToHost, g.stream)); CK(cudaStreamSynchronize(g.stream)); return GGML_B200_OK; }
int ggml_b200_synchronize(void) { CK(cudaStreamSynchronize(g.stream)); return GGML_B200_OK; }
// a few marks on the stream: wait for what was queued up to a mark although more has been queued behind it
int ggml_b200_mark(int slot) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (slot < 0 || slot >= GGML_B200_MARKS) { set_error("mark: bad slot"); return GGML_B200_ERR_ARG; }
    if (!g_marks[slot]) CK(cudaEventCreateWithFlags(&g_marks[slot], cudaEventDisableTiming));
    CK(cudaEventRecord(g_marks[slot], g.stream));
    return GGML_B200_OK;
}
int ggml_b200_wait_mark(int slot) {
    if (slot < 0 || slot >= GGML_B200_MARKS || !g_marks[slot]) { set_error("wait_mark: no such mark"); return GGML_B200_ERR_ARG; }
    CK(cudaEventSynchronize(g_marks[slot]));
    return GGML_B200_OK;
}

// ---- upload -----------------------------------------------------------------------------------
ggml_b200_weight * ggml_b200_upload_q4_0(const void * blocks, int64_t M, int64_t K, size_t nb1, int64_t row_lo, int64_t row_hi) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return nullptr; }
    // ggml.c:2343-2344: n % 32 == 0 and an even number of blocks per row
    if (!blocks || M <= 0 || K <= 0 || (K % 64) != 0 || K > INT32_MAX) { set_error("bad shape M=%lld K=%lld (K must be a positive multiple of 64)", (long long)M, (long long)K); return nullptr; }
    const int64_t nbk = K / QK;
    const size_t row_bytes = (size_t)nbk * Q4_0_BLOCK_BYTES;
    if (nb1 < row_bytes) { set_error("row stride %zu < %zu", nb1, row_bytes); return nullptr; }
    if (row_lo < 0 || row_hi < row_lo || row_hi > M) { set_error("bad row range [%lld,%lld) of %lld", (long long)row_lo, (long long)row_hi, (long long)M); return nullptr; }

    ggml_b200_weight * w = new ggml_b200_weight();
    w->M = M; w->K = K; w->row_lo = row_lo; w->row_hi = row_hi;
    w->layout = g.repack ? GGML_B200_LAYOUT_PLANES : GGML_B200_LAYOUT_RAW;
    const int64_t rows = row_hi - row_lo;
    const size_t raw_bytes = (size_t)rows * row_bytes;
    if (rows == 0) return w;

    void * raw = nullptr;
    cudaError_t e = cudaMalloc(&raw, raw_bytes);
    if (e != cudaSuccess) { cuda_fail(e, "cudaMalloc(raw)", __FILE__, __LINE__); delete w; return nullptr; }
    const char * src = (const char *)blocks + (size_t)row_lo * nb1;
    // blocking copy like the reference (ggml-cuda.cu:2797): the caller may free `blocks` on return
    e = cudaMemcpy2DAsync(raw, row_bytes, src, nb1, row_bytes, (size_t)rows, cudaMemcpyDefault, g.stream);   // host or device source
    if (e == cudaSuccess) e = cudaStreamSynchronize(g.stream);
    if (e != cudaSuccess) { cuda_fail(e, "upload copy", __FILE__, __LINE__); cudaFree(raw); delete w; return nullptr; }

    if (w->layout == GGML_B200_LAYOUT_RAW) {
        w->raw = raw;
        w->dev_bytes = raw_bytes;
        return w;
    }
    const long long nblocks = (long long)rows * nbk;
    e = cudaMalloc(&w->wq, (size_t)nblocks * sizeof(uint4));
    if (e == cudaSuccess) e = cudaMalloc(&w->wd, (size_t)nblocks * sizeof(__half) + 16);
    if (e != cudaSuccess) { cuda_fail(e, "cudaMalloc(planes)", __FILE__, __LINE__); cudaFree(raw); ggml_b200_free_weight(w); return nullptr; }
    const int threads = 256;
    repack_q4_0_to_planes_kernel<<<(unsigned)((nblocks + threads - 1) / threads), threads, 0, g.stream>>>((const unsigned short *)raw, w->wq, w->wd, nblocks);
    g.launches++;
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(g.stream);
    cudaFree(raw);
    if (e != cudaSuccess) { cuda_fail(e, "repack", __FILE__, __LINE__); ggml_b200_free_weight(w); return nullptr; }
    w->dev_bytes = (size_t)nblocks * (sizeof(uint4) + sizeof(__half));
    if (gemm_prepare_weight(w) != GGML_B200_OK) { ggml_b200_free_weight(w); return nullptr; }
    return w;
}

void ggml_b200_free_weight(ggml_b200_weight * w) {
    B200_API_LOCK;
    if (!w) return;
    if (g.run.head == w) { g.run.head = nullptr; g.run.count = 0; }
    gemm_release_weight(w);
    if (w->raw) cudaFree(w->raw);
    if (w->wq) cudaFree(w->wq);
    if (w->wd) cudaFree(w->wd);
    delete w;
}

int64_t ggml_b200_weight_rows(const ggml_b200_weight * w) { return w ? w->rows() : 0; }
int64_t ggml_b200_weight_row_lo(const ggml_b200_weight * w) { return w ? w->row_lo : 0; }
int64_t ggml_b200_weight_M(const ggml_b200_weight * w) { return w ? w->M : 0; }
int64_t ggml_b200_weight_K(const ggml_b200_weight * w) { return w ? w->K : 0; }
int     ggml_b200_weight_layout(const ggml_b200_weight * w) { return w ? w->layout : -1; }
size_t  ggml_b200_weight_device_bytes(const ggml_b200_weight * w) { return w ? w->dev_bytes : 0; }

int ggml_b200_download_q4_0(const ggml_b200_weight * w, void * out) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!w || !out) { set_error("null argument"); return GGML_B200_ERR_ARG; }
    const int64_t rows = w->rows();
    if (rows == 0) return GGML_B200_OK;
    const long long nblocks = (long long)rows * (w->K / QK);
    const size_t bytes = (size_t)nblocks * Q4_0_BLOCK_BYTES;
    if (w->layout == GGML_B200_LAYOUT_RAW) {
        CK(cudaMemcpyAsync(out, w->raw, bytes, cudaMemcpyDeviceToHost, g.stream));
        CK(cudaStreamSynchronize(g.stream));
        return GGML_B200_OK;
    }
    void * raw = nullptr;
    CK(cudaMalloc(&raw, bytes));
    const int threads = 256;
    unpack_planes_to_q4_0_kernel<<<(unsigned)((nblocks + threads - 1) / threads), threads, 0, g.stream>>>(w->wq, w->wd, (unsigned short *)raw, nblocks);
    g.launches++;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, raw, bytes, cudaMemcpyDeviceToHost, g.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(g.stream);
    cudaFree(raw);
    if (e != cudaSuccess) return cuda_fail(e, "download", __FILE__, __LINE__);
    return GGML_B200_OK;
}

// ---- weight-side codec (codec_q4_0.cuh) -----------------------------------------------------------
static unsigned codec_grid(long long threads_needed) {
    long long blocks = (threads_needed + 255) / 256;
    const long long cap = (long long)g.sm_count * 16;
    return (unsigned)(blocks < 1 ? 1 : blocks > cap ? cap : blocks);
}

int ggml_b200_quantize_q4_0(const float * x_dev, int64_t nrows, int64_t K, size_t nb1, void * blocks_dev) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!x_dev || !blocks_dev || nrows < 0 || K <= 0 || (K % QK) != 0 || K > INT32_MAX || nb1 < (size_t)K * 4 || (nb1 & 3)) { set_error("bad quantize_q4_0 arguments"); return GGML_B200_ERR_ARG; }
    if (nrows == 0) return GGML_B200_OK;
    quantize_q4_0_kernel<<<codec_grid(nrows * (K / QK) * 4), 256, 0, g.stream>>>(x_dev, nrows, (int)K, nb1, (unsigned short *)blocks_dev);
    g.launches++;
    CK(cudaGetLastError());
    return GGML_B200_OK;
}

int ggml_b200_dequantize_q4_0(const void * blocks_dev, int64_t nblocks, float * y_dev) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!blocks_dev || !y_dev || nblocks < 0) { set_error("bad dequantize_q4_0 arguments"); return GGML_B200_ERR_ARG; }
    if (nblocks == 0) return GGML_B200_OK;
    dequantize_q4_0_kernel<<<codec_grid(nblocks * 8), 256, 0, g.stream>>>((const unsigned short *)blocks_dev, nblocks, y_dev);
    g.launches++;
    CK(cudaGetLastError());
    return GGML_B200_OK;
}

int ggml_b200_get_rows_q4_0(const ggml_b200_weight * w, const int32_t * ids_dev, int64_t n, float * y_dev, size_t nb1) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!w || !ids_dev || !y_dev || n < 0 || nb1 < (size_t)w->K * 4 || (nb1 & 3)) { set_error("bad get_rows arguments"); return GGML_B200_ERR_ARG; }
    if (n == 0) return GGML_B200_OK;
    int rc = ensure_stage(&g.sstage, &g.sstage_cap, 16);
    if (rc) return rc;
    int * bad = (int *)g.sstage;
    CK(cudaMemsetAsync(bad, 0, 4, g.stream));
    const int nbk = (int)(w->K / QK);
    const unsigned grid = codec_grid(n * nbk * 8);
    if (w->layout == GGML_B200_LAYOUT_PLANES)
        get_rows_q4_0_kernel<true><<<grid, 256, 0, g.stream>>>(nullptr, w->wq, w->wd, ids_dev, n, nbk, w->row_lo, w->rows(), y_dev, (long long)(nb1 / 4), bad);
    else
        get_rows_q4_0_kernel<false><<<grid, 256, 0, g.stream>>>((const unsigned short *)w->raw, nullptr, nullptr, ids_dev, n, nbk, w->row_lo, w->rows(), y_dev, (long long)(nb1 / 4), bad);
    g.launches++;
    CK(cudaGetLastError());
    int h_bad = 0;
    CK(cudaMemcpyAsync(&h_bad, bad, 4, cudaMemcpyDeviceToHost, g.stream));
    CK(cudaStreamSynchronize(g.stream));
    if (h_bad) { set_error("get_rows: a row id is outside this shard [%lld,%lld)", (long long)w->row_lo, (long long)w->row_hi); return GGML_B200_ERR_ARG; }
    return GGML_B200_OK;
}

// ---- the element-wise neighbours of the op (elementwise.cuh) ----------------------------------------
static int eltwise_check(const void * a, const void * y, int64_t n) {
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!a || !y || n < 0) { set_error("bad element-wise arguments"); return GGML_B200_ERR_ARG; }
    return GGML_B200_OK;
}
static unsigned eltwise_grid(int64_t n) {
    const int64_t b = (n + 255) / 256;
    return (unsigned)(b < 1 ? 1 : b > (int64_t)g.sm_count * 8 ? (int64_t)g.sm_count * 8 : b);
}
// rows of up to 16384 floats, everything 16-byte aligned, weights one row or one per element: the register-resident kernel
static int rms_norm_launch(const float * x, float * y, int64_t ncols, int64_t nrows, float eps, const float * w, float * y2, int64_t nw) {
    const int wmode = w == nullptr ? 0 : nw == ncols ? 1 : nw == ncols * nrows ? 2 : -1;
    const bool aligned = ((((uintptr_t)x) | ((uintptr_t)y) | ((uintptr_t)w) | ((uintptr_t)y2)) & 15) == 0;
    if (wmode >= 0 && aligned && (ncols & 3) == 0 && ncols <= 16384) {
        const int threads = (int)(((ncols / 4) + 31) / 32 * 32 > 1024 ? 1024 : ((ncols / 4) + 31) / 32 * 32);
        if (ncols <= 4096)
            return launch_k(rms_norm_f32_row_kernel<1>, dim3((unsigned)nrows), dim3(threads), 0, x, y, (int)ncols, eps, w, y2, wmode, (unsigned)g.launches);
        if (ncols <= 8192)
            return launch_k(rms_norm_f32_row_kernel<2>, dim3((unsigned)nrows), dim3(threads), 0, x, y, (int)ncols, eps, w, y2, wmode, (unsigned)g.launches);
        return launch_k(rms_norm_f32_row_kernel<4>, dim3((unsigned)nrows), dim3(threads), 0, x, y, (int)ncols, eps, w, y2, wmode, (unsigned)g.launches);
    }
    return launch_k(rms_norm_f32_kernel, dim3((unsigned)nrows), dim3(256), 0, x, y, (int)ncols, eps, w, y2, (int)(w ? nw : 1), (unsigned)g.launches);
}
int ggml_b200_rms_norm_f32(const float * x_dev, float * y_dev, int64_t ncols, int64_t nrows, float eps) {
    B200_API_LOCK;
    int rc = eltwise_check(x_dev, y_dev, ncols);
    if (rc) return rc;
    if (nrows <= 0 || ncols == 0) return GGML_B200_OK;
    if (ncols > INT32_MAX || nrows > INT32_MAX) { set_error("rms_norm: shape too large"); return GGML_B200_ERR_ARG; }
    return rms_norm_launch(x_dev, y_dev, ncols, nrows, eps, nullptr, nullptr, 1);
}
int ggml_b200_rms_norm_mul_f32(const float * x_dev, float * y_norm_dev, const float * w_dev, float * y_dev, int64_t ncols, int64_t nrows, int64_t nw, float eps) {
    B200_API_LOCK;
    int rc = eltwise_check(x_dev, y_dev, ncols);
    if (rc) return rc;
    if (!y_norm_dev || !w_dev || nw <= 0 || (ncols * nrows) % nw) { set_error("rms_norm_mul: bad arguments"); return GGML_B200_ERR_ARG; }
    if (nrows <= 0 || ncols == 0) return GGML_B200_OK;
    if (ncols > INT32_MAX || nrows > INT32_MAX || nw > INT32_MAX) { set_error("rms_norm_mul: shape too large"); return GGML_B200_ERR_ARG; }
    return rms_norm_launch(x_dev, y_norm_dev, ncols, nrows, eps, w_dev, y_dev, nw);
}
int ggml_b200_rope_f32(const float * x_dev, float * y_dev, const float * cos_sin_dev, int64_t ne0, int64_t rows_per_pos, int64_t npos) {
    B200_API_LOCK;
    const int64_t n = ne0 * rows_per_pos * npos;
    int rc = eltwise_check(x_dev, y_dev, n);
    if (rc) return rc;
    if (!cos_sin_dev || ne0 <= 0 || (ne0 & 1) || rows_per_pos <= 0 || npos < 0 || ne0 > INT32_MAX || rows_per_pos > INT32_MAX ||
        ((((uintptr_t)x_dev) | ((uintptr_t)y_dev) | ((uintptr_t)cos_sin_dev)) & 7)) { set_error("rope: bad arguments"); return GGML_B200_ERR_ARG; }
    if (n == 0) return GGML_B200_OK;
    return launch_k(rope_f32_kernel, dim3(eltwise_grid(n / 2)), dim3(256), 0, x_dev, y_dev, (const float2 *)cos_sin_dev, (int)ne0, (int)rows_per_pos, (long long)(n / 2),
                    (unsigned)g.launches);
}
int ggml_b200_mul_f32(const float * a_dev, const float * b_dev, float * y_dev, int64_t n, int64_t nb) {
    B200_API_LOCK;
    int rc = eltwise_check(a_dev, y_dev, n);
    if (rc) return rc;
    if (!b_dev || nb <= 0 || n % nb) { set_error("mul: src1 must tile src0"); return GGML_B200_ERR_ARG; }
    if (n == 0) return GGML_B200_OK;
    return launch_k(eltwise_f32_kernel<0>, dim3(eltwise_grid(n)), dim3(256), 0, a_dev, b_dev, y_dev, (long long)n, (long long)nb, (float *)nullptr, (unsigned)g.launches);
}
int ggml_b200_add_f32(const float * a_dev, const float * b_dev, float * y_dev, int64_t n, int64_t nb) {
    B200_API_LOCK;
    int rc = eltwise_check(a_dev, y_dev, n);
    if (rc) return rc;
    if (!b_dev || nb <= 0 || n % nb) { set_error("add: src1 must tile src0"); return GGML_B200_ERR_ARG; }
    if (n == 0) return GGML_B200_OK;
    return launch_k(eltwise_f32_kernel<1>, dim3(eltwise_grid(n)), dim3(256), 0, a_dev, b_dev, y_dev, (long long)n, (long long)nb, (float *)nullptr, (unsigned)g.launches);
}
int ggml_b200_silu_f32(const float * x_dev, float * y_dev, int64_t n) {
    B200_API_LOCK;
    int rc = eltwise_check(x_dev, y_dev, n);
    if (rc) return rc;
    if (n == 0) return GGML_B200_OK;
    return launch_k(eltwise_f32_kernel<2>, dim3(eltwise_grid(n)), dim3(256), 0, x_dev, (const float *)nullptr, y_dev, (long long)n, (long long)n, (float *)nullptr, (unsigned)g.launches);
}
int ggml_b200_silu_mul_f32(const float * x_dev, const float * b_dev, float * y_dev, int64_t n, float * silu_dev) {
    B200_API_LOCK;
    int rc = eltwise_check(x_dev, y_dev, n);
    if (rc) return rc;
    if (!b_dev) { set_error("silu_mul: null src1"); return GGML_B200_ERR_ARG; }
    if (n == 0) return GGML_B200_OK;
    return launch_k(eltwise_f32_kernel<3>, dim3(eltwise_grid(n)), dim3(256), 0, x_dev, b_dev, y_dev, (long long)n, (long long)n, silu_dev, (unsigned)g.launches);
}

// ---- a recorded sequence of this library's launches and copies as ONE submission (the glue replays the device nodes
// between two CPU nodes of a ggml graph this way: the host then pays one call instead of one per node) -------------
struct GraphExec { cudaGraphExec_t exec; uint64_t kernels; };
static uint64_t g_capture_launches0 = 0;
int ggml_b200_graph_begin(void) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    g_capture_launches0 = g.launches;
    CK(cudaStreamBeginCapture(g.stream, cudaStreamCaptureModeThreadLocal));
    return GGML_B200_OK;
}
int ggml_b200_graph_end(void ** exec_out) {
    B200_API_LOCK;
    cudaGraph_t graph = nullptr;
    CK(cudaStreamEndCapture(g.stream, &graph));
    GraphExec * ge = new GraphExec();
    ge->kernels = g.launches - g_capture_launches0;
    g.launches = g_capture_launches0;                 /* counted when the sequence actually runs */
    cudaError_t e = cudaGraphInstantiate(&ge->exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) { delete ge; return cuda_fail(e, "cudaGraphInstantiate", __FILE__, __LINE__); }
    *exec_out = ge;
    return GGML_B200_OK;
}
int ggml_b200_graph_launch(void * exec) {
    B200_API_LOCK;
    GraphExec * ge = (GraphExec *)exec;
    if (!ge) { set_error("null graph"); return GGML_B200_ERR_ARG; }
    CK(cudaGraphLaunch(ge->exec, g.stream));
    g.launches += ge->kernels;
    return GGML_B200_OK;
}
void ggml_b200_graph_free(void * exec) {
    GraphExec * ge = (GraphExec *)exec;
    if (!ge) return;
    cudaGraphExecDestroy(ge->exec);
    delete ge;
}
// A decode graph moves rows of 16-64 KB between the host's compute buffer and HBM at every CPU node (q/k/v down, the
// attention result up).  As copy-engine transfers each costs 8-11 us on the stream's critical path
// (profiles/r02_trace_seam_before.log: the idle time between wq and wk); when the host side lies in a block from
// ggml_b200_host_malloc (pinned AND mapped) a kernel moves the row over PCIe with 16-byte accesses instead -- posted
// writes on the way down, a handful of read round trips on the way up -- and takes the programmatic launch edge like
// every other kernel here.  Larger or unaligned copies, and pageable host memory, take cudaMemcpyAsync.
constexpr size_t SMALL_COPY_MAX = (size_t)256 << 10;
static bool small_copy_ok(const void * host, const void * dev, size_t n) {
    return g.tune.small_copy && n > 0 && n <= SMALL_COPY_MAX && (n & 15) == 0 && ((((uintptr_t)host) | ((uintptr_t)dev)) & 15) == 0 && in_pinned_block(host, n);
}
static int small_copy(void * d, const void * s, size_t n) {
    const int n16 = (int)(n / 16);
    return launch_k(copy_rows_kernel, dim3((unsigned)((n16 + 127) / 128)), dim3(128), 0, (uint4 *)d, (const uint4 *)s, n16, (unsigned)g.launches);
}
int ggml_b200_memcpy_h2d_async(void * d, const void * s, size_t n) {
    B200_API_LOCK;
    if (small_copy_ok(s, d, n)) return small_copy(d, s, n);
    CK(cudaMemcpyAsync(d, s, n, cudaMemcpyHostToDevice, g.stream)); return GGML_B200_OK;
}
int ggml_b200_memcpy_d2h_async(void * d, const void * s, size_t n) {
    B200_API_LOCK;
    if (small_copy_ok(d, s, n)) return small_copy(d, s, n);
    CK(cudaMemcpyAsync(d, s, n, cudaMemcpyDeviceToHost, g.stream)); return GGML_B200_OK;
}

// ---- kernel 1 ---------------------------------------------------------------------------------
int ggml_b200_quantize_q8_0(const float * x_dev, int64_t n, int64_t K, size_t nb11, void * blocks34_dev) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!x_dev || n < 0 || K <= 0 || (K % QK) != 0 || nb11 < (size_t)K * 4 || (nb11 & 3)) { set_error("bad quantize arguments"); return GGML_B200_ERR_ARG; }
    return launch_quantize(x_dev, n, K, nb11, blocks34_dev);
}

// ---- the op -----------------------------------------------------------------------------------
int ggml_b200_mul_mat_q4_0(const ggml_b200_weight * w, const float * x_dev, int64_t n, size_t nb11, float * y_dev, size_t nb1) {
    B200_API_LOCK;
    int rc = check_op_args(w, x_dev, n, nb11, y_dev, nb1);
    if (rc) return rc;
    return mul_mat_dev(w, x_dev, n, nb11, y_dev, nb1);
}

// ---- the op with its immediate neighbour folded into the store (gemv_q4_0_neigh.cuh) -------------
// n = 1, this GPU holds every row of the weight, plane layout; anything else is GGML_B200_ERR_ARG (callers ask
// ggml_b200_neigh_ok first and issue the nodes one by one otherwise).
int ggml_b200_neigh_ok(const ggml_b200_weight * w) {
    B200_API_LOCK;
    return g.inited && w != nullptr && w->rows() == w->M && fused_ok(w, 1) ? 1 : 0;
}

static int launch_neigh(int EPI, NeighArgs & a, const ggml_b200_weight * w) {
    a.nbk = (int)(w->K / QK); a.rows = (int)w->M; a.x_aligned = (((uintptr_t)a.x) & 15) == 0; a.seq = (unsigned)g.launches;
    g.run.count = 0; g.run.x = nullptr;                 // reads more than x: never a member of a same-src1 run
    const size_t smem = gemv_smem_bytes(a.nbk, 1);
    if (EPI == NEIGH_ADD) {
        auto kern = gemv_q4_0_neigh_kernel<4, 1024, NEIGH_ADD>;
        B200_SET_SMEM_ATTR(kern, MAX_DYN_SMEM);
        return launch_k(kern, g.sm_count, 1024, smem, a);
    }
    auto kern = gemv_q4_0_neigh_kernel<4, 512, NEIGH_SWIGLU>;
    B200_SET_SMEM_ATTR(kern, MAX_DYN_SMEM);
    return launch_k(kern, 2 * g.sm_count, 512, smem, a);
}

int ggml_b200_mul_mat_q4_0_add(const ggml_b200_weight * w, const float * x_dev, float * y_dev, const float * r_dev, float * y2_dev) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!w || !x_dev || !y_dev || !r_dev || !y2_dev) { set_error("mul_mat_q4_0_add: null argument"); return GGML_B200_ERR_ARG; }
    if (w->rows() != w->M || !fused_ok(w, 1)) { set_error("mul_mat_q4_0_add: needs the whole weight on this GPU in the plane layout"); return GGML_B200_ERR_ARG; }
    NeighArgs a{};
    a.wq[0] = (const uint4 *)w->wq; a.wd[0] = (const __half *)w->wd; a.x = x_dev; a.r = r_dev; a.y = y_dev; a.y2 = y2_dev;
    return launch_neigh(NEIGH_ADD, a, w);
}

int ggml_b200_mul_mat_q4_0_swiglu(const ggml_b200_weight * w1, const ggml_b200_weight * w3, const float * x_dev, float * y1_dev, float * y3_dev,
                                  float * silu_dev, float * y2_dev) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!w1 || !w3 || !x_dev || !y1_dev || !y3_dev || !y2_dev) { set_error("mul_mat_q4_0_swiglu: null argument"); return GGML_B200_ERR_ARG; }
    if (w1->K != w3->K || w1->M != w3->M) { set_error("mul_mat_q4_0_swiglu: the two weights differ in shape"); return GGML_B200_ERR_ARG; }
    if (w1->rows() != w1->M || w3->rows() != w3->M || !fused_ok(w1, 1) || !fused_ok(w3, 1)) {
        set_error("mul_mat_q4_0_swiglu: needs both whole weights on this GPU in the plane layout"); return GGML_B200_ERR_ARG;
    }
    NeighArgs a{};
    a.wq[0] = (const uint4 *)w1->wq; a.wd[0] = (const __half *)w1->wd; a.wq[1] = (const uint4 *)w3->wq; a.wd[1] = (const __half *)w3->wd;
    a.x = x_dev; a.y = y1_dev; a.yb = y3_dev; a.ys = silu_dev; a.y2 = y2_dev;
    return launch_neigh(NEIGH_SWIGLU, a, w1);
}

// ---- the op without any allocation (no counterpart in the reference) ----------------------
// ggml_b200_mul_mat_q4_0 keeps the quantized activations (kernel 1's output: xq, xds, xdT) in a workspace that grows
// with cudaMalloc.  Callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid it)
// get the same op with that scratch in a file-scope device array of fixed size.  Every GPU that loads the library
// holds the array until exit.  All calls share it and it keeps the last call's contents (nothing in it must start at
// zero: kernel 1 writes every byte kernels 2/3 read, padded scale columns only feed columns that are never stored);
// calls that overlap on the GPU (two streams) race on it.
constexpr size_t STATIC_MAX_ELEMS = (size_t)4 << 20;                  // n * K of one call
constexpr size_t STATIC_XQ_BYTES = STATIC_MAX_ELEMS;                  // int8 [n][K]
constexpr size_t STATIC_XDS_BYTES = STATIC_MAX_ELEMS / 4;             // int2 [n][K/32]
constexpr size_t STATIC_XDT_BYTES = (size_t)1 << 20;                  // float [K/32][n_pad], n_pad <= n + 63
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_XQ_BYTES + STATIC_XDS_BYTES + STATIC_XDT_BYTES];
static unsigned char * g_static_ptr = nullptr;        // DEVICE address of the array on g.device (the symbol's name is only a host-side stand-in)
static unsigned g_static_epoch = 0;

// ---- n = 1 host entry point: one launch, no driver copies, no stream sync (gemv_q4_0_hostio.cuh) -----------
static int hostio_grow_mapped(float ** p, size_t * cap, size_t need) {
    if (need <= *cap) return GGML_B200_OK;
    if (*p) { CK(cudaStreamSynchronize(g.stream)); CK(cudaFreeHost(*p)); *p = nullptr; *cap = 0; }
    need = (need + 65535) & ~(size_t)65535;
    CK(cudaHostAlloc((void **)p, need, cudaHostAllocMapped | cudaHostAllocPortable));
    memset(*p, 0, need);   // no stale word may carry a future op number
    *cap = need;
    return GGML_B200_OK;
}

static void hostio_release() {
    auto & h = g.hio;
    if (h.hx) cudaFreeHost(h.hx);
    if (h.hy) cudaFreeHost(h.hy);
    if (h.xdev) cudaFree(h.xdev);
    if (h.ydev) cudaFree(h.ydev);
    if (h.counters) cudaFree(h.counters);
    h.hx = h.hy = h.xdev = h.ydev = nullptr; h.counters = nullptr;
    h.hx_cap = h.hy_cap = h.xdev_cap = h.ydev_cap = 0; h.seq = 0;
}

static inline uint64_t now_ns() { return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

static size_t hostio_smem(const ggml_b200_weight * w) {
    return gemv_smem_bytes((int)(w->K / QK), 1) + (size_t)((w->rows() + g.sm_count - 1) / g.sm_count) * 4 + 16;
}
// (the kernel has a little static shared memory: its attribute is capped at MAX_DYN_SMEM - 1024; what does not fit
// takes the staged cudaMemcpyAsync path below)
static bool hostio_ok(const ggml_b200_weight * w, int64_t n) { return g.tune.host_fast && fused_ok(w, n) && hostio_smem(w) <= (size_t)MAX_DYN_SMEM - 1024; }

static int mul_mat_hostio(const ggml_b200_weight * w, const float * x_host, float * y_host) {
    auto & h = g.hio;
    const int64_t rows = w->rows();
    const int nbk = (int)(w->K / QK);
    int rc;
    if ((rc = hostio_grow_mapped(&h.hx, &h.hx_cap, (size_t)w->K * 4))) return rc;
    if ((rc = hostio_grow_mapped(&h.hy, &h.hy_cap, (size_t)rows * 8))) return rc;
    if ((rc = ensure_stage((void **)&h.xdev, &h.xdev_cap, (size_t)w->K * 4))) return rc;
    if (!h.counters) {
        CK(cudaMalloc((void **)&h.counters, 64));
        CK(cudaMemsetAsync(h.counters, 0, 64, g.stream));
        h.seq = 0;
    }
    const uint64_t ts0 = now_ns();
    memcpy(h.hx, x_host, (size_t)w->K * 4);
    const uint64_t ts1 = now_ns();
    HostIO io;
    CK(cudaHostGetDevicePointer((void **)&io.x_host, h.hx, 0));
    CK(cudaHostGetDevicePointer((void **)&io.y_host, h.hy, 0));
    io.x_dev = h.xdev; io.counters = h.counters; io.seq = ++h.seq;
    static const int verbose = getenv("GGML_B200_VERBOSE") != nullptr;
    io.profile = verbose;
    const size_t smem = hostio_smem(w);
    auto kern = gemv_q4_0_hostio_kernel<4, 1024>;
    B200_SET_SMEM_ATTR(kern, MAX_DYN_SMEM - 1024);
    rc = launch_k(kern, g.sm_count, 1024, smem, (const uint4 *)w->wq, (const __half *)w->wd, (int)rows, nbk, io);
    if (rc) { h.seq--; return rc; }
    // copy y out as it lands: a word is valid once it carries this op's number.  Look at the stream now and then
    // so that a faulted kernel is reported, not awaited.
    const uint64_t ts2 = now_ns();
    const volatile unsigned long long * words = reinterpret_cast<const volatile unsigned long long *>(h.hy);
    float * yo = y_host + w->row_lo;
    const unsigned long long tag = (unsigned long long)io.seq << 32;
    unsigned spins = 0;
    uint64_t ts3 = 0;
    for (int64_t i = 0; i < rows;) {
        if (i + 4 <= rows) {
            const unsigned long long a = words[i], b = words[i + 1], c = words[i + 2], d = words[i + 3];
            if ((((a ^ tag) | (b ^ tag) | (c ^ tag) | (d ^ tag)) >> 32) == 0) {
                const unsigned ua = (unsigned)a, ub = (unsigned)b, uc = (unsigned)c, ud = (unsigned)d;
                memcpy(yo + i, &ua, 4); memcpy(yo + i + 1, &ub, 4); memcpy(yo + i + 2, &uc, 4); memcpy(yo + i + 3, &ud, 4);
                if (!ts3) ts3 = now_ns();
                i += 4;
                continue;
            }
        }
        const unsigned long long a = words[i];
        if ((a >> 32) == io.seq) { const unsigned ua = (unsigned)a; memcpy(yo + i, &ua, 4); i++; continue; }
        if ((++spins & 0xFFFF) == 0) {
            const cudaError_t e = cudaStreamQuery(g.stream);
            if (e != cudaSuccess && e != cudaErrorNotReady) return cuda_fail(e, "host-path kernel", __FILE__, __LINE__);
            if (now_ns() - ts2 > 20000000000ull) { set_error("host-path kernel did not deliver y in 20 s"); return GGML_B200_ERR_CUDA; }
        }
#if defined(__x86_64__)
        __builtin_ia32_pause();
#endif
    }
    if (!ts3) ts3 = now_ns();
    const uint64_t ts4 = now_ns();
    g_hio_stats.ops++; g_hio_stats.copy_in += ts1 - ts0; g_hio_stats.launch += ts2 - ts1; g_hio_stats.spin += ts3 - ts2; g_hio_stats.copy_out += ts4 - ts3;
    return GGML_B200_OK;
}

int ggml_b200_mul_mat_q4_0_host(const ggml_b200_weight * w, const float * x_host, int64_t n, size_t nb11, float * y_host, size_t nb1) {
    B200_API_LOCK;
    int rc = check_op_args(w, x_host, n, nb11, y_host, nb1);
    if (rc) return rc;
    const int64_t rows = w->rows();
    if (n == 0 || rows == 0) return GGML_B200_OK;
    if (n == 1 && hostio_ok(w, n)) return mul_mat_hostio(w, x_host, y_host);
    const size_t xrow = (size_t)w->K * 4;
    rc = ensure_stage((void **)&g.xstage, &g.xstage_cap, (size_t)n * xrow);
    if (rc) return rc;
    rc = ensure_stage((void **)&g.ystage, &g.ystage_cap, (size_t)n * rows * 4);
    if (rc) return rc;
    if (nb11 == xrow) CK(cudaMemcpyAsync(g.xstage, x_host, (size_t)n * xrow, cudaMemcpyHostToDevice, g.stream));
    else              CK(cudaMemcpy2DAsync(g.xstage, xrow, x_host, nb11, xrow, (size_t)n, cudaMemcpyHostToDevice, g.stream));
    if (fused_ok(w, n)) {
        rc = launch_fused(w, g.xstage, g.ystage, nullptr, nullptr);
    } else {
        rc = launch_quantize(g.xstage, n, w->K, xrow, nullptr);
        if (rc) return rc;
        rc = launch_shard(w, g.ystage, rows, (int)n);
    }
    if (rc) return rc;
    CK(cudaMemcpy2DAsync(y_host + w->row_lo, nb1, g.ystage, (size_t)rows * 4, (size_t)rows * 4, (size_t)n, cudaMemcpyDeviceToHost, g.stream));
    CK(cudaStreamSynchronize(g.stream));
    return GGML_B200_OK;
}

size_t ggml_b200_static_max_elems(void) { return STATIC_MAX_ELEMS; }

int ggml_b200_mul_mat_q4_0_static(const ggml_b200_weight * w, const float * x_dev, int64_t n, size_t nb11, float * y_dev, size_t nb1) {
    B200_API_LOCK;
    int rc = check_op_args(w, x_dev, n, nb11, y_dev, nb1);
    if (rc) return rc;
    if (n == 0 || w->rows() == 0) return GGML_B200_OK;
    const uint64_t elems = (uint64_t)n * (uint64_t)w->K;
    const uint64_t n_pad = ((uint64_t)n + 63) & ~63ull;
    if (elems > STATIC_MAX_ELEMS || (uint64_t)(w->K / QK) * n_pad * 4 > STATIC_XDT_BYTES) {
        set_error("ggml_b200_mul_mat_q4_0_static: n * K = %llu activations exceed the static scratch (max %llu); use ggml_b200_mul_mat_q4_0",
                  (unsigned long long)elems, (unsigned long long)STATIC_MAX_ELEMS);
        return GGML_B200_ERR_ARG;
    }
    if (g_static_epoch != g_init_epoch) {             // set-up, once per device (per init), with that device current
        CK(cudaSetDevice(g.device));
        void * p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_static_scratch));
        g_static_ptr = (unsigned char *)p;
        g_static_epoch = g_init_epoch;
    }
    // run the unchanged dispatch with the workspace pointing at the static array: sized above, so nothing grows
    struct Saved { int8_t * xq; size_t xq_cap; int2 * xds; size_t xds_cap; float * xdT; size_t xdT_cap; long long n_pad; } sv =
        {g.xq, g.xq_cap, g.xds, g.xds_cap, g.xdT, g.xdT_cap, g.xdT_n_pad};
    g.xq = (int8_t *)g_static_ptr;                                          g.xq_cap = STATIC_XQ_BYTES;
    g.xds = (int2 *)(g_static_ptr + STATIC_XQ_BYTES);                       g.xds_cap = STATIC_XDS_BYTES;
    g.xdT = (float *)(g_static_ptr + STATIC_XQ_BYTES + STATIC_XDS_BYTES);   g.xdT_cap = STATIC_XDT_BYTES;
    g.runN.count = 0;      // "same src1 as the previous op": its quantized activations live in the other workspace
    rc = mul_mat_dev(w, x_dev, n, nb11, y_dev, nb1);
    g.runN.count = 0;
    g.xq = sv.xq; g.xq_cap = sv.xq_cap; g.xds = sv.xds; g.xds_cap = sv.xds_cap; g.xdT = sv.xdT; g.xdT_cap = sv.xdT_cap; g.xdT_n_pad = sv.n_pad;
    return rc;
}

int ggml_b200_block_dots_q4_0(const ggml_b200_weight * w, const float * x_dev, int64_t n, size_t nb11, int32_t * S_dev, int path) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!w || !x_dev || !S_dev || n < 0 || nb11 < (size_t)w->K * 4 || (nb11 & 3)) { set_error("bad block_dots arguments"); return GGML_B200_ERR_ARG; }
    if (n == 0 || w->rows() == 0) return GGML_B200_OK;
    int rc = ensure_stage((void **)&g.ystage, &g.ystage_cap, (size_t)n * w->rows() * 4);
    if (rc) return rc;
    if (path == GGML_B200_PATH_GEMV && fused_ok(w, n)) return launch_fused(w, x_dev, g.ystage, S_dev, nullptr);
    rc = launch_quantize(x_dev, n, w->K, nb11, nullptr);
    if (rc) return rc;
    if (path == GGML_B200_PATH_GEMM) {
        if (!gemm_available(w)) { set_error("tensor-core GEMM path not available for this weight (layout %d)", w->layout); return GGML_B200_ERR_ARG; }
        return launch_gemm(w, g.ystage, w->rows(), (int)n, S_dev);
    }
    return launch_gemv_dump(w, g.ystage, w->rows(), (int)n, S_dev);
}

// per-CTA timeline trace of the n = 1 kernels (profiling aid; see common.cuh CtaTrace)
int ggml_b200_trace_begin(unsigned int capacity) {
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    CK(cudaStreamSynchronize(g.stream));
    if (h_trace_dev) { cudaFree(h_trace_dev); h_trace_dev = nullptr; }
    CK(cudaMalloc(&h_trace_dev, (size_t)capacity * 64));
    CK(cudaMemset(h_trace_dev, 0, (size_t)capacity * 64));
    h_trace_cap = capacity;
    unsigned int zero = 0;
    CK(cudaMemcpyToSymbol(g_trace_cap, &capacity, sizeof(capacity)));
    CK(cudaMemcpyToSymbol(g_trace_pos, &zero, sizeof(zero)));
    CK(cudaMemcpyToSymbol(g_trace, &h_trace_dev, sizeof(h_trace_dev)));
    return GGML_B200_OK;
}

int ggml_b200_trace_end(unsigned long long * out, unsigned int max_records) {
    if (!h_trace_dev) return 0;
    cudaDeviceSynchronize();
    unsigned long long * null_ptr = nullptr;
    unsigned int pos = 0;
    cudaMemcpyFromSymbol(&pos, g_trace_pos, sizeof(pos));
    cudaMemcpyToSymbol(g_trace, &null_ptr, sizeof(null_ptr));
    if (pos > h_trace_cap) pos = h_trace_cap;
    if (pos > max_records) pos = max_records;
    cudaMemcpy(out, h_trace_dev, (size_t)pos * 64, cudaMemcpyDeviceToHost);
    cudaFree(h_trace_dev);
    h_trace_dev = nullptr;
    return (int)pos;
}

int ggml_b200_fused_checkpoint_q4_0(const ggml_b200_weight * w, const float * x_dev, int32_t * S_dev, void * blocks34_dev) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!w || !x_dev) { set_error("null argument"); return GGML_B200_ERR_ARG; }
    if (w->layout != GGML_B200_LAYOUT_PLANES) { set_error("the fused kernel needs the plane layout"); return GGML_B200_ERR_ARG; }
    if (w->rows() == 0) return GGML_B200_OK;
    int rc = ensure_stage((void **)&g.ystage, &g.ystage_cap, (size_t)w->rows() * 4);
    if (rc) return rc;
    int * S = S_dev;
    if (!S) {   // the DUMP kernel always stores S_b
        rc = ensure_stage((void **)&g.sstage, &g.sstage_cap, (size_t)w->rows() * (w->K / QK) * 4);
        if (rc) return rc;
        S = (int *)g.sstage;
    }
    return launch_fused(w, x_dev, g.ystage, S, (unsigned char *)blocks34_dev);
}

}  // extern "C"
