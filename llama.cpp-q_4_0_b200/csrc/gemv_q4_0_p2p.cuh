// gemv_q4_0_p2p.cuh -- kernel 2 (n = 1) for a row shard with the multi-GPU gather fused in: ONE launch per
// mul_mat computes this rank's rows, delivers them to every peer over NVLink and assembles the peers' rows.
//
// Same arithmetic as gemv_q4_0_fused.cuh (quantize_row_q8_0 in the prologue, exact S_b on IDP.4A, per-lane fp32
// partials + xor-shuffle tree).  What differs:
//   * a CTA owns a CONTIGUOUS slice of the shard's rows (warps take rows of the slice round-robin), so its
//     results form one contiguous run of dst;
//   * results go to shared memory and to this rank's dst; when the slice is done the CTA stores it, tagged with
//     the op number, into every peer's arena (p2p.cuh) -- coalesced 8-byte words, no fence, and the stores of
//     early CTAs travel while later CTAs still stream weights;
//   * the CTA then polls this rank's arena for its share of every peer's rows and writes them to dst.
// Against GEMV + stand-alone gather kernel this removes one dependent-launch boundary (~1.4 us) per matmul and
// hides the remote-store latency under the weight stream.  Reference mechanism replaced: GGML_BACKEND_GPU_SPLIT
// row slices + per-column cudaMemcpyAsync gather (ggml-cuda.cu:2514-2540, :2780-2797).
// Roofline: HBM; algorithmic bytes per launch 18*rows*K/32 + 4*K + 4*M.
#pragma once
#include "common.cuh"
#include "gemv_q4_0.cuh"
#include "gemv_q4_0_fused.cuh"
#include "p2p.cuh"

namespace b200 {

// dynamic smem: quantized-x planes (40*nbk B) + one float per row of the CTA's slice
__host__ __device__ inline size_t gemv_p2p_smem_bytes(int nbk, int rows, int grid) {
    return gemv_smem_bytes(nbk, 1) + (size_t)((rows + grid - 1) / grid) * 4 + 16;
}

template <int UNROLL, int MAXT>
__global__ void __launch_bounds__(MAXT, MAXT <= 512 ? 2 : 1)
gemv_q4_0_fused_p2p_kernel(const uint4 * __restrict__ wq, const __half * __restrict__ wd,
                           const float * __restrict__ x, float * __restrict__ y /* whole dst column */,
                           int rows, int nbk, int x_aligned, unsigned seq_trace, const __grid_constant__ P2PExchange ex,
                           int run_index /* 0: dependent op; j > 0: j-th op that only shares src1 with its predecessors */) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    GemvSmem<1> s(smem_raw, nbk);
    float * yres = reinterpret_cast<float *>(smem_raw + gemv_smem_bytes(nbk, 1));
    __shared__ unsigned s_seq;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int nwarps = blockDim.x >> 5;
    CtaTrace tr;
    if (threadIdx.x == 0) tr.begin(seq_trace);
    if (run_index && threadIdx.x == 0) pdl_launch_dependents();      // independent of the previous op (gemv_q4_0_fused.cuh): no wait at the top
    unsigned seq_ld = 0;
    if (threadIdx.x == 0) seq_ld = p2p_counter(ex) + (run_index ? (unsigned)run_index : (unsigned)P2P_MAX_RUN);   // stable before the wait (p2p.cuh)

    // this CTA's contiguous slice of the shard
    int j0, j1;
    p2p_cta_slice(rows, j0, j1);
    const unsigned short * wd16 = reinterpret_cast<const unsigned short *>(wd);

    // 1. first row of this warp (weights are never produced by the previous kernel).  Warp 0 of a dependent op
    // holds back: its thread 0 publishes the op number behind a __threadfence right after the wait, and a fence
    // with weight loads in flight would stall on them.
    uint4 q[UNROLL];
    unsigned short dh[UNROLL];
    bool prefetched = run_index != 0 || warp != 0;
    if (prefetched) {
        const int row = j0 + warp;
        if (row < j1) {
#pragma unroll
            for (int u = 0; u < UNROLL; u++) {
                const int b = lane + 32 * u;
                if (b < nbk) { q[u] = ldg_stream_u4(wq + (size_t)row * nbk + b); dh[u] = ldg_stream_u16(wd16 + (size_t)row * nbk + b); }
            }
        }
    }

    // 2. x and dst may belong to the previous kernel.  A dependent op publishes its op number for the followers of
    // its run and releases its dependents right after the wait.
    if (!run_index) {
        pdl_wait();
        if (threadIdx.x == 0) { tr.mark(1); p2p_head_publish_and_release(ex, seq_ld); }
    }

    // 3. quantize_row_q8_0 into the smem planes
    quantize_x_to_smem<false>(s, x, nbk, x_aligned, warp, lane, nullptr);
    if (threadIdx.x == 0) s_seq = seq_ld;
    __syncthreads();
    if (threadIdx.x == 0) tr.mark(2);
    const unsigned seq = s_seq;
    int lo_me, n_me;
    p2p_rows_of(ex, ex.me, lo_me, n_me);

    // 4. dot loop over the slice
    for (int row = j0 + warp; row < j1; row += nwarps) {
        float acc = 0.0f;
        for (int b0 = lane; b0 < nbk; b0 += 32 * UNROLL) {
            if (!prefetched) {
#pragma unroll
                for (int u = 0; u < UNROLL; u++) {
                    const int b = b0 + 32 * u;
                    if (b < nbk) { q[u] = ldg_stream_u4(wq + (size_t)row * nbk + b); dh[u] = ldg_stream_u16(wd16 + (size_t)row * nbk + b); }
                }
            }
            prefetched = false;
#pragma unroll
            for (int u = 0; u < UNROLL; u++) {
                const int b = b0 + 32 * u;
                if (b < nbk) {
                    const int2 ds = s.ds[b];
                    const int sb = block_dot_q4_q8(q[u], s.lo[b], s.hi[b], ds.y);
                    acc = fmaf((float)sb, half_bits_to_float(dh[u]) * __int_as_float(ds.x), acc);
                }
            }
        }
        const float t = warp_sum(acc);
        if (lane == 0) { yres[row - j0] = t; y[lo_me + row] = t; }
    }
    __syncthreads();

    // 5. slice -> every peer; peers' rows -> dst
    p2p_send(ex, seq, yres, lo_me + j0, j1 - j0);
    p2p_receive(ex, seq, y);
    if (threadIdx.x == 0) tr.mark(3);
    if (run_index) pdl_wait();                   // completion of this op must imply completion of its predecessors
}

}  // namespace b200
