// gemm_q4_0_tc.cuh -- kernel 3: prefill GEMM, Q4_0 weights x q8_0 activations on tcgen05.
//
// Device replacement of the COMPUTE phase of ggml_compute_forward_mul_mat_q_f32 for many
// columns (reference ggml.c:11373-11411, a loop of ggml_vec_dot_q4_0_q8_0 per column;
// ggml.c:2339-2607):
//     Y[c][m] = sum_b fp32(dw[m,b]) * fp32(dx[c,b]) * S_b[m,c],  S_b = sum_j (q4_j - 8) * q8_j
// One `tcgen05.mma.kind::i8` (M = 128 weight rows, N = 128 activation columns, K = 32 = one
// Q4_0 block) forms the 128x128 int32 S_b exactly in TMEM; because every block has its own
// fp16 x fp16 scale pair the int32 accumulator cannot run across blocks, so each MMA targets a
// fresh TMEM buffer (4-deep ring) and the epilogue warps fold S_b * dw * dx into fp32 registers.
//
// CTA = one 128 (rows) x 128 (columns) output tile, 16 warps:
//   warp 0      TMA producer: per k-stage (4 blocks = 128 K-bytes) loads the 128x64 B nibble tile
//               (plane layout), the 128x128 B int8 activation tile (SWIZZLE_128B, K-major) and the
//               4x128 block-major activation scales, each into its own mbarrier ring
//   warp 1      TMEM allocator + MMA issuer (one thread): 4 MMAs per stage, tcgen05.commit per MMA
//   warps 4-7   unpack: nibbles -> signed int8 (q - 8), written K-major with the 128-byte swizzle
//               the UMMA descriptor expects; (x + 0x78) ^ 0x80 per byte = x - 8 without carries
//   warps 8-15  epilogue: tcgen05.ld of S_b, exact int->float by the 1.5*2^23 bias trick folded
//               into an FMA with dw, second FMA with dx into 64 fp32 accumulators per thread;
//               warps 8-11 own columns 0-63, warps 12-15 columns 64-127 of the same 128 TMEM lanes
// Roofline: the north star judges it against the dense int8 tensor peak; the practical bound is the
// epilogue (2 FMA-pipe ops + 1 ALU op per S_b element), not the tensor pipe.
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "ptx.cuh"

namespace b200 {

constexpr int GEMM_BM = 128;            // weight rows per tile  (UMMA M, TMEM lanes)
constexpr int GEMM_BN = 128;            // activation columns per tile (UMMA N, TMEM columns per buffer)
constexpr int GEMM_KB = 4;              // Q4_0 blocks per k-stage (128 K-bytes of int8 = one swizzle row)
constexpr int GEMM_NS_RAW = 4;          // nibble-tile ring
constexpr int GEMM_NS_OP = 4;           // int8 operand ring (A unpacked + B)
constexpr int GEMM_NS_X = 8;            // activation-scale ring
constexpr int GEMM_NS_D = 4;            // TMEM accumulator ring (4 x 128 columns = 512)
constexpr int GEMM_THREADS = 512;
constexpr int GEMM_UNPACK_WARP0 = 4, GEMM_EPI_WARP0 = 8;

constexpr int GEMM_RAW_BYTES = GEMM_BM * GEMM_KB * 16;          // 8 KB
constexpr int GEMM_A_BYTES = GEMM_BM * GEMM_KB * 32;            // 16 KB
constexpr int GEMM_B_BYTES = GEMM_BN * GEMM_KB * 32;            // 16 KB
constexpr int GEMM_X_BYTES = GEMM_KB * GEMM_BN * 4;             // 2 KB
constexpr int GEMM_SMEM_OP = 0;                                                    // 1024-aligned
constexpr int GEMM_SMEM_RAW = GEMM_SMEM_OP + GEMM_NS_OP * (GEMM_A_BYTES + GEMM_B_BYTES);
constexpr int GEMM_SMEM_X = GEMM_SMEM_RAW + GEMM_NS_RAW * GEMM_RAW_BYTES;
constexpr int GEMM_SMEM_BAR = GEMM_SMEM_X + GEMM_NS_X * GEMM_X_BYTES;
constexpr int GEMM_SMEM_BYTES = GEMM_SMEM_BAR + 512 + 1024;                        // + alignment slack

struct GemmBars {   // byte offsets of the mbarriers inside the barrier block
    static constexpr int raw_full = 0, raw_empty = 8 * GEMM_NS_RAW;
    static constexpr int op_full = raw_empty + 8 * GEMM_NS_RAW, op_empty = op_full + 8 * GEMM_NS_OP;
    static constexpr int x_full = op_empty + 8 * GEMM_NS_OP, x_empty = x_full + 8 * GEMM_NS_X;
    static constexpr int d_full = x_empty + 8 * GEMM_NS_X, d_empty = d_full + 8 * GEMM_NS_D;
    static constexpr int tmem_slot = d_empty + 8 * GEMM_NS_D;
};
static_assert(GemmBars::tmem_slot + 8 <= 512, "barrier block too small");

// 8 nibbles of one 32-bit word -> two words of signed int8 (q - 8): low nibbles, high nibbles
__device__ __forceinline__ void unpack_q4_word(uint32_t w, uint32_t & lo, uint32_t & hi) {
    lo = ((w & 0x0F0F0F0Fu) + 0x78787878u) ^ 0x80808080u;
    hi = (((w >> 4) & 0x0F0F0F0Fu) + 0x78787878u) ^ 0x80808080u;
}

template <bool DUMP>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_q4_0_tc_kernel(const __grid_constant__ CUtensorMap tm_nib,   // nibble plane  [rows][nbk*16 B], box 64 B x 128
                    const __grid_constant__ CUtensorMap tm_xq,    // activations   [n][K] int8,      box 128 B x 128, SW128
                    const __grid_constant__ CUtensorMap tm_xd,    // act. scales   [nbk][n_pad] f32, box 128 x 4
                    const __half * __restrict__ wd,               // weight scales [rows][nbk]
                    float * __restrict__ y, long long ldy, int rows, int n, int nbk,
                    int * __restrict__ S_out) {
    extern __shared__ unsigned char smem_unaligned[];
    unsigned char * smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_unaligned) + 1023) & ~(uintptr_t)1023);
    const uint32_t sbase = smem_u32(smem);
    const uint32_t bars = sbase + GEMM_SMEM_BAR;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * GEMM_BM;       // first weight row of the tile
    const int c0 = blockIdx.y * GEMM_BN;       // first activation column of the tile
    const int nstages = (nbk + GEMM_KB - 1) / GEMM_KB;

    if (threadIdx.x == 0) {
        for (int i = 0; i < GEMM_NS_RAW; i++) { mbar_init(bars + GemmBars::raw_full + 8 * i, 1); mbar_init(bars + GemmBars::raw_empty + 8 * i, 4); }
        for (int i = 0; i < GEMM_NS_OP; i++)  { mbar_init(bars + GemmBars::op_full + 8 * i, 1 + 4); mbar_init(bars + GemmBars::op_empty + 8 * i, 1); }
        for (int i = 0; i < GEMM_NS_X; i++)   { mbar_init(bars + GemmBars::x_full + 8 * i, 1); mbar_init(bars + GemmBars::x_empty + 8 * i, 8); }
        for (int i = 0; i < GEMM_NS_D; i++)   { mbar_init(bars + GemmBars::d_full + 8 * i, 1); mbar_init(bars + GemmBars::d_empty + 8 * i, 8); }
        mbar_fence_init();
        tma_prefetch_desc(&tm_nib); tma_prefetch_desc(&tm_xq); tma_prefetch_desc(&tm_xd);
    }
    if (warp == 1) {
        tmem_alloc(bars + GemmBars::tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    pdl_launch_dependents();
    pdl_wait();   // the quantized activations come from the previous kernel
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t *>(smem + GEMM_SMEM_BAR + GemmBars::tmem_slot);

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            for (int s = 0; s < nstages; s++) {
                const int sr = s % GEMM_NS_RAW, so = s % GEMM_NS_OP, sx = s % GEMM_NS_X;
                if (s >= GEMM_NS_RAW) mbar_wait(bars + GemmBars::raw_empty + 8 * sr, ((s / GEMM_NS_RAW) - 1) & 1);
                mbar_expect_tx(bars + GemmBars::raw_full + 8 * sr, GEMM_RAW_BYTES);
                tma_load_2d(sbase + GEMM_SMEM_RAW + sr * GEMM_RAW_BYTES, &tm_nib, s * GEMM_KB * 16, m0, bars + GemmBars::raw_full + 8 * sr);
                if (s >= GEMM_NS_X) mbar_wait(bars + GemmBars::x_empty + 8 * sx, ((s / GEMM_NS_X) - 1) & 1);
                mbar_expect_tx(bars + GemmBars::x_full + 8 * sx, GEMM_X_BYTES);
                tma_load_2d(sbase + GEMM_SMEM_X + sx * GEMM_X_BYTES, &tm_xd, c0, s * GEMM_KB, bars + GemmBars::x_full + 8 * sx);
                if (s >= GEMM_NS_OP) mbar_wait(bars + GemmBars::op_empty + 8 * so, ((s / GEMM_NS_OP) - 1) & 1);
                mbar_expect_tx(bars + GemmBars::op_full + 8 * so, GEMM_B_BYTES);
                tma_load_2d(sbase + GEMM_SMEM_OP + so * (GEMM_A_BYTES + GEMM_B_BYTES) + GEMM_A_BYTES, &tm_xq, s * GEMM_KB * 32, c0,
                            bars + GemmBars::op_full + 8 * so);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_i8(GEMM_BM, GEMM_BN);
            int blk = 0;
            for (int s = 0; s < nstages; s++) {
                const int so = s % GEMM_NS_OP;
                mbar_wait(bars + GemmBars::op_full + 8 * so, (s / GEMM_NS_OP) & 1);
                tc_fence_after_sync();
                const uint32_t a_addr = sbase + GEMM_SMEM_OP + so * (GEMM_A_BYTES + GEMM_B_BYTES);
                const uint32_t b_addr = a_addr + GEMM_A_BYTES;
                const int nb = min(GEMM_KB, nbk - s * GEMM_KB);
                for (int j = 0; j < nb; j++, blk++) {
                    const int d = blk % GEMM_NS_D;
                    if (blk >= GEMM_NS_D) mbar_wait(bars + GemmBars::d_empty + 8 * d, ((blk / GEMM_NS_D) - 1) & 1);
                    tc_fence_after_sync();
                    mma_i8_ss(tmem_base + d * GEMM_BN, umma_desc_k_sw128(a_addr + j * 32), umma_desc_k_sw128(b_addr + j * 32), idesc, 0u);
                    mma_commit(bars + GemmBars::d_full + 8 * d);
                }
                mma_commit(bars + GemmBars::op_empty + 8 * so);   // operands of this stage are free again
            }
        }
    } else if (warp >= GEMM_UNPACK_WARP0 && warp < GEMM_EPI_WARP0) {
        // ===== unpack: nibble tile -> K-major int8 A tile with the 128-byte swizzle =====
        const int t = threadIdx.x - GEMM_UNPACK_WARP0 * 32;      // 0..127
        const int j = t & 3;                                       // block within the stage
        for (int s = 0; s < nstages; s++) {
            const int sr = s % GEMM_NS_RAW, so = s % GEMM_NS_OP;
            mbar_wait(bars + GemmBars::raw_full + 8 * sr, (s / GEMM_NS_RAW) & 1);
            if (s >= GEMM_NS_OP) mbar_wait(bars + GemmBars::op_empty + 8 * so, ((s / GEMM_NS_OP) - 1) & 1);
            const unsigned char * raw = smem + GEMM_SMEM_RAW + sr * GEMM_RAW_BYTES;
            unsigned char * a = smem + GEMM_SMEM_OP + so * (GEMM_A_BYTES + GEMM_B_BYTES);
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int r = (t >> 2) + 32 * i;                  // tile row
                const uint4 q = *reinterpret_cast<const uint4 *>(raw + r * 64 + j * 16);
                uint4 lo, hi;
                unpack_q4_word(q.x, lo.x, hi.x);
                unpack_q4_word(q.y, lo.y, hi.y);
                unpack_q4_word(q.z, lo.z, hi.z);
                unpack_q4_word(q.w, lo.w, hi.w);
                // K bytes [32j, 32j+16) = elements 0..15 (low nibbles), [32j+16, 32j+32) = elements 16..31
                unsigned char * row = a + r * 128;
                *reinterpret_cast<uint4 *>(row + (((2 * j) ^ (r & 7)) << 4)) = lo;
                *reinterpret_cast<uint4 *>(row + (((2 * j + 1) ^ (r & 7)) << 4)) = hi;
            }
            fence_proxy_async_smem();      // generic-proxy stores -> tensor-core (async proxy) reads
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(bars + GemmBars::op_full + 8 * so);
                mbar_arrive(bars + GemmBars::raw_empty + 8 * sr);
            }
        }
    } else if (warp >= GEMM_EPI_WARP0) {
        // ===== epilogue: S_b (TMEM) * dw * dx -> fp32 accumulators =====
        const int ew = warp - GEMM_EPI_WARP0;          // 0..7
        const int quad = ew & 3;                        // TMEM lane quadrant (== warp % 4)
        const int half = ew >> 2;                       // column half: 0 -> 0..63, 1 -> 64..127
        const int m = m0 + quad * 32 + lane;            // this thread's weight row
        const bool row_ok = m < rows;
        const unsigned short * drow = reinterpret_cast<const unsigned short *>(wd) + (size_t)(row_ok ? m : 0) * nbk;
        const uint32_t t_lane = tmem_base + ((uint32_t)(quad * 32) << 16) + half * 64;
        const float MAGIC = 12582912.0f;                // 1.5 * 2^23: int_as_float(S + bits(MAGIC)) == MAGIC + S exactly

        float acc[64];
#pragma unroll
        for (int i = 0; i < 64; i++) acc[i] = 0.0f;

        unsigned short dnext[GEMM_KB];
#pragma unroll
        for (int jj = 0; jj < GEMM_KB; jj++) dnext[jj] = (row_ok && jj < nbk) ? ldg_stream_u16(drow + jj) : (unsigned short)0;

        int blk = 0;
        for (int s = 0; s < nstages; s++) {
            const int sx = s % GEMM_NS_X;
            unsigned short dcur[GEMM_KB];
#pragma unroll
            for (int jj = 0; jj < GEMM_KB; jj++) dcur[jj] = dnext[jj];
            if (s + 1 < nstages) {
#pragma unroll
                for (int jj = 0; jj < GEMM_KB; jj++) {
                    const int b = (s + 1) * GEMM_KB + jj;
                    dnext[jj] = (row_ok && b < nbk) ? ldg_stream_u16(drow + b) : (unsigned short)0;
                }
            }
            mbar_wait(bars + GemmBars::x_full + 8 * sx, (s / GEMM_NS_X) & 1);
            const float * xs = reinterpret_cast<const float *>(smem + GEMM_SMEM_X + sx * GEMM_X_BYTES) + half * 64;
            const int nb = min(GEMM_KB, nbk - s * GEMM_KB);
#pragma unroll
            for (int jj = 0; jj < GEMM_KB; jj++) {
                if (jj < nb) {
                    const int d = blk % GEMM_NS_D;
                    mbar_wait(bars + GemmBars::d_full + 8 * d, (blk / GEMM_NS_D) & 1);
                    tc_fence_after_sync();
                    const float dw = half_bits_to_float(dcur[jj]);
                    const float nc = -MAGIC * dw;                         // exact: dw has 11 significant bits
                    const float * xrow = xs + jj * GEMM_BN;
#pragma unroll
                    for (int ch = 0; ch < 2; ch++) {
                        uint32_t sreg[32];
                        tmem_ld_32x32(t_lane + d * GEMM_BN + ch * 32, sreg);
                        tmem_ld_wait();
                        if (ch == 1) {
                            // both halves of this thread's 64 columns are in registers: free the buffer
                            tc_fence_before_sync();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(bars + GemmBars::d_empty + 8 * d);
                        }
#pragma unroll
                        for (int i = 0; i < 32; i += 4) {
                            const float4 dx4 = *reinterpret_cast<const float4 *>(xrow + ch * 32 + i);   // warp-uniform broadcast
                            const float dxv[4] = {dx4.x, dx4.y, dx4.z, dx4.w};
#pragma unroll
                            for (int k = 0; k < 4; k++) {
                                const int sb = (int)sreg[i + k];
                                if (DUMP) {
                                    const int c = c0 + half * 64 + ch * 32 + i + k;
                                    if (row_ok && c < n) S_out[((size_t)c * rows + m) * nbk + (s * GEMM_KB + jj)] = sb;
                                }
                                const float f = __int_as_float(sb + 0x4B400000);      // MAGIC + S_b, exact
                                const float tt = fmaf(f, dw, nc);                     // == rn(S_b * dw)
                                acc[ch * 32 + i + k] = fmaf(tt, dxv[k], acc[ch * 32 + i + k]);
                            }
                        }
                    }
                    blk++;
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bars + GemmBars::x_empty + 8 * sx);
        }
        // store: for a fixed column the 32 lanes of a warp hold 32 consecutive rows -> 128-byte stores
        if (row_ok) {
#pragma unroll
            for (int i = 0; i < 64; i++) {
                const int c = c0 + half * 64 + i;
                if (c < n) y[(size_t)c * ldy + m] = acc[i];
            }
        }
    }

    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace b200
