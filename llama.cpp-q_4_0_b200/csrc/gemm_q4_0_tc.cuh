// gemm_q4_0_tc.cuh -- kernel 3: prefill GEMM, Q4_0 weights x q8_0 activations on tcgen05.
//
// Device replacement of the COMPUTE phase of ggml_compute_forward_mul_mat_q_f32 for many
// columns (reference ggml.c:11373-11411, a loop of ggml_vec_dot_q4_0_q8_0 per column;
// ggml.c:2339-2607):
//     Y[c][m] = sum_b fp32(dw[m,b]) * fp32(dx[c,b]) * S_b[m,c],  S_b = sum_j (q4_j - 8) * q8_j
// One `tcgen05.mma.kind::i8` (M = 128 weight rows, N = 128 activation columns, K = 32 = one
// Q4_0 block) forms the 128x128 int32 S_b exactly in TMEM; because every block has its own
// fp16 x fp16 scale pair the int32 accumulator cannot run across blocks, so each block's MMA
// targets a fresh TMEM buffer (4-deep ring) and the epilogue warps fold S_b * dw * dx into fp32
// registers.  That epilogue (2 FP32 multiply-adds per S_b element on the CUDA cores), not the
// tensor pipe, is the bound of this formulation: 128x128 elements per block at 128 packed
// FMA lanes per clock per SM = 256 clk against 64 clk of MMA (see DESIGN.md "Kernel 3").
//
// CTA = one 128 (rows) x 128 (columns) output tile, 20 warps (640 threads x 96 registers):
//   warp 18     TMA producer: per k-stage (4 blocks = 128 K-bytes) loads the 128x64 B nibble tile
//               (plane layout), the 128x128 B int8 activation tile (SWIZZLE_128B, K-major) and the
//               4x128 block-major activation scales, each into its own mbarrier ring
//   warp 19     TMEM allocator + MMA issuer (one thread): per block one MMA (+ one constant-tile
//               MMA that adds 32768 to every S_b when BIAS), tcgen05.commit per block
//   warps 16-17 unpack: nibbles -> signed int8 (q - 8), written K-major with the 128-byte swizzle
//               the UMMA descriptor expects; (x + 0x78) ^ 0x80 per byte = x - 8 without carries.
//               8 (row, block) items per thread and k-stage, loads batched ahead of the expansion.
//               The same threads stage the items' weight scales (fp16 -> fp32) for the epilogue.
//   warps 0-15  epilogue: warp = (TMEM lane quadrant, 32-column group).  tcgen05.ld of S_b, then
//               per pair of columns two packed FMAs (fma.rn.f32x2): t = rn(S_b * dw), acc += t * dx.
//               int -> float without a conversion instruction:
//                 BIAS = false: f = int_as_float(S_b + 0x4B400000) = 1.5*2^23 + S_b (one IADD), and
//                               t = fma(f, dw, -1.5*2^23*dw) = rn(S_b * dw) exactly;
//                 BIAS = true : the constant-tile MMA made the accumulator S_b + 32768 >= 0, whose
//                               bits read as a float are the DENORMAL (S_b + 32768) * 2^-149, exact;
//                               t' = fma(f, dw*2^100, -2^-134 * dw*2^100) = rn(S_b * dw) * 2^-49, the
//                               accumulators run scaled by 2^-49 and are rescaled once at the end.
//               Same fp32 operation order either way: y = sum_b rn(rn(S_b*dw) * dx + acc).
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
constexpr int GEMM_THREADS = 384;
// Warp roles, lowest id first.  The SM's warp schedulers favour the HIGHEST warp id among eligible warps, so the
// producers sit on top: the single MMA-issuing thread and the TMA thread must never queue behind 16 warps of FMAs
// (measured: with the issuer as warp 1 every block cost it ~400 clk of issue-slot starvation).
constexpr int GEMM_EPI_WARP0 = 4, GEMM_EPI_WARPS = 8, GEMM_UNPACK_WARP0 = 2, GEMM_UNPACK_WARPS = 2;
constexpr int GEMM_TMA_WARP = 0, GEMM_MMA_WARP = 1;
constexpr int GEMM_UNPACK_ROWS = GEMM_BM / (GEMM_UNPACK_WARPS * 8);   // rows per unpack thread (8)
// 640 threads x 96 registers (65536 / 640 = 102 -> 96): the epilogue (32 accumulators + 32 S_b + operands)
// fits without spills, so no setmaxnreg re-split is needed.

constexpr int GEMM_RAW_BYTES = GEMM_BM * GEMM_KB * 16;          // 8 KB
constexpr int GEMM_A_BYTES = GEMM_BM * GEMM_KB * 32;            // 16 KB
constexpr int GEMM_B_BYTES = GEMM_BN * GEMM_KB * 32;            // 16 KB
constexpr int GEMM_DW_STRIDE = 136;     // floats per block row of staged weight scales (== 8 mod 32: conflict-free stores)
constexpr int GEMM_XD_BYTES = GEMM_KB * GEMM_BN * 4;            // 2 KB   activation scales [4][128] f32 (TMA)
constexpr int GEMM_DW_BYTES = GEMM_KB * GEMM_DW_STRIDE * 4;     // 2176 B weight scales [4][136] f32 (unpack warps)
constexpr int GEMM_X_BYTES = GEMM_XD_BYTES + GEMM_DW_BYTES;     // 4224 B (33 * 128)
constexpr int GEMM_SMEM_OP = 0;                                                    // 1024-aligned
constexpr int GEMM_SMEM_CONST = GEMM_SMEM_OP + GEMM_NS_OP * (GEMM_A_BYTES + GEMM_B_BYTES);   // 16 KB of bytes = 32
constexpr int GEMM_SMEM_RAW = GEMM_SMEM_CONST + GEMM_A_BYTES;
constexpr int GEMM_SMEM_X = GEMM_SMEM_RAW + GEMM_NS_RAW * GEMM_RAW_BYTES;
constexpr int GEMM_SMEM_BAR = GEMM_SMEM_X + GEMM_NS_X * GEMM_X_BYTES;
constexpr int GEMM_SMEM_BYTES = GEMM_SMEM_BAR + 512;
static_assert(GEMM_SMEM_BYTES <= 227 * 1024, "shared memory budget");

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

constexpr int GEMM_EPI_COLS = 64;       // tile columns per epilogue thread

// parity checkpoint plumbing of the epilogue (DUMP instantiations only)
struct GemmDump {
    int * S; int rows, n, nbk, m, c_first; bool row_ok;
};

// one chunk of the epilogue = 32 rows x 16 columns of the warp's slice, held in the accumulator-fragment layout
// (ptx.cuh tmem_ld_16x256b_x2): lo[] = TMEM lanes 0..15 of the quadrant, hi[] = lanes 16..31.  Thread i owns rows
// (i/4) + 8j, j = 0..3, and columns 8k + 2(i%4) + {0,1}, k = 0,1: per chunk 4 row scales (registers) and two
// 8-byte column-scale loads whose addresses differ per lane -> 1 shared-memory wavefront each.  (With the
// one-row-per-thread layout every dx was a 32-way broadcast: 512 LSU wavefronts per block, the measured limiter.)
//   acc[((CH * 2 + k) * 4 + j)] = (row j, column pair k of chunk CH)
template <int CH, bool DUMP, bool BIAS>
__device__ __forceinline__ void gemm_epi_chunk(uint32_t (&lo)[8], uint32_t (&hi)[8], uint64_t (&acc)[GEMM_EPI_COLS / 2],
                                               uint32_t xa, const uint64_t (&dw2)[4], const uint64_t (&nc2)[4],
                                               const GemmDump & dp, int blk, int lane) {
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const uint64_t dx2 = lds_b64(xa + (CH * 16 + 8 * k) * 4);    // xa already carries + 2(i%4) columns
#pragma unroll
        for (int j = 0; j < 4; j++) {
            uint32_t f0 = (j < 2) ? lo[4 * k + 2 * (j & 1)] : hi[4 * k + 2 * (j & 1)];
            uint32_t f1 = (j < 2) ? lo[4 * k + 2 * (j & 1) + 1] : hi[4 * k + 2 * (j & 1) + 1];
            if (DUMP) {
                const int c = dp.c_first + CH * 16 + 8 * k + 2 * (lane & 3);
                const int mm = dp.m + 8 * j;
                const int off = BIAS ? 32768 : 0;
                if (mm < dp.rows) {
                    int * dst = dp.S + ((size_t)c * dp.rows + mm) * dp.nbk + blk;
                    if (c < dp.n)     dst[0] = (int)f0 - off;
                    if (c + 1 < dp.n) dst[(size_t)dp.rows * dp.nbk] = (int)f1 - off;
                }
            }
            if (!BIAS) { f0 += 0x4B400000u; f1 += 0x4B400000u; }
            const uint64_t t = fma_f32x2(pack_u32x2(f0, f1), dw2[j], nc2[j]);   // == rn(S_b * dw) (x 2^-49 when BIAS)
            fma_acc_f32x2(acc[(CH * 2 + k) * 4 + j], t, dx2);
        }
    }
}

// TRACE: CTA (0,0) stamps clock64 at its pipeline hand-offs into trc (tools/trace_gemm.py); compiled out otherwise
template <bool DUMP, bool BIAS, bool TRACE>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_q4_0_tc_kernel(const __grid_constant__ CUtensorMap tm_nib,   // nibble plane  [rows][nbk*16 B], box 64 B x 128
                    const __grid_constant__ CUtensorMap tm_xq,    // activations   [n][K] int8,      box 128 B x 128, SW128
                    const __grid_constant__ CUtensorMap tm_xd,    // act. scales   [nbk][n_pad] f32, box 128 x 4
                    const __half * __restrict__ wd,               // weight scales [rows][nbk]
                    float * __restrict__ y, long long ldy, int rows, int n, int nbk,
                    int * __restrict__ S_out, long long * __restrict__ trc,
                    int independent /* reads the activations its predecessor read, writes its own dst: no wait at the top */) {
#define GEMM_TR(role, idx) do { if (TRACE && trc != nullptr && blockIdx.x == 0 && blockIdx.y == 0) trc[(role) * 2048 + (idx)] = clock64(); } while (0)
    // no static shared memory in this kernel: the dynamic window starts 1024-byte aligned (checked below),
    // which the 128-byte swizzle atoms need; a computed alignment would be rematerialised all over the epilogue
    extern __shared__ __align__(1024) unsigned char smem[];
    const uint32_t sbase = smem_u32(smem);
    if ((sbase & 1023u) != 0) __trap();
    const uint32_t bars = sbase + GEMM_SMEM_BAR;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * GEMM_BM;       // first weight row of the tile
    const int c0 = blockIdx.y * GEMM_BN;       // first activation column of the tile
    const int nstages = (nbk + GEMM_KB - 1) / GEMM_KB;

    if (threadIdx.x == 0) {
        for (int i = 0; i < GEMM_NS_RAW; i++) { mbar_init(bars + GemmBars::raw_full + 8 * i, 1); mbar_init(bars + GemmBars::raw_empty + 8 * i, GEMM_UNPACK_WARPS); }
        for (int i = 0; i < GEMM_NS_OP; i++)  { mbar_init(bars + GemmBars::op_full + 8 * i, 1 + GEMM_UNPACK_WARPS); mbar_init(bars + GemmBars::op_empty + 8 * i, 1); }
        for (int i = 0; i < GEMM_NS_X; i++)   { mbar_init(bars + GemmBars::x_full + 8 * i, 1 + GEMM_UNPACK_WARPS); mbar_init(bars + GemmBars::x_empty + 8 * i, GEMM_EPI_WARPS); }
        for (int i = 0; i < GEMM_NS_D; i++)   { mbar_init(bars + GemmBars::d_full + 8 * i, 1); mbar_init(bars + GemmBars::d_empty + 8 * i, GEMM_EPI_WARPS); }
        mbar_fence_init();
        tma_prefetch_desc(&tm_nib); tma_prefetch_desc(&tm_xq); tma_prefetch_desc(&tm_xd);
    }
    if (warp == GEMM_MMA_WARP) {
        tmem_alloc(bars + GemmBars::tmem_slot, 512);
        tmem_relinquish();
    }
    if (BIAS) {
        // constant operand tile: every byte = 32, so a K = 32 MMA of it with itself adds 32*32*32 = 32768
        uint4 * cst = reinterpret_cast<uint4 *>(smem + GEMM_SMEM_CONST);
        for (int i = threadIdx.x; i < GEMM_A_BYTES / 16; i += GEMM_THREADS) cst[i] = make_uint4(0x20202020u, 0x20202020u, 0x20202020u, 0x20202020u);
        fence_proxy_async_smem();
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    // the quantized activations come from the previous kernel -- unless this op only shares them with its predecessor
    // (wq/wk/wv, w3/w1 of one layer; the host skipped kernel 1): then its tiles fill the SMs the predecessor's last wave
    // leaves idle, and it waits at its END so that its completion still implies the predecessor's
    if (!independent) pdl_wait();
    pdl_launch_dependents();   // after the wait: see gemv_q4_0_fused.cuh (independent ops)
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t *>(smem + GEMM_SMEM_BAR + GemmBars::tmem_slot);

    if (warp < GEMM_EPI_WARP0) {
      if (warp < GEMM_UNPACK_WARP0) {
        if (warp == GEMM_TMA_WARP && lane == 0) {
            // ===== TMA producer =====
            for (int s = 0; s < nstages; s++) {
                const int sr = s % GEMM_NS_RAW, so = s % GEMM_NS_OP, sx = s % GEMM_NS_X;
                if (s >= GEMM_NS_RAW) mbar_wait(bars + GemmBars::raw_empty + 8 * sr, ((s / GEMM_NS_RAW) - 1) & 1);
                GEMM_TR(0, 3 * s);
                mbar_expect_tx(bars + GemmBars::raw_full + 8 * sr, GEMM_RAW_BYTES);
                tma_load_2d(sbase + GEMM_SMEM_RAW + sr * GEMM_RAW_BYTES, &tm_nib, s * GEMM_KB * 16, m0, bars + GemmBars::raw_full + 8 * sr);
                if (s >= GEMM_NS_OP) mbar_wait(bars + GemmBars::op_empty + 8 * so, ((s / GEMM_NS_OP) - 1) & 1);
                mbar_expect_tx(bars + GemmBars::op_full + 8 * so, GEMM_B_BYTES);
                tma_load_2d(sbase + GEMM_SMEM_OP + so * (GEMM_A_BYTES + GEMM_B_BYTES) + GEMM_A_BYTES, &tm_xq, s * GEMM_KB * 32, c0,
                            bars + GemmBars::op_full + 8 * so);
                GEMM_TR(0, 3 * s + 1);
                if (s >= GEMM_NS_X) mbar_wait(bars + GemmBars::x_empty + 8 * sx, ((s / GEMM_NS_X) - 1) & 1);
                GEMM_TR(0, 3 * s + 2);
                mbar_expect_tx(bars + GemmBars::x_full + 8 * sx, GEMM_XD_BYTES);
                tma_load_2d(sbase + GEMM_SMEM_X + sx * GEMM_X_BYTES, &tm_xd, c0, s * GEMM_KB, bars + GemmBars::x_full + 8 * sx);
            }
        } else if (warp == GEMM_MMA_WARP && lane == 0) {
            // ===== MMA issuer =====
            const uint32_t idesc = umma_idesc_i8(GEMM_BM, GEMM_BN);
            const uint64_t cdesc = umma_desc_k_sw128(sbase + GEMM_SMEM_CONST);
            int blk = 0;
            for (int s = 0; s < nstages; s++) {
                const int so = s % GEMM_NS_OP;
                mbar_wait(bars + GemmBars::op_full + 8 * so, (s / GEMM_NS_OP) & 1);
                GEMM_TR(1, 512 + s);
                tc_fence_after_sync();
                const uint32_t a_addr = sbase + GEMM_SMEM_OP + so * (GEMM_A_BYTES + GEMM_B_BYTES);
                const uint32_t b_addr = a_addr + GEMM_A_BYTES;
                const int nb = min(GEMM_KB, nbk - s * GEMM_KB);
                for (int j = 0; j < nb; j++, blk++) {
                    const int d = blk % GEMM_NS_D;
                    if (blk >= GEMM_NS_D) mbar_wait(bars + GemmBars::d_empty + 8 * d, ((blk / GEMM_NS_D) - 1) & 1);
                    GEMM_TR(1, blk);
                    tc_fence_after_sync();
                    mma_i8_ss(tmem_base + d * GEMM_BN, umma_desc_k_sw128(a_addr + j * 32), umma_desc_k_sw128(b_addr + j * 32), idesc, 0u);
                    if (BIAS) mma_i8_ss(tmem_base + d * GEMM_BN, cdesc, cdesc, idesc, 1u);
                    mma_commit(bars + GemmBars::d_full + 8 * d);
                }
                mma_commit(bars + GemmBars::op_empty + 8 * so);   // operands of this stage are free again
            }
        }
      } else {
        // ===== unpack: nibble tile -> K-major int8 A tile with the 128-byte swizzle =====
        // 64 threads x 8 (row, block) items per k-stage; all 8 shared-memory loads are issued before the
        // first is expanded so the two warps run at issue rate, not at LDS + ALU-chain latency
        const int t = threadIdx.x - GEMM_UNPACK_WARP0 * 32;      // 0..63
        const int j = t & 3;                                       // block within the stage
        constexpr int RSTEP = GEMM_UNPACK_WARPS * 8;              // tile rows between this thread's items (16)
        // the same threads stage the weight scales of their items (fp16 -> fp32, x 2^100 when BIAS) for the
        // epilogue; the global loads run one k-stage ahead
        const unsigned short * wd16 = reinterpret_cast<const unsigned short *>(wd);
        const float dscale = BIAS ? 1.2676506002282294e30f : 1.0f;
        unsigned short dnext[GEMM_UNPACK_ROWS];
#pragma unroll
        for (int i = 0; i < GEMM_UNPACK_ROWS; i++) {
            const int m = m0 + (t >> 2) + RSTEP * i;
            dnext[i] = (m < rows && j < nbk) ? ldg_stream_u16(wd16 + (size_t)m * nbk + j) : (unsigned short)0;
        }
        for (int s = 0; s < nstages; s++) {
            const int sr = s % GEMM_NS_RAW, so = s % GEMM_NS_OP, sx = s % GEMM_NS_X;
            if (t == 0) GEMM_TR(3, 4 * s);
            mbar_wait(bars + GemmBars::raw_full + 8 * sr, (s / GEMM_NS_RAW) & 1);
            if (t == 0) GEMM_TR(3, 4 * s + 1);
            const uint32_t raw = sbase + GEMM_SMEM_RAW + sr * GEMM_RAW_BYTES + (t >> 2) * 64 + j * 16;
            uint4 q[GEMM_UNPACK_ROWS];
#pragma unroll
            for (int i = 0; i < GEMM_UNPACK_ROWS; i++) q[i] = lds_u32x4(raw + RSTEP * i * 64);
            if (s >= GEMM_NS_OP) mbar_wait(bars + GemmBars::op_empty + 8 * so, ((s / GEMM_NS_OP) - 1) & 1);
            if (s >= GEMM_NS_X) mbar_wait(bars + GemmBars::x_empty + 8 * sx, ((s / GEMM_NS_X) - 1) & 1);
            if (t == 0) GEMM_TR(3, 4 * s + 2);
            const uint32_t a = sbase + GEMM_SMEM_OP + so * (GEMM_A_BYTES + GEMM_B_BYTES);
            const uint32_t dws = sbase + GEMM_SMEM_X + sx * GEMM_X_BYTES + GEMM_XD_BYTES;
#pragma unroll
            for (int i = 0; i < GEMM_UNPACK_ROWS; i++)
                sts_f32(dws + (j * GEMM_DW_STRIDE + (t >> 2) + RSTEP * i) * 4, half_bits_to_float(dnext[i]) * dscale);
            if (s + 1 < nstages) {
                const int b = (s + 1) * GEMM_KB + j;
#pragma unroll
                for (int i = 0; i < GEMM_UNPACK_ROWS; i++) {
                    const int m = m0 + (t >> 2) + RSTEP * i;
                    dnext[i] = (m < rows && b < nbk) ? ldg_stream_u16(wd16 + (size_t)m * nbk + b) : (unsigned short)0;
                }
            }
#pragma unroll
            for (int i = 0; i < GEMM_UNPACK_ROWS; i++) {
                const int r = (t >> 2) + RSTEP * i;               // tile row
                uint4 lo, hi;
                unpack_q4_word(q[i].x, lo.x, hi.x);
                unpack_q4_word(q[i].y, lo.y, hi.y);
                unpack_q4_word(q[i].z, lo.z, hi.z);
                unpack_q4_word(q[i].w, lo.w, hi.w);
                // K bytes [32j, 32j+16) = elements 0..15 (low nibbles), [32j+16, 32j+32) = elements 16..31
                const uint32_t row = a + r * 128;
                sts_u32x4(row + (((2 * j) ^ (r & 7)) << 4), lo);
                sts_u32x4(row + (((2 * j + 1) ^ (r & 7)) << 4), hi);
            }
            fence_proxy_async_smem();      // generic-proxy stores -> tensor-core (async proxy) reads
            __syncwarp();
            if (t == 0) GEMM_TR(3, 4 * s + 3);
            if (lane == 0) {
                mbar_arrive(bars + GemmBars::op_full + 8 * so);
                mbar_arrive(bars + GemmBars::raw_empty + 8 * sr);
                mbar_arrive(bars + GemmBars::x_full + 8 * sx);
            }
        }
      }
    } else {
        // ===== epilogue: S_b (TMEM) * dw * dx -> fp32 accumulators =====
        // 8 warps = 4 TMEM lane quadrants x 2 column halves; a thread owns one weight row x 64 columns (64 fp32
        // accumulators).  Few fat warps instead of many thin ones: with <= 168 registers the tcgen05.ld of the
        // NEXT 16-column chunk is in flight while the packed FMAs of the current chunk issue, so TMEM latency
        // hides behind the warp's own math (the warps of one scheduler run in lockstep on the same barriers and
        // cannot hide it for each other).
        const int ew = warp - GEMM_EPI_WARP0;          // 0..7 (GEMM_EPI_WARP0 is a multiple of 4)
        const int quad = ew & 3;                        // TMEM lane quadrant (== warp % 4)
        const int col0 = (ew >> 2) * GEMM_EPI_COLS;     // first tile column of this warp
        const int r0 = quad * 32 + (lane >> 2);         // this thread's tile rows: r0 + 8j, j = 0..3
        const int m = m0 + r0;
        const bool cols_live = c0 + col0 < n;           // warp-uniform: any real column in this half?
        const uint32_t t_lane = tmem_base + ((uint32_t)(quad * 32) << 16) + col0;
        // BIAS: accumulator bits = denormal (S_b + 32768) * 2^-149;  2^-134 = 32768 * 2^-149
        const float NEG_OFFSET = BIAS ? -__int_as_float(32768) : -12582912.0f;
        uint64_t acc[GEMM_EPI_COLS / 2];
#pragma unroll
        for (int i = 0; i < GEMM_EPI_COLS / 2; i++) acc[i] = 0ull;
        const GemmDump dump = {S_out, rows, n, nbk, m, c0 + col0, true};

        // block j of a k-stage always lands in TMEM buffer j (GEMM_KB == GEMM_NS_D)
        static_assert(GEMM_KB == GEMM_NS_D, "block j of a k-stage uses TMEM buffer j");
        int sx = 0;
        uint32_t xph = 0;
        if (cols_live) {
            uint32_t al[8], ah[8], bl[8], bh[8];       // two chunk buffers (lanes 0..15 / 16..31 of the quadrant)
            mbar_wait(bars + GemmBars::x_full, 0);
            mbar_wait(bars + GemmBars::d_full, 0);
            tc_fence_after_sync();
            tmem_ld_16x256b_x2(t_lane, al);
            tmem_ld_16x256b_x2(t_lane + (16u << 16), ah);
#pragma unroll 1
            for (int blk = 0; blk < nbk; blk++) {
                const int jj = blk & (GEMM_KB - 1);
                const uint32_t slot = sbase + GEMM_SMEM_X + sx * GEMM_X_BYTES;
                const uint32_t xa = slot + (jj * GEMM_BN + col0 + 2 * (lane & 3)) * 4;
                const uint32_t tb = t_lane + jj * GEMM_BN;
                const uint32_t da = slot + GEMM_XD_BYTES + (jj * GEMM_DW_STRIDE + r0) * 4;
                uint64_t dw2[4], nc2[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const float dw = lds_f32(da + 8 * j * 4);
                    const float nc = NEG_OFFSET * dw;                 // exact: dw has 11 significant bits
                    dw2[j] = pack_f32x2(dw, dw);
                    nc2[j] = pack_f32x2(nc, nc);
                }
                tmem_ld_wait();                                       // a = columns 0..15
                tmem_ld_16x256b_x2(tb + 16, bl);
                tmem_ld_16x256b_x2(tb + 16 + (16u << 16), bh);
                gemm_epi_chunk<0, DUMP, BIAS>(al, ah, acc, xa, dw2, nc2, dump, blk, lane);
                tmem_ld_wait();                                       // b = columns 16..31
                tmem_ld_16x256b_x2(tb + 32, al);
                tmem_ld_16x256b_x2(tb + 32 + (16u << 16), ah);
                gemm_epi_chunk<1, DUMP, BIAS>(bl, bh, acc, xa, dw2, nc2, dump, blk, lane);
                tmem_ld_wait();                                       // a = columns 32..47
                tmem_ld_16x256b_x2(tb + 48, bl);
                tmem_ld_16x256b_x2(tb + 48 + (16u << 16), bh);
                gemm_epi_chunk<2, DUMP, BIAS>(al, ah, acc, xa, dw2, nc2, dump, blk, lane);
                tmem_ld_wait();                                       // b = columns 48..63
                // this warp's slice of the buffer is in registers: hand it back to the MMA issuer
                tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive(bars + GemmBars::d_empty + 8 * jj);
                const bool stage_done = jj == GEMM_KB - 1 || blk + 1 == nbk;
                int sx1 = sx;
                uint32_t xph1 = xph;
                if (blk + 1 < nbk) {
                    if (jj == GEMM_KB - 1) {
                        if (++sx1 == GEMM_NS_X) { sx1 = 0; xph1 ^= 1; }
                        mbar_wait(bars + GemmBars::x_full + 8 * sx1, xph1);   // next stage's scales have landed
                    }
                    const int j1 = (blk + 1) & (GEMM_KB - 1);
                    mbar_wait(bars + GemmBars::d_full + 8 * j1, ((blk + 1) / GEMM_KB) & 1);
                    tc_fence_after_sync();
                    tmem_ld_16x256b_x2(t_lane + j1 * GEMM_BN, al);
                    tmem_ld_16x256b_x2(t_lane + j1 * GEMM_BN + (16u << 16), ah);
                }
                gemm_epi_chunk<3, DUMP, BIAS>(bl, bh, acc, xa, dw2, nc2, dump, blk, lane);
                if (stage_done) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bars + GemmBars::x_empty + 8 * sx);
                }
                sx = sx1;
                xph = xph1;
            }
        } else {
            // no real column in this half: only keep the rings moving (wait before every arrive so an arrival can
            // never land in an earlier phase of the barrier)
            for (int st = 0; st < nstages; st++) {
                mbar_wait(bars + GemmBars::x_full + 8 * sx, xph);
                const int nb = min(GEMM_KB, nbk - st * GEMM_KB);
                for (int jj = 0; jj < nb; jj++) {
                    mbar_wait(bars + GemmBars::d_full + 8 * jj, st & 1);
                    if (lane == 0) mbar_arrive(bars + GemmBars::d_empty + 8 * jj);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(bars + GemmBars::x_empty + 8 * sx);
                if (++sx == GEMM_NS_X) { sx = 0; xph ^= 1; }
            }
        }
        // store: a thread holds rows m + 8j and column pairs; for one (j, pair) the 8 threads with equal i%4 cover
        // 8 consecutive rows = one 32-byte sector per column
        if (cols_live) {
            const float rescale = BIAS ? 562949953421312.0f : 1.0f;   // 2^49
#pragma unroll
            for (int q = 0; q < GEMM_EPI_COLS / 2; q++) {
                const int j = q & 3, k = (q >> 2) & 1, ch = q >> 3;
                float a0, a1;
                unpack_f32x2(acc[q], a0, a1);
                const int c = c0 + col0 + ch * 16 + 8 * k + 2 * (lane & 3);
                const int mm = m + 8 * j;
                if (mm < rows) {
                    if (c < n)     y[(size_t)c * ldy + mm] = a0 * rescale;
                    if (c + 1 < n) y[(size_t)(c + 1) * ldy + mm] = a1 * rescale;
                }
            }
        }
    }

    tc_fence_before_sync();
    __syncthreads();
    if (warp == GEMM_MMA_WARP) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, 512);
    }
    if (independent) pdl_wait();
}

}  // namespace b200
