// gemm_q4_0_tc3.cuh -- kernel 3: prefill GEMM, Q4_0 weights x q8_0 activations on tcgen05.
//
// Device replacement of the COMPUTE phase of ggml_compute_forward_mul_mat_q_f32 for many columns (reference
// ggml.c:11373-11411, a loop of ggml_vec_dot_q4_0_q8_0 per column; ggml.c:2339-2607):
//     Y[c][m] = sum_b fp32(dw[m,b]) * fp32(dx[c,b]) * S_b[m,c],  S_b = sum_j (q4_j - 8) * q8_j
// One `tcgen05.mma.kind::i8` (M = 128 weight rows, N = 128 activation columns, K = 32 = one Q4_0 block) forms a
// 128x128 int32 tile exactly in TMEM; every block has its own fp16 x fp16 scale pair, so the int32 accumulator cannot
// run across blocks: each block's MMA targets a fresh TMEM buffer (4-deep ring) and the epilogue warps fold
// S_b * dw * dx into fp32 registers.  What bounds this formulation, measured (profiles/r02_ubench_epi_rate.log,
// r02_ubench_mma_rate.log):
//   * the CUDA cores: 2 FP32 FMAs per S_b = 32768 per block; FFMA2 runs at 98-124 fp32 FMA/clk/SM depending on operand
//     shapes: 300 clk per block together with
//   * the TMEM read port: 255 B/clk/SM -> 64 KB of S_b per block = 257 clk,
//   against 128 clk for the block's MMA pair (64 clk per 128x128x32 MMA at any N <= 128): at most ~40 % of the int8 rate
//   for the MMA pair, 25 % counting the data MMA alone.  (One FMA per S_b with the rank-1 scale tile dw x dx formed by an
//   extra MMA needs the scale tile from TMEM too: 532 clk per block, measured -- rejected.)  With the kernel's scale loads
//   from shared memory, barrier arrivals and the unpack warps beside it the same epilogue takes 463 clk per block in
//   isolation; the kernel runs at ~640: ncu shows no saturated pipe (tensor 17 %, FMA 28 %, issue 53 % with 8 epilogue
//   warps) -- it is latency bound on the per-warp chain TMEM load -> FMA -> hand-off, and tcgen05.wait::ld waits for ALL of
//   a warp's loads, so only more warps hide it.  Hence:
//   * 16 epilogue warps (4 per SM sub-partition), each 32 lanes x 32 columns, half blocks of 16 registers
//     (tcgen05.ld.16x256b.x4), the load of the next half in flight under the 16 FFMA2 pairs of this one; 96 registers per
//     thread at launch, setmaxnreg 104 (epilogue) / 56 (service warps).  +11-13 % over 8 warps of 64 columns.
//   * the products rn(S_b * dw) are formed IN PLACE in the fragment's own registers (epi_half): +3-12 %.
//   * A operand = 16 * (q - 8): a nibble moved to the HIGH half of its byte with the top bit flipped IS that signed byte, so
//     the nibble -> int8 expansion is 3 instructions per 8 weights instead of 7; the MMA yields 16 * S_b
//     (|16 S_b| <= 520192 < 2^23) and the factor 2^-4 goes into the weight scale for free.
//   * int -> float costs no instruction: a constant-tile MMA (every byte -128: 32 * 128 * 128 = 2^19) makes the
//     accumulator 16 S_b + 2^19 >= 0, whose bits read as fp32 are the DENORMAL (16 S_b + 2^19) * 2^-149, exactly;
//     t = fma(f, dw * 2^96, -2^-130 * dw * 2^96) = rn(S_b * dw) * 2^-49 in ONE rounding; the accumulators run scaled by
//     2^-49 and are rescaled once.  fp32 operation order: y = sum_b rn(rn(S_b * dw) * dx + acc).  (Resetting the
//     accumulators to a float magic number with tcgen05.st instead of the second MMA was measured: the stores contend with
//     the loads, 307 TOPS against 364 -- profiles/r02_bench_gemm_v3_tmem_reset.log.)
//   * scales arrive in the epilogue's own order: kernel 1 writes the activation scales with the columns of every 32-group
//     permuted so that a thread's 8 scales of a block are contiguous (2 LDS.128), the unpack warps stage the weight
//     scales so that a thread's 4 rows are contiguous (1 LDS.128);
//   * the mbarrier probes for the NEXT block are non-blocking (test_wait) at the top of a block and only looked at after
//     its first half; a miss falls back to try_wait;
//   * the MMA warp walks its loop converged and ONE ELECTED lane issues (elect.sync): inside `if (lane == 0) { loop }` ptxas
//     wraps every UTCIMMA in an ELECT / BRA.U.ANY loop;
//   * the service warps have the highest warp ids; the unpack warps fetch their rows' weight scales three k-stages ahead.
//
// CTA = one 128 (rows) x 128 (columns) output tile, 20 warps:
//   warps 0-15  epilogue: warp = (TMEM lane quadrant w % 4, 32-column slice w / 4); a thread owns rows r0 + 8j (j = 0..3) x 4
//               column pairs = 32 fp32 accumulators in the tcgen05 accumulator-fragment layout
//   warp 16     TMA producer (one thread): per k-stage (4 blocks = 128 K-bytes) the 128x64 B nibble tile (plane layout),
//               the 128x128 B int8 activation tile (SWIZZLE_128B, K-major) and the 4x128 activation scales
//   warp 17     TMEM allocator + MMA issuer: per block one MMA + the constant-tile MMA, commit per block
//   warps 18-19 unpack: nibbles -> 16 * (q - 8), K-major with the 128-byte swizzle of the UMMA descriptor; weight scales
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "ptx.cuh"

namespace b200 {
namespace g3 {

constexpr int BM = 128;                 // weight rows per tile  (UMMA M, TMEM lanes)
constexpr int BN = 128;                 // activation columns per tile (UMMA N, TMEM columns per buffer)
constexpr int KB = 4;                   // Q4_0 blocks per k-stage (128 K-bytes of int8 = one swizzle row)
constexpr int NS_RAW = 4;               // nibble-tile ring
constexpr int NS_OP = 4;                // int8 operand ring (A unpacked + B)
constexpr int NS_X = 6;                 // scale ring (the epilogue trails the MMA by up to one stage, the unpack warps lead it by NS_OP)
constexpr int NS_D = 4;                 // TMEM accumulator ring (4 x 128 columns = 512)
constexpr int THREADS = 640;
constexpr int EPI_WARPS = 16, TMA_WARP = 16, MMA_WARP = 17, UNPACK_WARP0 = 18, UNPACK_WARPS = 2;
constexpr int WCOLS = BN / (EPI_WARPS / 4);                // tile columns per epilogue warp (32)
constexpr int NKG = WCOLS / 8;                             // 8-column groups per warp = the .xN of its TMEM loads (4)
constexpr int EPI_REGS = 104, SVC_REGS = 56;              // setmaxnreg: 512 * 104 + 128 * 56 <= 640 * 96 (the launch pool)
constexpr int UNPACK_ROWS = BM / (UNPACK_WARPS * 8);      // (row, block) items per unpack thread and k-stage (8)
static_assert(KB == NS_D, "block j of a k-stage uses TMEM buffer j");

constexpr int RAW_BYTES = BM * KB * 16;          // 8 KB
constexpr int A_BYTES = BM * KB * 32;            // 16 KB
constexpr int B_BYTES = BN * KB * 32;            // 16 KB
constexpr int XD_BYTES = KB * BN * 4;            // 2 KB   activation scales [4][128] f32 (TMA; columns permuted per 64-group)
constexpr int DW_BYTES = KB * BM * 4;            // 2 KB   weight scales [4][quadrant 4][lane/4 8][j 4] f32 (unpack warps)
constexpr int X_BYTES = XD_BYTES + DW_BYTES;     // 4 KB
constexpr int SMEM_OP = 0;                                                       // 1024-aligned
constexpr int SMEM_CONST = SMEM_OP + NS_OP * (A_BYTES + B_BYTES);                // 16 KB of bytes = -128
constexpr int SMEM_RAW = SMEM_CONST + A_BYTES;
constexpr int SMEM_X = SMEM_RAW + NS_RAW * RAW_BYTES;
constexpr int SMEM_BAR = SMEM_X + NS_X * X_BYTES;
constexpr int SMEM_BYTES = SMEM_BAR + 512;
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");

struct Bars {   // byte offsets of the mbarriers inside the barrier block
    static constexpr int raw_full = 0, raw_empty = 8 * NS_RAW;
    static constexpr int op_full = raw_empty + 8 * NS_RAW, op_empty = op_full + 8 * NS_OP;
    static constexpr int x_full = op_empty + 8 * NS_OP, x_empty = x_full + 8 * NS_X;
    static constexpr int d_full = x_empty + 8 * NS_X, d_empty = d_full + 8 * NS_D;
    static constexpr int tmem_slot = d_empty + 8 * NS_D;
};
static_assert(Bars::tmem_slot + 8 <= 512, "barrier block too small");

constexpr int S_SHIFT = 4;                       // the MMA forms 16 * S_b
constexpr int S_BIAS = 1 << 19;                  // constant-tile MMA: 32 * (-128) * (-128)

// where kernel 1 puts the scale of activation column c inside a row of xdT (quantize_q8_0.cuh has the same formula):
// within every group of 32 columns, the 8 columns {8k + 2t, 8k + 2t + 1 : k = 0..3} that epilogue thread t (= lane % 4)
// owns become contiguous
__host__ __device__ constexpr int xd_perm(int c) { return (c & ~31) + (((c & 7) >> 1) << 3) + (((c & 31) >> 3) << 1) + (c & 1); }

// 16 TMEM lanes x 64 columns in the accumulator-fragment layout (16x256b.x8): thread i gets, for column group
// k = 0..7: r[4k+0], r[4k+1] = lane (i/4),     columns 8k + 2(i%4), +1
//           r[4k+2], r[4k+3] = lane (i/4) + 8, the same two columns.
__device__ __forceinline__ void tmem_ld_frag(uint32_t taddr, uint32_t (&r)[4 * NKG]) {
    static_assert(NKG == 4, "one tcgen05.ld.16x256b.x4 per half block and warp");
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x4.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}

// parity checkpoint plumbing of the epilogue (DUMP instantiations only)
struct Dump {
    int * S; int rows, n, nbk, m, c_first;
};

// Multi-GPU prefill (row shards): the finished tile also goes to every peer's copy of dst over NVLink, straight from this
// kernel (comm.cu: "prefill exchange") -- the all-gather that used to follow the GEMM rides under the tiles still being
// computed.  dst[p] = the peer's landing buffer for THIS rank's rows (column stride ld floats); when the CTA is done, one
// release-add per peer on ctr[p] tells the peer's collect kernel that one more tile has landed.
struct PeerOut {
    float * dst[8];
    unsigned * ctr[8];
    long long ld;
    int ndst;
};
constexpr int T_STRIDE = BM + 4;        // words between two columns of the staged tile: 8 rows x 4 column pairs hit 32 banks

// One half block: H = 0 -> the thread's rows r0, r0 + 8 (TMEM lanes 0..15 of the quadrant), H = 1 -> rows r0 + 16, r0 + 24;
// 8 column pairs each.  acc[k * 4 + j] = (column pair k, row j).
template <int H, bool DUMP, int WFMT>
__device__ __forceinline__ void epi_half(uint32_t (&r)[4 * NKG], uint64_t (&acc)[4 * NKG], const uint64_t (&dx2)[NKG],
                                         const float (&dw)[4], const float (&nc)[4], const Dump & dp, int blk, int lane) {
    // The products rn(S_b * dw) are formed IN PLACE in the fragment's own register pairs, all of them before the first
    // accumulation: with a separate temporary per product ptxas parks the temporaries in the OTHER fragment's (dead)
    // registers and then has to sink that fragment's tcgen05.ld below these FMAs -- its latency lands on the critical path.
#pragma unroll
    for (int k = 0; k < NKG; k++) {
#pragma unroll
        for (int jj = 0; jj < 2; jj++) {
            const int j = 2 * H + jj;
            const uint32_t f0 = r[4 * k + 2 * jj], f1 = r[4 * k + 2 * jj + 1];
            if (DUMP) {
                const int c = dp.c_first + 8 * k + 2 * (lane & 3);
                const int mm = dp.m + 8 * j;
                if (mm < dp.rows) {
                    int * dst = dp.S + ((size_t)c * dp.rows + mm) * dp.nbk + blk;
                    if (c < dp.n)     dst[0] = ((int)f0 - S_BIAS) >> (WFMT == 8 ? 0 : S_SHIFT);
                    if (c + 1 < dp.n) dst[(size_t)dp.rows * dp.nbk] = ((int)f1 - S_BIAS) >> (WFMT == 8 ? 0 : S_SHIFT);
                }
            }
            fma_inplace_pair_f32x2(r[4 * k + 2 * jj], r[4 * k + 2 * jj + 1], pack_f32x2(dw[j], dw[j]), pack_f32x2(nc[j], nc[j]));   // == rn(S_b * dw) * 2^-49
        }
    }
#pragma unroll
    for (int k = 0; k < NKG; k++) {
#pragma unroll
        for (int jj = 0; jj < 2; jj++) fma_acc_f32x2(acc[k * 4 + 2 * H + jj], pack_u32x2(r[4 * k + 2 * jj], r[4 * k + 2 * jj + 1]), dx2[k]);
    }
}

__device__ __forceinline__ float4 lds_f4(uint32_t addr) { return lds_f32x4(addr); }

// The epilogue of one warp over the whole K loop.
template <bool DUMP, bool TRACE, int WFMT, bool PEERS>
__device__ __forceinline__ void epilogue(uint32_t sbase, uint32_t tmem_base, int ew, int lane, int m0, int c0, int rows, int n, int nbk,
                                         float * __restrict__ y, long long ldy, int * S_out, long long * trc, const PeerOut & po) {
#define G3_TR(idx) do { if (TRACE && trc != nullptr && ew == 0 && lane == 0 && blockIdx.x == 0 && blockIdx.y == 0) trc[2 * 2048 + (idx)] = clock64(); } while (0)
    const uint32_t bars = sbase + SMEM_BAR;
    const int quad = ew & 3;                        // TMEM lane quadrant (== warp % 4)
    const int col0 = (ew >> 2) * WCOLS;             // first tile column of this warp
    const int r0 = quad * 32 + (lane >> 2);         // this thread's tile rows: r0 + 8j, j = 0..3
    const int m = m0 + r0;
    const bool cols_live = c0 + col0 < n;           // warp-uniform: any real column in this half?
    uint32_t t_lane = tmem_base + ((uint32_t)(quad * 32) << 16) + col0;
    // per-thread base addresses, made opaque: ptxas otherwise re-derives them from %tid in every unrolled block
    uint32_t xa0 = sbase + SMEM_X + (col0 + 2 * NKG * (lane & 3)) * 4;                       // this thread's 16 (permuted) column scales of block 0
    uint32_t da0 = sbase + SMEM_X + XD_BYTES + (quad * 32 + (lane >> 2) * 4) * 4;       // its 4 row scales of block 0
    asm volatile("" : "+r"(t_lane), "+r"(xa0), "+r"(da0));
    const int nstages = (nbk + KB - 1) / KB;
    const float NEG_OFFSET = -__int_as_float(S_BIAS);    // -(2^19 * 2^-149) = -2^-130
    uint64_t acc[4 * NKG];
#pragma unroll
    for (int i = 0; i < 4 * NKG; i++) acc[i] = 0ull;
    const Dump dump = {S_out, rows, n, nbk, m, c0 + col0};

    if (cols_live) {
        uint32_t ra[4 * NKG], rb[4 * NKG];                    // the two half blocks (TMEM lanes 0..15 / 16..31 of the quadrant)
        mbar_wait(bars + Bars::x_full, 0);
        mbar_wait(bars + Bars::d_full, 0);
        tc_fence_after_sync();
        tmem_ld_frag(t_lane, ra);
        int sx = 0;
        uint32_t xph = 0;
#pragma unroll 1
        for (int s = 0; s < nstages; s++) {
            const int nb = min(KB, nbk - s * KB);
            const uint32_t xa = xa0 + sx * X_BYTES;
            const uint32_t da = da0 + sx * X_BYTES;
            int sx1 = sx + 1;
            uint32_t xph1 = xph;
            if (sx1 == NS_X) { sx1 = 0; xph1 ^= 1; }
#pragma unroll
            for (int jj = 0; jj < KB; jj++) {
                if (jj < nb) {
                    const int blk = s * KB + jj;
                    const bool has_next = blk + 1 < nbk;
                    constexpr int dummy = 0; (void)dummy;
                    const int j1 = (jj + 1) & (KB - 1);
                    // the tests for what the NEXT block needs, issued now and looked at after this block's first half
                    const uint32_t d_bar = bars + Bars::d_full + 8 * j1, d_par = (jj == KB - 1 ? s + 1 : s) & 1;
                    const uint32_t x_bar = bars + Bars::x_full + 8 * sx1;
                    bool ok_d = true, ok_x = true;
                    if (has_next) {
                        ok_d = mbar_test_wait(d_bar, d_par);
                        if (jj == KB - 1) ok_x = mbar_test_wait(x_bar, xph1);
                    }
                    G3_TR(8 * blk);
                    if (TRACE && trc != nullptr && lane == 0 && blockIdx.x == 0 && blockIdx.y == 0 && (ew == 3 || ew == 5)) trc[2 * 2048 + 8 * blk + (ew == 3 ? 6 : 7)] = clock64();
                    float dw[4], nc[4];
                    {
                        const float4 w4 = lds_f4(da + jj * (BM * 4));
                        dw[0] = w4.x; dw[1] = w4.y; dw[2] = w4.z; dw[3] = w4.w;
#pragma unroll
                        for (int j = 0; j < 4; j++) nc[j] = NEG_OFFSET * dw[j];      // exact: dw has 11 significant bits
                    }
                    uint64_t dx2[NKG];
#pragma unroll
                    for (int q = 0; q < NKG / 2; q++) {
                        const float4 v = lds_f4(xa + jj * (BN * 4) + 16 * q);
                        dx2[2 * q] = pack_f32x2(v.x, v.y);
                        dx2[2 * q + 1] = pack_f32x2(v.z, v.w);
                    }
                    tmem_ld_wait();                                       // ra = rows r0, r0 + 8
                    G3_TR(8 * blk + 1);
                    tmem_ld_frag(t_lane + jj * BN + (16u << 16), rb);
                    epi_half<0, DUMP, WFMT>(ra, acc, dx2, dw, nc, dump, blk, lane);
                    G3_TR(8 * blk + 2);
                    tmem_ld_wait();                                       // rb = rows r0 + 16, r0 + 24
                    G3_TR(8 * blk + 3);
                    // this warp's slice of the TMEM buffer is in registers: hand it back to the MMA issuer
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bars + Bars::d_empty + 8 * jj);
                    if (has_next) {
                        if (!ok_x) mbar_wait(x_bar, xph1);                // next stage's scales have landed
                        if (!ok_d) mbar_wait(d_bar, d_par);
                        tc_fence_after_sync();
                        tmem_ld_frag(t_lane + j1 * BN, ra);
                    }
                    G3_TR(8 * blk + 4);
                    epi_half<1, DUMP, WFMT>(rb, acc, dx2, dw, nc, dump, blk, lane);
                    G3_TR(8 * blk + 5);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bars + Bars::x_empty + 8 * sx);
            sx = sx1;
            xph = xph1;
        }
    } else {
        // no real column in this half: only keep the rings moving (wait before every arrive so an arrival can never land
        // in an earlier phase of the barrier)
        int sx = 0;
        uint32_t xph = 0;
        for (int st = 0; st < nstages; st++) {
            mbar_wait(bars + Bars::x_full + 8 * sx, xph);
            const int nb = min(KB, nbk - st * KB);
            for (int jj = 0; jj < nb; jj++) {
                mbar_wait(bars + Bars::d_full + 8 * jj, st & 1);
                if (lane == 0) mbar_arrive(bars + Bars::d_empty + 8 * jj);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bars + Bars::x_empty + 8 * sx);
            if (++sx == NS_X) { sx = 0; xph ^= 1; }
        }
    }
    // store, multi-GPU: the tile goes through shared memory (the operand rings are idle now: the last block's commit has arrived),
    // [column][row] with a padded column stride, so that every destination receives whole 512-byte column segments --
    // 16-byte stores, 4 x 128 B per column -- instead of 32-byte pieces: what NVLink wants for the peers' copies.
    if (!PEERS) {
        // single GPU: straight from the registers -- a thread holds rows m + 8j and column pairs; for one (j, pair) the 8
        // threads with equal i%4 cover 8 consecutive rows = one 32-byte sector per column (the staged path below costs 2-6 %
        // when there is nobody to send to: profiles/r02_bench_gemm_staged.log)
        if (cols_live) {
            const float rescale = 562949953421312.0f;   // 2^49
#pragma unroll
            for (int q = 0; q < 4 * NKG; q++) {
                const int j = q & 3, k = q >> 2;
                float v0, v1;
                unpack_f32x2(acc[q], v0, v1);
                const int c = c0 + col0 + 8 * k + 2 * (lane & 3);
                const int mm = m + 8 * j;
                if (mm < rows) {
                    if (c < n)     y[(size_t)c * ldy + mm] = v0 * rescale;
                    if (c + 1 < n) y[(size_t)(c + 1) * ldy + mm] = v1 * rescale;
                }
            }
        }
        return;
    }
    const uint32_t tbase = sbase + SMEM_OP;
    if (cols_live) {
        const float rescale = 562949953421312.0f;   // 2^49
#pragma unroll
        for (int q = 0; q < 4 * NKG; q++) {
            const int j = q & 3, k = q >> 2;
            float v0, v1;
            unpack_f32x2(acc[q], v0, v1);
            const int cl = col0 + 8 * k + 2 * (lane & 3);
            const int rl = r0 + 8 * j;
            sts_f32(tbase + (cl * T_STRIDE + rl) * 4, v0 * rescale);
            sts_f32(tbase + ((cl + 1) * T_STRIDE + rl) * 4, v1 * rescale);
        }
    }
    asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32) : "memory");      // the 16 epilogue warps
    {
        const int te = ew * 32 + lane;
        // (every CTA starts its round of the peers somewhere else, so that no peer is everybody's first target)
        const int rot = po.ndst > 0 ? (int)((blockIdx.x + blockIdx.y) % (unsigned)po.ndst) : 0;
        for (int dd = -1; dd < po.ndst; dd++) {
            const int d = dd < 0 ? -1 : (dd + rot) % po.ndst;
            float * __restrict__ base = d < 0 ? y : po.dst[d];
            const long long ld = d < 0 ? ldy : po.ld;
            const bool vec = ((ld & 3) == 0) && ((reinterpret_cast<uintptr_t>(base + m0) & 15) == 0);
            for (int idx = te; idx < BN * (BM / 4); idx += EPI_WARPS * 32) {
                const int cl = idx >> 5, r4 = (idx & 31) * 4;
                const int c = c0 + cl, mm = m0 + r4;
                if (c >= n || mm >= rows) continue;
                const float4 v = lds_f32x4(tbase + (cl * T_STRIDE + r4) * 4);
                float * out = base + (size_t)c * ld + mm;
                if (vec && mm + 3 < rows) *reinterpret_cast<float4 *>(out) = v;
                else {
                    out[0] = v.x;
                    if (mm + 1 < rows) out[1] = v.y;
                    if (mm + 2 < rows) out[2] = v.z;
                    if (mm + 3 < rows) out[3] = v.w;
                }
            }
        }
        if (po.ndst > 0) __threadfence_system();      // this thread's peer stores are out before the CTA signals (kernel end)
    }
#undef G3_TR
}

// TRACE: CTA (0,0) stamps clock64 at its pipeline hand-offs into trc (tools/trace_gemm.py); compiled out otherwise
// WFMT = 4: Q4_0 weights (nibble plane, unpacked by warps 18-19).  WFMT = 8: Q8_0 weights (a sibling format
// through the same kernel): the int8 plane [rows][K] IS the A operand, so the TMA warp drops the 128 x 128 B tile straight
// into the operand ring (SWIZZLE_128B, like the activations), the MMA forms S_b itself (no factor 16: the staged weight scale
// carries 2^4 instead) and warps 18-19 only stage the weight scales.  |S_b| <= 32 * 128 * 127 < 2^19 as before.
// PEERS: the multi-GPU instantiation (tile staged in shared memory, peer stores, arrival signal).  A template parameter and not
// a run-time branch: with both store paths in one kernel ptxas schedules the K loop 2-7 % slower (A/B on one box, 394 -> 378 TOPS
// at 4096^2), which the single-GPU kernel must not pay.
template <bool DUMP, bool TRACE, int WFMT = 4, bool PEERS = false>
__global__ void __launch_bounds__(THREADS, 1)
gemm_q4_0_tc3_kernel(const __grid_constant__ CUtensorMap tm_nib,   // nibble plane  [rows][nbk*16 B], box 64 B x 128   (WFMT 8: int8 plane [rows][K], box 128 B x 128, SW128)
                     const __grid_constant__ CUtensorMap tm_xq,    // activations   [n][K] int8,      box 128 B x 128, SW128
                     const __grid_constant__ CUtensorMap tm_xd,    // act. scales   [nbk][n_pad] f32 (columns permuted: xd_perm), box 128 x 4
                     const __half * __restrict__ wd,               // weight scales [rows][nbk]
                     float * __restrict__ y, long long ldy, int rows, int n, int nbk,
                     int * __restrict__ S_out, long long * __restrict__ trc,
                     int independent /* reads the activations its predecessor read, writes its own dst: no wait at the top */,
                     const __grid_constant__ PeerOut po /* ndst = 0: single GPU */) {
#define G3K_TR(role, idx) do { if (TRACE && trc != nullptr && blockIdx.x == 0 && blockIdx.y == 0) trc[(role) * 2048 + (idx)] = clock64(); } while (0)
    // no static shared memory in this kernel: the dynamic window starts 1024-byte aligned (checked below), which the
    // 128-byte swizzle atoms need
    extern __shared__ __align__(1024) unsigned char smem[];
    const uint32_t sbase = smem_u32(smem);
    if ((sbase & 1023u) != 0) __trap();
    const uint32_t bars = sbase + SMEM_BAR;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * BM;       // first weight row of the tile
    const int c0 = blockIdx.y * BN;       // first activation column of the tile
    const int nstages = (nbk + KB - 1) / KB;

    if (threadIdx.x == 0) {
        for (int i = 0; i < NS_RAW; i++) { mbar_init(bars + Bars::raw_full + 8 * i, 1); mbar_init(bars + Bars::raw_empty + 8 * i, UNPACK_WARPS); }
        for (int i = 0; i < NS_OP; i++)  { mbar_init(bars + Bars::op_full + 8 * i, WFMT == 8 ? 1 : 1 + UNPACK_WARPS); mbar_init(bars + Bars::op_empty + 8 * i, 1); }
        for (int i = 0; i < NS_X; i++)   { mbar_init(bars + Bars::x_full + 8 * i, 1 + UNPACK_WARPS); mbar_init(bars + Bars::x_empty + 8 * i, EPI_WARPS); }
        for (int i = 0; i < NS_D; i++)   { mbar_init(bars + Bars::d_full + 8 * i, 1); mbar_init(bars + Bars::d_empty + 8 * i, EPI_WARPS); }
        mbar_fence_init();
        tma_prefetch_desc(&tm_nib); tma_prefetch_desc(&tm_xq); tma_prefetch_desc(&tm_xd);
    }
    if (warp == MMA_WARP) {
        tmem_alloc(bars + Bars::tmem_slot, 512);
        tmem_relinquish();
    }
    {
        // constant operand tile: every byte = -128, so a K = 32 MMA of it with itself adds 32 * 128 * 128 = 2^19
        uint4 * cst = reinterpret_cast<uint4 *>(smem + SMEM_CONST);
        for (int i = threadIdx.x; i < A_BYTES / 16; i += THREADS) cst[i] = make_uint4(0x80808080u, 0x80808080u, 0x80808080u, 0x80808080u);
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
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t *>(smem + SMEM_BAR + Bars::tmem_slot);

    // register budget: setmaxnreg at the top of each role's branch (the branches only meet again at the final barrier)
    if (warp < EPI_WARPS) {
        setmaxnreg_inc<EPI_REGS>();
        epilogue<DUMP, TRACE, WFMT, PEERS>(sbase, tmem_base, warp, lane, m0, c0, rows, n, nbk, y, ldy, S_out, trc, po);
    } else if (warp == TMA_WARP) {
        setmaxnreg_dec<SVC_REGS>();
        if (lane == 0) {
            for (int s = 0; s < nstages; s++) {
                const int sr = s % NS_RAW, so = s % NS_OP, sx = s % NS_X;
                if (WFMT == 8) {
                    (void)sr;
                    if (s >= NS_OP) mbar_wait(bars + Bars::op_empty + 8 * so, ((s / NS_OP) - 1) & 1);
                    G3K_TR(0, 3 * s + 1);
                    mbar_expect_tx(bars + Bars::op_full + 8 * so, A_BYTES + B_BYTES);
                    tma_load_2d(sbase + SMEM_OP + so * (A_BYTES + B_BYTES), &tm_nib, s * KB * 32, m0, bars + Bars::op_full + 8 * so);
                    tma_load_2d(sbase + SMEM_OP + so * (A_BYTES + B_BYTES) + A_BYTES, &tm_xq, s * KB * 32, c0, bars + Bars::op_full + 8 * so);
                } else {
                if (s >= NS_RAW) mbar_wait(bars + Bars::raw_empty + 8 * sr, ((s / NS_RAW) - 1) & 1);
                G3K_TR(0, 3 * s);
                mbar_expect_tx(bars + Bars::raw_full + 8 * sr, RAW_BYTES);
                tma_load_2d(sbase + SMEM_RAW + sr * RAW_BYTES, &tm_nib, s * KB * 16, m0, bars + Bars::raw_full + 8 * sr);
                if (s >= NS_OP) mbar_wait(bars + Bars::op_empty + 8 * so, ((s / NS_OP) - 1) & 1);
                G3K_TR(0, 3 * s + 1);
                mbar_expect_tx(bars + Bars::op_full + 8 * so, B_BYTES);
                tma_load_2d(sbase + SMEM_OP + so * (A_BYTES + B_BYTES) + A_BYTES, &tm_xq, s * KB * 32, c0, bars + Bars::op_full + 8 * so);
                }
                if (s >= NS_X) mbar_wait(bars + Bars::x_empty + 8 * sx, ((s / NS_X) - 1) & 1);
                G3K_TR(0, 3 * s + 2);
                mbar_expect_tx(bars + Bars::x_full + 8 * sx, XD_BYTES);
                tma_load_2d(sbase + SMEM_X + sx * X_BYTES, &tm_xd, c0, s * KB, bars + Bars::x_full + 8 * sx);
            }
        }
    } else if (warp == MMA_WARP) {
        setmaxnreg_dec<SVC_REGS>();
        // The whole warp walks the loop converged and one elected lane issues: every operand stays in uniform registers.
        // (Inside `if (lane == 0) { loop }` ptxas wraps each UTCIMMA in an ELECT / BRA.U.ANY loop; the issuing thread's own
        // scalar instructions are what paces this pipeline -- tools/ubench/mma_rate.cu: the MMA pair itself takes 128 clk.)
        const uint32_t idesc = umma_idesc_i8(BM, BN);
        const uint64_t cdesc = umma_desc_k_sw128(sbase + SMEM_CONST);
        int blk = 0;
        for (int s = 0; s < nstages; s++) {
            const int so = s % NS_OP;
            if (lane == 0) G3K_TR(1, 1024 + s);
            mbar_wait(bars + Bars::op_full + 8 * so, (s / NS_OP) & 1);
            if (lane == 0) G3K_TR(1, 512 + s);
            tc_fence_after_sync();
            const uint32_t a_addr = sbase + SMEM_OP + so * (A_BYTES + B_BYTES);
            const uint64_t adesc = umma_desc_k_sw128(a_addr), bdesc = umma_desc_k_sw128(a_addr + A_BYTES);
            const int nb = min(KB, nbk - s * KB);
#pragma unroll
            for (int j = 0; j < KB; j++) {
                if (j < nb) {
                    if (s > 0) mbar_wait(bars + Bars::d_empty + 8 * j, (s - 1) & 1);
                    if (lane == 0) G3K_TR(1, blk + j);
                    tc_fence_after_sync();
                    if (elect_one()) {
                        mma_i8_ss(tmem_base + j * BN, adesc + 2 * j, bdesc + 2 * j, idesc, 0u);     // + 32 bytes >> 4 in the address field
                        mma_i8_ss(tmem_base + j * BN, cdesc, cdesc, idesc, 1u);
                        mma_commit(bars + Bars::d_full + 8 * j);
                    }
                    __syncwarp();
                }
            }
            blk += nb;
            if (elect_one()) mma_commit(bars + Bars::op_empty + 8 * so);   // operands of this stage are free again
            __syncwarp();
        }
    } else {
        setmaxnreg_dec<SVC_REGS>();
        // ===== unpack: nibble tile -> K-major signed-byte A tile (16 * (q - 8)) with the 128-byte swizzle =====
        // 64 threads x 8 (row, block) items per k-stage, in two batches of 4 with the shared-memory loads of a batch
        // issued before its first expansion.
        const int t = threadIdx.x - UNPACK_WARP0 * 32;      // 0..63
        const int j = t & 3;                                  // block within the stage
        constexpr int RSTEP = UNPACK_WARPS * 8;              // tile rows between this thread's items (16)
        const float DSCALE = WFMT == 8 ? 1.2676506002282294e30f : 7.922816251426434e28f;          // 2^96 (x 2^4 when the MMA forms S_b, not 16 S_b)
        // Weight scales: thread t owns tile rows t and t + 64 and fetches their 4 fp16 scales of a k-stage as two
        // 32-bit words per row, THREE stages ahead.  Reads past a row's last block (tail stage) land in the next row or
        // in the plane's 16 bytes of slack and are never used.
        const uint32_t * wd32 = reinterpret_cast<const uint32_t *>(wd);
        const size_t wrow_a = (size_t)min(m0 + t, rows - 1) * (size_t)(nbk >> 1);
        const size_t wrow_b = (size_t)min(m0 + t + 64, rows - 1) * (size_t)(nbk >> 1);
        uint32_t p0[4] = {0, 0, 0, 0}, p1[4] = {0, 0, 0, 0}, p2[4] = {0, 0, 0, 0};
#define G3_LOAD_SCALES(st, v) do { if ((st) < nstages) { \
            (v)[0] = ldg_stream_u32(wd32 + wrow_a + 2 * (st)); (v)[1] = ldg_stream_u32(wd32 + wrow_a + 2 * (st) + 1); \
            (v)[2] = ldg_stream_u32(wd32 + wrow_b + 2 * (st)); (v)[3] = ldg_stream_u32(wd32 + wrow_b + 2 * (st) + 1); } } while (0)
        G3_LOAD_SCALES(0, p0); G3_LOAD_SCALES(1, p1); G3_LOAD_SCALES(2, p2);
        // where this thread's two rows sit in the staged layout [block][quadrant][lane / 4][j]: row r = 32 quad + g + 8 j
        const int ra_ = t, rb_ = t + 64;
        const int ia = (ra_ >> 5) * 32 + (ra_ & 7) * 4 + ((ra_ & 31) >> 3), ib = (rb_ >> 5) * 32 + (rb_ & 7) * 4 + ((rb_ & 31) >> 3);
        for (int s = 0; s < nstages; s++) {
            const int sr = s % NS_RAW, so = s % NS_OP, sx = s % NS_X;
            if (t == 0) G3K_TR(3, 4 * s);
            if (WFMT != 8) mbar_wait(bars + Bars::raw_full + 8 * sr, (s / NS_RAW) & 1);
            if (t == 0) G3K_TR(3, 4 * s + 1);
            const uint32_t raw = sbase + SMEM_RAW + sr * RAW_BYTES + (t >> 2) * 64 + j * 16;
            constexpr int HALF = UNPACK_ROWS / 2;
            uint4 q[HALF];
            if (WFMT != 8) {
#pragma unroll
            for (int i = 0; i < HALF; i++) q[i] = lds_u32x4(raw + RSTEP * i * 64);
            if (s >= NS_OP) mbar_wait(bars + Bars::op_empty + 8 * so, ((s / NS_OP) - 1) & 1);
            }
            if (s >= NS_X) mbar_wait(bars + Bars::x_empty + 8 * sx, ((s / NS_X) - 1) & 1);
            if (t == 0) G3K_TR(3, 4 * s + 2);
            const uint32_t a = sbase + SMEM_OP + so * (A_BYTES + B_BYTES);
            const uint32_t dws = sbase + SMEM_X + sx * X_BYTES + XD_BYTES;
#pragma unroll
            for (int jb = 0; jb < KB; jb++) {
                const unsigned short ha = (unsigned short)(p0[jb >> 1] >> (16 * (jb & 1)));
                const unsigned short hb = (unsigned short)(p0[2 + (jb >> 1)] >> (16 * (jb & 1)));
                sts_f32(dws + (jb * BM + ia) * 4, half_bits_to_float(ha) * DSCALE);
                sts_f32(dws + (jb * BM + ib) * 4, half_bits_to_float(hb) * DSCALE);
            }
#pragma unroll
            for (int i = 0; i < 4; i++) { p0[i] = p1[i]; p1[i] = p2[i]; }
            G3_LOAD_SCALES(s + 3, p2);
            if (WFMT != 8) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
#pragma unroll
                for (int i = 0; i < HALF; i++) {
                    const int r = (t >> 2) + RSTEP * (h * HALF + i);               // tile row
                    uint4 lo, hi;
                    // byte = q_hi << 4 | q_lo.  (nibble << 4) ^ 0x80 as a signed byte = 16 * (nibble - 8)
                    lo.x = ((q[i].x << 4) & 0xF0F0F0F0u) ^ 0x80808080u;  hi.x = (q[i].x & 0xF0F0F0F0u) ^ 0x80808080u;
                    lo.y = ((q[i].y << 4) & 0xF0F0F0F0u) ^ 0x80808080u;  hi.y = (q[i].y & 0xF0F0F0F0u) ^ 0x80808080u;
                    lo.z = ((q[i].z << 4) & 0xF0F0F0F0u) ^ 0x80808080u;  hi.z = (q[i].z & 0xF0F0F0F0u) ^ 0x80808080u;
                    lo.w = ((q[i].w << 4) & 0xF0F0F0F0u) ^ 0x80808080u;  hi.w = (q[i].w & 0xF0F0F0F0u) ^ 0x80808080u;
                    // K bytes [32j, 32j+16) = elements 0..15 (low nibbles), [32j+16, 32j+32) = elements 16..31
                    const uint32_t row = a + r * 128;
                    sts_u32x4(row + (((2 * j) ^ (r & 7)) << 4), lo);
                    sts_u32x4(row + (((2 * j + 1) ^ (r & 7)) << 4), hi);
                }
                if (h == 0) {
#pragma unroll
                    for (int i = 0; i < HALF; i++) q[i] = lds_u32x4(raw + RSTEP * (HALF + i) * 64);
                }
            }
            fence_proxy_async_smem();      // generic-proxy stores -> tensor-core (async proxy) reads
            }
            __syncwarp();
            if (lane == 0) {
                if (WFMT != 8) {
                    mbar_arrive(bars + Bars::op_full + 8 * so);
                    mbar_arrive(bars + Bars::raw_empty + 8 * sr);
                }
                mbar_arrive(bars + Bars::x_full + 8 * sx);
            }
            if (t == 0) G3K_TR(3, 4 * s + 3);
        }
#undef G3_LOAD_SCALES
    }

    tc_fence_before_sync();
    __syncthreads();
    if (warp == MMA_WARP) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, 512);
    }
    if (PEERS && (int)threadIdx.x < po.ndst) asm volatile("red.release.sys.global.add.u32 [%0], 1;" ::"l"(po.ctr[threadIdx.x]) : "memory");   // this tile has landed at peer threadIdx.x
    if (independent) pdl_wait();
#undef G3K_TR
}

}  // namespace g3
}  // namespace b200
