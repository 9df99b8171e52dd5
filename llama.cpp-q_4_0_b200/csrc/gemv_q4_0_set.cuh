// gemv_q4_0_set.cuh -- kernel 2 for a whole LIST of n = 1 mul_mats in ONE persistent launch
// ("one launch over several weights"; the decode step of a model is such a list).
//
// Why: one launch per mul_mat costs ~1.9 us of dependent-launch boundary + x quantize + first-load latency per op
// (profiles/r01_experiments.md), more than the 1.45 us of HBM traffic of a 4096x4096 op.  Here the weight stream is
// decoupled from the op boundaries:
//   * grid = one CTA per SM, resident for the whole list (cooperative launch);
//   * ONE producer thread per CTA walks the CTA's tiles of ALL ops in list order and copies them with
//     cp.async.bulk (TMA engine, UBLKCP) into a shared-memory ring of NS slots, completion on an mbarrier per slot.
//     Weights are immutable, so the producer never waits for an op to finish: while the consumers sit in a
//     dependency barrier or quantize the next x, up to ~190 KB per SM (28 MB per GPU) of the NEXT ops' weights
//     are already in flight, and HBM never idles across op boundaries;
//   * W consumer warps in groups of 4; slot j belongs to group j % NG.  A tile is R = 4*RPW consecutive weight rows
//     (nibble plane rows + scale plane rows, both contiguous in the plane layout); warp i of the group owns rows
//     i*RPW .. i*RPW+RPW-1 of the tile, lane l owns blocks l, l+32, ... of its rows: the arithmetic, its order and
//     therefore the result bits are those of gemv_q4_0_fused_kernel (S_b exact on IDP.4A, per-lane fmaf chain
//     over b, xor-shuffle tree);
//   * tiles are dealt round-robin over the CTAs CONTINUING across ops, so the per-op remainder does not idle SMs;
//   * ops that read the same x as their predecessor and write rows nobody of the run touches (wq/wk/wv, w3/w1)
//     run back to back; every other op starts with a grid-wide barrier (all its predecessors complete, results
//     visible) and a re-quantization of x into the shared-memory planes (quantize_row_q8_0 contract,
//     ggml.c:1192-1250, same code as the fused kernel).  Semantics == calling ggml_b200_mul_mat_q4_0 per op in order.
// Roofline: HBM; algorithmic bytes per launch = sum over ops of 18*M*K/32 + 4*K + 4*M.
#pragma once
#include "common.cuh"
#include "gemv_q4_0.cuh"
#include "gemv_q4_0_fused.cuh"
#include "ptx.cuh"

namespace b200 {

constexpr int SET_WPG = 4;            // consumer warps per group (= rows of a tile / RPW)
constexpr int SET_MAX_SLOTS = 16;

enum { SET_OP_BARRIER = 1, SET_OP_QUANT = 2 };

struct SetOp {                        // 64 bytes, device resident, read-only during the launch
    const uint4 * wq;                 // nibble plane of this shard [rows][nbk]
    const unsigned short * wd;        // scale plane [rows][nbk] (fp16 bits)
    const float * x;                  // K floats, 16-byte aligned
    float * y;                        // first output row of this shard
    int * S;                          // parity checkpoint: [rows][nbk] int32 block dots, or nullptr
    int rows, nbk;
    int rpw;                          // rows per warp per tile (1 or 2); tile = 4*rpw rows
    int ntiles;
    int first_mod;                    // (tiles of all earlier ops) % grid: CTA c takes tiles (c - first_mod) mod grid, + grid, ...
    int flags;                        // SET_OP_BARRIER | SET_OP_QUANT
};

struct SetSync {                      // device resident, zero between launches
    unsigned arrived;                 // dependency barriers: monotonically counted arrivals of this launch
    unsigned done;                    // CTAs that finished the list; the last one re-zeroes both words
};

__device__ __forceinline__ unsigned ld_acquire_gpu_u32(const unsigned * p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_gpu_add_u32(unsigned * p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <int NCONS>
__device__ __forceinline__ void set_consumer_sync() {
    asm volatile("bar.sync 1, %0;" ::"n"(NCONS) : "memory");
}

// all consumer threads of every CTA; `target` = arrivals that complete this barrier
template <int NCONS>
__device__ __forceinline__ void set_grid_barrier(SetSync * sync, unsigned target) {
    set_consumer_sync<NCONS>();       // this CTA's y stores of the finished ops are issued
    if (threadIdx.x == 0) {
        __threadfence();
        red_release_gpu_add_u32(&sync->arrived, 1u);
        unsigned long long t0 = 0;
        while ((int)(ld_acquire_gpu_u32(&sync->arrived) - target) < 0) {
            const unsigned long long t1 = gtime();
            if (t0 == 0) t0 = t1;
            if (t1 - t0 > 10000000000ull) __trap();   // a CTA of the grid is not resident / died: fail, do not hang
        }
        __threadfence();
    }
    set_consumer_sync<NCONS>();
}

template <int RPW, bool DUMP>
__device__ __forceinline__ void set_tile_rows(const unsigned char * __restrict__ nib, const unsigned char * __restrict__ sc,
                                              const GemvSmem<1> & xs, int nbk, int lane, int nvalid, float * __restrict__ yrow,
                                              int * __restrict__ Srow, uint32_t empty_bar) {
    float acc[RPW];
#pragma unroll
    for (int k = 0; k < RPW; k++) acc[k] = 0.0f;
#pragma unroll 2
    for (int b = lane; b < nbk; b += 32) {
        const int2 ds = xs.ds[b];
        const uint4 ylo = xs.lo[b], yhi = xs.hi[b];
        const float dxf = __int_as_float(ds.x);
#pragma unroll
        for (int k = 0; k < RPW; k++) {
            const uint4 q = *reinterpret_cast<const uint4 *>(nib + ((size_t)k * nbk + b) * 16);
            const unsigned short dh = *reinterpret_cast<const unsigned short *>(sc + ((size_t)k * nbk + b) * 2);
            const int sb = block_dot_q4_q8_fast(q, ylo, yhi, ds.y);
            if (DUMP) { if (Srow != nullptr && k < nvalid) Srow[(size_t)k * nbk + b] = sb; }
            acc[k] = fmaf((float)sb, half_bits_to_float(dh) * dxf, acc[k]);
        }
    }
    // the slot's bytes of this warp's rows are in registers: hand the slot back before the reduction
    __syncwarp();
    if (lane == 0) mbar_arrive(empty_bar);
#pragma unroll
    for (int k = 0; k < RPW; k++) {
        const float t = warp_sum(acc[k]);
        if (lane == 0 && k < nvalid) yrow[k] = t;
    }
}

template <int W, bool DUMP>
__global__ void __launch_bounds__(W * 32 + 32, 1)
gemv_q4_0_set_kernel(const SetOp * __restrict__ ops, int nops, SetSync * __restrict__ sync, int nslots, int slot_bytes,
                     int x_bytes, unsigned seq) {
    constexpr int NCONS = W * 32;
    constexpr int NG = W / SET_WPG;
    extern __shared__ __align__(128) unsigned char smem[];
    // [x planes: x_bytes][full barriers 16 x 8][empty barriers 16 x 8][ring: nslots x slot_bytes]
    GemvSmem<1> xs(smem, x_bytes / 40);
    uint64_t * bars = reinterpret_cast<uint64_t *>(smem + x_bytes);
    unsigned char * ring = smem + x_bytes + 2 * SET_MAX_SLOTS * 8;
    const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + SET_MAX_SLOTS);
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const int G = gridDim.x;

    if (tid == 0) {
        for (int s = 0; s < nslots; s++) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, SET_WPG);
        }
        mbar_fence_init();
    }
    __syncthreads();
    CtaTrace tr;
    if (tid == 0) tr.begin(seq);

    if (warp == W) {
        // ===== producer: streams this CTA's tiles of every op, in list order, through the ring =====
        if (lane == 0) {
            int j = 0;
            for (int o = 0; o < nops; o++) {
                const SetOp op = ops[o];
                const int R = SET_WPG * op.rpw;
                const unsigned char * nib_base = reinterpret_cast<const unsigned char *>(op.wq);
                const unsigned char * sc_base = reinterpret_cast<const unsigned char *>(op.wd);
                int first = (int)blockIdx.x - op.first_mod;
                if (first < 0) first += G;
                for (int t = first; t < op.ntiles; t += G, j++) {
                    const int s = j % nslots;
                    if (j >= nslots) mbar_wait(empty0 + 8 * s, ((j / nslots) - 1) & 1);
                    const int row0 = t * R;
                    const int nr = min(R, op.rows - row0);
                    const uint32_t nib_sz = (uint32_t)nr * op.nbk * 16;
                    const uint32_t sc_sz = ((uint32_t)nr * op.nbk * 2 + 15u) & ~15u;   // the scale plane is allocated with 16 B of slack
                    const uint32_t dst = smem_u32(ring + (size_t)s * slot_bytes);
                    mbar_expect_tx(full0 + 8 * s, nib_sz + sc_sz);
                    bulk_g2s(dst, nib_base + (size_t)row0 * op.nbk * 16, nib_sz, full0 + 8 * s);
                    bulk_g2s(dst + (uint32_t)R * op.nbk * 16, sc_base + (size_t)row0 * op.nbk * 2, sc_sz, full0 + 8 * s);
                }
            }
        }
        return;
    }

    // ===== consumers =====
    const int group = warp / SET_WPG;
    const int wig = warp - group * SET_WPG;
    int j = 0;
    unsigned nbar = 0;
    for (int o = 0; o < nops; o++) {
        const SetOp op = ops[o];
        if (op.flags & SET_OP_BARRIER) {
            nbar++;
            set_grid_barrier<NCONS>(sync, nbar * (unsigned)G);
        }
        if (op.flags & SET_OP_QUANT) {
            if (!(op.flags & SET_OP_BARRIER)) set_consumer_sync<NCONS>();   // everybody is done with the old planes
            // x may have been written by other SMs during this launch: L2-coherent loads (mode 2)
            quantize_x_to_smem<false>(xs, op.x, op.nbk, 2, warp, lane, nullptr, NCONS);
            set_consumer_sync<NCONS>();
        }
        const int R = SET_WPG * op.rpw;
        int first = (int)blockIdx.x - op.first_mod;
        if (first < 0) first += G;
        for (int t = first; t < op.ntiles; t += G, j++) {
            if (j % NG != group) continue;
            const int s = j % nslots;
            const int row0 = t * R + wig * op.rpw;
            const int nvalid = min(op.rpw, op.rows - row0);
            const unsigned char * slot = ring + (size_t)s * slot_bytes;
            const unsigned char * nib = slot + (size_t)wig * op.rpw * op.nbk * 16;
            const unsigned char * sc = slot + (size_t)R * op.nbk * 16 + (size_t)wig * op.rpw * op.nbk * 2;
            mbar_wait(full0 + 8 * s, (j / nslots) & 1);
            if (nvalid <= 0) {            // ragged last tile: nothing for this warp, still release the slot
                if (lane == 0) mbar_arrive(empty0 + 8 * s);
                continue;
            }
            int * Srow = (DUMP && op.S != nullptr) ? op.S + (size_t)row0 * op.nbk : nullptr;
            if (op.rpw == 2) set_tile_rows<2, DUMP>(nib, sc, xs, op.nbk, lane, nvalid, op.y + row0, Srow, empty0 + 8 * s);
            else             set_tile_rows<1, DUMP>(nib, sc, xs, op.nbk, lane, nvalid, op.y + row0, Srow, empty0 + 8 * s);
        }
    }
    if (tid == 0) tr.mark(3);

    // leave the sync words zero for the next launch: the last CTA to finish has seen every barrier complete
    set_consumer_sync<NCONS>();
    if (tid == 0) {
        __threadfence();
        const unsigned old = atomicAdd(&sync->done, 1u);
        if (old == (unsigned)G - 1u) {
            atomicExch(&sync->arrived, 0u);
            atomicExch(&sync->done, 0u);
        }
    }
}

}  // namespace b200
