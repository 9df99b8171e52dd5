// gemv_q4_0_stream.cuh -- kernel 2 for n = 1, streaming version: bulk-async (TMA engine) copies
// into a shared-memory ring, mbarrier producer/consumer pipeline, kernel 1 fused in.
//
// Why: the direct-to-register GEMV (gemv_q4_0_fused.cuh) streams at the measured HBM peak in
// steady state but pays ~2-3 us per launch of exposed DRAM latency, quantize bubble and tail,
// which is most of the time of a 4096x4096 (1.4 us of traffic).  Here the weight stream is
// decoupled from the math:
//   * each CTA owns a CONTIGUOUS row range [cta*M/G, (cta+1)*M/G) (balanced to one row), i.e. one
//     contiguous byte range of the nibble plane and one of the scale plane;
//   * one producer thread walks that range in stages of RS rows with cp.async.bulk (UBLKCP) into a
//     ring of up to 8 stages (<= 96 KB), completion on an mbarrier per stage (complete_tx::bytes);
//   * the ring starts filling BEFORE griddepcontrol.wait: weights are never written by a previous
//     kernel, so under programmatic dependent launch the next kernel's first ~96 KB per SM are
//     already in flight / resident while this kernel is still computing, and HBM never idles
//     across kernel boundaries (a CTA uses < half an SM, so two kernels co-reside);
//   * consumer warps work in TEAMS of WPR warps; a team owns whole rows (row = team, team+NT, ..
//     of the CTA's range), lane l of the team owns block columns l, l+32*WPR, ... (at most 4), so
//     the q8_0 quants of its columns live in registers for the whole kernel (quantized by the
//     lane itself from 32 floats of x: quantize_row_q8_0 contract, ggml.c:1192-1250, bit-exact)
//     and shared memory carries weights only: 18 B written + 18 B read per block;
//   * teams run decoupled: each waits on the full-barrier of the stage holding its row and
//     arrives on that stage's empty-barrier when done; there is no CTA-wide barrier.
// S_b is exact int32 (IDP.4A); only the fp32 summation order differs from the CPU
// (ggml.c:2412-2435): per-lane partials over its columns, xor-shuffle tree, then the WPR warp
// partials in a fixed order (deterministic).
// Roofline: HBM; algorithmic bytes per launch 18*M*K/32 + 4*K + 4*M.
#pragma once
#include "common.cuh"
#include "ptx.cuh"
#include "quantize_q8_0.cuh"

namespace b200 {

constexpr int STREAM_WARPS = 8;                        // consumer warps
constexpr int STREAM_CONSUMERS = STREAM_WARPS * 32;
constexpr int STREAM_THREADS = STREAM_CONSUMERS + 32;  // + 1 producer warp
constexpr int STREAM_RING_BYTES = 96 * 1024;
constexpr int STREAM_MAX_STAGES = 8;
constexpr int STREAM_MAX_NCL = 4;

struct StreamPlan {
    int rs;            // rows per stage
    int wpr;           // warps per row (team size): 1, 2, 4, 8
    int ncl;           // block columns per lane (1..4)
    int stages;        // ring depth
    int stage_bytes;   // bytes of one stage buffer (nibbles + scales + slack, 128-byte multiple)
    size_t smem;       // total dynamic shared memory
    bool ok;
};

__host__ inline StreamPlan stream_plan(int nbk) {
    StreamPlan p = {};
    p.wpr = 1;
    while (p.wpr < STREAM_WARPS && (nbk + 32 * p.wpr - 1) / (32 * p.wpr) > STREAM_MAX_NCL) p.wpr <<= 1;
    p.ncl = (nbk + 32 * p.wpr - 1) / (32 * p.wpr);
    p.ok = p.ncl <= STREAM_MAX_NCL;
    const int nt = STREAM_WARPS / p.wpr;      // teams
    int rs = 8;
    while (rs > 1 && rs * nbk * 18 > 21 * 1024) rs >>= 1;
    p.rs = rs;
    p.stage_bytes = (rs * nbk * 18 + 32 + 127) / 128 * 128;
    // two kernels co-reside on an SM under PDL: ring + quantized-x planes (40*nbk B) <= ~110 KB
    int ring = 110 * 1024 - 40 * nbk - 256;
    if (ring > STREAM_RING_BYTES) ring = STREAM_RING_BYTES;
    p.stages = ring / p.stage_bytes;
    if (p.stages > STREAM_MAX_STAGES) p.stages = STREAM_MAX_STAGES;
    // every team must be able to hold a different stage: teams * rows-in-flight fit the ring
    if (p.stages < 2 || p.stages * rs < nt) p.ok = false;
    // ring + barriers (2 * 8 * 8 B) + reduction scratch (2 parities * 8 warps * 4 B, padded)
    p.smem = (size_t)p.stages * p.stage_bytes + 128 + 128 + (size_t)40 * nbk;
    return p;
}

template <int NCL, bool DUMP>
__global__ void __launch_bounds__(STREAM_THREADS, 2)
gemv_q4_0_stream_kernel(const uint4 * __restrict__ wq, const __half * __restrict__ wd,
                        const float * __restrict__ x, float * __restrict__ y, int rows, int nbk,
                        int rs, int wpr, int stages, int stage_bytes, int x_aligned,
                        int * __restrict__ S_out, unsigned char * __restrict__ blocks34, unsigned seq) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    unsigned char * ring = smem;
    uint64_t * bars = reinterpret_cast<uint64_t *>(smem + (size_t)stages * stage_bytes);   // full[0..8), empty[8..16)
    float * red = reinterpret_cast<float *>(smem + (size_t)stages * stage_bytes + 128);    // [2][8 warps]
    unsigned char * xs = smem + (size_t)stages * stage_bytes + 256;                        // quantized x planes, 40*nbk B
    const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + STREAM_MAX_STAGES);

    // this CTA's contiguous row range
    const int r_lo = (int)(((long long)blockIdx.x * rows) / gridDim.x);
    const int r_hi = (int)(((long long)(blockIdx.x + 1) * rows) / gridDim.x);
    const int my_rows = r_hi - r_lo;
    const int nstage_iters = (my_rows + rs - 1) / rs;

    if (tid == 0) {
        for (int s = 0; s < stages; s++) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, rs * wpr);   // every warp of every row of the stage arrives once
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    CtaTrace tr;
    if (tid == 0) tr.begin(seq);
    pdl_launch_dependents();

    if (warp == STREAM_WARPS) {
        // ===== producer: one thread streams [r_lo, r_hi) through the ring; it has no dependency
        // on the previous kernel (weights are immutable), so it does not wait for it =====
        if (lane == 0) {
            const unsigned char * nib_base = reinterpret_cast<const unsigned char *>(wq);
            const unsigned char * sc_base = reinterpret_cast<const unsigned char *>(wd);
            for (int it = 0; it < nstage_iters; it++) {
                const int s = it % stages;
                if (it >= stages) mbar_wait(empty0 + 8 * s, ((it / stages) - 1) & 1);
                const int row0 = r_lo + it * rs;
                const int nr = min(rs, r_hi - row0);
                const size_t nib_off = (size_t)row0 * nbk * 16;
                const uint32_t nib_sz = (uint32_t)nr * nbk * 16;
                const size_t sc_a = (size_t)row0 * nbk * 2, sc_b = sc_a + (size_t)nr * nbk * 2;
                const size_t sc_a16 = sc_a & ~(size_t)15;
                const uint32_t sc_sz = (uint32_t)(((sc_b + 15) & ~(size_t)15) - sc_a16);
                const uint32_t dst = smem_u32(ring + (size_t)s * stage_bytes);
                mbar_expect_tx(full0 + 8 * s, nib_sz + sc_sz);
                bulk_g2s(dst, nib_base + nib_off, nib_sz, full0 + 8 * s);
                bulk_g2s(dst + rs * nbk * 16, sc_base + sc_a16, sc_sz, full0 + 8 * s);
            }
        }
        return;
    }

    // ===== consumers =====
    const int team = warp / wpr;               // wpr is a power of two <= 8
    const int wit = warp - team * wpr;         // warp in team
    const int nteams = STREAM_WARPS / wpr;
    const int tl = 32 * wpr;                   // lanes in a team
    const int tlane = wit * 32 + lane;

    pdl_wait();   // x (and y) may belong to the previous kernel
    if (tid == 0) tr.mark(1);

    // kernel 1, once per CTA: the 8 consumer warps quantize x into shared-memory planes (4 threads
    // per block, 8 floats per thread), then every lane copies ITS columns into registers
    {
        uint4 * slo = reinterpret_cast<uint4 *>(xs);
        uint4 * shi = slo + nbk;
        int2 * sds = reinterpret_cast<int2 *>(shi + nbk);
        uint32_t * lo32 = reinterpret_cast<uint32_t *>(slo);
        uint32_t * hi32 = reinterpret_cast<uint32_t *>(shi);
        const int total = nbk * 4;
        for (int base = warp * 32; base < total; base += STREAM_CONSUMERS) {
            int item = base + lane;
            const bool active = item < total;
            if (!active) item = total - 1;
            const float * src = x + 8 * item;
            float v[8];
            if (x_aligned) {
                const float4 a = __ldg(reinterpret_cast<const float4 *>(src));
                const float4 c = __ldg(reinterpret_cast<const float4 *>(src) + 1);
                v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
            } else {
#pragma unroll
                for (int i = 0; i < 8; i++) v[i] = __ldg(src + i);
            }
            float amax = 0.0f;
#pragma unroll
            for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v[i]));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
            const float d = __fdiv_rn(amax, 127.0f);
            const float id = (amax != 0.0f) ? __fdiv_rn(127.0f, amax) : 0.0f;
            const __half dhalf = __float2half_rn(d);
            int qq[8];
            int sum = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) { qq[i] = q8_round(v[i], id); sum += qq[i]; }
            sum += __shfl_xor_sync(0xffffffffu, sum, 1);
            sum += __shfl_xor_sync(0xffffffffu, sum, 2);
            const uint32_t w0 = (uint32_t)(qq[0] & 0xFF) | ((uint32_t)(qq[1] & 0xFF) << 8) |
                                ((uint32_t)(qq[2] & 0xFF) << 16) | ((uint32_t)(qq[3] & 0xFF) << 24);
            const uint32_t w1 = (uint32_t)(qq[4] & 0xFF) | ((uint32_t)(qq[5] & 0xFF) << 8) |
                                ((uint32_t)(qq[6] & 0xFF) << 16) | ((uint32_t)(qq[7] & 0xFF) << 24);
            if (active) {
                const int b = item >> 2, part = item & 3;
                uint32_t * dst = (part & 2) ? hi32 : lo32;
                *reinterpret_cast<uint2 *>(dst + 4 * b + 2 * (part & 1)) = make_uint2(w0, w1);
                if (part == 0) sds[b] = make_int2(__float_as_int(__half2float(dhalf)), -8 * sum);
                if (DUMP) {
                    if (blocks34 != nullptr && blockIdx.x == 0) {
                        unsigned char * blk = blocks34 + (size_t)b * Q8_0_BLOCK_BYTES;
                        if (part == 0) *reinterpret_cast<unsigned short *>(blk) = __half_as_ushort(dhalf);
                        unsigned short * qs = reinterpret_cast<unsigned short *>(blk + 2 + 8 * part);
                        qs[0] = (unsigned short)(w0 & 0xFFFF); qs[1] = (unsigned short)(w0 >> 16);
                        qs[2] = (unsigned short)(w1 & 0xFFFF); qs[3] = (unsigned short)(w1 >> 16);
                    }
                }
            }
        }
        asm volatile("bar.sync 9, %0;" ::"n"(STREAM_CONSUMERS) : "memory");
    }
    uint4 xlo[NCL], xhi[NCL];
    float dx[NCL];
    int s8[NCL];
#pragma unroll
    for (int j = 0; j < NCL; j++) {
        const int col = tlane + tl * j;
        xlo[j] = make_uint4(0, 0, 0, 0); xhi[j] = make_uint4(0, 0, 0, 0); dx[j] = 0.0f; s8[j] = 0;
        if (col < nbk) {
            const uint4 * slo = reinterpret_cast<const uint4 *>(xs);
            xlo[j] = slo[col];
            xhi[j] = slo[nbk + col];
            const int2 ds = reinterpret_cast<const int2 *>(slo + 2 * nbk)[col];
            dx[j] = __int_as_float(ds.x);
            s8[j] = ds.y;
        }
    }

    if (tid == 0) tr.mark(2);
    int nrow = 0;   // rows this team has finished (parity of the reduction scratch)
    for (int rl = team; rl < my_rows; rl += nteams, nrow++) {
        const int it = rl / rs;
        const int s = it % stages;
        const int r = rl - it * rs;
        const int row0 = r_lo + it * rs;
        const unsigned char * nib = ring + (size_t)s * stage_bytes + (size_t)r * nbk * 16;
        const unsigned char * sc = ring + (size_t)s * stage_bytes + (size_t)rs * nbk * 16 +
                                   (((size_t)row0 * nbk * 2) & 15) + (size_t)r * nbk * 2;
        mbar_wait(full0 + 8 * s, (it / stages) & 1);

        uint4 q[NCL];
        unsigned short dh[NCL];
#pragma unroll
        for (int j = 0; j < NCL; j++) {
            const int col = tlane + tl * j;
            if (col < nbk) {
                q[j] = *reinterpret_cast<const uint4 *>(nib + (size_t)col * 16);
                dh[j] = *reinterpret_cast<const unsigned short *>(sc + (size_t)col * 2);
            }
        }
        float acc = 0.0f;
#pragma unroll
        for (int j = 0; j < NCL; j++) {
            const int col = tlane + tl * j;
            if (col < nbk) {
                const int sb = block_dot_q4_q8_fast(q[j], xlo[j], xhi[j], s8[j]);
                if (DUMP) S_out[(size_t)(r_lo + rl) * nbk + col] = sb;
                acc = fmaf((float)sb, half_bits_to_float(dh[j]) * dx[j], acc);
            }
        }
        // this warp is done with the stage's bytes of this row
        __syncwarp();
        if (lane == 0) mbar_arrive(empty0 + 8 * s);

        acc = warp_sum(acc);
        if (wpr == 1) {
            if (lane == 0) y[r_lo + rl] = acc;
        } else {
            float * rbuf = red + (nrow & 1) * STREAM_WARPS + team * wpr;
            if (lane == 0) rbuf[wit] = acc;
            asm volatile("bar.sync %0, %1;" ::"r"(1 + team), "r"(tl) : "memory");
            if (tlane == 0) {
                float t = 0.0f;
                for (int w = 0; w < wpr; w++) t += rbuf[w];
                y[r_lo + rl] = t;
            }
        }
    }
    if (tid == 0) tr.mark(3);
}

}  // namespace b200
