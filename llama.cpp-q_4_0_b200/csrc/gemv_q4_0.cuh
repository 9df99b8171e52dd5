// gemv_q4_0.cuh -- kernel 2: batch-1 (and small-n) Q4_0 x Q8_0 GEMV on dp4a + warp shuffles.
//
// Device replacement of the COMPUTE phase of ggml_compute_forward_mul_mat_q_f32 for small n
// (reference ggml.c:11373-11411: for each weight row, vec_dot_q against each q8_0 column), i.e.
// of ggml_vec_dot_q4_0_q8_0 (ggml.c:2339-2607):
//     y[c][row] = sum_b fp32(dw[row,b]) * fp32(dx[c,b]) * S_b,   S_b = sum_j (q4_j - 8) * q8_j
// S_b is formed exactly in int32 (IDP.4A over the unsigned nibbles, started from -8*sum(q8));
// only the fp32 accumulation order differs from the CPU (lane-strided partial sums + xor
// shuffle tree instead of 8 AVX lanes + hsum).
//
// Work split: one warp per weight row, lane l owns blocks l, l+32, ... of that row, so a warp
// reads 512 contiguous bytes of nibbles + 64 contiguous bytes of scales per step (plane layout)
// and keeps UNROLL independent 16-byte loads per lane in flight.  Rows are dealt round-robin
// over CTAs first (row = cta + grid*slot) so every SM gets floor/ceil(M/148) rows.  The
// quantized activations (K bytes + 8 bytes per block, per column) are staged once per CTA in
// shared memory, split into a low-half plane and a high-half plane so the per-lane 16-byte LDS
// are conflict-free.
//
// Roofline: HBM.  Algorithmic bytes per launch = 18*M*K/32 + 4*K + 4*M (n = 1).
#pragma once
#include "common.cuh"

namespace b200 {

// shared-memory bytes for NC staged columns
__host__ __device__ inline size_t gemv_smem_bytes(int nbk, int nc) { return (size_t)nc * nbk * 40; }

template <int NC>
struct GemvSmem {
    uint4 * lo;   // [NC][nbk]  q8 elements 0..15 of each block
    uint4 * hi;   // [NC][nbk]  q8 elements 16..31
    int2  * ds;   // [NC][nbk]  {bits of d, -8*sum}
    __device__ GemvSmem(unsigned char * base, int nbk) {
        lo = reinterpret_cast<uint4 *>(base);
        hi = lo + (size_t)NC * nbk;
        ds = reinterpret_cast<int2 *>(hi + (size_t)NC * nbk);
    }
};

template <int NC>
__device__ __forceinline__ void gemv_stage_x(GemvSmem<NC> & s, const int8_t * __restrict__ xq,
                                             const int2 * __restrict__ xds, int nbk, int col0, int ncols) {
    const int K = nbk * QK;
    for (int c = 0; c < NC; c++) {
        const bool live = c < ncols;
        const uint4 * src = reinterpret_cast<const uint4 *>(xq + (size_t)(col0 + (live ? c : 0)) * K);
        for (int i = threadIdx.x; i < 2 * nbk; i += blockDim.x) {
            const uint4 v = live ? src[i] : make_uint4(0, 0, 0, 0);
            ((i & 1) ? s.hi : s.lo)[c * nbk + (i >> 1)] = v;
        }
        const int2 * dsrc = xds + (size_t)(col0 + (live ? c : 0)) * nbk;
        for (int i = threadIdx.x; i < nbk; i += blockDim.x) {
            s.ds[c * nbk + i] = live ? dsrc[i] : make_int2(0, 0);
        }
    }
}

// NC columns per pass, UNROLL blocks per lane in flight.  DUMP: also store S_b (parity checkpoint).
template <int NC, int UNROLL, bool DUMP>
__global__ void __launch_bounds__(512)
gemv_q4_0_planes_kernel(const uint4 * __restrict__ wq, const __half * __restrict__ wd,
                        const int8_t * __restrict__ xq, const int2 * __restrict__ xds,
                        float * __restrict__ y, int rows, int nbk, long long ldy, int n,
                        int * __restrict__ S_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    GemvSmem<NC> s(smem_raw, nbk);
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int nwarps = blockDim.x >> 5;
    pdl_wait();   // the quantized activations come from the previous kernel
    pdl_launch_dependents();   // after the wait: see gemv_q4_0_fused.cuh (independent ops)

    for (int col0 = 0; col0 < n; col0 += NC) {
        const int ncols = min(NC, n - col0);
        if (col0 > 0) __syncthreads();
        gemv_stage_x<NC>(s, xq, xds, nbk, col0, ncols);
        __syncthreads();

        for (int row = blockIdx.x + gridDim.x * warp; row < rows; row += gridDim.x * nwarps) {
            const uint4 * wrow = wq + (size_t)row * nbk;
            const unsigned short * drow = reinterpret_cast<const unsigned short *>(wd) + (size_t)row * nbk;
            float acc[NC];
#pragma unroll
            for (int c = 0; c < NC; c++) acc[c] = 0.0f;

            for (int b0 = lane; b0 < nbk; b0 += 32 * UNROLL) {
                uint4 q[UNROLL];
                unsigned short dh[UNROLL];
#pragma unroll
                for (int u = 0; u < UNROLL; u++) {
                    const int b = b0 + 32 * u;
                    if (b < nbk) {
                        q[u]  = ldg_stream_u4(wrow + b);
                        dh[u] = ldg_stream_u16(drow + b);
                    }
                }
#pragma unroll
                for (int u = 0; u < UNROLL; u++) {
                    const int b = b0 + 32 * u;
                    if (b < nbk) {
                        const float dw = half_bits_to_float(dh[u]);
#pragma unroll
                        for (int c = 0; c < NC; c++) {
                            const int2 ds = s.ds[c * nbk + b];
                            const int sb = block_dot_q4_q8(q[u], s.lo[c * nbk + b], s.hi[c * nbk + b], ds.y);
                            if (DUMP) {
                                if (c < ncols) S_out[((size_t)(col0 + c) * rows + row) * nbk + b] = sb;
                            }
                            acc[c] = fmaf((float)sb, dw * __int_as_float(ds.x), acc[c]);
                        }
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < NC; c++) {
                const float r = warp_sum(acc[c]);
                if (lane == 0 && c < ncols) y[(size_t)(col0 + c) * ldy + row] = r;
            }
        }
    }
}

// RAW layout (repack off): the reference's 18-byte blocks, rows contiguous and 4-byte aligned.
// A lane owns a PAIR of blocks = 36 bytes = 9 aligned words:
//   w0 = d0 | qs0[0..1]<<16, w1..w3, w4 = qs0[14..15] | d1<<16, w5..w8 = qs1[0..15]
// so qs0 word j = funnelshift(w[j], w[j+1], 16) and qs1 word j = w[5+j].
template <int NC, bool DUMP>
__global__ void __launch_bounds__(512)
gemv_q4_0_raw_kernel(const uint32_t * __restrict__ wraw,
                     const int8_t * __restrict__ xq, const int2 * __restrict__ xds,
                     float * __restrict__ y, int rows, int nbk, long long ldy, int n,
                     int * __restrict__ S_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    GemvSmem<NC> s(smem_raw, nbk);
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int nwarps = blockDim.x >> 5;
    const int npairs = nbk >> 1;
    const size_t row_words = (size_t)nbk * Q4_0_BLOCK_BYTES / 4;
    pdl_wait();
    pdl_launch_dependents();   // after the wait: see gemv_q4_0_fused.cuh (independent ops)

    for (int col0 = 0; col0 < n; col0 += NC) {
        const int ncols = min(NC, n - col0);
        if (col0 > 0) __syncthreads();
        gemv_stage_x<NC>(s, xq, xds, nbk, col0, ncols);
        __syncthreads();

        for (int row = blockIdx.x + gridDim.x * warp; row < rows; row += gridDim.x * nwarps) {
            const uint32_t * wrow = wraw + (size_t)row * row_words;
            float acc[NC];
#pragma unroll
            for (int c = 0; c < NC; c++) acc[c] = 0.0f;

            for (int p = lane; p < npairs; p += 32) {
                uint32_t w[9];
#pragma unroll
                for (int i = 0; i < 9; i++) w[i] = ldg_stream_u32(wrow + (size_t)p * 9 + i);
                uint4 q0, q1;
                q0.x = __funnelshift_r(w[0], w[1], 16);
                q0.y = __funnelshift_r(w[1], w[2], 16);
                q0.z = __funnelshift_r(w[2], w[3], 16);
                q0.w = __funnelshift_r(w[3], w[4], 16);
                q1 = make_uint4(w[5], w[6], w[7], w[8]);
                const float d0 = half_bits_to_float((unsigned short)(w[0] & 0xFFFF));
                const float d1 = half_bits_to_float((unsigned short)(w[4] >> 16));
                const int b = 2 * p;
#pragma unroll
                for (int c = 0; c < NC; c++) {
                    const int2 ds0 = s.ds[c * nbk + b];
                    const int2 ds1 = s.ds[c * nbk + b + 1];
                    const int s0 = block_dot_q4_q8(q0, s.lo[c * nbk + b], s.hi[c * nbk + b], ds0.y);
                    const int s1 = block_dot_q4_q8(q1, s.lo[c * nbk + b + 1], s.hi[c * nbk + b + 1], ds1.y);
                    if (DUMP) {
                        if (c < ncols) {
                            int * dst = S_out + ((size_t)(col0 + c) * rows + row) * nbk + b;
                            dst[0] = s0;
                            dst[1] = s1;
                        }
                    }
                    acc[c] = fmaf((float)s0, d0 * __int_as_float(ds0.x), acc[c]);
                    acc[c] = fmaf((float)s1, d1 * __int_as_float(ds1.x), acc[c]);
                }
            }
#pragma unroll
            for (int c = 0; c < NC; c++) {
                const float r = warp_sum(acc[c]);
                if (lane == 0 && c < ncols) y[(size_t)(col0 + c) * ldy + row] = r;
            }
        }
    }
}

}  // namespace b200
