// gemv_q4_0_neigh.cuh -- kernel 2 (n = 1, kernel 1 fused in) with the op's immediate neighbours in a decode graph
// folded into its store, so that the element-wise node behind a mul_mat costs neither a launch nor a
// trip of the row through L2:
//   NEIGH_ADD     y = W x;  y2 = y + r                 the residual add behind wo / w2 (llama.cpp:1462, :1504;
//                                                       ggml_compute_forward_add_f32, ggml.c:8480-8560) or a bias row
//   NEIGH_SWIGLU  ya = W1 x;  yb = W3 x;  ys = silu(ya);  y2 = ys * yb
//                                                       the LLaMA FFN gate (llama.cpp:1479-1497): ONE kernel reads x once
//                                                       and streams both weights, row i of W1 and of W3 in the same warp;
//                                                       silu with the reference's fp16-table semantics (ggml.c:3530-3538)
// Every intermediate result the graph names is still written (4 bytes per row each), so a node that reads one of them
// later finds it.  Same q8_0 quants, same S_b, same fp32 accumulation order as gemv_q4_0_fused.cuh: y / ya / yb are
// bit-identical to the unfused op, and y2 applies the same single fp32 operation the element-wise kernels apply.
// Roofline: HBM; algorithmic bytes NW * 18*M*K/32 + 4*K + (2 | 4) * 4*M per launch.
#pragma once
#include "common.cuh"
#include "gemv_q4_0.cuh"
#include "gemv_q4_0_fused.cuh"

namespace b200 {

enum { NEIGH_ADD = 1, NEIGH_SWIGLU = 2 };

struct NeighArgs {
    const uint4 * wq[2]; const __half * wd[2];   // weight planes ([1] only for NEIGH_SWIGLU)
    const float * x;                              // activation row (K floats)
    const float * r;                              // NEIGH_ADD: the addend (M floats)
    float * y;                                    // W x        (NEIGH_SWIGLU: W1 x)
    float * yb;                                   // NEIGH_SWIGLU: W3 x
    float * ys;                                   // NEIGH_SWIGLU: silu(W1 x), may be nullptr
    float * y2;                                   // the fused result
    int rows, nbk, x_aligned;
    unsigned seq;
};

__device__ __forceinline__ float silu_f16_table(float v) {      // as eltwise_f32_kernel<2> (elementwise.cuh)
    const float xf = __half2float(__float2half_rn(v));
    return __half2float(__float2half_rn(__fdiv_rn(xf, 1.0f + expf(-xf))));
}

template <int UNROLL, int MAXT, int EPI>
__global__ void __launch_bounds__(MAXT, (2048 / MAXT) / 2)
gemv_q4_0_neigh_kernel(const NeighArgs a) {
    constexpr int NW = EPI == NEIGH_SWIGLU ? 2 : 1;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    GemvSmem<1> s(smem_raw, a.nbk);
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int nwarps = blockDim.x >> 5;
    const int nbk = a.nbk, rows = a.rows;
    CtaTrace tr;
    if (threadIdx.x == 0) tr.begin(a.seq);

    // 1. first chunk of this warp's first row of every weight (no dependency)
    const int rstride = gridDim.x * nwarps;
    const int row_first = blockIdx.x + gridDim.x * warp;
    uint4 q[NW][UNROLL];
    unsigned short dh[NW][UNROLL];
    if (row_first < rows) {
#pragma unroll
        for (int m = 0; m < NW; m++) {
            const uint4 * wrow = a.wq[m] + (size_t)row_first * nbk;
            const unsigned short * drow = reinterpret_cast<const unsigned short *>(a.wd[m]) + (size_t)row_first * nbk;
#pragma unroll
            for (int u = 0; u < UNROLL; u++) {
                const int b = lane + 32 * u;
                if (b < nbk) { q[m][u] = ldg_stream_u4(wrow + b); dh[m][u] = ldg_stream_u16(drow + b); }
            }
        }
    }

    // 2. x, r and the results may belong to the previous kernel
    pdl_wait();
    pdl_launch_dependents();
    if (threadIdx.x == 0) tr.mark(1);

    // 3. quantize_row_q8_0 into the smem planes
    quantize_x_to_smem<false>(s, a.x, nbk, a.x_aligned, warp, lane, nullptr);
    __syncthreads();
    if (threadIdx.x == 0) tr.mark(2);

    // 4. dot loop
    bool prefetched = true;
    for (int row = row_first; row < rows; row += rstride) {
        float acc[NW];
#pragma unroll
        for (int m = 0; m < NW; m++) acc[m] = 0.0f;
        float rv = 0.0f;
        if (EPI == NEIGH_ADD && lane == 0) rv = __ldcg(a.r + row);
        for (int b0 = lane; b0 < nbk; b0 += 32 * UNROLL) {
            if (!prefetched) {
#pragma unroll
                for (int m = 0; m < NW; m++) {
                    const uint4 * wrow = a.wq[m] + (size_t)row * nbk;
                    const unsigned short * drow = reinterpret_cast<const unsigned short *>(a.wd[m]) + (size_t)row * nbk;
#pragma unroll
                    for (int u = 0; u < UNROLL; u++) {
                        const int b = b0 + 32 * u;
                        if (b < nbk) { q[m][u] = ldg_stream_u4(wrow + b); dh[m][u] = ldg_stream_u16(drow + b); }
                    }
                }
            }
            prefetched = false;
#pragma unroll
            for (int u = 0; u < UNROLL; u++) {
                const int b = b0 + 32 * u;
                if (b < nbk) {
                    const int2 ds = s.ds[b];
                    const uint4 ylo = s.lo[b], yhi = s.hi[b];
                    const float dxf = __int_as_float(ds.x);
#pragma unroll
                    for (int m = 0; m < NW; m++) {
                        const int sb = block_dot_q4_q8(q[m][u], ylo, yhi, ds.y);
                        acc[m] = fmaf((float)sb, half_bits_to_float(dh[m][u]) * dxf, acc[m]);
                    }
                }
            }
        }
        float t[NW];
#pragma unroll
        for (int m = 0; m < NW; m++) t[m] = warp_sum(acc[m]);
        if (lane == 0) {
            a.y[row] = t[0];
            if (EPI == NEIGH_ADD) a.y2[row] = __fadd_rn(t[0], rv);
            if (EPI == NEIGH_SWIGLU) {
                a.yb[row] = t[NW - 1];
                const float sv = silu_f16_table(t[0]);
                if (a.ys != nullptr) a.ys[row] = sv;
                a.y2[row] = __fmul_rn(sv, t[NW - 1]);
            }
        }
    }
    if (threadIdx.x == 0) tr.mark(3);
}

}  // namespace b200
