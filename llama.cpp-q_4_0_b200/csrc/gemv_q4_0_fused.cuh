// gemv_q4_0_fused.cuh -- kernel 2 for n = 1 with kernel 1 fused in: one launch per mul_mat.
//
// Same arithmetic as quantize_q8_0.cuh + gemv_q4_0.cuh (bit-identical q8_0 quants and S_b, same
// fp32 accumulation order as the unfused GEMV), but every CTA quantizes the activation row itself
// straight into the shared-memory planes the dot loop reads (the 4*K bytes of x come from L2), so
// the op is ONE kernel: no q8 round trip through HBM and no second launch.  Order inside a CTA:
//   1. every warp issues the 16-byte weight loads of its first row          (no dependency)
//   2. griddepcontrol.wait, then griddepcontrol.launch_dependents            (x may be produced
//      by the previous kernel; weights never are)
//   3. quantize x -> smem planes (quantize_row_q8_0 contract, ggml.c:1192-1250), __syncthreads
//   4. dot loop over the CTA's rows (row = cta + grid * slot), warp-shuffle reduce, store y
// Roofline: HBM; algorithmic bytes 18*M*K/32 + 4*K + 4*M per launch.
#pragma once
#include "common.cuh"
#include "gemv_q4_0.cuh"
#include "quantize_q8_0.cuh"

namespace b200 {

// kernel 1 inside a CTA: quantize_row_q8_0 (AVX2 contract, ggml.c:1192-1250) of the whole activation row into the
// shared-memory planes the dot loops read; 4 threads per block, 8 floats per thread.  Caller syncs the CTA.
template <bool DUMP>
__device__ __forceinline__ void quantize_x_to_smem(GemvSmem<1> & s, const float * __restrict__ x, int nbk, int x_aligned,
                                                   int warp, int lane, unsigned char * __restrict__ blocks34, int nthreads = 0) {
        const int total = nbk * 4;
        const int stride = nthreads ? nthreads : (int)blockDim.x;   // the threads that take part (warps 0 .. stride/32-1)
        uint32_t * lo32 = reinterpret_cast<uint32_t *>(s.lo);
        uint32_t * hi32 = reinterpret_cast<uint32_t *>(s.hi);
        for (int base = warp * 32; base < total; base += stride) {
            int item = base + lane;
            const bool active = item < total;
            if (!active) item = total - 1;
            const float * src = x + 8 * item;
            float v[8];
            if (x_aligned) {   // 2: x was written by other SMs during this kernel -> L2-coherent loads
                const float4 a = __ldcg(reinterpret_cast<const float4 *>(src));
                const float4 c = __ldcg(reinterpret_cast<const float4 *>(src) + 1);
                v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
            } else {
#pragma unroll
                for (int i = 0; i < 8; i++) v[i] = __ldcg(src + i);
            }
            float amax = 0.0f;
#pragma unroll
            for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v[i]));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
            const float d  = __fdiv_rn(amax, 127.0f);
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
                if (part == 0) s.ds[b] = make_int2(__float_as_int(__half2float(dhalf)), -8 * sum);
                if (DUMP) {
                    if (blocks34 != nullptr && blockIdx.x == 0) {   // reference-format copy for the parity test
                        unsigned char * blk = blocks34 + (size_t)b * Q8_0_BLOCK_BYTES;
                        if (part == 0) *reinterpret_cast<unsigned short *>(blk) = __half_as_ushort(dhalf);
                        unsigned short * qs = reinterpret_cast<unsigned short *>(blk + 2 + 8 * part);
                        qs[0] = (unsigned short)(w0 & 0xFFFF); qs[1] = (unsigned short)(w0 >> 16);
                        qs[2] = (unsigned short)(w1 & 0xFFFF); qs[3] = (unsigned short)(w1 >> 16);
                    }
                }
            }
        }
}

template <int UNROLL, int RIF, int MAXT, bool DUMP>
__global__ void __launch_bounds__(MAXT, (2048 / MAXT) / 2)
gemv_q4_0_fused_kernel(const uint4 * __restrict__ wq, const __half * __restrict__ wd,
                       const float * __restrict__ x, float * __restrict__ y, int rows, int nbk,
                       int x_aligned, int * __restrict__ S_out, unsigned char * __restrict__ blocks34, unsigned seq,
                       int independent) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    GemvSmem<1> s(smem_raw, nbk);
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int nwarps = blockDim.x >> 5;
    CtaTrace tr;
    if (threadIdx.x == 0) tr.begin(seq);
    // `independent`: the host established that this op reads the same x as the library's previous op on this
    // stream and writes rows nobody in flight touches (wq/wk/wv, w1/w3).  It then need not wait for that kernel:
    // everything older is complete, because every kernel of this library releases its dependents only AFTER its
    // own dependency wait.  It waits at its END instead, so that its completion still implies the completion of
    // everything before it (a later kernel's griddepcontrol.wait only covers its immediate predecessor).
    if (independent) pdl_launch_dependents();

    // 1. first chunk of this warp's first RIF rows (RIF rows in flight per warp)
    const int rstride = gridDim.x * nwarps;
    const int row_first = blockIdx.x + gridDim.x * warp;
    uint4 q[RIF][UNROLL];
    unsigned short dh[RIF][UNROLL];
#pragma unroll
    for (int r = 0; r < RIF; r++) {
        const int row = row_first + r * rstride;
        if (row < rows) {
            const uint4 * wrow = wq + (size_t)row * nbk;
            const unsigned short * drow = reinterpret_cast<const unsigned short *>(wd) + (size_t)row * nbk;
#pragma unroll
            for (int u = 0; u < UNROLL; u++) {
                const int b = lane + 32 * u;
                if (b < nbk) { q[r][u] = ldg_stream_u4(wrow + b); dh[r][u] = ldg_stream_u16(drow + b); }
            }
        }
    }

    // 2. activations (and y) may belong to the previous kernel
    if (!independent) { pdl_wait(); pdl_launch_dependents(); }
    if (threadIdx.x == 0) tr.mark(1);

    // 3. quantize_row_q8_0 into the smem planes
    quantize_x_to_smem<DUMP>(s, x, nbk, x_aligned, warp, lane, blocks34);
    __syncthreads();
    if (threadIdx.x == 0) tr.mark(2);

    // 4. dot loop, RIF rows at a time
    bool prefetched = true;
    for (int row0 = row_first; row0 < rows; row0 += RIF * rstride) {
        float acc[RIF];
#pragma unroll
        for (int r = 0; r < RIF; r++) acc[r] = 0.0f;
        for (int b0 = lane; b0 < nbk; b0 += 32 * UNROLL) {
            if (!prefetched) {
#pragma unroll
                for (int r = 0; r < RIF; r++) {
                    const int row = row0 + r * rstride;
                    if (row < rows) {
                        const uint4 * wrow = wq + (size_t)row * nbk;
                        const unsigned short * drow = reinterpret_cast<const unsigned short *>(wd) + (size_t)row * nbk;
#pragma unroll
                        for (int u = 0; u < UNROLL; u++) {
                            const int b = b0 + 32 * u;
                            if (b < nbk) { q[r][u] = ldg_stream_u4(wrow + b); dh[r][u] = ldg_stream_u16(drow + b); }
                        }
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
                    for (int r = 0; r < RIF; r++) {
                        const int row = row0 + r * rstride;
                        if (row < rows) {
                            const int sb = block_dot_q4_q8(q[r][u], ylo, yhi, ds.y);
                            if (DUMP) S_out[(size_t)row * nbk + b] = sb;
                            acc[r] = fmaf((float)sb, half_bits_to_float(dh[r][u]) * dxf, acc[r]);
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int r = 0; r < RIF; r++) {
            const int row = row0 + r * rstride;
            const float t = warp_sum(acc[r]);
            if (lane == 0 && row < rows) y[row] = t;
        }
    }
    if (threadIdx.x == 0) tr.mark(3);
    if (independent) pdl_wait();
}

}  // namespace b200
