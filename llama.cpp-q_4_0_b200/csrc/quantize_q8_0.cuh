// quantize_q8_0.cuh -- kernel 1: fused activation quantizer (amax + scale + round + pack + sum).
//
// Device replacement of quantize_row_q8_0 as the op's INIT phase runs it on every src1 row
// (reference ggml.c:11353-11367).  The arithmetic contract is the reference's AVX2 branch
// (ggml.c:1192-1250), NOT the scalar _reference (ggml.c:1097-1120):
//     amax = max |x_j|                         (ggml.c:1202-1211)
//     d    = amax / 127.f  -> fp16 RNE         (ggml.c:1214-1215)   IEEE divide, __fdiv_rn
//     id   = amax ? 127.f / amax : 0           (ggml.c:1216)        a second IEEE divide
//     q_j  = int(round_to_nearest_even(x_j*id)) (ggml.c:1220-1233)  __fmul_rn + cvt.rni
// so the 34-byte blocks are bit-identical to the CPU's.  (x_j*id is NaN only when x_j is inf,
// where cvtps_epi32 + saturating packs give -128; mirrored below.)
//
// One thread owns 8 consecutive floats (two 16-byte loads), four threads own a block; amax and
// sum(q) are reduced with two __shfl_xor steps.  Bound: HBM/L2 streaming, 4 B read + ~1.06 B
// written per element.
#pragma once
#include "common.cuh"

namespace b200 {

__device__ __forceinline__ int q8_round(float x, float id) {
    const float v = __fmul_rn(x, id);
    return (v != v) ? -128 : __float2int_rn(v);
}

// ALIGNED: rows are 16-byte aligned (true for every ggml tensor: GGML_MEM_ALIGN 16, K % 32 == 0)
template <bool ALIGNED, bool WRITE_BLOCKS34>
__global__ void __launch_bounds__(256)
quantize_q8_0_kernel(const float * __restrict__ x, long long n, int K, size_t nb11,
                     int8_t * __restrict__ xq, int2 * __restrict__ xds,
                     unsigned char * __restrict__ blocks34,
                     float * __restrict__ xdT, long long n_pad, int xd_perm /* kernel 3's (tc3) column order inside every 32-group (gemm_q4_0_tc3.cuh xd_perm) */) {
    pdl_wait();   // x may be produced, and xq/xds still be read, by the previous kernel
    pdl_launch_dependents();   // after the wait: see gemv_q4_0_fused.cuh (independent ops)
    const int per_row = K >> 3;
    const long long total = n * (long long)per_row;
    long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = t < total;
    if (!active) t = total - 1;  // keep the quad's shuffles convergent; results discarded
    const long long row = t / per_row;
    const int j = (int)(t - row * per_row);

    const float * src = reinterpret_cast<const float *>(reinterpret_cast<const char *>(x) + row * nb11) + 8 * j;
    float v[8];
    if (ALIGNED) {
        const float4 a = __ldcg(reinterpret_cast<const float4 *>(src));
        const float4 b = __ldcg(reinterpret_cast<const float4 *>(src) + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
        v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
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
    const __half dh = __float2half_rn(d);

    int q[8];
    int sum = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) { q[i] = q8_round(v[i], id); sum += q[i]; }
    sum += __shfl_xor_sync(0xffffffffu, sum, 1);
    sum += __shfl_xor_sync(0xffffffffu, sum, 2);

    const uint32_t w0 = (uint32_t)(q[0] & 0xFF) | ((uint32_t)(q[1] & 0xFF) << 8) |
                        ((uint32_t)(q[2] & 0xFF) << 16) | ((uint32_t)(q[3] & 0xFF) << 24);
    const uint32_t w1 = (uint32_t)(q[4] & 0xFF) | ((uint32_t)(q[5] & 0xFF) << 8) |
                        ((uint32_t)(q[6] & 0xFF) << 16) | ((uint32_t)(q[7] & 0xFF) << 24);
    if (!active) return;

    const int nbk = K >> 5;
    const int b = j >> 2;
    *reinterpret_cast<uint2 *>(xq + row * (long long)K + 8 * j) = make_uint2(w0, w1);
    if ((j & 3) == 0) {
        xds[row * nbk + b] = make_int2(__float_as_int(__half2float(dh)), -8 * sum);
        // block-major copy of the scales for the tensor-core GEMM's epilogue: xdT[b][row]
        if (xdT != nullptr) {
            const long long c = row;
            const long long pos = xd_perm ? (c & ~31ll) + (((c & 7) >> 1) << 3) + (((c & 31) >> 3) << 1) + (c & 1) : c;
            xdT[(long long)b * n_pad + pos] = __half2float(dh);
        }
    }
    if (WRITE_BLOCKS34) {
        unsigned char * blk = blocks34 + (row * nbk + b) * (long long)Q8_0_BLOCK_BYTES;
        if ((j & 3) == 0) *reinterpret_cast<unsigned short *>(blk) = __half_as_ushort(dh);
        unsigned short * qs = reinterpret_cast<unsigned short *>(blk + 2 + 8 * (j & 3));  // 2-byte aligned
        qs[0] = (unsigned short)(w0 & 0xFFFF);
        qs[1] = (unsigned short)(w0 >> 16);
        qs[2] = (unsigned short)(w1 & 0xFFFF);
        qs[3] = (unsigned short)(w1 >> 16);
    }
}

}  // namespace b200
