// codec_q4_0.cuh -- the weight-side Q4_0 codec on the device: quantize_row_q4_0, dequantize_row_q4_0
// and get_rows on a Q4_0 tensor.  HBM-bound streaming kernels; every output is bit-exact against the reference.
//
//   quantize   ggml.c:918-957 (quantize_row_q4_0_reference; quantize_row_q4_0 and ggml_quantize_q4_0 :19157-19178
//              call it): per block  max = the element of largest |x| (the FIRST one on ties, signed),
//              d = max / -8, id = d ? 1.0f/d : 0, q = MIN(15, (int8_t)(x*id + 8.5f)), qs[j] = q[j] | q[j+16] << 4,
//              block.d = fp16(d) (RNE).  All fp32, no contraction: __fdiv_rn / __fmul_rn / __fadd_rn.
//   dequantize ggml.c:1500-1518: y[j] = ((qs[j] & 0xF) - 8) * fp32(d), y[j+16] = ((qs[j] >> 4) - 8) * fp32(d).
//   get_rows   ggml.c:11868-11890 (ggml_compute_forward_get_rows_q): dequantize row ids[i] of the tensor into
//              dst row i; here straight from an uploaded weight in either HBM layout.
#pragma once
#include "common.cuh"

namespace b200 {

// 4 threads per block, 8 consecutive floats each (the pattern of quantize_q8_0.cuh); blocks dense [nrows][nbk]
__global__ void __launch_bounds__(256)
quantize_q4_0_kernel(const float * __restrict__ x, long long nrows, int K, size_t row_stride,
                     unsigned short * __restrict__ out) {
    const int nbk = K / QK;
    const long long total = nrows * nbk * 4;
    const unsigned lane = threadIdx.x & 31;
    for (long long base = (long long)blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < total; base += (long long)gridDim.x * blockDim.x) {
        long long item = base + lane;
        const bool active = item < total;
        if (!active) item = total - 1;
        const long long blk = item >> 2;
        const int part = (int)(item & 3);
        const long long row = blk / nbk;
        const int b = (int)(blk - row * nbk);
        const float * src = reinterpret_cast<const float *>(reinterpret_cast<const char *>(x) + row * row_stride) + b * QK + 8 * part;
        float v[8];
        if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
            const float4 a = __ldg(reinterpret_cast<const float4 *>(src));
            const float4 c = __ldg(reinterpret_cast<const float4 *>(src) + 1);
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = __ldg(src + i);
        }
        // largest |x|, first index on ties (ggml.c:928-934: `if (amax < fabsf(v))`, strict)
        float amax = 0.0f, vmax = 0.0f;
        int imax = 32;                      // "none yet": an all-zero block keeps max = 0
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const float a = fabsf(v[i]);
            if (amax < a) { amax = a; vmax = v[i]; imax = 8 * part + i; }
        }
#pragma unroll
        for (int off = 1; off <= 2; off <<= 1) {
            const float oa = __shfl_xor_sync(0xffffffffu, amax, off);
            const float ov = __shfl_xor_sync(0xffffffffu, vmax, off);
            const int oi = __shfl_xor_sync(0xffffffffu, imax, off);
            if (oa > amax || (oa == amax && oi < imax)) { amax = oa; vmax = ov; imax = oi; }
        }
        const float d = __fdiv_rn(vmax, -8.0f);
        const float id = (d != 0.0f) ? __fdiv_rn(1.0f, d) : 0.0f;
        uint32_t nib = 0;                   // this thread's 8 quants, 4 bits each
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const float t = __fadd_rn(__fmul_rn(v[i], id), 8.5f);
            int qi = (int)(signed char)(int)t;          // (int8_t)(x0 + 8.5f): truncation toward zero
            qi = qi < 15 ? qi : 15;
            nib |= (uint32_t)(qi & 0xF) << (4 * i);
        }
        // byte j = q[j] | q[j+16] << 4: parts 0,1 hold q[0..15], parts 2,3 hold q[16..31]
        const uint32_t hi = __shfl_down_sync(0xffffffffu, nib, 2);
        if (active && part < 2) {
            unsigned short * dst = out + blk * (Q4_0_BLOCK_BYTES / 2);
            if (part == 0) dst[0] = __half_as_ushort(__float2half_rn(d));
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const uint32_t l0 = (nib >> (8 * k)) & 0xF, l1 = (nib >> (8 * k + 4)) & 0xF;
                const uint32_t h0 = (hi >> (8 * k)) & 0xF, h1 = (hi >> (8 * k + 4)) & 0xF;
                dst[1 + 4 * part + k] = (unsigned short)((l0 | (h0 << 4)) | ((l1 | (h1 << 4)) << 8));
            }
        }
    }
}

// 8 threads per block: thread j owns bytes qs[2j], qs[2j+1] -> y[2j], y[2j+1], y[16+2j], y[17+2j]
__device__ __forceinline__ void dequant_store(float * __restrict__ y, int j, unsigned short two_bytes, float d) {
    const int b0 = two_bytes & 0xFF, b1 = two_bytes >> 8;
    const float2 lo = make_float2((float)((b0 & 0xF) - 8) * d, (float)((b1 & 0xF) - 8) * d);
    const float2 hi = make_float2((float)((b0 >> 4) - 8) * d, (float)((b1 >> 4) - 8) * d);
    if ((reinterpret_cast<uintptr_t>(y) & 7) == 0) {
        *reinterpret_cast<float2 *>(y + 2 * j) = lo;
        *reinterpret_cast<float2 *>(y + 16 + 2 * j) = hi;
    } else {
        y[2 * j] = lo.x; y[2 * j + 1] = lo.y; y[16 + 2 * j] = hi.x; y[17 + 2 * j] = hi.y;
    }
}

__global__ void __launch_bounds__(256)
dequantize_q4_0_kernel(const unsigned short * __restrict__ blocks, long long nblocks, float * __restrict__ y) {
    const long long total = nblocks * 8;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long blk = i >> 3;
        const int j = (int)(i & 7);
        const unsigned short * p = blocks + blk * (Q4_0_BLOCK_BYTES / 2);
        dequant_store(y + blk * QK, j, p[1 + j], __half2float(__ushort_as_half(p[0])));
    }
}

// rows ids[0..n) of an uploaded weight -> y[n][K] (row stride ldy floats); ids are rows of the WHOLE tensor, the
// shard must hold them.  RAW: 18-byte blocks; PLANES: nibbles uint4[rows][nbk] + scales half[rows][nbk].
template <bool PLANES>
__global__ void __launch_bounds__(256)
get_rows_q4_0_kernel(const unsigned short * __restrict__ raw, const uint4 * __restrict__ wq, const __half * __restrict__ wd,
                     const int * __restrict__ ids, long long n, int nbk, long long row_lo, long long rows,
                     float * __restrict__ y, long long ldy, int * __restrict__ bad) {
    const long long total = n * nbk * 8;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int j = (int)(i & 7);
        const long long bi = i >> 3;
        const long long r = bi / nbk;
        const int b = (int)(bi - r * nbk);
        const long long row = (long long)ids[r] - row_lo;
        if (row < 0 || row >= rows) { if (j == 0 && b == 0) atomicExch(bad, 1); continue; }
        unsigned short two;
        float d;
        if (PLANES) {
            const unsigned short * q16 = reinterpret_cast<const unsigned short *>(wq + row * nbk + b);
            two = q16[j];
            d = __half2float(wd[row * nbk + b]);
        } else {
            const unsigned short * p = raw + (row * nbk + b) * (Q4_0_BLOCK_BYTES / 2);
            two = p[1 + j];
            d = __half2float(__ushort_as_half(p[0]));
        }
        dequant_store(y + r * ldy + (long long)b * QK, j, two, d);
    }
}

}  // namespace b200
