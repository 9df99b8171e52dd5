// repack.cuh -- one-time re-layout of the reference's 18-byte block_q4_0 stream (ggml.c:870-875)
// into the plane layout (common.cuh), and its exact inverse.  Runs once per tensor at load time
// (the seam of ggml_cuda_transform_tensor, ggml-cuda.cu:2766-2809); lossless, bit-for-bit.
#pragma once
#include "common.cuh"

namespace b200 {

// one thread per block; raw rows are 2-byte aligned streams
__global__ void __launch_bounds__(256)
repack_q4_0_to_planes_kernel(const unsigned short * __restrict__ raw, uint4 * __restrict__ wq,
                             __half * __restrict__ wd, long long nblocks) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nblocks) return;
    const unsigned short * p = raw + i * (Q4_0_BLOCK_BYTES / 2);
    unsigned short h[9];
#pragma unroll
    for (int k = 0; k < 9; k++) h[k] = p[k];
    uint4 q;
    q.x = (uint32_t)h[1] | ((uint32_t)h[2] << 16);
    q.y = (uint32_t)h[3] | ((uint32_t)h[4] << 16);
    q.z = (uint32_t)h[5] | ((uint32_t)h[6] << 16);
    q.w = (uint32_t)h[7] | ((uint32_t)h[8] << 16);
    wq[i] = q;
    wd[i] = __ushort_as_half(h[0]);
}

__global__ void __launch_bounds__(256)
unpack_planes_to_q4_0_kernel(const uint4 * __restrict__ wq, const __half * __restrict__ wd,
                             unsigned short * __restrict__ raw, long long nblocks) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nblocks) return;
    const uint4 q = wq[i];
    unsigned short * p = raw + i * (Q4_0_BLOCK_BYTES / 2);
    p[0] = __half_as_ushort(wd[i]);
    p[1] = (unsigned short)(q.x & 0xFFFF); p[2] = (unsigned short)(q.x >> 16);
    p[3] = (unsigned short)(q.y & 0xFFFF); p[4] = (unsigned short)(q.y >> 16);
    p[5] = (unsigned short)(q.z & 0xFFFF); p[6] = (unsigned short)(q.z >> 16);
    p[7] = (unsigned short)(q.w & 0xFFFF); p[8] = (unsigned short)(q.w >> 16);
}

}  // namespace b200
