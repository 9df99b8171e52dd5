// gemv_q4_0_hostio.cuh -- kernel 2 (n = 1) for the reference-facing HOST entry point: one launch takes x from
// pinned host memory and leaves y in pinned host memory; the CPU neither issues copies nor synchronizes a stream.
//
// The ggml hook is synchronous: src1 lives in host memory, dst must be complete in host memory when the hook
// returns (ggml-cuda.cu:2514-2552 does copy / kernel / copy / cudaStreamSynchronize).  For a 1-15 us GEMV those four
// driver calls cost more than the kernel.  Here, per op:
//   CPU   memcpy x -> the library's mapped staging row; ONE kernel launch; spin on a word in mapped memory;
//         memcpy y out of the mapped staging row.
//   GPU   every CTA first requests its first weight row (no dependency on x);
//         CTAs 0..HOSTIO_FETCH-1 pull their slice of x over PCIe into an HBM row, publish it (release counter);
//         every CTA waits for the counter (3.8 us after kernel start, measured), quantizes x into shared memory (quantize_row_q8_0 contract) and runs
//         the dot loop of gemv_q4_0_fused.cuh over a CONTIGUOUS slice of the rows, so that its results are one
//         coalesced store into the mapped staging row.  Every result travels as an 8-byte word
//         (op number << 32 | bits of y): 8-byte words are single-copy atomic, so a word whose number matches IS
//         the datum and the kernel needs no system fence (3.8 us when measured), no ticket and no completion
//         word; the CPU checks the numbers while it copies y out.
// Counters are monotonic: op number `seq` (host-side, this path is never graph-captured) gives the targets.
// Same q8_0 quants, S_b and fp32 order as the fused kernel (bit-identical y).
#pragma once
#include "common.cuh"
#include "gemv_q4_0.cuh"
#include "gemv_q4_0_fused.cuh"

namespace b200 {

constexpr int HOSTIO_FETCH = 8;      // CTAs that fetch x

struct HostIO {
    const float * x_host;            // device view of the mapped staging row holding x
    float * x_dev;                   // HBM row: x as fetched
    unsigned long long * y_host;     // device view of the mapped staging row receiving (op number, y) words
    unsigned * counters;             // HBM: [0] x slices published; u64 [2] start stamp, [3..4] ns totals (profiling)
    unsigned seq;                    // this op's number (1, 2, ...)
    int profile;                     // GGML_B200_VERBOSE: accumulate phase times in counters[4..]
};

__device__ __forceinline__ unsigned ld_acquire_gpu_u32(const unsigned * p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_sys_f32x4(const float * p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}

template <int UNROLL, int MAXT>
__global__ void __launch_bounds__(MAXT, 1)
gemv_q4_0_hostio_kernel(const uint4 * __restrict__ wq, const __half * __restrict__ wd, int rows, int nbk, const __grid_constant__ HostIO io) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    GemvSmem<1> s(smem_raw, nbk);
    float * yres = reinterpret_cast<float *>(smem_raw + gemv_smem_bytes(nbk, 1));
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int nwarps = blockDim.x >> 5;
    const int K = nbk * QK;
    const unsigned short * wd16 = reinterpret_cast<const unsigned short *>(wd);

    // 1. first row of this warp: weights do not depend on x
    const int slice = (rows + gridDim.x - 1) / gridDim.x;
    const int j0 = min((int)blockIdx.x * slice, rows), j1 = min(j0 + slice, rows);
    const int rstride = nwarps;
    const int row_first = j0 + warp;
    uint4 q[UNROLL];
    unsigned short dh[UNROLL];
    if (row_first < j1) {
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            const int b = lane + 32 * u;
            if (b < nbk) { q[u] = ldg_stream_u4(wq + (size_t)row_first * nbk + b); dh[u] = ldg_stream_u16(wd16 + (size_t)row_first * nbk + b); }
        }
    }
    pdl_wait();                      // the HBM rows and counters may still be in use by the previous op's tail
    pdl_launch_dependents();
    unsigned long long * prof = reinterpret_cast<unsigned long long *>(io.counters) + 2;
    if (io.profile && blockIdx.x == 0 && threadIdx.x == 0) prof[0] = gtime();

    {
        // 2. x: pinned host -> HBM by the first HOSTIO_FETCH CTAs (K is a multiple of 64: 16-byte pieces)
        const int nfetch = min((int)gridDim.x, HOSTIO_FETCH);
        if ((int)blockIdx.x < nfetch) {
            const int pieces = K / 4;
            const int per = (pieces + nfetch - 1) / nfetch;
            const int p0 = blockIdx.x * per, p1 = min(p0 + per, pieces);
            for (int p = p0 + threadIdx.x; p < p1; p += blockDim.x)
                reinterpret_cast<float4 *>(io.x_dev)[p] = ld_sys_f32x4(io.x_host + 4 * p);
            __threadfence();
            __syncthreads();
            if (threadIdx.x == 0) atomicAdd(io.counters, 1u);
        }
        if (threadIdx.x == 0) {
            const unsigned target = io.seq * (unsigned)nfetch;
            while ((int)(ld_acquire_gpu_u32(io.counters) - target) < 0) { }
            if (io.profile && blockIdx.x == 0) atomicAdd(prof + 1, gtime() - prof[0]);   // start -> x published
        }
        __syncthreads();
        // 3. quantize_row_q8_0 into the smem planes (x_dev was written by other SMs: read through L2)
        quantize_x_to_smem<false>(s, io.x_dev, nbk, 2, warp, lane, nullptr);
    }
    __syncthreads();

    // 4. dot loop over the slice
    bool prefetched = true;
    for (int row = row_first; row < j1; row += rstride) {
        float acc = 0.0f;
        for (int b0 = lane; b0 < nbk; b0 += 32 * UNROLL) {
            if (!prefetched) {
#pragma unroll
                for (int u = 0; u < UNROLL; u++) {
                    const int b = b0 + 32 * u;
                    if (b < nbk) { q[u] = ldg_stream_u4(wq + (size_t)row * nbk + b); dh[u] = ldg_stream_u16(wd16 + (size_t)row * nbk + b); }
                }
            }
            prefetched = false;
#pragma unroll
            for (int u = 0; u < UNROLL; u++) {
                const int b = b0 + 32 * u;
                if (b < nbk) {
                    const int2 ds = s.ds[b];
                    const int sb = block_dot_q4_q8(q[u], s.lo[b], s.hi[b], ds.y);
                    acc = fmaf((float)sb, half_bits_to_float(dh[u]) * __int_as_float(ds.x), acc);
                }
            }
        }
        const float t = warp_sum(acc);
        if (lane == 0) yres[row - j0] = t;
    }

    // 5. slice -> pinned host, each word tagged with the op number
    __syncthreads();
    const unsigned long long tag = (unsigned long long)io.seq << 32;
    for (int i = threadIdx.x; i < j1 - j0; i += blockDim.x)
        asm volatile("st.relaxed.sys.global.b64 [%0], %1;" ::"l"(io.y_host + j0 + i), "l"(tag | __float_as_uint(yres[i])) : "memory");
    if (io.profile && threadIdx.x == 0) atomicMax(prof + 2, gtime() - *reinterpret_cast<volatile unsigned long long *>(prof));   // start -> stores issued (max over CTAs and ops)
}

}  // namespace b200
