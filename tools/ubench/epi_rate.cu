// Microbenchmark (GPU box): what bounds kernel 3's epilogue -- measured in isolation on every SM, no barriers, TMEM
// content arbitrary.  One Q4_0 block of a 128x128 tile = 16384 int32 S_b in TMEM that 8 epilogue warps (2 per SM
// sub-partition, each a 32-lane x 64-column slice) must read and fold into fp32 accumulators.
//   (1) tcgen05.ld throughput: 16x256b.x8 loads (4 KB per warp instruction) back to back, 4 and 8 warps;
//   (2) the two-FMA formulation (kernel 3): t = fma(S, dw, c); acc = fma(t, dx, acc), S read from TMEM;
//   (3) the one-FMA formulation VERDICT r1 asks to measure: acc = fma(S, D, acc) with BOTH S and the rank-1 scale tile
//       D = dw x dx read from TMEM (as an extra MMA would leave it there): half the FMAs, twice the TMEM traffic;
//   (4) FFMA2 operand shapes: (pair, scalar, scalar), three pairs, and the epilogue's alternation of the two.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I ../../llama.cpp-q_4_0_b200/csrc -o epi_rate epi_rate.cu
// Not part of the product.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#include "ptx.cuh"

using namespace b200;

__device__ __forceinline__ void ld_x8(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x8.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
}

__device__ __forceinline__ uint32_t tmem_setup(uint32_t * slot_smem, int warp) {
    if (warp == 0) { tmem_alloc(smem_u32(slot_smem), 512); tmem_relinquish(); }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    return *reinterpret_cast<volatile uint32_t *>(slot_smem);
}
__device__ __forceinline__ void tmem_teardown(uint32_t base, int warp) {
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) { tc_fence_after_sync(); tmem_dealloc(base, 512); }
}

template <int BATCH>
__global__ void __launch_bounds__(256, 1) ldtm_kernel(long long * cyc, unsigned * sink, int iters, int warps) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    const uint32_t base = tmem_setup(&slot, warp);
    unsigned acc = 0;
    long long t0 = 0, t1 = 0;
    if (warp < warps) {
        const uint32_t t_lane = base + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 64;
        t0 = clock64();
        for (int it = 0; it < iters; it++) {
            uint32_t r[BATCH][32];
#pragma unroll
            for (int b = 0; b < BATCH; b++) ld_x8(t_lane + (b & 3) * 128 + ((b & 4) ? (16u << 16) : 0u), r[b]);
            tmem_ld_wait();
#pragma unroll
            for (int b = 0; b < BATCH; b++) acc += r[b][0] ^ r[b][31];
        }
        t1 = clock64();
    }
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    sink[blockIdx.x * 256 + threadIdx.x] = acc;
    tmem_teardown(base, warp);
}

// (2) two FMAs per element, S from TMEM, scales from registers; half a block (32 registers) double buffered.
// mma_gap >= 0: a ninth warp issues kernel 3's MMA pair (128x128x32 int8, data + constant tile) into the buffers being read,
// one pair every max(128, mma_gap) clocks -- does the tensor core's accumulator traffic slow tcgen05.ld down?
__global__ void __launch_bounds__(288, 1) twofma_kernel(long long * cyc, float * y, int nbk, int mma_gap) {
    extern __shared__ __align__(1024) unsigned char opsmem[];
    __shared__ uint32_t slot;
    __shared__ volatile int done;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) done = 0;
    if (mma_gap >= 0) {
        for (int i = threadIdx.x; i < 32768 / 16; i += 288) reinterpret_cast<uint4 *>(opsmem)[i] = make_uint4(i, 0x80808080u, ~i, 0x01020304u);
        fence_proxy_async_smem();
    }
    const uint32_t base = tmem_setup(&slot, warp);
    if (warp == 8) {
        if (mma_gap >= 0) {
            const uint32_t idesc = umma_idesc_i8(128, 128);
            const uint64_t ad = umma_desc_k_sw128(smem_u32(opsmem)), bd = umma_desc_k_sw128(smem_u32(opsmem) + 16384);
            int i = 0;
            while (!done) {
                const long long t = clock64();
                if (elect_one()) {
                    mma_i8_ss(base + (i & 3) * 128, ad + 2 * (i & 3), bd + 2 * (i & 3), idesc, 0u);
                    mma_i8_ss(base + (i & 3) * 128, ad, ad, idesc, 1u);
                }
                __syncwarp();
                i++;
                while (clock64() - t < mma_gap && !done) { }
            }
        }
    } else {
    const uint32_t t_lane = base + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 64;
    uint64_t acc[32];
#pragma unroll
    for (int i = 0; i < 32; i++) acc[i] = 0ull;
    uint64_t dw2[4], nc2[4], dx2[8];
    for (int i = 0; i < 4; i++) { dw2[i] = pack_f32x2(1.0f + i + threadIdx.x, 1.0f + i + threadIdx.x); nc2[i] = pack_f32x2(-1e-30f * i, -1e-30f * i); }
    for (int i = 0; i < 8; i++) dx2[i] = pack_f32x2(0.5f + i, 0.25f + i);
    uint32_t ra[32], rb[32];
    long long t0 = clock64();
    ld_x8(t_lane, ra);
#pragma unroll 1
    for (int blk = 0; blk < nbk; blk++) {
        const uint32_t tb = t_lane + (blk & 3) * 128;
        tmem_ld_wait();
        ld_x8(tb + (16u << 16), rb);
#pragma unroll
        for (int k = 0; k < 8; k++)
#pragma unroll
            for (int jj = 0; jj < 2; jj++) {
                const uint64_t t = fma_f32x2(pack_u32x2(ra[4 * k + 2 * jj], ra[4 * k + 2 * jj + 1]), dw2[jj], nc2[jj]);
                fma_acc_f32x2(acc[k * 4 + jj], t, dx2[k]);
            }
        tmem_ld_wait();
        ld_x8(t_lane + ((blk + 1) & 3) * 128, ra);
#pragma unroll
        for (int k = 0; k < 8; k++)
#pragma unroll
            for (int jj = 0; jj < 2; jj++) {
                const uint64_t t = fma_f32x2(pack_u32x2(rb[4 * k + 2 * jj], rb[4 * k + 2 * jj + 1]), dw2[2 + jj], nc2[2 + jj]);
                fma_acc_f32x2(acc[k * 4 + 2 + jj], t, dx2[k]);
            }
    }
    tmem_ld_wait();
    long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 32; i++) { float a, b; unpack_f32x2(acc[i], a, b); s += a + b; }
    y[blockIdx.x * 256 + threadIdx.x] = s + (float)ra[0];
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    asm volatile("bar.sync 1, 256;" ::: "memory");     // the 8 reading warps
    if (threadIdx.x == 0) done = 1;
    }
    tmem_teardown(base, warp);
}

// (2b) the same epilogue plus, one at a time, what kernel 3 has around it (flags F):
//   1: the block's scales come from shared memory (5 LDS.128 per thread and block)
//   2: tcgen05 fences + __syncwarp + one mbarrier.arrive per warp and block
//   4: two mbarrier.test_wait per block (a completed phase), looked at after the first half
//   8: two more warps run an unpack-like loop (LDS.128, 24 logic ops, 2 STS.128 per item) flat out
template <int F>
__global__ void __launch_bounds__(320, 1) epi_plus_kernel(long long * cyc, float * y, int nbk) {
    extern __shared__ __align__(1024) unsigned char sm[];
    __shared__ uint32_t slot;
    __shared__ __align__(8) unsigned long long bar[2];
    __shared__ volatile int done;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t sb = smem_u32(sm);
    if (threadIdx.x == 0) { done = 0; mbar_init(smem_u32(&bar[0]), 8); mbar_init(smem_u32(&bar[1]), 1); mbar_fence_init(); }
    for (int i = threadIdx.x; i < 49152 / 4; i += 320) reinterpret_cast<float *>(sm)[i] = 1.0f + 1e-3f * (i & 63);
    const uint32_t base = tmem_setup(&slot, warp);
    if (threadIdx.x == 0) mbar_arrive(smem_u32(&bar[1]));
    __syncthreads();
    if (warp >= 8) {
        if (F & 8) {
            const uint32_t src = sb + 16384 + (warp - 8) * 8192 + lane * 16, dst = sb + 32768 + (warp - 8) * 8192 + lane * 32;
            int i = 0;
            while (!done) {
                const uint4 q = lds_u32x4(src + (i & 7) * 512);
                uint4 lo, hi;
                lo.x = ((q.x << 4) & 0xF0F0F0F0u) ^ 0x80808080u; hi.x = (q.x & 0xF0F0F0F0u) ^ 0x80808080u;
                lo.y = ((q.y << 4) & 0xF0F0F0F0u) ^ 0x80808080u; hi.y = (q.y & 0xF0F0F0F0u) ^ 0x80808080u;
                lo.z = ((q.z << 4) & 0xF0F0F0F0u) ^ 0x80808080u; hi.z = (q.z & 0xF0F0F0F0u) ^ 0x80808080u;
                lo.w = ((q.w << 4) & 0xF0F0F0F0u) ^ 0x80808080u; hi.w = (q.w & 0xF0F0F0F0u) ^ 0x80808080u;
                sts_u32x4(dst + (i & 7) * 1024, lo);
                sts_u32x4(dst + (i & 7) * 1024 + 16, hi);
                i++;
            }
        }
    } else {
    const uint32_t t_lane = base + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 64;
    uint32_t xa = sb + ((warp >> 2) * 64 + 16 * (lane & 3)) * 4, da = sb + 2048 + ((warp & 3) * 32 + (lane >> 2) * 4) * 4;
    asm volatile("" : "+r"(xa), "+r"(da));
    uint64_t acc[32];
#pragma unroll
    for (int i = 0; i < 32; i++) acc[i] = 0ull;
    float dw[4], nc[4];
    uint64_t dx2[8];
    for (int i = 0; i < 4; i++) { dw[i] = 1.0f + i + threadIdx.x; nc[i] = -1e-30f * i; }
    for (int i = 0; i < 8; i++) dx2[i] = pack_f32x2(0.5f + i, 0.25f + i);
    uint32_t ra[32], rb[32];
    unsigned int misses = 0;
    long long t0 = clock64();
    ld_x8(t_lane, ra);
#pragma unroll 1
    for (int blk = 0; blk < nbk; blk++) {
        const uint32_t tb = t_lane + (blk & 3) * 128;
        bool ok0 = true, ok1 = true;
        if (F & 4) { ok0 = mbar_test_wait(smem_u32(&bar[1]), 0); ok1 = mbar_test_wait(smem_u32(&bar[1]), 0); }
        if (F & 1) {
            const float4 w4 = lds_f32x4(da + (blk & 3) * 512);
            dw[0] = w4.x; dw[1] = w4.y; dw[2] = w4.z; dw[3] = w4.w;
#pragma unroll
            for (int j = 0; j < 4; j++) nc[j] = -7.3e-40f * dw[j];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const float4 v = lds_f32x4(xa + (blk & 3) * 512 + 16 * q);
                dx2[2 * q] = pack_f32x2(v.x, v.y);
                dx2[2 * q + 1] = pack_f32x2(v.z, v.w);
            }
        }
        tmem_ld_wait();
        ld_x8(tb + (16u << 16), rb);
#pragma unroll
        for (int k = 0; k < 8; k++)
#pragma unroll
            for (int jj = 0; jj < 2; jj++) {
                const uint64_t t = fma_f32x2(pack_u32x2(ra[4 * k + 2 * jj], ra[4 * k + 2 * jj + 1]), pack_f32x2(dw[jj], dw[jj]), pack_f32x2(nc[jj], nc[jj]));
                fma_acc_f32x2(acc[k * 4 + jj], t, dx2[k]);
            }
        tmem_ld_wait();
        if (F & 2) {
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&bar[0]));
        }
        if (F & 4) { if (!ok0 || !ok1) misses++; }
        if (F & 2) tc_fence_after_sync();
        ld_x8(t_lane + ((blk + 1) & 3) * 128, ra);
#pragma unroll
        for (int k = 0; k < 8; k++)
#pragma unroll
            for (int jj = 0; jj < 2; jj++) {
                const uint64_t t = fma_f32x2(pack_u32x2(rb[4 * k + 2 * jj], rb[4 * k + 2 * jj + 1]), pack_f32x2(dw[2 + jj], dw[2 + jj]), pack_f32x2(nc[2 + jj], nc[2 + jj]));
                fma_acc_f32x2(acc[k * 4 + 2 + jj], t, dx2[k]);
            }
    }
    tmem_ld_wait();
    long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 32; i++) { float a, b; unpack_f32x2(acc[i], a, b); s += a + b; }
    y[blockIdx.x * 256 + threadIdx.x] = s + (float)ra[0] + (float)misses;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (threadIdx.x == 0) done = 1;
    }
    tmem_teardown(base, warp);
}

template <int F>
static void run_plus(long long * cyc, float * y, const char * what) {
    const int G = 148, nbk = 512;
    cudaFuncSetAttribute(epi_plus_kernel<F>, cudaFuncAttributeMaxDynamicSharedMemorySize, 50176);
    for (int rep = 0; rep < 2; rep++) {
        epi_plus_kernel<F><<<G, 320, 50176>>>(cyc, y, nbk);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("CUDA error: %s (%s)\n", cudaGetErrorString(e), what); return; }
    }
    static long long h[148];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double c = 0;
    for (int i = 0; i < G; i++) c += (double)h[i];
    printf("  %-72s %6.1f clk per block\n", what, c / G / nbk);
}

// (3) one FMA per element: S and D both from TMEM
__global__ void __launch_bounds__(256, 1) onefma_kernel(long long * cyc, float * y, int nbk) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    const uint32_t base = tmem_setup(&slot, warp);
    const uint32_t t_lane = base + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 64;
    uint64_t acc[32];
#pragma unroll
    for (int i = 0; i < 32; i++) acc[i] = 0ull;
    uint32_t sa[32], da[32], sb[32], db[32];
    long long t0 = clock64();
    ld_x8(t_lane, sa);
    ld_x8(t_lane + 256, da);
#pragma unroll 1
    for (int blk = 0; blk < nbk; blk++) {
        const uint32_t tb = t_lane + (blk & 1) * 128;
        tmem_ld_wait();
        ld_x8(tb + (16u << 16), sb);
        ld_x8(tb + 256 + (16u << 16), db);
#pragma unroll
        for (int k = 0; k < 8; k++)
#pragma unroll
            for (int jj = 0; jj < 2; jj++)
                fma_acc_f32x2(acc[k * 4 + jj], pack_u32x2(sa[4 * k + 2 * jj], sa[4 * k + 2 * jj + 1]), pack_u32x2(da[4 * k + 2 * jj], da[4 * k + 2 * jj + 1]));
        tmem_ld_wait();
        ld_x8(t_lane + ((blk + 1) & 1) * 128, sa);
        ld_x8(t_lane + ((blk + 1) & 1) * 128 + 256, da);
#pragma unroll
        for (int k = 0; k < 8; k++)
#pragma unroll
            for (int jj = 0; jj < 2; jj++)
                fma_acc_f32x2(acc[k * 4 + 2 + jj], pack_u32x2(sb[4 * k + 2 * jj], sb[4 * k + 2 * jj + 1]), pack_u32x2(db[4 * k + 2 * jj], db[4 * k + 2 * jj + 1]));
    }
    tmem_ld_wait();
    long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 32; i++) { float a, b; unpack_f32x2(acc[i], a, b); s += a + b; }
    y[blockIdx.x * 256 + threadIdx.x] = s + (float)(sa[0] + da[0]);
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    tmem_teardown(base, warp);
}

template <int MODE>   // 0: d = a*s + s' (pair, scalar, scalar)   1: d = a*b + d (three pairs)   2: the epilogue's pair of FMAs
__global__ void __launch_bounds__(256, 1) ffma2_kernel(long long * cyc, float * out, int iters, float seed) {
    uint64_t a[16], b[16];
    for (int i = 0; i < 16; i++) { a[i] = pack_f32x2(seed + i + threadIdx.x, seed * 0.5f + i); b[i] = pack_f32x2(1.0f + 1e-7f * i, 1.0f - 1e-7f * i); }
    if (MODE == 4) for (int i = 0; i < 16; i++) b[i] = pack_u32x2((1u << 19) + 16u * (threadIdx.x + i), (1u << 19) - 16u * (threadIdx.x + 3 * i));   // denormals
    const float s0 = MODE >= 3 ? seed * 5e28f : 1.0000001f, s1 = 1e-9f;
    const uint64_t m2 = pack_f32x2(s0, s0), c2 = pack_f32x2(s1, s1);
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) {
            if (MODE == 0) a[i] = fma_f32x2(a[i], m2, c2);
            else if (MODE == 1) fma_acc_f32x2(a[i], b[i], b[(i + 5) & 15]);
            else if (MODE == 2) { const uint64_t t = fma_f32x2(b[i], m2, c2); fma_acc_f32x2(a[i], t, b[(i + 3) & 15]); }
            else fma_acc_f32x2(a[i], b[i], m2);      // modes 3 / 4: b normal / denormal

        }
    }
    long long t1 = clock64();
    float s = 0;
    for (int i = 0; i < 16; i++) { float lo, hi; unpack_f32x2(a[i], lo, hi); s += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

static double mean_cycles(long long * dcyc, int n) {
    static long long h[148];
    cudaMemcpy(h, dcyc, sizeof(long long) * n, cudaMemcpyDeviceToHost);
    double c = 0;
    for (int i = 0; i < n; i++) c += (double)h[i];
    return c / n;
}
#define CHECK() do { cudaError_t e = cudaDeviceSynchronize(); if (e != cudaSuccess) { printf("CUDA error: %s at line %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

int main() {
    const int G = 148;
    long long * cyc; unsigned * sink; float * y;
    cudaMalloc(&cyc, G * 8); cudaMalloc(&sink, G * 256 * 4); cudaMalloc(&y, (size_t)G * 256 * 4);
    printf("== (1) tcgen05.ld.16x256b.x8 (4 KB per warp instruction), %d CTAs\n", G);
    const int iters = 4000;
    for (int warps = 4; warps <= 8; warps += 4) {
        for (int batch = 1; batch <= 4; batch *= 2) {
            for (int rep = 0; rep < 2; rep++) {
                if (batch == 1) ldtm_kernel<1><<<G, 256>>>(cyc, sink, iters, warps);
                if (batch == 2) ldtm_kernel<2><<<G, 256>>>(cyc, sink, iters, warps);
                if (batch == 4) ldtm_kernel<4><<<G, 256>>>(cyc, sink, iters, warps);
                CHECK();
            }
            const double c = mean_cycles(cyc, G);
            printf("  %d warps, %d loads per wait: %7.1f clk per wait, %7.1f B/clk/SM  (a 128x128 int32 tile = 65536 B -> %5.0f clk)\n",
                   warps, batch, c / iters, (double)iters * batch * warps * 4096 / c, 65536.0 / ((double)iters * batch * warps * 4096 / c));
        }
    }
    printf("== (2) two-FMA formulation (kernel 3): S from TMEM, 8 warps, per Q4_0 block of a 128x128 tile  (FMA-pipe floor 256 clk, MMA 64 clk)\n");
    cudaFuncSetAttribute(twofma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 33792);
    const int gaps[5] = {-1, 1000, 500, 250, 0};
    for (int gi = 0; gi < 5; gi++) {
        const int nbk = 512;
        for (int rep = 0; rep < 2; rep++) { twofma_kernel<<<G, 288, 33792>>>(cyc, y, nbk, gaps[gi]); CHECK(); }
        const double c = mean_cycles(cyc, G);
        if (gaps[gi] < 0) printf("  nbk %4d, tensor core idle:                        %9.0f clk total, %6.1f clk per block\n", nbk, c, c / nbk);
        else printf("  nbk %4d, an MMA pair every max(128, %4d) clk:     %9.0f clk total, %6.1f clk per block\n", nbk, gaps[gi], c, c / nbk);
    }
    printf("== (2b) the two-FMA epilogue plus what kernel 3 has around it\n");
    run_plus<0>(cyc, y, "nothing added");
    run_plus<1>(cyc, y, "+ scales from shared memory (5 LDS.128 per thread and block)");
    run_plus<2>(cyc, y, "+ fences, __syncwarp, mbarrier.arrive per warp and block");
    run_plus<4>(cyc, y, "+ two mbarrier.test_wait per block");
    run_plus<7>(cyc, y, "+ all three");
    run_plus<8>(cyc, y, "+ two unpack-like warps (LDS.128 / logic / STS.128) flat out");
    run_plus<15>(cyc, y, "+ everything");
    printf("== (3) one-FMA formulation: S and D = dw x dx both from TMEM, 8 warps  (FMA-pipe floor 128 clk, TMEM-read floor 512 clk)\n");
    for (int nbk = 128; nbk <= 512; nbk *= 4) {
        for (int rep = 0; rep < 2; rep++) { onefma_kernel<<<G, 256>>>(cyc, y, nbk); CHECK(); }
        const double c = mean_cycles(cyc, G);
        printf("  nbk %4d: %9.0f clk total, %6.1f clk per block\n", nbk, c, c / nbk);
    }
    printf("== (4) FFMA2 operand shapes, 8 warps per SM, 16 chains per thread\n");
    const char * names[5] = {"d = a * s + s' (pair, scalar, scalar)", "d = a * b + d  (three pairs)", "t = b*s+s'; d = t*b'+d (t is loop invariant)",
                             "d = b * s + d, b normal", "d = b * s + d, b DENORMAL (the S_b + 2^19 bits)"};
    for (int mode = 0; mode < 5; mode++) {
        const int it2 = 20000;
        for (int rep = 0; rep < 2; rep++) {
            if (mode == 0) ffma2_kernel<0><<<G, 256>>>(cyc, y, it2, 1.5f);
            if (mode == 1) ffma2_kernel<1><<<G, 256>>>(cyc, y, it2, 1.5f);
            if (mode == 2) ffma2_kernel<2><<<G, 256>>>(cyc, y, it2, 1.5f);
            if (mode == 3) ffma2_kernel<3><<<G, 256>>>(cyc, y, it2, 1.5f);
            if (mode == 4) ffma2_kernel<4><<<G, 256>>>(cyc, y, it2, 1.5f);
            CHECK();
        }
        const double c = mean_cycles(cyc, G);
        const double n2 = (double)it2 * 16 * (mode == 2 ? 2 : 1) * 8;       // FFMA2 warp instructions per SM
        printf("  %-46s %6.2f clk per FFMA2 per SMSP, %6.1f fp32 FMA/clk/SM\n", names[mode], c / (n2 / 4), n2 * 64 / c);
    }
    return 0;
}
