// Microbenchmark (GPU box): the issue/execution rate of tcgen05.mma.kind::i8 M=128 K=32 in the patterns kernel 3 uses --
// one thread per SM issues, operands are arbitrary shared-memory tiles (K-major, SWIZZLE_128B), nobody reads TMEM.
//   mode 0: every MMA accumulates into ONE 128x128 buffer, one commit at the end              (the plain-GEMM pattern)
//   mode 1: per Q4_0 block a data MMA (overwrite) + the constant-tile MMA (accumulate), buffers rotate over 4, no commits
//   mode 2: as 1, plus a tcgen05.commit to an mbarrier after every block                         (kernel 3's pattern)
//   mode 3: one MMA per block (overwrite), rotating buffers, commit per block
//   mode 4: as 2 with N = 256 (two buffers): half as many, twice as wide instructions
//   (the mode is a template parameter: the issue loop is latency bound on its own scalar instructions, a runtime switch
//    inside it changed every number)
//   mode 5: as 2, the issuer WAITS for each block's commit before issuing the next (latency of MMA pair + commit)
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I ../../llama.cpp-q_4_0_b200/csrc -o mma_rate mma_rate.cu
// Not part of the product.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#include "ptx.cuh"

using namespace b200;

// A operand from TENSOR MEMORY (128 lanes x 8 columns of 4 int8), B from shared memory
__device__ __forceinline__ void mma_i8_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// the SS form without the disable_output_lane vector
__device__ __forceinline__ void mma_i8_ss_plain(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

// HOW = 0: one thread runs the whole issue loop (if (threadIdx.x == 0) { for ... }) -- what kernel 3 did;
// HOW = 1: the whole warp runs the loop converged, only the tcgen05 instructions sit under elect.sync (operands stay warp-uniform)
template <int HOW, int mode>
__global__ void __launch_bounds__(128, 1) mma_kernel(long long * cyc, int nblk) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint32_t slot;
    __shared__ __align__(8) unsigned long long bar[8];
    const int warp = threadIdx.x >> 5;
    const uint32_t sbase = smem_u32(smem);
    for (int i = threadIdx.x; i < 3 * 32768 / 16; i += 128) reinterpret_cast<uint4 *>(smem)[i] = make_uint4(0x01020304u * (i & 7), 0x80808080u, i, ~i);
    fence_proxy_async_smem();
    if (threadIdx.x == 0) { for (int i = 0; i < 8; i++) mbar_init(smem_u32(&bar[i]), 1); mbar_fence_init(); }
    if (warp == 0) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tm = *reinterpret_cast<volatile uint32_t *>(&slot);
    if (HOW == 0 ? threadIdx.x == 0 : warp == 0) {
#define ISSUE if (HOW == 0 || elect_one())
        const int N = mode == 4 ? 256 : mode == 7 ? 64 : mode == 8 ? 16 : 128;
        const uint32_t idesc = umma_idesc_i8(128, N);
        const uint32_t a = sbase, b = sbase + 32768, c = sbase + 65536;
        const uint64_t cdesc = umma_desc_k_sw128(c);
        const int nbuf = 512 / N;
        uint32_t ph[4] = {0, 0, 0, 0};
        const long long t0 = clock64();
        for (int blk = 0; blk < nblk; blk++) {
            const int j = blk % nbuf;
            const uint32_t d = tm + j * N;
            const uint64_t ad = umma_desc_k_sw128(a + (blk & 3) * 32), bd = umma_desc_k_sw128(b + (blk & 3) * 32);
            if (mode == 9) {
                // A from TMEM: 3 D buffers (columns 0..383), A tiles of 8 columns from column 384 on (16 data tiles + the constant tile)
                const uint32_t d3 = tm + (blk % 3) * 128;
                ISSUE { mma_i8_ts(d3, tm + 384 + (blk & 15) * 8, bd, idesc, 0u);
                mma_i8_ts(d3, tm + 384 + 128 - 8, cdesc, idesc, 1u);
                mma_commit(smem_u32(&bar[blk & 3])); }
            } else if (mode == 10) {
                ISSUE { mma_i8_ss_plain(d, ad, bd, idesc, 0u);
                mma_i8_ss_plain(d, cdesc, cdesc, idesc, 1u);
                mma_commit(smem_u32(&bar[j])); }
            } else if (mode == 11) {
                const uint32_t d3 = tm + (blk % 3) * 128;
                ISSUE { mma_i8_ts(d3, tm + 384 + (blk & 15) * 8, bd, idesc, 0u);
                mma_commit(smem_u32(&bar[blk & 3])); }
            } else if (mode == 6) {
                // two blocks per step: data MMA each (N = 128, overwrite), ONE constant-tile MMA of N = 256 over both buffers
                const uint32_t d2 = tm + (blk & 1) * 256;
                ISSUE { mma_i8_ss(d2, ad, bd, idesc, 0u);
                mma_i8_ss(d2 + 128, ad, bd, idesc, 0u);
                mma_i8_ss(d2, cdesc, cdesc, umma_idesc_i8(128, 256), 1u);
                mma_commit(smem_u32(&bar[blk & 1])); }
            } else if (mode == 0) {
                ISSUE mma_i8_ss(tm, ad, bd, idesc, blk > 0);
            } else if (mode == 3) {
                ISSUE { mma_i8_ss(d, ad, bd, idesc, 0u);
                mma_commit(smem_u32(&bar[j])); }
            } else {
                ISSUE { mma_i8_ss(d, ad, bd, idesc, 0u);
                mma_i8_ss(d, cdesc, cdesc, idesc, 1u);
                if (mode != 1) mma_commit(smem_u32(&bar[j])); }
                if (mode == 5) { mbar_wait(smem_u32(&bar[j]), ph[j]); ph[j] ^= 1; tc_fence_after_sync(); }
            }
        }
        const long long t1 = clock64();
        ISSUE mma_commit(smem_u32(&bar[7]));
        mbar_wait(smem_u32(&bar[7]), 0);
        const long long t2 = clock64();
        if (threadIdx.x == 0) { cyc[2 * blockIdx.x] = t1 - t0; cyc[2 * blockIdx.x + 1] = t2 - t0; }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) { tc_fence_after_sync(); tmem_dealloc(tm, 512); }
}

int main_proto(long long * cyc);
void main_prim(long long * cyc);
void main_after(long long * cyc);
int main() {
    const int G = 148;
    long long * cyc;
    cudaMalloc(&cyc, 32768);
    typedef void (*kern_t)(long long *, int);
#define K2(m) {mma_kernel<0, m>, mma_kernel<1, m>}
    kern_t kerns[12][2] = {K2(0), K2(1), K2(2), K2(3), K2(4), K2(5), K2(6), K2(7), K2(8), K2(9), K2(10), K2(11)};
    for (int m = 0; m < 12; m++) for (int h = 0; h < 2; h++) cudaFuncSetAttribute(kerns[m][h], cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 32768 + 1024);
    const char * names[12] = {"one buffer, accumulate, 1 MMA per step", "data + constant MMA, 4 buffers, no commit", "data + constant MMA, commit per block (kernel 3)",
                             "1 MMA per block, commit per block", "data + constant MMA, N=256, commit per block", "as kernel 3 but wait for each commit",
                             "2 blocks: 2 data MMAs + one N=256 constant MMA, 1 commit", "as mode 1 with N = 64", "as mode 1 with N = 16",
                             "A FROM TMEM: data + constant MMA, 3 buffers, commit per block", "as mode 2 without the disable_output_lane operand",
                             "A FROM TMEM: one MMA per block, commit per block"};
    for (int how = 0; how < 2; how++) {
    printf("== issue loop run by %s\n", how == 0 ? "ONE THREAD (if (threadIdx.x == 0) around the loop)" : "the WHOLE WARP, tcgen05 instructions under elect.sync");
    for (int mode = 0; mode < 12; mode++) {
        for (int nblk = 128; nblk <= 2048; nblk *= 16) {
            static long long h[2 * G];
            for (int rep = 0; rep < 2; rep++) {
                kerns[mode][how]<<<G, 128, 3 * 32768 + 1024>>>(cyc, nblk);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("CUDA error: %s (mode %d)\n", cudaGetErrorString(e), mode); return 1; }
            }
            cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
            double issue = 0, done = 0;
            for (int i = 0; i < G; i++) { issue += (double)h[2 * i]; done += (double)h[2 * i + 1]; }
            issue /= G; done /= G;
            const double cols = mode == 4 || mode == 6 ? 256 : 128;
            printf("mode %d %-52s steps %5d: issue %7.1f clk/step, complete %7.1f clk/step = %6.1f clk per 128x128x32 data MMA\n", mode, names[mode], nblk,
                   issue / nblk, done / nblk, done / nblk / (cols / 128));
        }
    }
    }
    main_proto(cyc);
    main_prim(cyc);
    main_after(cyc);
    return 0;
}

// ---- (B) the accumulator hand-off protocol of kernel 3 alone: warp 8 issues the MMA pair of block b into TMEM buffer b % NBUF
// after d_empty[b % NBUF] completed, commits to d_full[b % NBUF]; 8 "epilogue" warps wait for d_full and arrive on d_empty at
// once (no TMEM read, no math).  What does one block cost when nothing but the protocol runs?
//   V = 0: waits by mbarrier.try_wait loops (ptx.cuh mbar_wait)      V = 1: waits spin on mbarrier.test_wait
//   V = 2: as 0, epilogue warps probe the NEXT block's barrier with test_wait before waiting (kernel 3)
template <int V, int NBUF>
__global__ void __launch_bounds__(384, 1) proto_kernel(long long * cyc, int nblk) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint32_t slot;
    __shared__ __align__(8) unsigned long long bar[2 * NBUF];
    __shared__ unsigned int freed[NBUF];          // V = 6: epilogue warps count their releases here instead of d_empty
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t sbase = smem_u32(smem);
    if (threadIdx.x < NBUF) freed[threadIdx.x] = 0;
    for (int i = threadIdx.x; i < 3 * 32768 / 16; i += blockDim.x) reinterpret_cast<uint4 *>(smem)[i] = make_uint4(0x01020304u * (i & 7), 0x80808080u, i, ~i);
    fence_proxy_async_smem();
    const uint32_t d_full = smem_u32(&bar[0]), d_empty = smem_u32(&bar[NBUF]);
    if (threadIdx.x == 0) { for (int i = 0; i < NBUF; i++) { mbar_init(d_full + 8 * i, 1); mbar_init(d_empty + 8 * i, 8); } mbar_fence_init(); }
    if (warp == 0) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tm = *reinterpret_cast<volatile uint32_t *>(&slot);
    constexpr int N = 512 / NBUF;
    auto wait = [&](uint32_t b, uint32_t par) {
        if (V == 1) { while (!mbar_test_wait(b, par)) { } }
        else if (V == 16) { while (!mbar_test_wait(b, par)) __nanosleep(40); }
        else if (V == 17) { while (!mbar_test_wait(b, par)) __nanosleep(200); }
        else mbar_wait(b, par);
    };
    const long long t0 = clock64();
    if (warp > 8 && V != 13) {
        // only V = 13 uses warps 9..11
    } else if (warp >= 8) {
        const uint32_t idesc = umma_idesc_i8(128, N);
        const uint64_t ad = umma_desc_k_sw128(sbase), bd = umma_desc_k_sw128(sbase + 32768), cd = umma_desc_k_sw128(sbase + 65536);
        bool ok_next = true;
        // V = 13: FOUR issuing warps, warp 8 + j owns TMEM buffer j (blocks j, j + 4, ...): what a warp loses after its commit
        // (its next wait only returns once that commit has retired) overlaps the other three warps' issues
        for (int blk = (V == 13 ? warp - 8 : 0); blk < nblk; blk += (V == 13 ? 4 : 1)) {
            const int j = blk % NBUF, r = blk / NBUF;
            if (V == 3) {
                // the probe of block blk was issued one iteration ago: its latency ran under the previous block's issue
                if (!ok_next) mbar_wait(d_empty + 8 * j, (r - 1) & 1);
                const int b1 = blk + 1;
                ok_next = b1 / NBUF == 0 ? true : mbar_test_wait(d_empty + 8 * (b1 % NBUF), (b1 / NBUF - 1) & 1);
            } else if (V == 6 || V == 12) {
                if (r > 0) {
                    unsigned int v;
                    if (V == 6) do { asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"(smem_u32(&freed[j])) : "memory"); } while (v < 8u * r);
                    else        do { asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v) : "r"(smem_u32(&freed[j])) : "memory"); } while (v < 8u * r);
                }
            } else if (V == 14) {
                // no wait at all: does a WAITED-ON d_full alone slow the pipe down?
            } else {
            if (V == 0 && blockIdx.x == 0 && lane == 0 && blk >= 256 && blk < 272) cyc[512 + 4 * (blk - 256) + 3] = clock64() - t0;
            if (r > 0) wait(d_empty + 8 * j, (r - 1) & 1);
            }
            if (V != 7 && V != 9) tc_fence_after_sync();
            if (V == 0 && blockIdx.x == 0 && lane == 0 && blk >= 256 && blk < 272) cyc[512 + 4 * (blk - 256)] = clock64() - t0;
            if (V == 8 || V == 9) {
                if (lane == 0) {
                    mma_i8_ss(tm + j * N, ad + 2 * (blk & 3), bd + 2 * (blk & 3), idesc, 0u);
                    mma_i8_ss(tm + j * N, cd, cd, idesc, 1u);
                    mma_commit(d_full + 8 * j);
                }
            } else
            if (elect_one()) {
                mma_i8_ss(tm + j * N, ad + 2 * (blk & 3), bd + 2 * (blk & 3), idesc, 0u);
                mma_i8_ss(tm + j * N, cd, cd, idesc, 1u);
                mma_commit(d_full + 8 * j);
            }
            __syncwarp();
            if (V == 0 && blockIdx.x == 0 && lane == 0 && blk >= 256 && blk < 272) cyc[512 + 4 * (blk - 256) + 1] = clock64() - t0;
        }
    } else {
        bool ok = false;
        for (int blk = 0; blk < nblk; blk++) {
            const int j = blk % NBUF, r = blk / NBUF;
            if (V == 10) { if (lane == 0) wait(d_full + 8 * j, r & 1); __syncwarp(); }
            else if (V == 11) {
                // one warp polls, the other seven sleep on a hardware named barrier
                if (warp == 0) wait(d_full + 8 * j, r & 1);
                asm volatile("bar.sync 2, 256;" ::: "memory");
            }
            else if (!ok) wait(d_full + 8 * j, r & 1);
            if (V == 0 && blockIdx.x == 0 && threadIdx.x == 0 && blk >= 256 && blk < 272) cyc[512 + 4 * (blk - 256) + 2] = clock64() - t0;
            tc_fence_after_sync();
            if (V >= 2 && blk + 1 < nblk) ok = mbar_test_wait(d_full + 8 * ((blk + 1) % NBUF), ((blk + 1) / NBUF) & 1);
            tc_fence_before_sync();
            __syncwarp();
            if (V == 6 || V == 12) { if (lane == 0) asm volatile("red.release.cta.shared.add.u32 [%0], 1;" ::"r"(smem_u32(&freed[j])) : "memory"); }
            else if (lane == 0) mbar_arrive(d_empty + 8 * j);
        }
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) cyc[2 * blockIdx.x] = t1 - t0;
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) { tc_fence_after_sync(); tmem_dealloc(tm, 512); }
}

template <int V, int NBUF>
static void run_proto(long long * cyc, const char * what) {
    const int G = 148, nblk = 2048;
    static long long h[2 * G];
    cudaFuncSetAttribute(proto_kernel<V, NBUF>, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 32768 + 1024);
    for (int rep = 0; rep < 2; rep++) {
        proto_kernel<V, NBUF><<<G, V == 13 ? 384 : 288, 3 * 32768 + 1024>>>(cyc, nblk);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("CUDA error: %s (%s)\n", cudaGetErrorString(e), what); return; }
    }
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double c = 0;
    for (int i = 0; i < G; i++) c += (double)h[2 * i];
    printf("  %-70s %d buffers: %7.1f clk per block\n", what, NBUF, c / G / nblk);
    if (V == 0 && NBUF == 4) {
        static long long tr[64];
        cudaMemcpy(tr, cyc + 512, sizeof(tr), cudaMemcpyDeviceToHost);
        for (int b = 0; b < 16; b++) printf("      blk %d: MMA warp reaches the d_empty wait at %lld, passes it +%lld, has issued (3 tcgen05 instr.) +%lld; epilogue warp 0 sees d_full at %lld\n",
                                            256 + b, tr[4 * b + 3], tr[4 * b] - tr[4 * b + 3], tr[4 * b + 1] - tr[4 * b], tr[4 * b + 2]);
    }
}

int main_proto(long long * cyc) {
    printf("== (B) accumulator hand-off protocol alone (MMA pair -> commit -> 8 warps wait, arrive -> MMA warp waits)\n");
    run_proto<0, 4>(cyc, "try_wait loops");
    run_proto<1, 4>(cyc, "test_wait spins");
    run_proto<2, 4>(cyc, "try_wait loops + early test of the next block");
    run_proto<6, 4>(cyc, "d_empty as a shared-memory counter (red.release / ld.acquire), d_full an mbarrier");
    run_proto<6, 2>(cyc, "d_empty as a shared-memory counter (red.release / ld.acquire), d_full an mbarrier");
    run_proto<7, 4>(cyc, "try_wait loops, NO tcgen05.fence::after_thread_sync in the MMA warp");
    run_proto<8, 4>(cyc, "try_wait loops, lane 0 predicate instead of elect.sync");
    run_proto<9, 4>(cyc, "try_wait loops, lane 0 predicate, no fence");
    run_proto<16, 4>(cyc, "test_wait polls with __nanosleep(40) between them");
    run_proto<17, 4>(cyc, "test_wait polls with __nanosleep(200) between them");
    run_proto<13, 4>(cyc, "FOUR issuing warps, one per TMEM buffer (try_wait loops)");
    run_proto<12, 4>(cyc, "d_empty as a shared-memory counter polled with ld.volatile (no acquire)");
    run_proto<10, 4>(cyc, "only lane 0 of each epilogue warp polls d_full");
    run_proto<11, 4>(cyc, "only epilogue warp 0 polls d_full, bar.sync releases the other seven");
    run_proto<3, 4>(cyc, "both sides probe the next block early (test_wait), try_wait only on a miss");
    run_proto<3, 2>(cyc, "both sides probe the next block early (test_wait), try_wait only on a miss");
    run_proto<0, 2>(cyc, "try_wait loops");
    run_proto<0, 8>(cyc, "try_wait loops (N = 64 per buffer)");
    run_proto<1, 8>(cyc, "test_wait spins (N = 64 per buffer)");
    return 0;
}

// ---- (C) what the protocol's own instructions cost one warp, back to back (dependent: each result feeds a branch)
__global__ void __launch_bounds__(32, 1) prim_kernel(long long * out, int iters) {
    __shared__ __align__(8) unsigned long long bar[2];
    const uint32_t b0 = smem_u32(&bar[0]), b1 = smem_u32(&bar[1]);
    const int lane = threadIdx.x;
    if (lane == 0) { mbar_init(b0, 1); mbar_init(b1, 1); mbar_fence_init(); }
    __syncwarp();
    if (lane == 0) mbar_arrive(b0);       // phase 0 of b0 complete
    __syncwarp();
    int acc = 0;
    long long t[8];
    t[0] = clock64();
    for (int i = 0; i < iters; i++) acc += mbar_try_wait(b0, 0) ? 1 : 2;
    t[1] = clock64();
    for (int i = 0; i < iters; i++) acc += mbar_test_wait(b0, 0) ? 1 : 2;
    t[2] = clock64();
    for (int i = 0; i < iters; i++) { tc_fence_after_sync(); acc += i; tc_fence_before_sync(); }
    t[3] = clock64();
    for (int i = 0; i < iters; i++) { __syncwarp(); if (lane == 0) mbar_arrive(b1); }
    t[4] = clock64();
    for (int i = 0; i < iters; i++) { if (elect_one()) acc += i; __syncwarp(); }
    t[5] = clock64();
    // arrive then wait for that very phase (same warp): the arrive -> visible-to-try_wait latency
    uint32_t ph = (uint32_t)(iters & 1);
    for (int i = 0; i < iters; i++) { __syncwarp(); if (lane == 0) mbar_arrive(b1); mbar_wait(b1, ph); ph ^= 1; }
    t[6] = clock64();
    if (lane == 0) { for (int k = 0; k < 6; k++) out[k] = t[k + 1] - t[k]; out[7] = acc; }
}

void main_prim(long long * cyc) {
    const int iters = 2000;
    long long h[8];
    for (int rep = 0; rep < 2; rep++) { prim_kernel<<<1, 32>>>(cyc, iters); cudaDeviceSynchronize(); }
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    const char * names[6] = {"mbarrier.try_wait on a completed phase -> predicate", "mbarrier.test_wait on a completed phase -> predicate",
                             "tcgen05.fence::after_thread_sync + ::before_thread_sync", "__syncwarp + one lane mbarrier.arrive", "elect.sync + __syncwarp",
                             "__syncwarp + arrive + try_wait for that phase"};
    printf("== (C) one warp, back to back, clk per iteration\n");
    for (int k = 0; k < 6; k++) printf("  %-60s %7.1f\n", names[k], (double)h[k] / iters);
}


// ---- (D) one warp issues kernel 3's MMA pair + commit per step and, after the commit, does ONE more thing: what stalls?
// W = 6 / 7: eight more warps issue back-to-back conflict-free LDS.128 (W = 6: one per ~64 clk per warp, W = 7: flat out)
// while warp 0 issues: the MMA's operands come out of the same shared memory
template <int W>
__global__ void __launch_bounds__(384, 1) after_commit_kernel(long long * cyc, int nblk) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint32_t slot;
    __shared__ __align__(8) unsigned long long bar[8];
    __shared__ volatile unsigned int word;
    __shared__ volatile int done;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t sbase = smem_u32(smem);
    if (threadIdx.x == 0) done = 0;
    for (int i = threadIdx.x; i < 3 * 32768 / 16; i += blockDim.x) reinterpret_cast<uint4 *>(smem)[i] = make_uint4(0x01020304u * (i & 7), 0x80808080u, i, ~i);
    fence_proxy_async_smem();
    if (threadIdx.x == 0) { for (int i = 0; i < 8; i++) mbar_init(smem_u32(&bar[i]), 1); mbar_fence_init(); word = 0; }
    if (warp == 0) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tm = *reinterpret_cast<volatile uint32_t *>(&slot);
    if (threadIdx.x == 0) mbar_arrive(smem_u32(&bar[6]));      // bar[6]: phase 0 complete, never touched again
    __syncthreads();
    if (warp == 0) {
        const uint32_t idesc = umma_idesc_i8(128, 128);
        const uint64_t ad = umma_desc_k_sw128(sbase), bd = umma_desc_k_sw128(sbase + 32768), cd = umma_desc_k_sw128(sbase + 65536);
        unsigned int acc = 0;
        const long long t0 = clock64();
        for (int blk = 0; blk < nblk; blk++) {
            const int j = blk & 3;
            if (elect_one()) {
                mma_i8_ss(tm + j * 128, ad + 2 * j, bd + 2 * j, idesc, 0u);
                mma_i8_ss(tm + j * 128, cd, cd, idesc, 1u);
                mma_commit(smem_u32(&bar[j]));
            }
            __syncwarp();
            if (W == 1) acc += mbar_test_wait(smem_u32(&bar[6]), 0) ? 1 : 2;
            if (W == 2) acc += word;
            if (W == 3) tc_fence_after_sync();
            if (W == 4) { if (lane == 0) mbar_arrive(smem_u32(&bar[5])); }
            if (W == 5) acc += mbar_test_wait(smem_u32(&bar[(j + 2) & 3]), 0) ? 1 : 2;     // a barrier commits DO arrive on
        }
        const long long t1 = clock64();
        if (elect_one()) mma_commit(smem_u32(&bar[7]));
        __syncwarp();
        mbar_wait(smem_u32(&bar[7]), 0);
        const long long t2 = clock64();
        if (threadIdx.x == 0) { cyc[2 * blockIdx.x] = t1 - t0; cyc[2 * blockIdx.x + 1] = t2 - t0 + (acc == 12345u); done = 1; }
    } else if (warp >= 4 && (W == 6 || W == 7)) {
        uint4 a = make_uint4(0, 0, 0, 0);
        unsigned long long n = 0;
        while (!done) {
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const uint4 v = lds_u32x4(sbase + 65536 + 16384 + ((warp * 8 + k) & 31) * 512 + lane * 16);
                a.x ^= v.x; a.y += v.y; a.z ^= v.z; a.w += v.w;
                if (W == 6) { const long long t = clock64(); while (clock64() - t < 56) { } }
            }
            n += 8;
        }
        if (lane == 0) cyc[1024 + blockIdx.x * 8 + (warp - 4)] = (long long)n + (a.x + a.y + a.z + a.w == 77u);
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) { tc_fence_after_sync(); tmem_dealloc(tm, 512); }
}

template <int W>
static void run_after(long long * cyc, const char * what) {
    const int G = 148, nblk = 2048;
    static long long h[2 * G];
    cudaFuncSetAttribute(after_commit_kernel<W>, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 32768 + 1024);
    for (int rep = 0; rep < 2; rep++) {
        after_commit_kernel<W><<<G, (W == 6 || W == 7) ? 384 : 128, 3 * 32768 + 1024>>>(cyc, nblk);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("CUDA error: %s (%s)\n", cudaGetErrorString(e), what); return; }
    }
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double a = 0, b = 0;
    for (int i = 0; i < G; i++) { a += (double)h[2 * i]; b += (double)h[2 * i + 1]; }
    printf("  after the commit: %-58s issue %7.1f, complete %7.1f clk per block", what, a / G / nblk, b / G / nblk);
    if (W == 6 || W == 7) {
        static long long c8[8];
        cudaMemcpy(c8, cyc + 1024, sizeof(c8), cudaMemcpyDeviceToHost);
        double n = 0; for (int i = 0; i < 8; i++) n += (double)c8[i];
        printf("   (%.1f LDS.128 warp instructions per block = %.0f B/clk of shared-memory reads beside the MMA's 128)", n / nblk, n * 512 / (h[1]));
    }
    printf("\n");
}

void main_after(long long * cyc) {
    printf("== (D) one warp: MMA pair + commit per block, then one more instruction\n");
    run_after<0>(cyc, "nothing");
    run_after<1>(cyc, "mbarrier.test_wait on an unrelated, completed barrier");
    run_after<2>(cyc, "a volatile shared-memory load");
    run_after<3>(cyc, "tcgen05.fence::after_thread_sync");
    run_after<4>(cyc, "mbarrier.arrive on an unrelated barrier");
    run_after<5>(cyc, "mbarrier.test_wait on a barrier the commits arrive on");
    run_after<6>(cyc, "nothing; 8 other warps: one LDS.128 per ~64 clk each");
    run_after<7>(cyc, "nothing; 8 other warps: LDS.128 flat out");
}
