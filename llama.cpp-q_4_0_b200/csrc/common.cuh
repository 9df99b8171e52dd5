// common.cuh -- shared device helpers and the HBM data layouts of the Q4_0 x F32 mul_mat path.
//
// Data layout in HBM (see DESIGN.md "Data layout"):
//
//   weights, LAYOUT_PLANES (after the one-time repack of the reference's 18-byte block_q4_0
//   stream, ggml.c:870-875):
//       wq : uint4 [rows][nbk]   16 bytes of nibbles per 32-weight block, 16-byte aligned;
//                                byte j: low nibble = weight j, high nibble = weight j+16
//                                (ggml.c:942-951) -- the nibble order is NOT changed
//       wd : half  [rows][nbk]   the block scale d, bit-identical fp16
//     algorithmic bytes stay 18 per block (0.5625 B/weight).
//   weights, LAYOUT_RAW: the 18-byte blocks as uploaded, rows contiguous (18*nbk bytes, 4-byte
//     aligned because nbk is even, ggml.c:2344).
//
//   activations after kernel 1 (the device-side replacement of the op's q8_0 work buffer,
//   ggml.c:11353-11367):
//       xq  : int8 [n][K]        quants in natural element order (K-major rows: also the B
//                                operand of the tensor-core GEMM, loadable by TMA)
//       xds : int2 [n][nbk]      .x = bits of fp32(fp16(d))   .y = -8 * sum_j q8_j   (exact int)
//     The -8*sum term turns sum (q4-8)*q8 into a dp4a chain over the UNSIGNED nibbles that starts
//     from xds.y: the integer result S_b is identical (checked bit-exact in tests).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define QK 32
#define Q4_0_BLOCK_BYTES 18
#define Q8_0_BLOCK_BYTES 34

namespace b200 {

__device__ __forceinline__ uint4 ldg_stream_u4(const uint4 * p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ uint32_t ldg_stream_u32(const uint32_t * p) {
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}

__device__ __forceinline__ unsigned short ldg_stream_u16(const unsigned short * p) {
    unsigned short r;
    asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h"(r) : "l"(p));
    return r;
}

__device__ __forceinline__ float half_bits_to_float(unsigned short h) {
    return __half2float(__ushort_as_half(h));
}

// S_b for one block: 16 bytes of nibbles (4 words) against 32 int8 (lo = elements 0..15,
// hi = elements 16..31), starting from init = -8 * sum(q8).  ggml.c:2593-2600 restated as
// sum q4*q8 - 8*sum q8 on IDP.4A.
__device__ __forceinline__ int block_dot_q4_q8(const uint4 q, const uint4 ylo, const uint4 yhi, int init) {
    int s = init;
    s = __dp4a((int)(q.x & 0x0F0F0F0Fu), (int)ylo.x, s);
    s = __dp4a((int)(q.y & 0x0F0F0F0Fu), (int)ylo.y, s);
    s = __dp4a((int)(q.z & 0x0F0F0F0Fu), (int)ylo.z, s);
    s = __dp4a((int)(q.w & 0x0F0F0F0Fu), (int)ylo.w, s);
    s = __dp4a((int)((q.x >> 4) & 0x0F0F0F0Fu), (int)yhi.x, s);
    s = __dp4a((int)((q.y >> 4) & 0x0F0F0F0Fu), (int)yhi.y, s);
    s = __dp4a((int)((q.z >> 4) & 0x0F0F0F0Fu), (int)yhi.z, s);
    s = __dp4a((int)((q.w >> 4) & 0x0F0F0F0Fu), (int)yhi.w, s);
    return s;
}

// Same S_b with 8 instead of 12 unpack instructions and two independent IDP.4A chains: the high
// nibbles stay in place (byte = 16*q_hi, read as UNSIGNED by dp4a.u32.s32), so that chain yields
// 16 * sum(q_hi*q8) exactly and one arithmetic shift restores it.
__device__ __forceinline__ int dp4a_us(uint32_t a_unsigned, uint32_t b_signed, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a_unsigned), "r"(b_signed), "r"(c));
    return d;
}
__device__ __forceinline__ int block_dot_q4_q8_fast(const uint4 q, const uint4 ylo, const uint4 yhi, int init) {
    int lo = init, hi = 0;
    lo = dp4a_us(q.x & 0x0F0F0F0Fu, ylo.x, lo);
    hi = dp4a_us(q.x & 0xF0F0F0F0u, yhi.x, hi);
    lo = dp4a_us(q.y & 0x0F0F0F0Fu, ylo.y, lo);
    hi = dp4a_us(q.y & 0xF0F0F0F0u, yhi.y, hi);
    lo = dp4a_us(q.z & 0x0F0F0F0Fu, ylo.z, lo);
    hi = dp4a_us(q.z & 0xF0F0F0F0u, yhi.z, hi);
    lo = dp4a_us(q.w & 0x0F0F0F0Fu, ylo.w, lo);
    hi = dp4a_us(q.w & 0xF0F0F0F0u, yhi.w, hi);
    return lo + (hi >> 4);
}

__device__ __forceinline__ float warp_sum(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    return v;
}

}  // namespace b200

// ---- programmatic dependent launch (PDL) ------------------------------------------------------
// griddepcontrol.launch_dependents: lets the NEXT kernel in the stream start being scheduled as
// this grid's CTAs retire; griddepcontrol.wait: blocks until every prerequisite grid has fully
// completed and flushed.  Both are no-ops when the launch carries no programmatic edge.  Weights
// are never written by a previous kernel, so kernels prefetch them BEFORE the wait and the HBM
// stream keeps flowing across kernel boundaries.
namespace b200 {
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
}  // namespace b200

// ---- optional per-CTA timeline trace (debug / profiling only; off unless ggml_b200_trace_begin) ----
// record = 8 x u64: {kernel seq, blockIdx | smid << 32, t_start, t_after_wait, t_after_prologue, t_end, 0, 0}
namespace b200 {
static __device__ unsigned long long * g_trace = nullptr;   // per translation unit
static __device__ unsigned int g_trace_cap = 0;
static __device__ unsigned int g_trace_pos = 0;

__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
struct CtaTrace {
    unsigned long long * rec;
    __device__ __forceinline__ void begin(unsigned seq) {
        rec = nullptr;
        if (g_trace != nullptr) {
            const unsigned slot = atomicAdd(&g_trace_pos, 1u);
            if (slot < g_trace_cap) {
                unsigned smid;
                asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
                rec = g_trace + (size_t)slot * 8;
                rec[0] = seq;
                rec[1] = (unsigned long long)blockIdx.x | ((unsigned long long)smid << 32);
                rec[2] = gtime();
            }
        }
    }
    __device__ __forceinline__ void mark(int i) { if (rec) rec[2 + i] = gtime(); }
};
}  // namespace b200
