// siblings.cu -- SURVEY 8(f) row 4: the op's sibling weight formats, Q8_0, Q4_1, Q5_0 and Q5_1, through the same design as Q4_0.
//
// What the reference does for them (ggml_compute_forward_mul_mat_q_f32, ggml.c:11268-11411, is format blind; the table
// quantize_fns[], ggml.c:1619-1670, binds the per-format pieces):
//   Q8_0 weights (block_q8_0, 34 B: fp16 d + 32 int8, ggml.c:902-907) x Q8_0 activations:
//       y = sum_b (d_w * d_x) * S_b,  S_b = sum_j w_j * x_j                      ggml_vec_dot_q8_0_q8_0, ggml.c:3231-3330
//   Q4_1 weights (block_q4_1, 20 B: fp16 d, fp16 m + 16 nibble bytes, ggml.c:877-883) x Q8_1 activations
//   (block_q8_1: FLOAT d = amax / 127, s = d * sum_j x_j, ggml.c:909-915, quantize_row_q8_1 ggml.c:1318-1480):
//       y = sum_b (d_w * d_x) * S_b + m_w * s_x,  S_b = sum_j q_j * x_j  (q unsigned 0..15) ggml_vec_dot_q4_1_q8_1, ggml.c:2609-2734
//   Q5_0 (block_q5_0, 22 B: fp16 d, 32 fifth bits qh, 16 nibble bytes, ggml.c:885-891) x Q8_0: as Q8_0 with w_j = (nibble | bit << 4) - 16
//                                                                                 ggml_vec_dot_q5_0_q8_0, ggml.c:2735-2973
//   Q5_1 (block_q5_1, 24 B: + fp16 m, ggml.c:893-900) x Q8_1: as Q4_1 with q = nibble | bit << 4 (0..31)   ggml_vec_dot_q5_1_q8_1, ggml.c:2975-3229
// The activation quantizer is kernel 1's arithmetic (the AVX2 contract: two IEEE divides, round-half-even); Q8_1 keeps
// the scale in fp32 and adds s.
//
// Device layout ("planes", as for Q4_0): quants [rows][nbk] as 16-byte words (Q4_1: one uint4 of nibbles per block, Q8_0:
// two uint4 = 32 int8), scales [rows][nbk] (Q8_0: half d; Q4_1: half2 {d, m}), so that a warp's loads are full 512-byte /
// 1-KB lines.  One kernel: every CTA first quantizes the activation columns of the pass into shared memory (what the
// fused n = 1 Q4_0 kernel does), then its warps stream rows with dp4a, lane = block.  HBM-bound: 34 / 32 B (Q8_0) and
// 20 / 32 B (Q4_1) per weight.  n > SIB_NC columns are done in passes of SIB_NC (the weights are re-read per pass: this
// is the decode path; a tensor-core path for the siblings is not built).
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "quantize_q8_0.cuh"
#include "state.h"

struct ggml_b200_qweight {
    int     type = 0;              // GGML_B200_TYPE_Q8_0 / GGML_B200_TYPE_Q4_1 (the reference's enum values, ggml.h:230-242)
    int64_t M = 0, K = 0;
    uint4 * wq = nullptr;          // [M][nbk] (Q4_1) or [M][nbk][2] (Q8_0)
    void *  wd = nullptr;          // [M][nbk] half (Q8_0) or half2 {d, m} (Q4_1)
    size_t  dev_bytes = 0;
    void *  gemm = nullptr;        // Q8_0: TMA descriptor of the int8 plane for the tensor-core prefill kernel (gemm_q4_0_tc.cu)
};

namespace b200 {
namespace sib {

constexpr int SIB_NC = 4;          // activation columns per pass
constexpr int THREADS = 512;

__host__ __device__ constexpr int block_bytes(int type) {
    return type == GGML_B200_TYPE_Q8_0 ? 34 : type == GGML_B200_TYPE_Q4_1 ? 20 : type == GGML_B200_TYPE_Q5_0 ? 22 : 24;
}
// activations: q8_1 (fp32 d, s = d * sum) for the formats with a min term, q8_0 (fp16 d) otherwise (quantize_fns[].vec_dot_type)
__host__ __device__ constexpr bool act_q8_1(int type) { return type == GGML_B200_TYPE_Q4_1 || type == GGML_B200_TYPE_Q5_1; }
__host__ __device__ constexpr bool is_q5(int type) { return type == GGML_B200_TYPE_Q5_0 || type == GGML_B200_TYPE_Q5_1; }

// raw reference blocks (row stride nb1) -> planes.  One thread per block.
template <int TYPE>
__global__ void __launch_bounds__(256) repack_kernel(const unsigned char * __restrict__ raw, size_t nb1, long long M, int nbk,
                                                     uint4 * __restrict__ wq, void * __restrict__ wd) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= M * nbk) return;
    const long long r = t / nbk;
    const int b = (int)(t - r * nbk);
    const unsigned short * src = reinterpret_cast<const unsigned short *>(raw + r * nb1 + (size_t)b * block_bytes(TYPE));   // 2-byte aligned
    if (TYPE == GGML_B200_TYPE_Q8_0) {
        reinterpret_cast<unsigned short *>(wd)[t] = src[0];
        uint32_t w[8];
#pragma unroll
        for (int i = 0; i < 8; i++) w[i] = (uint32_t)src[1 + 2 * i] | ((uint32_t)src[2 + 2 * i] << 16);
        wq[2 * t]     = make_uint4(w[0], w[1], w[2], w[3]);
        wq[2 * t + 1] = make_uint4(w[4], w[5], w[6], w[7]);
    } else if (TYPE == GGML_B200_TYPE_Q4_1) {
        reinterpret_cast<uint32_t *>(wd)[t] = (uint32_t)src[0] | ((uint32_t)src[1] << 16);      // {d, m}
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; i++) w[i] = (uint32_t)src[2 + 2 * i] | ((uint32_t)src[3 + 2 * i] << 16);
        wq[t] = make_uint4(w[0], w[1], w[2], w[3]);
    } else {
        // Q5_0: {d, qh[4], qs[16]} (ggml.c:885-891); Q5_1: {d, m, qh[4], qs[16]} (ggml.c:893-900) -> scale plane {d | m << 16, qh}
        const int o = TYPE == GGML_B200_TYPE_Q5_1 ? 2 : 1;                                       // halves before qh
        const uint32_t dm = (uint32_t)src[0] | (TYPE == GGML_B200_TYPE_Q5_1 ? (uint32_t)src[1] << 16 : 0u);
        const uint32_t qh = (uint32_t)src[o] | ((uint32_t)src[o + 1] << 16);
        reinterpret_cast<uint2 *>(wd)[t] = make_uint2(dm, qh);
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; i++) w[i] = (uint32_t)src[o + 2 + 2 * i] | ((uint32_t)src[o + 3 + 2 * i] << 16);
        wq[t] = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// One activation block by one warp (lane = element): quants to xq, scale (and s) to xd / xs.  Kernel 1's arithmetic.
template <int TYPE>
__device__ __forceinline__ void quantize_block(float v, int lane, int8_t * xq_blk, float * xd, float * xs) {
    float amax = fabsf(v);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = (amax != 0.0f) ? __fdiv_rn(127.0f, amax) : 0.0f;
    const int q = q8_round(v, id);
    xq_blk[lane] = (int8_t)q;
    int sum = q;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) {
        if (!act_q8_1(TYPE)) *xd = __half2float(__float2half_rn(d));                  // block_q8_0.d is fp16 (ggml.c:1214-1215)
        else { *xd = d; *xs = __fmul_rn(d, (float)sum); }                             // block_q8_1: float d, s = d * sum (ggml.c:1419,1449)
    }
}

// int32 block dot of one weight block against one activation block (32 int8 in shared memory)
template <int TYPE>
__device__ __forceinline__ int block_dot(const uint4 & w0, const uint4 & w1, const uint4 & x0, const uint4 & x1) {
    int s = 0;
    if (TYPE == GGML_B200_TYPE_Q8_0) {
        s = __dp4a((int)w0.x, (int)x0.x, s); s = __dp4a((int)w0.y, (int)x0.y, s); s = __dp4a((int)w0.z, (int)x0.z, s); s = __dp4a((int)w0.w, (int)x0.w, s);
        s = __dp4a((int)w1.x, (int)x1.x, s); s = __dp4a((int)w1.y, (int)x1.y, s); s = __dp4a((int)w1.z, (int)x1.z, s); s = __dp4a((int)w1.w, (int)x1.w, s);
    } else if (is_q5(TYPE)) {
        // element j = nibble | bit j of qh << 4 (low nibbles: elements 0..15, high: 16..31; ggml.c:1024-1038).  Four bits of qh
        // go to bit 0 of four bytes by one multiply (bit m of x lands on bit 8m of x * 0x00204081).  Q5_1 uses the unsigned
        // 0..31 (ggml.c:3156-3159); Q5_0 the value - 16 = nibble | (bit ? 0 : 0xF0) as a signed byte (ggml.c:2903-2906).
        const uint32_t qh = w1.x;
        const uint32_t w[4] = {w0.x, w0.y, w0.z, w0.w};
        const uint32_t xl[4] = {x0.x, x0.y, x0.z, x0.w}, xh[4] = {x1.x, x1.y, x1.z, x1.w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t bl = (qh >> (4 * i)) & 0xFu, bh = (qh >> (16 + 4 * i)) & 0xFu;
            uint32_t lo = w[i] & 0x0F0F0F0Fu, hi = (w[i] >> 4) & 0x0F0F0F0Fu;
            if (TYPE == GGML_B200_TYPE_Q5_1) {
                lo |= ((bl * 0x00204081u) & 0x01010101u) << 4;
                hi |= ((bh * 0x00204081u) & 0x01010101u) << 4;
            } else {
                lo |= (((bl ^ 0xFu) * 0x00204081u) & 0x01010101u) * 0xF0u;
                hi |= (((bh ^ 0xFu) * 0x00204081u) & 0x01010101u) * 0xF0u;
            }
            s = __dp4a((int)lo, (int)xl[i], s);
            s = __dp4a((int)hi, (int)xh[i], s);
        }
    } else {
        // byte j of the 16 = elements j (low nibble) and j + 16 (high nibble), unsigned 0..15 (ggml.c:2722-2726)
        s = __dp4a((int)(w0.x & 0x0F0F0F0Fu), (int)x0.x, s); s = __dp4a((int)((w0.x >> 4) & 0x0F0F0F0Fu), (int)x1.x, s);
        s = __dp4a((int)(w0.y & 0x0F0F0F0Fu), (int)x0.y, s); s = __dp4a((int)((w0.y >> 4) & 0x0F0F0F0Fu), (int)x1.y, s);
        s = __dp4a((int)(w0.z & 0x0F0F0F0Fu), (int)x0.z, s); s = __dp4a((int)((w0.z >> 4) & 0x0F0F0F0Fu), (int)x1.z, s);
        s = __dp4a((int)(w0.w & 0x0F0F0F0Fu), (int)x0.w, s); s = __dp4a((int)((w0.w >> 4) & 0x0F0F0F0Fu), (int)x1.w, s);
    }
    return s;
}

// Eight consecutive activations per thread, four threads per block (kernel 1's thread shape): every load of a column is
// in flight at once.  Quants to xq, scale (and s) to xd / xs.
template <int TYPE>
__device__ __forceinline__ void quantize_column(const float * __restrict__ xc, int K, int8_t * xq, float * xd, float * xs) {
    for (int j0 = 0; j0 < (K >> 3); j0 += THREADS) {
        const int j = j0 + threadIdx.x;
        const bool active = j < (K >> 3);
        const int jj = active ? j : (K >> 3) - 1;                 // keep the quad's shuffles convergent
        const float4 a = __ldcg(reinterpret_cast<const float4 *>(xc + 8 * jj));
        const float4 c = __ldcg(reinterpret_cast<const float4 *>(xc + 8 * jj) + 1);
        const float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
        float amax = 0.0f;
#pragma unroll
        for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v[i]));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
        const float d  = __fdiv_rn(amax, 127.0f);
        const float id = (amax != 0.0f) ? __fdiv_rn(127.0f, amax) : 0.0f;
        int q[8], sum = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) { q[i] = q8_round(v[i], id); sum += q[i]; }
        sum += __shfl_xor_sync(0xffffffffu, sum, 1);
        sum += __shfl_xor_sync(0xffffffffu, sum, 2);
        if (active) {
            const uint32_t w0 = (uint32_t)(q[0] & 0xFF) | ((uint32_t)(q[1] & 0xFF) << 8) | ((uint32_t)(q[2] & 0xFF) << 16) | ((uint32_t)(q[3] & 0xFF) << 24);
            const uint32_t w1 = (uint32_t)(q[4] & 0xFF) | ((uint32_t)(q[5] & 0xFF) << 8) | ((uint32_t)(q[6] & 0xFF) << 16) | ((uint32_t)(q[7] & 0xFF) << 24);
            *reinterpret_cast<uint2 *>(xq + 8 * j) = make_uint2(w0, w1);
            if ((j & 3) == 0) {
                if (!act_q8_1(TYPE)) xd[j >> 2] = __half2float(__float2half_rn(d));                  // block_q8_0.d is fp16 (ggml.c:1214-1215)
                else { xd[j >> 2] = d; xs[j >> 2] = __fmul_rn(d, (float)sum); }                      // block_q8_1: float d, s = d * sum (ggml.c:1419,1449)
            }
        }
    }
}

template <int TYPE>
__device__ __forceinline__ void load_blocks(const uint4 * __restrict__ wq, const void * __restrict__ wd, long long rb, int b0, int lane, int nbk,
                                            uint4 (&w0)[4], uint4 (&w1)[4], float (&dw)[4], float (&mw)[4]) {
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int b = b0 + 32 * u + lane;
        if (b < nbk) {
            if (TYPE == GGML_B200_TYPE_Q8_0) {
                w0[u] = ldg_stream_u4(wq + 2 * (rb + b));
                w1[u] = ldg_stream_u4(wq + 2 * (rb + b) + 1);
                dw[u] = __half2float(__ushort_as_half(ldg_stream_u16(reinterpret_cast<const unsigned short *>(wd) + rb + b)));
                mw[u] = 0.0f;
            } else if (is_q5(TYPE)) {
                w0[u] = ldg_stream_u4(wq + rb + b);
                const uint32_t dm = ldg_stream_u32(reinterpret_cast<const uint32_t *>(wd) + 2 * (rb + b));
                w1[u] = make_uint4(ldg_stream_u32(reinterpret_cast<const uint32_t *>(wd) + 2 * (rb + b) + 1), 0u, 0u, 0u);    // qh
                dw[u] = __half2float(__ushort_as_half((unsigned short)(dm & 0xFFFF)));
                mw[u] = __half2float(__ushort_as_half((unsigned short)(dm >> 16)));
            } else {
                w0[u] = ldg_stream_u4(wq + rb + b);
                w1[u] = w0[u];
                const uint32_t dm = ldg_stream_u32(reinterpret_cast<const uint32_t *>(wd) + rb + b);
                dw[u] = __half2float(__ushort_as_half((unsigned short)(dm & 0xFFFF)));
                mw[u] = __half2float(__ushort_as_half((unsigned short)(dm >> 16)));
            }
        }
    }
}

// y[c][row] for NC activation columns (c0 .. c0 + nc - 1) and every row; S_out != nullptr: also S_b (parity checkpoint).
// The weights of a warp's first row are requested BEFORE the wait for the previous kernel and the quantizer prologue (they
// depend on neither): under programmatic dependent launch they stream while the predecessor drains.
template <int TYPE, int NC, bool DUMP>
__global__ void __launch_bounds__(THREADS, 2)
sib_gemv_kernel(const uint4 * __restrict__ wq, const void * __restrict__ wd, const float * __restrict__ x, size_t nb11,
                float * __restrict__ y, long long ldy, long long M, int K, int nc, int * __restrict__ S_out, long long n_total, long long c0) {
    extern __shared__ __align__(16) unsigned char sib_smem[];
    const int nbk = K >> 5;
    int8_t * xq = reinterpret_cast<int8_t *>(sib_smem);                          // [NC][K]
    float * xd = reinterpret_cast<float *>(sib_smem + (size_t)NC * K);           // [NC][nbk]
    float * xs = xd + NC * nbk;                                                  // [NC][nbk]  (Q4_1)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = THREADS / 32;
    const long long gwarps = (long long)gridDim.x * nwarps;
    const long long row_first = (long long)blockIdx.x * nwarps + warp;

    uint4 w0[4], w1[4];
    float dw[4], mw[4];
    if (row_first < M) load_blocks<TYPE>(wq, wd, row_first * nbk, 0, lane, nbk, w0, w1, dw, mw);

    pdl_wait();
    pdl_launch_dependents();
    // ---- the op's INIT phase, per CTA: quantize the pass's activation columns (ggml.c:11353-11367) ----
    for (int c = 0; c < NC; c++)
        if (c < nc)
            quantize_column<TYPE>(reinterpret_cast<const float *>(reinterpret_cast<const char *>(x) + (size_t)c * nb11), K, xq + (size_t)c * K, xd + c * nbk, xs + c * nbk);
    __syncthreads();

    // ---- COMPUTE: a warp per row, lane = block (4 blocks per lane in flight) ----
    for (long long row = row_first; row < M; row += gwarps) {
        float acc[NC], mins[NC];
#pragma unroll
        for (int c = 0; c < NC; c++) { acc[c] = 0.0f; mins[c] = 0.0f; }
        const long long rb = row * nbk;
        for (int b0 = 0; b0 < nbk; b0 += 128) {
            if (row != row_first || b0 != 0) load_blocks<TYPE>(wq, wd, rb, b0, lane, nbk, w0, w1, dw, mw);
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int b = b0 + 32 * u + lane;
                if (b < nbk) {
#pragma unroll
                    for (int c = 0; c < NC; c++) {
                        if (c < nc) {
                            const uint4 * xb = reinterpret_cast<const uint4 *>(xq + (size_t)c * K + 32 * b);
                            const int S = block_dot<TYPE>(w0[u], w1[u], xb[0], xb[1]);
                            if (DUMP) S_out[((c0 + c) * M + row) * nbk + b] = S;
                            // d = d_w * d_x rounded, then acc = fma(d, float(S), acc): ggml.c:3305-3313 / :2692-2706
                            acc[c] = fmaf(__fmul_rn(dw[u], xd[c * nbk + b]), (float)S, acc[c]);
                            if (act_q8_1(TYPE)) mins[c] = __fadd_rn(mins[c], __fmul_rn(mw[u], xs[c * nbk + b]));   // summs += m * s, ggml.c:2688 / :3153
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int c = 0; c < NC; c++) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], o);
                mins[c] += __shfl_xor_sync(0xffffffffu, mins[c], o);
            }
            if (lane == 0 && c < nc) y[(size_t)c * ldy + row] = act_q8_1(TYPE) ? __fadd_rn(acc[c], mins[c]) : acc[c];   // hsum(acc) + summs
        }
    }
    (void)n_total;
}

// reference-format activation blocks (block_q8_0 34 B / block_q8_1 40 B) for the parity tests: one warp per block
template <int TYPE>
__global__ void __launch_bounds__(256) sib_quantize_blocks_kernel(const float * __restrict__ x, size_t nb11, long long n, int K, unsigned char * __restrict__ out) {
    const int nbk = K >> 5;
    const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (gw >= n * nbk) return;
    const long long c = gw / nbk;
    const int b = (int)(gw - c * nbk);
    const float v = reinterpret_cast<const float *>(reinterpret_cast<const char *>(x) + (size_t)c * nb11)[32 * b + lane];
    __shared__ int8_t q[8][32];
    __shared__ float d[8], s[8];
    const int w = threadIdx.x >> 5;
    quantize_block<TYPE>(v, lane, q[w], &d[w], &s[w]);
    __syncwarp();
    if (!act_q8_1(TYPE)) {
        unsigned char * blk = out + (size_t)gw * 34;
        if (lane == 0) *reinterpret_cast<unsigned short *>(blk) = __half_as_ushort(__float2half_rn(d[w]));
        blk[2 + lane] = (unsigned char)q[w][lane];
    } else {
        unsigned char * blk = out + (size_t)gw * 40;
        if (lane == 0) { reinterpret_cast<float *>(blk)[0] = d[w]; reinterpret_cast<float *>(blk)[1] = s[w]; }
        blk[8 + lane] = (unsigned char)q[w][lane];
    }
}

}  // namespace sib
}  // namespace b200

// ================================= C ABI =================================
using namespace b200;

static bool sib_type_ok(int type) { return type == GGML_B200_TYPE_Q8_0 || type == GGML_B200_TYPE_Q4_1 || type == GGML_B200_TYPE_Q5_0 || type == GGML_B200_TYPE_Q5_1; }

extern "C" ggml_b200_qweight * ggml_b200_upload_quant(int type, const void * blocks, int64_t M, int64_t K, size_t nb1) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_upload_quant: library not initialised"); return nullptr; }
    if (!sib_type_ok(type)) { set_error("ggml_b200_upload_quant: type %d is not Q8_0 (8), Q4_1 (3), Q5_0 (6) or Q5_1 (7)", type); return nullptr; }
    if (!blocks || M <= 0 || K <= 0 || K % 32 != 0) { set_error("ggml_b200_upload_quant: bad shape M=%lld K=%lld", (long long)M, (long long)K); return nullptr; }
    const int64_t nbk = K / 32;
    const size_t row_bytes = (size_t)nbk * sib::block_bytes(type);
    if (nb1 == 0) nb1 = row_bytes;
    if (nb1 < row_bytes || nb1 % 2 != 0) { set_error("ggml_b200_upload_quant: bad row stride %zu", nb1); return nullptr; }
    ggml_b200_qweight * w = new ggml_b200_qweight();
    w->type = type; w->M = M; w->K = K;
    const size_t q_bytes = (size_t)M * nbk * (type == GGML_B200_TYPE_Q8_0 ? 32 : 16);
    const size_t d_bytes = (size_t)M * nbk * (type == GGML_B200_TYPE_Q8_0 ? 2 : type == GGML_B200_TYPE_Q4_1 ? 4 : 8);
    unsigned char * raw = nullptr;
    cudaError_t e = cudaMalloc(&w->wq, q_bytes + 16);
    if (e == cudaSuccess) e = cudaMalloc(&w->wd, d_bytes + 16);
    if (e == cudaSuccess) e = cudaMalloc(&raw, (size_t)M * nb1);
    if (e == cudaSuccess) e = cudaMemcpyAsync(raw, blocks, (size_t)M * nb1, cudaMemcpyHostToDevice, g.stream);
    if (e == cudaSuccess) {
        const long long total = (long long)M * nbk;
        const unsigned grid = (unsigned)((total + 255) / 256);
        if (type == GGML_B200_TYPE_Q8_0)      sib::repack_kernel<GGML_B200_TYPE_Q8_0><<<grid, 256, 0, g.stream>>>(raw, nb1, M, (int)nbk, w->wq, w->wd);
        else if (type == GGML_B200_TYPE_Q4_1) sib::repack_kernel<GGML_B200_TYPE_Q4_1><<<grid, 256, 0, g.stream>>>(raw, nb1, M, (int)nbk, w->wq, w->wd);
        else if (type == GGML_B200_TYPE_Q5_0) sib::repack_kernel<GGML_B200_TYPE_Q5_0><<<grid, 256, 0, g.stream>>>(raw, nb1, M, (int)nbk, w->wq, w->wd);
        else                                  sib::repack_kernel<GGML_B200_TYPE_Q5_1><<<grid, 256, 0, g.stream>>>(raw, nb1, M, (int)nbk, w->wq, w->wd);
        e = cudaGetLastError();
        g.launches++;
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(g.stream);
    if (raw) cudaFree(raw);
    if (e != cudaSuccess) {
        cuda_fail(e, "ggml_b200_upload_quant", __FILE__, __LINE__);
        if (w->wq) cudaFree(w->wq);
        if (w->wd) cudaFree(w->wd);
        delete w;
        return nullptr;
    }
    w->dev_bytes = q_bytes + d_bytes;
    if (type == GGML_B200_TYPE_Q8_0) w->gemm = gemm_prepare_q8(w->wq, M, K);      // (nullptr: no tensor-core path, column passes instead)
    return w;
}

extern "C" void ggml_b200_free_quant(ggml_b200_qweight * w) {
    B200_API_LOCK;
    if (!w) return;
    if (g.inited) cudaStreamSynchronize(g.stream);
    if (w->gemm) gemm_release_q8(w->gemm);
    if (w->wq) cudaFree(w->wq);
    if (w->wd) cudaFree(w->wd);
    delete w;
}

extern "C" int ggml_b200_quant_type(const ggml_b200_qweight * w) { return w ? w->type : 0; }
extern "C" size_t ggml_b200_quant_bytes(const ggml_b200_qweight * w) { return w ? w->dev_bytes : 0; }

template <int TYPE, bool DUMP, int NC>
static int sib_launch(const ggml_b200_qweight * w, const float * x, size_t nb11, float * y, long long ldy, int nc, int * S, long long n_total, long long c0) {
    const int K = (int)w->K, nbk = K / 32;
    const size_t smem = (size_t)NC * K + 2 * sizeof(float) * NC * nbk;
    if (smem > (size_t)MAX_DYN_SMEM) { set_error("sibling formats: K = %d needs %zu bytes of shared memory", K, smem); return GGML_B200_ERR_ARG; }
    auto kern = sib::sib_gemv_kernel<TYPE, NC, DUMP>;
    B200_SET_SMEM_ATTR(kern, MAX_DYN_SMEM);
    // two CTAs per SM (measured: 3.78 TB/s for a 4096x4096 Q8_0 op against 3.32 with one; profiles/r02_bench_siblings.log)
    const int ctas_per_sm = smem * 2 + 2048 <= (size_t)MAX_DYN_SMEM ? 2 : 1;
    const long long warps_needed = (w->M + (sib::THREADS / 32) - 1) / (sib::THREADS / 32);
    long long grid = (long long)g.sm_count * ctas_per_sm;
    if (grid > warps_needed) grid = warps_needed;
    if (grid < 1) grid = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(sib::THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = g.stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = g.tune.pdl ? 1 : 0;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    CK(cudaLaunchKernelEx(&cfg, kern, (const uint4 *)w->wq, (const void *)w->wd, x, nb11, y, ldy, (long long)w->M, K, nc, S, n_total, c0));
    g.launches++;
    return GGML_B200_OK;
}

static int sib_mul_mat(const ggml_b200_qweight * w, const float * x_dev, int64_t n, size_t nb11, float * y_dev, size_t nb1, int32_t * S_dev) {
    if (!g.inited) { set_error("sibling formats: library not initialised"); return GGML_B200_ERR_STATE; }
    if (!w || !sib_type_ok(w->type)) { set_error("sibling formats: bad weight handle"); return GGML_B200_ERR_ARG; }
    if (n == 0) return GGML_B200_OK;
    if (!x_dev || !y_dev || n < 0) { set_error("sibling formats: bad arguments"); return GGML_B200_ERR_ARG; }
    if (nb11 == 0) nb11 = (size_t)w->K * 4;
    if (nb1 == 0) nb1 = (size_t)w->M * 4;
    if (nb11 % 16 != 0 || ((uintptr_t)x_dev & 15) != 0) { set_error("sibling formats: activation rows must be 16-byte aligned (every ggml tensor is: GGML_MEM_ALIGN)"); return GGML_B200_ERR_ARG; }
    if (nb11 % 4 != 0 || nb1 % 4 != 0 || nb11 < (size_t)w->K * 4 || nb1 < (size_t)w->M * 4) { set_error("sibling formats: bad strides nb11=%zu nb1=%zu", nb11, nb1); return GGML_B200_ERR_ARG; }
    const long long ldy = (long long)(nb1 / 4);
    // prefill on the tensor cores: Q8_0's int8 plane is kernel 3's A operand as it lies (gemm_q4_0_tc3.cuh, WFMT = 8);
    // kernel 1 of the Q4_0 path quantizes the activations (the same q8_0 blocks) into kernel 3's layout
    if (w->type == GGML_B200_TYPE_Q8_0 && w->gemm != nullptr && n <= INT32_MAX && g.path != GGML_B200_PATH_GEMV &&
        (g.path == GGML_B200_PATH_GEMM || n >= g.gemm_min_n)) {
        int rc = launch_quantize(x_dev, n, w->K, nb11, nullptr);
        if (rc) return rc;
        return launch_gemm_q8(w->gemm, w->wd, w->M, w->K, y_dev, ldy, (int)n, S_dev);
    }
    for (int64_t c0 = 0; c0 < n; c0 += sib::SIB_NC) {
        const int nc = (int)(n - c0 < sib::SIB_NC ? n - c0 : sib::SIB_NC);
        const float * xc = reinterpret_cast<const float *>(reinterpret_cast<const char *>(x_dev) + (size_t)c0 * nb11);
        float * yc = y_dev + (size_t)c0 * ldy;
        int rc;
#define SIB_DISPATCH(T)                                                                                                   \
        do {                                                                                                              \
            if (S_dev)        rc = sib_launch<T, true, sib::SIB_NC>(w, xc, nb11, yc, ldy, nc, S_dev, n, c0);              \
            else if (nc == 1) rc = sib_launch<T, false, 1>(w, xc, nb11, yc, ldy, nc, nullptr, n, c0); /* decode */        \
            else              rc = sib_launch<T, false, sib::SIB_NC>(w, xc, nb11, yc, ldy, nc, nullptr, n, c0);           \
        } while (0)
        if (w->type == GGML_B200_TYPE_Q8_0)      SIB_DISPATCH(GGML_B200_TYPE_Q8_0);
        else if (w->type == GGML_B200_TYPE_Q4_1) SIB_DISPATCH(GGML_B200_TYPE_Q4_1);
        else if (w->type == GGML_B200_TYPE_Q5_0) SIB_DISPATCH(GGML_B200_TYPE_Q5_0);
        else                                     SIB_DISPATCH(GGML_B200_TYPE_Q5_1);
#undef SIB_DISPATCH
        if (rc != GGML_B200_OK) return rc;
    }
    return GGML_B200_OK;
}

extern "C" int ggml_b200_mul_mat_quant(const ggml_b200_qweight * w, const float * x_dev, int64_t n, size_t nb11, float * y_dev, size_t nb1) {
    B200_API_LOCK;
    return sib_mul_mat(w, x_dev, n, nb11, y_dev, nb1, nullptr);
}

extern "C" int ggml_b200_block_dots_quant(const ggml_b200_qweight * w, const float * x_dev, int64_t n, size_t nb11, int32_t * S_dev, float * y_dev, size_t nb1) {
    B200_API_LOCK;
    if (!S_dev) { set_error("ggml_b200_block_dots_quant: S_dev is null"); return GGML_B200_ERR_ARG; }
    return sib_mul_mat(w, x_dev, n, nb11, y_dev, nb1, S_dev);
}

extern "C" int ggml_b200_quantize_dot_blocks(int type, const float * x_dev, int64_t n, int64_t K, size_t nb11, void * blocks_dev) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_quantize_dot_blocks: library not initialised"); return GGML_B200_ERR_STATE; }
    if (!sib_type_ok(type) || !x_dev || !blocks_dev || n < 0 || K <= 0 || K % 32 != 0) { set_error("ggml_b200_quantize_dot_blocks: bad arguments"); return GGML_B200_ERR_ARG; }
    if (n == 0) return GGML_B200_OK;
    if (nb11 == 0) nb11 = (size_t)K * 4;
    const long long warps = (long long)n * (K / 32);
    const unsigned grid = (unsigned)((warps * 32 + 255) / 256);
    // the activation format of the weight type: q8_0 for Q8_0 / Q5_0, q8_1 for Q4_1 / Q5_1
    if (!sib::act_q8_1(type)) sib::sib_quantize_blocks_kernel<GGML_B200_TYPE_Q8_0><<<grid, 256, 0, g.stream>>>(x_dev, nb11, n, (int)K, (unsigned char *)blocks_dev);
    else                      sib::sib_quantize_blocks_kernel<GGML_B200_TYPE_Q4_1><<<grid, 256, 0, g.stream>>>(x_dev, nb11, n, (int)K, (unsigned char *)blocks_dev);
    CK(cudaGetLastError());
    g.launches++;
    return GGML_B200_OK;
}
