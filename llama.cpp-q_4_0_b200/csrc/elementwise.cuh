// elementwise.cuh -- the ops that sit BETWEEN consecutive Q4_0 mul_mats of a LLaMA layer, so that a
// graph's activations can stay in HBM from one matmul to the next instead of crossing PCIe per node:
//   rms_norm (ggml_compute_forward_rms_norm_f32, ggml.c:10400-10450), mul with a broadcast row (norm weights;
//   ggml_compute_forward_mul_f32, ggml.c:9160-9240), add (residual; ggml.c:8480-8560), silu (ggml_vec_silu_f32,
//   ggml.c:3530-3538: the reference looks silu up in an fp16 table, i.e. y = fp16(silu(fp16(x))) -- reproduced here).
// New kernels (the reference's CUDA versions, ggml-cuda.cu:217-281, compute silu and rms_norm in plain fp32 and differ
// from its own CPU code; these follow the CPU code, which is the oracle).  HBM/L2-bound streaming, 16-64 KB per call at
// n = 1: their cost is the launch; all take the programmatic-dependent-launch edge like the matmul kernels.
#pragma once
#include "common.cuh"

namespace b200 {

// one CTA per row; sum of squares in double like the reference (ggml_float), mean = (float)(sum / ncols).
// w != nullptr: the MUL by the norm weights that follows in every LLaMA layer fused in (y2 = y * w[i % nw]); both
// results are written, so whoever else reads the norm's own result still finds it.
__global__ void __launch_bounds__(256) rms_norm_f32_kernel(const float * __restrict__ x, float * __restrict__ y, int ncols, float eps,
                                                           const float * __restrict__ w, float * __restrict__ y2, int nw, unsigned seq) {
    CtaTrace tr;                       // (optional timeline, ggml_b200_trace_begin; element-wise kernels carry bit 31 in their number)
    if (threadIdx.x == 0) tr.begin(seq | 0x80000000u);
    pdl_wait();
    pdl_launch_dependents();
    if (threadIdx.x == 0) tr.mark(1);
    const float * xr = x + (size_t)blockIdx.x * ncols;
    float * yr = y + (size_t)blockIdx.x * ncols;
    double s = 0.0;
    for (int i = threadIdx.x; i < ncols; i += blockDim.x) { const float v = __ldcg(xr + i); s += (double)(v * v); }
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    __shared__ double part[8];
    __shared__ float s_scale;
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) t += part[w];
        const float mean = (float)(t / ncols);
        s_scale = __fdiv_rn(1.0f, __fsqrt_rn(mean + eps));
    }
    __syncthreads();
    const float scale = s_scale;
    for (int i = threadIdx.x; i < ncols; i += blockDim.x) {
        const float v = __fmul_rn(__ldcg(xr + i), scale);
        yr[i] = v;
        if (w != nullptr) y2[(size_t)blockIdx.x * ncols + i] = __fmul_rn(v, __ldcg(w + (((size_t)blockIdx.x * ncols + i) % nw)));
    }
    if (threadIdx.x == 0) tr.mark(3);
}

// The same op for the shapes a decode graph has (one row of a few thousand floats, 16-byte aligned, norm weights of one
// row): the single CTA is the whole critical path of the layer at that point, so the row is read ONCE with 16-byte
// loads into registers (NV float4 per thread, 1024 threads), reduced in double with shuffles, and written with 16-byte
// stores.  2.x us instead of 13 us for 4096 floats (profiles/r02_trace_seam*.log): the old kernel's 16 dependent
// L2 round trips per thread and pass were the cost, not the arithmetic.  Same sums in double (the order of a double
// tree vs the reference's serial loop changes the float mean by less than its rounding), same three fp32 operations per
// element.  wmode 0: no weights, 1: w has ncols entries (every row the same), 2: w has nrows * ncols entries.
template <int NV>
__global__ void __launch_bounds__(1024) rms_norm_f32_row_kernel(const float * __restrict__ x, float * __restrict__ y, int ncols, float eps,
                                                                const float * __restrict__ w, float * __restrict__ y2, int wmode, unsigned seq) {
    CtaTrace tr;
    if (threadIdx.x == 0) tr.begin(seq | 0x80000000u);
    const int nv = ncols >> 2;                           // float4 per row
    pdl_wait();
    pdl_launch_dependents();
    if (threadIdx.x == 0) tr.mark(1);
    const size_t row = (size_t)blockIdx.x * ncols;
    const float4 * xr = reinterpret_cast<const float4 *>(x + row);
    float4 v[NV];
    double s = 0.0;
#pragma unroll
    for (int j = 0; j < NV; j++) {
        const int i = threadIdx.x + j * 1024;
        if (i < nv) v[j] = __ldcg(xr + i);
    }
    float4 wv[NV];                                       // (requested together with x: one L2 round trip for both)
    if (wmode != 0) {
        const float4 * wr = reinterpret_cast<const float4 *>(wmode == 1 ? w : w + row);
#pragma unroll
        for (int j = 0; j < NV; j++) { const int i = threadIdx.x + j * 1024; if (i < nv) wv[j] = __ldcg(wr + i); }
    }
#pragma unroll
    for (int j = 0; j < NV; j++) {
        const int i = threadIdx.x + j * 1024;
        if (i < nv) s += (double)(v[j].x * v[j].x) + (double)(v[j].y * v[j].y) + (double)(v[j].z * v[j].z) + (double)(v[j].w * v[j].w);
    }
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    __shared__ double part[32];
    __shared__ float s_scale;
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        double t = threadIdx.x < (blockDim.x >> 5) ? part[threadIdx.x] : 0.0;
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (threadIdx.x == 0) {
            const float mean = (float)(t / ncols);
            s_scale = __fdiv_rn(1.0f, __fsqrt_rn(mean + eps));
        }
    }
    __syncthreads();
    const float scale = s_scale;
    float4 * yr = reinterpret_cast<float4 *>(y + row);
    float4 * y2r = reinterpret_cast<float4 *>(y2 + row);
#pragma unroll
    for (int j = 0; j < NV; j++) {
        const int i = threadIdx.x + j * 1024;
        if (i < nv) {
            float4 o = make_float4(__fmul_rn(v[j].x, scale), __fmul_rn(v[j].y, scale), __fmul_rn(v[j].z, scale), __fmul_rn(v[j].w, scale));
            yr[i] = o;
            if (wmode != 0) {
                const float4 ww = wv[j];
                y2r[i] = make_float4(__fmul_rn(o.x, ww.x), __fmul_rn(o.y, ww.y), __fmul_rn(o.z, ww.z), __fmul_rn(o.w, ww.w));
            }
        }
    }
    if (threadIdx.x == 0) tr.mark(3);
}

// OP 0: y = a * b[i % nb]   1: y = a + b[i % nb]   2: y = silu(a) (fp16 table semantics)   3: y = silu(a) * b (fused pair;
// s != nullptr also receives silu(a))
template <int OP>
__global__ void __launch_bounds__(256) eltwise_f32_kernel(const float * __restrict__ a, const float * __restrict__ b, float * __restrict__ y,
                                                          long long n, long long nb, float * __restrict__ s, unsigned seq) {
    CtaTrace tr;
    if (threadIdx.x == 0) tr.begin(seq | 0x80000000u);
    pdl_wait();
    pdl_launch_dependents();
    if (threadIdx.x == 0) tr.mark(1);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float av = __ldcg(a + i);
        float r;
        if (OP == 0) r = __fmul_rn(av, __ldcg(b + (nb == n ? i : i % nb)));
        else if (OP == 1) r = __fadd_rn(av, __ldcg(b + (nb == n ? i : i % nb)));
        else {
            const float xf = __half2float(__float2half_rn(av));
            const float sv = __half2float(__float2half_rn(__fdiv_rn(xf, 1.0f + expf(-xf))));
            r = OP == 2 ? sv : __fmul_rn(sv, __ldcg(b + i));
            if (OP == 3 && s != nullptr) s[i] = sv;
        }
        y[i] = r;
    }
    if (threadIdx.x == 0) tr.mark(3);
}

// rope, mode 0 (ggml_compute_forward_rope_f32, ggml.c:12714-12850, the `!is_neox` branch): adjacent pairs of every row
// are rotated by the angle of the row's position; tab[pos][k] = {cos, sin} of theta_k comes from the HOST (the glue runs
// the reference's own float recurrence through the host's libm, so the angles are the reference's bit for bit).  The node
// sits right behind wq / wk in the graph (llama.cpp:1342-1347); on the CPU it costs 4096 cosf + sinf per layer and token.
// x: [npos][rows_per_pos][ne0]; one thread per pair; y may alias x (rope_inplace).
__global__ void __launch_bounds__(256) rope_f32_kernel(const float * __restrict__ x, float * __restrict__ y, const float2 * __restrict__ tab,
                                                       int ne0, int rows_per_pos, long long npairs, unsigned seq) {
    CtaTrace tr;
    if (threadIdx.x == 0) tr.begin(seq | 0x80000000u);
    pdl_wait();
    pdl_launch_dependents();
    if (threadIdx.x == 0) tr.mark(1);
    const int half = ne0 >> 1;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < npairs; i += (long long)gridDim.x * blockDim.x) {
        const int k = (int)(i % half);
        const long long row = i / half;
        const long long pos = row / rows_per_pos;
        const float2 cs = __ldcg(tab + pos * half + k);
        const float2 v = __ldcg(reinterpret_cast<const float2 *>(x) + i);
        float2 o;
        o.x = __fsub_rn(__fmul_rn(v.x, cs.x), __fmul_rn(v.y, cs.y));
        o.y = __fadd_rn(__fmul_rn(v.x, cs.y), __fmul_rn(v.y, cs.x));
        reinterpret_cast<float2 *>(y)[i] = o;
    }
    if (threadIdx.x == 0) tr.mark(3);
}

// a small row between HBM and MAPPED pinned host memory (either direction), 16 bytes per thread (ggml-b200.cu: small_copy)
__global__ void __launch_bounds__(128) copy_rows_kernel(uint4 * __restrict__ dst, const uint4 * __restrict__ src, int n16, unsigned seq) {
    CtaTrace tr;
    if (threadIdx.x == 0) tr.begin(seq | 0x80000000u);
    pdl_wait();
    pdl_launch_dependents();
    if (threadIdx.x == 0) tr.mark(1);
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (i < n16) {
        uint4 v;
        asm volatile("ld.global.cv.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(src + i) : "memory");
        dst[i] = v;
    }
    if (threadIdx.x == 0) tr.mark(3);
}

}  // namespace b200
