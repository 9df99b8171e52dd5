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
                                                           const float * __restrict__ w, float * __restrict__ y2, int nw) {
    pdl_wait();
    pdl_launch_dependents();
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
}

// OP 0: y = a * b[i % nb]   1: y = a + b[i % nb]   2: y = silu(a) (fp16 table semantics)   3: y = silu(a) * b (fused pair;
// s != nullptr also receives silu(a))
template <int OP>
__global__ void __launch_bounds__(256) eltwise_f32_kernel(const float * __restrict__ a, const float * __restrict__ b, float * __restrict__ y,
                                                          long long n, long long nb, float * __restrict__ s = nullptr) {
    pdl_wait();
    pdl_launch_dependents();
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
}

}  // namespace b200
