// gemm_q4_0_tc.cu -- host side of kernel 3 (gemm_q4_0_tc3.cuh): TMA tensor maps, launch, tuning.
//
// cuTensorMapEncodeTiled is fetched through cudaGetDriverEntryPoint so the library keeps no
// link-time dependency on libcuda (it must load, and export its ABI, on a GPU-less build host).
#include <cuda.h>
#include <stdlib.h>
#include <cuda_runtime.h>
#include <string.h>

#include "gemm_q4_0_tc3.cuh"
#include "state.h"

namespace b200 {

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static bool g_gemm_ready = false;

struct GemmWeight {
    CUtensorMap tm_nib;
};

static int encode_2d(CUtensorMap * tm, CUtensorMapDataType dt, void * base, uint64_t inner, uint64_t outer,
                     uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_outer, CUtensorMapSwizzle sw) {
    const cuuint64_t dims[2] = {inner, outer};
    const cuuint64_t strides[1] = {row_stride_bytes};
    const cuuint32_t box[2] = {box_inner, box_outer};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(tm, dt, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d): inner %llu outer %llu stride %llu box %ux%u", (int)r,
                  (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)row_stride_bytes, box_inner, box_outer);
        return GGML_B200_ERR_CUDA;
    }
    return GGML_B200_OK;
}

int gemm_init() {
    void * fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || fn == nullptr) {
        cudaGetLastError();
        set_error("cuTensorMapEncodeTiled is not available from the driver");
        return GGML_B200_ERR_CUDA;
    }
    g_encode = (EncodeTiledFn)fn;
    CK(cudaFuncSetAttribute(g3::gemm_q4_0_tc3_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, g3::SMEM_BYTES));
    CK(cudaFuncSetAttribute(g3::gemm_q4_0_tc3_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, g3::SMEM_BYTES));
    CK(cudaFuncSetAttribute(g3::gemm_q4_0_tc3_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, g3::SMEM_BYTES));
    CK(cudaFuncSetAttribute(g3::gemm_q4_0_tc3_kernel<false, false, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, g3::SMEM_BYTES));
    CK(cudaFuncSetAttribute(g3::gemm_q4_0_tc3_kernel<false, false, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, g3::SMEM_BYTES));
    CK(cudaFuncSetAttribute(g3::gemm_q4_0_tc3_kernel<true, false, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, g3::SMEM_BYTES));
    g_gemm_ready = true;
    return GGML_B200_OK;
}

void gemm_shutdown() { g_gemm_ready = false; }

// kernel 1 writes the block-major activation scales in kernel 3's column order (gemm_q4_0_tc3.cuh xd_perm)
int gemm_xd_perm() { return 1; }
long long * g_gemm_trace = nullptr;   // debug timeline buffer (tools/trace_gemm.py)
int gemm_set_tuning(const char * key, int value) {
    if (!strcmp(key, "gemm_trace")) {   // value != 0: allocate and arm the 4 x 2048 clock64 timeline buffer
        if (g_gemm_trace) { cudaFree(g_gemm_trace); g_gemm_trace = nullptr; }
        if (value) { CK(cudaMalloc(&g_gemm_trace, 4 * 2048 * 8)); CK(cudaMemset(g_gemm_trace, 0, 4 * 2048 * 8)); }
        return GGML_B200_OK;
    }
    return GGML_B200_ERR_ARG;
}

bool gemm_available(const ggml_b200_weight * w) {
    return g_gemm_ready && w && w->layout == GGML_B200_LAYOUT_PLANES && w->gemm != nullptr;
}

int gemm_prepare_weight(ggml_b200_weight * w) {
    if (!g_gemm_ready || w->layout != GGML_B200_LAYOUT_PLANES || w->rows() == 0) return GGML_B200_OK;
    GemmWeight * gw = new GemmWeight();
    const uint64_t nbk = (uint64_t)(w->K / QK);
    // nibble plane as bytes: [rows][nbk * 16], box = 4 blocks (64 B) x 128 rows, no swizzle
    int rc = encode_2d(&gw->tm_nib, CU_TENSOR_MAP_DATA_TYPE_UINT8, w->wq, nbk * 16, (uint64_t)w->rows(), nbk * 16,
                       g3::KB * 16, g3::BM, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) { delete gw; return rc; }
    w->gemm = gw;
    return GGML_B200_OK;
}

void gemm_release_weight(ggml_b200_weight * w) {
    if (w->gemm) { delete (GemmWeight *)w->gemm; w->gemm = nullptr; }
}

// Q8_0 weights (siblings.cu): the int8 plane [M][K] is the A operand as it lies -- box = 128 K-bytes x 128 rows, SW128
void * gemm_prepare_q8(const void * wq, int64_t M, int64_t K) {
    if (!g_gemm_ready || M <= 0) return nullptr;
    GemmWeight * gw = new GemmWeight();
    if (encode_2d(&gw->tm_nib, CU_TENSOR_MAP_DATA_TYPE_UINT8, const_cast<void *>(wq), (uint64_t)K, (uint64_t)M, (uint64_t)K, g3::KB * 32, g3::BM,
                  CU_TENSOR_MAP_SWIZZLE_128B)) { delete gw; return nullptr; }
    return gw;
}
void gemm_release_q8(void * h) { delete (GemmWeight *)h; }

static int launch_gemm_any(const CUtensorMap & tm_w, const __half * wd, int64_t M, int64_t Kw, int wfmt, float * y, long long ldy, int n, int * S, int independent,
                           const g3::PeerOut * peers = nullptr);

int launch_gemm_q8(const void * h, const void * wd, int64_t M, int64_t K, float * y, long long ldy, int n, int * S) {
    if (!g_gemm_ready || !h) { set_error("tensor-core GEMM path not available for this weight"); return GGML_B200_ERR_STATE; }
    return launch_gemm_any(((const GemmWeight *)h)->tm_nib, (const __half *)wd, M, K, 8, y, ldy, n, S, 0);
}

// multi-GPU prefill: the tiles also go to the peers' landing buffers (comm.cu); dst[i] / ctr[i] for the ndst peers
int launch_gemm_peers(const ggml_b200_weight * w, float * y, long long ldy, int n, float * const * dst, unsigned * const * ctr, int ndst, long long ld) {
    if (!gemm_available(w)) { set_error("tensor-core GEMM path not available for this weight"); return GGML_B200_ERR_STATE; }
    g3::PeerOut po = {};
    if (ndst > 8) { set_error("too many peers"); return GGML_B200_ERR_ARG; }
    for (int i = 0; i < ndst; i++) { po.dst[i] = dst[i]; po.ctr[i] = ctr[i]; }
    po.ndst = ndst; po.ld = ld;
    return launch_gemm_any(((const GemmWeight *)w->gemm)->tm_nib, w->wd, w->rows(), w->K, 4, y, ldy, n, nullptr, 0, &po);
}

int launch_gemm(const ggml_b200_weight * w, float * y, long long ldy, int n, int * S, int independent) {
    if (!gemm_available(w)) { set_error("tensor-core GEMM path not available for this weight"); return GGML_B200_ERR_STATE; }
    return launch_gemm_any(((const GemmWeight *)w->gemm)->tm_nib, w->wd, w->rows(), w->K, 4, y, ldy, n, S, independent);
}

static int launch_gemm_any(const CUtensorMap & tm_w, const __half * wd, int64_t M, int64_t Kw, int wfmt, float * y, long long ldy, int n, int * S, int independent,
                           const g3::PeerOut * peers) {
    if (!g.xdT || g.xdT_n_pad < n) { set_error("activation scales were not laid out for the GEMM path"); return GGML_B200_ERR_STATE; }
    const uint64_t K = (uint64_t)Kw, nbk = K / QK;
    CUtensorMap tm_xq, tm_xd;
    // activations int8 [n][K], box = 128 K-bytes x 128 columns, 128-byte swizzle (UMMA K-major SW128)
    int rc = encode_2d(&tm_xq, CU_TENSOR_MAP_DATA_TYPE_UINT8, g.xq, K, (uint64_t)n, K, g3::KB * 32, g3::BN, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    // activation scales block-major [nbk][n_pad] f32, box = 128 columns x 4 blocks
    rc = encode_2d(&tm_xd, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, g.xdT, (uint64_t)g.xdT_n_pad, nbk, (uint64_t)g.xdT_n_pad * 4,
                   g3::BN, g3::KB, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;

    const dim3 grid((unsigned)((M + g3::BM - 1) / g3::BM), (unsigned)((n + g3::BN - 1) / g3::BN));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = dim3(g3::THREADS); cfg.dynamicSmemBytes = g3::SMEM_BYTES; cfg.stream = g.stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = g.tune.pdl ? 1 : 0;
    const int rows = (int)M, nbk_i = (int)nbk;
    g3::PeerOut po = {};
    if (peers) po = *peers;
    if (peers && peers->ndst > 0 && wfmt == 4 && !S) {
        CK(cudaLaunchKernelEx(&cfg, g3::gemm_q4_0_tc3_kernel<false, false, 4, true>, tm_w, tm_xq, tm_xd, wd, y, ldy, rows, n, nbk_i, S, g_gemm_trace, independent, po));
    }
    else if (wfmt == 8) {
        if (S) CK(cudaLaunchKernelEx(&cfg, g3::gemm_q4_0_tc3_kernel<true, false, 8>, tm_w, tm_xq, tm_xd, wd, y, ldy, rows, n, nbk_i, S, g_gemm_trace, independent, po));
        else   CK(cudaLaunchKernelEx(&cfg, g3::gemm_q4_0_tc3_kernel<false, false, 8>, tm_w, tm_xq, tm_xd, wd, y, ldy, rows, n, nbk_i, S, g_gemm_trace, independent, po));
    }
    else if (g_gemm_trace && !S) CK(cudaLaunchKernelEx(&cfg, g3::gemm_q4_0_tc3_kernel<false, true>, tm_w, tm_xq, tm_xd, wd, y, ldy, rows, n, nbk_i, S, g_gemm_trace, independent, po));
    else if (S)             CK(cudaLaunchKernelEx(&cfg, g3::gemm_q4_0_tc3_kernel<true, false>, tm_w, tm_xq, tm_xd, wd, y, ldy, rows, n, nbk_i, S, g_gemm_trace, independent, po));
    else                    CK(cudaLaunchKernelEx(&cfg, g3::gemm_q4_0_tc3_kernel<false, false>, tm_w, tm_xq, tm_xd, wd, y, ldy, rows, n, nbk_i, S, g_gemm_trace, independent, po));
    g.launches++;
    return GGML_B200_OK;
}

}  // namespace b200

// debug: copy the kernel-3 timeline (armed with set_tuning("gemm_trace", 1)) to the host: 4 roles x 2048 clock64 stamps
extern "C" int ggml_b200_gemm_trace_read(long long * out, int count) {
    if (!b200::g_gemm_trace) { b200::set_error("gemm trace not armed"); return GGML_B200_ERR_STATE; }
    if (count > 4 * 2048) count = 4 * 2048;
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaMemcpy(out, b200::g_gemm_trace, (size_t)count * 8, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return b200::cuda_fail(e, "trace read", __FILE__, __LINE__);
    return GGML_B200_OK;
}
