// set_q4_0.cu -- host side of the persistent multi-op GEMV (gemv_q4_0_set.cuh): compile a list of n = 1
// Q4_0 x F32 mul_mats (the decode step of a model) into a device-resident op table, run it as ONE cooperative
// launch.  Replaces, for such a list, the per-node loop of ggml_graph_compute over GGML_OP_MUL_MAT nodes
// (ggml.c:17541-17720 -> ggml_compute_forward_mul_mat_q_f32, ggml.c:11226-11424); result bits are those of
// ggml_b200_mul_mat_q4_0 called per op in list order.
#include "../../include/ggml-b200.h"

#include <cuda_runtime.h>
#include <string.h>

#include <vector>

#include "common.cuh"
#include "gemv_q4_0_set.cuh"
#include "state.h"

struct ggml_b200_set {
    std::vector<b200::SetOp> host_ops;
    b200::SetOp * dev_ops = nullptr;
    b200::SetOp * dev_ops_dump = nullptr;   // scratch table for the checkpoint launch
    b200::SetSync * sync = nullptr;
    int nslots = 0, slot_bytes = 0, x_bytes = 0;
    size_t smem = 0;
    int barriers = 0;
    int warps = 16;
    int grid = 0;
    uint64_t bytes = 0;                     // algorithmic bytes of one run
};

namespace b200 {

static bool set_ranges_overlap(const float * a_lo, const float * a_hi, const float * b_lo, const float * b_hi) { return a_lo < b_hi && b_lo < a_hi; }

template <int W, bool DUMP>
static int set_launch_t(const ggml_b200_set * s, const SetOp * table) {
    auto kern = gemv_q4_0_set_kernel<W, DUMP>;
    static bool attr_set = false;
    if (!attr_set) {
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_DYN_SMEM));
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)s->grid); cfg.blockDim = dim3(W * 32 + 32); cfg.dynamicSmemBytes = s->smem; cfg.stream = g.stream;
    // every CTA of the grid must be resident at once (dependency barriers spin on peers): cooperative launch makes
    // the driver check that instead of trusting it
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr; cfg.numAttrs = g.tune.set_coop ? 1 : 0;
    CK(cudaLaunchKernelEx(&cfg, kern, table, (int)s->host_ops.size(), s->sync, s->nslots, s->slot_bytes, s->x_bytes, (unsigned)g.launches));
    g.launches++;
    return GGML_B200_OK;
}

template <bool DUMP>
static int set_launch(const ggml_b200_set * s, const SetOp * table) {
    switch (s->warps) {
        case 8:  return set_launch_t<8, DUMP>(s, table);
        case 24: return set_launch_t<24, DUMP>(s, table);
        default: return set_launch_t<16, DUMP>(s, table);
    }
}

}  // namespace b200

using namespace b200;

extern "C" {

ggml_b200_set * ggml_b200_set_create(const ggml_b200_set_op * ops, int n_ops) {
    if (!g.inited) { set_error("ggml_b200_init was not called"); return nullptr; }
    if (!ops || n_ops <= 0) { set_error("empty op list"); return nullptr; }
    ggml_b200_set * s = new ggml_b200_set();
    s->grid = g.sm_count;
    s->warps = g.tune.set_warps;
    const int NG = s->warps / SET_WPG;
    int nbk_max = 0;
    long long tiles = 0;
    // the current run of mutually independent ops (same x, disjoint dst): same rule as overlap_same_src
    const float * run_x = nullptr; int64_t run_K = 0;
    std::vector<std::pair<const float *, const float *>> run_y;
    for (int i = 0; i < n_ops; i++) {
        const ggml_b200_weight * w = ops[i].w;
        const float * x = ops[i].x_dev;
        float * y = ops[i].y_dev;
        if (!w || !x || !y) { set_error("set op %d: null argument", i); delete s; return nullptr; }
        if (w->layout != GGML_B200_LAYOUT_PLANES) { set_error("set op %d: the persistent kernel needs the plane layout (repack on)", i); delete s; return nullptr; }
        if (((uintptr_t)x & 15) || ((uintptr_t)y & 3)) { set_error("set op %d: x must be 16-byte aligned (ggml tensors are, GGML_MEM_ALIGN)", i); delete s; return nullptr; }
        if (set_ranges_overlap(y + w->row_lo, y + w->row_hi, x, x + w->K)) { set_error("set op %d: dst overlaps src1 (ggml never does that: dst is a fresh tensor, ggml.c:11226)", i); delete s; return nullptr; }
        s->bytes += (uint64_t)(18 * (w->rows() * (w->K / QK)) + 4 * w->K + 4 * w->rows());
        if (w->rows() == 0) continue;
        SetOp o = {};
        o.wq = w->wq; o.wd = reinterpret_cast<const unsigned short *>(w->wd); o.x = x; o.y = y + w->row_lo; o.S = nullptr;
        o.rows = (int)w->rows(); o.nbk = (int)(w->K / QK);
        o.rpw = (g.tune.set_rpw == 1 || 8 * o.nbk * 18 > 20 * 1024) ? 1 : 2;
        const int R = SET_WPG * o.rpw;
        o.ntiles = (o.rows + R - 1) / R;
        o.first_mod = (int)(tiles % s->grid);
        tiles += o.ntiles;
        const float * y_lo = o.y, * y_hi = o.y + o.rows;
        bool independent = !s->host_ops.empty() && run_x == x && run_K == w->K;
        for (size_t k = 0; independent && k < run_y.size(); k++) independent = !set_ranges_overlap(y_lo, y_hi, run_y[k].first, run_y[k].second);
        if (!independent) {
            o.flags = SET_OP_QUANT | (s->host_ops.empty() ? 0 : SET_OP_BARRIER);
            if (o.flags & SET_OP_BARRIER) s->barriers++;
            run_x = x; run_K = w->K; run_y.clear();
        }
        run_y.push_back({y_lo, y_hi});
        if (o.nbk > nbk_max) nbk_max = o.nbk;
        const int stage = R * o.nbk * 16 + ((R * o.nbk * 2 + 15) & ~15);
        if (stage > s->slot_bytes) s->slot_bytes = stage;
        s->host_ops.push_back(o);
    }
    if (s->host_ops.empty()) { set_error("every op of the set has an empty row shard"); delete s; return nullptr; }
    s->slot_bytes = (s->slot_bytes + 127) & ~127;
    s->x_bytes = 40 * nbk_max;
    const long long ring = (long long)MAX_DYN_SMEM - s->x_bytes - 2 * SET_MAX_SLOTS * 8 - 128;
    s->nslots = (int)(ring / s->slot_bytes);
    if (s->nslots > SET_MAX_SLOTS) s->nslots = SET_MAX_SLOTS;
    if (g.tune.set_slots > 0 && g.tune.set_slots < s->nslots) s->nslots = g.tune.set_slots;
    if (s->nslots < 2) {
        set_error("K = %d: a 4-row tile (%d B) does not fit the shared-memory ring twice; use ggml_b200_mul_mat_q4_0 per op", nbk_max * QK, s->slot_bytes);
        delete s; return nullptr;
    }
    (void)NG;
    s->smem = (size_t)s->x_bytes + 2 * SET_MAX_SLOTS * 8 + (size_t)s->nslots * s->slot_bytes;
    cudaError_t e = cudaMalloc(&s->dev_ops, s->host_ops.size() * sizeof(SetOp));
    if (e == cudaSuccess) e = cudaMalloc(&s->sync, sizeof(SetSync));
    if (e == cudaSuccess) e = cudaMemcpy(s->dev_ops, s->host_ops.data(), s->host_ops.size() * sizeof(SetOp), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemset(s->sync, 0, sizeof(SetSync));
    if (e != cudaSuccess) { cuda_fail(e, "set_create", __FILE__, __LINE__); ggml_b200_set_free(s); return nullptr; }
    return s;
}

void ggml_b200_set_free(ggml_b200_set * s) {
    if (!s) return;
    if (s->dev_ops) cudaFree(s->dev_ops);
    if (s->dev_ops_dump) cudaFree(s->dev_ops_dump);
    if (s->sync) cudaFree(s->sync);
    delete s;
}

int ggml_b200_set_run(ggml_b200_set * s) {
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!s) { set_error("null set"); return GGML_B200_ERR_ARG; }
    return set_launch<false>(s, s->dev_ops);
}

int ggml_b200_set_checkpoint(ggml_b200_set * s, int32_t * const * S_dev, int n_ops) {
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!s || !S_dev || n_ops != (int)s->host_ops.size()) { set_error("set_checkpoint: need one S pointer per non-empty op (%d)", s ? (int)s->host_ops.size() : 0); return GGML_B200_ERR_ARG; }
    std::vector<SetOp> t = s->host_ops;
    for (int i = 0; i < n_ops; i++) t[i].S = S_dev[i];
    if (!s->dev_ops_dump) CK(cudaMalloc(&s->dev_ops_dump, t.size() * sizeof(SetOp)));
    CK(cudaMemcpyAsync(s->dev_ops_dump, t.data(), t.size() * sizeof(SetOp), cudaMemcpyHostToDevice, g.stream));
    CK(cudaStreamSynchronize(g.stream));   // t is a local
    return set_launch<true>(s, s->dev_ops_dump);
}

int      ggml_b200_set_barriers(const ggml_b200_set * s) { return s ? s->barriers : -1; }
int      ggml_b200_set_ops(const ggml_b200_set * s) { return s ? (int)s->host_ops.size() : -1; }
int      ggml_b200_set_slots(const ggml_b200_set * s) { return s ? s->nslots : -1; }
uint64_t ggml_b200_set_bytes(const ggml_b200_set * s) { return s ? s->bytes : 0; }

}  // extern "C"
