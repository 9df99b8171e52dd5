// set_q4_0.cu -- host side of the persistent multi-op GEMV (gemv_q4_0_set.cuh): compile a list of n = 1
// Q4_0 x F32 mul_mats (the decode step of a model) into a device-resident op table, run it as ONE cooperative
// launch.  Replaces, for such a list, the per-node loop of ggml_graph_compute over GGML_OP_MUL_MAT nodes
// (ggml.c:17541-17720 -> ggml_compute_forward_mul_mat_q_f32, ggml.c:11226-11424); result bits are those of
// ggml_b200_mul_mat_q4_0 called per op in list order.
#include "../../include/ggml-b200.h"

#include <cuda_runtime.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "common.cuh"
#include "gemv_q4_0_set.cuh"
#include "state.h"

struct ggml_b200_set {
    std::vector<b200::SetOp> host_ops;
    b200::SetOp * dev_ops = nullptr;
    b200::SetOp * dev_ops_dump = nullptr;   // scratch table for the checkpoint launch
    b200::SetSync * sync = nullptr;
    unsigned long long * staging = nullptr;   // tagged dst rows of the ops a later op reads as x
    int linked_runs = 0;
    bool big = false;                       // some op has K > 16384: the kernel build with the 6-blocks-per-lane plan
    int nslots = 0, slot_bytes = 0, x_bytes = 0, x_bufs = 1, tbl_bytes = 0;      // x_bytes: all x buffers together
    size_t smem = 0;
    int barriers = 0;
    int warps = 16;
    int grid = 0;
    uint64_t bytes = 0;                     // algorithmic bytes of one run
    // multi-GPU list (ggml_b200_set_create_sharded): this rank's arena and the peers' as mapped here
    bool gathered = false;
    unsigned long long * arena = nullptr;
    void * mapped[b200::P2P_MAX_WORLD] = {nullptr};
    b200::SetGather gth = {};
};

namespace b200 {

static bool set_ranges_overlap(const float * a_lo, const float * a_hi, const float * b_lo, const float * b_hi) { return a_lo < b_hi && b_lo < a_hi; }

template <int W, bool DUMP, bool PROF, bool BIG, bool GATHER = false>
static int set_launch_tb(const ggml_b200_set * s, const SetOp * table, long long * prof) {
    auto kern = gemv_q4_0_set_kernel<W, DUMP, PROF, BIG, GATHER>;
    B200_SET_SMEM_ATTR(kern, MAX_DYN_SMEM - 1024);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)s->grid); cfg.blockDim = dim3(set_threads<W, GATHER>()); cfg.dynamicSmemBytes = s->smem; cfg.stream = g.stream;
    // every CTA of the grid must be resident at once (dependency barriers spin on peers): cooperative launch makes
    // the driver check that instead of trusting it
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr; cfg.numAttrs = g.tune.set_coop ? 1 : 0;
    CK(cudaLaunchKernelEx(&cfg, kern, table, (int)s->host_ops.size(), s->sync, s->nslots, s->slot_bytes, s->x_bytes, s->x_bufs, s->tbl_bytes, g.tune.set_fly > 0 ? g.tune.set_fly : 1 << 30, (unsigned)g.launches, prof, s->gth));
    g.launches++;
    return GGML_B200_OK;
}

template <int W, bool DUMP, bool PROF>
static int set_launch_t(const ggml_b200_set * s, const SetOp * table, long long * prof) {
    if constexpr (!DUMP && !PROF && W == SET_GATHER_WARPS_CONSUMER) {
        if (s->gathered) return s->big ? set_launch_tb<W, false, false, true, true>(s, table, prof) : set_launch_tb<W, false, false, false, true>(s, table, prof);
    }
    if (s->gathered) { set_error("multi-GPU lists run %d consumer warps; checkpoint / profile launches are single-GPU only", SET_GATHER_WARPS_CONSUMER); return GGML_B200_ERR_ARG; }
    return s->big ? set_launch_tb<W, DUMP, PROF, true>(s, table, prof) : set_launch_tb<W, DUMP, PROF, false>(s, table, prof);
}

template <bool DUMP, bool PROF>
static int set_launch(const ggml_b200_set * s, const SetOp * table, long long * prof = nullptr) {
    switch (s->warps) {
        case 8:  return set_launch_t<8, DUMP, PROF>(s, table, prof);
        case 12: return set_launch_t<12, DUMP, PROF>(s, table, prof);
        case 20: return set_launch_t<20, DUMP, PROF>(s, table, prof);
        default: return set_launch_t<16, DUMP, PROF>(s, table, prof);
    }
}


// ---- hazard analysis of a list (pure host code; ggml_b200_set_plan exposes it to the CPU-only tests) ----------------
// The list must behave like its ops run one after the other.  Per op (non-empty shards only), against the ops still
// "live" (not yet covered by a barrier):
//   * x is exactly the dst of ONE live earlier op          -> the run reads x out of that op's tagged staging rows
//     (gemv_q4_0_set.cuh): the wait is per word, no barrier;
//   * x overlaps live dsts in any other way, or dst overlaps a live x or dst (WAR / WAW)   -> dependency barrier;
//   * otherwise x is not produced inside the list (or was before a barrier that covers it): plain loads.
// Ops that share x with the head of the current run and need no barrier join it (no re-quantization).
struct SetPlanOp { const float * x, * x_hi, * y_lo, * y_hi; long long K; bool whole; };
struct SetPlan {
    std::vector<int> flags;       // SET_OP_* | arrivals << SET_OP_ARRIVE_SHIFT, per op
    std::vector<int> link;        // per op: index of the op whose staging rows its x is read from (heads of linked runs), or -1
    int barriers = 0, linked_runs = 0;
};
// returns -1, or the index of the op behind which too many barriers would arrive
static int set_plan_hazards(const std::vector<SetPlanOp> & ops, const Tune & tune, SetPlan & out) {
    const int n = (int)ops.size();
    out.flags.assign(n, 0);
    out.link.assign(n, -1);
    struct Range { const float * lo, * hi; int op; bool whole; };
    std::vector<Range> live_x, live_y;
    const float * run_x = nullptr; long long run_K = 0; int run_link = -1;
    int arrive_mono = -1;
    // hazards of op j against the ops still live: does it need a barrier, behind which op, is its x one op's dst
    bool raw_bar = false;                  // the barrier being planned orders a read of x behind its writers (not only WAR / WAW)
    auto hazards = [&](int j, bool & bar_, int & after_, int & link_) {
        const SetPlanOp & o = ops[j];
        for (const Range & r : live_x) if (set_ranges_overlap(o.y_lo, o.y_hi, r.lo, r.hi)) { bar_ = true; after_ = std::max(after_, r.op); }      // WAR
        for (const Range & r : live_y) if (set_ranges_overlap(o.y_lo, o.y_hi, r.lo, r.hi)) { bar_ = true; after_ = std::max(after_, r.op); }      // WAW
        int writers = 0, last_writer = -1;
        link_ = -1;
        for (const Range & r : live_y) {
            if (!set_ranges_overlap(o.x, o.x_hi, r.lo, r.hi)) continue;
            writers++;
            last_writer = std::max(last_writer, r.op);
            if (r.lo == o.x && r.hi == o.x_hi && r.whole) link_ = r.op; else { bar_ = true; raw_bar = true; }      // RAW
        }
        if (writers > 1 || (link_ >= 0 && !tune.set_link)) { bar_ = true; raw_bar = true; }
        if (bar_) after_ = std::max(after_, last_writer);
    };
    for (int i = 0; i < n; i++) {
        const SetPlanOp & o = ops[i];
        bool barrier = false;
        int after = -1;                       // last earlier op this one conflicts with
        int link = -1;
        raw_bar = false;
        hazards(i, barrier, after, link);
        if (barrier) {
            // Split barrier: every warp ARRIVES when it has finished the last op this run conflicts with and WAITS here,
            // usually several ops later, when the count is long complete.  Arrival points are kept monotonic (barriers
            // are numbered in the order of their arrival points); ops after the arrival point stay live (the barrier
            // says nothing about them).  The ops that will share this run's x wait here too (a barrier of their own would
            // split the run), and ops up to `set_barrier_slack` before the head are waited for as well.
            for (int j = i + 1; j < n && ops[j].x == o.x && ops[j].K == o.K; j++) {
                bool b2 = false; int l2 = -1;
                hazards(j, b2, after, l2);
            }
            const int prev = i - 1;
            after = std::max(std::max(after, arrive_mono), prev - std::max(tune.set_barrier_slack, 0));
            if (after > prev || !tune.set_split_barrier) after = prev;
            while (after >= 0 && after < prev && ((out.flags[after] >> SET_OP_ARRIVE_SHIFT) & SET_OP_ARRIVE_MASK) >= SET_OP_ARRIVE_MASK) after++;
            arrive_mono = after;
            if (link >= 0) raw_bar = true;        // x IS one op's dst, but the run reads it with plain loads behind this barrier
            auto dead = [&](const Range & r) { return r.op <= after; };
            live_x.erase(std::remove_if(live_x.begin(), live_x.end(), dead), live_x.end());
            live_y.erase(std::remove_if(live_y.begin(), live_y.end(), dead), live_y.end());
            link = -1;
        }
        const bool same_run = i > 0 && !barrier && run_x == o.x && run_K == o.K && run_link == link;
        if (!same_run) {
            const bool bar = barrier && i > 0;
            out.flags[i] = SET_OP_QUANT | (bar ? SET_OP_BARRIER : 0);
            if (bar) {
                out.barriers++;
                if (((out.flags[after] >> SET_OP_ARRIVE_SHIFT) & SET_OP_ARRIVE_MASK) >= SET_OP_ARRIVE_MASK) return after;
                out.flags[after] += 1 << SET_OP_ARRIVE_SHIFT;
                if (raw_bar) out.flags[after] |= SET_OP_ARRIVE_GATHER;      // (multi-GPU lists: the peers' rows must be in dst too)
            }
            if (link >= 0) out.linked_runs++;
            run_x = o.x; run_K = o.K; run_link = link;
        }
        out.link[i] = same_run ? -1 : link;
        live_x.push_back({o.x, o.x_hi, i, false});
        live_y.push_back({o.y_lo, o.y_hi, i, o.whole});
    }
    // x that no op of the list writes: the kernel may cache it and fetch it ahead of the run that needs it
    for (int i = 0; i < n; i++) {
        if (!(out.flags[i] & SET_OP_QUANT) || !tune.set_xext) continue;
        bool written = false;
        for (const SetPlanOp & q : ops) if (set_ranges_overlap(ops[i].x, ops[i].x_hi, q.y_lo, q.y_hi)) written = true;
        if (!written) out.flags[i] |= SET_OP_XEXT;
    }
    return -1;
}

}  // namespace b200

using namespace b200;

extern "C" {

int ggml_b200_set_plan(const ggml_b200_set_plan_op * ops, int n_ops, int * flags_out, int * link_out, int * linked_runs_out) {
    if (!ops || n_ops <= 0) { set_error("empty op list"); return GGML_B200_ERR_ARG; }
    std::vector<SetPlanOp> po;
    std::vector<int> where;             // list index of every op with a non-empty shard
    for (int i = 0; i < n_ops; i++) {
        const ggml_b200_set_plan_op & o = ops[i];
        if (o.K <= 0 || o.K % QK || o.row_lo < 0 || o.row_hi < o.row_lo || o.row_hi > o.M) { set_error("set_plan op %d: bad shape", i); return GGML_B200_ERR_ARG; }
        if (flags_out) flags_out[i] = 0;
        if (link_out) link_out[i] = -1;
        if (o.row_hi == o.row_lo) continue;
        const float * x = reinterpret_cast<const float *>((uintptr_t)o.x);
        const float * y = reinterpret_cast<const float *>((uintptr_t)o.y);
        po.push_back({x, x + o.K, y + o.row_lo, y + o.row_hi, (long long)o.K, o.row_lo == 0 && o.row_hi == o.M});
        where.push_back(i);
    }
    SetPlan plan;
    if (set_plan_hazards(po, g.tune, plan) >= 0) { set_error("too many dependency barriers behind one op of the set"); return GGML_B200_ERR_ARG; }
    for (size_t k = 0; k < where.size(); k++) {
        if (flags_out) flags_out[where[k]] = plan.flags[k];
        if (link_out && plan.link[k] >= 0) link_out[where[k]] = where[plan.link[k]];
    }
    if (linked_runs_out) *linked_runs_out = plan.linked_runs;
    return plan.barriers;
}

static ggml_b200_set * set_create_impl(const ggml_b200_set_op * ops, int n_ops, bool gathered) {
    if (!g.inited) { set_error("ggml_b200_init was not called"); return nullptr; }
    if (!ops || n_ops <= 0) { set_error("empty op list"); return nullptr; }
    ggml_b200_set * s = new ggml_b200_set();
    s->grid = g.sm_count;
    s->warps = gathered ? SET_GATHER_WARPS_CONSUMER : g.tune.set_warps;
    int NG = s->warps / SET_WPG;
    int nbk_max = 0;
    long long tiles = 0;
    std::vector<SetPlanOp> plan_ops;
    for (int i = 0; i < n_ops; i++) {
        const ggml_b200_weight * w = ops[i].w;
        const float * x = ops[i].x_dev;
        float * y = ops[i].y_dev;
        if (!w || !x || !y) { set_error("set op %d: null argument", i); delete s; return nullptr; }
        if (w->layout != GGML_B200_LAYOUT_PLANES) { set_error("set op %d: the persistent kernel needs the plane layout (repack on)", i); delete s; return nullptr; }
        if (((uintptr_t)x & 15) || ((uintptr_t)y & 3)) { set_error("set op %d: x must be 16-byte aligned (ggml tensors are, GGML_MEM_ALIGN)", i); delete s; return nullptr; }
        if (set_ranges_overlap(y + w->row_lo, y + w->row_hi, x, x + w->K)) { set_error("set op %d: dst overlaps src1 (ggml never does that: dst is a fresh tensor, ggml.c:11226)", i); delete s; return nullptr; }
        s->bytes += (uint64_t)(18 * (w->rows() * (w->K / QK)) + 4 * w->K + 4 * w->rows());
        if (w->rows() == 0 && !gathered) continue;      // (a multi-GPU list keeps it: the peers' rows are still gathered into dst)
        SetOp o = {};
        o.wq = w->wq; o.wd = reinterpret_cast<const unsigned short *>(w->wd); o.x = x; o.y = y + w->row_lo; o.S = nullptr;
        o.gM = (int)w->M; o.grow0 = (int)w->row_lo;
        o.rows = (int)w->rows(); o.nbk = (int)(w->K / QK);
        if (o.nbk > SET_MAX_NBK) { set_error("set op %d: K = %lld > %d is beyond the persistent kernel's register plan; use ggml_b200_mul_mat_q4_0", i, (long long)w->K, SET_MAX_NBK * QK); delete s; return nullptr; }
        o.tile_rows = o.nbk <= SET_MAX_NBK_WPR1 ? 8 : 4;
        const int R = o.tile_rows;
        o.ntiles = (o.rows + R - 1) / R;
        o.first_mod = (int)(tiles % s->grid);
        tiles += o.ntiles;
        const float * y_lo = o.y, * y_hi = o.y + o.rows;
        const float * x_hi = x + w->K;
        if (gathered) plan_ops.push_back({x, x_hi, y, y + w->M, (long long)w->K, true});      // every rank ends up with the whole dst
        else plan_ops.push_back({x, x_hi, y_lo, y_hi, (long long)w->K, w->row_lo == 0 && w->row_hi == w->M});
        if (o.nbk > nbk_max) nbk_max = o.nbk;
        const int stage = R * o.nbk * 16 + ((R * o.nbk * 2 + 15) & ~15);
        if (stage > s->slot_bytes) s->slot_bytes = stage;
        s->host_ops.push_back(o);
    }
    if (s->host_ops.empty()) { set_error("every op of the set has an empty row shard"); delete s; return nullptr; }
    SetPlan plan;
    {
        const int bad = set_plan_hazards(plan_ops, g.tune, plan);
        if (bad >= 0) { set_error("too many dependency barriers behind op %d of the set", bad); delete s; return nullptr; }
    }
    for (size_t i = 0; i < s->host_ops.size(); i++) s->host_ops[i].flags = plan.flags[i];
    s->barriers = plan.barriers;
    s->linked_runs = plan.linked_runs;
    const std::vector<int> & link_of = plan.link;
    s->slot_bytes = (s->slot_bytes + 127) & ~127;
    s->x_bytes = (40 * nbk_max + 127) & ~127;      // per buffer; everything behind it stays 128-byte aligned
    s->big = nbk_max > 512;
    s->tbl_bytes = (int)((s->host_ops.size() * sizeof(SetOpS) + 127) & ~(size_t)127);
    if (s->tbl_bytes > SET_MAX_TABLE_BYTES || g.tune.set_table == 0) s->tbl_bytes = 0;      // long list: descriptors stay in global memory
    for (const SetOp & o : s->host_ops) if (o.first_mod > 0xFFFF) s->tbl_bytes = 0;
    // ring depth: what fits, capped.  Measured (profiles/r01_set_sweep_v5_slots.log, LLaMA-7B list, 12 warps): 3 slots 832 us,
    // 4: 757, 5: 793, 6: 801, 8: 802 -- a ring one slot deeper than the number of consumer groups keeps the stream busy, and
    // what the ring does not take stays L1.  set_slots overrides.  A second x buffer (the x warps a whole run ahead)
    // is taken when it does not cost the ring a slot.
    auto depth = [&](int bufs) {
        const long long ring = (long long)MAX_DYN_SMEM - 1024 - (long long)s->x_bytes * bufs - SET_BAR_BYTES - SET_RED_BYTES - s->tbl_bytes - 128;
        int n = (int)(ring / s->slot_bytes);
        if (n > SET_MAX_SLOTS) n = SET_MAX_SLOTS;
        if (g.tune.set_slots > 0) { if (g.tune.set_slots < n) n = g.tune.set_slots; }
        else if (n > NG + 1) n = NG + 1;
        return n;
    };
    s->nslots = depth(1);
    // The ring must be at least as deep as there are consumer groups: the kernel's parity waits are alias-free only if a
    // group's previous tile was issued after the previous use of its next slot (gemv_q4_0_set.cuh, tile loop) -- and a
    // group without a slot of its own would idle anyway.  Huge K with 16 or 20 warps: fewer consumer warps.
    if (s->nslots >= 2 && s->nslots < NG) {
        s->warps = SET_WPG * s->nslots;
        NG = s->nslots;
        s->nslots = depth(1);
    }
    if (g.tune.set_xbufs > 1 && depth(2) == s->nslots) { s->x_bufs = 2; s->x_bytes *= 2; }
    if (s->nslots < 2) {
        set_error("K = %d: a 4-row tile (%d B) does not fit the shared-memory ring twice; use ggml_b200_mul_mat_q4_0 per op", nbk_max * QK, s->slot_bytes);
        delete s; return nullptr;
    }
    s->smem = (size_t)s->x_bytes + SET_BAR_BYTES + SET_RED_BYTES + s->tbl_bytes + (size_t)s->nslots * s->slot_bytes;
    if (gathered) {
        // arena: [2 parities][sum of M, each padded to 8 words]; every rank computes the same layout from the same list
        s->gathered = true;
        size_t words = 0;
        std::vector<size_t> goff(s->host_ops.size());
        for (size_t i = 0; i < s->host_ops.size(); i++) { goff[i] = words; words += ((size_t)s->host_ops[i].gM + 7) & ~(size_t)7; }
        cudaError_t e2 = cudaMalloc(&s->arena, 2 * words * 8);
        if (e2 == cudaSuccess) e2 = cudaMemset(s->arena, 0, 2 * words * 8);      // launch numbers start at 1
        if (e2 != cudaSuccess) { cuda_fail(e2, "set_create arena", __FILE__, __LINE__); ggml_b200_set_free(s); return nullptr; }
        if (comm_share(s->arena, s->mapped) != GGML_B200_OK) { ggml_b200_set_free(s); return nullptr; }
        s->gth.stride = (long long)words; s->gth.me = ggml_b200_comm_rank(); s->gth.world = ggml_b200_comm_world();
        for (int r = 0; r < P2P_MAX_WORLD; r++) s->gth.arena[r] = (unsigned long long *)s->mapped[r];
        for (size_t i = 0; i < s->host_ops.size(); i++) {
            s->host_ops[i].ytag = s->arena + goff[i] + s->host_ops[i].grow0;                 // this shard's first row (parity 0)
            if (link_of[i] >= 0) s->host_ops[i].xtag = s->arena + goff[link_of[i]];           // the whole gathered vector
        }
    } else {
        size_t words = 0;
        std::vector<long long> off(s->host_ops.size(), -1);
        for (size_t i = 0; i < link_of.size(); i++) {
            const int k = link_of[i];
            if (k >= 0 && off[k] < 0) { off[k] = (long long)words; words += ((size_t)s->host_ops[k].rows + 7) & ~(size_t)7; }
        }
        if (words) {
            cudaError_t e2 = cudaMalloc(&s->staging, words * 8);
            if (e2 == cudaSuccess) e2 = cudaMemset(s->staging, 0, words * 8);      // launch numbers start at 1
            if (e2 != cudaSuccess) { cuda_fail(e2, "set_create staging", __FILE__, __LINE__); ggml_b200_set_free(s); return nullptr; }
            for (size_t i = 0; i < s->host_ops.size(); i++) if (off[i] >= 0) s->host_ops[i].ytag = s->staging + off[i];
            for (size_t i = 0; i < link_of.size(); i++) if (link_of[i] >= 0) s->host_ops[i].xtag = s->staging + off[link_of[i]];
        }
    }
    cudaError_t e = cudaMalloc(&s->dev_ops, s->host_ops.size() * sizeof(SetOp));
    if (e == cudaSuccess) e = cudaMalloc(&s->sync, sizeof(SetSync) + 4 * (size_t)(s->barriers + 1));      // + one arrival counter per barrier
    if (e == cudaSuccess) e = cudaMemcpy(s->dev_ops, s->host_ops.data(), s->host_ops.size() * sizeof(SetOp), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemset(s->sync, 0, sizeof(SetSync) + 4 * (size_t)(s->barriers + 1));
    if (e != cudaSuccess) { cuda_fail(e, "set_create", __FILE__, __LINE__); ggml_b200_set_free(s); return nullptr; }
    return s;
}

void ggml_b200_set_free(ggml_b200_set * s) {
    B200_API_LOCK;
    if (!s) return;
    if (s->gathered) {
        cudaDeviceSynchronize();       // this rank's last launch has seen every peer's rows; peers may only be as far
        comm_unshare(s->mapped);
        if (s->arena) cudaFree(s->arena);
    }
    if (s->dev_ops) cudaFree(s->dev_ops);
    if (s->dev_ops_dump) cudaFree(s->dev_ops_dump);
    if (s->sync) cudaFree(s->sync);
    if (s->staging) cudaFree(s->staging);
    delete s;
}

ggml_b200_set * ggml_b200_set_create(const ggml_b200_set_op * ops, int n_ops) {
    B200_API_LOCK;
    return set_create_impl(ops, n_ops, false);
}

ggml_b200_set * ggml_b200_set_create_sharded(const ggml_b200_set_op * ops, int n_ops) {
    B200_API_LOCK;
    if (ggml_b200_comm_world() <= 1) return set_create_impl(ops, n_ops, false);
    if (!comm_p2p_ready()) { set_error("set_create_sharded: the peers' memory is not mapped (ggml_b200_comm_init failed to set up peer access)"); return nullptr; }
    return set_create_impl(ops, n_ops, true);
}

int ggml_b200_set_run(ggml_b200_set * s) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!s) { set_error("null set"); return GGML_B200_ERR_ARG; }
    return set_launch<false, false>(s, s->dev_ops);
}

int ggml_b200_set_checkpoint(ggml_b200_set * s, int32_t * const * S_dev, int n_ops) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!s || !S_dev || n_ops != (int)s->host_ops.size()) { set_error("set_checkpoint: need one S pointer per non-empty op (%d)", s ? (int)s->host_ops.size() : 0); return GGML_B200_ERR_ARG; }
    std::vector<SetOp> t = s->host_ops;
    for (int i = 0; i < n_ops; i++) t[i].S = S_dev[i];
    if (!s->dev_ops_dump) CK(cudaMalloc(&s->dev_ops_dump, t.size() * sizeof(SetOp)));
    CK(cudaMemcpyAsync(s->dev_ops_dump, t.data(), t.size() * sizeof(SetOp), cudaMemcpyHostToDevice, g.stream));
    CK(cudaStreamSynchronize(g.stream));   // t is a local
    return set_launch<true, false>(s, s->dev_ops_dump);
}

int ggml_b200_set_profile(ggml_b200_set * s, long long * out_host, int max_ctas) {
    B200_API_LOCK;
    if (!g.inited) { set_error("ggml_b200_init was not called"); return GGML_B200_ERR_STATE; }
    if (!s || !out_host || max_ctas <= 0) { set_error("bad set_profile arguments"); return GGML_B200_ERR_ARG; }
    long long * dev = nullptr;
    const size_t bytes = (size_t)s->grid * 16 * sizeof(long long);
    CK(cudaMalloc(&dev, bytes));
    cudaError_t e = cudaMemsetAsync(dev, 0, bytes, g.stream);
    int rc = e == cudaSuccess ? set_launch<false, true>(s, s->dev_ops, dev) : cuda_fail(e, "memset", __FILE__, __LINE__);
    const int n = max_ctas < s->grid ? max_ctas : s->grid;
    if (rc == GGML_B200_OK) {
        e = cudaMemcpyAsync(out_host, dev, (size_t)n * 16 * sizeof(long long), cudaMemcpyDeviceToHost, g.stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(g.stream);
        if (e != cudaSuccess) rc = cuda_fail(e, "set_profile", __FILE__, __LINE__);
    }
    cudaFree(dev);
    return rc == GGML_B200_OK ? n : rc;
}

int      ggml_b200_set_barriers(const ggml_b200_set * s) { return s ? s->barriers : -1; }
int      ggml_b200_set_linked_runs(const ggml_b200_set * s) { return s ? s->linked_runs : -1; }
int      ggml_b200_set_ops(const ggml_b200_set * s) { return s ? (int)s->host_ops.size() : -1; }
int      ggml_b200_set_slots(const ggml_b200_set * s) { return s ? s->nslots : -1; }
int      ggml_b200_set_warps(const ggml_b200_set * s) { return s ? s->warps : -1; }
uint64_t ggml_b200_set_bytes(const ggml_b200_set * s) { return s ? s->bytes : 0; }

}  // extern "C"
