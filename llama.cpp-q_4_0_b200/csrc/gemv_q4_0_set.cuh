// gemv_q4_0_set.cuh -- kernel 2 for a whole LIST of n = 1 mul_mats in ONE persistent launch
// ("one launch over several weights"; the decode step of a model is such a list).
//
// Why: one launch per mul_mat costs ~1.9 us of dependent-launch boundary + x quantize + first-load latency per op
// (profiles/r01_experiments.md), more than the 1.45 us of HBM traffic of a 4096x4096 op.  Here the weight stream is
// decoupled from the op boundaries:
//   * grid = one CTA per SM, resident for the whole list (cooperative launch);
//   * ONE producer thread per CTA walks the CTA's tiles of ALL ops in list order and copies them with
//     cp.async.bulk (TMA engine, UBLKCP) into a shared-memory ring of NS slots, completion on an mbarrier per slot.
//     Weights are immutable, so the producer never waits for an op to finish: while consumers wait on a dependency
//     or change x, the NEXT ops' weights are already in flight, and HBM never idles across op boundaries.  Ring
//     depth = consumer groups + 1 (measured best; what the ring does not take stays L1);
//   * W consumer warps in groups of 4; slot j belongs to group j % NG.  A tile is R consecutive weight rows
//     (nibble plane rows + scale plane rows, both contiguous in the plane layout):
//       K <= 4096 (WPR = 1): R = 8, warp i of the group owns rows 2i, 2i+1, lane l owns blocks l, l+32, ... of them
//                            -- the arithmetic, its order and therefore the result BITS are those of
//                            gemv_q4_0_fused_kernel (S_b exact on IDP.4A, per-lane fmaf chain over b, xor-shuffle tree);
//       K >  4096 (WPR = 4): R = 4, the group's 128 lanes share every row (lane t owns blocks t, t+128, ...), the
//                            four warp partials are added in a fixed order (deterministic; S_b exact; fp32 order
//                            differs from the per-op kernel like every GPU order differs from the CPU's);
//     in both shapes a lane's slice of the quantized x (<= 6 blocks) lives in REGISTERS for the whole run of ops
//     that share x, so shared memory carries the weights only (18 B written by the TMA engine + 18 B read per
//     block): the consumers were bound by the shared-memory pipe when x was re-read per row (v1, 40 B more per
//     block and row pair; profiles/r01_set_sweep_v1.log);
//   * tiles are dealt round-robin over the CTAs CONTINUING across ops, so the per-op remainder does not idle SMs;
//   * ops that read the same x as their predecessor (wq/wk/wv, w3/w1) form a RUN and go back to back.  A change of x
//     costs a consumer warp one shared-memory wait: two x warps quantize the next run's x (quantize_row_q8_0 contract,
//     ggml.c:1192-1250, same arithmetic as the fused kernel) into the planes while the consumers work from registers;
//   * ordering follows the data hazards the host found (set_q4_0.cu, set_plan_hazards): x that is exactly one earlier
//     op's dst is read out of that op's tagged 8-byte words (the wait is per word); any other overlap is a split
//     dependency barrier -- CTAs arrive when they have finished the last conflicting op, a sentinel thread per CTA
//     forwards arrivals and waits ahead of the consumers.  Semantics == calling ggml_b200_mul_mat_q4_0 per op in order.
//   * warps of a CTA: W consumers, 1 producer, 1 sentinel, SET_XWARPS x warps.
// Roofline: HBM; algorithmic bytes per launch = sum over ops of 18*M*K/32 + 4*K + 4*M.
#pragma once
#include "common.cuh"
#include "gemv_q4_0.cuh"
#include "p2p.cuh"
#include "ptx.cuh"
#include "quantize_q8_0.cuh"

namespace b200 {

constexpr int SET_WPG = 4;            // consumer warps per group
constexpr int SET_XWARPS = 2;         // warps that quantize the NEXT run's x while the consumers work on this one
constexpr int SET_EXTRA_THREADS = 32 /* producer */ + 32 /* sentinel */ + 32 * SET_XWARPS;
// Multi-GPU lists (GATHER): every rank quantizes the whole replicated x of every run itself while it streams only 1/N
// of the weights, so the x warps, not the weight stream, become the limiter as N grows (measured at N = 2 with one x
// warp: 4.48 ms per 65B step against 2.98 ms of HBM time).  The gather build therefore runs 8 consumer warps, 5 x warps
// and 1 gather warp: still 512 threads / 128 registers.
constexpr int SET_XWARPS_GATHER = 5;
constexpr int SET_GATHER_WARPS_CONSUMER = 8;
template <bool GATHER> __host__ __device__ constexpr int set_xwarps() { return GATHER ? SET_XWARPS_GATHER : SET_XWARPS; }
template <int W, bool GATHER> __host__ __device__ constexpr int set_threads() { return W * 32 + 64 + 32 * set_xwarps<GATHER>() + (GATHER ? 32 : 0); }
constexpr int SET_MAX_SLOTS = 16;
constexpr int SET_MAX_GROUPS = 6;
constexpr int SET_RED_BYTES = 2 * SET_MAX_GROUPS * 16 * 4;   // cross-warp partials: [parity][group][row 4][warp 4] floats
constexpr int SET_BAR_BYTES = 2 * SET_MAX_SLOTS * 8;
constexpr int SET_MAX_NBK_WPR1 = 128; // blocks per row a single warp keeps x for (4 per lane)
constexpr int SET_MAX_NBK = 768;      // 6 per lane of a 4-warp team (K <= 24576)

enum { SET_OP_BARRIER = 1, SET_OP_QUANT = 2, SET_OP_XEXT = 4, SET_OP_ARRIVE_GATHER = 1 << 15 };
// ARRIVE_GATHER (multi-GPU lists): a barrier that arrives behind this op also waits until the peers' rows of every op
// up to it are in dst (a later run reads such a dst with plain loads)
// flags bits 8..14: how many dependency barriers every consumer warp arrives at when it has finished this op
constexpr int SET_OP_ARRIVE_SHIFT = 8, SET_OP_ARRIVE_MASK = 0x7F;
// XEXT: no op of the list writes x (it may be cached / fetched early)

struct SetOp {                        // 80 bytes, device resident, read-only during the launch
    const uint4 * wq;                 // nibble plane of this shard [rows][nbk]
    const unsigned short * wd;        // scale plane [rows][nbk] (fp16 bits)
    const float * x;                  // K floats, 16-byte aligned
    float * y;                        // first output row of this shard
    int * S;                          // parity checkpoint: [rows][nbk] int32 block dots, or nullptr
    unsigned long long * ytag;        // this op's rows as tagged words (launch number << 32 | bits of y), when a later op of the list reads them as x
    const unsigned long long * xtag;  // ... the tagged rows this op's x is read from, or nullptr: plain x
    int rows, nbk;
    int tile_rows;                    // R: 8 (nbk <= 128) or 4
    int ntiles;
    int first_mod;                    // (tiles of all earlier ops) % grid: CTA c takes tiles (c - first_mod) mod grid, + grid, ...
    int flags;                        // SET_OP_BARRIER | SET_OP_QUANT
    // multi-GPU lists (row shards, SetGather below): the whole tensor has gM rows, this shard starts at grow0; ytag then
    // points at this shard's first row inside this rank's ARENA (parity 0) and every peer has the same layout
    int gM, grow0;
};

// Multi-GPU lists: one rank per GPU runs the same list on its row shards (ggml_b200_set_create_sharded).  Every rank
// owns an arena [2 launch parities][sum of the ops' M] of tagged 8-byte words (launch number << 32 | bits of y),
// exported over CUDA IPC and mapped by every peer.  A consumer warp that finishes a row stores it into its own dst AND
// as a tagged word into EVERY rank's arena (NVLink stores, one lane per rank); 8-byte words are single-copy atomic, so a
// word whose number matches IS the datum: no fence, no flag, no collective call.  A run whose x is another op's dst
// reads the whole gathered vector out of the local arena (the linked-run mechanism of the single-GPU list), and one
// gather warp per CTA copies the peers' rows into dst as they land, so that every op ends with the full dst on every
// rank -- the contract of ggml_b200_mul_mat_q4_0_sharded, with NO launch and NO exchange step per mul_mat.
// The arena is double-buffered by launch parity: a peer can be at most one launch ahead (it cannot finish launch e + 1
// without this rank's rows of launch e + 1, which are produced after this rank's launch e has completed).
struct SetGather {
    unsigned long long * arena[P2P_MAX_WORLD];   // every rank's arena as mapped here; arena[me] = this rank's
    long long stride;                            // words per parity
    int me, world;
};

struct SetSync {                      // device resident, zero between launches
    unsigned arrived;                 // dependency barriers: monotonically counted arrivals of this launch
    unsigned done;                    // CTAs that finished the list; the last one re-zeroes both words
    unsigned epoch;                   // launches completed: launch number e = epoch + 1 tags the staging words
    unsigned pad;
};

__device__ __forceinline__ void st_relaxed_gpu_b64(unsigned long long * p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_gpu_b64(const unsigned long long * p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void ld_relaxed_gpu_v2b64(const unsigned long long * p, unsigned long long & a, unsigned long long & b) {
    asm volatile("ld.relaxed.gpu.global.v2.b64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}

__device__ __forceinline__ unsigned ld_relaxed_gpu_u32(const unsigned * p) {
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_gpu_add_u32(unsigned * p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

// the producer's wait for a free slot: it is usually a long wait (the ring is full most of the time), and a hot
// try_wait loop would take issue slots from the consumer warps of its scheduler
__device__ __forceinline__ void mbar_wait_backoff(uint32_t bar, uint32_t parity) {
    unsigned long long t0 = 0;
    while (!mbar_try_wait(bar, parity)) {
        __nanosleep(128);
        const unsigned long long t1 = gtime();
        if (t0 == 0) t0 = t1;
        if (t1 - t0 > 10000000000ull) __trap();
    }
}

template <int NCONS>
__device__ __forceinline__ void set_consumer_sync() {
    asm volatile("bar.sync 1, %0;" ::"n"(NCONS) : "memory");
}

// Dependency barrier, split in two.  ARRIVE: a consumer warp that has finished the last op the waiting run conflicts
// with counts itself in shared memory (CTA-scope release: its y stores -- lane 0's, or the lanes' before the
// __syncwarp -- are ordered before the count); when all warps of the CTA have, the sentinel adds the CTA to the
// grid-wide counter with a GPU-scope release, which by cumulativity covers those stores.  No consumer thread ever
// executes a GPU-scope fence or atomic.  WAIT (thread 0 of the CTA, at the head of the run): relaxed polls until every
// consumer warp of the grid has arrived, then one acquire fence; the CTA barrier that follows the quantization of x
// hands the result to the other threads.  The arrival point is usually several ops before the wait (a layer's wq
// against the next layer's wq: four ops), so the poll finds the count complete and a barrier costs two L2 round
// trips of one thread that has nothing else to do (the sentinel) instead of a grid-wide rendezvous.
// Each consumer warp publishes ITS OWN monotone count of arrival points passed (one shared word per warp): warps of
// different groups are coupled only through the ring and may be several arrival points apart, so a single CTA-wide
// count would let two fast groups (2-4 arrival points ahead on lists with less than a tile per CTA and op) stand in
// for a slow one.  The sentinel forwards arrival e only when the MINIMUM over all W warps has reached it.
__device__ __forceinline__ void set_barrier_arrive(uint32_t arr_slot, unsigned count, int lane) {
    __syncwarp();
    if (lane == 0) asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(arr_slot), "r"(count) : "memory");
}
// WAIT, by the CTA's sentinel thread (a warp of its own, so the two L2 round trips -- the poll that finds the count
// complete and the acquire fence -- are off the consumers' path): it polls the barriers in list order ahead of the
// consumers and publishes the number it has seen complete in shared memory (kernel body, `warp == W + 1`).
__device__ __forceinline__ void st_release_cta_shared_u32(uint32_t addr, unsigned v) {
    asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_cta_shared_u32(uint32_t addr) {
    unsigned v;
    asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}

// kernel 1 inside the CTA (quantize_row_q8_0, AVX2 contract ggml.c:1192-1250; arithmetic of quantize_x_to_smem):
// 4 threads per block, 8 floats per thread, up to 3 items per thread with all their loads issued first.  x may have
// been written by other SMs during this launch: ld.global.cg.
// 8 consecutive elements of x out of tagged words: a word whose upper half is this launch's number IS the datum
// (8-byte accesses are single-copy atomic), so the reader needs neither a flag nor a fence -- one L2 round trip
// after the producer's store has landed.
__device__ __forceinline__ void set_load_tagged8(const unsigned long long * __restrict__ w, unsigned long long tag, float (&v)[8]) {
    unsigned long long t0 = 0;
    for (;;) {
        // all eight words per poll: the poll that finds them complete has already fetched them (no second round trip)
        unsigned long long a[8];
#pragma unroll
        for (int i = 0; i < 4; i++) ld_relaxed_gpu_v2b64(w + 2 * i, a[2 * i], a[2 * i + 1]);
        unsigned long long bad = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) { bad |= a[i] ^ tag; v[i] = __uint_as_float((unsigned)a[i]); }
        if ((bad >> 32) == 0) return;
        __nanosleep(20);
        const unsigned long long t1 = gtime();
        if (t0 == 0) t0 = t1;
        if (t1 - t0 > 10000000000ull) __trap();      // the producer rows never arrive: fail, do not hang
    }
}

template <int NCONS>
__device__ __forceinline__ void set_quantize_x(GemvSmem<1> & s, const float * __restrict__ x, const unsigned long long * __restrict__ xtag,
                                               unsigned long long tag, int nbk, int tid, bool ext) {
    const int total = nbk * 4;
    uint32_t * lo32 = reinterpret_cast<uint32_t *>(s.lo);
    uint32_t * hi32 = reinterpret_cast<uint32_t *>(s.hi);
    for (int base = 0; base < total; base += 3 * NCONS) {
        float4 a[3], c[3];
#pragma unroll
        for (int p = 0; p < 3; p++) {
            int item = base + p * NCONS + tid;
            if (item >= total) item = total - 1;
            a[p] = make_float4(0.f, 0.f, 0.f, 0.f); c[p] = a[p];
            if (xtag == nullptr) {
                if (ext) {       // immutable during the launch: L1 may hold it (prefetched there during the previous run)
                    a[p] = __ldca(reinterpret_cast<const float4 *>(x + 8 * item));
                    c[p] = __ldca(reinterpret_cast<const float4 *>(x + 8 * item) + 1);
                } else {
                    a[p] = __ldcg(reinterpret_cast<const float4 *>(x + 8 * item));
                    c[p] = __ldcg(reinterpret_cast<const float4 *>(x + 8 * item) + 1);
                }
            }
        }
#pragma unroll
        for (int p = 0; p < 3; p++) {
            const int item = base + p * NCONS + tid;
            if (base + p * NCONS >= total) break;                       // warp-uniform (NCONS and total are multiples of 4; see below)
            const bool active = item < total;
            float v[8] = {a[p].x, a[p].y, a[p].z, a[p].w, c[p].x, c[p].y, c[p].z, c[p].w};
            if (xtag != nullptr) set_load_tagged8(xtag + 8 * (active ? item : total - 1), tag, v);      // (one item at a time: the wait is here)
            float amax = 0.0f;
#pragma unroll
            for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v[i]));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
            const float d  = __fdiv_rn(amax, 127.0f);
            const float id = (amax != 0.0f) ? __fdiv_rn(127.0f, amax) : 0.0f;
            const __half dhalf = __float2half_rn(d);
            int qq[8];
            int sum = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) { qq[i] = q8_round(v[i], id); sum += qq[i]; }
            sum += __shfl_xor_sync(0xffffffffu, sum, 1);
            sum += __shfl_xor_sync(0xffffffffu, sum, 2);
            const uint32_t w0 = (uint32_t)(qq[0] & 0xFF) | ((uint32_t)(qq[1] & 0xFF) << 8) |
                                ((uint32_t)(qq[2] & 0xFF) << 16) | ((uint32_t)(qq[3] & 0xFF) << 24);
            const uint32_t w1 = (uint32_t)(qq[4] & 0xFF) | ((uint32_t)(qq[5] & 0xFF) << 8) |
                                ((uint32_t)(qq[6] & 0xFF) << 16) | ((uint32_t)(qq[7] & 0xFF) << 24);
            if (active) {
                const int b = item >> 2, part = item & 3;
                uint32_t * dst = (part & 2) ? hi32 : lo32;
                *reinterpret_cast<uint2 *>(dst + 4 * b + 2 * (part & 1)) = make_uint2(w0, w1);
                if (part == 0) s.ds[b] = make_int2(__float_as_int(__half2float(dhalf)), -8 * sum);
            }
        }
    }
}

// The same arithmetic with ONE thread per 32-element block (x warps: few threads, off the consumers' path): no
// shuffles, the two divisions once per block instead of once per quarter block.  Plain x only (not tagged words).
template <int NT>
__device__ __forceinline__ void set_quantize_x_blocks(GemvSmem<1> & s, const float * __restrict__ x, int nbk, int t, bool ext) {
    for (int b = t; b < nbk; b += NT) {
        const float4 * p = reinterpret_cast<const float4 *>(x + 32 * b);
        float4 f[8];
#pragma unroll
        for (int i = 0; i < 8; i++) f[i] = ext ? __ldca(p + i) : __ldcg(p + i);
        float amax = 0.0f;
#pragma unroll
        for (int i = 0; i < 8; i++) amax = fmaxf(fmaxf(fmaxf(amax, fabsf(f[i].x)), fmaxf(fabsf(f[i].y), fabsf(f[i].z))), fabsf(f[i].w));
        const float d  = __fdiv_rn(amax, 127.0f);
        const float id = (amax != 0.0f) ? __fdiv_rn(127.0f, amax) : 0.0f;
        uint32_t w[8];
        int sum = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int q0 = q8_round(f[i].x, id), q1 = q8_round(f[i].y, id), q2 = q8_round(f[i].z, id), q3 = q8_round(f[i].w, id);
            sum += (q0 + q1) + (q2 + q3);
            w[i] = (uint32_t)(q0 & 0xFF) | ((uint32_t)(q1 & 0xFF) << 8) | ((uint32_t)(q2 & 0xFF) << 16) | ((uint32_t)(q3 & 0xFF) << 24);
        }
        s.lo[b] = make_uint4(w[0], w[1], w[2], w[3]);
        s.hi[b] = make_uint4(w[4], w[5], w[6], w[7]);
        s.ds[b] = make_int2(__float_as_int(__half2float(__float2half_rn(d))), -8 * sum);
    }
}

// what a consumer needs of an op; loaded one op AHEAD (an L2 round trip under a saturated memory system costs
// ~0.7 us, and the ops of a list are walked serially)
struct SetOpC {
    const float * x;
    float * y;
    int * S;
    unsigned long long * ytag;
    const unsigned long long * xtag;
    int rows, nbk, ntiles, first_mod, flags;
};
// the same, 48 bytes, as staged in shared memory at kernel start when the list fits (SET_MAX_TABLE_BYTES): the
// consumers then walk the list without touching global memory at all
struct __align__(16) SetOpS {
    const float * x;
    float * y;
    unsigned long long * ytag;
    const unsigned long long * xtag;
    int rows, nbk, ntiles, fm_flags;   // first_mod | flags << 16
};
constexpr int SET_MAX_TABLE_BYTES = 32 * 1024;

template <bool DUMP>
__device__ __forceinline__ SetOpC set_load_op(const SetOp * __restrict__ ops, const SetOpS * tbl, int o) {
    SetOpC r;
    if (tbl != nullptr) {
        const SetOpS t = tbl[o];
        r.x = t.x; r.y = t.y; r.ytag = t.ytag; r.xtag = t.xtag; r.rows = t.rows; r.nbk = t.nbk; r.ntiles = t.ntiles;
        r.first_mod = t.fm_flags & 0xFFFF; r.flags = (int)((unsigned)t.fm_flags >> 16);
        r.S = DUMP ? ops[o].S : nullptr;
        return r;
    }
    const SetOp * p = ops + o;
    r.x = p->x; r.y = p->y; r.S = DUMP ? p->S : nullptr; r.ytag = p->ytag; r.xtag = p->xtag;
    r.rows = p->rows; r.nbk = p->nbk; r.ntiles = p->ntiles; r.first_mod = p->first_mod; r.flags = p->flags;
    return r;
}

struct SetCtx {
    const SetOp * ops;
    const SetOpS * tbl;
    int nops;
    unsigned char * ring;
    float * red;
    uint32_t full0, empty0;
    int nslots, slot_bytes;
    int G, group, wig, lane;
    unsigned long long tag;           // launch number << 32
    uint32_t xempty;                  // mbarrier: every consumer warp has the run's x in registers (the planes are free)
    uint32_t arr_slot;                // this warp's shared-memory word: arrival points it has passed
    const SetGather * gth;            // multi-GPU lists: the ranks' arenas (else nullptr)
    long long gpar;                   // word offset of this launch's parity inside an arena
};

// ring position of the CTA's next tile: slot index and the parity of its current use
struct SetRing {
    int s, phase;
    __device__ __forceinline__ void next(int nslots) { if (++s == nslots) { s = 0; phase ^= 1; } }
};

// One run of ops that share x, for one consumer warp.  WPR warps cooperate on a row; a lane keeps the quantized x of
// its NCL blocks in registers.  Returns the index of the op that starts the next run.
// `op` = descriptor of op o on entry, of the returned op on exit (when there is one).
template <int NG, int WPR, int NCL, bool DUMP, bool PROF, bool GATHER>
__device__ __forceinline__ int set_consume_run(const SetCtx & c, const GemvSmem<1> & xs, int o, SetOpC & op, int & j, SetRing & rp,
                                               unsigned & ntile_team, unsigned & narrived, long long & t_full, long long & t_comp, long long & n_tiles, long long * tx) {
    constexpr int RPT = WPR == 1 ? 2 : 4;               // rows of a tile per team
    constexpr int R = (SET_WPG / WPR) * RPT;            // 8 or 4
    constexpr int STRIDE = 32 * WPR;
    const int tl = WPR == 1 ? c.lane : c.wig * 32 + c.lane;
    const int nbk = op.nbk;
    uint4 xlo[NCL], xhi[NCL];
    float dx[NCL];
    int s8[NCL];
    unsigned vmask = 0;                                 // bit u: this lane's u-th block exists in the row
#pragma unroll
    for (int u = 0; u < NCL; u++) {
        const int b = tl + STRIDE * u;
        xlo[u] = make_uint4(0, 0, 0, 0); xhi[u] = make_uint4(0, 0, 0, 0); dx[u] = 0.0f; s8[u] = 0;
        if (b < nbk) {
            vmask |= 1u << u;
            xlo[u] = xs.lo[b]; xhi[u] = xs.hi[b];
            const int2 ds = xs.ds[b];
            dx[u] = __int_as_float(ds.x); s8[u] = ds.y;
        }
    }
    __syncwarp();
    if (c.lane == 0) mbar_arrive(c.xempty);           // the planes may now be refilled with the next run's x (x warps)
    for (;;) {
        int first = (int)blockIdx.x - op.first_mod;
        if (first < 0) first += c.G;
        long long tl0 = 0;
        if (PROF) tl0 = clock64();
        for (int t = first; t < op.ntiles; t += c.G, j++, rp.next(c.nslots)) {
            if (PROF) { const long long now = clock64(); tx[0] += now - tl0; tl0 = now; tx[1]++; }      // since the previous iteration's last stamp
            if (j % NG != c.group) continue;
            const int rit = WPR == 1 ? c.wig * RPT : 0;           // first row of this team inside the tile
            const int row0 = t * R + rit;
            const int nvalid = min(RPT, op.rows - row0);
            const unsigned char * slot = c.ring + (size_t)rp.s * c.slot_bytes;
            const unsigned char * nib = slot + (size_t)rit * nbk * 16;
            const unsigned char * sc = slot + (size_t)R * nbk * 16 + (size_t)rit * nbk * 2;
            const uint32_t full = c.full0 + 8 * rp.s, empty = c.empty0 + 8 * rp.s;
            long long tf0 = 0;
            if (PROF) tf0 = clock64();
            // A parity wait can only tell the current phase of a barrier from the one before it.  The slot's previous use
            // may belong to ANOTHER group and its copy may still be in flight although later copies have landed (bulk
            // copies complete in any order): `full` would then be one phase behind what this warp assumes and the wait
            // would fall through on the phase before that.  The previous use having been RELEASED (its `empty` phase;
            // that wait cannot alias: this tile's predecessors were issued after the use before that was released)
            // pins `full` to the phase waited for.  (The two tests must not be overlapped: issued back to back they failed.)
            // Needs nslots >= NG, which the host guarantees: this group's previous tile j - NG is then not older than the
            // slot's previous use j - nslots, so that use has been issued, i.e. the one before it released.
            if (j >= c.nslots) mbar_wait(empty, rp.phase ^ 1);
            mbar_wait(full, rp.phase);
            long long tc0 = 0;
            if (PROF) { tc0 = clock64(); t_full += tc0 - tf0; n_tiles++; }
            if (nvalid <= 0) {            // ragged last tile (WPR == 1 only): nothing for this warp, still release the slot
                if (c.lane == 0) mbar_arrive(empty);
                continue;
            }
            float acc[RPT];
#pragma unroll
            for (int r = 0; r < RPT; r++) acc[r] = 0.0f;
            // Straight-line over the lane's NCL blocks: a lane whose block index is past the row re-reads the row's last
            // block (valid shared memory, finite or not) and only its final accumulate is predicated off, so there
            // is no branch between the 2*NCL*RPT loads and the dots and the scheduler can overlap all of them.
#pragma unroll
            for (int u = 0; u < NCL; u++) {
                const int b = tl + STRIDE * u;
                const int ba = min(b, nbk - 1);
                uint4 q[RPT];
                unsigned short dh[RPT];
#pragma unroll
                for (int r = 0; r < RPT; r++) {
                    q[r] = *reinterpret_cast<const uint4 *>(nib + ((size_t)r * nbk + ba) * 16);
                    dh[r] = *reinterpret_cast<const unsigned short *>(sc + ((size_t)r * nbk + ba) * 2);
                }
#pragma unroll
                for (int r = 0; r < RPT; r++) {
                    const int sb = block_dot_q4_q8_fast(q[r], xlo[u], xhi[u], s8[u]);
                    if (DUMP) { if (op.S != nullptr && r < nvalid && b < nbk) op.S[(size_t)(row0 + r) * nbk + b] = sb; }
                    const float t = fmaf((float)sb, half_bits_to_float(dh[r]) * dx[u], acc[r]);
                    if ((vmask >> u) & 1u) acc[r] = t;
                }
            }
            // the slot's bytes of this warp are in registers: hand the slot back before the reduction
            __syncwarp();
            if (c.lane == 0) mbar_arrive(empty);
#pragma unroll
            for (int r = 0; r < RPT; r++) acc[r] = warp_sum(acc[r]);
            if (WPR == 1) {
                if (GATHER) {
                    // every lane holds the sums (xor-shuffle tree): lane p < world sends the rows to rank p's arena
                    if (c.lane < c.gth->world) {
                        unsigned long long * dst = c.gth->arena[c.lane] + (op.ytag - c.gth->arena[c.gth->me]) + c.gpar + row0;
#pragma unroll
                        for (int r = 0; r < RPT; r++) if (r < nvalid) st_sys_b64(dst + r, c.tag | __float_as_uint(acc[r]));
                    }
                    if (c.lane == 0) {
#pragma unroll
                        for (int r = 0; r < RPT; r++) if (r < nvalid) op.y[row0 + r] = acc[r];
                    }
                } else if (c.lane == 0) {
#pragma unroll
                    for (int r = 0; r < RPT; r++) if (r < nvalid) {
                        op.y[row0 + r] = acc[r];
                        if (op.ytag != nullptr) st_relaxed_gpu_b64(op.ytag + row0 + r, c.tag | __float_as_uint(acc[r]));
                    }
                }
            } else {
                // four warp partials per row, added in a fixed order by the group's first warp
                float * red = c.red + ((ntile_team & 1u) * SET_MAX_GROUPS + c.group) * 16;
                if (c.lane == 0) {
#pragma unroll
                    for (int r = 0; r < RPT; r++) red[r * 4 + c.wig] = acc[r];
                }
                asm volatile("bar.sync %0, 128;" ::"r"(2 + c.group) : "memory");
                if (GATHER) {
                    // lane l of the group's first warp: row l % 4 to rank l / 4
                    if (c.wig == 0 && (c.lane & 3) < nvalid && (c.lane >> 2) < c.gth->world) {
                        const int r = c.lane & 3, pr = c.lane >> 2;
                        const float4 p = *reinterpret_cast<const float4 *>(red + r * 4);
                        const float v = ((p.x + p.y) + p.z) + p.w;
                        st_sys_b64(c.gth->arena[pr] + (op.ytag - c.gth->arena[c.gth->me]) + c.gpar + row0 + r, c.tag | __float_as_uint(v));
                        if (pr == 0) op.y[row0 + r] = v;
                    }
                } else if (c.wig == 0 && c.lane < nvalid) {
                    const float4 p = *reinterpret_cast<const float4 *>(red + c.lane * 4);
                    const float v = ((p.x + p.y) + p.z) + p.w;
                    op.y[row0 + c.lane] = v;
                    if (op.ytag != nullptr) st_relaxed_gpu_b64(op.ytag + row0 + c.lane, c.tag | __float_as_uint(v));
                }
                ntile_team++;
            }
            if (PROF) { tl0 = clock64(); t_comp += tl0 - tc0; }
        }
        if (PROF) { const long long now = clock64(); tx[2] += now - tl0; tl0 = now; }
        {
            const int narr = (op.flags >> SET_OP_ARRIVE_SHIFT) & SET_OP_ARRIVE_MASK;
            if (narr) set_barrier_arrive(c.arr_slot, ++narrived, c.lane);      // a later run waits for this op (and everything before it)
        }
        o++;
        if (o >= c.nops) break;
        op = set_load_op<DUMP>(c.ops, c.tbl, o);      // shared-memory table (lists beyond SET_MAX_TABLE_BYTES: an L2 round trip per op)
        if (PROF) tx[3] += clock64() - tl0;
        if (op.flags & SET_OP_QUANT) break;
    }
    return o;
}

// PROF (profiling aid, ggml_b200_set_profile): clock64 totals per CTA -> prof[cta*8 + ..] =
// {whole kernel, consumer warp 0 waiting for full slots, in dependency barriers, quantizing x, producer waiting for empty slots}
// BIG: also the 6-blocks-per-lane plan (K > 16384); the kernel without it needs fewer registers
template <int W, bool DUMP, bool PROF, bool BIG, bool GATHER>
__global__ void __launch_bounds__(set_threads<W, GATHER>(), 1)
gemv_q4_0_set_kernel(const SetOp * __restrict__ ops, int nops, SetSync * __restrict__ sync, int nslots, int slot_bytes,
                     int x_bytes, int x_bufs, int tbl_bytes, int max_fly, unsigned seq, long long * __restrict__ prof,
                     const __grid_constant__ SetGather gth) {
    constexpr int NCONS = W * 32;
    constexpr int NG = W / SET_WPG;
    static_assert(NG <= SET_MAX_GROUPS, "named barriers 2 .. 2+NG-1 and the partials buffer are sized for 6 groups");
    extern __shared__ __align__(128) unsigned char smem[];
    // [x planes: x_bytes][full barriers 16 x 8][empty barriers 16 x 8][team partials][op table: tbl_bytes][ring: nslots x slot_bytes]
    // x_bufs = 2: the x warps may be a whole run ahead (they fill one buffer while the consumers still copy the other)
    const int x_per = x_bytes / x_bufs;
    GemvSmem<1> xs0(smem, x_per / 40), xs1(smem + (x_bufs > 1 ? x_per : 0), x_per / 40);
    uint64_t * bars = reinterpret_cast<uint64_t *>(smem + x_bytes);
    SetOpS * tbl = tbl_bytes ? reinterpret_cast<SetOpS *>(smem + x_bytes + SET_BAR_BYTES + SET_RED_BYTES) : nullptr;
    unsigned char * ring = smem + x_bytes + SET_BAR_BYTES + SET_RED_BYTES + tbl_bytes;
    const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + SET_MAX_SLOTS);
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const int G = gridDim.x;

    // one arrival counter per barrier behind the sync words: a CTA may be several arrival points ahead of another, so a
    // single running count would not say WHO has arrived
    unsigned * bar_cnt = reinterpret_cast<unsigned *>(sync + 1);
    __shared__ unsigned s_launch;
    __shared__ unsigned s_bar_passed;                         // dependency barriers the sentinel has seen complete
    __shared__ unsigned s_arrived[W];                         // per consumer warp: arrival points passed
    __shared__ unsigned s_gathered;                           // multi-GPU lists: ops whose peer rows this CTA's gather warp has copied into dst
    __shared__ __align__(8) uint64_t s_xbar[4];      // per x buffer: full [0..1], empty [2..3]
    const uint32_t xfull0 = smem_u32(&s_xbar[0]), xempty0 = smem_u32(&s_xbar[2]);
    const uint32_t bar_passed = smem_u32(&s_bar_passed), arrived0 = smem_u32(&s_arrived[0]);
    if (tid < W) s_arrived[tid] = 0u;
    if (tid == 0) {
        s_bar_passed = 0u;
        s_gathered = 0u;
        for (int b = 0; b < 2; b++) {
            mbar_init(xfull0 + 8 * b, 1);       // the x warps filled the planes with a run's quantized x (or handed them over: tagged x)
            mbar_init(xempty0 + 8 * b, W);      // every consumer warp copied the run's x to registers
        }
        for (int s = 0; s < nslots; s++) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, SET_WPG);
        }
        mbar_fence_init();
        s_launch = ld_relaxed_gpu_u32(&sync->epoch) + 1u;     // stable: the last CTA of the previous launch wrote it
    }
    if (tbl != nullptr) {
        for (int i = tid; i < nops; i += blockDim.x) {
            const SetOp * p = ops + i;
            SetOpS t;
            t.x = p->x; t.y = p->y; t.ytag = p->ytag; t.xtag = p->xtag; t.rows = p->rows; t.nbk = p->nbk; t.ntiles = p->ntiles; t.fm_flags = p->first_mod | (p->flags << 16);
            tbl[i] = t;
        }
    }
    __syncthreads();
    CtaTrace tr;
    if (tid == 0) tr.begin(seq);
    long long t_begin = 0, t_full = 0, t_bar = 0, t_quant = 0, t_empty = 0, t_comp = 0, n_tiles = 0, t_run = 0;
    if (PROF) t_begin = clock64();

    if (warp == W) {
        // ===== producer: streams this CTA's tiles of every op, in list order, through the ring =====
        if (lane == 0) {
            SetRing rp = {0, 0};
            // At most max_fly copies in flight: the TMA engine's requests and the CTA's ordinary loads (dependency
            // polls, x) leave the SM through the same queue, and a load issued behind 8 x 18 KB of bulk requests
            // waits ~2400 clk for its turn (measured: profiles/r01_experiments.md).  `lag` trails the issue position by
            // max_fly slots; the copy issued there must have landed before the next one goes out.
            SetRing lag = {0, 0};
            int issued = 0;
            bool wrapped = false;
            SetOp op = ops[0];
            for (int o = 0; o < nops; o++) {
                SetOp nxt = op;
                if (o + 1 < nops) nxt = ops[o + 1];
                const int R = op.tile_rows;
                const unsigned char * nib_base = reinterpret_cast<const unsigned char *>(op.wq);
                const unsigned char * sc_base = reinterpret_cast<const unsigned char *>(op.wd);
                int first = (int)blockIdx.x - op.first_mod;
                if (first < 0) first += G;
                for (int t = first; t < op.ntiles; t += G) {
                    if (wrapped) {
                        long long t0 = 0;
                        if (PROF) t0 = clock64();
                        mbar_wait_backoff(empty0 + 8 * rp.s, rp.phase ^ 1);      // the consumers released the slot's previous use
                        if (PROF) t_empty += clock64() - t0;
                    }
                    if (issued >= max_fly) {
                        mbar_wait_backoff(full0 + 8 * lag.s, lag.phase);
                        lag.next(nslots);
                    }
                    issued++;
                    const int row0 = t * R;
                    const int nr = min(R, op.rows - row0);
                    const uint32_t nib_sz = (uint32_t)nr * op.nbk * 16;
                    const uint32_t sc_sz = ((uint32_t)nr * op.nbk * 2 + 15u) & ~15u;   // the scale plane is allocated with 16 B of slack
                    const uint32_t dst = smem_u32(ring + (size_t)rp.s * slot_bytes);
                    const uint32_t full = full0 + 8 * rp.s;
                    mbar_expect_tx(full, nib_sz + sc_sz);
                    bulk_g2s(dst, nib_base + (size_t)row0 * op.nbk * 16, nib_sz, full);
                    bulk_g2s(dst + (uint32_t)R * op.nbk * 16, sc_base + (size_t)row0 * op.nbk * 2, sc_sz, full);
                    if (++rp.s == nslots) { rp.s = 0; rp.phase ^= 1; wrapped = true; }
                }
                op = nxt;
            }
            if (PROF) prof[blockIdx.x * 16 + 4] = t_empty;
        }
        return;
    }

    if (warp == W + 1) {
        // ===== sentinel (one thread, never blocks on one duty while the other is due) =====
        //  arrive: when every consumer warp of this CTA is past an op that barriers wait for, the CTA arrives at them;
        //  wait:   polls the list's barriers, in order, as far ahead of the consumers as the arrivals allow, and
        //          publishes how many it has seen complete.
        if (lane == 0) {
            int oa = 0, ow = 0;                 // list positions of the next arrival point / the next barrier head
            unsigned ev = 0, kb = 0, k = 0;
            unsigned long long t0 = 0;
            auto flags_of = [&](int o) { return tbl != nullptr ? (int)((unsigned)tbl[o].fm_flags >> 16) : ops[o].flags; };
            while (oa < nops && !((flags_of(oa) >> SET_OP_ARRIVE_SHIFT) & SET_OP_ARRIVE_MASK)) oa++;
            while (ow < nops && !(flags_of(ow) & SET_OP_BARRIER)) ow++;
            while (oa < nops || ow < nops) {
                bool progress = false;
                bool all_arrived = oa < nops;
                for (int w = 0; all_arrived && w < W; w++) all_arrived = (int)(ld_acquire_cta_shared_u32(arrived0 + 4 * w) - (ev + 1)) >= 0;
                if (GATHER && all_arrived && (flags_of(oa) & SET_OP_ARRIVE_GATHER))
                    all_arrived = (int)(ld_acquire_cta_shared_u32(smem_u32(&s_gathered)) - (unsigned)(oa + 1)) >= 0;
                if (all_arrived) {
                    const int narr = (flags_of(oa) >> SET_OP_ARRIVE_SHIFT) & SET_OP_ARRIVE_MASK;
                    for (int i = 0; i < narr; i++) red_release_gpu_add_u32(bar_cnt + kb++, 1u);      // barriers are numbered in the order of their arrival points
                    ev++;
                    oa++;
                    while (oa < nops && !((flags_of(oa) >> SET_OP_ARRIVE_SHIFT) & SET_OP_ARRIVE_MASK)) oa++;
                    progress = true;
                }
                if (ow < nops && ld_relaxed_gpu_u32(bar_cnt + k) >= (unsigned)G) {
                    fence_acq_rel_gpu();
                    k++;
                    st_release_cta_shared_u32(bar_passed, k);
                    ow++;
                    while (ow < nops && !(flags_of(ow) & SET_OP_BARRIER)) ow++;
                    progress = true;
                }
                if (progress) { t0 = 0; continue; }
                __nanosleep(64);
                const unsigned long long t1 = gtime();
                if (t0 == 0) t0 = t1;
                if (t1 - t0 > 10000000000ull) __trap();   // a CTA of the grid is not resident / died: fail, do not hang
            }
        }
        return;
    }
    constexpr int XW = set_xwarps<GATHER>();
    if (warp > W + 1 && warp < W + 2 + XW) {
        // ===== x warps: kernel 1 (quantize_row_q8_0) of every run, one run AHEAD of the consumers =====
        // The consumers keep a run's x in registers, so the planes are free as soon as every warp has copied them: the
        // next run's x is quantized into them during this run's tiles, and a change of x costs a consumer warp one
        // shared-memory wait and the register load -- no CTA-wide rendezvous, no L2 round trip, no quantization.  Runs
        // whose x is another op's result (tagged words) are left to the consumers: all of them poll and quantize, which
        // is the shorter critical path; the x warps only hand the planes over.
        constexpr int NX = XW * 32;
        const int xt = tid - (W + 2) * 32;
        unsigned r = 0, kx = 0;
        for (int o = 0; o < nops; o++) {
            const SetOpC op = set_load_op<false>(ops, tbl, o);
            if (!(op.flags & SET_OP_QUANT)) continue;
            const unsigned b = x_bufs > 1 ? (r & 1u) : 0u, u = x_bufs > 1 ? (r >> 1) : r;      // buffer, and its how-manieth use
            if (u > 0) mbar_wait_backoff(xempty0 + 8 * b, (u - 1) & 1u);
            if (op.flags & SET_OP_BARRIER) kx++;
            // x that an op of the list writes may only be read behind every barrier up to this run: the hazard analysis
            // drops an op from the live sets once a barrier covers it, so a LATER run that reads its dst carries no flag
            // of its own -- the consumers are behind that barrier by program order, the x warps (up to two runs ahead,
            // and not held by barrier heads whose own x is external) must check the count themselves
            if (!(op.flags & SET_OP_XEXT) && op.xtag == nullptr) while ((int)(ld_acquire_cta_shared_u32(bar_passed) - kx) < 0) __nanosleep(64);
            if (op.xtag == nullptr) set_quantize_x_blocks<NX>(b ? xs1 : xs0, op.x, op.nbk, xt, (op.flags & SET_OP_XEXT) != 0);
            asm volatile("bar.sync 15, %0;" ::"n"(NX) : "memory");
            if (xt == 0) mbar_arrive(xfull0 + 8 * b);
            r++;
        }
        return;
    }

    if (GATHER && warp == W + 2 + XW) {
        // ===== gather warp (multi-GPU lists): the peers' rows of every op, out of this rank's arena into dst =====
        // In list order; this CTA takes a fixed slice of every op's rows, so the same thread handles the same address
        // of a dst buffer that several ops reuse (program order = list order).  Dst buffers that an EARLIER op still
        // reads as x or writes are protected by the list's barriers: the warp waits for every barrier up to the op.
        const unsigned launch = s_launch;
        const unsigned long long * mine = gth.arena[gth.me] + (long long)(launch & 1u) * gth.stride;
        unsigned kg = 0;
        unsigned long long t0 = 0;
        for (int o = 0; o < nops; o++) {
            const SetOp * p = ops + o;
            const int flags = p->flags;
            if (flags & SET_OP_BARRIER) {
                kg++;
                while ((int)(ld_acquire_cta_shared_u32(bar_passed) - kg) < 0) __nanosleep(64);
            }
            const int M = p->gM, lo_me = p->grow0, hi_me = p->grow0 + p->rows;
            const unsigned long long * src = mine + (p->ytag - gth.arena[gth.me]) - lo_me;      // row 0 of the whole tensor
            float * yfull = p->y - lo_me;
            const int per_cta = (M + G - 1) / G;
            const int m0 = min((int)blockIdx.x * per_cta, M), m1 = min(m0 + per_cta, M);
            for (int base = m0; base < m1; base += 128) {
                unsigned long long a[4];
                bool need[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int m = base + 32 * u + lane;
                    need[u] = m < m1 && (m < lo_me || m >= hi_me);
                    a[u] = 0;
                }
                for (;;) {
                    bool again = false;
#pragma unroll
                    for (int u = 0; u < 4; u++) if (need[u] && (unsigned)(a[u] >> 32) != launch) a[u] = ld_sys_b64(src + base + 32 * u + lane);
#pragma unroll
                    for (int u = 0; u < 4; u++) again |= need[u] && (unsigned)(a[u] >> 32) != launch;
                    if (!again) break;
                    __nanosleep(200);
                    const unsigned long long t1 = gtime();
                    if (t0 == 0) t0 = t1;
                    if (t1 - t0 > P2P_TIMEOUT_NS) __trap();      // a peer never delivered: fail, do not hang
                }
                t0 = 0;
#pragma unroll
                for (int u = 0; u < 4; u++) if (need[u]) yfull[base + 32 * u + lane] = __uint_as_float((unsigned)a[u]);
            }
            __syncwarp();
            if (lane == 0) st_release_cta_shared_u32(smem_u32(&s_gathered), (unsigned)(o + 1));
        }
        return;
    }

    // ===== consumers =====
    SetCtx c;
    c.ops = ops; c.tbl = tbl; c.nops = nops; c.ring = ring; c.red = reinterpret_cast<float *>(smem + x_bytes + SET_BAR_BYTES);
    c.full0 = full0; c.empty0 = empty0; c.nslots = nslots; c.slot_bytes = slot_bytes;
    c.G = G; c.group = warp / SET_WPG; c.wig = warp % SET_WPG; c.lane = lane;
    c.tag = (unsigned long long)s_launch << 32;
    c.arr_slot = arrived0 + 4 * warp;
    c.gth = GATHER ? &gth : nullptr;
    c.gpar = GATHER ? (long long)(s_launch & 1u) * gth.stride : 0;
    c.xempty = xempty0;
    SetRing rp = {0, 0};
    int j = 0;
    unsigned nbar = 0, ntile_team = 0, nrun = 0, narrived = 0;
    SetOpC op = set_load_op<DUMP>(ops, tbl, 0);
    long long tp[4] = {0, 0, 0, 0};
    long long tx[4] = {0, 0, 0, 0};
    for (int o = 0; o < nops;) {
        // Every run starts with an op that (re)quantizes x.  Where the host found a hazard it cannot express as
        // "x is the dst of ONE earlier op" the run waits for everything before it (grid barrier); otherwise x is
        // either untouched by the list (plain loads) or read out of its producer's tagged words, which is the wait.
        const int flags = op.flags;
        const int nbk = op.nbk;
        const float * x = op.x;
        long long tq0 = 0;
        if (PROF) tq0 = clock64();
        if (flags & SET_OP_BARRIER) {
            // every thread sees for itself that the sentinel has passed this barrier (normally long ago: the arrival
            // point lies ops back); this run's y stores (and reads of x that the list writes) come after it
            nbar++;
            while ((int)(ld_acquire_cta_shared_u32(bar_passed) - nbar) < 0) { }
        }
        if (PROF) { const long long t1 = clock64(); t_bar += t1 - tq0; tq0 = t1; }
        // the x warps have the run's quantized x in the planes (they started on it when the previous run's x went to
        // registers) -- or, for x that is another op's result, have handed the free planes over
        const unsigned xb = x_bufs > 1 ? (nrun & 1u) : 0u, xu = x_bufs > 1 ? (nrun >> 1) : nrun;
        mbar_wait(xfull0 + 8 * xb, xu & 1u);
        nrun++;
        GemvSmem<1> & xs = xb ? xs1 : xs0;
        c.xempty = xempty0 + 8 * xb;
        if (op.xtag != nullptr) {
            set_quantize_x<NCONS>(xs, x, op.xtag + c.gpar, c.tag, nbk, tid, false);
            set_consumer_sync<NCONS>();
        }
        long long tr0 = 0;
        if (PROF) { tr0 = clock64(); t_quant += tr0 - tq0; }
        if (nbk <= SET_MAX_NBK_WPR1) {
            if (nbk <= 32)       o = set_consume_run<NG, 1, 1, DUMP, PROF, GATHER>(c, xs, o, op, j, rp, ntile_team, narrived, t_full, t_comp, n_tiles, tx);
            else if (nbk <= 64)  o = set_consume_run<NG, 1, 2, DUMP, PROF, GATHER>(c, xs, o, op, j, rp, ntile_team, narrived, t_full, t_comp, n_tiles, tx);
            else if (nbk <= 96)  o = set_consume_run<NG, 1, 3, DUMP, PROF, GATHER>(c, xs, o, op, j, rp, ntile_team, narrived, t_full, t_comp, n_tiles, tx);
            else                 o = set_consume_run<NG, 1, 4, DUMP, PROF, GATHER>(c, xs, o, op, j, rp, ntile_team, narrived, t_full, t_comp, n_tiles, tx);
        } else {
            if (nbk <= 256)      o = set_consume_run<NG, 4, 2, DUMP, PROF, GATHER>(c, xs, o, op, j, rp, ntile_team, narrived, t_full, t_comp, n_tiles, tx);
            else if (nbk <= 384) o = set_consume_run<NG, 4, 3, DUMP, PROF, GATHER>(c, xs, o, op, j, rp, ntile_team, narrived, t_full, t_comp, n_tiles, tx);
            else if (nbk <= 512 || !BIG) o = set_consume_run<NG, 4, 4, DUMP, PROF, GATHER>(c, xs, o, op, j, rp, ntile_team, narrived, t_full, t_comp, n_tiles, tx);
            else { if constexpr (BIG) o = set_consume_run<NG, 4, 6, DUMP, PROF, GATHER>(c, xs, o, op, j, rp, ntile_team, narrived, t_full, t_comp, n_tiles, tx); }
        }
        if (PROF) t_run += clock64() - tr0;
    }
    if (tid == 0) tr.mark(3);
    if (PROF) {
        if (tid == 0) {
            prof[blockIdx.x * 16 + 0] = clock64() - t_begin;
            prof[blockIdx.x * 16 + 1] = t_full;
            prof[blockIdx.x * 16 + 2] = t_bar;
            prof[blockIdx.x * 16 + 3] = t_quant;
            prof[blockIdx.x * 16 + 5] = t_comp;
            prof[blockIdx.x * 16 + 6] = n_tiles;
            prof[blockIdx.x * 16 + 7] = t_run;
            for (int i = 0; i < 4; i++) prof[blockIdx.x * 16 + 8 + i] = tp[i];
            for (int i = 0; i < 4; i++) prof[blockIdx.x * 16 + 12 + i] = tx[i];
        }
    }

    // leave the sync words zero for the next launch: the last CTA to finish has seen every barrier complete
    set_consumer_sync<NCONS>();
    if (tid == 0) {
        __threadfence();
        const unsigned old = atomicAdd(&sync->done, 1u);
        if (old == (unsigned)G - 1u) {
            for (unsigned i = 0; i < nbar; i++) atomicExch(bar_cnt + i, 0u);      // (every arrival has been observed: all barriers passed everywhere)
            atomicExch(&sync->done, 0u);
            atomicExch(&sync->epoch, s_launch);
        }
    }
}

}  // namespace b200
