// p2p.cuh -- n = 1 multi-GPU exchange over NVLink peer memory, shared by comm.cu (arena setup, stand-alone
// gather kernel) and gemv_q4_0_p2p.cuh (GEMV with the gather fused in).
//
// An NCCL all-gather of a few KB costs ~12 us per matmul, several times the GEMV it follows.  Instead every rank
// owns an "arena" in its HBM, exported with cudaIpcGetMemHandle and mapped by every peer at comm_init:
//     [P2P_REGIONS][P2P_MAX_M] 8-byte words (bits of y[m], op number); op s uses region s % P2P_REGIONS
//     P2P_CTR_SLOTS op counters             one per CTA index, touched only by the owning rank
// A rank stores its rows into every peer's arena with each word carrying the op number next to the value: 8-byte
// words are single-copy atomic, so a word whose number matches IS the datum -- no fence, no separate flag, no
// round trip, one NVLink one-way latency per op.  It then polls its own arena for the peers' rows.
//   * A peer can run ahead by at most one run of independent ops (<= P2P_MAX_RUN; the op after a run needs this
//     rank's rows of every op of the run), so P2P_REGIONS > P2P_MAX_RUN regions never collide.
//   * Op numbers live on the device, so a captured CUDA graph replays correctly, and every rank numbers its ops
//     identically (same op sequence).  Only DEPENDENT ops (heads of runs) write the counters: CTA b reads ctr[b]
//     BEFORE its dependency wait (the value is stable: the previous head wrote it before it released anything),
//     and after the wait one thread stores number = old + P2P_MAX_RUN to ctr[b, b + grid, ...], fences, and only
//     then releases the dependents (griddepcontrol.launch_dependents counts per CTA as soon as ONE thread issues
//     it: tools/ubench/pdl_trigger.cu).  An op that runs independently of its predecessor (same src1: wq/wk/wv)
//     does not wait; the host tells it its position j in the run and it takes ctr[b] + j, which the head stored
//     before this launch could be scheduled.  Heads always advance by P2P_MAX_RUN, whatever the run's length, so
//     numbers never repeat even when a captured graph is replayed after a different history.
//   * Ranks are separate GPUs driven by separate processes: a poll cannot starve the peer it waits for.
#pragma once
#include <stdint.h>

namespace b200 {

constexpr int P2P_MAX_WORLD = 8;
constexpr int P2P_MAX_M = 65536;
constexpr int P2P_REGIONS = 8;
constexpr int P2P_MAX_RUN = 4;                                 // ops of one same-src1 run that may be in flight together
constexpr int P2P_CTR_SLOTS = 256;                             // >= any grid that takes part
constexpr size_t P2P_CTR_OFF = (size_t)P2P_REGIONS * P2P_MAX_M * 8;
constexpr size_t P2P_ARENA_BYTES = P2P_CTR_OFF + P2P_CTR_SLOTS * 4;
constexpr unsigned long long P2P_TIMEOUT_NS = 60000000000ull;  // a peer that never arrives: trap, do not hang

struct P2PPeers { unsigned char * arena[P2P_MAX_WORLD]; };

// what a kernel needs to trade rows: y = the whole dst column (this rank's rows already in place or about to be)
struct P2PExchange {
    P2PPeers peers;
    int me, world;
    int per;      // rows of a full shard (ggml_b200_shard_rows): rank r owns [r*per, min(M, (r+1)*per))
    int M;
};

__device__ __forceinline__ void st_sys_b64(unsigned long long * p, unsigned long long a) {
    asm volatile("st.relaxed.sys.global.b64 [%0], %1;" ::"l"(p), "l"(a) : "memory");
}
__device__ __forceinline__ unsigned long long ld_sys_b64(const unsigned long long * p) {
    unsigned long long a;
    asm volatile("ld.relaxed.sys.global.b64 %0, [%1];" : "=l"(a) : "l"(p) : "memory");
    return a;
}
__device__ __forceinline__ unsigned long long p2p_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }

__device__ __forceinline__ unsigned * p2p_ctr(const P2PExchange & e) { return reinterpret_cast<unsigned *>(e.peers.arena[e.me] + P2P_CTR_OFF); }

// the counter as this CTA sees it: a head may read it BEFORE its dependency wait, a follower any time
__device__ __forceinline__ unsigned p2p_counter(const P2PExchange & e) {
    return reinterpret_cast<volatile unsigned *>(p2p_ctr(e))[blockIdx.x % P2P_CTR_SLOTS];
}
// head, ONE thread, after the dependency wait: store the run's number, make it visible, release the dependents.
// No other thread of the CTA may issue griddepcontrol.launch_dependents.
__device__ __forceinline__ void p2p_head_publish_and_release(const P2PExchange & e, unsigned seq) {
    volatile unsigned * ctr = reinterpret_cast<volatile unsigned *>(p2p_ctr(e));
    for (unsigned slot = blockIdx.x; slot < (unsigned)P2P_CTR_SLOTS; slot += gridDim.x) ctr[slot] = seq;
    __threadfence();
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// rows [lo, lo+n) of rank r
__device__ __forceinline__ void p2p_rows_of(const P2PExchange & e, int r, int & lo, int & n) {
    lo = min(r * e.per, e.M);
    n = min(lo + e.per, e.M) - lo;
}

// this CTA's slice [j0, j1) of an n-row range split evenly over the grid
__device__ __forceinline__ void p2p_cta_slice(int n, int & j0, int & j1) {
    const int r = (n + gridDim.x - 1) / gridDim.x;
    j0 = min((int)blockIdx.x * r, n);
    j1 = min(j0 + r, n);
}

// send vals[0..cnt) = y[lo_me + j0 ...] to every peer (all threads of the CTA; consecutive threads -> consecutive words)
__device__ __forceinline__ void p2p_send(const P2PExchange & e, unsigned seq, const float * vals, int first_row, int cnt) {
    const unsigned long long tag = (unsigned long long)seq << 32;
    const size_t half = (size_t)(seq % P2P_REGIONS) * P2P_MAX_M;
    const int total = (e.world - 1) * cnt;
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
        const int pi = i / cnt, j = i - pi * cnt;
        const int p = pi < e.me ? pi : pi + 1;
        st_sys_b64(reinterpret_cast<unsigned long long *>(e.peers.arena[p]) + half + first_row + j, tag | __float_as_uint(vals[j]));
    }
}

// receive this CTA's slice of every peer's rows into y as they land
__device__ __forceinline__ void p2p_receive(const P2PExchange & e, unsigned seq, float * __restrict__ y) {
    const size_t half = (size_t)(seq % P2P_REGIONS) * P2P_MAX_M;
    const unsigned long long * in = reinterpret_cast<const unsigned long long *>(e.peers.arena[e.me]) + half;
    const int rs = (e.per + gridDim.x - 1) / gridDim.x;        // slice length of a full shard
    const int total = (e.world - 1) * rs;
    const unsigned long long t0 = p2p_now();
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
        const int pi = i / rs, j = i - pi * rs;
        const int p = pi < e.me ? pi : pi + 1;
        int lo, n, j0, j1;
        p2p_rows_of(e, p, lo, n);
        p2p_cta_slice(n, j0, j1);
        if (j0 + j < j1) {
            const int m = lo + j0 + j;
            unsigned long long a;
            while (((a = ld_sys_b64(in + m)) >> 32) != seq)
                if (p2p_now() - t0 > P2P_TIMEOUT_NS) __trap();
            y[m] = __uint_as_float((unsigned)a);
        }
    }
}

}  // namespace b200
