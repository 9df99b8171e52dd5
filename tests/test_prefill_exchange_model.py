"""A model of the multi-GPU prefill exchange (comm.cu `ggml_b200_mul_mat_q4_0_sharded`, n >= gemm_min_n; kernel 3's peer
stores in gemm_q4_0_tc3.cuh; `pf_collect_kernel`) under an adversarial scheduler: ranks run at any relative speed, a rank's
tiles reach its peers in any order and at any later time.

Protocol: every rank owns TWO landing regions and PF_SLOTS arrival counters.  Op s (same number on every rank) uses region
s % 2 and counter slot s % PF_SLOTS; a rank's GEMM stores every finished tile into every peer's region and adds 1 to the
peer's counter slot; the rank's collect -- stream-ordered behind its own GEMM -- waits until its counter slot has reached
the slot's running target (sum of the peers' tile counts of every op that used the slot so far) and copies the region out;
op s + 1 starts behind it.  Pinned properties:
  * a collect never copies a region that lacks a tile of its op or holds a tile of another op;
  * no rank can be more than one op ahead of a peer's collect, so two regions are enough (with ONE region the model finds
    the overwrite at once);
  * counters never have to be reset: the monotonic target tells op s from op s + PF_SLOTS."""
import random

import pytest

PF_SLOTS = 4          # (64 in comm.cu; small here so that slots are reused within a run)


def shard_rows(M, r, world):
    per = -(-M // world)
    lo = min(r * per, M)
    return lo, min(lo + per, M)


def tiles_of(M, n, r, world):
    lo, hi = shard_rows(M, r, world)
    return (-(-(hi - lo) // 128)) * (-(-n // 128))


def run(world, ops, seed, regions=2):
    """ops: [(M, n)].  Returns None when every collect saw exactly its op's tiles, else a description."""
    rng = random.Random(seed)
    # per rank: landing regions (set of (op, sender, tile)), counters, host-side targets
    region = [[set() for _ in range(regions)] for _ in range(world)]
    ctr = [[0] * PF_SLOTS for _ in range(world)]
    target = [[0] * PF_SLOTS for _ in range(world)]
    want = [None] * world                 # (op, slot, target value) the rank's collect is waiting for
    step = [0] * world                    # next op a rank launches; its stream: gemm(s), collect(s), gemm(s+1), ...
    phase = ["gemm"] * world
    in_flight = []                        # (dst rank, op, sender, tile): stored by the sender, not yet landed (tile + its counter add)
    pending_tiles = [[] for _ in range(world)]      # tiles the rank's running GEMM still has to produce
    for r in range(world):
        pending_tiles[r] = None
    for _ in range(200000):
        acts = []
        for r in range(world):
            s = step[r]
            if s >= len(ops):
                continue
            M, n = ops[s]
            if phase[r] == "gemm":
                if pending_tiles[r] is None:
                    pending_tiles[r] = list(range(tiles_of(M, n, r, world)))
                    slot = s % PF_SLOTS
                    target[r][slot] += sum(tiles_of(M, n, p, world) for p in range(world) if p != r)
                    want[r] = (s, slot, target[r][slot])
                if pending_tiles[r]:
                    acts.append(("tile", r))
                else:
                    acts.append(("gemm_done", r))
            else:
                sw, slot, tv = want[r]
                if ctr[r][slot] >= tv:
                    acts.append(("collect", r))
        for i in range(len(in_flight)):
            acts.append(("land", i))
        if not acts:
            return None if all(s >= len(ops) for s in step) else f"deadlock at steps {step}"
        a = rng.choice(acts)
        if a[0] == "tile":
            r = a[1]
            t = pending_tiles[r].pop(rng.randrange(len(pending_tiles[r])))
            for p in range(world):
                if p != r:
                    in_flight.append((p, step[r], r, t))
        elif a[0] == "land":
            p, s, snd, t = in_flight.pop(a[1])
            reg = region[p][s % regions]
            other = [e for e in reg if e[0] != s]
            if other:
                return f"rank {snd}'s tile of op {s} lands in rank {p}'s region while it still holds op {other[0][0]}"
            reg.add((s, snd, t))
            ctr[p][s % PF_SLOTS] += 1                 # (release-add behind the tile's stores)
        elif a[0] == "gemm_done":
            phase[a[1]] = "collect"
        else:
            r = a[1]
            s, slot, tv = want[r]
            M, n = ops[s]
            reg = region[r][s % regions]
            need = {(s, p, t) for p in range(world) if p != r for t in range(tiles_of(M, n, p, world))}
            if reg != need:
                return f"rank {r} collects op {s} with {len(reg & need)} of {len(need)} tiles and {len(reg - need)} foreign ones"
            reg.clear()
            pending_tiles[r] = None
            phase[r] = "gemm"
            step[r] += 1
    return "did not finish"


SHAPES = [(8192, 512), (22016, 512), (32000, 512), (300, 64), (128, 33), (4096, 200)]


@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_two_regions_and_running_targets_are_enough(world):
    for seed in range(12):
        rng = random.Random(1000 * world + seed)
        ops = [rng.choice(SHAPES) for _ in range(3 * PF_SLOTS + 2)]
        assert run(world, ops, seed) is None


def test_one_region_is_not_enough():
    bad = 0
    for seed in range(20):
        ops = [(8192, 512)] * 6
        if run(2, ops, seed, regions=1) is not None:
            bad += 1
    assert bad > 0


def test_ranks_without_rows_still_collect():
    # M = 128 on 8 ranks: only rank 0 has rows; everybody else sends nothing and still has to wait for rank 0's tile
    for seed in range(6):
        assert run(8, [(128, 512), (128, 64), (8192, 512), (128, 512)], seed) is None
