"""A model of the persistent kernel's weight ring (gemv_q4_0_set.cuh: producer thread, NG consumer groups, `full` /
`empty` mbarriers per slot, parity waits) under an adversarial scheduler -- bulk copies land in ANY order, groups run at
any relative speed.  It pins the protocol's two conditions, both found the hard way on the GPU (launch failures with
12 warps / 4 slots, DESIGN.md kernel 2s):
  * before waiting on `full`, a consumer waits for the slot's previous use to have been released (`empty`);
  * the ring is at least as deep as there are consumer groups (ggml_b200_set_create guarantees it).
A parity wait can only tell an mbarrier's current phase from the one before it; without these a wait falls through
one phase early and the group consumes a slot that still holds (or is receiving) another tile."""
import random

import pytest


class Mbar:
    """mbarrier with one arrival per phase; try_wait.parity(P) is true when the phase with parity P is the most
    recently COMPLETED one, i.e. the current (incomplete) phase has the other parity."""

    def __init__(self):
        self.phase = 0

    def arrive(self):
        self.phase += 1

    def try_wait(self, parity):
        return (self.phase & 1) != parity


def run(ns, ng, tiles, seed, wait_released=True):
    """Returns None when every tile was consumed from a slot holding exactly that tile, else a description."""
    rng = random.Random(seed)
    full = [Mbar() for _ in range(ns)]
    empty = [Mbar() for _ in range(ns)]
    content = [None] * ns
    issued = 0
    in_flight = []
    nxt = list(range(ng))                 # next tile of every group
    done = 0
    for _ in range(100 * tiles + 1000):
        acts = []
        if issued < tiles:
            s, use = issued % ns, issued // ns
            if use == 0 or empty[s].try_wait((use - 1) & 1):
                acts.append(("issue",))
        for j in in_flight:
            acts.append(("land", j))
        for g in range(ng):
            j = nxt[g]
            if j >= tiles:
                continue
            s, use = j % ns, j // ns
            if wait_released and use > 0 and not empty[s].try_wait((use - 1) & 1):
                continue
            if full[s].try_wait(use & 1):
                acts.append(("consume", g))
        if not acts:
            return None if done == tiles else f"deadlock after {done} tiles"
        a = rng.choice(acts)
        if a[0] == "issue":
            in_flight.append(issued)
            issued += 1
        elif a[0] == "land":
            j = a[1]
            in_flight.remove(j)
            content[j % ns] = j
            full[j % ns].arrive()
        else:
            g = a[1]
            j = nxt[g]
            if content[j % ns] != j:
                return f"group {g} consumed slot {j % ns} for tile {j} while it held tile {content[j % ns]}"
            empty[j % ns].arrive()
            nxt[g] += ng
            done += 1
    return "no progress"


@pytest.mark.parametrize("ns,ng", [(4, 3), (3, 2), (8, 3), (5, 4), (3, 3), (4, 4), (4, 2), (7, 3), (16, 5)])
def test_protocol_holds_when_ring_is_at_least_as_deep_as_the_groups(ns, ng):
    for seed in range(300):
        assert run(ns, ng, 60, seed) is None, (ns, ng, seed)


@pytest.mark.parametrize("ns,ng", [(4, 3), (3, 2)])
def test_without_the_released_wait_the_parity_wait_falls_through(ns, ng):
    """the configurations that failed on the GPU: ring depth = groups + 1"""
    assert any(run(ns, ng, 60, seed, wait_released=False) is not None for seed in range(300))


@pytest.mark.parametrize("ns,ng", [(3, 4), (2, 3), (4, 5)])
def test_a_ring_shallower_than_the_groups_is_not_safe_even_with_it(ns, ng):
    """why ggml_b200_set_create lowers the consumer warps instead"""
    assert any(run(ns, ng, 60, seed) is not None for seed in range(500))


# ---- split dependency barriers: why every barrier has its own arrival counter -----------------------------------
def run_barriers(n_cta, arrive_at, wait_at, n_ops, seed, per_barrier):
    """CTAs walk ops 0..n_ops-1 at any relative speed.  Barrier k: a CTA arrives when it has finished op arrive_at[k]
    and must not start op wait_at[k] before EVERY CTA has finished op arrive_at[k].  Returns a violation or None."""
    rng = random.Random(seed)
    pos = [0] * n_cta                     # next op of every CTA
    single = 0
    cnt = [0] * len(arrive_at)
    for _ in range(100 * n_cta * n_ops):
        movable = []
        for c in range(n_cta):
            o = pos[c]
            if o >= n_ops:
                continue
            blocked = False
            for k, w in enumerate(wait_at):
                if w == o:
                    ok = cnt[k] >= n_cta if per_barrier else single >= (k + 1) * n_cta
                    if not ok:
                        blocked = True
                    elif min(pos) <= arrive_at[k]:
                        return f"CTA {c} passed barrier {k} while a CTA was still at op {min(pos)} <= {arrive_at[k]}"
            if not blocked:
                movable.append(c)
        if not movable:
            return None if min(pos) >= n_ops else "deadlock"
        c = rng.choice(movable)
        o = pos[c]
        pos[c] += 1                       # op o finished
        for k, a in enumerate(arrive_at):
            if a == o:
                cnt[k] += 1
                single += 1
    return "no progress"


def test_per_barrier_counters_hold_and_one_running_counter_does_not():
    # the decode list's pattern: arrival points early (several before the first wait), waits ops later
    arrive_at, wait_at, n_ops = [2, 3, 5, 6], [7, 10, 11, 13], 14
    for seed in range(300):
        assert run_barriers(4, arrive_at, wait_at, n_ops, seed, per_barrier=True) is None
    assert any(run_barriers(4, arrive_at, wait_at, n_ops, seed, per_barrier=False) is not None for seed in range(300))


# ---- inside a CTA: why the arrival is counted per WARP (advisor r1) ---------------------------------------------------
def run_cta_arrivals(n_groups, warps_per_group, ops_with_tile, arrive_ops, max_skew, seed, per_warp):
    """One CTA.  Group g has a tile only in the ops of ops_with_tile[g]; a warp that has no tile in an op walks through
    it.  Every warp arrives (counts itself) at the ops in arrive_ops.  The sentinel forwards arrival e when
      per_warp:  min over warps of (arrival points passed) >= e + 1
      else:      the CTA-wide sum of arrivals >= (e + 1) * W          (the round-1 code)
    Warps of different groups are coupled only by the ring: a warp may be at most max_skew ops ahead of the slowest.
    Returns a violation (arrival e forwarded while some warp has not finished op arrive_ops[e]) or None."""
    rng = random.Random(seed)
    W = n_groups * warps_per_group
    n_ops = max(arrive_ops) + 2
    pos = [0] * W                         # ops finished per warp
    arrived = [0] * W
    forwarded = 0
    for _ in range(200 * W * n_ops):
        while forwarded < len(arrive_ops):
            ok = min(arrived) >= forwarded + 1 if per_warp else sum(arrived) >= (forwarded + 1) * W
            if not ok:
                break
            if min(pos) <= arrive_ops[forwarded]:
                return f"arrival {forwarded} forwarded while a warp is still at op {min(pos)} <= {arrive_ops[forwarded]}"
            forwarded += 1
        movable = [w for w in range(W) if pos[w] < n_ops and pos[w] - min(pos) < max_skew]
        if not movable:
            return None
        # a group with a tile in its current op is slow (it waits for the bulk copy); empty-handed groups run ahead
        fast = [w for w in movable if pos[w] not in ops_with_tile[w // warps_per_group]]
        w = rng.choice(fast if fast and rng.random() < 0.9 else movable)
        if pos[w] in arrive_ops:
            arrived[w] += 1
        pos[w] += 1
    return None


def test_arrivals_counted_per_warp_hold_and_a_cta_wide_count_does_not():
    # 3 groups x 4 warps (W = 12); group 0 holds the only tile of op 0, the others have none in ops 0..3;
    # every op is an arrival point; the ring lets a group run 4 ops ahead: 8 warps x 2 arrivals = 16 >= 12
    tiles = [{0}, set(), set()]
    arrive_ops = [0, 1, 2, 3]
    for seed in range(300):
        assert run_cta_arrivals(3, 4, tiles, arrive_ops, 4, seed, per_warp=True) is None
    assert any(run_cta_arrivals(3, 4, tiles, arrive_ops, 4, seed, per_warp=False) is not None for seed in range(300))
