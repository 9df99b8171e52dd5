"""GPU parity tests of the persistent multi-op GEMV (ggml_b200_set_*, gemv_q4_0_set.cuh): a LIST of n = 1
Q4_0 x F32 mul_mats in one launch.  Contract: result bits == ggml_b200_mul_mat_q4_0 called per op in list order
(which test_gpu_parity.py pins against the oracle); S_b bit-exact against the oracle; outputs within 1e-4."""
import ctypes as C

import numpy as np
import pytest

from test_gpu_parity import assert_close, make_weights, make_x, oracle_Y

pytestmark = pytest.mark.gpu


def _per_op(b200, ops):
    for (w, xp, yp) in ops:
        b200.mul_mat_dev(w, xp, 1, yp)
    b200.synchronize()


def _chain(b200, port, ref, shapes, seed):
    """A decode-like chain: groups of ops that share x; the first result of a group is the next group's x.
    shapes: list of groups, each a list of (M, K); M of a group's first op == K of the next group."""
    Ws, ws, ops_desc = [], [], []
    for gi, grp in enumerate(shapes):
        for oi, (M, K) in enumerate(grp):
            W = make_weights(port, ref, M, K, seed + 17 * gi + oi)
            Ws.append(W)
            ws.append(b200.Weight(W))
            ops_desc.append((gi, oi, M, K))
    return Ws, ws, ops_desc


def _buffers(b200, port, shapes, ops_desc, x0):
    xd0 = b200.DevBuf.from_numpy(x0)
    yd = [b200.DevBuf(max(4, M * 4)) for (_, _, M, _) in ops_desc]
    first_of_group = {}
    for i, (gi, oi, _, _) in enumerate(ops_desc):
        if oi == 0:
            first_of_group[gi] = i
    xs = []
    for (gi, oi, M, K) in ops_desc:
        xs.append(xd0.ptr if gi == 0 else yd[first_of_group[gi - 1]].ptr)
    return xd0, yd, xs, first_of_group


CHAIN_SMALL_K = [[(1024, 1024), (1003, 1024), (600, 1024)], [(2048, 1024)], [(1024, 2048), (1021, 2048)], [(2052, 1024)]]
CHAIN_MIXED_K = [[(1024, 1024), (1003, 1024), (600, 1024)], [(5120, 1024)], [(1024, 5120), (1021, 5120)], [(11008, 1024)], [(515, 11008)]]


@pytest.mark.parametrize("warps,shapes,exact", [(16, CHAIN_SMALL_K, True), (8, CHAIN_SMALL_K, True), (12, CHAIN_SMALL_K, True),
                                                (20, CHAIN_SMALL_K, True), (16, CHAIN_MIXED_K, False), (12, CHAIN_MIXED_K, False)])
def test_set_chain_bit_identical_to_per_op_calls(b200, port, ref, warps, shapes, exact):
    """qkv-like run -> op on its first result -> w3/w1-like run -> down projection -> ...; ragged tiles (M % 8 != 0),
    more tiles than ring slots, 30 runs back to back (the sync words must be left zero by every launch).
    K <= 4096: one warp per row, result bits identical to the per-op kernel.  K > 4096 (5120, 11008): four warps per
    row, partials added in a fixed order -- checked against the oracle along the chain."""
    Ws, ws, desc = _chain(b200, port, ref, shapes, 900)
    x0 = make_x(port, 1, 1024, 41)
    b200.set_tuning("set_warps", warps)
    try:
        xd0, yd, xs, fog = _buffers(b200, port, shapes, desc, x0)
        ops = [(ws[i], xs[i], yd[i].ptr) for i in range(len(ws))]
        _per_op(b200, ops)
        want = [yd[i].to_numpy(np.float32, (1, desc[i][2])) for i in range(len(ws))]
        for y in yd:
            b200.lib().ggml_b200_memcpy_h2d(y.ptr, np.zeros(y.nbytes, np.uint8).ctypes.data_as(C.c_void_p), y.nbytes)
        s = b200.OpSet(ops)
        assert s.n_ops == len(ops)
        assert s.barriers == 0 and s.linked_runs == len(shapes) - 1     # every change of x is "the dst of one earlier op"
        assert s.slots >= 3             # one more than the consumer groups (8 warps = 2 groups) or what fits
        c0 = b200.launch_count()
        for _ in range(30):
            s.run()
        b200.synchronize()
        assert b200.launch_count() - c0 == 30          # ONE launch per list
        got = [yd[i].to_numpy(np.float32, (1, desc[i][2])) for i in range(len(ws))]
        for i in range(len(ws)):
            if exact:
                assert np.array_equal(got[i], want[i]), f"op {i} {desc[i]}: set result differs from the per-op call"
        # and against the CPU truth, following the chain
        xin = {0: x0}
        for i, (gi, oi, M, K) in enumerate(desc):
            assert_close(got[i], oracle_Y(port, ref, Ws[i], xin[gi]))
            if oi == 0:
                xin[gi + 1] = got[i]
        s.free()
    finally:
        b200.set_tuning("set_warps", 12)
    for w in ws:
        w.free()


def test_set_block_dots_bit_exact(b200, port, ref):
    """S_b as the persistent kernel forms them == the oracle's int32 block dots (ggml.c:2593-2600)."""
    # every register plan of the kernel: blocks per lane 1..4 with one warp per row, 2/3/4/6 with four
    shapes = [(520, 2048), (64, 2048), (300, 5120), (33, 1024), (70, 3072), (129, 4096), (70, 11008), (50, 16384), (41, 22016)]
    Ws = [make_weights(port, ref, M, K, 950 + i) for i, (M, K) in enumerate(shapes)]
    ws = [b200.Weight(W) for W in Ws]
    xh = [make_x(port, 1, K, 5 + i) for i, (M, K) in enumerate(shapes)]
    xh[1] = xh[0]
    xdev = [b200.DevBuf.from_numpy(x) for x in xh]
    xdev[1] = xdev[0]
    yd = [b200.DevBuf(W.shape[0] * 4) for W in Ws]
    Sd = [b200.DevBuf(W.shape[0] * W.shape[1] * 4) for W in Ws]
    ops = [(ws[i], xdev[i].ptr, yd[i].ptr) for i in range(len(ws))]
    s = b200.OpSet(ops)
    assert s.barriers == 0 and s.linked_runs == 0      # every x comes from outside the list
    s.checkpoint([b.ptr for b in Sd])
    b200.synchronize()
    for i, (W, x) in enumerate(zip(Ws, xh)):
        S = Sd[i].to_numpy(np.int32, (W.shape[0], W.shape[1]))
        want = port.block_dots_matrix(W, port.quantize_q8_0(x))[0]
        assert np.array_equal(S, want), f"op {i}: int32 block dots differ"
        assert_close(yd[i].to_numpy(np.float32, (1, W.shape[0])), oracle_Y(port, ref, W, x))
    s.free()
    for w in ws:
        w.free()


def test_set_in_place_and_repeated_dst_are_ordered(b200, port, ref):
    """An op that writes a buffer an earlier op of the list reads or writes must wait for it: y -> x chains
    through the SAME buffer (square weights), and two ops with the same dst (the later one wins)."""
    M = K = 1024
    Ws = [make_weights(port, ref, M, K, 970 + i) for i in range(4)]
    ws = [b200.Weight(W) for W in Ws]
    x0 = make_x(port, 1, K, 99)
    buf = b200.DevBuf.from_numpy(x0)          # x of op 0, overwritten by op 1
    t = b200.DevBuf(M * 4)
    ops = [(ws[0], buf.ptr, t.ptr),           # t = W0 x0
           (ws[1], t.ptr, buf.ptr),           # buf = W1 t        (overwrites op 0's x)
           (ws[2], buf.ptr, t.ptr),           # t = W2 buf        (overwrites op 1's x)
           (ws[3], buf.ptr, t.ptr)]           # t = W3 buf        (same dst as op 2: must run after it)
    s = b200.OpSet(ops)
    assert s.barriers == 3
    s.run()
    b200.synchronize()
    got_t, got_buf = t.to_numpy(np.float32, (1, M)), buf.to_numpy(np.float32, (1, M))
    b200.lib().ggml_b200_memcpy_h2d(buf.ptr, x0.ctypes.data_as(C.c_void_p), x0.nbytes)
    _per_op(b200, ops)
    assert np.array_equal(got_t, t.to_numpy(np.float32, (1, M)))
    assert np.array_equal(got_buf, buf.to_numpy(np.float32, (1, M)))
    y0 = oracle_Y(port, ref, Ws[0], x0)
    y1 = oracle_Y(port, ref, Ws[1], y0)
    assert_close(got_buf, y1)
    s.free()
    for w in ws:
        w.free()


def test_set_row_shards_and_empty_shard(b200, port, ref):
    """A shard writes only [row_lo,row_hi) of the full dst; an empty shard is skipped (multi-GPU row split)."""
    M, K = 1100, 1024
    W = make_weights(port, ref, M, K, 990)
    x = make_x(port, 1, K, 7)
    parts = [b200.Weight(W, 0, 384), b200.Weight(W, 384, 384), b200.Weight(W, 384, 1100)]
    xd = b200.DevBuf.from_numpy(x)
    yd = b200.DevBuf.from_numpy(np.full((1, M), -7.0, np.float32))
    s = b200.OpSet([(p, xd.ptr, yd.ptr) for p in parts])
    assert s.n_ops == 2 and s.barriers == 0
    s.run()
    b200.synchronize()
    assert_close(yd.to_numpy(np.float32, (1, M)), oracle_Y(port, ref, W, x))
    s.free()
    for p in parts:
        p.free()


def test_set_rejects_what_it_cannot_run(b200, port, ref):
    W = make_weights(port, ref, 64, 1024, 995)
    x = b200.DevBuf(1024 * 4 + 16)
    y = b200.DevBuf(64 * 4)
    b200.set_repack(False)
    try:
        raw = b200.Weight(W)
    finally:
        b200.set_repack(True)
    with pytest.raises(b200.B200Error):
        b200.OpSet([(raw, x.ptr, y.ptr)])                 # RAW layout
    w = b200.Weight(W)
    with pytest.raises(b200.B200Error):
        b200.OpSet([(w, x.ptr + 4, y.ptr)])               # x not 16-byte aligned
    with pytest.raises(b200.B200Error):
        b200.OpSet([])
    raw.free(); w.free()


def test_set_replays_in_a_cuda_graph(b200, port, ref):
    """The launch is capturable: device-side sync words need no host reset between replays."""
    torch = pytest.importorskip("torch")
    M = K = 1024
    Ws = [make_weights(port, ref, M, K, 1010 + i) for i in range(3)]
    ws = [b200.Weight(W) for W in Ws]
    x0 = make_x(port, 1, K, 3)
    xd = b200.DevBuf.from_numpy(x0)
    yd = [b200.DevBuf(M * 4) for _ in range(3)]
    ops = [(ws[0], xd.ptr, yd[0].ptr), (ws[1], yd[0].ptr, yd[1].ptr), (ws[2], yd[1].ptr, yd[2].ptr)]
    s = b200.OpSet(ops)
    assert s.linked_runs == 2
    stream = torch.cuda.Stream()
    old = b200.lib().ggml_b200_get_stream()
    try:
        b200.set_stream(stream.cuda_stream)
        with torch.cuda.stream(stream):
            s.run()
        stream.synchronize()
        want = yd[2].to_numpy(np.float32, (1, M))
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            b200.set_stream(torch.cuda.current_stream().cuda_stream)
            s.run()
        b200.set_stream(stream.cuda_stream)
        b200.lib().ggml_b200_memcpy_h2d(yd[2].ptr, np.zeros(M * 4, np.uint8).ctypes.data_as(C.c_void_p), M * 4)
        with torch.cuda.stream(stream):
            for _ in range(5):
                graph.replay()
        stream.synchronize()
        assert np.array_equal(yd[2].to_numpy(np.float32, (1, M)), want)
        del graph
    finally:
        b200.set_stream(None)
        assert b200.lib().ggml_b200_get_stream() == old or True
    s.free()
    for w in ws:
        w.free()


@pytest.mark.parametrize("warps,slots", [(12, 0), (12, 4), (8, 3), (16, 5), (12, 8)])
def test_set_tight_ring_many_tiles_per_cta(b200, port, ref, warps, slots):
    """Regression: with a ring one slot deeper than the consumer groups, a group's next tile lives in the slot another
    group is still being served from; a bulk copy that lands before an older one must not let the parity wait fall
    through on the phase before (K = 8192: four warps per row, 14 tiles per CTA, natural ring depth; launch failure
    before the fix).  30 launches back to back, results against the oracle and bit-identical from launch to launch."""
    M, K = 2048, 8192
    Ws = [make_weights(port, ref, M, K, 1100 + i) for i in range(4)]
    ws = [b200.Weight(W) for W in Ws]
    x = make_x(port, 1, K, 77)
    xd = b200.DevBuf.from_numpy(x)
    yd = [b200.DevBuf(M * 4) for _ in range(4)]
    b200.set_tuning("set_warps", warps)
    b200.set_tuning("set_slots", slots)
    try:
        s = b200.OpSet([(ws[i], xd.ptr, yd[i].ptr) for i in range(4)])
        assert s.barriers == 0 and s.linked_runs == 0
        s.run()
        b200.synchronize()
        first = [y.to_numpy(np.float32, (1, M)) for y in yd]
        for _ in range(30):
            s.run()
        b200.synchronize()
        for i in range(4):
            got = yd[i].to_numpy(np.float32, (1, M))
            assert np.array_equal(got, first[i])
            assert_close(got, oracle_Y(port, ref, Ws[i], x))
        s.free()
    finally:
        b200.set_tuning("set_warps", 12)
        b200.set_tuning("set_slots", 0)
    for w in ws:
        w.free()


def test_set_reused_dst_buffers_layer_after_layer(b200, port, ref):
    """The decode graph's buffer pattern: x comes from outside the list, every position of a layer has one dst that the
    next layer overwrites (write-after-write).  The barriers are split -- a CTA arrives when it has finished the op the
    later run conflicts with and the wait happens ops later -- and the last writer must win, launch after launch."""
    M = K = 1024
    L, P = 6, 4                                    # layers x positions; positions 0,1 share x
    Ws = [make_weights(port, ref, M + 8 * (i % P), K, 1200 + i) for i in range(L * P)]
    ws = [b200.Weight(W) for W in Ws]
    xh = [make_x(port, 1, K, 300 + g) for g in range(3)]
    xd = [b200.DevBuf.from_numpy(x) for x in xh]
    grp = [0, 0, 1, 2]
    yd = [b200.DevBuf((M + 8 * p) * 4) for p in range(P)]
    ops = [(ws[i], xd[grp[i % P]].ptr, yd[i % P].ptr) for i in range(L * P)]
    s = b200.OpSet(ops)
    assert s.linked_runs == 0 and 0 < s.barriers <= 3 * (L - 1) + 1
    for _ in range(20):
        s.run()
    b200.synchronize()
    for p in range(P):
        i = (L - 1) * P + p
        assert_close(yd[p].to_numpy(np.float32, (1, M + 8 * p)), oracle_Y(port, ref, Ws[i], xh[grp[p]]))
    got = [yd[p].to_numpy(np.float32, (1, M + 8 * p)) for p in range(P)]
    _per_op(b200, ops)
    for p in range(P):
        assert np.array_equal(got[p], yd[p].to_numpy(np.float32, (1, M + 8 * p)))
    s.free()
    for w in ws:
        w.free()


def test_set_x_written_by_two_row_shards_waits_for_both(b200, port, ref):
    """x of the last op is the dst of TWO earlier ops (row shards of one matrix): not a single producer, so the run
    waits at a barrier whose arrival point is the later shard; an unrelated op sits between them and the reader."""
    M, K = 1024, 1024
    W0 = make_weights(port, ref, M, K, 1300)
    W1 = make_weights(port, ref, 520, K, 1301)
    W2 = make_weights(port, ref, 777, M, 1302)
    x0 = make_x(port, 1, K, 5)
    xd = b200.DevBuf.from_numpy(x0)
    y0 = b200.DevBuf(M * 4)
    y1 = b200.DevBuf(520 * 4)
    y2 = b200.DevBuf(777 * 4)
    parts = [b200.Weight(W0, 0, 400), b200.Weight(W0, 400, M)]
    w1, w2 = b200.Weight(W1), b200.Weight(W2)
    ops = [(parts[0], xd.ptr, y0.ptr), (parts[1], xd.ptr, y0.ptr), (w1, xd.ptr, y1.ptr), (w2, y0.ptr, y2.ptr)]
    s = b200.OpSet(ops)
    assert s.barriers == 1 and s.linked_runs == 0
    for _ in range(10):
        s.run()
    b200.synchronize()
    h = oracle_Y(port, ref, W0, x0)
    assert_close(y0.to_numpy(np.float32, (1, M)), h)
    assert_close(y1.to_numpy(np.float32, (1, 520)), oracle_Y(port, ref, W1, x0))
    got = y2.to_numpy(np.float32, (1, 777))
    assert_close(got, oracle_Y(port, ref, W2, y0.to_numpy(np.float32, (1, M))))
    _per_op(b200, ops)
    assert np.array_equal(got, y2.to_numpy(np.float32, (1, 777)))
    s.free()
    for w in parts + [w1, w2]:
        w.free()


def test_set_long_list_walks_the_table_in_global_memory(b200, port, ref):
    """More ops than the shared-memory op table holds (32 KB / 48 B): consumers, producer, sentinel and x warps read
    the descriptors from global memory.  800 small ops on three weights, three x, four dst buffers reused round-robin
    (write-after-write barriers all along); the last writer of every dst must win."""
    shapes = [(72, 1024), (64, 2048), (40, 1024)]
    Ws = [make_weights(port, ref, M, K, 1400 + i) for i, (M, K) in enumerate(shapes)]
    ws = [b200.Weight(W) for W in Ws]
    xh = {1024: make_x(port, 1, 1024, 11), 2048: make_x(port, 1, 2048, 12)}
    xd = {K: b200.DevBuf.from_numpy(x) for K, x in xh.items()}
    yd = [b200.DevBuf(72 * 4) for _ in range(4)]
    ops, last = [], {}
    for i in range(800):
        k = (i * 7 + i // 5) % 3
        d = i % 4
        ops.append((ws[k], xd[shapes[k][1]].ptr, yd[d].ptr))
        last[d] = k
    s = b200.OpSet(ops)
    assert s.n_ops == 800 and s.barriers > 0
    for _ in range(5):
        s.run()
    b200.synchronize()
    for d, k in last.items():
        M, K = shapes[k]
        assert_close(yd[d].to_numpy(np.float32, (1, 72))[:, :M], oracle_Y(port, ref, Ws[k], xh[K]))
    s.free()
    for w in ws:
        w.free()


def test_set_ring_never_shallower_than_the_consumer_groups(b200, port, ref):
    """K = 22016 with 20 consumer warps: four 49.5 KB slots fit, five groups were asked for.  The parity waits of the
    tile loop need ring depth >= groups, so create takes 16 warps; many tiles per CTA, results against the oracle."""
    M, K = 600, 22016
    Ws = [make_weights(port, ref, M, K, 1500 + i) for i in range(3)]
    ws = [b200.Weight(W) for W in Ws]
    x = make_x(port, 1, K, 21)
    xd = b200.DevBuf.from_numpy(x)
    yd = [b200.DevBuf(M * 4) for _ in range(12)]
    b200.set_tuning("set_warps", 20)
    try:
        s = b200.OpSet([(ws[i % 3], xd.ptr, yd[i].ptr) for i in range(12)])     # 12 x 150 tiles over 148 CTAs
        assert s.slots >= 2 and s.warps == 4 * s.slots and s.warps < 20
        for _ in range(10):
            s.run()
        b200.synchronize()
        want = [oracle_Y(port, ref, W, x) for W in Ws]
        for i in range(12):
            assert_close(yd[i].to_numpy(np.float32, (1, M)), want[i % 3])
        s.free()
    finally:
        b200.set_tuning("set_warps", 12)
    for w in ws:
        w.free()


@pytest.mark.parametrize("warps", [8, 12, 16])
def test_set_many_tiny_ops_barrier_after_each(b200, port, ref, warps):
    """Advisor r1 (high): with about one tile or fewer per CTA and op, consumer groups without a tile walk straight
    through the arrival points; a CTA-wide arrival COUNT let two fast groups stand in for a slow one and released a
    barrier before every writer had stored.  The arrival is now per warp (minimum over the CTA's warps).  120 ops of
    8-300 rows (1-38 tiles over 148 CTAs): even ops write one of four scratch buffers round-robin (write-after-write
    barriers), odd ops read a scratch buffer written earlier through a partial overlap (read-after-write barriers).
    Every odd op's result and the final scratch contents are compared with the per-op path, 6 launches."""
    K = 1024
    rows = [24, 300, 64, 152, 40, 96, 8, 256]
    Ws = [make_weights(port, ref, M, K, 1600 + i) for i, M in enumerate(rows)]
    ws = [b200.Weight(W) for W in Ws]
    x0 = make_x(port, 1, K, 31)
    xd = b200.DevBuf.from_numpy(x0)
    init = [np.ascontiguousarray(make_x(port, 1, K + 16, 77 + j).reshape(-1)) for j in range(4)]
    scratch = [b200.DevBuf.from_numpy(v) for v in init]          # dst of the even ops; first K floats = x of the odd ops
    outs = [b200.DevBuf(300 * 4) for _ in range(120)]
    ops = []
    for i in range(120):
        k = (5 * i + i // 7) % len(rows)
        if i % 2 == 0:
            ops.append((ws[k], xd.ptr, scratch[(i // 2) % 4].ptr))
        else:
            ops.append((ws[k], scratch[((i // 2) + 2) % 4].ptr, outs[i].ptr))

    def reset():
        for sb, v in zip(scratch, init):
            b200.lib().ggml_b200_memcpy_h2d(sb.ptr, v.ctypes.data_as(C.c_void_p), v.nbytes)

    b200.set_tuning("set_warps", warps)
    try:
        s = b200.OpSet(ops)
        assert s.barriers >= 50 and s.linked_runs == 0
        reset()
        _per_op(b200, ops)
        want = [o.to_numpy(np.float32, (300,)).copy() for o in outs]
        want_s = [sb.to_numpy(np.float32, (K + 16,)).copy() for sb in scratch]
        for rep in range(6):
            reset()
            for o in outs:
                b200.lib().ggml_b200_memcpy_h2d(o.ptr, np.zeros(300, np.float32).ctypes.data_as(C.c_void_p), 1200)
            s.run()
            b200.synchronize()
            for i in range(1, 120, 2):
                M = rows[(5 * i + i // 7) % len(rows)]
                assert np.array_equal(outs[i].to_numpy(np.float32, (300,))[:M], want[i][:M]), f"rep {rep}: op {i} read its x before the writer was done"
            for j, sb in enumerate(scratch):
                assert np.array_equal(sb.to_numpy(np.float32, (K + 16,)), want_s[j]), f"rep {rep}: scratch {j}: the wrong writer won"
        s.free()
    finally:
        b200.set_tuning("set_warps", 12)
    for w in ws:
        w.free()


def test_set_x_of_an_op_a_barrier_already_covers(b200, port, ref):
    """Advisor r1 (medium): op3 reads the dst of op0, but a write-after-write barrier between them (op2 rewrites op1's
    dst; its own x is external) already covers op0, so the hazard analysis gives op3's run no flag of its own.  The x
    warps run up to two runs ahead of the consumers and used to wait only at barrier heads with non-external x: they
    could quantize y_A before the barrier that guarantees it.  Big op0 (slow to finish), tiny ops behind it."""
    K = 1024
    W0 = make_weights(port, ref, 1024, K, 1700)       # y_A, K floats: the x of op3
    W1 = make_weights(port, ref, 64, K, 1701)
    W2 = make_weights(port, ref, 64, K, 1702)
    W3 = make_weights(port, ref, 200, 1024, 1703)
    ws = [b200.Weight(W) for W in (W0, W1, W2, W3)]
    x0, x1 = make_x(port, 1, K, 41), make_x(port, 1, K, 42)
    xd0, xd1 = b200.DevBuf.from_numpy(x0), b200.DevBuf.from_numpy(x1)
    yA, yB, yC = b200.DevBuf(1024 * 4), b200.DevBuf(64 * 4), b200.DevBuf(200 * 4)
    ops = [(ws[0], xd0.ptr, yA.ptr), (ws[1], xd0.ptr, yB.ptr), (ws[2], xd1.ptr, yB.ptr), (ws[3], yA.ptr, yC.ptr)]
    s = b200.OpSet(ops)
    assert s.barriers == 1 and s.linked_runs == 0      # the WAW barrier covers op0; op3's run has no flag of its own
    want_A = oracle_Y(port, ref, W0, x0)
    for rep in range(20):
        b200.lib().ggml_b200_memcpy_h2d(yA.ptr, np.full(1024, 1e30, np.float32).ctypes.data_as(C.c_void_p), 4096)   # stale x would be loud
        s.run()
        b200.synchronize()
        got_A = yA.to_numpy(np.float32, (1, 1024))
        assert_close(got_A, want_A)
        assert_close(yC.to_numpy(np.float32, (1, 200)), oracle_Y(port, ref, W3, got_A))
        assert_close(yB.to_numpy(np.float32, (1, 64)), oracle_Y(port, ref, W2, x1))
    s.free()
    for w in ws:
        w.free()
