"""CPU tests of the hazard analysis behind ggml_b200_set_create (ggml_b200_set_plan: pure host code).  The list must
behave like its ops run one after the other (the per-node loop of ggml_graph_compute, ggml.c:17541-17720); the plan
says where the persistent kernel re-quantizes x, which runs read x out of their producer's tagged rows, where a
dependency barrier is waited for and behind which op every CTA arrives at it."""
import pytest

from conftest import load_binding

X0, X1, X2, X3 = 0x10000000, 0x10100000, 0x10200000, 0x10300000
Y = [0x20000000 + 0x100000 * i for i in range(64)]


@pytest.fixture(scope="module")
def b200():
    return load_binding()


def test_ops_that_share_x_form_one_run_and_external_x_is_marked(b200):
    nb, linked, p = b200.set_plan([(X0, Y[0], 4096, 4096), (X0, Y[1], 4096, 4096), (X0, Y[2], 4096, 4096), (X1, Y[3], 4096, 4096)])
    assert nb == 0 and linked == 0
    assert [o["new_run"] for o in p] == [True, False, False, True]
    assert p[0]["ext"] and p[3]["ext"] and not any(o["barrier"] for o in p)


def test_x_that_is_one_ops_dst_links_instead_of_a_barrier(b200):
    # wq, wk -> wo reads wq's result -> w1 reads wo's result
    nb, linked, p = b200.set_plan([(X0, Y[0], 1024, 1024), (X0, Y[1], 1024, 1024), (Y[0], Y[2], 2048, 1024), (Y[2], Y[3], 512, 2048)])
    assert nb == 0 and linked == 2
    assert [o["link"] for o in p] == [-1, -1, 0, 2]
    assert not p[2]["ext"] and not p[3]["ext"] and p[0]["ext"]


def test_reused_dst_buffers_layer_after_layer_split_barriers(b200):
    """bench.py's list: x from outside, one dst per position of a layer, reused by every layer.  Every run of layer L+1
    waits for the ops of layer L that wrote its dsts, and the arrival point is the LAST of them, not the op before
    the run: wq(L+1) waits behind wv(L), four ops earlier."""
    layer = [(X0, 0), (X0, 1), (X0, 2), (X1, 3), (X2, 4), (X2, 5), (X3, 6)]
    ops = []
    for L in range(3):
        for (x, pos) in layer:
            ops.append((x, Y[pos], 4096, 11008 if pos == 6 else 4096))
    nb, linked, p = b200.set_plan(ops)
    assert linked == 0 and nb == 2 * 4                  # four runs per layer, none in the first
    heads = [i for i, o in enumerate(p) if o["new_run"]]
    assert heads == [0, 3, 4, 6, 7, 10, 11, 13, 14, 17, 18, 20]
    assert [i for i, o in enumerate(p) if o["barrier"]] == [7, 10, 11, 13, 14, 17, 18, 20]
    # arrival points: behind wv, wo, w1, w2 of the previous layer
    assert [i for i, o in enumerate(p) for _ in range(o["arrivals"])] == [2, 3, 5, 6, 9, 10, 12, 13]
    assert all(o["ext"] for i, o in enumerate(p) if o["new_run"])


def test_a_barrier_covers_everything_up_to_its_arrival_point(b200):
    # op 3 conflicts with op 0, op 4 with op 2; op 5 conflicts with op 1, which op 4's barrier (arrival behind op 2) already
    # covers: no third barrier, and arrival points come out in list order, each before its wait
    ops = [(X0, Y[0], 256, 1024), (X0, Y[1], 256, 1024), (X0, Y[2], 256, 1024),
           (X1, Y[0], 256, 1024), (X2, Y[2], 256, 1024), (X3, Y[1], 256, 1024)]
    nb, linked, p = b200.set_plan(ops)
    assert nb == 2
    arr = [i for i, o in enumerate(p) for _ in range(o["arrivals"])]
    assert arr == [0, 2]
    waits = [i for i, o in enumerate(p) if o["barrier"]]
    assert waits == [3, 4] and all(a < w for a, w in zip(arr, waits))
    # ... while a conflict with an op BEHIND the arrival point still gets its own barrier
    nb, linked, p = b200.set_plan(ops + [(X0, Y[0], 256, 1024)])      # rewrites op 3's dst
    assert nb == 3 and p[6]["barrier"] and [i for i, o in enumerate(p) for _ in range(o["arrivals"])] == [0, 2, 3]


def test_in_place_chain_and_repeated_dst(b200):
    buf, t = X0, Y[0]
    ops = [(buf, t, 1024, 1024), (t, buf, 1024, 1024), (buf, t, 1024, 1024), (buf, t, 1024, 1024)]
    nb, linked, p = b200.set_plan(ops)
    assert nb == 3 and [o["barrier"] for o in p] == [False, True, True, True]
    assert [i for i, o in enumerate(p) for _ in range(o["arrivals"])] == [0, 1, 2]


def test_two_row_shards_written_then_read_need_a_barrier_behind_the_later_one(b200):
    M = 1024
    ops = [(X0, Y[0], M, 1024, 0, 400), (X0, Y[0], M, 1024, 400, M), (X0, Y[1], 520, 1024), (Y[0], Y[2], 777, M)]
    nb, linked, p = b200.set_plan(ops)
    assert nb == 1 and linked == 0
    assert p[3]["barrier"] and p[3]["link"] == -1 and not p[3]["ext"]
    assert [o["arrivals"] for o in p] == [0, 1, 0, 0]


def test_empty_shards_are_skipped_and_a_shard_of_a_dst_does_not_link(b200):
    ops = [(X0, Y[0], 1088, 1024, 0, 384), (X0, Y[0], 1088, 1024, 384, 384), (X1, Y[1], 64, 1088)]
    nb, linked, p = b200.set_plan(ops)
    assert nb == 0 and p[1] == dict(barrier=False, new_run=False, ext=False, arrivals=0, link=-1)
    # x = a buffer of which the list wrote only rows [0, 384): not "the dst of one op" -> barrier, never a link
    ops = [(X0, Y[0], 1088, 1024, 0, 384), (Y[0], Y[1], 64, 1088)]
    nb, linked, p = b200.set_plan(ops)
    assert nb == 1 and linked == 0 and p[1]["barrier"] and p[1]["link"] == -1


def test_bad_arguments_are_rejected(b200):
    with pytest.raises(b200.B200Error):
        b200.set_plan([])
    with pytest.raises(b200.B200Error):
        b200.set_plan([(X0, Y[0], 64, 1000)])            # K not a multiple of 32


def test_tuning_keys_change_the_plan_not_the_semantics(b200):
    chain = [(X0, Y[0], 1024, 1024), (X0, Y[1], 1024, 1024), (Y[0], Y[2], 2048, 1024), (Y[2], Y[3], 512, 2048)]
    layer = [(X0, Y[0], 256, 1024), (X1, Y[1], 256, 1024), (X2, Y[2], 256, 1024)]
    try:
        b200.set_tuning("set_link", 0)                      # no tagged words: every dependency is a barrier
        nb, linked, p = b200.set_plan(chain)
        assert linked == 0 and nb == 2 and [o["barrier"] for o in p] == [False, False, True, True]
        assert [i for i, o in enumerate(p) for _ in range(o["arrivals"])] == [0, 2]      # behind the producer, not behind op 1
        b200.set_tuning("set_link", 1)
        b200.set_tuning("set_split_barrier", 0)             # arrive right before the wait: the classic grid barrier
        nb, linked, p = b200.set_plan(layer * 3)
        waits = [i for i, o in enumerate(p) if o["barrier"]]
        arr = [i for i, o in enumerate(p) for _ in range(o["arrivals"])]
        assert arr == [w - 1 for w in waits]
        # a full barrier covers everything before it: one per layer instead of one per run
        assert waits == [3, 6]
        b200.set_tuning("set_split_barrier", 1)
        nb, linked, p = b200.set_plan(layer * 3)
        assert [i for i, o in enumerate(p) if o["barrier"]] == [3, 4, 5, 6, 7, 8]
        assert [i for i, o in enumerate(p) for _ in range(o["arrivals"])] == [0, 1, 2, 3, 4, 5]
        b200.set_tuning("set_xext", 0)
        assert not any(o["ext"] for o in b200.set_plan(layer)[2])
    finally:
        for k, v in (("set_link", 1), ("set_split_barrier", 1), ("set_xext", 1)):
            b200.set_tuning(k, v)
