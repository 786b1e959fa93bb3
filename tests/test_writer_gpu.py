"""GPU: vtn_writer_order (BCFOrderedWriter's output order computed on the device) through the C ABI against the
oracle's literal list emulation (pinned to the reference by tests/test_writer_oracle.py) and the reference's golden vectors."""
import os

import numpy as np
import pytest

import helpers
import test_writer_oracle as W

pytestmark = pytest.mark.gpu
NONE = np.iinfo(np.int32).min


def _check(ctx, oracle, rid, pos, window):
    order, late = ctx.writer_order(rid, pos, window)
    eo, el = oracle.writer_order(rid, pos, window)
    np.testing.assert_array_equal(order, eo)
    np.testing.assert_array_equal(late, el)


@pytest.mark.parametrize("name,window", W.GOLD)
def test_writer_reference_golden(ctx, name, window):
    rid, pos, _ = W.read_vcf(os.path.join(helpers.GOLDEN, name + "_in.vcf"))
    _, _, ks = W.read_vcf(os.path.join(helpers.GOLDEN, name + "_out.vcf"))
    order, late = ctx.writer_order(rid, pos, window)
    np.testing.assert_array_equal(order, ks)
    n_warn = open(os.path.join(helpers.GOLDEN, name + "_out.stderr")).read().count("Might not be sorted")
    assert int((late != NONE).sum()) == n_warn


@pytest.mark.parametrize("seed,n,window,jitter", [(1, 5000, 50, 10), (2, 5000, 5, 10), (3, 4000, 1, 3), (4, 20000, 200, 40), (5, 30000, 10000, 60),
                                                   (6, 3000, 2, 0), (7, 2500, 1000000, 30)])
def test_writer_matches_oracle(ctx, oracle, seed, n, window, jitter):
    rid, pos = W.gen_jittered(np.random.default_rng(seed), n, jitter)
    _check(ctx, oracle, rid, pos, window)


def test_writer_edge_cases(ctx, oracle):
    z = lambda *a: np.array(a, np.int32)
    order, late = ctx.writer_order(z(), z(), 10)
    assert len(order) == 0
    _check(ctx, oracle, z(0), z(5), 10)                                   # one record
    _check(ctx, oracle, z(0, 0, 0, 0), z(1, 1, 1, 1), 10)                 # every position 1: the front is kept back (size > 1 rule)
    _check(ctx, oracle, z(0, 0, 0, 0, 0), z(1, 0, 1, 0, 2), 10)           # positions 0 and 1 at a contig start
    _check(ctx, oracle, z(0, 0, 0, 0, 0, 0), z(0, 0, 1, 1, 0, 50), 3)
    _check(ctx, oracle, z(0, 1, 0, 1, 0, 1), z(9, 9, 8, 8, 7, 7), 100)    # rid changes at every record
    _check(ctx, oracle, z(0, 0, 0, 0, 0, 0), z(100, 90, 80, 70, 60, 50), 5)  # descending: each arrival is late
    _check(ctx, oracle, z(0, 0, 0, 0, 0, 0), z(100, 90, 80, 70, 60, 50), 500)
    _check(ctx, oracle, z(0, 0, 0), z(2147483000, 5, 2147483647), 10000)  # pos + window beyond int32
    # window 0: straight through
    order, late = ctx.writer_order(z(0, 0, 0), z(9, 3, 5), 0)
    assert list(order) == [0, 1, 2] and (late == NONE).all()
    # long equal-position runs and one very large flush (a gap far beyond the window)
    pos = np.concatenate([np.repeat(np.arange(10, 2010, dtype=np.int32), 3), np.array([10_000_000], np.int32)])
    rng = np.random.default_rng(3)
    pos[:-1] = pos[:-1] - rng.integers(0, 40, len(pos) - 1).astype(np.int32)
    _check(ctx, oracle, np.zeros(len(pos), np.int32), np.maximum(pos, 1), 10000)


def test_writer_fullsize_property(ctx):
    """4 M records whose disorder (<= 64 bp, like left-alignment shifts) stays inside the 10000 bp window: the writer's order is then
    the stable sort by position inside every contig run, and nothing is late."""
    rng = np.random.default_rng(11)
    n = 4_000_000
    base = np.sort(rng.integers(100, 50_000_000, n)).astype(np.int32)
    pos = base - (rng.integers(0, 65, n) * (rng.random(n) < 0.15)).astype(np.int32)
    rid = (np.arange(n) // (n // 4 + 1)).astype(np.int32)
    order, late = ctx.writer_order(rid, pos, 10000)
    assert (late == NONE).all()
    exp = np.lexsort((np.arange(n), pos, rid)).astype(np.uint32)
    np.testing.assert_array_equal(order, exp)


@pytest.mark.parametrize("seed,window,chunk", [(1, 50, 700), (2, 5, 333), (3, 10000, 1000), (4, 1, 97)])
def test_writer_streaming_equals_whole(ctx, oracle, seed, window, chunk):
    """chunks + carried-over pending records == one call over the whole stream (order and warnings of the new records)"""
    rid, pos = W.gen_jittered(np.random.default_rng(100 + seed), 6000, 12)
    whole, late_whole = oracle.writer_order(rid, pos, window)
    out, warned = [], []
    carry = np.zeros(0, np.int64)  # global indices still buffered, in arrival order
    for lo in range(0, len(rid), chunk):
        new = np.arange(lo, min(lo + chunk, len(rid)))
        idx = np.concatenate([carry, new])
        order, late, pend = ctx.writer_order(rid[idx], pos[idx], window, want_pending=True)
        done = len(idx) - pend
        out.extend(idx[order[:done]])
        carry = np.sort(idx[order[done:]])
        fresh = np.arange(len(carry_prev := idx[:len(idx) - len(new)]), len(idx))
        warned.extend(int(idx[k]) for k in fresh if late[k] != NONE)
    out.extend(carry[np.lexsort((carry, pos[carry]))])  # closing flush: ascending (pos, arrival); all of one contig run
    np.testing.assert_array_equal(np.array(out, np.uint32), whole)
    assert warned == [int(v) for v in np.nonzero(late_whole != NONE)[0]]
