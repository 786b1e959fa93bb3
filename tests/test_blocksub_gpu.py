"""GPU: vtn_blocksub (vt decompose_blocksub's fan-out plan on the device) through the C ABI: bit-exact against the oracle,
and against the reference's fixtures / vt_ref outputs (tests/golden/blocksub_*)."""
import numpy as np
import pytest

import blocksub_helpers as B
import helpers

pytestmark = pytest.mark.gpu


def _check(ctx, oracle, batch, keep, dist):
    got = ctx.blocksub(batch, keep, dist)
    exp = oracle.blocksub_plan(batch, keep, dist)
    for g, e, name in zip(got[:3], exp[:3], ("src_rec", "off", "len")):
        np.testing.assert_array_equal(g, e, err_msg=name)
    assert got[3] == exp[3]
    return got


@pytest.mark.parametrize("stem,flags,tag", B.CASES)
def test_blocksub_reference_golden(ctx, stem, flags, tag):
    vin, vout, verr = B.paths(stem, tag)
    sites = B.read_sites(vin)
    keep, dist = B.mode(flags)
    src, off, ln, cnt = ctx.blocksub(B.batch_of(sites), keep, dist)
    assert B.stable_by_pos(B.apply_plan(sites, src, off, ln)) == B.read_sites(vout)
    assert cnt == B.stats_of(open(verr).read())


def _random_batch(rng, n, bs_frac=0.3, maxlen=40):
    recs, pos = [], 10
    for _ in range(n):
        pos += int(rng.integers(0, 30))
        u = rng.random()
        L = int(rng.integers(2, maxlen))
        ref = "".join(rng.choice(list("ACGTN"), L))
        if u < bs_frac:
            p = rng.choice([0.0, 0.1, 0.5, 0.9, 1.0])
            al = [ref, "".join(c if rng.random() < p else rng.choice(list("ACGTacgt")) for c in ref)]
        elif u < bs_frac + 0.1:
            al = [ref, ref[::-1], ref[:1]]          # multiallelic: passes through
        elif u < bs_frac + 0.2:
            al = [ref]                              # ALT "."
        elif u < bs_frac + 0.4:
            al = [ref, ref[: L // 2]]               # deletion
        else:
            al = [ref[:1], "ACGT"[int(rng.integers(0, 4))]]  # one base each (also REF == ALT)
        recs.append((0, pos, al))
    return helpers.Batch.from_records(recs)


@pytest.mark.parametrize("seed,n,keep,dist", [(1, 3000, False, 2), (2, 3000, True, 2), (3, 20000, True, 1), (4, 5000, True, 5), (5, 4000, True, 0),
                                              (6, 4000, True, -1), (7, 2100, False, 2)])
def test_blocksub_matches_oracle(ctx, oracle, seed, n, keep, dist):
    _check(ctx, oracle, _random_batch(np.random.default_rng(seed), n), keep, dist)


def test_blocksub_edge_cases(ctx, oracle):
    src, off, ln, cnt = ctx.blocksub(helpers.Batch.from_records([]))
    assert len(src) == 0 and cnt == [0, 0, 0, 0]
    # REF == ALT block: emits nothing; a 5000-base block with every base different; tile-boundary record counts
    long_ref, long_alt = "A" * 5000, "C" * 5000
    b = helpers.Batch.from_records([(0, 5, ["ACGT", "ACGT"]), (0, 9, [long_ref, long_alt]), (0, 9000, ["AC", "CA"])])
    for keep in (False, True):
        _check(ctx, oracle, b, keep, 2)
    for n in (2047, 2048, 2049):
        _check(ctx, oracle, _random_batch(np.random.default_rng(n), n, bs_frac=0.6, maxlen=8), True, 2)


def test_blocksub_fullsize_properties(ctx):
    """1 M records, a tenth of them block substitutions: every output lies inside its source, the differences are conserved"""
    rng = np.random.default_rng(3)
    n = 1_000_000
    is_bs = rng.random(n) < 0.1
    L = np.where(is_bs, rng.integers(2, 9, n), 1)
    aoff = np.zeros(2 * n + 1, np.int64)
    aoff[1::2] = np.cumsum(2 * L) - L
    aoff[2::2] = np.cumsum(2 * L)
    data = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, int(aoff[-1]))]
    b = helpers.Batch(np.arange(1, 20 * n + 1, 20, dtype=np.int32), np.zeros(n, np.int32), np.arange(0, 2 * n + 1, 2, dtype=np.uint32),
                      aoff.astype(np.uint32), data)
    src, off, ln, cnt = ctx.blocksub(b)
    ref_i = np.concatenate([np.arange(aoff[2 * r], aoff[2 * r + 1]) for r in np.nonzero(is_bs)[0][:0]] or [np.zeros(0, np.int64)])
    # differences per record, computed with numpy on the flat byte array
    starts_r, starts_a = aoff[0:-1:2], aoff[1::2]
    diff_total = 0
    for l in range(2, 9):
        rr = np.nonzero(is_bs & (L == l))[0]
        if len(rr):
            R = data[starts_r[rr][:, None] + np.arange(l)]
            A = data[starts_a[rr][:, None] + np.arange(l)]
            diff_total += int((R != A).sum())
    n_pass = int((~is_bs).sum())
    assert cnt == [n, int(is_bs.sum()), diff_total, diff_total + n_pass]
    assert len(src) == diff_total + n_pass
    assert np.all(np.diff(src.astype(np.int64)) >= 0)            # input order
    assert np.all((ln == 0) == (~is_bs[src])) and np.all(off + ln <= np.maximum(L[src], 1))
    snp = ln == 1
    assert np.all(data[starts_r[src[snp]] + off[snp]] != data[starts_a[src[snp]] + off[snp]])


@pytest.mark.parametrize("stem,flags,tag", B.ALIGN_CASES)
def test_blocksub_align_reference_golden(ctx, stem, flags, tag):
    vin, vout, verr = B.paths(stem, tag)
    sites = B.read_sites(vin)
    keep, dist = B.mode(flags)
    batch = B.batch_of(sites)
    src, off, ln, cnt = ctx.blocksub(batch, keep, dist)
    align = ctx.blocksub_align(batch)
    assert B.stable_by_pos(B.apply_plan_align(sites, src, off, ln, align)) == B.read_sites(vout)
    assert B.align_counters(cnt, align) == B.stats_of(open(verr).read())


@pytest.mark.parametrize("seed,n,maxlen", [(11, 4000, 30), (12, 300, 700), (13, 2000, 3)])
def test_blocksub_align_matches_oracle(ctx, oracle, seed, n, maxlen):
    """random REF / ALT pairs of every shape (equal, one base, empty ALT, long, lower case, multiallelic): chunk for chunk equal"""
    rng = np.random.default_rng(seed)
    recs = []
    for k in range(n):
        L = int(rng.integers(1, maxlen + 1))
        ref = "".join(rng.choice(list("ACGT"), L))
        alt = "".join((c if rng.random() < 0.7 else "") + ("" if rng.random() < 0.8 else str(rng.choice(list("ACGTa")))) for c in ref) or "G"
        al = [ref, alt]
        if k % 9 == 0:
            al.append(ref[:1])
        if k % 17 == 0:
            al = [ref]
        recs.append((0, 10 + k, [a.encode() for a in al]))
    b = helpers.Batch.from_records(recs)
    got, exp = ctx.blocksub_align(b), oracle.blocksub_align(b)
    assert got == exp
    assert sum(1 for a in exp if a) > n // 8
    assert ctx.blocksub_align(helpers.Batch.from_records([])) == []
