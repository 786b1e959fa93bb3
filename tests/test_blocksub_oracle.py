"""CPU: the oracle's restatement of vt decompose_blocksub (oracle/vt_oracle.c vto_blocksub_plan; decompose_blocksub.cpp:185-421
without -a) against the reference's own fixtures (01 default, 03 -p) and against outputs of the unmodified reference binary on the
test_mnv.sh inputs and a synthetic input, default / -m / -m -d N (tests/golden/blocksub_*, made by make_blocksub_golden.py)."""
import pytest

import blocksub_helpers as B
import helpers


@pytest.mark.parametrize("stem,flags,tag", B.CASES)
def test_blocksub_oracle_matches_reference(stem, flags, tag):
    vin, vout, verr = B.paths(stem, tag)
    sites = B.read_sites(vin)
    keep, dist = B.mode(flags)
    src, off, ln, cnt = helpers.Oracle().blocksub_plan(B.batch_of(sites), keep, dist)
    # the reference writes through a 1000-bp ordered writer: positions of one input record's outputs are increasing and the
    # inputs are sorted, so the file order is the stable sort of the plan by position
    assert B.stable_by_pos(B.apply_plan(sites, src, off, ln)) == B.read_sites(vout)
    assert cnt == B.stats_of(open(verr).read())


@pytest.mark.parametrize("stem,flags,tag", B.ALIGN_CASES)
def test_blocksub_align_oracle_matches_reference(stem, flags, tag):
    """vto_blocksub_align (needle.cpp + decompose_blocksub.cpp:190-303 restated) against the reference's fixture 02 and vt_ref -a outputs"""
    vin, vout, verr = B.paths(stem, tag)
    sites = B.read_sites(vin)
    keep, dist = B.mode(flags)
    batch = B.batch_of(sites)
    o = helpers.Oracle()
    src, off, ln, cnt = o.blocksub_plan(batch, keep, dist)
    align = o.blocksub_align(batch)
    assert sum(1 for a in align if a) > 0
    assert B.stable_by_pos(B.apply_plan_align(sites, src, off, ln, align)) == B.read_sites(vout)
    assert B.align_counters(cnt, align) == B.stats_of(open(verr).read())
