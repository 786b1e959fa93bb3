"""GPU: vtn_blocksub_noalloc (include/vtn.h) -- the entry point that allocates nothing (scratch in a fixed device array inside the
library) -- through the C ABI against the existing vtn_blocksub on the same batches: every output array and counter bit-exact (the
tolerance of all blocksub tests), on two calls in a row, the second on a smaller batch so that the stale contents of the shared
scratch would show.  Collected last (file name).  Device buffers come from torch (test
plumbing only); the entry point itself takes plain device pointers."""
import ctypes as C

import numpy as np
import pytest

import helpers
from vt_b200 import _lib as L

pytestmark = pytest.mark.gpu
MAX_REC = 1 << 20  # VTN_BLOCKSUB_NOALLOC_MAX_REC


def _batch(rng, n):
    recs, pos = [], 10
    for _ in range(n):
        pos += int(rng.integers(0, 30))
        u = rng.random()
        Ln = int(rng.integers(2, 40))
        ref = "".join(rng.choice(list("ACGTN"), Ln))
        if u < 0.35:
            p = rng.choice([0.0, 0.1, 0.5, 0.9, 1.0])
            al = [ref, "".join(c if rng.random() < p else rng.choice(list("ACGTacgt")) for c in ref)]
        elif u < 0.45:
            al = [ref, ref[::-1], ref[:1]]
        elif u < 0.55:
            al = [ref]
        elif u < 0.7:
            al = [ref, ref[: Ln // 2]]
        else:
            al = [ref[:1], "ACGT"[int(rng.integers(0, 4))]]
        recs.append((0, pos, al))
    return helpers.Batch.from_records(recs)


def _noalloc(ctx, b, keep, dist):
    import torch
    dev = torch.device("cuda", 0)
    cap = int(ctx.lib.vtn_blocksub_out_bound(C.byref(b.as_c())))
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).copy()).to(dev)
    d_ab, d_ao = t(b.allele_beg), t(b.allele_off)
    d_by = torch.zeros(b.n_bytes + 16, dtype=torch.uint8, device=dev)  # readable 16 bytes past the end
    d_by[:b.n_bytes] = torch.from_numpy(b.data.copy()).to(dev)
    d_src, d_off, d_len = (torch.full((max(cap, 1),), 0x7fffffff, dtype=torch.int32, device=dev) for _ in range(3))
    d_n = torch.full((1,), -1, dtype=torch.int64, device=dev)
    d_cnt = torch.full((4,), -1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()  # the library runs on its own stream
    rc = ctx.lib.vtn_blocksub_noalloc(ctx.h, b.n_rec, d_ab.data_ptr(), d_ao.data_ptr(), d_by.data_ptr(), 1 if keep else 0, dist, d_src.data_ptr(),
                                      d_off.data_ptr(), d_len.data_ptr(), cap, d_n.data_ptr(), d_cnt.data_ptr())
    assert rc == 0, ctx.lib.vtn_last_error(ctx.h)
    ctx.sync()
    k = int(d_n.cpu()[0])
    assert 0 <= k <= cap
    u32 = lambda x: x.cpu().numpy().view(np.uint32)
    return u32(d_src)[:k], u32(d_off)[:k], u32(d_len)[:k], [int(x) for x in u32(d_cnt)]


def test_noalloc_matches_vtn_blocksub_twice_in_a_row(ctx):
    rng = np.random.default_rng(123)
    for (n, keep, dist) in ((9000, True, 2), (2500, False, 2), (0, True, 2)):
        b = _batch(rng, n)
        got = _noalloc(ctx, b, keep, dist)
        exp = ctx.blocksub(b, keep, dist)
        for g, e, name in zip(got[:3], exp[:3], ("src_rec", "off", "len")):
            np.testing.assert_array_equal(g, e, err_msg="%s (n=%d)" % (name, n))
        assert got[3] == exp[3], (n, got[3], exp[3])


def test_noalloc_rejects_more_than_the_fixed_scratch(ctx):
    rc = ctx.lib.vtn_blocksub_noalloc(ctx.h, MAX_REC + 1, None, None, None, 0, 2, None, None, None, 0, None, None)
    assert rc == -2  # VTN_E_ARG
    msg = ctx.lib.vtn_last_error(ctx.h).decode()
    assert "VTN_BLOCKSUB_NOALLOC_MAX_REC" in msg and str(MAX_REC) in msg
