"""Sharding of a record batch over the GPUs of one box and the ordered re-merge of the results.
The path has no exchange step: records are independent and the reference is replicated, so each rank normalizes
a contiguous record range cut on contig boundaries where possible; only counters (summed) and the per-rank
outputs (concatenated in rank order, which IS input order) come back.  torch.distributed is used for that
plumbing alone (nccl on the GPUs, gloo in the CPU tests)."""
import numpy as np


def plan_shards(rid, n_ranks, weights=None):
    """Contiguous record ranges [lo, hi) per rank with ~equal total weight (weight 1 per record by default; C4 uses
    an expected-shift estimate).  A cut moves to the nearest contig boundary when one lies within 2% of the ideal
    cut, otherwise it splits the contig (ordering needs no halo: the merge is a stable sort on POS)."""
    n = int(rid.shape[0])
    if n_ranks <= 1 or n == 0:
        return [(0, n)] + [(n, n)] * (max(n_ranks, 1) - 1)
    w = np.ones(n, np.float64) if weights is None else np.asarray(weights, np.float64)
    cum = np.concatenate([[0.0], np.cumsum(w)])
    bounds = np.flatnonzero(np.diff(rid) != 0) + 1  # record indices where a new contig starts
    cuts = [0]
    for r in range(1, n_ranks):
        ideal = int(np.searchsorted(cum, cum[-1] * r / n_ranks))
        ideal = min(max(ideal, cuts[-1]), n)
        if bounds.size:
            k = int(np.argmin(np.abs(bounds - ideal)))
            if abs(int(bounds[k]) - ideal) <= 0.02 * n and int(bounds[k]) >= cuts[-1]:
                ideal = int(bounds[k])
        cuts.append(ideal)
    cuts.append(n)
    return [(cuts[r], cuts[r + 1]) for r in range(n_ranks)]


def slice_batch(batch, lo, hi):
    """Sub-batch of records [lo, hi) (views where possible; offsets rebased)."""
    from .api import Batch
    a0, a1 = int(batch.allele_beg[lo]), int(batch.allele_beg[hi])
    b0, b1 = int(batch.allele_off[a0]), int(batch.allele_off[a1])
    return Batch(batch.pos1[lo:hi].copy(), batch.rid[lo:hi].copy(), (batch.allele_beg[lo:hi + 1] - a0).astype(np.uint32),
                 (batch.allele_off[a0:a1 + 1] - b0).astype(np.uint32), batch.data[b0:b1].copy())


def merge_results(parts):
    """parts: per rank, in rank order, dicts with status, vtype, chg_rec (shard-local), chg_pos1, chg_lt, chg_le,
    chg_rt, chg_off, chg_bytes, counters (12), lo.  -> one dict indexed by global record number."""
    out = {"status": np.concatenate([p["status"] for p in parts]), "vtype": np.concatenate([p["vtype"] for p in parts])}
    for k in ("chg_pos1", "chg_lt", "chg_le", "chg_rt"):
        out[k] = np.concatenate([p[k] for p in parts])
    out["chg_rec"] = np.concatenate([p["chg_rec"].astype(np.int64) + p["lo"] for p in parts])
    offs, base = [], 0
    for p in parts:
        offs.append(p["chg_off"][:-1].astype(np.int64) + base)
        base += int(p["chg_off"][-1])
    out["chg_off"] = np.concatenate(offs + [np.array([base], np.int64)])
    out["chg_bytes"] = np.concatenate([p["chg_bytes"][:int(p["chg_off"][-1])] for p in parts])
    out["counters"] = [int(sum(p["counters"][c] for p in parts)) for c in range(12)]
    return out


def result_part(res, lo):
    """numpy views of a NormResult-like object (vt_b200.api.NormResult or the test oracle's) as a merge part"""
    k = res.n_changed
    n = res.status.shape[0] if hasattr(res.status, "shape") else len(res.status)
    return {"status": np.asarray(res.status[:n]).copy(), "vtype": np.asarray(res.vtype[:n]).copy(),
            "chg_rec": np.asarray(res.chg_rec[:k]).copy(), "chg_pos1": np.asarray(res.chg_pos1[:k]).copy(),
            "chg_lt": np.asarray(res.chg_lt[:k]).copy(), "chg_le": np.asarray(res.chg_le[:k]).copy(),
            "chg_rt": np.asarray(res.chg_rt[:k]).copy(), "chg_off": np.asarray(res.chg_off[:k + 1]).copy(),
            "chg_bytes": np.asarray(res.chg_bytes[:res.n_changed_bytes]).copy(), "counters": list(res.counters), "lo": lo}


def gather_parts(part, dist):
    """all ranks -> rank 0: list of parts in rank order (None elsewhere)"""
    world = dist.get_world_size()
    out = [None] * world if dist.get_rank() == 0 else None
    dist.gather_object(part, out, dst=0)
    return out
