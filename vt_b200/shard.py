"""Process-per-GPU plumbing around the same sharding that vtn_multi_* does inside one process (vt_b200/csrc/vtn_multi.cu, SURVEY.md
8e): each rank takes the record range vtn_shard_plan gives it; only counters (summed) and the per-rank outputs (concatenated in
rank order, which IS input order because the ranges are contiguous) come back.  torch.distributed is used for that plumbing alone
(nccl on the GPUs, gloo in the CPU tests).  The product path for several GPUs is the C ABI: vtn_multi_normalize_bi /
vtn_multi_decompose_normalize."""
import numpy as np


def plan_shards(rid, n_ranks):
    """Contiguous record ranges [lo, hi) per rank from libvtn's planner (vtn_shard_plan, the one vtn_multi_* uses inside one
    process): equal record counts, a cut within a sixteenth of a shard of a contig change moves onto it, otherwise the contig is
    split.  Ranges are contiguous and in rank order, so putting results back together is concatenation in rank order."""
    from .api import shard_plan
    n = max(int(n_ranks), 1)
    cut = shard_plan(np.ascontiguousarray(rid, np.int32), n)
    return [(int(cut[r]), int(cut[r + 1])) for r in range(n)]


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
