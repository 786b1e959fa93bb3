"""Synthetic GRCh38-shaped references and VCF-shaped record batches.
Tooling for bench.py and the full-size property tests; deterministic from the seeds below."""
import numpy as np

from .api import Batch

# GRCh38 primary assembly lengths chr1..22, X, Y
GRCH38 = [
    ("chr1", 248956422), ("chr2", 242193529), ("chr3", 198295559), ("chr4", 190214555), ("chr5", 181538259),
    ("chr6", 170805979), ("chr7", 159345973), ("chr8", 145138636), ("chr9", 138394717), ("chr10", 133797422),
    ("chr11", 135086622), ("chr12", 133275309), ("chr13", 114364328), ("chr14", 107043718), ("chr15", 101991189),
    ("chr16", 90338345), ("chr17", 83257441), ("chr18", 80373285), ("chr19", 58617616), ("chr20", 64444167),
    ("chr21", 46709983), ("chr22", 50818468), ("chrX", 156040895), ("chrY", 57227415),
]
SEED_REF_C2 = 20260101
SEED_VCF_C2 = 20260102
SEED_VCF_C3 = 20260103
SEED_REF_C4 = 20260104
SEED_VCF_C4 = 20260105

_ACGT = np.frombuffer(b"ACGT", np.uint8)


def _base_lut():
    lut = np.empty(256, np.uint8)  # GC 41%: A/T 29.7% each, C/G 20.3% each (quantised to 1/256)
    lut[:76] = 65
    lut[76:128] = 67
    lut[128:180] = 71
    lut[180:] = 84
    return lut


def make_contig(rng, length, str_frac=0.0, n_telomere=10000, n_block=True, tracts_out=None):
    """One contig: i.i.d. ACGT (GC 41%), N runs at both ends and one 1-3 Mb (scaled) N block; optionally
    str_frac of the contig covered by tandem repeats (C4: unit 1-6, 40% homopolymers, tract 10-2000 bp
    log-uniform, 0-2% interruptions)."""
    lut = _base_lut()
    s = lut[rng.integers(0, 256, size=length, dtype=np.uint8)]
    if str_frac > 0:
        mean_tract = 200.0
        n_tr = int(length * str_frac / mean_tract)
        starts = np.sort(rng.integers(0, max(length - 2100, 1), size=n_tr))
        tl = np.exp(rng.uniform(np.log(10), np.log(2000), size=n_tr)).astype(np.int64)
        ul = np.where(rng.random(n_tr) < 0.4, 1, rng.integers(2, 7, size=n_tr))
        units = _ACGT[rng.integers(0, 4, size=(n_tr, 6))]
        total = int(tl.sum())
        seg = np.repeat(np.arange(n_tr), tl)
        within = np.arange(total) - np.repeat(np.cumsum(tl) - tl, tl)
        vals = units[seg, within % ul[seg]]
        inter = rng.random(total) < np.repeat(rng.uniform(0, 0.02, size=n_tr), tl)
        vals = np.where(inter, _ACGT[rng.integers(0, 4, size=total)], vals)
        idx = np.minimum(np.repeat(starts, tl) + within, length - 1)
        s[idx] = vals
        if tracts_out is not None:
            tracts_out.append((starts.astype(np.int64), tl, ul.astype(np.int64)))
    t = min(n_telomere, length // 20)
    s[:t] = ord("N")
    s[length - t:] = ord("N")
    if n_block and length > 1000:
        bl = int(rng.integers(1_000_000, 3_000_001) * (length / 100e6)) if length < 40_000_000 else int(rng.integers(1_000_000, 3_000_001))
        bl = min(bl, length // 10)
        b0 = int(rng.integers(length // 3, length // 2))
        s[b0:b0 + bl] = ord("N")
    return s


def make_reference(seed=SEED_REF_C2, scale=1.0, str_frac=0.0, contigs=None, tracts_out=None):
    """-> (names, [uint8 arrays]).  scale < 1 shrinks every contig (tests).  tracts_out (a list) receives one
    (starts, lengths, unit_lengths) triple per contig when str_frac > 0 (C4)."""
    names, seqs = [], []
    for i, (name, ln) in enumerate(contigs or GRCH38):
        rng = np.random.default_rng([seed, i])
        names.append(name)
        seqs.append(make_contig(rng, max(int(ln * scale), 1000), str_frac=str_frac, tracts_out=tracts_out))
    return names, seqs


def _gather(seqs, base, gpos):
    """global position -> base, contig by contig (keeps peak memory low)"""
    out = np.empty(gpos.shape[0], np.uint8)
    c = np.searchsorted(base, gpos, side="right") - 1
    for ci in np.unique(c):
        m = c == ci
        out[m] = seqs[ci][gpos[m] - base[ci]]
    return out


def _ragged(lens):
    """(segment id, index within segment) for segments of the given lengths"""
    seg = np.repeat(np.arange(lens.shape[0]), lens)
    within = np.arange(int(lens.sum())) - np.repeat(np.cumsum(lens) - lens, lens)
    return seg, within


def make_str_indels(seqs, tracts, n, seed=SEED_VCF_C4):
    """C4 records: each an insertion/deletion of k in 1..4 whole repeat units, 90% at the RIGHTMOST
    copy of a tandem-repeat tract (maximal left shift), 5% mid-tract, 5% outside tracts.  Both alleles are
    reference substrings: deletion REF=ref[a..a+D] ALT=ref[a]; insertion REF=ref[a] ALT=ref[a]+ref[a+1-D..a]."""
    rng = np.random.default_rng(seed)
    lens = np.array([s.shape[0] for s in seqs], np.int64)
    base = np.concatenate([[0], np.cumsum(lens)])[:-1]
    n_tr = np.array([t[0].shape[0] for t in tracts], np.int64)
    tcum = np.concatenate([[0], np.cumsum(n_tr)])
    starts = np.concatenate([t[0] for t in tracts])
    tls = np.concatenate([t[1] for t in tracts])
    uls = np.concatenate([t[2] for t in tracts])
    tcontig = np.repeat(np.arange(len(tracts)), n_tr)
    pick = rng.integers(0, tcum[-1], size=n)
    c = tcontig[pick]
    ul = uls[pick]
    k = rng.integers(1, 5, size=n)
    D = np.minimum(k * ul, np.maximum(tls[pick] - 2, 1))
    e = np.minimum(starts[pick] + tls[pick], lens[c] - 64)          # exclusive end of the tract
    where = rng.random(n)
    e = np.where(where < 0.05, e - tls[pick] // 2, e)                # mid-tract
    e = np.where(where > 0.95, e + rng.integers(60, 5000, size=n), e)  # outside
    e = np.clip(e, 128, lens[c] - 64)
    is_ins = rng.random(n) < 0.5
    a = np.where(is_ins, e - 1, e - D - 1)                            # 0-based anchor
    # no variants inside / next to the N blocks (a run of N is a homopolymer millions of bases long: real call sets
    # have no records there, and the per-base reference loop would never finish on them)
    g_chk = base[c] + a
    ok = (_gather(seqs, base, g_chk) != 78) & (_gather(seqs, base, g_chk - 100) != 78) & (_gather(seqs, base, g_chk + 60) != 78)
    c, a, D, is_ins = c[ok], a[ok], D[ok], is_ins[ok]
    n = int(c.shape[0])
    o = np.lexsort((a, c))
    c, a, D, is_ins = c[o], a[o], D[o], is_ins[o]
    ref_len = np.where(is_ins, 1, 1 + D)
    alt_len = np.where(is_ins, 1 + D, 1)
    tot = ref_len + alt_len
    off = np.concatenate([[0], np.cumsum(tot)])
    data = np.empty(int(off[-1]), np.uint8)
    g0 = base[c] + a
    seg, within = _ragged(ref_len)
    data[np.repeat(off[:-1], ref_len) + within] = _gather(seqs, base, g0[seg] + within)
    seg, within = _ragged(alt_len)
    src = np.where(within == 0, 0, within - D[seg])                  # insertion body = the D bases ending at the anchor
    data[np.repeat(off[:-1] + ref_len, alt_len) + within] = _gather(seqs, base, g0[seg] + src)
    allele_off = np.empty(2 * n + 1, np.uint32)
    allele_off[0:2 * n:2] = off[:-1]
    allele_off[1:2 * n:2] = off[:-1] + ref_len
    allele_off[2 * n] = off[-1]
    return Batch((a + 1).astype(np.int32), c.astype(np.int32), (np.arange(n + 1) * 2).astype(np.uint32), allele_off, data)


def _sample_positions(rng, seqs, n, margin=64):
    """n positions (contig, 0-based p0), density proportional to contig length, sorted, avoiding N runs."""
    lens = np.array([s.shape[0] for s in seqs], np.int64)
    cum = np.concatenate([[0], np.cumsum(lens)])
    out_c, out_p = [], []
    need = n
    while need > 0:
        g = np.sort(rng.integers(0, cum[-1], size=int(need * 1.15) + 16))
        c = np.searchsorted(cum, g, side="right") - 1
        p = g - cum[c]
        ok = (p >= margin) & (p < lens[c] - margin)
        c, p = c[ok], p[ok]
        keep = np.ones(c.shape[0], bool)
        for ci in np.unique(c):
            m = c == ci
            s = seqs[ci]
            pp = p[m]
            keep[m] = (s[pp] != 78) & (s[pp - margin + 1] != 78) & (s[pp + margin - 1] != 78)
        out_c.append(c[keep])
        out_p.append(p[keep])
        need -= int(keep.sum())
    c = np.concatenate(out_c)
    p = np.concatenate(out_p)
    if c.shape[0] > n:
        sel = np.sort(rng.choice(c.shape[0], size=n, replace=False))
        c, p = c[sel], p[sel]
    o = np.lexsort((p, c))
    return c[o].astype(np.int32), p[o].astype(np.int64)


def make_biallelic(seqs, n, seed=SEED_VCF_C2, snv_frac=0.8, unnormalized_frac=0.5, max_indel=50, rightmost_in_tract=False):
    """C2-style records: snv_frac SNVs, the rest half deletions / half insertions with geometric(0.4) length
    capped at max_indel; REF always from the reference; unnormalized_frac of the indels carry 1-3 shared
    trailing bases (right-trim needed, and a left shift when the indel sits in a repeat)."""
    rng = np.random.default_rng(seed)
    rid, p0 = _sample_positions(rng, seqs, n)
    lens = np.array([s.shape[0] for s in seqs], np.int64)
    base = np.concatenate([[0], np.cumsum(lens)])[:-1]
    u = rng.random(n)
    kind = np.where(u < snv_frac, 0, np.where(u < snv_frac + (1 - snv_frac) / 2, 1, 2))  # 0 snv 1 del 2 ins
    L = np.minimum(rng.geometric(0.4, size=n), max_indel).astype(np.int64)
    L[kind == 0] = 0
    pad = np.where((kind != 0) & (rng.random(n) < unnormalized_frac), rng.integers(1, 4, size=n), 0).astype(np.int64)
    ins_copy = rng.random(n) < 0.3  # inserted bases copy the following reference (shiftable)
    # allele lengths
    ref_len = np.where(kind == 1, 1 + L + pad, 1 + pad)
    alt_len = np.where(kind == 2, 1 + L + pad, 1 + pad)
    alt_len[kind == 0] = 1
    ref_len[kind == 0] = 1
    tot = ref_len + alt_len
    off = np.concatenate([[0], np.cumsum(tot)])
    data = np.empty(int(off[-1]), np.uint8)

    def gather(gpos):
        # global position -> base, contig by contig (keeps peak memory low)
        out = np.empty(gpos.shape[0], np.uint8)
        c = np.searchsorted(base, gpos, side="right") - 1
        for ci in np.unique(c):
            m = c == ci
            out[m] = seqs[ci][gpos[m] - base[ci]]
        return out

    g0 = base[rid] + p0
    # REF = ref[p0 : p0+ref_len]
    seg = np.repeat(np.arange(n), ref_len)
    within = np.arange(int(ref_len.sum())) - np.repeat(np.cumsum(ref_len) - ref_len, ref_len)
    dst = np.repeat(off[:-1], ref_len) + within
    data[dst] = gather(g0[seg] + within)
    # ALT
    seg = np.repeat(np.arange(n), alt_len)
    within = np.arange(int(alt_len.sum())) - np.repeat(np.cumsum(alt_len) - alt_len, alt_len)
    dst = np.repeat(off[:-1] + ref_len, alt_len) + within
    k = kind[seg]
    Ls = L[seg]
    # deletions: alt[j] = ref[p0 + (j==0 ? 0 : j + L)] ; insertions: j==0 anchor, 1..L inserted, then ref[p0 + j - L]
    src = np.where(k == 1, np.where(within == 0, 0, within + Ls), np.where((k == 2) & (within > Ls), within - Ls, within))
    vals = gather(g0[seg] + src)
    is_ins_body = (k == 2) & (within >= 1) & (within <= Ls)
    lit = _ACGT[rng.integers(0, 4, size=vals.shape[0])]
    vals = np.where(is_ins_body & ~ins_copy[seg], lit, vals)
    # SNV alt: a different base
    snv = k == 0
    code = np.searchsorted(_ACGT, vals)  # A0 C1 G2 T3 (vals are ACGT here)
    vals = np.where(snv, _ACGT[(code + 1 + rng.integers(0, 3, size=vals.shape[0])) % 4], vals)
    data[dst] = vals
    allele_off = np.empty(2 * n + 1, np.uint32)
    allele_off[0:2 * n:2] = off[:-1]
    allele_off[1:2 * n:2] = off[:-1] + ref_len
    allele_off[2 * n] = off[-1]
    allele_beg = (np.arange(n + 1) * 2).astype(np.uint32)
    return Batch((p0 + 1).astype(np.int32), rid, allele_beg, allele_off, data)


def algorithmic_bytes(batch, res):
    """SURVEY.md 8d canonical bytes for a biallelic batch given its normalize result (numpy, exact):
    16 in-hdr + allele bytes + ceil((C*rl + S)/4) packed reference + 16 out-hdr + 12*ch + ch*(rl'+al')."""
    n = batch.n_rec
    lens = np.diff(batch.allele_off.astype(np.int64))
    na = np.diff(batch.allele_beg.astype(np.int64))
    tot_allele = np.add.reduceat(lens, batch.allele_beg[:-1].astype(np.int64)) if n else np.zeros(0, np.int64)
    rl = lens[batch.allele_beg[:-1].astype(np.int64)] if n else np.zeros(0, np.int64)
    C = (res.vtype[:n] != 1).astype(np.int64)
    S = np.zeros(n, np.int64)
    k = res.n_changed
    S[res.chg_rec[:k]] = res.chg_le[:k]
    total = 16 * n + int(tot_allele.sum()) + int(((C * rl + S + 3) // 4).sum()) + 16 * n + 12 * k + res.n_changed_bytes
    total += int((2 * (na - 2)).sum())  # multiallelic input header: 16 + 2n
    return int(total)


SEED_C5 = 20260106


def make_multiallelic(seqs, n, seed=SEED_VCF_C3, max_indel=50):
    """C3 records: every record multiallelic, n_alt = clip(1 + geometric(0.5), 2, 6); each ALT is an
    SNV (80%), deletion (10%) or insertion (10%) with geometric(0.4) length capped at max_indel, all sharing one REF
    that spans the longest deletion: SNV = REF with a different first base, deletion of L = REF[0] + REF[1+L:],
    insertion of L = REF[0] + L literal bases + REF[1:]."""
    rng = np.random.default_rng(seed)
    rid, p0 = _sample_positions(rng, seqs, n, margin=max_indel + 8)
    lens = np.array([s.shape[0] for s in seqs], np.int64)
    base = np.concatenate([[0], np.cumsum(lens)])[:-1]
    n_alt = np.clip(1 + rng.geometric(0.5, size=n), 2, 6).astype(np.int64)
    A = int(n_alt.sum())
    arec = np.repeat(np.arange(n), n_alt)
    aidx = np.arange(A) - np.repeat(np.cumsum(n_alt) - n_alt, n_alt)
    u = rng.random(A)
    akind = np.where(u < 0.8, 0, np.where(u < 0.9, 1, 2))  # 0 snv 1 del 2 ins
    aL = np.minimum(rng.geometric(0.4, size=A), max_indel).astype(np.int64)
    aL[akind == 0] = 0
    dmax = np.maximum.reduceat(np.where(akind == 1, aL, 0), np.cumsum(n_alt) - n_alt)
    ref_len = 1 + dmax
    alt_len = np.where(akind == 0, ref_len[arec], np.where(akind == 1, ref_len[arec] - aL, ref_len[arec] + aL))
    # allele slots in record order: REF, ALT1..ALTk
    n_slot = n + A
    slot_of_ref = np.cumsum(np.concatenate([[0], 1 + n_alt]))[:-1]
    slot_of_alt = slot_of_ref[arec] + 1 + aidx
    slen = np.empty(n_slot, np.int64)
    slen[slot_of_ref] = ref_len
    slen[slot_of_alt] = alt_len
    skind = np.full(n_slot, -1, np.int64)
    skind[slot_of_alt] = akind
    sL = np.zeros(n_slot, np.int64)
    sL[slot_of_alt] = aL
    srec = np.empty(n_slot, np.int64)
    srec[slot_of_ref] = np.arange(n)
    srec[slot_of_alt] = arec
    sidx = np.zeros(n_slot, np.int64)
    sidx[slot_of_alt] = aidx
    off = np.concatenate([[0], np.cumsum(slen)])
    seg, within = _ragged(slen)
    k, L = skind[seg], sL[seg]
    src = np.where(k == 1, np.where(within == 0, 0, within + L), np.where((k == 2) & (within > L), within - L, within))
    g0 = base[rid] + p0
    vals = _gather(seqs, base, g0[srec[seg]] + src)
    body = (k == 2) & (within >= 1) & (within <= L)
    vals = np.where(body, _ACGT[rng.integers(0, 4, size=vals.shape[0])], vals)
    snv0 = (k == 0) & (within == 0)
    code = np.searchsorted(_ACGT, np.where(vals == 78, 65, vals))
    vals = np.where(snv0, _ACGT[(code + 1 + sidx[seg] % 3) % 4], vals)
    allele_beg = np.concatenate([slot_of_ref, [n_slot]]).astype(np.uint32)
    return Batch((p0 + 1).astype(np.int32), rid, allele_beg, off.astype(np.uint32), vals.astype(np.uint8))
