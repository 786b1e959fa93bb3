#!/usr/bin/env python3
"""bench.py -- vt normalize hot path on B200 (BASELINE.json metric: M records/s decompose+normalize).

A "step" = one pass of the normalize path over one batch of 10 M synthetic biallelic SNV+indel records
(config C2) against the synthetic 3.1 Gb GRCh38-shaped reference resident in HBM.
  value : records/s with the batch already resident in HBM (CUDA events on the library's stream)
  e2e   : the same through the reference-facing C-ABI call vtn_normalize() with pinned HOST buffers,
          H2D + kernel + D2H inside the timed region
  roofline / cpu_baseline : see DESIGN.md "Measurement"
  --impl reference : the UNMODIFIED reference `vt normalize` (oracle/_ref/vt_ref) on the host cores.
Launch: python bench.py --gpus N --steps K --warmup W   (N>1: torchrun, one rank per GPU, weak scaling)
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "normalize_records_per_s"
UNIT = "M records/s"
N_RECORDS = int(os.environ.get("VTN_BENCH_RECORDS", 10_000_000))
REF_SCALE = float(os.environ.get("VTN_BENCH_REF_SCALE", 1.0))
CPU_SAMPLE_RECORDS = int(os.environ.get("VTN_BENCH_CPU_SAMPLE", 4_000_000))
CHECK_RECORDS = int(os.environ.get("VTN_BENCH_CHECK_RECORDS", 100_000))
VT_REF = os.path.join(ROOT, "oracle", "_ref", "vt_ref")


# The contract is ONE JSON line on stdout.  Libraries write to fd 1 behind Python's back (NCCL prints its version
# banner there under torchrun), so keep a private handle to the real stdout and point fd 1 at stderr.
_REAL_STDOUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def emit(line):
    _REAL_STDOUT.write(json.dumps(line) + "\n")
    _REAL_STDOUT.flush()


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.p, self.t = [], None, None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "20"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.p.terminate()
        sm, mx, reasons = [], None, set()
        rows = [r for (t, r) in self.rows if t0 <= t <= t1 + 0.05] or [r for (_, r) in self.rows[-3:]]
        for r in rows:
            f = [x.strip() for x in r.split(",")]
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def write_vcf(path, names, seqs, batch, lo, hi):
    """records [lo, hi) of a biallelic batch as a sites-only VCF (vectorised: one bytes join per column)"""
    with open(path, "wb") as f:
        f.write(b"##fileformat=VCFv4.2\n")
        for n, s in zip(names, seqs):
            f.write(("##contig=<ID=%s,length=%d>\n" % (n, s.shape[0])).encode())
        f.write(b"#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n")
        data = batch.data.tobytes()
        ao = batch.allele_off
        bnames = [n.encode() for n in names]
        rid, pos = batch.rid, batch.pos1
        CH = 200000
        for c0 in range(lo, hi, CH):
            c1 = min(hi, c0 + CH)
            o = ao[2 * c0:2 * c1 + 1].tolist()
            rr = rid[c0:c1].tolist()
            pp = pos[c0:c1].tolist()
            f.write(b"".join(b"%s\t%d\t.\t%s\t%s\t.\t.\t.\n" % (bnames[rr[i]], pp[i], data[o[2 * i]:o[2 * i + 1]], data[o[2 * i + 1]:o[2 * i + 2]])
                             for i in range(c1 - c0)))


def _run_procs(cmds):
    t0 = time.perf_counter()
    procs = [subprocess.Popen(c, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL) for c in cmds]
    rcs = [p.wait() for p in procs]
    dt = time.perf_counter() - t0
    if any(rcs):
        raise RuntimeError("reference vt failed: %s" % rcs)
    return dt


CLI_RECORDS = int(os.environ.get("VTN_BENCH_CLI_RECORDS", 4_000_000))
VTN_CLI = os.path.join(ROOT, "vt_b200", "vtn")


def cli_wall(tmp, fa, names, seqs, batch, n):
    """The I/O-bound end-to-end wall, parse -> write, on ONE file for both: `vtn normalize` (this repo's command, GPU) and the
    unmodified `vt normalize` (one process, its native mode), plus `vt view` (parse + write floor); outputs compared byte for byte
    in the same run.  VCF text and FASTA on tmpfs."""
    import filecmp
    vcf = os.path.join(tmp, "cli_in.vcf")
    write_vcf(vcf, names, seqs, batch, 0, n)
    o_vtn, o_ref = os.path.join(tmp, "cli_vtn.vcf"), os.path.join(tmp, "cli_ref.vcf")
    walls, phases = [], {}
    for rep in range(4):  # CUDA context creation alone varies between 0.25 and 2 s from run to run on these boxes: best of four, its phases
        t0 = time.perf_counter()
        p = subprocess.run([VTN_CLI, "normalize", vcf, "-r", fa, "-o", o_vtn, "-q"], stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True,
                           env=dict(os.environ, VTN_CLI_TIMING="1"))
        walls.append(time.perf_counter() - t0)
        if p.returncode not in (0, 1):
            raise RuntimeError("vtn normalize failed (%d): %s" % (p.returncode, p.stderr[-500:]))
        if walls[-1] == min(walls):
            phases = {l.split("]")[1].rsplit(None, 2)[0].strip(): float(l.split()[-2]) for l in p.stderr.splitlines() if l.startswith("[vtn timing]")}
    t_ref = _run_procs([[VT_REF, "normalize", vcf, "-r", fa, "-o", o_ref, "-q"]])
    t_view = _run_procs([[VT_REF, "view", vcf, "-o", os.path.join(tmp, "cli_view.vcf")]])
    same = filecmp.cmp(o_vtn, o_ref, shallow=False)
    w = min(walls)
    # the part of both walls that does not depend on the file: a one-record input (process start, CUDA context / index load)
    one = os.path.join(tmp, "cli_one.vcf")
    write_vcf(one, names, seqs, batch, 0, 1)
    f_vtn = min(_run_procs([[VTN_CLI, "normalize", one, "-r", fa, "-o", os.path.join(tmp, "one_vtn.vcf"), "-q"]]) for _ in range(4))
    f_ref = min(_run_procs([[VT_REF, "normalize", one, "-r", fa, "-o", os.path.join(tmp, "one_ref.vcf"), "-q"]]) for _ in range(2))
    fixed = {"vtn_fixed_s": f_vtn, "vt_fixed_s": f_ref,
             "vtn_marginal_m_records_per_s": n / max(w - f_vtn, 1e-9) / 1e6, "vt_marginal_m_records_per_s": n / max(t_ref - f_ref, 1e-9) / 1e6,
             "marginal_speedup": (t_ref - f_ref) / max(w - f_vtn, 1e-9),
             "fixed_note": "fixed = wall on a one-record file (for vtn: process start + CUDA context creation + the first contig made resident); marginal = (wall - fixed) per record"}
    return {**fixed, "what": "parse -> write wall of `vtn normalize` vs the unmodified `vt normalize` (one process) on the SAME %d-record VCF (first records of the C2 batch) and the same %.2f Gb FASTA, both on tmpfs" % (n, sum(x.shape[0] for x in seqs) / 1e9),
            "records": n, "vtn_wall_s": w, "vtn_wall_all_runs_s": [round(x, 3) for x in walls], "vtn_m_records_per_s": n / w / 1e6, "vtn_phases_s": phases,
            "vt_wall_s": t_ref, "vt_m_records_per_s": n / t_ref / 1e6, "vt_view_wall_s": t_view, "speedup": t_ref / w, "outputs_identical": bool(same)}


def cpu_reference_run(n_sample, steps, warmup, names=None, seqs=None, batch=None, cli_out=None):
    """Times the unmodified reference `vt normalize` (oracle/_ref/vt_ref) on a bounded sample OF THE SAME WORKLOAD: the first
    n_sample records of the C2 batch (sorted by contig, position) against the same full synthetic reference as a FASTA + .fai
    on tmpfs, cut into P = host cores consecutive range shards, one single-threaded `vt normalize` process per shard with its
    own output file (vt's own way to go parallel is one process per interval, program.cpp:152-190);
    throughput = n_sample / slowest wall.  Also: one process alone on one shard (its native mode), and `vt view` on that
    shard (the parse + write floor)."""
    from vt_b200 import synth
    import helpers
    if not os.path.exists(VT_REF):
        raise RuntimeError("oracle/_ref/vt_ref missing: run __graft_entry__.build() in the build container")
    if ORIG_AFFINITY is not None:
        os.sched_setaffinity(0, ORIG_AFFINITY)  # the GPU legs ran on the GPU's NUMA node; the reference gets every core of the box
    cores = len(os.sched_getaffinity(0))
    P = max(1, min(cores, 64))
    if seqs is None:
        names, seqs, batch = load_workload("c2", 0)
    n_sample = min(n_sample, batch.n_rec)
    tmp = tempfile.mkdtemp(prefix="vtn_cpu_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    try:
        fa = os.path.join(tmp, "ref.fa")
        helpers.write_fasta(fa, names, seqs)
        cuts = [n_sample * i // P for i in range(P + 1)]
        for i in range(P):
            write_vcf(os.path.join(tmp, "in%d.vcf" % i), names, seqs, batch, cuts[i], cuts[i + 1])
        cmds = [[VT_REF, "normalize", os.path.join(tmp, "in%d.vcf" % i), "-r", fa, "-o", os.path.join(tmp, "out%d.vcf" % i), "-q"] for i in range(P)]
        times = []
        for it in range(warmup + steps):
            dt = _run_procs(cmds)
            if it >= warmup:
                times.append(dt)
        one = _run_procs(cmds[:1])
        view = _run_procs([[VT_REF, "view", os.path.join(tmp, "in0.vcf"), "-o", os.path.join(tmp, "view0.vcf")]])
        if cli_out is not None:
            try:
                cli_out.update(cli_wall(tmp, fa, names, seqs, batch, min(n_sample, CLI_RECORDS)))
            except Exception as e:
                cli_out["error"] = str(e)[:400]
    finally:
        import shutil
        shutil.rmtree(tmp, ignore_errors=True)
    ms = 1e3 * float(np.mean(times))
    val = n_sample / (ms / 1e3) / 1e6
    n0 = cuts[1] - cuts[0]
    sample = ("the first %d records of the C2 batch (80%% SNV/10%% del/10%% ins) against the same %.2f Gb synthetic reference (FASTA + .fai on tmpfs), "
              "cut into %d consecutive range shards, one single-threaded unmodified `vt normalize` process per shard, VCF text in and out, wall = slowest; "
              "one process alone on one shard: %.3f M records/s; `vt view` on that shard (parse + write floor): %.3f M records/s"
              % (n_sample, sum(s.shape[0] for s in seqs) / 1e9, P, n0 / one / 1e6, n0 / view / 1e6))
    return val, ms, P, sample


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    val, ms, P, sample = cpu_reference_run(CPU_SAMPLE_RECORDS, args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": "C2: 10M synthetic biallelic SNV+indel records, synthetic 3.1 Gb GRCh38-shaped reference (bounded sample, see cpu_baseline.sample)"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": P, "kind": "reference", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


C5_SITES = int(os.environ.get("VTN_BENCH_C5_SITES", 20000))
C5_CPU_SAMPLE = int(os.environ.get("VTN_BENCH_C5_CPU_SAMPLE", 800))
C5_REF_SCALE = float(os.environ.get("VTN_BENCH_C5_REF_SCALE", 0.02))
# warp instructions per reference proc_comp in the DP inner loop, counted from the SASS of vtn_hmm_kernel
# (profiles/r1h_hmm_sass_count.txt, dynamic count profiles/r1h_ncu_hmm.txt): the issue-rate roofline of SURVEY.md 8d
HMM_WARP_INSTR_PER_PROC_COMP = float(os.environ.get("VTN_HMM_IPP", 7.65))


def normalized_unique(batch, res):
    """the batch after vt normalize | vt uniq, sorted: (rid, pos1, [alleles]) records (tooling for C5, not timed)"""
    k = res.n_changed
    new = {}
    for j in range(k):
        r = int(res.chg_rec[j])
        new[r] = (int(res.chg_pos1[j]), res.changed_alleles(j, [len(a) for a in batch.alleles(r)]))
    out = set()
    for r in range(batch.n_rec):
        p, al = new.get(r, (int(batch.pos1[r]), batch.alleles(r)))
        out.add((int(batch.rid[r]), p, tuple(al)))
    return sorted(out)


def run_c5(args, rank, world, local):
    """C5: `vt annotate_indels` on synthetic STR indels (C4's generator, normalized first) through vtn_annotate_indels -- the pinned
    driver (tests/test_annot_*: identical alignment tuples and INFO values to the unmodified reference) decides which (model,
    sequence) pairs exist; value = all of them re-timed as one resident batch on the HMM kernel; e2e = the whole call."""
    import torch
    import torch.distributed as dist
    from vt_b200 import Context, synth, api as A
    from vt_b200.api import Batch
    import helpers
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = Context(local)
    t_setup = time.time()
    tracts = []
    names, seqs = synth.make_reference(seed=synth.SEED_REF_C4, scale=C5_REF_SCALE, str_frac=0.3, tracts_out=tracts)
    raw = synth.make_str_indels(seqs, tracts, C5_SITES, seed=synth.SEED_C5 + 1000 * rank)
    ctx.load_reference(names, seqs)
    recs = normalized_unique(raw, ctx.normalize(raw))
    batch = Batch.from_records(recs)
    tmpd = tempfile.mkdtemp(prefix="vtn_c5_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    trace = os.path.join(tmpd, "trace.txt")
    res = ctx.annotate_indels(batch, trace=trace, raw=True)
    n_indel = int(res["n_indel"])
    cases = []
    kind_of = {"A": A.AHMM, "L": A.LFHMM, "R": A.RFHMM}
    for l in open(trace):
        f = l.split()
        cases.append((kind_of[f[2]], f[3], "" if f[4] == "-" else f[4], f[5]))
    os.remove(trace)
    n_ahmm = sum(1 for c in cases if c[0] == A.AHMM)
    hb = A.HmmBatch(cases)
    log("[rank %d] C5 setup %.1fs: %d records (%d indels annotated) -> %d alignments (%d AHMM), %d sequence bytes" %
        (rank, time.time() - t_setup, batch.n_rec, n_indel, hb.n, n_ahmm, hb.n_bytes))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.sync()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    db = ctx.hmm_upload(hb)
    cells, pcs = ctx.hmm_work(db)
    for _ in range(args.warmup):
        ctx.hmm_align_resident(db)
    ctx.sync()
    sampler = ClockSampler(local)
    barrier()
    l0 = ctx.launch_count()
    t0 = time.time()
    ctx.timer_start()
    for _ in range(args.steps):
        ctx.hmm_align_resident(db)
    ms_total = ctx.timer_stop()
    barrier()
    launches = ctx.launch_count() - l0
    ms_step = max_over_ranks(ms_total / args.steps)
    value = world * hb.n / (ms_step / 1e3) / 1e6
    # end to end: the whole vtn_annotate_indels call (host buffers in, 34 values per indel out; regions, fetches, every round)
    e2e_steps = max(2, min(args.steps, 3))
    barrier()
    w0 = time.perf_counter()
    for _ in range(e2e_steps):
        r2 = ctx.annotate_indels(batch, raw=True)
    barrier()
    e2e_ms = max_over_ranks(1e3 * (time.perf_counter() - w0) / e2e_steps)
    t1 = time.time()
    clocks = sampler.stop(t0, t1)
    assert int(r2["n_indel"]) == n_indel and int(r2["n_alignments"]) == hb.n
    ctx.hmm_free(db)
    if rank != 0:
        return
    sm_mhz = clocks.get("sm_mhz") or 1965.0
    peak_pc = 148 * 4 * 32 * sm_mhz * 1e6 / HMM_WARP_INSTR_PER_PROC_COMP / 1e9  # G proc_comp/s at full issue rate
    ach_pc = pcs / (ms_total / args.steps / 1e3) / 1e9
    h2d = 8 * batch.n_rec + 4 * (batch.n_rec + 1) + 4 * (batch.n_allele + 1) + batch.n_bytes + hb.n + 4 * (3 * hb.n + 1) + hb.n_bytes
    d2h = 96 * hb.n + 160 * n_indel
    line = {
        "metric": "hmm_alignments_per_s", "value": value, "unit": "M alignments/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {
            "workload": "C5: vt annotate_indels on %d normalized synthetic STR indels per GPU (C4 generator on a %.3f Gb reference with 30%% tandem repeats; sample of the 5M-indel config) -> %d AHMM/LFHMM/RFHMM alignments issued by the FlankDetector rounds of vtn_annotate_indels (%d RFHMM/LFHMM + %d AHMM), reads up to 1006 bp" % (n_indel, sum(x.shape[0] for x in seqs) / 1e9, hb.n, hb.n - n_ahmm, n_ahmm),
            "indels_per_gpu": n_indel, "alignments_per_gpu": hb.n, "dp_cells_per_step": int(cells), "reference_proc_comps_per_step": int(pcs),
            "g_cell_updates_per_s": world * cells / (ms_step / 1e3) / 1e9,
            "annotate_indels_e2e_indels_per_s": world * n_indel / (e2e_ms / 1e3),
            "l2": "back-pointer scratch of the resident warps (>= 148*16 warps x up to 1.1 MB) exceeds the 126 MB L2 for long reads; no explicit flush",
            "sharding": "independent records per rank, no collective",
            "tolerance": "optimal_score bit-identical to the reference's float32 (stated tolerance 1e-4 abs), paths and integer results exact; the driver issues the reference's exact alignment list (tests/test_annot_gpu.py)",
        },
        "roofline": {
            "bound": "issue", "achieved": ach_pc, "peak": peak_pc, "unit": "G proc_comp/s", "frac": ach_pc / peak_pc,
            "traffic": None, "kernel": "vtn_hmm_kernel",
            "peak_source": "148 SM x 4 sub-cores x 32 lanes x %.0f MHz (median SM clock sampled under load) / %.2f warp instructions per reference proc_comp (SASS count, profiles/r1h_hmm_sass_count.txt): issue-slot utilisation against the kernel's own instruction mix, not an algorithmic bound" % (sm_mhz, HMM_WARP_INSTR_PER_PROC_COMP),
        },
        "e2e": {"value": world * hb.n / (e2e_ms / 1e3) / 1e6, "unit": "M alignments/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_ms, "steps": e2e_steps},
        "gpu_launches": int(launches), "clocks": clocks,
    }
    if not args.no_cpu_baseline:
        try:
            # the unmodified `vt annotate_indels` on a bounded sample of the same records: P range shards of C5_CPU_SAMPLE indels in all
            cores = len(os.sched_getaffinity(0))
            P = max(1, min(cores, 64))
            fa = os.path.join(tmpd, "ref.fa")
            helpers.write_fasta(fa, names, seqs)
            ns = min(C5_CPU_SAMPLE, batch.n_rec)
            cuts = [ns * i // P for i in range(P + 1)]
            for i in range(P):
                with open(os.path.join(tmpd, "in%d.vcf" % i), "w") as f:
                    f.write("##fileformat=VCFv4.2\n" + "".join("##contig=<ID=%s,length=%d>\n" % (n_, q.shape[0]) for n_, q in zip(names, seqs)) + "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n")
                    for (c, p_, al) in recs[cuts[i]:cuts[i + 1]]:
                        f.write("%s\t%d\t.\t%s\t%s\t.\t.\t.\n" % (names[c], p_, al[0].decode(), ",".join(x.decode() for x in al[1:])))
            dt = _run_procs([[VT_REF, "annotate_indels", os.path.join(tmpd, "in%d.vcf" % i), "-r", fa, "-o", os.path.join(tmpd, "out%d.vcf" % i)] for i in range(P)])
            frac_al = hb.n / max(batch.n_rec, 1)
            line["cpu_baseline"] = {"value": ns * frac_al / dt / 1e6, "unit": "M alignments/s", "cores": P, "kind": "reference",
                                    "sample": "the first %d records of this batch in %d range shards, one unmodified single-threaded `vt annotate_indels` process per shard (VCF + FASTA on tmpfs), %.1f s wall = %.1f records/s; alignments counted at this batch's %.2f alignments per record" % (ns, P, dt, ns / dt, frac_al)}
        except Exception as e:
            line["cpu_baseline"] = {"value": None, "unit": "M alignments/s", "cores": 0, "kind": "reference", "sample": "unavailable: %s" % e}
    import shutil
    shutil.rmtree(tmpd, ignore_errors=True)
    emit(line)


U1_RECORDS = int(os.environ.get("VTN_BENCH_U1_RECORDS", 20_000_000))


def run_u1(args, rank, world, local):
    """U1 (its own metric): `vt uniq` over sorted biallelic records with natural position collisions --
    which duplicate survives and the reader's output order, decided on the device (vtn_uniq_kernel)."""
    import torch
    import torch.distributed as dist
    from vt_b200 import Context
    from vt_b200.api import Batch
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rng = np.random.default_rng(20260106 + 1000 * rank)
    n = U1_RECORDS
    pos = np.sort(rng.integers(1, 4 * n, n)).astype(np.int32)
    rid = np.zeros(n, np.int32)
    lens = np.ones(2 * n, np.int64)
    indel = rng.random(n) < 0.2
    lens[1::2][indel] = rng.integers(2, 5, int(indel.sum()))
    aoff = np.zeros(2 * n + 1, np.uint32)
    aoff[1:] = np.cumsum(lens)
    data = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, int(aoff[-1]))]
    batch = Batch(pos, rid, np.arange(0, 2 * n + 1, 2, dtype=np.uint32), aoff, data)
    ctx = Context(local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.sync()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    db = ctx.upload(batch)
    for _ in range(args.warmup):
        ctx.uniq_resident(db)
    ctx.sync()
    sampler = ClockSampler(local)
    barrier()
    l0 = ctx.launch_count()
    t0 = time.time()
    ctx.timer_start()
    for _ in range(args.steps):
        ctx.uniq_resident(db)
    ms_total = ctx.timer_stop()
    barrier()
    launches = ctx.launch_count() - l0
    ms_step = max_over_ranks(ms_total / args.steps)
    order, first, cnt, bad = ctx.fetch_uniq(db, n)
    assert bad == -1 and cnt[0] == n
    # end to end: host buffers in, order / first out, through vtn_uniq
    e2e_steps = max(3, min(args.steps, 10))
    ctx.uniq(batch)
    barrier()
    w0 = time.perf_counter()
    for _ in range(e2e_steps):
        o2, f2, c2, _ = ctx.uniq(batch)
    barrier()
    e2e_ms = max_over_ranks(1e3 * (time.perf_counter() - w0) / e2e_steps)
    t1 = time.time()
    clocks = sampler.stop(t0, t1)
    assert c2 == cnt and np.array_equal(o2, order) and np.array_equal(f2, first)
    ctx.free(db)
    if rank != 0:
        return
    # algorithmic bytes: every record reads pos + rid (8 B) and writes order + first (5 B); a record of a position group of
    # >= 2 also reads its allele_beg (4 B), allele_off (8 B) and allele bytes
    same_prev = np.zeros(n, bool)
    same_prev[1:] = pos[1:] == pos[:-1]
    grouped = same_prev.copy()
    grouped[:-1] |= same_prev[1:]
    rec_bytes = (aoff[2::2].astype(np.int64) - aoff[:-1:2].astype(np.int64))
    alg = 13 * n + int((12 + rec_bytes[grouped]).sum())
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = alg / (ms_total / args.steps / 1e3) / 1e9
    h2d = 8 * n + 4 * (n + 1) + 4 * (2 * n + 1) + batch.n_bytes
    d2h = 5 * n + 8
    emit({
        "metric": "uniq_records_per_s", "value": world * n / (ms_step / 1e3) / 1e6, "unit": "M records/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "U1: vt uniq over %d sorted synthetic biallelic records per GPU (80%% SNV, 20%% short insertions), %d in position groups of >= 2, %d unique" % (n, int(grouped.sum()), cnt[1]),
                   "records_per_gpu": n, "l2": "inputs+outputs per step (%d MB) exceed the 126 MB L2; no explicit flush" % ((8 * n + 5 * n) // 1000000),
                   "sharding": "independent batch per rank (a shard boundary must fall between position groups), no collective"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                     "kernel": "vtn_uniq_kernel", "algorithmic_bytes_per_launch": alg,
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)"},
        "e2e": {"value": world * n / (e2e_ms / 1e3) / 1e6, "unit": "M records/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_ms, "steps": e2e_steps},
        "gpu_launches": int(launches), "clocks": clocks,
    })


B1_RECORDS = int(os.environ.get("VTN_BENCH_B1_RECORDS", 20_000_000))


def run_b1(args, rank, world, local):
    """B1 (its own metric): vt decompose_blocksub's fan-out plan on the device (vtn_blocksub: three launches) over
    biallelic records of which 10 % are block substitutions of 2..8 bases."""
    import torch
    import torch.distributed as dist
    from vt_b200 import Context
    from vt_b200.api import Batch
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rng = np.random.default_rng(20260109 + 1000 * rank)
    n = B1_RECORDS
    is_bs = rng.random(n) < 0.1
    L = np.where(is_bs, rng.integers(2, 9, n), 1).astype(np.int64)
    aoff = np.zeros(2 * n + 1, np.int64)
    aoff[1::2] = np.cumsum(2 * L) - L
    aoff[2::2] = np.cumsum(2 * L)
    data = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, int(aoff[-1]))]
    batch = Batch(np.arange(1, 20 * n + 1, 20, dtype=np.int64).astype(np.int32), np.zeros(n, np.int32), np.arange(0, 2 * n + 1, 2, dtype=np.uint32),
                  aoff.astype(np.uint32), data)
    ctx = Context(local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.sync()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    db = ctx.upload(batch)
    for _ in range(args.warmup):
        ctx.blocksub_resident(db)
    ctx.sync()
    sampler = ClockSampler(local)
    barrier()
    l0 = ctx.launch_count()
    t0 = time.time()
    ctx.timer_start()
    for _ in range(args.steps):
        ctx.blocksub_resident(db)
    ms_total = ctx.timer_stop()
    barrier()
    launches = ctx.launch_count() - l0
    ms_step = max_over_ranks(ms_total / args.steps)
    e2e_steps = max(3, min(args.steps, 10))
    ctx.blocksub(batch)
    barrier()
    w0 = time.perf_counter()
    for _ in range(e2e_steps):
        src, off, ln, cnt = ctx.blocksub(batch)
    barrier()
    e2e_ms = max_over_ranks(1e3 * (time.perf_counter() - w0) / e2e_steps)
    t1 = time.time()
    clocks = sampler.stop(t0, t1)
    n_out = len(src)
    assert cnt[0] == n and cnt[1] == int(is_bs.sum()) and n_out == cnt[3] and np.all(np.diff(src.astype(np.int64)) >= 0)
    ctx.free(db)
    if rank != 0:
        return
    # algorithmic bytes: allele_beg (4 B) + the output triple (12 B per output) for everyone; allele_off (12 B) for biallelic records;
    # the allele bytes of the block substitutions twice (count and emit walk them)
    alg = 4 * n + 12 * n_out + 12 * n + 2 * int((2 * L[is_bs]).sum())
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = alg / (ms_total / args.steps / 1e3) / 1e9
    h2d = 8 * n + 4 * (n + 1) + 4 * (2 * n + 1) + batch.n_bytes
    emit({
        "metric": "blocksub_records_per_s", "value": world * n / (ms_step / 1e3) / 1e6, "unit": "M records/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "B1: vt decompose_blocksub plan over %d biallelic records per GPU, %d block substitutions of 2..8 bases -> %d output records" % (n, int(is_bs.sum()), n_out),
                   "records_per_gpu": n, "l2": "inputs+outputs per step (%d MB) exceed the 126 MB L2; no explicit flush" % (alg // 1000000),
                   "sharding": "independent batch per rank, no collective"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                     "kernel": "vtb_count_kernel + vtb_emit_kernel", "algorithmic_bytes_per_launch": alg,
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)"},
        "e2e": {"value": world * n / (e2e_ms / 1e3) / 1e6, "unit": "M records/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 12 * n_out + 24,
                "ms_per_step": e2e_ms, "steps": e2e_steps},
        "gpu_launches": int(launches), "clocks": clocks,
    })


T1_RECORDS = int(os.environ.get("VTN_BENCH_T1_RECORDS", 4_000_000))


def run_t1(args, rank, world, local):
    """T1 (its own metric): VCF record text -> SoA batch on the device (vtn_vcf_tokenize: eight launches), sites-only
    lines shaped like C2 (80 % SNV, 20 % short indels / multiallelic)."""
    import torch
    import torch.distributed as dist
    from vt_b200 import Context
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rng = np.random.default_rng(20260108 + 1000 * rank)
    n = T1_RECORDS
    pos = np.sort(rng.integers(1, 248_000_000, n))
    b = np.array(list("ACGT"))
    ref = b[rng.integers(0, 4, n)]
    alt = b[rng.integers(0, 4, n)]
    kind = rng.random(n)
    ext = np.array(["", "AC", "GTT", ",T", ",TAA,G"])[np.where(kind < 0.8, 0, rng.integers(1, 5, n))]
    lines = ["chr1\t%d\t.\t%s\t%s%s\t.\tPASS\tAC=1;AN=2" % (pos[i], ref[i], alt[i], ext[i]) for i in range(n)]
    text = ("\n".join(lines) + "\n").encode()
    del lines
    ctx = Context(local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.sync()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    h, caps = ctx.vcf_upload(text)
    for _ in range(args.warmup):
        ctx.vcf_tokenize_resident(h)
    ctx.sync()
    sampler = ClockSampler(local)
    barrier()
    l0 = ctx.launch_count()
    t0 = time.time()
    ctx.timer_start()
    for _ in range(args.steps):
        ctx.vcf_tokenize_resident(h)
    ms_total = ctx.timer_stop()
    barrier()
    launches = ctx.launch_count() - l0
    ms_step = max_over_ranks(ms_total / args.steps)
    f = ctx.vcf_fetch(h, caps)
    assert f["n_rec"] == n and not f["status"].any() and np.array_equal(f["pos1"], pos.astype(np.int32))
    e2e_steps = max(3, min(args.steps, 5))
    ctx.vcf_tokenize(text)
    barrier()
    w0 = time.perf_counter()
    for _ in range(e2e_steps):
        f2 = ctx.vcf_tokenize(text)
    barrier()
    e2e_ms = max_over_ranks(1e3 * (time.perf_counter() - w0) / e2e_steps)
    t1 = time.time()
    clocks = sampler.stop(t0, t1)
    assert all(np.array_equal(f[k], f2[k]) for k in ("line_off", "pos1", "allele_beg", "allele_off", "bytes", "tail_off", "chrom_len"))
    ctx.vcf_free(h)
    if rank != 0:
        return
    d2h = 25 * n + 8 + 4 * (f["n_allele"] + 1) + f["n_bytes"]
    alg = len(text) + d2h
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = alg / (ms_total / args.steps / 1e3) / 1e9
    emit({
        "metric": "vcf_tokenize_records_per_s", "value": world * n / (ms_step / 1e3) / 1e6, "unit": "M records/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "T1: %d sites-only VCF record lines per GPU (%.1f MB of text, %.1f B per line; 80%% SNV, 20%% indel / multiallelic) -> pos1, allele_beg/off/bytes, line and tail offsets" % (n, len(text) / 1e6, len(text) / n),
                   "records_per_gpu": n, "text_gb_per_s": world * len(text) / (ms_step / 1e3) / 1e9,
                   "l2": "text + outputs per step (%d MB) exceed the 126 MB L2; no explicit flush" % (alg // 1000000),
                   "host_sync": "one 4-byte readback of the line count per step sizes the record-level grids",
                   "sharding": "independent text chunk per rank (cut at a line end), no collective"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                     "kernel": "vtk_* (eight launches)", "algorithmic_bytes_per_launch": alg,
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)"},
        "e2e": {"value": world * n / (e2e_ms / 1e3) / 1e6, "unit": "M records/s", "h2d_bytes_per_step": len(text), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_ms, "steps": e2e_steps},
        "gpu_launches": int(launches), "clocks": clocks,
    })


W1_RECORDS = int(os.environ.get("VTN_BENCH_W1_RECORDS", 20_000_000))


def run_w1(args, rank, world, local):
    """W1 (its own metric): the order in which BCFOrderedWriter (window 10000) writes a normalized record stream,
    computed on the device (vtn_writer_order: nine streaming launches).  Positions are sorted per contig, then 12.5 % of them are
    moved left by 1..50 bases like left-aligned indels."""
    import torch
    import torch.distributed as dist
    from vt_b200 import Context
    from vt_b200.api import Batch
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rng = np.random.default_rng(20260107 + 1000 * rank)
    n = W1_RECORDS
    n_contig = 24
    rid = np.sort(rng.integers(0, n_contig, n)).astype(np.int32)
    pos = rng.integers(100, 200_000_000, n).astype(np.int64)
    o = np.lexsort((pos, rid))
    pos = pos[o]
    moved = rng.random(n) < 0.125
    pos[moved] -= rng.integers(1, 51, int(moved.sum()))
    pos = pos.astype(np.int32)
    window = 10000
    data = np.frombuffer(b"AC", np.uint8)[np.tile(np.arange(2), n)]
    batch = Batch(pos, rid, np.arange(0, 2 * n + 1, 2, dtype=np.uint32), np.arange(0, 2 * n + 1, dtype=np.uint32), data)
    ctx = Context(local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.sync()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    db = ctx.upload(batch)
    for _ in range(args.warmup):
        ctx.writer_order_resident(db, window)
    ctx.sync()
    sampler = ClockSampler(local)
    barrier()
    l0 = ctx.launch_count()
    t0 = time.time()
    ctx.timer_start()
    for _ in range(args.steps):
        ctx.writer_order_resident(db, window)
    ms_total = ctx.timer_stop()
    barrier()
    launches = ctx.launch_count() - l0
    ms_step = max_over_ranks(ms_total / args.steps)
    order, late, pend = ctx.fetch_writer_order(db, n)
    e2e_steps = max(3, min(args.steps, 10))
    ctx.writer_order(rid, pos, window)
    barrier()
    w0 = time.perf_counter()
    for _ in range(e2e_steps):
        o2, l2 = ctx.writer_order(rid, pos, window)
    barrier()
    e2e_ms = max_over_ranks(1e3 * (time.perf_counter() - w0) / e2e_steps)
    t1 = time.time()
    clocks = sampler.stop(t0, t1)
    assert np.array_equal(o2, order) and np.array_equal(l2, late)
    # size-independent check: the disorder stays inside the window, so the order is the stable sort by (contig run, position)
    assert np.array_equal(order, np.lexsort((np.arange(n), pos, rid)).astype(np.uint32)) and (late == np.iinfo(np.int32).min).all()
    ctx.free(db)
    if rank != 0:
        return
    alg = 16 * n  # rid + pos in, order + late_front out
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = alg / (ms_total / args.steps / 1e3) / 1e9
    emit({
        "metric": "writer_order_records_per_s", "value": world * n / (ms_step / 1e3) / 1e6, "unit": "M records/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": {"workload": "W1: BCFOrderedWriter order (window %d) of %d records per GPU on %d contigs, 12.5%% moved left by 1..50 bases; %d records out of arrival order" % (window, n, n_contig, int((order != np.arange(n)).sum())),
                   "records_per_gpu": n, "l2": "the intermediate arrays (52 B per record) exceed the 126 MB L2; no explicit flush",
                   "sharding": "independent stream per rank (a shard boundary at a contig change needs no carry), no collective"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                     "kernel": "vto_* (nine launches, all streaming)", "algorithmic_bytes_per_launch": alg,
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)"},
        "e2e": {"value": world * n / (e2e_ms / 1e3) / 1e6, "unit": "M records/s", "h2d_bytes_per_step": 8 * n, "d2h_bytes_per_step": 8 * n,
                "ms_per_step": e2e_ms, "steps": e2e_steps},
        "gpu_launches": int(launches), "clocks": clocks,
    })


C3_RECORDS = int(os.environ.get("VTN_BENCH_C3_RECORDS", 12_500_000))


def run_c3(args, rank, world, local):
    """C3: multiallelic records, `decompose -s`-style allele fan-out fused with normalize on the device (the biallelic
    batch is materialised in HBM by the emit kernel, no host round trip).  100 M records over 8 GPUs = 12.5 M per GPU."""
    import torch
    import torch.distributed as dist
    from vt_b200 import Context, synth
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    names, seqs = synth.make_reference(scale=REF_SCALE)
    batch = synth.make_multiallelic(seqs, C3_RECORDS, seed=synth.SEED_VCF_C3 + 1000 * rank)
    ctx = Context(local)
    ctx.load_reference(names, seqs)
    ref_bases = sum(x.shape[0] for x in seqs)
    del seqs

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.sync()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # one host-buffer call: the result is needed for the byte accounting (and is the e2e path timed below)
    src, alt, dcnt, res = ctx.decompose(batch, normalize=True)
    n_out = int(src.shape[0])
    lens = np.diff(batch.allele_off.astype(np.int64))
    na = np.diff(batch.allele_beg.astype(np.int64))
    ref_len = lens[batch.allele_beg[:-1].astype(np.int64)]
    rl = ref_len[src.astype(np.int64)]
    Cc = (res.vtype[:n_out] != 1).astype(np.int64)
    S = np.zeros(n_out, np.int64)
    k = res.n_changed
    S[res.chg_rec[:k]] = res.chg_le[:k]
    alg_bytes = int((16 + 2 * na).sum() + lens.sum() + ((Cc * rl + S + 3) // 4).sum() + (16 + 6) * n_out + 12 * k + res.n_changed_bytes)
    db = ctx.upload(batch)
    for _ in range(args.warmup):
        ctx.decompose_resident(db, True)
    ctx.sync()
    sampler = ClockSampler(local)
    barrier()
    l0 = ctx.launch_count()
    t0 = time.time()
    ctx.timer_start()
    for _ in range(args.steps):
        ctx.decompose_resident(db, True)
    ms_total = ctx.timer_stop()
    barrier()
    launches = ctx.launch_count() - l0
    ms_step = max_over_ranks(ms_total / args.steps)
    # end to end through vtn_decompose_normalize with pinned host buffers allocated once (inputs and results)
    import ctypes as C
    from vt_b200 import _lib as L
    from vt_b200.api import NormResult, _p
    hb = pinned_batch(ctx, batch)
    nb_bound = int((np.maximum(na - 1, 1) * ref_len).sum() + lens.sum())
    hres = NormResult(n_out, nb_bound, ctx)
    hsrc = ctx.pinned_empty(n_out + 1, np.uint32)
    halt = ctx.pinned_empty(n_out + 1, np.uint16)
    hn = C.c_uint64(0)
    hdc = L.VtnDecompCounters()
    hcb = hb.as_c()

    def e2e_call():
        ctx._chk(ctx.lib.vtn_decompose_normalize(ctx.h, C.byref(hcb), _p(hsrc, L.c_u32p), _p(halt, L.c_u16p), C.byref(hn), C.byref(hdc),
                                                 C.byref(hres.c)), "vtn_decompose_normalize")
    e2e_call()
    assert int(hn.value) == n_out and hres.n_changed == k
    e2e_steps = max(3, min(args.steps, 5))
    barrier()
    w0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_call()
    barrier()
    e2e_ms = max_over_ranks(1e3 * (time.perf_counter() - w0) / e2e_steps)
    clocks = sampler.stop(t0, time.time())
    ctx.free(db)
    if rank != 0:
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = alg_bytes / (ms_total / args.steps / 1e3) / 1e9
    h2d = 8 * batch.n_rec + 4 * (batch.n_rec + 1) + 4 * (batch.n_allele + 1) + batch.n_bytes
    d2h = 6 * n_out + 2 * n_out + 20 * k + 8 * (k + 1) + res.n_changed_bytes + 64
    line = {
        "metric": METRIC, "value": world * n_out / (ms_step / 1e3) / 1e6, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {
            "workload": "C3: %d synthetic multiallelic records per GPU (2-6 ALTs, 80%% SNV/10%% del/10%% ins per ALT) -> %d biallelic records, allele fan-out fused with normalize on device, synthetic %.2f Gb reference resident in HBM; value counts the biallelic OUTPUT records" % (batch.n_rec, n_out, ref_bases / 1e9),
            "input_records_per_gpu": batch.n_rec, "output_records_per_gpu": n_out, "changed_per_gpu": int(k),
            "input_m_records_per_s": world * batch.n_rec / (ms_step / 1e3) / 1e6,
            "decompose_counters": [int(x) for x in dcnt],
            "l2": "inputs+outputs per step (%d MB) exceed the 126 MB L2; no explicit flush" % ((h2d + d2h) // 1000000),
            "sharding": "independent batch per rank (100 M records / 8 GPUs), reference replicated, no collective",
        },
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                     "kernel": "vtd_count_kernel + vtd_emit_kernel + vtn_normalize_kernel + vtn_compact_kernel",
                     "algorithmic_bytes_per_launch": alg_bytes,
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)"},
        "e2e": {"value": world * n_out / (e2e_ms / 1e3) / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_ms, "steps": e2e_steps},
        "gpu_launches": int(launches), "clocks": clocks,
    }
    emit(line)


def pinned_batch(ctx, batch):
    """Copy a Batch into pinned host memory obtained from the library (vtn_host_alloc)."""
    import ctypes as C
    from vt_b200.api import Batch

    def pin(a):
        out = ctx.pinned_empty(a.size, a.dtype)
        out[:] = a
        return out
    return Batch(pin(batch.pos1), pin(batch.rid), pin(batch.allele_beg), pin(batch.allele_off), pin(batch.data))


NUMA_NOTE = None
ORIG_AFFINITY = None  # the cores the process started with: the CPU baseline gets all of them back


def bind_to_gpu_numa_node(dev):
    """Runs this rank on the cores of the NUMA node its GPU hangs off, BEFORE any page-locked buffer exists: the pinned batches are then
    allocated on that node (first touch, local policy) and the H2D / D2H DMA does not cross the socket link (round 1's SCALE had all
    eight ranks' pinned buffers on node 0).  Returns what was done, for the bench line."""
    if os.environ.get("VTN_BENCH_NO_NUMA"):
        return "not bound (VTN_BENCH_NO_NUMA set)"
    try:
        try:
            from cuda.bindings import runtime as cudart
        except ImportError:
            from cuda import cudart
        err, bus = cudart.cudaDeviceGetPCIBusId(32, dev)
        bus = (bus.decode(errors="replace") if isinstance(bus, bytes) else str(bus)).split("\x00")[0].strip().lower()
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bus).read())
        if node < 0:
            return "gpu %s: no NUMA node reported, not bound" % bus
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        global ORIG_AFFINITY
        ORIG_AFFINITY = os.sched_getaffinity(0)
        cpus &= ORIG_AFFINITY
        if not cpus:
            return "gpu %s on node %d: none of its cores are available to this process, not bound" % (bus, node)
        os.sched_setaffinity(0, cpus)
        return "gpu %s on NUMA node %d: process bound to its %d cores before the pinned buffers were allocated" % (bus, node, len(cpus))
    except Exception as e:  # no sysfs entry, no cuda-python: run unbound
        return "not bound (%s)" % str(e)[:80]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="vtn", choices=["vtn", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--layout", default="bi", choices=["bi", "general"], help="c2 / c4: compact biallelic batch (vtn_normalize_bi, default) or the general vtn_batch")
    ap.add_argument("--resident-only", action="store_true", help="c2 / c4: only the device-resident loop (profiling passes); prints a reduced line")
    ap.add_argument("--no-sharded", action="store_true", help="c2: skip the sharded multi-device C3 measurement (vtn_multi_decompose_normalize) added to the line")
    ap.add_argument("--no-compare", action="store_true", help="c2 / c4: skip the side measurement of the general layout")
    ap.add_argument("--workload", default="c2", choices=["c2", "c3", "c4", "c5", "u1", "w1", "t1", "b1"],
                    help="c2 (default, the metric's configuration), c3 = multiallelic decompose+normalize fused, c4 = STR/homopolymer reference with maximal left shifts, "
                         "c5 = AHMM/LFHMM/RFHMM realignment (its own metric)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "vtn" else args.warmup
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    global NUMA_NOTE
    NUMA_NOTE = bind_to_gpu_numa_node(local)
    if args.workload == "c5":
        run_c5(args, rank, world, local)
        return
    if args.workload == "c3":
        run_c3(args, rank, world, local)
        return
    if args.workload == "u1":
        run_u1(args, rank, world, local)
        return
    if args.workload == "w1":
        run_w1(args, rank, world, local)
        return
    if args.workload == "t1":
        run_t1(args, rank, world, local)
        return
    if args.workload == "b1":
        run_b1(args, rank, world, local)
        return

    run_c2_c4(args, rank, world, local)


def load_workload(workload, rank):
    rank = int(os.environ.get("VTN_BENCH_SEED_RANK", rank))  # which rank's batch to generate (explains per-rank time differences)
    """(names, seqs, batch) of C2 / C4 for this rank.  VTN_BENCH_CACHE=<dir> keeps the generated arrays as .npy files there so
    that several bench invocations of one session (variants, the ncu passes) do not regenerate 3.1 Gb each."""
    from vt_b200 import synth
    from vt_b200.api import Batch
    cache = os.environ.get("VTN_BENCH_CACHE")
    tag = "%s_%d_%g_%d" % (workload, N_RECORDS, REF_SCALE, rank)
    if cache and os.path.exists(os.path.join(cache, tag + ".done")):
        names = open(os.path.join(cache, tag + ".names")).read().split()
        seqs = [np.load(os.path.join(cache, "%s_seq%d.npy" % (tag, i))) for i in range(len(names))]
        b = [np.load(os.path.join(cache, "%s_b%d.npy" % (tag, i))) for i in range(5)]
        return names, seqs, Batch(*b)
    if workload == "c4":
        tracts = []
        names, seqs = synth.make_reference(seed=synth.SEED_REF_C4, scale=REF_SCALE, str_frac=0.3, tracts_out=tracts)
        batch = synth.make_str_indels(seqs, tracts, N_RECORDS, seed=synth.SEED_VCF_C4 + 1000 * rank)
        del tracts
    else:
        names, seqs = synth.make_reference(scale=REF_SCALE)
        batch = synth.make_biallelic(seqs, N_RECORDS, seed=synth.SEED_VCF_C2 + 1000 * rank)
    if cache:
        os.makedirs(cache, exist_ok=True)
        open(os.path.join(cache, tag + ".names"), "w").write(" ".join(names))
        for i, x in enumerate(seqs):
            np.save(os.path.join(cache, "%s_seq%d.npy" % (tag, i)), x)
        for i, x in enumerate((batch.pos1, batch.rid, batch.allele_beg, batch.allele_off, batch.data)):
            np.save(os.path.join(cache, "%s_b%d.npy" % (tag, i)), x)
        open(os.path.join(cache, tag + ".done"), "w").write("ok")
    return names, seqs, batch


C3S_TOTAL = int(os.environ.get("VTN_BENCH_C3S_TOTAL", 12_500_000))


def sharded_c3(n_dev, names, seqs, steps):
    """Configuration C3 as north_star states it, through the product's multi-device entry (vtn_multi_decompose_normalize): ONE set of
    multiallelic records, sharded by contig / record range over the n_dev GPUs of the box inside one process (one host thread per
    device, reference packed once and replicated by peer copies, no NCCL), decompose fused with normalize on every device, results
    copied by every device straight to their place in the caller's arrays (input order).  STRONG scaling: the total is fixed
    (VTN_BENCH_C3S_TOTAL records, a sample of the 100 M-record configuration).  Timed region (wall clock, pinned host buffers):
    shard -> H2D -> kernels -> D2H of the merged result.  The merged result is compared with the single-device call (untimed)."""
    from vt_b200 import synth
    from vt_b200.api import MultiContext, NormResult
    import ctypes as C
    from vt_b200 import _lib as L
    t0 = time.time()
    batch = synth.make_multiallelic(seqs, C3S_TOTAL, seed=synth.SEED_VCF_C3)
    lens = np.diff(batch.allele_off.astype(np.int64))
    na = np.diff(batch.allele_beg.astype(np.int64))
    ref_len = lens[batch.allele_beg[:-1].astype(np.int64)]
    bound = int(np.maximum(na - 1, 1).sum())
    nb = int((np.maximum(na - 1, 1) * ref_len).sum() + lens.sum())
    slots = max(1, int(os.environ.get("VTN_BENCH_C3S_SLOTS", 3)))  # contexts per GPU: chunks of one device overlap their H2D / kernels / D2H
    with MultiContext(list(range(n_dev)) * slots) as m:
        m.load_reference(names, seqs)
        hb = pinned_batch(m.ctx0, batch)
        bufs = (m.ctx0.pinned_empty(bound + 1, np.uint32), m.ctx0.pinned_empty(bound + 1, np.uint16), NormResult(bound, nb, m.ctx0))
        log("sharded C3 setup %.1fs: %d records -> %d biallelic, %d devices" % (time.time() - t0, batch.n_rec, bound, n_dev))
        for _ in range(2):
            src, alt, dcnt, res = m.decompose_normalize(hb, bufs)
        w0 = time.perf_counter()
        for _ in range(steps):
            src, alt, dcnt, res = m.decompose_normalize(hb, bufs)
        ms = 1e3 * (time.perf_counter() - w0) / steps
        n_out, k = int(src.shape[0]), res.n_changed
        one = (np.zeros(bound + 1, np.uint32), np.zeros(bound + 1, np.uint16), NormResult(bound, nb))
        n1, d1 = C.c_uint64(0), L.VtnDecompCounters()
        cb = hb.as_c()
        m.ctx0._chk(m.ctx0.lib.vtn_decompose_normalize(m.ctx0.h, C.byref(cb), one[0].ctypes.data_as(L.c_u32p), one[1].ctypes.data_as(L.c_u16p), C.byref(n1), C.byref(d1), C.byref(one[2].c)), "vtn_decompose_normalize")
        r1 = one[2]
        same = (int(n1.value) == n_out and r1.n_changed == k and np.array_equal(one[0][:n_out], src) and np.array_equal(one[1][:n_out], alt)
                and np.array_equal(r1.status[:n_out], res.status[:n_out]) and np.array_equal(r1.vtype[:n_out], res.vtype[:n_out])
                and np.array_equal(r1.chg_rec[:k], res.chg_rec[:k]) and np.array_equal(r1.chg_pos1[:k], res.chg_pos1[:k])
                and np.array_equal(r1.chg_off[:k + 1], res.chg_off[:k + 1]) and np.array_equal(r1.chg_bytes[:r1.n_changed_bytes], res.chg_bytes[:res.n_changed_bytes])
                and r1.counters == res.counters)
        assert same, "sharded C3: merged result differs from the single-device call"
    h2d = 8 * batch.n_rec + 4 * (batch.n_rec + 1) + 4 * (batch.n_allele + 1) + batch.n_bytes
    d2h = 6 * n_out + 2 * n_out + 20 * k + 8 * (k + 1) + res.n_changed_bytes
    return {"workload": "C3 sharded in one process: %d multiallelic records (sample of the 100 M config) -> %d biallelic, contig/range chunks of 2 M records over %d GPU(s) x %d contexts each (a chunk's copy-out waits only for the earlier chunks' kernels), decompose fused with normalize per device, results in input order" % (batch.n_rec, n_out, n_dev, slots),
            "slots_per_gpu": slots,
            "scaling": "strong", "n_gpus": n_dev, "value": n_out / (ms / 1e3) / 1e6, "unit": UNIT, "input_m_records_per_s": batch.n_rec / (ms / 1e3) / 1e6,
            "ms_per_step": ms, "steps": steps, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "changed": int(k),
            "merged_equals_single_device": True, "decompose_counters": [int(x) for x in dcnt]}


def slice_batch(batch, a, b):
    """records [a, b) of a Batch as a Batch of their own (checker leg)"""
    from vt_b200.api import Batch
    s0, s1 = int(batch.allele_beg[a]), int(batch.allele_beg[b])
    o0, o1 = int(batch.allele_off[s0]), int(batch.allele_off[s1])
    return Batch(batch.pos1[a:b], batch.rid[a:b], batch.allele_beg[a:b + 1] - np.uint32(s0), batch.allele_off[s0:s1 + 1] - np.uint32(o0), batch.data[o0:o1])


def oracle_check_slice(seqs, batch, status, vtype, res, a, b):
    """Untimed checker leg (the only use of oracle/ in this arm): records [a, b) of the timed batch through the CPU oracle,
    compared bit for bit with what the device left for them.  -> number of records checked."""
    import helpers
    sub = slice_batch(batch, a, b)
    exp = helpers.Oracle().normalize(seqs, sub)
    n = b - a
    assert np.array_equal(status[a:b], exp.status[:n]), "bench: status differs from the oracle on the checked slice"
    assert np.array_equal(vtype[a:b], exp.vtype[:n]), "bench: vtype differs from the oracle on the checked slice"
    k = res.n_changed
    rec = res.chg_rec[:k]
    lo, hi = int(np.searchsorted(rec, a)), int(np.searchsorted(rec, b))
    ke = exp.n_changed
    assert hi - lo == ke, "bench: changed-record count differs from the oracle on the checked slice (%d vs %d)" % (hi - lo, ke)
    assert np.array_equal(rec[lo:hi] - np.uint32(a), exp.chg_rec[:ke])
    for name in ("chg_pos1", "chg_lt", "chg_le", "chg_rt"):
        assert np.array_equal(getattr(res, name)[lo:hi], getattr(exp, name)[:ke]), "bench: %s differs from the oracle" % name
    b0, b1 = int(res.chg_off[lo]), int(res.chg_off[hi])
    assert b1 - b0 == exp.n_changed_bytes and np.array_equal(res.chg_bytes[b0:b1], exp.chg_bytes[:exp.n_changed_bytes]), "bench: new allele bytes differ from the oracle"
    return n


def kernel_source_hash():
    """sha1 over the sources of the normalize kernels: keys the ncu traffic record in profiles/traffic.json"""
    import hashlib
    h = hashlib.sha1()
    for f in ("vtn_normalize.cu", "vtn_fast.cuh", "vtn_ref_dev.cuh", "vtn_internal.h", "vtn_scan.cuh"):
        h.update(open(os.path.join(ROOT, "vt_b200", "csrc", f), "rb").read())
    return h.hexdigest()[:16]


def measured_traffic(workload, layout):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel from an ncu --set full capture made with
    THESE kernel sources (profiles/traffic.json, written by profiles/ncu_traffic.py); None when no capture matches."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:
        return None, None
    e = t.get("%s:%s:%s" % (kernel_source_hash(), workload, layout))
    if not e or e.get("records") != N_RECORDS or e.get("ref_scale") != REF_SCALE:
        return None, None
    return float(e["dram_bytes"]), e.get("source")


def run_c2_c4(args, rank, world, local):
    import torch
    import torch.distributed as dist
    from vt_b200 import Context, synth
    from vt_b200.api import NormResult, NormResultBi
    torch.cuda.set_device(local)
    cpu_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        cpu_group = dist.new_group(backend="gloo")  # waits that must not put a spinning kernel on a GPU (the end of this function)
        import atexit
        atexit.register(lambda: dist.is_initialized() and dist.destroy_process_group())

    t_setup = time.time()
    names, seqs, batch = load_workload(args.workload, rank)
    ctx = Context(local)
    ctx.load_reference(names, seqs)
    ref_bases = sum(s.shape[0] for s in seqs)
    if rank != 0:
        del seqs
    log("[rank %d] setup %.1fs: %d records, %d allele bytes, reference %d bases" % (rank, time.time() - t_setup, batch.n_rec, batch.n_bytes, ref_bases))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.sync()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def all_ranks(x):
        if world == 1:
            return [x]
        t = torch.zeros(world, dtype=torch.float64, device="cuda")
        t[rank] = x
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return [float(v) for v in t.tolist()]

    bi_layout = args.layout == "bi"
    bi = ctx.bi_pack(batch) if bi_layout else None

    def timed_resident(db, steps):
        for _ in range(args.warmup):
            ctx.normalize_resident(db)
        ctx.sync()
        barrier()
        l0 = ctx.launch_count()
        ctx.timer_start()
        for _ in range(steps):
            ctx.normalize_resident(db)
        ms = ctx.timer_stop()
        barrier()
        return ms / steps, ctx.launch_count() - l0

    # ---------------- device-resident: value + roofline ----------------
    db = ctx.upload_bi(bi) if bi_layout else ctx.upload(batch)
    ctx.normalize_resident(db)
    ctx.sync()
    if bi_layout:
        res = ctx.fetch_result_bi(db, bi)
        status, vtype = res.expand()
    else:
        res = ctx.fetch_result(db, batch)
        status, vtype = res.status, res.vtype

    class _R:  # what synth.algorithmic_bytes reads
        pass
    ar = _R()
    ar.vtype, ar.n_changed, ar.chg_rec, ar.chg_le, ar.n_changed_bytes = vtype, res.n_changed, res.chg_rec, res.chg_le, res.n_changed_bytes
    alg_bytes = synth.algorithmic_bytes(batch, ar)
    sampler = ClockSampler(local)
    t0 = time.time()
    ms_local, launches = timed_resident(db, args.steps)
    ms_step = max_over_ranks(ms_local)
    per_rank_ms = all_ranks(ms_local)
    value = world * batch.n_rec / (ms_step / 1e3) / 1e6
    # the timed result once more, after the timed loop: this is what the checker leg below looks at
    if bi_layout:
        res = ctx.fetch_result_bi(db, bi)
        status, vtype = res.expand()
    else:
        res = ctx.fetch_result(db, batch)
        status, vtype = res.status, res.vtype

    if args.resident_only:
        if rank == 0:
            emit({"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
                  "resident_only": True, "layout": args.layout, "gpu_launches": int(launches), "algorithmic_bytes_per_launch": alg_bytes, "clocks": sampler.stop(t0, time.time())})
        ctx.free(db)
        return
    # ---------------- end to end through the C ABI with pinned host buffers ----------------
    e2e_steps = max(3, min(args.steps, 20))
    if bi_layout:
        hres = NormResultBi(batch.n_rec, batch.n_bytes, ctx)
        call = lambda: ctx.normalize_bi(bi, hres)
    else:
        hb = pinned_batch(ctx, batch)
        hres = NormResult(batch.n_rec, batch.n_bytes, ctx)
        call = lambda: ctx.normalize(hb, hres)
    for _ in range(3):
        call()
    barrier()
    w0 = time.perf_counter()
    for _ in range(e2e_steps):
        call()
    barrier()
    e2e_local = 1e3 * (time.perf_counter() - w0) / e2e_steps
    e2e_ms = max_over_ranks(e2e_local)
    per_rank_e2e = all_ranks(e2e_local)
    t1 = time.time()
    clocks = sampler.stop(t0, t1)
    e2e_val = world * batch.n_rec / (e2e_ms / 1e3) / 1e6
    k = hres.n_changed
    if bi_layout:
        h2d = bi.h2d_bytes
        d2h = batch.n_rec + 16 + 48 + 4 + 20 * k + 8 * (k + 1) + hres.n_changed_bytes + 5 * hres.n_notes
    else:
        h2d = 4 * batch.n_rec * 2 + 4 * (batch.n_rec + 1) + 4 * (batch.n_allele + 1) + batch.n_bytes
        d2h = 2 * batch.n_rec + 16 + 48 + 20 * k + 8 * (k + 1) + hres.n_changed_bytes
    assert hres.n_changed == res.n_changed and hres.counters == res.counters
    assert np.array_equal(hres.chg_rec[:k], res.chg_rec[:k]) and np.array_equal(hres.chg_pos1[:k], res.chg_pos1[:k])

    # the general layout beside it (same records, vtn_batch): what the compact layout is measured against
    other = None
    if bi_layout and not args.no_compare:
        db2 = ctx.upload(batch)
        o_ms, _ = timed_resident(db2, max(3, min(args.steps, 10)))
        r2 = ctx.fetch_result(db2, batch)
        assert np.array_equal(r2.status[:batch.n_rec], status) and np.array_equal(r2.vtype[:batch.n_rec], vtype) and r2.n_changed == res.n_changed
        assert np.array_equal(r2.chg_bytes[:r2.n_changed_bytes], res.chg_bytes[:res.n_changed_bytes])
        ctx.free(db2)
        other = {"layout": "general (vtn_batch)", "ms_per_step": max_over_ranks(o_ms)}

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = alg_bytes / (ms_local / 1e3) / 1e9

    if rank != 0:
        ctx.free(db)
        if world > 1:
            # rank 0 is still measuring (sharded C3 over ALL GPUs, CPU baseline): wait on the CPU (gloo) -- an NCCL barrier would leave a
            # spinning kernel on this rank's GPU while rank 0's contexts run on it
            torch.cuda.synchronize()
            dist.barrier(group=cpu_group)
        return
    # ---------------- checker leg (untimed): a fixed slice of the timed batch against the CPU oracle ----------------
    n_chk = min(batch.n_rec, CHECK_RECORDS)
    a_chk = (batch.n_rec - n_chk) // 2
    checked = oracle_check_slice(seqs, batch, status, vtype, res, a_chk, a_chk + n_chk)
    traffic, traffic_src = measured_traffic(args.workload, args.layout)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {
            "workload": ("C2: %d synthetic biallelic records per GPU (80%% SNV, 10%% del, 10%% ins, half of the indels unnormalized), synthetic %.2f Gb GRCh38-shaped reference resident in HBM" if args.workload == "c2" else
                         "C4: %d synthetic STR/homopolymer indels per GPU (k=1..4 whole units, 90%% at the rightmost copy of the tract), synthetic %.2f Gb reference with 30%% tandem repeats resident in HBM") % (batch.n_rec, ref_bases / 1e9),
            "layout": "compact biallelic batch (vtn_batch_bi: pos1 + 2 length bytes + allele bytes per record, contig runs)" if bi_layout else "general batch (vtn_batch)",
            "records_per_gpu": batch.n_rec, "changed_per_gpu": res.n_changed,
            "l2": "inputs+outputs per step (%d MB) exceed the 126 MB L2; no explicit flush" % ((h2d + d2h) // 1000000),
            "sharding": "independent batch per rank, reference replicated, no collective",
            "numa": "every rank: " + str(NUMA_NOTE) + " (rank 0's note)",
            "per_rank_ms_per_step": per_rank_ms, "per_rank_e2e_ms_per_step": per_rank_e2e,
            "checked_against_oracle": "%d records [%d, %d) of the timed batch: status, vtype, changed records and new allele bytes bit-exact (untimed checker leg)" % (checked, a_chk, a_chk + n_chk),
        },
        "roofline": {
            "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": traffic, "traffic_source": traffic_src,
            "kernel": "vtn_normalize_bi_kernel" if bi_layout else "vtn_normalize_kernel", "algorithmic_bytes_per_launch": alg_bytes,
            "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)",
        },
        "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_ms, "steps": e2e_steps},
        "gpu_launches": int(launches),
        "clocks": clocks,
    }
    if other:
        line["config"]["general_layout"] = other
    if args.workload == "c4":
        k_ = res.n_changed
        le = np.asarray(res.chg_le[:k_], dtype=np.int64)
        hist = np.bincount(np.minimum(np.floor(np.log2(np.maximum(le, 1))).astype(np.int64), 14), minlength=15)
        line["config"]["left_shift_histogram_log2_bins"] = [int(x) for x in hist]
        line["config"]["left_shift_mean"] = float(le.mean()) if k_ else 0.0
    if args.workload == "c2" and not args.no_sharded:
        # the other ranks wait (on the CPU) at the end of this function while rank 0 drives all `world` GPUs through vtn_multi_*
        try:
            line["sharded_c3"] = sharded_c3(world, names, seqs, 3)
        except Exception as e:
            line["sharded_c3"] = {"error": str(e)[:300]}
    if not args.no_cpu_baseline and world >= 1 and args.workload == "c2":
        try:
            cli = {}
            cv, cms, P, sample = cpu_reference_run(CPU_SAMPLE_RECORDS, 2, 1, names, seqs, batch, cli_out=cli)
            line["cpu_baseline"] = {"value": cv, "unit": UNIT, "cores": P, "kind": "reference", "sample": sample}
            line["e2e_cli"] = cli
        except Exception as e:  # the checker binary is missing on this box
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "unavailable: %s" % e}
    emit(line)
    ctx.free(db)
    if world > 1:
        dist.barrier(group=cpu_group)


if __name__ == "__main__":
    main()
