#!/usr/bin/env python
"""bench.py — haplotype materialisation throughput (simulated haplotype Gbp/s) on B200.

Workload (BASELINE.json configs[1]): chr1-sized synthetic reference (248 956 422 bp), 1 000 haplotypes,
per-haplotype SV edit lists -> segment scan -> tiled copy kernel. One "step" = one batch of `--batch` haplotypes
(plan kernels + materialisation kernel) with all inputs resident in HBM; K steps walk through the 1 000 haplotypes.
Inputs per step (>= 31 GB written, 249 MB reference) are far larger than the 126 MB L2, so no flush is needed.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Under torchrun each rank owns its own shard of haplotypes (individuals are sharded, no data-path collective):
weak scaling. `--impl reference` times the CPU restatement of the reference path (oracle/, kind "port": no JVM in
this image) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "population-sv-simulator_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

CHR1 = 248_956_422
N_HAPLOTYPES = 1000
ALGO_BYTES_PER_BP = 2.0  # 1 B read (reference/payload) + 1 B written per output base


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def measured_traffic(batch: int, L: int, bases_per_launch: float):
    """dram__bytes_read.sum + dram__bytes_write.sum of one materialise_kernel launch of THIS step shape, from the committed ncu
    capture of `bench.py` itself (profiles/r02_traffic.json, made by tools/capture_traffic.py). A capture whose launch shape
    differs (batch, reference length, or bases per launch off by more than 1 %) is not quoted: returns None."""
    p = os.path.join(ROOT, "profiles", "r02_traffic.json")
    if os.path.exists(p):
        t = json.load(open(p))
        if (t.get("batch") == batch and t.get("reference_length") == L and
                abs(t.get("bases_per_launch", 0) - bases_per_launch) <= 0.01 * bases_per_launch):
            return float(t["dram_bytes_per_launch"]), t
    return None, None


def synth_population(L: int, H: int, seed: int = 20261104):
    """Synthetic variant table + per-haplotype edit lists of the shape SimulateHaplotypes produces
    (V ~ 11k SVs, ~1.1k per haplotype, type mix DEL .52 / DUP .15 / INS .328 / INV .002, globally non-overlapping so
    every subset is collision-free). Used until row G/C/S1 feed the bench from the real sampler."""
    rng = np.random.default_rng(seed)
    V = min(11000, max(8, L // 4000))
    types = rng.choice(4, size=V, p=[0.52, 0.15, 0.328, 0.002]).astype(np.int8)
    slot = (L - 2000) // V
    maxlen = min(20000, slot // 2)
    length = np.minimum(np.exp(rng.uniform(np.log(50), np.log(maxlen), size=V)), maxlen).astype(np.int32)
    pos = (1000 + np.arange(V, dtype=np.int64) * slot + rng.integers(0, slot - maxlen - 1, size=V)).astype(np.int32)
    plen = np.where(types == 2, length, 0).astype(np.int64)
    payload_off = np.concatenate(([0], np.cumsum(plen))).astype(np.int64)
    payload = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=int(plen.sum()), dtype=np.uint8)].copy()
    af = np.exp(rng.uniform(np.log(4.6e-5), 0.0, size=V))
    carr = rng.random((H, V)) < af[None, :]
    lists = [np.nonzero(carr[h])[0].astype(np.int32) for h in range(H)]
    list_off = np.concatenate(([0], np.cumsum([l.size for l in lists]))).astype(np.int64)
    ids = np.concatenate(lists).astype(np.int32)
    return dict(sv_type=types, position=pos, length=length, payload_off=payload_off, payload=payload, lists=lists,
                list_off=list_off, ids=ids)


def random_reference(L: int, seed: int = 20261104) -> np.ndarray:
    rng = np.random.default_rng(seed)
    return np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=L, dtype=np.uint8)].copy()


CCS_CLI = ["--length-min", "5000", "--length-max", "30000", "--accuracy-min", "0.99", "--accuracy-max", "1.0",
           "--difference-ratio", "6:50:54", "--length-mean", "15000", "--length-sd", "2000", "--accuracy-mean", "0.995"]
READ_ALGO_BYTES_PER_BASE = 2.0  # 1 B of haplotype read + 1 B of read written per emitted base


def reads_leg(ctx, splib, synth, work, first_hap, n_haps, depth, peak, torch, quality="phred", resample=True):
    """BASELINE configs[3] on this GPU's shard: pbsim2-model CCS reads at `depth`x over `n_haps` chr1-sized haplotypes, taken
    individual by individual as docker/haplotype2reads.sh does: the two haplotypes are materialised on the device (they never
    leave it), sequenced, and their reads mixed into one file (cat + seeded shuf + tr) that stays in HBM. The timed region
    covers everything from the genotype kernels of an individual to its mixed file. Returns the `reads` object."""
    model = os.path.join(work, "P6C4_synthetic.model")
    synth.write_hmm_model(model, quality=quality)
    args, _ = splib.ReadArgs.from_options(CCS_CLI + ["--depth", str(depth), "--hmm_model", model, "--seed", "42"],
                                          stream_mode=splib.READS_PER_READ_STREAMS)
    ctx.reads_configure(args)
    n_ind = n_haps // 2
    tot = dict(bases=0, reads=0, dropped=0, k_ms=0.0, p_ms=0.0, mix_copy_ms=0.0, mix_order_ms=0.0, mixed_bytes=0, mixed_reads=0)
    used = last = None
    one_by_one_ms = []

    def individual(h0, timed, pair=True):
        nonlocal used, last
        if resample:
            ctx.population_genotypes(h0, h0 + 1)   # rows G/C/N for this individual
        ctx.plan_haplotypes(h0, h0 + 1, False)
        ctx.materialise()
        ctx.reads_mix_begin()
        if pair:
            # both read kernels of the individual in flight (sp_reads_mix_add_pair): the kernel time is the device span from
            # the first kernel's start to the second one's end
            res, span = ctx.reads_mix_add_pair(h0, h0 + 1, 1, 1)
            if timed:
                for r in res:
                    tot["bases"] += r["bases"]; tot["reads"] += r["n_reads"]; tot["dropped"] += r["n_dropped"]
                tot["k_ms"] += span; tot["p_ms"] += ctx.reads_timing()["plan_ms"]
        else:   # warm-up: one by one (also what the draw statistics are taken from)
            for h in (h0, h0 + 1):
                r = ctx.reads_simulate_haplotype(h, 1)
                last = r
                _, _, used = ctx.reads_log(r["n_reads"])   # rand() calls of this haplotype's reads
                one_by_one_ms.append(ctx.reads_timing()["kernel_ms"])
                ctx.reads_mix_add()
        n_mixed, mixed_bytes = ctx.reads_mix_finish(42 + h0)
        if timed:
            mt = ctx.reads_mix_timing()
            tot["mix_copy_ms"] += mt["copy_ms"]; tot["mix_order_ms"] += mt["order_ms"]
            tot["mixed_bytes"] += mixed_bytes; tot["mixed_reads"] += n_mixed
        return n_mixed, mixed_bytes

    individual(first_hap, False, pair=False)   # warm-up: allocations, table upload, the three buffer sets the mix rotates
    individual(first_hap, False, pair=True)    # and the second batch state of the pair path
    torch.cuda.synchronize()
    l0 = ctx.kernel_launches
    t0 = time.perf_counter()
    for k in range(n_ind):
        n_mixed, mixed_bytes = individual(first_hap + 2 * k, True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    bases, k_ms = tot["bases"], tot["k_ms"]
    draws_per_base = float(used.astype(np.int64).sum()) / max(1, last["bases"])
    # docker/haplotype2reads.sh:59-63 (cat, shuf, tr) and the 10x/20x/30x prefixes of reads2svs.sh:65-128, last individual
    sweep = {f"{c}x": ctx.reads_mix_coverage_prefix(depth, c)["bytes"] for c in sorted({max(1, depth // 3), max(1, 2 * depth // 3), depth})}
    mix_obj = {"what": "cat + seeded shuf + tr 'Z' '\\n' of each individual's reads on the device, coverage prefixes as byte ranges",
               "individuals": n_ind, "reads": tot["mixed_reads"], "file_bytes": tot["mixed_bytes"], "order_ms": tot["mix_order_ms"], "copy_ms": tot["mix_copy_ms"],
               "copy_GBps_algorithmic": 2.0 * tot["mixed_bytes"] / (tot["mix_copy_ms"] / 1e3) / 1e9,
               "copy_frac_of_peak": 2.0 * tot["mixed_bytes"] / (tot["mix_copy_ms"] / 1e3) / 1e9 / peak,
               "coverage_prefix_bytes_last_individual": sweep}
    achieved = READ_ALGO_BYTES_PER_BASE * bases / (k_ms / 1e3) / 1e9
    # the bound that applies: warp instructions issued per second against 4 per clock per SM, with the
    # instructions per base of the committed ncu capture of the default kernel (profiles/r02m_reads_final_raw.csv: 46.86 G for 7.469 G bases)
    props = torch.cuda.get_device_properties(torch.cuda.current_device())
    ipb = 46856658240 / 7468692660
    peak_issue = props.multi_processor_count * 4 * 1.965e9
    rate = ipb * bases / (k_ms / 1e3)
    issue = {"bound": "issue", "achieved": rate / 1e12, "peak": peak_issue / 1e12, "unit": "T warp-inst/s",
             "frac": rate / peak_issue, "warp_instructions_per_base": round(ipb, 3), "note": "peak = SMs x 4 schedulers x 1965 MHz"}
    return {"metric": "CCS read Gbp/s", "value": bases / dt / 1e9, "unit": "Gbp/s",
            "workload": f"pbsim2 model mode, {depth}x over {2 * n_ind} materialised chr1-sized haplotypes ({n_ind} individuals, mixed per individual), "
                        f"length 15000+-2000 in [5000,30000], accuracy-mean 0.995, 6:50:54, synthetic 5-state HMM quality model ({quality} emission centres), per-read glibc streams",
            "timed": "per individual: genotype kernels, plan, materialisation, two read launches in flight together, mix; wall clock around all individuals",
            "draws_per_base": round(draws_per_base, 2),
            "reads": tot["reads"], "dropped_NNNNN": tot["dropped"], "bases": bases, "wall_s": dt, "kernel_ms": k_ms, "host_plan_ms": tot["p_ms"],
            "kernel_gbp_s": bases / (k_ms / 1e3) / 1e9, "gpu_launches": ctx.kernel_launches - l0,
            "kernel_ms_what": "per individual, the span of its two read kernels in flight together (sp_reads_mix_add_pair)",
            "one_by_one_kernel_ms_warmup_individual": one_by_one_ms,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": 20.9 * bases, "traffic_source": "20.9 B per base (126.9 GB read + 29.5 GB written for 7.469 Gbases), ncu of one launch of the default kernel (profiles/r02m_reads_final_raw.csv)",
                         "kernel": "reads_kernel", "algorithmic_bytes_per_base": READ_ALGO_BYTES_PER_BASE},
            "issue_roofline": issue, "mix": mix_obj}


def length_bins_leg(ctx, synth, work, peak, torch, cpu_baseline=True, genome=25_000_000, cov=30.0):
    """Row D (SampleReadsFromLengthBins, src/SampleReadsFromLengthBins.java:19-81) with the bins resident in HBM: draws by
    jump-ahead, in-bin ranks, cut at 2*coverage*genome bases, gather-copy of the records into one FASTQ (2 B per byte)."""
    n_bins, bl = 70, 500
    cum = synth.cumulative_bimodal(n_bins, bl, 9000, 1500, 15000, 2000, 0.3)
    prob = np.diff(np.concatenate(([0.0], cum)))
    target = int(genome * (cov * 2))
    mean_len = float(sum(prob[b] * (b + 0.5) * bl for b in range(n_bins)))
    per = [int(prob[b] * target / mean_len * 1.2) + 64 for b in range(n_bins)]
    # synthetic bins in BuildReadLengthBins' format (4-line records, "@r" names), assembled from one random pool
    rng = np.random.default_rng(7)
    pool_n = 1 << 24
    pool_seq = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, size=pool_n, dtype=np.uint8)]
    pool_q = rng.integers(33 + 20, 33 + 60, size=pool_n).astype(np.uint8)
    lens = [rng.integers(max(1, b * bl), (b + 1) * bl, size=per[b]) for b in range(n_bins)]
    total = sum(int((2 * l + 7).sum()) for l in lens)
    blob = np.empty(total, np.uint8)
    off = np.zeros(n_bins + 1, np.int64)
    head, mid = np.frombuffer(b"@r\n", np.uint8), np.frombuffer(b"\n+\n", np.uint8)
    o = 0
    for b in range(n_bins):
        for ln in lens[b]:
            ln = int(ln)
            a = int(rng.integers(0, pool_n - ln))
            blob[o:o + 3] = head; o += 3
            blob[o:o + ln] = pool_seq[a:a + ln]; o += ln
            blob[o:o + 3] = mid; o += 3
            blob[o:o + ln] = pool_q[a:a + ln]; o += ln
            blob[o] = 10; o += 1
        off[b + 1] = o
    ctx.length_bins_set(blob, off)
    res = None
    times = []
    for it in range(4):   # first pass = warm-up (allocations)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = ctx.length_bins_sample(cum, target)
        torch.cuda.synchronize()
        times.append(time.perf_counter() - t0)
        tm = ctx.length_bins_timing()
    wall = min(times[1:])
    achieved = 2.0 * res["fastq_bytes"] / (tm["copy_ms"] / 1e3) / 1e9
    out = {"metric": "length-bin read Gbp/s", "value": res["bases"] / wall / 1e9, "unit": "Gbp/s",
           "workload": f"SampleReadsFromLengthBins, 70 bins of 500 bp resident in HBM ({blob.nbytes} bytes), bimodal N(9000,1500)/N(15000,2000) "
                       f"weight 0.3, genome {genome} bp, coverage {cov} per haplotype",
           "reads": res["n_reads"], "bases": res["bases"], "fastq_bytes": res["fastq_bytes"], "attempts": res["attempts"],
           "wall_s": wall, "sample_ms": tm["sample_ms"], "copy_ms": tm["copy_ms"],
           "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                        "kernel": "materialise_kernel (record gather)", "algorithmic_bytes_per_output_byte": 2.0}}
    if cpu_baseline:
        import oracle_helpers as oh
        oh.ensure_oracle_built()
        exe = os.path.join(ROOT, "oracle", "_build", "synthpop_oracle")
        g2 = 2_000_000
        prefix = os.path.join(work, "bins_")
        synth.write_length_bins(prefix, n_bins, bl, [int(prob[b] * g2 * cov * 2 / mean_len * 1.3) + 8 for b in range(n_bins)], with_names=False)
        cdf = os.path.join(work, "cdf.txt")
        with open(cdf, "w") as f:
            for v in cum:
                f.write(repr(float(v)) + "\n")
        argv = ["9000", "1500", "15000", "2000", "0.3", str(bl), str(n_bins * bl), prefix, str(g2), str(cov), os.path.join(work, "cpu.fastq")]
        t0 = time.perf_counter()
        rc = subprocess.run([exe, "--seed", "42", "--cdf", cdf, "SampleReadsFromLengthBins"] + argv, stderr=subprocess.DEVNULL).returncode
        dt = time.perf_counter() - t0
        if rc == 0:
            out["cpu_baseline"] = {"value": g2 * cov * 2 / dt / 1e9, "unit": "Gbp/s", "cores": 1, "kind": "port",
                                   "sample": f"CPU restatement of SampleReadsFromLengthBins.main (files on local disk), genome {g2} bp at {cov}x per haplotype in {dt:.1f} s"}
    return out


def reads_cpu_baseline(synth, work, L=1_500_000, depth=30):
    """The reference's own docker/pbsim2.cpp (compiled unmodified into oracle/_ref/pbsim2) on a bounded sample: one core,
    as the reference runs it (one pbsim process per haplotype)."""
    exe = os.path.join(ROOT, "oracle", "_ref", "pbsim2")
    if not os.path.exists(exe):
        return None
    bases, _, _ = synth.random_reference(L, 7)
    fa = os.path.join(work, "cpu_reads.fa")
    synth.write_fasta(fa, ">chr1_0", bases)
    model = os.path.join(work, "P6C4_synthetic.model")
    if not os.path.exists(model):
        synth.write_hmm_model(model, quality="phred")
    t0 = time.perf_counter()
    subprocess.run([exe] + CCS_CLI + ["--depth", str(depth), "--hmm_model", model, "--seed", "42", "--prefix",
                                      os.path.join(work, "cpu_reads"), fa], check=True, stderr=subprocess.DEVNULL)
    dt = time.perf_counter() - t0
    n = depth * L
    return {"value": n / dt / 1e9, "unit": "Gbp/s", "cores": 1, "kind": "reference",
            "sample": f"oracle/_ref/pbsim2 (the reference's docker/pbsim2.cpp, unmodified), {depth}x over a {L} bp sequence in {dt:.1f} s"}


class ClockSampler(threading.Thread):
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md clocks line)."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self.stop_flag = False

    def run(self):
        try:   # NVML in-process: a sample every 20 ms (nvidia-smi takes ~0.4 s per call on these boxes)
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            bits = [(getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8), 2), (getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40), 3),
                    (getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20), 4), (getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4), 5)]
            while not self.stop_flag:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                row = [str(sm), str(mx), "", "", "", ""]
                for bit, col in bits:
                    row[col] = "Active" if (r & bit) else "Not Active"
                self.rows.append(row)
                time.sleep(0.02)
            return
        except Exception:
            pass
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        mx = [int(r[1]) for r in self.rows if r[1].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def cpu_baseline_sample(ref, pop, budget_s: float = 12.0, max_haps: int = 64):
    """The oracle's restatement of buildHaplotypes (single thread, as the reference is) on a bounded sample."""
    import oracle_helpers as oh
    olib = oh.load_oracle()
    t0 = time.perf_counter()
    bases = 0
    n = 0
    for h in range(max_haps):
        out = oh.build_haplotype(olib, ref, pop["sv_type"], pop["position"], pop["length"], pop["payload_off"],
                                 pop["payload"], None, pop["lists"][h])
        bases += out.size
        n += 1
        if time.perf_counter() - t0 > budget_s and n >= 2:
            break
    dt = time.perf_counter() - t0
    return bases / dt / 1e9, n, dt


_REF_ARM = None


def _ref_arm_one(h):
    """one haplotype of the reference arm in a forked worker (the arrays are inherited, not pickled)"""
    import oracle_helpers as oh
    ref, pop = _REF_ARM
    out = oh.build_haplotype(oh.load_oracle(), ref, pop["sv_type"], pop["position"], pop["length"], pop["payload_off"],
                             pop["payload"], None, pop["lists"][h])
    return int(out.size)


def workload_config(L: int, batch: int, n_variants: int, entries: int):
    """The `config` object: identical in both arms (the population is the same SimulateHaplotypes run: same generator seed for
    the inputs, same java.util.Random seed)."""
    return {"workload": f"chr1-sized reference ({L} bp), {N_HAPLOTYPES} haplotypes, genotype sampling + materialisation in batches of {batch}",
            "population": "SimulateHaplotypes (seed 42) on the seeded synthetic GRCh38-chr1-shaped inputs of synthpop.synth "
                          "(generator seed 20261104): AF threshold 0.01, 5000 frequent variants, 500 individuals",
            "n_variants": int(n_variants), "variants_per_haplotype": round(entries / N_HAPLOTYPES, 1),
            "l2": "inputs larger than L2 (>=31 GB written per step of 125 haplotypes, 249 MB per haplotype)"}


def parse_population_files(out_dir: str):
    """variants.txt + haplotype2variants.txt (src/SimulateHaplotypes.java:1123-1144) -> arrays for buildHaplotypes."""
    with open(os.path.join(out_dir, "variants.txt")) as f:
        rows = f.read().split("\n")
    n = int(rows[0]) + 1
    sv_type = np.zeros(n, np.int8); position = np.zeros(n, np.int32); length = np.zeros(n, np.int32)
    payload_off = np.zeros(n + 1, np.int64)
    chunks = []
    for i in range(n):
        r = rows[1 + 10 * i: 11 + 10 * i]
        sv_type[i] = int(r[0]); length[i] = int(r[2]); position[i] = int(r[4])
        seq = r[6]
        if seq != "null":
            chunks.append(seq.encode())
            payload_off[i + 1] = payload_off[i] + len(seq)
        else:
            payload_off[i + 1] = payload_off[i]
    payload = np.frombuffer(b"".join(chunks), np.uint8).copy() if chunks else np.zeros(1, np.uint8)
    lists = []
    with open(os.path.join(out_dir, "haplotype2variants.txt")) as f:
        H = int(f.readline())
        for _ in range(H):
            t = f.readline().rstrip("\n").split(",")
            lists.append(np.array([int(x) for x in t[1:2 + int(t[0])]], np.int32))
    return dict(sv_type=sv_type, position=position, length=length, payload_off=payload_off, payload=payload, lists=lists)


def reference_arm(args, L):
    """--impl reference: the reference's own CPU implementation of the path on the host cores. No JVM exists in this image,
    so it is the C++ restatement (oracle/, kind "port"): SimulateHaplotypes once (single process, as the reference runs
    it) on the SAME inputs and seed as the b200 arm, then buildHaplotypes in one process per core (the WDL scatters one
    JVM per individual, wdl/SimulateHaplotypes.wdl:263). A step = `cores` haplotypes; the genotype sampling + collision
    fixing the reference does for all 1 000 haplotypes in one sequential pass is charged per haplotype."""
    import multiprocessing as mp
    import tempfile
    import shutil
    from synthpop import synth
    import oracle_helpers as oh
    oh.ensure_oracle_built()
    cores = len(os.sched_getaffinity(0))
    work = tempfile.mkdtemp(prefix="synthpop_ref_arm_")
    inp = synth.generate_population_inputs(os.path.join(work, "in"), L, n_insertions=30000)
    out_dir = os.path.join(work, "out"); os.makedirs(out_dir)
    log = os.path.join(work, "draw.log")
    t0 = time.perf_counter()
    subprocess.run([os.path.join(ROOT, "oracle", "_build", "synthpop_oracle"), "--seed", "42", "--quiet", "--draw-log", log,
                    "SimulateHaplotypes"] + inp.simulate_argv(out_dir, N_HAPLOTYPES // 2, n_frequent=5000), check=True)
    t_sim = time.perf_counter() - t0
    phases = {k[len("seconds_"):]: float(v) for k, v in (l.split("=") for l in open(log).read().split()) if k.startswith("seconds_")}
    t_pop_per_hap = phases.get("buildPopulation", 0.0) / N_HAPLOTYPES
    pop = parse_population_files(out_dir)
    ref = inp.extra["bases"]
    entries = sum(l.size for l in pop["lists"])
    global _REF_ARM
    _REF_ARM = (ref, pop)
    vals, t_all = [], []
    with mp.get_context("fork").Pool(cores) as pool:
        h = 0
        for s in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            bases = sum(pool.map(_ref_arm_one, [(h + k) % N_HAPLOTYPES for k in range(cores)], chunksize=1))
            h += cores
            dt = time.perf_counter() - t0 + cores * t_pop_per_hap
            if s >= args.warmup:
                vals.append(bases / dt / 1e9)
                t_all.append(dt)
    v = float(np.mean(vals))
    line = {"metric": "simulated haplotype Gbp/s", "value": v, "unit": "Gbp/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(t_all)), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "impl": "reference",
            "config": workload_config(L, args.batch, pop["sv_type"].size, entries),
            "cpu_baseline": {"value": v, "unit": "Gbp/s", "cores": cores, "kind": "port",
                             "sample": f"{cores} chr1-sized haplotypes per step, one buildHaplotypes process per host core, plus "
                                       f"{1e3 * t_pop_per_hap:.2f} ms per haplotype of the single-threaded buildPopulation "
                                       f"(CPU restatement; the reference is Java and no JVM exists in this image)"},
            "e2e": {"value": v, "unit": "Gbp/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "prologue_s": {"total_SimulateHaplotypes": round(t_sim, 2), **{k: round(x, 3) for k, x in phases.items()}}}
    print(json.dumps(line))
    shutil.rmtree(work, ignore_errors=True)


def pcie_d2h_peak(torch, nbytes=1 << 30):
    """Measured device-to-host copy rate of this box into pinned memory (the bound of the end-to-end leg)."""
    d = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    h = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    best = 0.0
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); h.copy_(d, non_blocking=True); e1.record(); torch.cuda.synchronize()
        best = max(best, nbytes / (e0.elapsed_time(e1) / 1e3) / 1e9)
    return best


def file_emission_leg(splib, synth, work, local_rank):
    """BASELINE configs[0] (chr21-sized reference, 10 individuals) THROUGH THE FILES: SimulateHaplotypes with its six outputs
    and PrintHaplotypes of all 20 haplotypes (src/SimulateHaplotypes.java:113-115, src/PrintHaplotypes.java:6-18)."""
    Lc = 46_709_983
    t0 = time.perf_counter()
    inp = synth.generate_population_inputs(os.path.join(work, "c1_in"), Lc, n_insertions=30000)
    t_gen = time.perf_counter() - t0
    out = os.path.join(work, "c1_out"); os.makedirs(out)
    with splib.Context(local_rank, 42) as ctx:
        t0 = time.perf_counter()
        ctx.simulate_haplotypes(inp.simulate_argv(out, 10, n_frequent=5000))
        t_sim = time.perf_counter() - t0
        rows = ctx.population_timing()
        vcf_bytes = sum(os.path.getsize(os.path.join(out, f)) for f in os.listdir(out) if f.endswith(".vcf"))
        t0 = time.perf_counter()
        ctx.print_haplotypes(0, 19, inp.reference_file, os.path.join(out, "haplotype2variants.txt"), os.path.join(out, "variants.txt"), out)
        t_print = time.perf_counter() - t0
    fa_bytes = sum(os.path.getsize(os.path.join(out, f)) for f in os.listdir(out) if f.endswith(".fa"))
    return {"workload": "configs[0]: chr21-sized reference (46709983 bp), 10 individuals, files on local disk",
            "generate_inputs_s": round(t_gen, 2), "simulate_haplotypes_s": round(t_sim, 3), "rows_s": {k: round(v, 4) for k, v in rows.items()},
            "vcf_bytes": vcf_bytes, "vcf_MBps": vcf_bytes / max(rows["W_buildVCF"], 1e-9) / 1e6,
            "print_haplotypes_s": round(t_print, 3), "fasta_bytes": fa_bytes, "print_haplotypes_MBps": fa_bytes / t_print / 1e6}


def next_rows_leg(splib, synth, work, local_rank, emission_dir, cpu_baseline=True):
    """The rows either side of the path on the same box: AnnotateVCF over the truth VCFs
    the file-emission leg wrote, BuildReadLengthBins over a synthetic CCS-like FASTQ. Wall seconds through the C ABI with files
    on local disk, the device share beside them, the CPU restatement timed on the same inputs."""
    import oracle_helpers as oh
    oh.ensure_oracle_built()
    oracle = os.path.join(ROOT, "oracle", "_build", "synthpop_oracle")
    out = {}
    inp_dir = os.path.join(work, "c1_in")
    vcfs = sorted(os.path.join(emission_dir, f) for f in os.listdir(emission_dir) if f.endswith(".vcf") and "_annotated" not in f)
    lst = os.path.join(work, "annot_list.txt")
    open(lst, "w").write("".join(v + "\n" for v in vcfs))
    argv = [lst, "10", "100", os.path.join(inp_dir, "reference.fa"), os.path.join(inp_dir, "repeatIntervals.txt"), os.path.join(inp_dir, "segdups.txt")]
    with splib.Context(local_rank, 42) as ctx:
        t0 = time.perf_counter()
        ctx.annotate_vcf(argv)
        t = time.perf_counter() - t0
        tm = ctx.annotate_timing()
    nbytes = sum(os.path.getsize(v) for v in vcfs)
    out["annotate_vcf"] = {"workload": f"AnnotateVCF over the {len(vcfs)} truth VCFs of configs[0] ({nbytes} bytes), chr21-sized reference",
                           "wall_s": round(t, 3), "load_positions_s": round(tm["load_s"], 3), "records": tm["records"],
                           "records_per_s": tm["records"] / max(t - tm["load_s"], 1e-9), "last_file_kernel_ms": tm["kernel_ms"]}
    if cpu_baseline:
        t0 = time.perf_counter()
        subprocess.run([oracle, "AnnotateVCF"] + argv, check=True, stderr=subprocess.DEVNULL)
        out["annotate_vcf"]["cpu_port_wall_s"] = round(time.perf_counter() - t0, 3)
    # BuildReadLengthBins: 60 000 records around two modes, sequence + quality from a random pool
    rng = np.random.default_rng(11)
    pool = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, size=1 << 22, dtype=np.uint8)].tobytes()
    qual = rng.integers(33 + 20, 33 + 60, size=1 << 22).astype(np.uint8).tobytes()
    fq = os.path.join(work, "bins_in.fastq")
    n_rec = 60_000
    with open(fq, "wb") as f:
        for k in range(n_rec):
            ln = max(50, int(rng.normal(9000, 1500) if rng.random() < 0.3 else rng.normal(15000, 2000)))
            a = int(rng.integers(0, (1 << 22) - ln))
            f.write(b"@m64/%d/ccs\n" % k + pool[a:a + ln] + b"\n+\n" + qual[a:a + ln] + b"\n")
    lst2 = os.path.join(work, "bins_list.txt")
    open(lst2, "w").write(fq + "\n")
    argv = [lst2, "500", "35000", "3100000000", os.path.join(work, "bb_")]
    with splib.Context(local_rank, 42) as ctx:
        ctx.build_read_length_bins(argv)   # warm-up (allocations)
        t0 = time.perf_counter()
        ctx.build_read_length_bins(argv)
        t = time.perf_counter() - t0
        tm = ctx.build_bins_timing()
    fq_bytes = os.path.getsize(fq)
    out["build_read_length_bins"] = {"workload": f"BuildReadLengthBins, {n_rec} records, {fq_bytes} bytes of FASTQ, 70 bins of 500 bp",
                                     "wall_s": round(t, 3), "fastq_MBps_wall": fq_bytes / t / 1e6, "device_ms": {k: round(tm[k], 3) for k in ("index_ms", "shuffle_ms", "copy_ms")},
                                     "gather_GBps_algorithmic": 2.0 * tm["bytes"] / max(tm["copy_ms"], 1e-9) / 1e6}
    if cpu_baseline:
        argv_o = argv[:4] + [os.path.join(work, "bbo_")]
        t0 = time.perf_counter()
        subprocess.run([oracle, "--seed", "42", "BuildReadLengthBins"] + argv_o, check=True, stderr=subprocess.DEVNULL)
        out["build_read_length_bins"]["cpu_port_wall_s"] = round(time.perf_counter() - t0, 3)
    # BuildModel (rank 4): synthetic gnomAD-SV-shaped records and annotation tracks over the GRCh38 karyotype
    bm = synth.write_build_model_inputs(os.path.join(work, "bm_in"), n_records=120_000, n_repeats=300_000, n_trf=60_000, n_segdups=5_000)
    argv = [str(x) for x in bm["argv"]]
    with splib.Context(local_rank, 42) as ctx:
        t0 = time.perf_counter()
        ctx.build_model(argv)
        t = time.perf_counter() - t0
        tm = ctx.build_model_timing()
    out["build_model"] = {"workload": f"BuildModel, {bm['n_vcf_records']} VCF records, {bm['n_repeatmasker']} RepeatMasker + {bm['n_trf']} TRF + {bm['n_segdups']} segdup rows",
                          "wall_s": round(t, 3), "load_and_scan_s": round(tm["parse_s"], 3), "device_ms": round(tm["device_ms"], 3),
                          "float_passes_and_files_s": round(tm["finish_s"], 3), "records_counted": tm["records"], "allele_frequencies_sorted": tm["allele_frequencies"]}
    if cpu_baseline:
        odir = os.path.join(work, "bm_oracle")
        os.makedirs(odir, exist_ok=True)
        t0 = time.perf_counter()
        subprocess.run([oracle, "BuildModel"] + argv[:10] + [odir, "0"], check=True, stderr=subprocess.DEVNULL)
        out["build_model"]["cpu_port_wall_s"] = round(time.perf_counter() - t0, 3)
        out["build_model"]["files_identical_to_cpu_port"] = all(
            open(os.path.join(odir, f), "rb").read() == open(os.path.join(bm["out_dir"], f), "rb").read() for f in synth.BUILD_MODEL_OUTPUTS)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=160)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--batch", type=int, default=125, help="haplotypes per step (per GPU)")
    ap.add_argument("--length", type=int, default=CHR1)
    ap.add_argument("--e2e-haplotypes", type=int, default=16)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-overlap", action="store_true", help="genotype/collision/plan kernels on the materialisation stream")
    ap.add_argument("--read-haplotypes", type=int, default=-1,
                    help="haplotypes of the CCS read leg per GPU (0 = skip; default = configs[3]: 100 haplotypes over the GPUs of the run)")
    ap.add_argument("--read-depth", type=int, default=30)
    ap.add_argument("--length-bins", type=int, default=1, help="row D leg (SampleReadsFromLengthBins), 0 = skip")
    ap.add_argument("--file-emission", type=int, default=1, help="configs[0] through the files (rank 0, N=1), 0 = skip")
    ap.add_argument("--read-quality", default="phred", choices=["phred", "high"], help="emission centres of the synthetic HMM quality model")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    L = args.length
    if args.read_haplotypes < 0:
        args.read_haplotypes = max(2, (100 // max(world, 1)) // 2 * 2)   # configs[3]: 100 haplotypes, whole individuals per rank

    if args.impl == "reference":
        if rank == 0:
            reference_arm(args, L)
        return

    import tempfile
    import torch
    from synthpop import lib as splib, synth, sharded

    # cores (and, by first touch, pinned host buffers) on the GPU's NUMA node: the e2e leg's D2H streams must not cross sockets
    bound_cores = sharded.bind_rank_to_gpu(local_rank) if os.environ.get("SYNTHPOP_NO_BIND") is None else 0
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # synthetic GRCh38-chr1-shaped inputs in the reference's own file formats (seeded). One rank of the box writes them,
    # every rank reads them: the files are identical by construction and the host cores are shared.
    shared = os.path.join(tempfile.gettempdir(), f"synthpop_bench_{os.environ.get('MASTER_PORT', 'solo')}_{os.getppid() if world > 1 else os.getpid()}")
    work = tempfile.mkdtemp(prefix=f"synthpop_bench_r{rank}_")
    t_gen = time.perf_counter()
    if local_rank == 0:
        inp = synth.generate_population_inputs(os.path.join(shared, "in"), L, n_insertions=30000)
    barrier()
    if local_rank != 0:
        inp = synth.generate_population_inputs(os.path.join(shared, "in"), L, n_insertions=30000, reuse=True)
    t_gen = time.perf_counter() - t_gen
    out_dir = os.path.join(work, "out"); os.makedirs(out_dir)
    argv = inp.simulate_argv(out_dir, N_HAPLOTYPES // 2, n_frequent=5000)
    ctx = splib.Context(local_rank, 42)
    stream = torch.cuda.Stream()
    ctx.set_stream(stream.cuda_stream)
    pop_stream = torch.cuda.Stream(priority=-1)   # high priority, or its CTAs queue behind the 1.8 M tiles of the copy kernel

    # ---- (1) SimulateHaplotypes sharded by individual over the ranks, with the path's ONE exchange (all-gather of the
    # carrier counts, then of the lists for the joint outputs), checked against the same run done by this rank alone
    collective = None
    if dist is not None:
        warm = torch.zeros(world * 16, dtype=torch.int32, device="cuda")   # communicator set-up is not the exchange
        dist.all_gather_into_tensor(warm, warm[:16].clone())
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = sharded.simulate_population_sharded(ctx, argv, dist, write_files=False)
        t_sharded = time.perf_counter() - t0
        collective = dict(res["collective"], sharded_simulate_s=round(t_sharded, 3), shard=[int(res["first"]), int(res["last"])])
    # SimulateHaplotypes over all 1 000 haplotypes on this rank alone, untimed: loaders, variant sampling, genotype matrix +
    # collision fixer, insertion strings, repeat annotation (no file emission)
    t_pop = time.perf_counter()
    ctx.population_begin(argv, 0, -1)
    ctx.population_finish(None, None, None, 2)
    t_pop = time.perf_counter() - t_pop
    prologue_rows = ctx.population_timing()
    info = ctx.population_info()
    if dist is not None:
        counts_alone = ctx.population_local_counts()
        off_alone, ids_alone = ctx.get_haplotype_lists()
        ok = (np.array_equal(res["counts"], counts_alone) and np.array_equal(res["all_off"], off_alone) and
              np.array_equal(res["all_ids"], ids_alone))
        flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        collective["check"] = "ok" if int(flag[0]) == 1 else "MISMATCH"
        collective["what"] = ("ncclAllGather of int32 carrier counts [V] per rank (summed -> AF, isActive), then of the CSR lists; "
                              "global counts and every haplotype list equal the single-rank run of the same seed")
        if collective["check"] != "ok":
            print(json.dumps({"error": "sharded SimulateHaplotypes differs from the single-rank run", "collective": collective}))
            sys.exit(1)

    # ---- (2) the timed region: each rank owns its own batches of haplotypes (weak scaling: `batch` haplotypes per step per GPU)
    nb = N_HAPLOTYPES // args.batch

    def batch_range(step):
        b = (step + rank * 3) % nb
        return b * args.batch, b * args.batch + args.batch - 1

    launches0 = 0
    step_events = []
    bases_total = 0
    sampler = ClockSampler(local_rank)
    ctx.set_population_stream(None if args.no_overlap else pop_stream.cuda_stream)
    ev_region = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    t_geno = t_plan_host = 0.0
    last = None
    for s in range(args.warmup + args.steps):
        if s == args.warmup:
            barrier()
            sampler.start()
            launches0 = ctx.kernel_launches
            t_wall0 = time.perf_counter()
            ev_region[0].record(stream)
        a, b = batch_range(s)
        with torch.cuda.stream(stream):
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            tg = time.perf_counter()
            ctx.population_genotypes(a, b)            # rows G/C/N: genotype sampling for this batch (returns complete)
            tp = time.perf_counter()
            lens, stride = ctx.plan_haplotypes(a, b, False)   # row M: segment scan + tile table (planning stream)
            tq = time.perf_counter()
            ev[0].record(stream)
            ctx.materialise()                          # row M: tiled copy kernel
            ev[1].record(stream)
        if s >= args.warmup:
            step_events.append(ev)
            bases_total += int(lens.sum())
            t_geno += tp - tg
            t_plan_host += tq - tp
        last = (a, b, lens, stride)
    ev_region[1].record(stream)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    sampler.stop_flag = True
    launches = ctx.kernel_launches - launches0
    # the timed region: first event before step W's genotype call, last after step W+K-1's materialisation kernel
    t_step = ev_region[0].elapsed_time(ev_region[1]) / 1e3
    t_kernel = sum(e[0].elapsed_time(e[1]) for e in step_events) / 1e3
    mine = torch.tensor([t_step, t_kernel, float(bases_total), t_geno, t_plan_host], dtype=torch.float64, device="cuda")
    if dist is not None:
        allr = torch.empty(world * mine.numel(), dtype=torch.float64, device="cuda")
        dist.all_gather_into_tensor(allr, mine)
        allr = allr.view(world, -1).cpu().numpy()
    else:
        allr = mine.view(1, -1).cpu().numpy()
    t_step_max, t_kernel_max, bases_all = float(allr[:, 0].max()), float(allr[:, 1].max()), float(allr[:, 2].sum())
    value = bases_all / t_step_max / 1e9

    # ---- (3) parity spot check on the LAST timed batch: one haplotype back to the host, against the CPU restatement of
    # buildHaplotypes on the sampler's own lists (the batch's lists are still on the device)
    import oracle_helpers as oh
    a, b, lens, stride = last
    ctx.synchronize()
    pop = ctx.get_variants()
    off_b, ids_b = ctx.population_local_lists(b - a + 1)
    hsel = a + (rank * 7 + 3) % (b - a + 1)
    lst = ids_b[off_b[hsel - a]:off_b[hsel - a + 1]]
    got = ctx.copy_haplotype(hsel, stride)
    ref = inp.extra["bases"]
    want = oh.build_haplotype(oh.load_oracle(), ref, pop["sv_type"], pop["position"], pop["length"], pop["payload_off"],
                              pop["payload"], None, lst)
    spot_ok = got.size == want.size and np.array_equal(got, want)
    flag = torch.tensor([1 if spot_ok else 0], dtype=torch.int32, device="cuda")
    if dist is not None:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if int(flag[0]) != 1:
        if rank == 0:
            print(json.dumps({"error": "parity spot check failed: a haplotype of the last timed batch differs from the CPU restatement",
                              "haplotype": int(hsel), "rank_local_ok": bool(spot_ok)}))
        sys.exit(1)
    ctx.set_population_stream(None)
    pop["list_off"], pop["ids"] = ctx.get_haplotype_lists()
    pop["lists"] = [pop["ids"][pop["list_off"][h]:pop["list_off"][h + 1]] for h in range(64)]

    # CCS read leg (row P): every rank simulates reads over its own haplotypes; no collective
    reads_obj = None
    if args.read_haplotypes > 0:
        peak, peak_kind = measured_peak_gbs()
        a0 = (rank * args.read_haplotypes) % (N_HAPLOTYPES - args.read_haplotypes)
        reads_obj = reads_leg(ctx, splib, synth, work, a0, args.read_haplotypes, args.read_depth, peak, torch, args.read_quality)
        if dist is not None:
            tr = torch.tensor([reads_obj["wall_s"], reads_obj["kernel_ms"]], dtype=torch.float64, device="cuda")
            dist.all_reduce(tr, op=dist.ReduceOp.MAX)
            tb = torch.tensor([float(reads_obj["bases"])], dtype=torch.float64, device="cuda")
            dist.all_reduce(tb, op=dist.ReduceOp.SUM)
            reads_obj["value"] = float(tb[0]) / float(tr[0]) / 1e9
            reads_obj["wall_s"] = float(tr[0])

    # BASELINE configs[4] (row T, an extension: the reference's trios are real families): a child of two simulated parents,
    # its truth VCF, its 30x reads mixed and cut into the 10x/20x/30x prefixes
    trio_obj = None
    if args.read_haplotypes > 0 and rank == 0 and world == 1:
        peak, _ = measured_peak_gbs()
        t0 = time.perf_counter()
        child = ctx.trio_add_child(3, 8, 42, write_vcf=True)
        t_child = time.perf_counter() - t0
        h0 = child["child_first_haplotype"]
        r = reads_leg(ctx, splib, synth, work, h0, 2, args.read_depth, peak, torch, args.read_quality, resample=False)
        trio_obj = {"what": "sp_trio_add_child (one nextBoolean() per parent, truth VCF by the buildVCF merge) + the child's reads as in `reads` (one individual)",
                    "father_individual": 3, "mother_individual": 8, "child_haplotypes": [h0, h0 + 1], "inherited": [child["from_father_haplotype"], child["from_mother_haplotype"]],
                    "add_child_and_vcf_s": round(t_child, 4), "child_vcf_bytes": os.path.getsize(os.path.join(out_dir, f"groundTruth_individual_{h0}.vcf")),
                    "reads_gbp_s": r["value"], "reads_kernel_gbp_s": r["kernel_gbp_s"], "reads": r["reads"], "bases": r["bases"],
                    "mix": r["mix"], "parity": "unpinned by construction (nothing in the reference derives a child); every piece is the oracle-pinned row it reuses"}

    # row D leg (rank 0 only: the same work would run on every GPU)
    bins_obj = None
    if args.length_bins and rank == 0:
        peak, _ = measured_peak_gbs()
        bins_obj = length_bins_leg(ctx, synth, work, peak, torch, cpu_baseline=(world == 1 and not args.no_cpu_baseline))

    # e2e: the reference-facing call with HOST buffers — reference + variant table + lists H2D, plan, materialise, D2H
    ne = args.e2e_haplotypes
    d2h_peak = pcie_d2h_peak(torch) if ne > 0 else None
    host_out = torch.empty(int((L + (4 << 20)) * max(ne, 1)) if ne > 0 else 16, dtype=torch.uint8, pin_memory=True)
    host_np = host_out.numpy()
    ref_pinned = torch.from_numpy(ref).pin_memory().numpy()
    e2e_t, e2e_bases, h2d, d2h = [], 0, 0, 0
    for s in range((2 + 3) if ne > 0 else 0):
        barrier()
        t0 = time.perf_counter()
        ctx.set_reference(ref_pinned)
        ctx.set_variants(pop["sv_type"], pop["position"], pop["length"], pop["payload_off"], pop["payload"])
        ctx.set_haplotype_lists(pop["list_off"], pop["ids"])
        offs = ctx.build_haplotypes_host(s * ne, s * ne + ne - 1, host_np, False)
        barrier()
        if s >= 2:
            e2e_t.append(time.perf_counter() - t0)
            e2e_bases += int(offs[-1])
            h2d = int(ref.nbytes + pop["sv_type"].nbytes + pop["position"].nbytes + pop["length"].nbytes +
                      pop["payload_off"].nbytes + pop["payload"].nbytes + pop["list_off"].nbytes + pop["ids"].nbytes)
            d2h = int(offs[-1])
    e2e_val = torch.tensor([e2e_bases / max(sum(e2e_t), 1e-9) / 1e9], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(e2e_val, op=dist.ReduceOp.SUM)

    emission = next_rows = None
    if args.file_emission and rank == 0 and world == 1:
        emission = file_emission_leg(splib, synth, work, local_rank)
        next_rows = next_rows_leg(splib, synth, work, local_rank, os.path.join(work, "c1_out"), cpu_baseline=not args.no_cpu_baseline)

    if rank == 0:
        peak, peak_kind = measured_peak_gbs()
        per_launch_bases = bases_total / args.steps
        per_launch_s = t_kernel / args.steps
        algorithmic = ALGO_BYTES_PER_BP * per_launch_bases / per_launch_s / 1e9
        traffic, cap = measured_traffic(args.batch, L, per_launch_bases)
        # DRAM side of the launch: what ncu measured for this launch shape; without a matching capture, the bytes the kernel
        # cannot avoid (every output byte written once + every reference byte read once from HBM)
        dram_bytes = traffic if traffic is not None else per_launch_bases + L
        achieved = dram_bytes / per_launch_s / 1e9
        cfg = workload_config(L, args.batch, info["n_variants"], info["local_entries"])
        line = {"metric": "simulated haplotype Gbp/s", "value": value, "unit": "Gbp/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": 1e3 * t_step_max / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": cfg,
                "arm": {"haplotypes_per_step_per_gpu": args.batch,
                        "step": "genotype kernel + collision fixer + segment scan + tile table + materialisation kernel for one batch",
                        "overlap": "none" if args.no_overlap else "rows G/C/N and the plan of batch s+1 on a second stream under the materialisation kernel of batch s",
                        "sharding": "by individual; the one all-gather (carrier counts) runs before the timed region, see `collective`"},
                "gpu_launches": int(launches),
                "parity_spot_check": "ok",
                "collective": collective,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": traffic, "traffic_source": ("ncu capture " + cap["capture"]) if cap else "no capture of this launch shape: output bytes + reference bytes",
                             "kernel": "materialise_kernel", "peak_kind": peak_kind,
                             "what": "DRAM bytes per launch / average launch duration (CUDA events on the launching stream) / measured copy peak",
                             "algorithmic_bytes_per_bp": ALGO_BYTES_PER_BP, "algorithmic_achieved": algorithmic,
                             "algorithmic_frac": algorithmic / peak,
                             "l2_reuse": f"every reference byte is read by {args.batch} haplotypes of the launch; region-major tile order serves {args.batch - 1} of them from L2",
                             "kernel_ms_per_step": 1e3 * per_launch_s, "bases_per_launch": per_launch_bases},
                "per_rank": {"ms_per_step": [1e3 * x / args.steps for x in allr[:, 0]],
                             "kernel_ms_per_step": [1e3 * x / args.steps for x in allr[:, 1]],
                             "host_ms_in_population_genotypes": [1e3 * x / args.steps for x in allr[:, 3]],
                             "host_ms_in_plan_haplotypes": [1e3 * x / args.steps for x in allr[:, 4]],
                             "kernel_ms_per_step_max": 1e3 * t_kernel_max / args.steps},
                "prologue": {"generate_inputs_s": round(t_gen, 2), "simulate_population_once_s": round(t_pop, 3),
                             "rows_s": {k: round(v, 4) for k, v in prologue_rows.items()},
                             "note": "rows A1-A3, S1, I, R1 are the replicated sequential prologue (host); G/C/N is the device part over all 1000 haplotypes"},
                "e2e": {"value": float(e2e_val[0]), "unit": "Gbp/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "haplotypes_per_call": ne, "d2h_peak_GBps_measured": d2h_peak,
                        "frac_of_d2h_peak": (float(e2e_val[0]) / world / d2h_peak) if d2h_peak else None,
                        "cores_bound_to_gpu_numa_node_rank0": bound_cores},
                "clocks": sampler.summary(), "wall_s_timed_region": t_wall}
        if reads_obj is not None:
            line["reads"] = reads_obj
        if bins_obj is not None:
            line["length_bins"] = bins_obj
        if trio_obj is not None:
            line["trio"] = trio_obj
        if emission is not None:
            line["file_emission"] = emission
        if next_rows is not None:
            line["next_rows"] = next_rows
        if world == 1 and not args.no_cpu_baseline:
            if reads_obj is not None:
                rb = reads_cpu_baseline(synth, work)
                if rb is not None:
                    reads_obj["cpu_baseline"] = rb
            v, n, dt = cpu_baseline_sample(ref, pop)
            line["cpu_baseline"] = {"value": v, "unit": "Gbp/s", "cores": 1, "kind": "port",
                                    "sample": f"{n} chr1-sized haplotypes in {dt:.1f} s (CPU restatement of buildHaplotypes, 1 thread)"}
        print(json.dumps(line))
    ctx.close()
    import shutil
    shutil.rmtree(work, ignore_errors=True)
    barrier()
    if local_rank == 0:
        shutil.rmtree(shared, ignore_errors=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
