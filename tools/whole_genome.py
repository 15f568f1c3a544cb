"""BASELINE config 3: a whole-genome GRCh38-shaped population (24 chromosomes, 3.09 Gbp), 100 diploid individuals
sharded by individual over the GPUs of one box.

The reference simulates one sequence per run (single FASTA record, `int` coordinates, src/SimulateHaplotypes.java:
120-135), so a genome is 24 independent runs of SimulateHaplotypes + PrintHaplotypes; the WDL scatters individuals
(wdl/SimulateHaplotypes.wdl:142-163). Here every rank runs the 24 chromosomes in turn on its own individuals
(synthpop.sharded: 13/13/13/13/12/12/12/12 at 8 ranks), with the one all-gather of carrier counts per chromosome.

  python tools/whole_genome.py                                  one GPU plays rank 0 of 8 (its 13 individuals)
  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/whole_genome.py

Prints one JSON line: per-chromosome device times (CUDA events on the library's stream) and the whole-genome
haplotype Gbp/s = bases all ranks materialised / max over ranks of the summed device time; beside it the WALL seconds of the
whole job after input generation (the chromosomes' host prologues run in threads, the exchanges in chromosome order).
The inputs are written once per box (rank r writes chromosomes r, r+world, ...).
"""
import argparse
import json
import os
import shutil
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "population-sv-simulator_b200"))

GRCH38 = [("chr1", 248956422), ("chr2", 242193529), ("chr3", 198295559), ("chr4", 190214555), ("chr5", 181538259),
          ("chr6", 170805979), ("chr7", 159345973), ("chr8", 145138636), ("chr9", 138394717), ("chr10", 133797422),
          ("chr11", 135086622), ("chr12", 133275309), ("chr13", 114364328), ("chr14", 107043718), ("chr15", 101991189),
          ("chr16", 90338345), ("chr17", 83257441), ("chr18", 80373285), ("chr19", 58617616), ("chr20", 64444167),
          ("chr21", 46709983), ("chr22", 50818468), ("chrX", 156040895), ("chrY", 57227415)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--individuals", type=int, default=100)
    ap.add_argument("--as-world", type=int, default=8, help="single process: play rank --as-rank of this many ranks")
    ap.add_argument("--as-rank", type=int, default=0)
    ap.add_argument("--scale", type=float, default=1.0, help="shrink every chromosome (smoke runs)")
    ap.add_argument("--chromosomes", type=int, default=len(GRCH38))
    ap.add_argument("--only", default="", help="comma-separated chromosome names (profiling one launch shape)")
    ap.add_argument("--repeats", type=int, default=3, help="materialisation passes per chromosome (first is warm-up)")
    ap.add_argument("--threads", type=int, default=0, help="host threads for the chromosomes' prologues (0 = cores / ranks, at most 8)")
    ap.add_argument("--keep-inputs", action="store_true")
    ap.add_argument("--mode", choices=("owner", "sharded"), default="owner",
                    help="multi-rank runs: 'owner' = each chromosome's SimulateHaplotypes runs on one rank, which broadcasts the reference "
                         "and the tables (sharded.share_population); 'sharded' = every rank runs every prologue on its own columns")
    args = ap.parse_args()

    import torch
    from synthpop import lib as splib, sharded, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    sharded.bind_rank_to_gpu(local_rank)
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    stream = torch.cuda.Stream()
    rows = []
    dev_ms = 0.0
    bases_total = 0
    launches = 0
    only = set(filter(None, args.only.split(",")))
    chroms = [(name, max(200_000, int(full * args.scale))) for name, full in GRCH38[: args.chromosomes] if not only or name in only]

    # ---- inputs: every file once per box (rank r writes chromosomes r, r+world, ...), not once per rank
    shared = os.path.join(tempfile.gettempdir(), f"synthpop_wg_shared_{os.environ.get('MASTER_PORT', '0')}_{args.scale}")
    t0 = time.perf_counter()
    for k, (name, L) in enumerate(chroms):
        if k % world == rank:
            synth.generate_population_inputs(os.path.join(shared, name), L, n_insertions=30000)
    if dist is not None:
        dist.barrier()
    t_gen_total = time.perf_counter() - t0

    def inputs_of(name, L):   # the paths generate_population_inputs wrote (no arrays are rebuilt)
        root = os.path.join(shared, name)
        return synth.SynthInputs(root=root, reference_file=os.path.join(root, "reference.fa"), model_dir=os.path.join(root, "model"),
                                 repeat_intervals_file=os.path.join(root, "repeatIntervals.txt"), segdups_file=os.path.join(root, "segdups.txt"),
                                 insertions_file=os.path.join(root, "insertions.txt"), n_insertions=30000, reference_length=L)

    if dist is not None:
        first, last = sharded.shard_of_haplotypes(args.individuals, rank, world)
    else:
        first, last = sharded.shard_of_haplotypes(args.individuals, args.as_rank, args.as_world)

    def begin(name, L, seed0=42):
        """The prologue of one chromosome's SimulateHaplotypes (rows A1-A3, S1, G/C/N of this rank's columns): host work
        with no collective in it, so the chromosomes' prologues run side by side in threads (the C ABI releases the GIL)."""
        argv = inputs_of(name, L).simulate_argv(os.path.join(shared, name, f"out_r{rank}"), args.individuals, n_frequent=5000)
        t0 = time.perf_counter()
        # The reference throws on some (inputs, seed) pairs (e.g. ArrayIndexOutOfBoundsException in its interval
        # intersection when the pair buffer grows, src/SimulateHaplotypes.java:548-555) and the library reproduces
        # that; the prologue is the same on every rank, so every rank moves on to the next seed together.
        for seed in range(seed0, 50):
            ctx = splib.Context(local_rank, seed)
            try:
                if dist is not None:
                    ctx.population_begin(argv, first, last)
                else:
                    # one process cannot all-gather with absent ranks: draw the whole population here (the genotype
                    # kernel over all 200 haplotypes is microseconds) and materialise only the shard's individuals
                    ctx.population_begin(argv, 0, -1)
                ctx.synchronize()
                return ctx, argv, seed, time.perf_counter() - t0
            except splib.SynthpopError as e:
                ctx.close()
                if e.code != 6 or seed == 49:
                    raise

    from concurrent.futures import ThreadPoolExecutor
    n_threads = args.threads if args.threads > 0 else max(1, (os.cpu_count() or 8) // max(1, world if dist is not None else 1))
    n_threads = min(n_threads, 8)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    t_wall0 = time.perf_counter()
    t_pop_total = 0.0
    owner_mode = dist is not None and args.mode == "owner"

    def begin_whole(name, L):
        """Owner mode: the whole of SimulateHaplotypes for one chromosome on this rank (all haplotypes, no files)."""
        argv = inputs_of(name, L).simulate_argv(os.path.join(shared, name, f"out_r{rank}"), args.individuals, n_frequent=5000)
        t0 = time.perf_counter()
        for seed in range(42, 50):
            ctx = splib.Context(local_rank, seed)
            try:
                ctx.population_begin(argv, 0, -1)
                ctx.population_finish(None, None, None, 2)
                ctx.synchronize()
                return ctx, argv, seed, time.perf_counter() - t0
            except splib.SynthpopError as e:
                ctx.close()
                if e.code != 6 or seed == 49:
                    raise

    def materialise_passes(ctx, f, l):
        best = None
        for it in range(args.repeats):
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            with torch.cuda.stream(stream):
                ev[0].record(stream)
                lens, stride = ctx.plan_haplotypes(f, l, False)
                ev[1].record(stream)
                ctx.materialise()
                ev[2].record(stream)
            torch.cuda.synchronize()
            t = (ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]))
            if it > 0 or args.repeats == 1:
                best = t if best is None or sum(t) < sum(best) else best
        return best, lens, stride

    moved_bytes = 0
    if owner_mode:
        recv = splib.Context(local_rank, 42)   # one receiving context for all the chromosomes this rank does not own
        recv.set_stream(stream.cuda_stream)
        with ThreadPoolExecutor(n_threads) as pool:
            futures = {k: pool.submit(begin_whole, name, L) for k, (name, L) in enumerate(chroms) if k % world == rank}
            for k, (name, L) in enumerate(chroms):   # broadcasts and launches in chromosome order on every rank
                owner = k % world
                t_begin, seed = 0.0, -1
                t0 = time.perf_counter()
                if owner == rank:
                    ctx, argv, seed, t_begin = futures[k].result()
                    t0 = time.perf_counter()
                    ctx.set_stream(stream.cuda_stream)
                else:
                    ctx = recv
                f, l, moved = sharded.share_population(ctx, dist, owner, first, last)
                moved_bytes += moved["tables_bytes"] + moved["reference_bytes"]
                t_share = time.perf_counter() - t0
                l0 = ctx.kernel_launches
                best, lens, stride = materialise_passes(ctx, f, l)
                launches += (ctx.kernel_launches - l0) // args.repeats
                assert lens.size == last - first + 1 and int(lens.min()) > 0 and stride >= int(lens.max())
                bases = int(lens.sum())
                rows.append({"chromosome": name, "length": L, "variants": moved["n_variants"], "owner": owner, "seed": seed, "haplotypes": int(lens.size),
                             "bases": bases, "plan_ms": round(best[0], 3), "copy_ms": round(best[1], 3),
                             "prologue_s": round(t_begin, 3), "share_s": round(t_share, 3)})
                dev_ms += sum(best)
                bases_total += bases
                t_pop_total += t_begin + t_share
                if owner == rank:
                    ctx.close()
        recv.close()
    else:
        with ThreadPoolExecutor(n_threads) as pool:
            futures = [pool.submit(begin, name, L) for name, L in chroms]
            for (name, L), fut in zip(chroms, futures):   # exchanges and launches in chromosome order on every rank
                ctx, argv, seed, t_begin = fut.result()
                t0 = time.perf_counter()
                while True:
                    try:
                        ctx.set_stream(stream.cuda_stream)
                        if dist is not None:
                            sharded.simulate_population_sharded(ctx, argv, dist, write_files=False, begun=True)
                        else:
                            ctx.population_finish(None, None, None, 2)
                        break
                    except splib.SynthpopError as e:   # rows I/R1 can throw too; same on every rank: next seed, together
                        ctx.close()
                        if e.code != 6 or seed >= 49:
                            raise
                        ctx, argv, seed, _ = begin(name, L, seed + 1)
                ctx.synchronize()
                t_pop = t_begin + time.perf_counter() - t0
                info = ctx.population_info()
                best = None
                l0 = ctx.kernel_launches
                for it in range(args.repeats):
                    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
                    with torch.cuda.stream(stream):
                        ev[0].record(stream)
                        lens, stride = ctx.plan_haplotypes(first, last, False)
                        ev[1].record(stream)
                        ctx.materialise()
                        ev[2].record(stream)
                    torch.cuda.synchronize()
                    t = (ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]))
                    if it > 0 or args.repeats == 1:
                        best = t if best is None or sum(t) < sum(best) else best
                launches += (ctx.kernel_launches - l0) // args.repeats
                # size-independent property: every haplotype is the reference plus its own net edit, and differs from it
                assert lens.size == last - first + 1 and int(lens.min()) > 0 and stride >= int(lens.max())
                bases = int(lens.sum())
                rows.append({"chromosome": name, "length": L, "variants": info["n_variants"], "seed": seed, "haplotypes": int(lens.size),
                             "bases": bases, "plan_ms": round(best[0], 3), "copy_ms": round(best[1], 3),
                             "prologue_s": round(t_begin, 3), "exchange_and_finish_s": round(t_pop - t_begin, 3)})
                dev_ms += sum(best)
                bases_total += bases
                t_pop_total += t_pop
                ctx.close()
    torch.cuda.synchronize()
    wall_s = time.perf_counter() - t_wall0
    if dist is not None:
        dist.barrier()
    if rank == 0 and not args.keep_inputs:
        shutil.rmtree(shared, ignore_errors=True)

    if dist is not None:
        t = torch.tensor([dev_ms, t_pop_total, wall_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        b = torch.tensor([bases_total], dtype=torch.int64, device="cuda")
        dist.all_reduce(b, op=dist.ReduceOp.SUM)
        dev_ms, t_pop_total, wall_s, bases_total = float(t[0]), float(t[1]), float(t[2]), int(b[0])
    if rank == 0:
        n_ranks = world if dist is not None else 1
        line = {"metric": "simulated haplotype Gbp/s", "value": bases_total / dev_ms / 1e6, "unit": "Gbp/s", "n_gpus": n_ranks,
                "config": {"workload": f"whole genome, {len(rows)} GRCh38-sized chromosomes (scale {args.scale}), "
                                       f"{args.individuals} diploid individuals sharded by individual over "
                                       f"{world if dist is not None else args.as_world} ranks"
                                       + ("" if dist is not None else f"; this process is rank {args.as_rank} alone")},
                "bases": bases_total, "device_ms": dev_ms,
                "wall_s": wall_s, "wall_what": f"max over ranks, from the barrier after input generation to the last chromosome materialised "
                                               f"({args.repeats} plan+copy passes each), prologues in {n_threads} host threads per rank",
                "wall_gbp_s": bases_total / wall_s / 1e9,
                "simulate_population_s_summed_over_chromosomes": t_pop_total, "prologue_threads": n_threads,
                "mode": ("owner" if owner_mode else "sharded"), "broadcast_bytes_per_rank": moved_bytes,
                "generate_inputs_s": t_gen_total, "gpu_launches": launches, "timing": "CUDA events, max over ranks of "
                "the summed plan+copy time, best of the passes after the warm-up one", "chromosomes": rows}
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
