"""torchrun worker of tests/test_sharded_multi_gpu.py: one SimulateHaplotypes run sharded over the ranks (NCCL all-gather
of the carrier counts), then every rank prints its own individuals' haplotypes. argv: out_dir, then the 11 positional
arguments of SimulateHaplotypes."""
import os
import sys

import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "population-sv-simulator_b200"))
from synthpop import lib, sharded  # noqa: E402


def main():
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    hap_dir, argv = sys.argv[1], sys.argv[2:]
    ctx = lib.Context(local_rank, 42)
    if os.environ.get("SYNTHPOP_TEST_OWNER_MODE"):
        # one owner (the LAST rank) runs the whole of SimulateHaplotypes and writes its files; the others receive the
        # reference and the tables (sharded.share_population) and materialise their own individuals
        rank, world = dist.get_rank(), dist.get_world_size()
        owner = world - 1
        first, last = sharded.shard_of_haplotypes(int(argv[9]), rank, world)
        if rank == owner:
            ctx.simulate_haplotypes(argv)
        f, l, moved = sharded.share_population(ctx, dist, owner, first, last)
        assert moved["reference_bytes"] > 0 and moved["n_variants"] > 0
        lens, stride = ctx.plan_haplotypes(f, l, True)
        ctx.materialise()
        for h in range(first, last + 1):
            data = ctx.copy_haplotype(f + (h - first), stride).tobytes()
            if rank != owner:   # a receiver numbers its haplotypes from 0: the FASTA header says so too
                head, _, rest = data.partition(b"\n")
                assert head == b">chr1_%d" % (f + (h - first))
                data = b">chr1_%d\n" % h + rest
            with open(os.path.join(hap_dir, f"haplotype_{h}.fa"), "wb") as fh:
                fh.write(data)
        dist.barrier()
        ctx.close()
        dist.destroy_process_group()
        return
    res = sharded.simulate_population_sharded(ctx, argv, dist)
    # each rank materialises its own haplotypes from the lists still resident on its GPU
    first, last = res["first"], res["last"]
    lens, stride = ctx.plan_haplotypes(first, last, True)
    ctx.materialise()
    for h in range(first, last + 1):
        with open(os.path.join(hap_dir, f"haplotype_{h}.fa"), "wb") as f:
            f.write(ctx.copy_haplotype(h, stride).tobytes())
    dist.barrier()
    ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
