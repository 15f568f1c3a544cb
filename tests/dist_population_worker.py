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
