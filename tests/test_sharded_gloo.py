"""Host-side logic of the multi-GPU path (synthpop/sharded.py) on CPU: world_size-2 gloo processes exchange fake
shards (deterministic per-rank counts and CSR lists) through the same all-gather helpers the NCCL path uses."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from synthpop import sharded


def test_shard_ranges_cover_everything_once():
    for n, w in [(100, 8), (10, 3), (8, 8), (1000, 7), (5, 1)]:
        seen = []
        for r in range(w):
            a, b = sharded.shard_of_individuals(n, r, w)
            seen += list(range(a, b))
            f, l = sharded.shard_of_haplotypes(n, r, w)
            assert (f, l) == ((2 * a, 2 * b - 1) if b > a else (0, -1))
        assert seen == list(range(n))
    sizes = [sharded.shard_of_individuals(100, r, 8) for r in range(8)]
    assert [b - a for a, b in sizes] == [13, 13, 13, 13, 12, 12, 12, 12]


def _fake_shard(rank, world, n_individuals=7, V=50):
    rng = np.random.default_rng(1234)  # same stream on every rank: the "replicated prologue"
    lists = [np.sort(rng.choice(V, size=int(rng.integers(0, 12)), replace=False)).astype(np.int32) for _ in range(2 * n_individuals)]
    f, l = sharded.shard_of_haplotypes(n_individuals, rank, world)
    mine = lists[f:l + 1]
    off = np.concatenate(([0], np.cumsum([x.size for x in mine]))).astype(np.int64)
    ids = np.concatenate(mine).astype(np.int32) if mine else np.zeros(0, np.int32)
    counts = np.bincount(ids, minlength=V).astype(np.int32)
    return lists, off, ids, counts


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lists, off, ids, counts = _fake_shard(rank, world)
    g = sharded.allgather_counts(torch.from_numpy(counts), dist).numpy()
    all_off, all_ids = sharded.allgather_lists(off, ids, dist, torch.device("cpu"))
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), g=g, off=all_off, ids=all_ids)
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_allgather_of_counts_and_lists(tmp_path, world):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    lists, _, _, _ = _fake_shard(0, world)
    want_ids = np.concatenate(lists)
    want_off = np.concatenate(([0], np.cumsum([x.size for x in lists])))
    want_counts = np.bincount(want_ids, minlength=50)
    for r in range(world):
        z = np.load(tmp_path / f"r{r}.npz")
        assert np.array_equal(z["g"], want_counts)
        assert np.array_equal(z["off"], want_off) and np.array_equal(z["ids"], want_ids)
