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


# ---- one owner per run: pack/unpack of the tables and the three broadcasts of share_population -------------------------

class _FakeContext:
    """The seven Context methods share_population uses, on host arrays."""

    def __init__(self, population=None):
        self.pop = population
        self.received = {}

    def get_haplotype_lists(self):
        return self.pop["list_off"], self.pop["list_ids"]

    def get_variants(self):
        return {k: self.pop[k] for k in ("sv_type", "position", "length", "payload_off", "payload")}

    @property
    def reference_length(self):
        return int(self.pop["ref"].size)

    @property
    def host_reference(self):
        return self.pop["ref"]

    def set_reference(self, bases):
        self.received["ref"] = np.array(bases, copy=True)

    def set_variants(self, sv_type, position, length, payload_off, payload):
        self.received.update(sv_type=sv_type, position=position, length=length, payload_off=payload_off, payload=payload)

    def set_haplotype_lists(self, off, ids):
        self.received.update(list_off=off, list_ids=ids)


def _fake_population(seed=5, V=40, H=10):
    rng = np.random.default_rng(seed)
    plen = rng.integers(0, 30, size=V)
    lists = [np.sort(rng.choice(V, size=int(rng.integers(0, 9)), replace=False)).astype(np.int32) for _ in range(H)]
    return dict(ref=rng.integers(65, 85, size=5000).astype(np.uint8), sv_type=rng.integers(0, 5, size=V).astype(np.int8),
                position=np.sort(rng.integers(0, 5000, size=V)).astype(np.int32), length=rng.integers(1, 300, size=V).astype(np.int32),
                payload_off=np.concatenate(([0], np.cumsum(plen))).astype(np.int64), payload=rng.integers(65, 85, size=int(plen.sum())).astype(np.uint8),
                list_off=np.concatenate(([0], np.cumsum([x.size for x in lists]))).astype(np.int64),
                list_ids=(np.concatenate(lists) if lists else np.zeros(0)).astype(np.int32))


def test_pack_tables_round_trip():
    p = _fake_population()
    v, off, ids = sharded.unpack_tables(sharded.pack_tables({k: p[k] for k in ("sv_type", "position", "length", "payload_off", "payload")}, p["list_off"], p["list_ids"]))
    for k in ("sv_type", "position", "length", "payload_off", "payload"):
        assert v[k].dtype == p[k].dtype and np.array_equal(v[k], p[k]), k
    assert np.array_equal(off, p["list_off"]) and np.array_equal(ids, p["list_ids"])
    o, i = sharded.take_lists(p["list_off"], p["list_ids"], 4, 7)
    assert o[0] == 0 and o.size == 5 and np.array_equal(i, p["list_ids"][p["list_off"][4]:p["list_off"][8]])
    o, i = sharded.take_lists(p["list_off"], p["list_ids"], 0, -1)
    assert o.size == 1 and i.size == 0


def _share_worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    got = {}
    for run, owner in enumerate([0, 1, 0]):   # three runs, owners alternate, broadcasts in run order on every rank
        pop = _fake_population(seed=100 + run)
        ctx = _FakeContext(pop if rank == owner else None)
        first, last = sharded.shard_of_haplotypes(5, rank, world)
        f, l, moved = sharded.share_population(ctx, dist, owner, first, last)
        got[f"run{run}"] = np.array([f, l, moved["tables_bytes"], moved["reference_bytes"]])
        for k, v in ctx.received.items():
            got[f"run{run}_{k}"] = np.asarray(v)
    np.savez(os.path.join(out_dir, f"share_r{rank}.npz"), **got)
    dist.destroy_process_group()


def test_share_population_broadcasts_under_gloo(tmp_path):
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_share_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    for rank in range(world):
        z = np.load(tmp_path / f"share_r{rank}.npz")
        first, last = sharded.shard_of_haplotypes(5, rank, world)
        for run, owner in enumerate([0, 1, 0]):
            pop = _fake_population(seed=100 + run)
            f, l, tb, rb = z[f"run{run}"]
            assert rb == pop["ref"].size and tb > 0
            if rank == owner:
                assert (f, l) == (first, last) and f"run{run}_ref" not in z.files   # the owner keeps its population
                continue
            assert (f, l) == (0, last - first)
            assert np.array_equal(z[f"run{run}_ref"], pop["ref"])
            for k in ("sv_type", "position", "length", "payload_off", "payload"):
                assert np.array_equal(z[f"run{run}_{k}"], pop[k]), (rank, run, k)
            off, ids = sharded.take_lists(pop["list_off"], pop["list_ids"], first, last)
            assert np.array_equal(z[f"run{run}_list_off"], off) and np.array_equal(z[f"run{run}_list_ids"], ids)
