"""Sharding of one SimulateHaplotypes run across the GPUs of a box: one process per GPU,
`torch.distributed` for the plumbing.

Individuals (pairs of consecutive haplotype ids) are split in contiguous blocks, the way the reference's WDL scatters
them (wdl/SimulateHaplotypes.wdl:142-163). The prologue (loadReference, model, loadPositions, sampleVariants) is a pure
function of (inputs, seed), so every rank computes the same variant table and nothing is broadcast. The genotype cells
are index-addressable (src/SimulateHaplotypes.java:621-634: cell (j,i) uses LCG steps S1+2(j*H+i)), so a rank draws only
its own columns. The ONE exchange is after the collision fixer: per-variant carrier counts (int32[V]) are all-gathered
and summed (AF_j := carriers_j / H, :986; isActive, :693-706) and the per-haplotype lists are all-gathered for the joint
VCF's HAPLOTYPES= field and haplotype2variants.txt, which rank 0 writes. Materialisation and reads need no traffic.
"""
from __future__ import annotations

import numpy as np


def shard_of_individuals(n_individuals: int, rank: int, world: int) -> tuple[int, int]:
    """Individuals [a, b) owned by `rank`: the first (n mod world) ranks get one more (100 over 8 -> 13,13,13,13,12,...)."""
    base, extra = divmod(n_individuals, world)
    a = rank * base + min(rank, extra)
    return a, a + base + (1 if rank < extra else 0)


def shard_of_haplotypes(n_individuals: int, rank: int, world: int) -> tuple[int, int]:
    """Inclusive haplotype range [first, last] of `rank` (haplotypes 2k and 2k+1 belong to individual k); (0, -1) if empty."""
    a, b = shard_of_individuals(n_individuals, rank, world)
    return (2 * a, 2 * b - 1) if b > a else (0, -1)


def allgather_counts(local_counts, dist, group=None):
    """Sum over ranks of the per-variant carrier counts through ONE all-gather (a tensor on the backend's device:
    CUDA for nccl, CPU for gloo). Returns the global counts as a tensor on the same device."""
    import torch
    world = dist.get_world_size(group)
    gathered = torch.empty(world * local_counts.numel(), dtype=local_counts.dtype, device=local_counts.device)
    dist.all_gather_into_tensor(gathered, local_counts.contiguous().view(-1), group=group)
    return gathered.view(world, -1).sum(dim=0, dtype=torch.int32)


def allgather_lists(local_off: np.ndarray, local_ids: np.ndarray, dist, device, group=None):
    """All-gather of the ranks' CSR haplotype lists (rank order = haplotype order). Returns (all_off, all_ids) as numpy."""
    import torch
    world = dist.get_world_size(group)
    n_h = int(local_off.size - 1)
    meta = torch.tensor([n_h, int(local_ids.size)], dtype=torch.int64, device=device)
    metas = torch.empty(world * 2, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(metas, meta, group=group)
    metas = metas.view(world, 2).cpu().numpy()
    max_h, max_ids = int(metas[:, 0].max()), int(metas[:, 1].max())
    lens = torch.zeros(max(max_h, 1), dtype=torch.int64, device=device)
    if n_h:
        lens[:n_h] = torch.from_numpy(np.diff(local_off).astype(np.int64)).to(device)
    ids = torch.zeros(max(max_ids, 1), dtype=torch.int32, device=device)
    if local_ids.size:
        ids[:local_ids.size] = torch.from_numpy(np.ascontiguousarray(local_ids, dtype=np.int32)).to(device)
    all_lens = torch.empty(world * lens.numel(), dtype=torch.int64, device=device)
    all_ids = torch.empty(world * ids.numel(), dtype=torch.int32, device=device)
    dist.all_gather_into_tensor(all_lens, lens, group=group)
    dist.all_gather_into_tensor(all_ids, ids, group=group)
    all_lens, all_ids = all_lens.view(world, -1).cpu().numpy(), all_ids.view(world, -1).cpu().numpy()
    lens_cat = np.concatenate([all_lens[r, : metas[r, 0]] for r in range(world)]) if world else np.zeros(0, np.int64)
    ids_cat = np.concatenate([all_ids[r, : metas[r, 1]] for r in range(world)]) if world else np.zeros(0, np.int32)
    off = np.concatenate(([0], np.cumsum(lens_cat))).astype(np.int64)
    return off, ids_cat.astype(np.int32)


class _DeviceInt32:
    """__cuda_array_interface__ view of a device buffer the library owns (no copy)."""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i4", "data": (ptr, False), "version": 3, "strides": None}


def simulate_population_sharded(ctx, argv, dist=None, write_files: bool = True, begun: bool = False):
    """SimulateHaplotypes.main (src/SimulateHaplotypes.java:75-116) across the ranks of `dist` (None = single process).
    Every rank calls this with the same argv and a context created with the same seed. Returns a dict with the rank's
    haplotype range and the global counts. Afterwards the context holds the rank's haplotype lists, ready for
    sp_plan_haplotypes / the read kernel on the rank's own individuals.
    begun=True: the caller has already run ctx.population_begin(argv, first, last) for this rank's shard (the prologue has no
    collective in it, so several runs' prologues may proceed in host threads while the exchanges stay in one order)."""
    import torch
    rank = dist.get_rank() if dist is not None else 0
    world = dist.get_world_size() if dist is not None else 1
    n_individuals = int(argv[9])
    if n_individuals < world:
        raise ValueError(f"{n_individuals} individuals cannot be sharded over {world} ranks")
    first, last = shard_of_haplotypes(n_individuals, rank, world)
    if not begun:
        ctx.population_begin(argv, first, last)
    info = ctx.population_info()
    n_local = max(0, last - first + 1)
    if dist is None or world == 1:
        ctx.population_finish(None, None, None, 1 if write_files else 2)
        return dict(first=first, last=last, counts=ctx.population_local_counts(), n_variants=info["n_variants"])
    nccl = dist.get_backend() == "nccl"
    device = torch.device("cuda", torch.cuda.current_device()) if nccl else torch.device("cpu")
    V = info["n_variants"]
    if nccl:
        ctx.synchronize()
        local = torch.as_tensor(_DeviceInt32(ctx.population_counts_device(), V), device=device)  # zero-copy: the library's buffer
    else:
        local = torch.from_numpy(ctx.population_local_counts().copy())
    # the ONE exchange of the path (north_star): all-gather of the per-variant carrier counts, timed on the device
    ev = None
    if nccl:
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        ev[0].record()
    counts_t = allgather_counts(local, dist)
    if nccl:
        ev[1].record()
    counts = counts_t.cpu().numpy().astype(np.int32)
    off, ids = ctx.population_local_lists(n_local)
    all_off, all_ids = allgather_lists(off, ids, dist, device)
    timing = dict(counts_bytes_per_rank=int(V) * 4, lists_bytes_total=int(all_ids.size) * 4 + int(all_off.size) * 8)
    if nccl:
        ev[2].record()
        torch.cuda.synchronize()
        timing.update(counts_allgather_ms=ev[0].elapsed_time(ev[1]), lists_allgather_ms=ev[1].elapsed_time(ev[2]))
    if write_files:
        ctx.population_finish(counts, all_off if rank == 0 else None, all_ids if rank == 0 else None, 1 if rank == 0 else 0)
    else:
        ctx.population_finish(counts, None, None, 2)
    return dict(first=first, last=last, counts=counts, n_variants=V, all_off=all_off, all_ids=all_ids, collective=timing)


# ---- one owner per run (whole genome: 24 independent runs of SimulateHaplotypes, BASELINE configs[2]) ----------------
# A genome is 24 runs, one per chromosome (the reference takes one FASTA record per run, src/SimulateHaplotypes.java:120-135).
# Sharding the individuals of every run over the ranks (above) makes every rank repeat every run's sequential prologue
# (loadReference, the model, loadPositions, sampleVariants): 15 s of host time per rank for GRCh38. Here a run has ONE owner
# that does the whole of SimulateHaplotypes for it (the genotype kernel over all haplotypes is microseconds) and then
# broadcasts what materialisation needs (rows A1, V0 and the lists) to the other GPUs: the reference stays on the device and
# goes over NVLink straight into the receivers' contexts, the tables travel as one packed byte tensor.

_TABLE_FIELDS = (("sv_type", np.int8), ("position", np.int32), ("length", np.int32), ("payload_off", np.int64),
                 ("payload", np.uint8), ("list_off", np.int64), ("list_ids", np.int32))


def pack_tables(variants: dict, list_off: np.ndarray, list_ids: np.ndarray) -> np.ndarray:
    """Variant table (Context.get_variants) + CSR haplotype lists as one uint8 array: a header of 7 int64 element counts,
    then the arrays, each padded to 8 bytes."""
    arrays = dict(variants, list_off=list_off, list_ids=list_ids)
    parts = [np.array([np.asarray(arrays[k]).size for k, _ in _TABLE_FIELDS], dtype=np.int64).view(np.uint8)]
    for k, dt in _TABLE_FIELDS:
        b = np.ascontiguousarray(arrays[k], dtype=dt).view(np.uint8).ravel()
        parts += [b, np.zeros((-b.size) % 8, dtype=np.uint8)]
    return np.concatenate(parts)


def unpack_tables(buf: np.ndarray):
    """Inverse of pack_tables: (variants dict, list_off, list_ids)."""
    buf = np.ascontiguousarray(buf, dtype=np.uint8)
    counts = buf[: 8 * len(_TABLE_FIELDS)].view(np.int64)
    at = 8 * len(_TABLE_FIELDS)
    out = {}
    for (k, dt), n in zip(_TABLE_FIELDS, counts):
        nb = int(n) * np.dtype(dt).itemsize
        out[k] = buf[at: at + nb].view(dt).copy()
        at += nb + (-nb) % 8
    list_off, list_ids = out.pop("list_off"), out.pop("list_ids")
    return out, list_off, list_ids


def take_lists(list_off: np.ndarray, list_ids: np.ndarray, first: int, last: int):
    """The CSR rows [first, last] renumbered from 0."""
    if last < first:
        return np.zeros(1, np.int64), np.zeros(0, np.int32)
    a, b = int(list_off[first]), int(list_off[last + 1])
    return (list_off[first: last + 2] - a).astype(np.int64), list_ids[a:b].astype(np.int32)


class _DeviceBytes:
    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}


def share_population(ctx, dist, owner: int, first: int, last: int):
    """After the owner's ctx finished a whole-population SimulateHaplotypes (population_begin(argv, 0, -1) +
    population_finish), give every other rank's (empty) ctx what it needs to materialise ITS haplotypes [first, last]:
    three broadcasts from `owner` (sizes, the packed tables, the reference device-to-device). On return a non-owner holds
    the reference, the variant table and its own lists as haplotypes 0..last-first; the owner keeps its population and
    addresses its shard as [first, last]. Returns (first, last) to pass to plan_haplotypes on this rank, and the bytes moved."""
    import torch
    rank = dist.get_rank()
    nccl = dist.get_backend() == "nccl"
    device = torch.device("cuda", torch.cuda.current_device()) if nccl else torch.device("cpu")
    if rank == owner:
        off, ids = ctx.get_haplotype_lists()
        packed = pack_tables(ctx.get_variants(), off, ids)
        ref_ptr, ref_len = ctx.reference_device() if nccl else (0, ctx.reference_length)
        meta = torch.tensor([packed.size, ref_len], dtype=torch.int64, device=device)
    else:
        meta = torch.zeros(2, dtype=torch.int64, device=device)
    dist.broadcast(meta, src=owner)
    n_packed, ref_len = (int(x) for x in meta.cpu())
    tables = torch.from_numpy(packed).to(device) if rank == owner else torch.empty(n_packed, dtype=torch.uint8, device=device)
    dist.broadcast(tables, src=owner)
    if nccl:
        ptr = ref_ptr if rank == owner else ctx.reference_device_buffer(ref_len)
        ref_t = torch.as_tensor(_DeviceBytes(ptr, ref_len), device=device)   # the library's own buffer on both sides
        dist.broadcast(ref_t, src=owner)
        torch.cuda.current_stream().synchronize()
    else:   # gloo (CPU tests of this logic): the reference travels as host bytes
        ref_t = torch.from_numpy(np.ascontiguousarray(ctx.host_reference)) if rank == owner else torch.empty(ref_len, dtype=torch.uint8)
        dist.broadcast(ref_t, src=owner)
        if rank != owner:
            ctx.set_reference(ref_t.numpy())
    moved = dict(tables_bytes=n_packed, reference_bytes=ref_len)
    if rank == owner:
        moved["n_variants"] = int(packed[:8].view(np.int64)[0])
        return first, last, moved
    variants, off, ids = unpack_tables(tables.cpu().numpy())
    moved["n_variants"] = int(variants["sv_type"].size)
    ctx.set_variants(variants["sv_type"], variants["position"], variants["length"], variants["payload_off"], variants["payload"])
    my_off, my_ids = take_lists(off, ids, first, last)
    ctx.set_haplotype_lists(my_off, my_ids)
    return 0, last - first, moved


def bind_rank_to_gpu(local_rank: int):
    """Pin this process to the CPU cores the driver reports as local to its GPU (NVML cpu affinity), so that pinned host
    buffers allocated afterwards are on the GPU's NUMA node and the D2H streams of the ranks of one box do not cross sockets.
    Returns the number of cores bound, or 0 when NVML is not usable (nothing is changed then)."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        visible = os.environ.get("CUDA_VISIBLE_DEVICES")
        index = int(visible.split(",")[local_rank]) if visible and visible.split(",")[local_rank].isdigit() else local_rank
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        allowed = os.sched_getaffinity(0)
        cpus = {64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1} & allowed
        if not cpus or cpus == allowed:
            return 0
        os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception:
        return 0
