"""Data-parallel sharding of variants / genes over the GPUs of one box.

Every window is independent and the decay transforms couple only the shifts of ONE variant / gene, so the
path shards by row with no collective inside the forward.  The reference shards by hand the same way
(``--chunk_i/--chunk_size``, chromatin.py:213-214; per-chromosome shell loops).  One process per GPU
(``torch.distributed``: NCCL on GPUs, gloo in the CPU tests); ranks take contiguous, order-preserving row
ranges so that concatenating rank outputs reproduces the reference's row order.  The only communication is
the end-of-shift gather of the ref/alt/diff (or feature) rows to rank 0 — or none at all when every rank
pwrite()s its own row slab of the contiguous HDF5 datasets (``H5Writer.write_rows``).
"""
from __future__ import annotations

import os

import numpy as np
import torch
import torch.distributed as dist


def world_info() -> tuple[int, int, int]:
    """(rank, world_size, local_rank) from the torchrun environment (1-process defaults)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def init_distributed(backend: str | None = None) -> tuple[int, int, int]:
    """Join the process group when launched under torchrun; no-op for a single process."""
    rank, world, local = world_info()
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank, world, local


def shard_bounds(n: int, world: int) -> np.ndarray:
    """Row boundaries [world+1] of the contiguous equal-count split (first ``n % world`` ranks get one more)."""
    base, extra = divmod(int(n), int(world))
    counts = np.full(world, base, dtype=np.int64)
    counts[:extra] += 1
    return np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    b = shard_bounds(n, world)
    return int(b[rank]), int(b[rank + 1])


def gather_rows(local: torch.Tensor, n_total: int, dst: int = 0) -> torch.Tensor | None:
    """Gather the ranks' contiguous row shards ``[n_local, F]`` into ``[n_total, F]`` on ``dst``.
    Shards are padded to the largest count so one fixed-size gather (NCCL send/recv group over NVLink, or
    gloo) moves them; returns None on the other ranks."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        assert local.shape[0] == n_total
        return local
    world, rank = dist.get_world_size(), dist.get_rank()
    bounds = shard_bounds(n_total, world)
    counts = np.diff(bounds)
    if local.shape[0] != counts[rank]:
        raise ValueError(f"rank {rank} holds {local.shape[0]} rows, expected {counts[rank]}")
    cmax = int(counts.max())
    feat = tuple(local.shape[1:])
    padded = local
    if local.shape[0] != cmax:
        padded = torch.zeros((cmax,) + feat, dtype=local.dtype, device=local.device)
        padded[:local.shape[0]] = local
    padded = padded.contiguous()
    if rank == dst:
        bufs = [torch.empty_like(padded) for _ in range(world)]
        dist.gather(padded, bufs, dst=dst)
        return torch.cat([bufs[r][:counts[r]] for r in range(world)], dim=0)
    dist.gather(padded, None, dst=dst)
    return None


def gather_strand_rows(local: torch.Tensor, n_total: int, dst: int = 0) -> torch.Tensor | None:
    """chromatin.py's [2N, F] layout (forward rows 0..N-1, reverse-complement rows N..2N-1,
    chromatin.py:170-171) from per-rank ``[2*n_local, F]`` tensors laid out the same way: the two strand
    halves are gathered separately so that rank order == row order inside each half."""
    n_local = local.shape[0] // 2
    fwd = gather_rows(local[:n_local], n_total, dst)
    rc = gather_rows(local[n_local:], n_total, dst)
    if fwd is None:
        return None
    return torch.cat([fwd, rc], dim=0)


def allreduce_sum_ints(values, device=None) -> list[int]:
    """Sum small integer counters (the allele-match diagnostics of chromatin.py:256-260) over ranks."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return [int(v) for v in values]
    t = torch.tensor(list(values), dtype=torch.int64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [int(v) for v in t.tolist()]


def barrier() -> None:
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


def load_genome(path: str, device=None):
    """``PackedGenome.from_fasta`` for every rank without packing the FASTA ``world`` times: rank 0 packs it first
    (on its GPU when ``device`` is given) and writes the ``.bx2.npz`` cache next to it, atomically; the other ranks
    wait at a barrier and then read the cache — or, where the cache could not be written (read-only directory), pack
    the FASTA themselves."""
    from .genome import PackedGenome

    if not dist.is_initialized() or dist.get_world_size() == 1:
        return PackedGenome.from_fasta(path, device=device)
    genome = PackedGenome.from_fasta(path, device=device) if dist.get_rank() == 0 else None
    barrier()
    return genome if genome is not None else PackedGenome.from_fasta(path, device=device)


def finalize() -> None:
    """End of a script's ``main()``: leave the process group cleanly (NCCL warns about leaked resources otherwise)."""
    if dist.is_initialized():
        dist.destroy_process_group()
