"""The N>1 host path on CPU: world_size-2 gloo processes exercising the shard split, the strand-aware gather
to rank 0 and the no-collective alternative (each rank pwrite()s its row slabs of one HDF5 file)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from expecto_b200 import hdf5, shard

F = 2002


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fake_rows(n):
    """[2n, F] in chromatin.py's layout: value encodes (strand, variant, feature)."""
    v = np.arange(n, dtype=np.float32)[:, None] * 10 + np.arange(F, dtype=np.float32)[None, :] / 4096
    return np.concatenate([v, -v - 1], axis=0)


def _worker(rank, world, port, n, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    r, w, _ = shard.init_distributed("gloo")
    assert (r, w) == (rank, world)
    full = _fake_rows(n)
    lo, hi = shard.shard_range(n, rank, world)
    local = torch.from_numpy(np.concatenate([full[lo:hi], full[n + lo:n + hi]], axis=0))
    got = shard.gather_strand_rows(local, n)
    if rank == 0:
        assert got is not None and np.array_equal(got.numpy(), full)
    else:
        assert got is None
    plain = shard.gather_rows(torch.from_numpy(full[lo:hi].copy()), n)
    if rank == 0:
        assert np.array_equal(plain.numpy(), full[:n])
    assert shard.allreduce_sum_ints([rank + 1, 10, hi - lo]) == [sum(range(1, world + 1)), 10 * world, n]
    # no-collective mode: rank 0 creates, the others attach and write their own slabs
    path = os.path.join(tmp, "slabs.h5")
    shapes = {"diff": (2 * n, F)}
    if rank == 0:
        wtr = hdf5.H5Writer(path, shapes)
    shard.barrier()
    if rank != 0:
        wtr = hdf5.H5Writer(path, shapes, create=False)
    wtr.write_rows("diff", lo, full[lo:hi])
    wtr.write_rows("diff", n + lo, full[n + lo:n + hi])
    wtr.close()
    shard.barrier()
    if rank == 0:
        with hdf5.File(path) as f:
            assert np.array_equal(np.asarray(f["diff"]), full)
    dist.destroy_process_group()


@pytest.mark.parametrize("n", [7, 8, 1])
def test_two_rank_gather_and_slab_write(tmp_path, n):
    mp.spawn(_worker, args=(2, _free_port(), n, str(tmp_path)), nprocs=2, join=True)


def test_shard_bounds_cover_rows_in_order():
    for n in (0, 1, 7, 8, 1000003):
        for world in (1, 2, 4, 8):
            b = shard.shard_bounds(n, world)
            assert b[0] == 0 and b[-1] == n and np.all(np.diff(b) >= 0)
            assert np.diff(b).max() - np.diff(b).min() <= 1
            assert [shard.shard_range(n, r, world) for r in range(world)] == list(zip(b[:-1], b[1:]))


def test_attach_rejects_mismatched_file(tmp_path):
    p = str(tmp_path / "a.h5")
    hdf5.H5Writer(p, {"diff": (4, F)}).close()
    with pytest.raises(ValueError):
        hdf5.H5Writer(p, {"diff": (6, F)}, create=False)


def _features_worker(rank, world, port, n_genes, tmp, gather="nccl"):
    """compute_expecto_features.run under two gloo ranks with the GPU part stubbed: gene shards, gather, one .npy."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from expecto_b200 import compute_expecto_features as cef

    genes = [(f"ENSG{i:05d}", "chr1", 1000 + 37 * i, 1 if i % 2 else -1) for i in range(n_genes)]

    def fake_features(engine, genome, part, shifts, weights, genes_per_batch, mode="segments"):
        tss = np.array([g[2] for g in part], dtype=np.float64)
        return (tss[:, None] + np.arange(20020, dtype=np.float64)[None, :] / 32).astype(np.float32).astype(np.float64)

    class Genome:
        offsets = {"chr1": 0}

    class Engine:
        device = None
        _genome = None

    g, e = Genome(), Engine()
    e._genome = g
    cef.gene_features = fake_features
    cef.read_genes = lambda a, t, h=False: genes
    shard.init_distributed("gloo")
    args = cef.build_parser().parse_args(["anno.csv", "tss.tsv", "-o", tmp, "--gather", gather])
    path = cef.run(args, engine=e, genome=g)
    if rank == 0:
        got = np.load(path)
        want = fake_features(None, None, genes, None, None, 0)
        assert got.dtype == np.float64 and got.shape == (n_genes, 20020)
        assert np.array_equal(got, want)
        ref = os.path.join(tmp, "np_save.npy")          # the bytes np.save gives for the same matrix
        np.save(ref, want)
        assert open(path, "rb").read() == open(ref, "rb").read()
    dist.destroy_process_group()


@pytest.mark.parametrize("gather", ["nccl", "file"])
@pytest.mark.parametrize("n_genes", [5, 1])
def test_two_rank_feature_shards_gather_to_one_file(tmp_path, n_genes, gather):
    mp.spawn(_features_worker, args=(2, _free_port(), n_genes, str(tmp_path), gather), nprocs=2, join=True)


def _genome_worker(rank, world, port, fa):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    shard.init_distributed("gloo")
    g = shard.load_genome(fa)
    # every rank ends up with the same packed arrays; only rank 0 parsed the FASTA (the others found the cache)
    sums = shard.allreduce_sum_ints([int(g.bases2.sum() % (1 << 31)), int(g.nmask.sum() % (1 << 31)), g.n_bases])
    assert sums == [world * int(g.bases2.sum() % (1 << 31)), world * int(g.nmask.sum() % (1 << 31)), world * g.n_bases]
    assert os.path.exists(fa + ".bx2.npz")
    dist.destroy_process_group()


def test_genome_is_packed_once_and_shared_through_the_cache(tmp_path):
    from expecto_b200 import synth
    fa = str(tmp_path / "g.fa")
    synth.write_fasta(fa, synth.make_genome([("chr1", 50000), ("chr2", 30000)], seed=4), width=60)
    mp.spawn(_genome_worker, args=(2, _free_port(), fa), nprocs=2, join=True)


class _FakeEngine:
    """Stands in for BelugaEngine's variant-plan calls on CPU tensors: value = f(tensor, shift, strand, global variant)."""

    def __init__(self, lo, chunk=5):
        self.device, self.lo, self.chunk = torch.device("cpu"), lo, chunk

    def chromatin_open(self, cols, maxshift, delta=True, chunk_variants=0):
        self.n = cols["n"]
        if chunk_variants:
            self.chunk = min(self.chunk, chunk_variants)
        return (self.n + self.chunk - 1) // self.chunk, self.chunk

    def chromatin_chunk_info(self, k):
        return {"v0": k * self.chunk, "m": min(self.chunk, self.n - k * self.chunk), "delta": 0, "full": 0, "window": 0}

    def chromatin_submit_device(self, k, ref, alt, diff, shift_stride, strand_stride, row0):
        info = self.chromatin_chunk_info(k)
        for j, t in enumerate((ref, alt, diff)):
            for si in range(3):
                for strand in range(2):
                    for v in range(info["m"]):
                        t[si * shift_stride + strand * strand_stride + row0 + v] = _fake_value(j, si, strand, self.lo + info["v0"] + v)

    def chromatin_close(self):
        pass


def _fake_value(j, si, strand, g):
    return float(1000 * j + 100 * si + 10 * strand) + g / 64.0


def _gather_worker(rank, world, port, n, tmp):
    from expecto_b200 import chromatin
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    shard.init_distributed("gloo")
    lo, hi = shard.shard_range(n, rank, world)
    eng = _FakeEngine(lo, chunk=5 if rank == 0 else 4)      # ranks must settle on the smaller chunk size
    got = {s: {k: np.full((2 * n, F), np.nan, np.float32) for k in ("ref", "alt", "diff")} for s in (0, -200, 200)}
    sinks = None
    if rank == 0:
        def make(s):
            def sink(name, row0, arr):
                got[s][name][row0:row0 + len(arr)] = arr
            return sink
        sinks = {s: make(s) for s in got}
    info = chromatin.predict_all_shifts_gathered(eng, {"n": hi - lo}, 200, n, sinks)
    assert info["chunk_variants"] == 4
    if rank == 0:
        for si, s in enumerate((0, -200, 200)):
            for j, name in enumerate(("ref", "alt", "diff")):
                for strand in range(2):
                    want = np.array([_fake_value(j, si, strand, g) for g in range(n)], np.float32)
                    assert np.array_equal(got[s][name][strand * n:(strand + 1) * n, 0], want), (s, name, strand)
                    assert np.array_equal(got[s][name][strand * n:(strand + 1) * n, F - 1], want)
    shard.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n", [23, 8, 1])
def test_streaming_gather_places_every_rank_chunk_in_global_row_order(tmp_path, n):
    """predict_all_shifts_gathered's host logic with world_size 2 over gloo: uneven shards (one rank may even be
    empty), different chunk sizes on the ranks, partial last chunks, forward rows at g and reverse-complement rows at
    n + g of the whole job (chromatin.py:281-286's layout)."""
    mp.spawn(_gather_worker, args=(2, _free_port(), n, str(tmp_path)), nprocs=2, join=True)
