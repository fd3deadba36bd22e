"""Drop-in for the reference's ``replicate_expecto_features.py`` (same CLI, same per-gene ``.npy`` dumps) on B200s.

    python -m expecto_b200.replicate_expecto_features resources/geneanno.csv [-o out_dir]

Reference behaviour reproduced (file:line into /root/reference/replicate_expecto_features.py):
  * arguments ``annoFile --windowsize --cuda -o``                                                   :17-24
  * every line of annoFile must split into 7 comma-separated fields, the first line is the header   :39-42
  * 200 shifts -20000..19800 around ``CAGE_representative_TSS``, window ``tss + shift*strand - 999 .. + 1000``
    (1-based inclusive, not reverse complemented for '-' genes)                                     :44-45,65-71
  * ``0.5 * (prediction + rc_prediction)`` in float32, saved as ``{out}/{gene_id}.npy`` [200,2002]   :75-86
Same windows as compute_expecto_features, so the conv stack runs once per gene and strand (bx_forward_segments,
bit-identical to the per-window path).  Under torchrun the genes are split into contiguous per-rank shards and every
rank writes the files of its own genes (no collective).
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Replicate ExPecto chromatin features")
    p.add_argument("annoFile")
    p.add_argument("--windowsize", action="store", dest="windowsize", type=int, default=2000)
    p.add_argument("--cuda", action="store_true", help="accepted for compatibility (always on)")
    p.add_argument("-o", dest="out_dir", type=str, default="temp_replicate_expecto_features")
    # extensions
    p.add_argument("--genome", default="./resources/hg19.fa")
    p.add_argument("--weights", default="./resources/deepsea.beluga.pth")
    p.add_argument("--dtype", default="f16", choices=["f16", "bf16"])
    p.add_argument("--device", type=int, default=0)
    p.add_argument("--genes_per_batch", type=int, default=32, help="genes per GPU batch (x400 windows)")
    p.add_argument("--mode", default="segments", choices=["segments", "windows"])
    return p


def read_genes(anno_file: str):
    """:39-42 -> list of (gene_id, chrom, CAGE TSS, strand +-1)."""
    genes = []
    for i, line in enumerate(open(anno_file)):
        gene_id, _symbol, chrom, strand, _tss, cage_tss, _type = line.rstrip().split(",")
        if i > 0:
            genes.append((gene_id, chrom, int(cage_tss), 1 if strand == "+" else -1))
    return genes


def gene_predictions(engine, genome, genes, shifts, genes_per_batch: int = 32, mode: str = "segments"):
    """Yields (index, float32 [S,2002]) per gene in order: ``0.5 * (forward + reverse-complement prediction)``."""
    import torch

    from .windows import check_valid_bases, tss_descs, tss_segment_descs

    S = len(shifts)
    no_alleles = np.zeros(1, dtype=np.uint8)
    steps = np.diff(np.asarray(shifts))
    use_segments = mode == "segments" and 1 < S <= 200 and bool(np.all(steps == 200))
    B = max(1, min(genes_per_batch, max(1, len(genes))))
    pinned = [torch.empty((B, S, 2002), dtype=torch.float32, pin_memory=True) for _ in range(2)]
    pending = None

    def collect(p):
        ev, hb, at, m = p
        ev.synchronize()
        for j in range(m):
            yield at + j, hb[j].numpy().copy()

    for bi, g0 in enumerate(range(0, len(genes), B)):
        part = genes[g0:g0 + B]
        cols = ([g[1] for g in part], [g[2] for g in part], [g[3] for g in part])
        if use_segments:
            descs, k_of, seg_len = tss_segment_descs(genome, *cols, shifts)
            check_valid_bases(genome, descs)
            probs = engine.forward_segments(descs, no_alleles, len(part), 2, seg_len, k_of, shift_major=False)
        else:
            descs = tss_descs(genome, *cols, shifts)
            check_valid_bases(genome, descs)
            probs = engine.forward_windows(descs, no_alleles)
        probs = probs.view(len(part), S, 2, 2002)
        mean = 0.5 * (probs[:, :, 0] + probs[:, :, 1])
        hb = pinned[bi % 2]
        hb[:len(part)].copy_(mean, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(engine.device))
        if pending is not None:
            yield from collect(pending)
        pending = (ev, hb, g0, len(part))
        del probs, mean
    if pending is not None:
        yield from collect(pending)


def run(args, engine=None, genome=None) -> int:
    """The whole script; returns the number of files this process wrote."""
    from . import shard
    from .transform import TSS_SHIFTS

    if args.windowsize != 2000:
        raise ValueError("Beluga's architecture fixes --windowsize at 2000 (Beluga.py:43)")
    rank, world, local = shard.init_distributed()
    os.makedirs(args.out_dir, exist_ok=True)
    if genome is None:
        genome = shard.load_genome(args.genome, device=engine.device if engine is not None else
                                   (local if world > 1 else args.device))
    if engine is None:
        from .engine import BelugaEngine
        engine = BelugaEngine(local if world > 1 else args.device, dtype=args.dtype, capacity=4096)
        engine.load_weights_file(args.weights)
    if engine._genome is not genome:
        engine.set_genome(genome)
    genes = read_genes(args.annoFile)
    missing = sorted({g[1] for g in genes if g[1] not in genome.offsets})
    if missing:
        raise KeyError(f"contig(s) {missing} are not in the genome")
    lo, hi = shard.shard_range(len(genes), rank, world)
    mine = genes[lo:hi]
    n = 0
    for i, pred in gene_predictions(engine, genome, mine, TSS_SHIFTS, args.genes_per_batch, args.mode):
        np.save(f"{args.out_dir}/{mine[i][0]}", pred)
        n += 1
    shard.barrier()
    return n


def main(argv=None) -> int:
    from . import shard
    run(build_parser().parse_args(argv))
    shard.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
