"""Drop-in for the reference's ``compute_expecto_features.py`` (same CLI, same ``.npy`` out) on B200s.

    python -m expecto_b200.compute_expecto_features resources/geneanno.csv tss.tsv [-o out_dir]

Reference behaviour reproduced (file:line into /root/reference/compute_expecto_features.py):
  * arguments ``annoFile tss_file --windowsize -o --cuda``                               :18-27
  * annoFile parsed by ``line.rstrip().split(",")`` (7 columns, header skipped)            :51-54
  * optional per-gene TSS override from ``tss_file`` (strand taken from that table)       :43-44,55-69
  * 200 shifts -20000..19800, window ``tss + shift*strand - 999 .. + 1000``, forward and
    ``[:, ::-1, ::-1]`` predictions averaged, ``pos_weights[10,200]`` reduction            :88-124
  * ``{out}/Xreducedall.2002.representative_tss_top.npy`` float64 [n_genes, 20020]         :126-128
The hg38->hg19 conversion (:45,58-66) needs the third-party ``liftover`` package and chain files that are not
available offline: rows of ``tss_file`` that would trigger it are refused unless ``--tss_is_hg19`` declares
their coordinates already lifted.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Compute ExPecto chromatin features for TSS list")
    p.add_argument("annoFile")
    p.add_argument("tss_file")
    p.add_argument("--windowsize", action="store", dest="windowsize", type=int, default=2000)
    p.add_argument("-o", dest="out_dir", type=str, default="temp_compute_expecto_features")
    p.add_argument("--cuda", action="store_true", help="accepted for compatibility (always on)")
    # extensions
    p.add_argument("--genome", default="./resources/hg19.fa")
    p.add_argument("--weights", default="./resources/deepsea.beluga.pth")
    p.add_argument("--dtype", default="f16", choices=["f16", "bf16", "f16x3"])
    p.add_argument("--device", type=int, default=0)
    p.add_argument("--genes_per_batch", type=int, default=32, help="genes per GPU batch (x400 windows)")
    p.add_argument("--gather", default="nccl", choices=["nccl", "file"],
                   help="multi-GPU (torchrun) only: gather the feature rows on rank 0 over NCCL, or every rank writes "
                        "its own rows of the .npy (no collective; the widening to float64 and the write run in parallel)")
    p.add_argument("--mode", default="segments", choices=["segments", "windows"],
                   help="segments: conv stack once per gene and strand (shared activations, bit-identical)")
    p.add_argument("--tss_is_hg19", action="store_true", help="tss_file coordinates are already hg19 (no liftover)")
    return p


def read_genes(anno_file: str, tss_file: str, tss_is_hg19: bool = False):
    """compute_expecto_features.py:43-69 -> list of (gene_id, chrom, tss, strand +-1)."""
    import pandas as pd

    tss_df = pd.read_csv(tss_file, sep="\t", index_col=0).set_index("ens_id")
    genes, found = [], 0
    for i, line in enumerate(open(anno_file)):
        if i == 0:
            continue
        gene_id, _, chrom, strand, _, rep_tss, _ = line.rstrip().split(",")
        if gene_id in tss_df.index:
            found += 1
            chrom_new, tss_new, strand, _, is_default = tss_df.loc[gene_id]
            if not is_default:
                if not tss_is_hg19:
                    raise NotImplementedError(
                        f"gene {gene_id}: the TSS override needs an hg38->hg19 liftover (third-party `liftover` + chain "
                        "files, unavailable offline); pre-lift the table and pass --tss_is_hg19")
                chrom, rep_tss = chrom_new, tss_new
        genes.append((gene_id, chrom, int(rep_tss), 1 if strand == "+" else -1))
    print(f"Found {found} genes in geneAnno file that match a TSS in provided TSS file...")
    return genes


def gene_features(engine, genome, genes, shifts, weights, genes_per_batch: int = 32, mode: str = "segments") -> np.ndarray:
    """float64 [len(genes), 20020] for (id, chrom, tss, strand) tuples — the loop at :105-124, batched.
    ``mode="segments"``: the 200 windows of a gene are 200-bp-apart substrings of one 41.8-kb sequence per strand,
    so the conv stack runs once per gene and strand (bx_forward_segments; bit-identical to ``"windows"``)."""
    import torch

    from .transform import decay_transform_fixed
    from .windows import check_valid_bases, tss_descs, tss_segment_descs

    S = len(shifts)
    out = np.empty((len(genes), 10 * 2002), dtype=np.float64)
    no_alleles = np.zeros(1, dtype=np.uint8)
    steps = np.diff(np.asarray(shifts))
    use_segments = mode == "segments" and S > 1 and S <= 200 and bool(np.all(steps == 200))
    # two pinned result buffers: batch k's features are fetched after batch k+1 has been enqueued, so the GPU never
    # waits for the host's descriptor building or the float64 widening
    pinned = [torch.empty((genes_per_batch, 10 * 2002), dtype=torch.float32, pin_memory=True) for _ in range(2)]
    pending = None

    def collect(p):
        ev, hb, at, m = p
        ev.synchronize()
        out[at:at + m] = hb[:m].numpy()

    for bi, g0 in enumerate(range(0, len(genes), genes_per_batch)):
        part = genes[g0:g0 + genes_per_batch]
        cols = ([g[1] for g in part], [g[2] for g in part], [g[3] for g in part])
        if use_segments:
            descs, k_of, seg_len = tss_segment_descs(genome, *cols, shifts)
            check_valid_bases(genome, descs)
            probs = engine.forward_segments(descs, no_alleles, len(part), 2, seg_len, k_of, shift_major=False)
        else:
            descs = tss_descs(genome, *cols, shifts)
            check_valid_bases(genome, descs)
            probs = engine.forward_windows(descs, no_alleles).view(len(part), S, 2, 2002)
        feats = decay_transform_fixed(probs, weights)
        hb = pinned[bi % 2]
        hb[:len(part)].copy_(feats, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(engine.device))
        if pending is not None:
            collect(pending)
        pending = (ev, hb, g0, len(part))
        del probs, feats
    if pending is not None:
        collect(pending)
    torch.cuda.synchronize(engine.device)
    return out


def gather_features(local: np.ndarray, n_total: int, device=None):
    """Per-rank float64 [n_local, 20020] blocks (contiguous gene shards in rank order) -> the full matrix on rank 0,
    None elsewhere.  The values are float32 results widened to float64 (as the reference's ``np.sum`` over float64
    weights stores them), so they travel as float32 — half the bytes over NVLink — and are widened again on rank 0."""
    import torch

    from . import shard

    t = torch.from_numpy(np.ascontiguousarray(local, dtype=np.float32))
    if device is not None:
        t = t.to(device)
    full = shard.gather_rows(t, n_total)
    return None if full is None else full.cpu().numpy().astype(np.float64)


def write_feature_slabs(npy_path: str, local: np.ndarray, row0: int, n_total: int, rank: int) -> None:
    """``--gather file``: rank 0 creates the ``.npy`` (the header ``np.save`` writes for a float64 [n_total, 20020]
    C-order array, compute_expecto_features.py:128), every rank then writes its own contiguous row range in place."""
    from . import shard

    shape = (int(n_total), local.shape[1])
    if rank == 0:
        mm = np.lib.format.open_memmap(npy_path, mode="w+", dtype=np.float64, shape=shape)
        del mm
    shard.barrier()
    mm = np.lib.format.open_memmap(npy_path, mode="r+")
    if mm.shape != shape or mm.dtype != np.float64:
        raise ValueError(f"{npy_path}: found {mm.dtype}{mm.shape}, expected float64{shape}")
    mm[row0:row0 + local.shape[0]] = local
    mm.flush()
    del mm


def run(args, engine=None, genome=None) -> str:
    """The whole script.  Under torchrun (one process per GPU) the genes are split into contiguous per-rank shards
    (expecto_b200/shard.py: no collective on the forward path); ``--gather nccl`` collects the feature rows on rank 0
    over NCCL and rank 0 writes the ``.npy``, ``--gather file`` has every rank write its own rows of it."""
    from . import shard
    from .transform import TSS_SHIFTS, fixed_pos_weights

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
    genes = read_genes(args.annoFile, args.tss_file, args.tss_is_hg19)
    missing = sorted({g[1] for g in genes if g[1] not in genome.offsets})
    if missing:
        raise KeyError(f"contig(s) {missing} are not in the genome")
    lo, hi = shard.shard_range(len(genes), rank, world)
    feats = gene_features(engine, genome, genes[lo:hi], TSS_SHIFTS, fixed_pos_weights(TSS_SHIFTS), args.genes_per_batch,
                          mode=args.mode)
    path = f"{args.out_dir}/Xreducedall.2002.representative_tss_top"
    if world > 1 and args.gather == "file":
        write_feature_slabs(path + ".npy", feats, lo, len(genes), rank)
    else:
        if world > 1:
            feats = gather_features(feats, len(genes), getattr(engine, "device", None))
        if rank == 0:
            np.save(path, feats)
    shard.barrier()
    return path + ".npy"


def main(argv=None) -> int:
    from . import shard
    run(build_parser().parse_args(argv))
    shard.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
