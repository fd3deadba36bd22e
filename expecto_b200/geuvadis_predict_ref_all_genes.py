"""Drop-in for the reference's ``geuvadis_predict_ref_all_genes.py`` (same CLI, same ``ref_preds.csv``) on B200s.

    python -m expecto_b200.geuvadis_predict_ref_all_genes model.save consensus_dir genes.csv [-o out_dir]

For every gene of ``genes_file`` the single record of ``{consensus_dir}/{gene.lower()}/ref.fa`` is the 393,216-bp
reference-genome sequence around the TSS (TSS at ``len // 2``).  Reference behaviour reproduced (file:line into
/root/reference/geuvadis_predict_ref_all_genes.py):
  * arguments ``expecto_model consensus_dir genes_file --beluga_model --batch_size -o`` (default output directory
    ``temp_sed_for_top_eqtls``, as the reference has it)                                              :24-33
  * genes in file order; the symbol column falls back to the ens_id and keeps its case in the output,
    the directory name is its lower case                                                              :57-59,62-67
  * exactly one FASTA record per ``ref.fa`` (AssertionError otherwise), upper-cased, N-padded to 393,216
    (front when the interval starts with ``-``, else end)                                             :113-146
  * 200 shifts -20000..19800, forward + reverse-complement predictions averaged                       :70-78,149-173
  * ``pos_weights[10,200]`` reduction, a zero column in front of each basis function (20030 wide),
    gblinear score                                                                                    :80-98
  * ``{out}/ref_preds.csv`` with columns ``genes, ref_preds`` written by ``DataFrame.to_csv(header=True,
    index=False)`` from ``np.array(list of [1]-arrays).squeeze()``                                    :100-103

How it runs here: the record's +-21 kb around the TSS is packed to 2 bits, the 200 windows of a gene are substrings
of one 41.8-kb segment per strand, so the conv stack runs once per gene and strand (``bx_forward_segments`` through
``geuvadis_predict_consensus.record_chromatin``); genes are processed ``--genes_per_batch`` at a time, grouped by
strand.  ``--batch_size`` is accepted and ignored (the reference's loop, ``i * batch_size`` with ``i`` stepping by
``batch_size``, is only correct when one batch holds all 400 windows — the default 1024 does).
Under torchrun the genes are split over the ranks in contiguous runs and rank 0 collects the scores and writes the file.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

from .geuvadis_predict_consensus import (gblinear_scores, legacy_features, record_chromatin, sample_seqs_and_ids,
                                         strand_sign)


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Predict expression for consensus sequences using ExPecto")
    p.add_argument("expecto_model")
    p.add_argument("consensus_dir")
    p.add_argument("genes_file")
    p.add_argument("--beluga_model", type=str, default="./resources/deepsea.beluga.pth")
    p.add_argument("--batch_size", action="store", dest="batch_size", type=int, default=1024,
                   help="accepted for compatibility (see module docstring)")
    p.add_argument("-o", dest="out_dir", type=str, default="temp_sed_for_top_eqtls", help="Output directory")
    # extensions
    p.add_argument("--dtype", default="f16", choices=["f16", "bf16"])
    p.add_argument("--device", type=int, default=0)
    p.add_argument("--genes_per_batch", type=int, default=64, help="ref.fa records held in memory at a time")
    p.add_argument("--mode", default="segments", choices=["segments", "windows"])
    return p


def read_genes(genes_file: str):
    """:57-59 — DataFrame indexed by gene symbol (ens_id where the symbol is missing), original case."""
    import pandas as pd

    df = pd.read_csv(genes_file, names=["ens_id", "chrom", "bp", "gene_symbol", "strand"], index_col=False)
    df["gene_symbol"] = df["gene_symbol"].fillna(df["ens_id"])
    return df.set_index("gene_symbol")


def read_ref_sequence(fasta_file: str) -> str:
    """get_1_id_and_seq_from_fasta, :113-146: the one record of ``fasta_file``, upper-cased and N-padded."""
    records = list(sample_seqs_and_ids([fasta_file]))
    assert len(records) == 1, f"Expected 1 record in fasta file {fasta_file}, but got {len(records)} records"
    return records[0][0]


def ref_scores(engine, consensus_dir: str, genes, strands, model, genes_per_batch: int = 64, mode: str = "segments"):
    """float32 [len(genes)]: the ExPecto score of every gene's reference sequence (:61-98)."""
    from .transform import TSS_SHIFTS

    scores = np.zeros(len(genes), dtype=np.float32)
    for g0 in range(0, len(genes), max(1, genes_per_batch)):
        idx = list(range(g0, min(len(genes), g0 + max(1, genes_per_batch))))
        seqs = {i: read_ref_sequence(f"{consensus_dir}/{genes[i].lower()}/ref.fa") for i in idx}
        for sgn in (1, -1):   # record_chromatin takes one strand per call
            sel = [i for i in idx if strand_sign(strands[i]) == sgn]
            if not sel:
                continue
            preds = record_chromatin(engine, [seqs[i] for i in sel], sgn, TSS_SHIFTS, records_per_batch=32, mode=mode)
            scores[sel] = gblinear_scores(legacy_features(preds, TSS_SHIFTS), model)
    return scores


def run(args, engine=None):
    """The whole script; returns the DataFrame written (rank 0) or None (other ranks)."""
    import pandas as pd
    import torch.distributed as dist

    from . import gblinear, shard

    rank, world, local = shard.init_distributed()
    os.makedirs(args.out_dir, exist_ok=True)
    model = gblinear.load_model(args.expecto_model.strip())
    genes_df = read_genes(args.genes_file)
    genes = [str(g) for g in genes_df.index]
    strands = [genes_df["strand"].iloc[i] for i in range(len(genes))]
    lo, hi = shard.shard_range(len(genes), rank, world)
    if engine is None and hi > lo:
        from .engine import BelugaEngine
        engine = BelugaEngine(local if world > 1 else args.device, dtype=args.dtype, capacity=4096)
        engine.load_weights_file(args.beluga_model)
    mine = ref_scores(engine, args.consensus_dir, genes[lo:hi], strands[lo:hi], model, args.genes_per_batch, args.mode)
    if world > 1:
        parts = [None] * world
        dist.all_gather_object(parts, mine)
        mine = np.concatenate(parts)
    df = None
    if rank == 0:
        # :100-103 — the reference stacks one [1]-array per gene and squeezes
        expecto_ref_preds = np.array([np.array([v], dtype=np.float32) for v in mine]).squeeze()
        df = pd.DataFrame({"genes": np.array(genes_df.index.values), "ref_preds": expecto_ref_preds})
        df.to_csv(f"{args.out_dir}/ref_preds.csv", header=True, index=False)
    shard.barrier()
    return df


def main(argv=None) -> int:
    from . import shard
    run(build_parser().parse_args(argv))
    shard.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
