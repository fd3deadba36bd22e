"""Drop-in for the reference's ``geuvadis_predict_consensus.py`` (same CLI, same per-gene ``.h5`` files) on B200s.

    python -m expecto_b200.geuvadis_predict_consensus model.save consensus_dir genes.csv [-o out_dir]

Per gene directory ``{consensus_dir}/{gene}/samples/*.fa`` every FASTA record is one individual's 393,216-bp
consensus sequence with the TSS at ``len // 2``.  Reference behaviour reproduced (file:line into
/root/reference/geuvadis_predict_consensus.py):
  * arguments ``expecto_model consensus_dir genes_file --beluga_model --batch_size --overwrite --exp_only
    --num_chunks --chunk_i -o``                                                                   :26-40
  * genes = natural-sorted directory names, optional ``np.array_split`` chunking                   :58,66-69
  * strand looked up by lower-cased gene symbol (ens_id where the symbol is missing)               :61-63,74
  * records upper-cased; short records padded with N at the front (interval starting with ``-``)
    or at the end; record id ``{record.id}|{file stem}``                                           :138-168
  * 200 shifts -20000..19800, window ``seq[tss_i + shift*strand - 999 : tss_i + shift*strand + 1001]``,
    forward and reverse-complement predictions averaged in float64                                 :91-103,210-243
  * ``pos_weights[10,200]`` reduction, one zero column prepended per basis function (the legacy 20030-wide
    layout), gblinear score                                                                        :108-126
  * ``{out}/{gene}/{gene}_chromatin.h5`` (``chromatin_preds`` float64 [n,200,2002], ``record_ids`` S) and
    ``{out}/{gene}/{gene}.h5`` (``expecto_preds`` float32 [n], ``record_ids``)                     :127-133
  * ``--exp_only`` re-scores a stored ``{gene}_chromatin.h5``; genes whose ``{gene}.h5`` exists are skipped
    unless ``--overwrite``                                                                          :79-88

How it runs here: the ±21 kb around the TSS of every record of a gene is packed to 2 bits and becomes the engine's
"genome" for that gene; the 200 windows of a record are 200-bp-apart substrings of one 41.8-kb segment per strand,
so the conv stack runs once per record and strand (bx_forward_segments, bit-identical to the per-window path).
``--batch_size`` is accepted and ignored: it only sizes the reference's mini-batches, and its loop (:95-98, index
``i * batch_size`` with ``i`` already stepping by ``batch_size``) is correct only when one batch holds all 400
windows, which the default 1024 does — that is the behaviour reproduced.
Under torchrun (one process per GPU) the genes (of the selected chunk) are split over the ranks with the same
``np.array_split``; every rank writes the files of its own genes, no collective.
"""
from __future__ import annotations

import argparse
import glob
import os
import re
import sys
from pathlib import Path

import numpy as np

ENFORMER_SEQ_LENGTH = 393216   # geuvadis_predict_consensus.py:23
HALF_SPAN = 21000              # bases kept on either side of the TSS: covers every shift +-20000 and +-1000
LEGACY_WIDTH = 10 * 2003       # 20030: one zero column in front of each basis function's 2002 features (:123-124)


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Predict expression for consensus sequences using ExPecto")
    p.add_argument("expecto_model")
    p.add_argument("consensus_dir")
    p.add_argument("genes_file")
    p.add_argument("--beluga_model", type=str, default="./resources/deepsea.beluga.pth")
    p.add_argument("--batch_size", action="store", dest="batch_size", type=int, default=1024,
                   help="accepted for compatibility (see module docstring)")
    p.add_argument("--overwrite", action="store_true", dest="overwrite", default=False)
    p.add_argument("--exp_only", action="store_true", dest="exp_only", default=False)
    p.add_argument("--num_chunks", action="store", dest="num_chunks", type=int, default=None)
    p.add_argument("--chunk_i", action="store", dest="chunk_i", type=int, default=None)
    p.add_argument("-o", dest="out_dir", type=str, default="temp_predict_consensus")
    # extensions
    p.add_argument("--dtype", default="f16", choices=["f16", "bf16"])
    p.add_argument("--device", type=int, default=0)
    p.add_argument("--records_per_batch", type=int, default=32, help="records per GPU batch (x400 windows)")
    p.add_argument("--mode", default="segments", choices=["segments", "windows"])
    return p


def natural_key(name: str):
    """Sort key of ``natsort.natsorted`` with its default algorithm (unsigned integers): digit runs compare as
    numbers, everything else as text; the split always starts with a (possibly empty) text piece, so keys compare
    piece by piece without mixing types."""
    parts = re.split(r"(\d+)", name)
    return tuple(int(p) if i % 2 else p for i, p in enumerate(parts))


def read_fasta_records(path: str):
    """(record.id, sequence) pairs as ``Bio.SeqIO.parse(path, 'fasta')`` yields them: the id is the header up to the
    first whitespace, the sequence is the record's lines with whitespace removed."""
    rid, chunks = None, []
    with open(path) as f:
        for line in f:
            if line.startswith(">"):
                if rid is not None:
                    yield rid, "".join(chunks)
                head = line[1:].split()
                rid, chunks = (head[0] if head else ""), []
            elif rid is not None:
                chunks.append("".join(line.split()))
    if rid is not None:
        yield rid, "".join(chunks)


def sample_seqs_and_ids(fasta_files):
    """geuvadis_predict_consensus.py:138-168: upper-case, N-pad to 393,216 (front when the interval starts with
    ``-``, else end), check the interval spans 393,216 bp; yields (sequence, ``{record.id}|{file stem}``)."""
    for fasta_file in fasta_files:
        for rid, seq in read_fasta_records(fasta_file):
            seq = seq.upper()
            interval = rid.split(":")[1]
            if interval.startswith("-"):
                bp_start = -int(interval.split("-")[-2])
                bp_end = int(interval.split("-")[-1])
                assert bp_end - bp_start + 1 == ENFORMER_SEQ_LENGTH
                seq = "N" * (ENFORMER_SEQ_LENGTH - len(seq)) + seq
            else:
                bp_start, bp_end = map(int, interval.split("-"))
                assert bp_end - bp_start + 1 == ENFORMER_SEQ_LENGTH
                if len(seq) < ENFORMER_SEQ_LENGTH:
                    seq = seq + "N" * (ENFORMER_SEQ_LENGTH - len(seq))
            assert len(seq) == ENFORMER_SEQ_LENGTH, f"Sequence length is {len(seq)} for {rid}"
            yield seq, f"{rid}|{Path(fasta_file).stem}"


def read_gene_strands(genes_file: str):
    """:61-63 — DataFrame indexed by lower-cased gene symbol (ens_id where the symbol is empty)."""
    import pandas as pd

    df = pd.read_csv(genes_file, names=["ens_id", "chrom", "bp", "gene_symbol", "strand"], index_col=False)
    df["gene_symbol"] = df["gene_symbol"].fillna(df["ens_id"]).str.lower()
    return df.set_index("gene_symbol")


def strand_sign(strand) -> int:
    if isinstance(strand, str) and strand == "+":
        return 1
    if isinstance(strand, str) and strand == "-":
        return -1
    raise AssertionError(f"strand {strand} not recognized")


def record_chromatin(engine, seqs, strand: int, shifts, records_per_batch: int = 32, mode: str = "segments"):
    """float64 [len(seqs), S, 2002]: per record and shift the mean of the forward and reverse-complement prediction
    (:91-103).  ``seqs`` are full-length consensus strings; only the +-21 kb around ``len // 2`` is packed."""
    import torch

    from .genome import PackedGenome
    from .windows import check_valid_bases, tss_descs, tss_segment_descs

    S = len(shifts)
    n = len(seqs)
    tss_i = ENFORMER_SEQ_LENGTH // 2
    for s in seqs:
        if len(s) != ENFORMER_SEQ_LENGTH:
            raise AssertionError(f"Expected seq of length {ENFORMER_SEQ_LENGTH} but got {len(s)}")
    names = [f"r{i}" for i in range(n)]
    genome = PackedGenome.from_sequences({nm: s[tss_i - HALF_SPAN: tss_i + HALF_SPAN + 2] for nm, s in zip(names, seqs)})
    engine.set_genome(genome)
    out = np.empty((n, S, 2002), dtype=np.float64)
    no_alleles = np.zeros(1, dtype=np.uint8)
    steps = np.diff(np.asarray(shifts))
    use_segments = mode == "segments" and 1 < S <= 200 and bool(np.all(steps == 200))
    B = max(1, min(records_per_batch, n))
    pinned = [torch.empty((B, S, 2002), dtype=torch.float64, pin_memory=True) for _ in range(2)]
    pending = None

    def collect(p):
        ev, hb, at, m = p
        ev.synchronize()
        out[at:at + m] = hb[:m].numpy()

    for bi, r0 in enumerate(range(0, n, B)):
        m = min(B, n - r0)
        cols = (names[r0:r0 + m], [HALF_SPAN + 1] * m, [strand] * m)      # 1-based TSS inside the packed slice
        if use_segments:
            descs, k_of, seg_len = tss_segment_descs(genome, *cols, shifts)
            check_valid_bases(genome, descs)
            probs = engine.forward_segments(descs, no_alleles, m, 2, seg_len, k_of, shift_major=False)
        else:
            descs = tss_descs(genome, *cols, shifts)
            check_valid_bases(genome, descs)
            probs = engine.forward_windows(descs, no_alleles)
        probs = probs.view(m, S, 2, 2002)
        mean = (probs[:, :, 0].double() + probs[:, :, 1].double()) / 2
        hb = pinned[bi % 2]
        hb[:m].copy_(mean, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(engine.device))
        if pending is not None:
            collect(pending)
        pending = (ev, hb, r0, m)
        del probs, mean
    if pending is not None:
        collect(pending)
    return out


def legacy_features(preds: np.ndarray, shifts) -> np.ndarray:
    """:108-124 — float64 [n, 20030]: exponential-decay sums over the shifts, a zero column in front of each of the
    ten basis functions ("backwards compatibility" with models trained on 2003-wide blocks)."""
    from .transform import fixed_pos_weights

    w = fixed_pos_weights(shifts)                                   # float64 [10, S]
    feats = np.matmul(w[None, :, :], np.asarray(preds, dtype=np.float64))   # [n, 10, 2002]
    out = np.zeros((feats.shape[0], 10, 2003), dtype=np.float64)
    out[:, :, 1:] = feats
    return out.reshape(-1, LEGACY_WIDTH)


def gblinear_scores(features: np.ndarray, model) -> np.ndarray:
    """``bst.predict(xgb.DMatrix(features))`` (:125-126): DMatrix stores float32 entries, predictions are float32."""
    if model.weights.size != features.shape[1]:
        raise ValueError(f"model has {model.weights.size} features, the consensus layout has {features.shape[1]}")
    x = np.asarray(features).astype(np.float32).astype(np.float64)
    return (model.offset + x @ model.weights.astype(np.float64)).astype(np.float32)


def write_gene(preds_dir: str, gene: str, preds: np.ndarray, record_ids, model, shifts) -> np.ndarray:
    """:108-133 for one gene: features, scores, the two ``.h5`` files.  Returns the scores."""
    from . import hdf5

    ids = np.array(record_ids, "S")
    scores = gblinear_scores(legacy_features(preds, shifts), model)
    hdf5.write_datasets(f"{preds_dir}/{gene}_chromatin.h5", {"chromatin_preds": preds, "record_ids": ids})
    hdf5.write_datasets(f"{preds_dir}/{gene}.h5", {"expecto_preds": scores, "record_ids": ids})
    return scores


def select_genes(consensus_dir: str, num_chunks, chunk_i, rank: int = 0, world: int = 1):
    """:58,66-69 plus the per-rank split."""
    genes = sorted((os.path.basename(f) for f in glob.glob(f"{consensus_dir}/*")), key=natural_key)
    if num_chunks is not None:
        genes = np.array_split(genes, num_chunks)[chunk_i]
        assert len(genes) > 0, "Gene split resulted in empty list"
    if world > 1:
        genes = np.array_split(genes, world)[rank]
    return [str(g) for g in genes]


def run(args, engine=None) -> list[str]:
    """The whole script; returns the genes this process handled.  Post-processing and file writes of gene k run on
    a worker thread while the GPU computes gene k+1."""
    from concurrent.futures import ThreadPoolExecutor

    from . import gblinear, hdf5, shard
    from .transform import TSS_SHIFTS

    rank, world, local = shard.init_distributed()
    os.makedirs(args.out_dir, exist_ok=True)
    model = gblinear.load_model(args.expecto_model.strip())
    shifts = TSS_SHIFTS
    genes_df = read_gene_strands(args.genes_file)
    genes = select_genes(args.consensus_dir, args.num_chunks, args.chunk_i, rank, world)
    if engine is None and not args.exp_only and genes:
        from .engine import BelugaEngine
        engine = BelugaEngine(local if world > 1 else args.device, dtype=args.dtype, capacity=4096)
        engine.load_weights_file(args.beluga_model)
    print("Predicting chromatin for all samples for all genes...")
    done, job = [], None
    with ThreadPoolExecutor(max_workers=1) as pool:
        for gene in genes:
            fasta_files = glob.glob(f"{args.consensus_dir}/{gene}/samples/*.fa")
            strand = genes_df.loc[gene, "strand"]
            preds_dir = f"{args.out_dir}/{gene}"
            os.makedirs(preds_dir, exist_ok=True)
            if not args.overwrite and os.path.exists(f"{preds_dir}/{gene}.h5"):
                print(f"Skipping gene {gene} since h5 is already present.")
                continue
            if args.exp_only:
                with hdf5.File(f"{preds_dir}/{gene}_chromatin.h5") as f:
                    preds = np.array(f["chromatin_preds"][...], copy=True)   # the file is rewritten below
                    record_ids = [x.decode("utf-8") for x in np.asarray(f["record_ids"])]
            else:
                pairs = list(sample_seqs_and_ids(fasta_files))
                if not pairs:
                    raise ValueError(f"gene {gene}: need at least one consensus record")
                record_ids = [p[1] for p in pairs]
                preds = record_chromatin(engine, [p[0] for p in pairs], strand_sign(strand), shifts,
                                         args.records_per_batch, args.mode)
            if job is not None:
                job.result()
            job = pool.submit(write_gene, preds_dir, gene, preds, record_ids, model, shifts)
            done.append(gene)
        if job is not None:
            job.result()
    shard.barrier()
    return done


def main(argv=None) -> int:
    from . import shard
    run(build_parser().parse_args(argv))
    shard.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
