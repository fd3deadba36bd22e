"""Drop-in for the reference's ``make_closest_genes_file.py`` (same CLI, same two files out), vectorised.

    python -m expecto_b200.make_closest_genes_file out/snps_hg19.vcf [--all_in_receptive_field] [--add_chr_prefix]
        [--geneanno_file ./resources/geneanno.csv] [-o temp_closest_gene_file]

The reference walks the VCF with ``iterrows`` and, per variant, filters the 24 k-row gene table and takes
``np.argmin(|TSS - pos|)`` (make_closest_genes_file.py:46-66,69-79) or every gene whose receptive field contains
the variant (:82-108) — O(variants x genes) pandas work that dominates config 5 once chromatin is fast
(SURVEY §8f rank 3).  Here the TSSs of each contig are sorted once and every variant is resolved with two binary
searches (closest gene) or one interval query (receptive field); row order, tie-breaking (first gene in
annotation order wins, as ``argmin`` does) and both output files are identical to the reference's.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

# is_in_receptive_field (make_closest_genes_file.py:111-116) with shifts = range(-20000, 20000, 200), windowsize 1000:
# '+' genes: -20000-499 <= pos - tss <= 19800+500 ; '-' genes: -19800-499 <= pos - tss <= 20000+500
RF_BOUNDS = {1: (-20499, 20300), -1: (-20299, 20500)}


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Make closest gene file required by predict.py")
    p.add_argument("hg19_snps_file")
    p.add_argument("--all_in_receptive_field", action="store_true")
    p.add_argument("--add_chr_prefix", action="store_true")
    p.add_argument("--geneanno_file", dest="geneanno_file", type=str, default="./resources/geneanno.csv")
    p.add_argument("-o", dest="out_dir", type=str, default="temp_closest_gene_file", help="Output directory")
    return p


class GeneIndex:
    """Per-contig TSS arrays sorted by (position, annotation order)."""

    def __init__(self, seqnames, tss, strand):
        self.by_chrom = {}
        seqnames = np.asarray(seqnames).astype(str)
        tss = np.asarray(tss, dtype=np.int64)
        sgn = np.where(np.asarray(strand).astype(str) == "+", 1, -1)
        for c in np.unique(seqnames):
            idx = np.flatnonzero(seqnames == c)
            order = np.lexsort((idx, tss[idx]))
            self.by_chrom[c] = (tss[idx][order], idx[order], sgn[idx][order])

    def closest(self, chrom: str, pos: np.ndarray) -> np.ndarray:
        """Annotation row of ``argmin |tss - pos|`` among the contig's genes (first row wins ties)."""
        t, g, _ = self.by_chrom[chrom]
        pos = np.asarray(pos, dtype=np.int64)
        r = np.searchsorted(t, pos, side="left")                 # first tss >= pos
        rc = np.minimum(r, len(t) - 1)
        lc = np.maximum(r - 1, 0)
        lc = np.searchsorted(t, t[lc], side="left")              # first row of the run of equal tss on the left
        dr = np.where(r < len(t), np.abs(t[rc] - pos), np.iinfo(np.int64).max)
        dl = np.where(r > 0, np.abs(t[lc] - pos), np.iinfo(np.int64).max)
        take_left = (dl < dr) | ((dl == dr) & (g[lc] < g[rc]))
        return np.where(take_left, g[lc], g[rc])

    def in_receptive_field(self, chrom: str, pos: int) -> np.ndarray:
        """Annotation rows (in annotation order) whose receptive field contains ``pos``."""
        t, g, s = self.by_chrom[chrom]
        lo = np.searchsorted(t, pos - 20500, side="left")
        hi = np.searchsorted(t, pos + 20499, side="right")
        d = pos - t[lo:hi]
        ok = np.where(s[lo:hi] > 0, (d >= RF_BOUNDS[1][0]) & (d <= RF_BOUNDS[1][1]),
                      (d >= RF_BOUNDS[-1][0]) & (d <= RF_BOUNDS[-1][1]))
        return np.sort(g[lo:hi][ok])


def match_genes(vcf_chrom, vcf_pos, index: GeneIndex, all_in_receptive_field: bool):
    """(variant row, annotation row) pairs in the reference's output order."""
    chrom = np.asarray(vcf_chrom).astype(str)
    pos = np.asarray(vcf_pos, dtype=np.int64)
    closest = np.full(len(pos), -1, dtype=np.int64)
    for c in np.unique(chrom):
        if c not in index.by_chrom:
            # the reference's argmin over an empty selection raises ValueError
            raise ValueError(f"attempt to get argmin of an empty sequence (no gene on contig {c!r})")
        m = chrom == c
        closest[m] = index.closest(c, pos[m])
    if not all_in_receptive_field:
        return np.arange(len(pos)), closest
    vi, gi = [], []
    for i in range(len(pos)):
        g = index.in_receptive_field(chrom[i], int(pos[i]))
        if g.size == 0:
            g = closest[i:i + 1]
        vi.append(np.full(g.size, i, dtype=np.int64))
        gi.append(g)
    return (np.concatenate(vi), np.concatenate(gi)) if vi else (np.zeros(0, np.int64), np.zeros(0, np.int64))


def run(args) -> dict:
    import pandas as pd

    os.makedirs(args.out_dir, exist_ok=True)
    vcf = pd.read_csv(args.hg19_snps_file, sep="\t", header=None, comment="#")
    if args.add_chr_prefix:
        vcf[0] = "chr" + vcf[0].astype(str)
    geneanno = pd.read_csv(args.geneanno_file, index_col=0)
    index = GeneIndex(geneanno["seqnames"].values, geneanno["CAGE_representative_TSS"].values,
                      geneanno["strand"].values)
    vi, gi = match_genes(vcf[0].values, vcf[1].values, index, args.all_in_receptive_field)

    vcf_out_path = f"{args.out_dir}/snps_hg19.vcf"
    with open(vcf_out_path, "w") as f:
        print("##fileformat=VCFv4.3", file=f)
        print("#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO", file=f)
    snp_chrom = vcf[0].values.astype(str)[vi]
    snp_pos = vcf[1].values.astype(np.int64)[vi]
    tss_chrom = geneanno["seqnames"].values.astype(str)[gi]
    tss_pos = geneanno["CAGE_representative_TSS"].values.astype(np.int64)[gi]
    out = pd.DataFrame({
        "snp_chrom": [c[3:] for c in snp_chrom], "snp_pos_start": snp_pos - 1, "snp_pos": snp_pos,
        "ref": vcf[3].values[vi], "alt": vcf[4].values[vi], "tss_chrom": [c[3:] for c in tss_chrom],
        "tss_pos_start": tss_pos - 1, "tss_pos": tss_pos, "tss_strand": geneanno["strand"].values[gi],
        "ens_id": geneanno.index.values[gi], "dist_to_tss": tss_pos - snp_pos})
    out.to_csv(f"{args.out_dir}/closest_genes.tsv", sep="\t", index=False, header=False)
    vcf.iloc[vi].to_csv(vcf_out_path, sep="\t", header=False, index=False, mode="a")
    return {"rows": len(out), "variants": len(vcf)}


def main(argv=None) -> int:
    run(build_parser().parse_args(argv))
    return 0


if __name__ == "__main__":
    sys.exit(main())
