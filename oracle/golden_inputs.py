"""Seeded inputs shared by oracle/make_golden.py (which runs the reference on them) and tests/ (which feed the
same inputs to the CUDA path and the oracle).  TEST INFRASTRUCTURE ONLY — see oracle/beluga_oracle.py."""
from __future__ import annotations

import numpy as np

from expecto_b200 import synth

WEIGHT_SEED = 0
BELUGA_N, BELUGA_SEED = 6, 2
GBLINEAR_SEED = 5
SMALL_CONTIGS = [("chr1", 60000), ("chr2", 40000), ("chrX", 30000)]
SMALL_GENOME_SEED = 11
ENCODE_SHIFTS = [0, -800, 800, 200]
EXTREME_SHIFTS = [1000, -1000, 1040, 1049, 1050, 1051, 1200, 2000, -1049, -1050, -1051, -1100, -1200, -2000, 3000, -3000]
CFG1_KEEP_SHIFTS = [0, 800]
CFG1_CONTIGS = [("chr1", 11_200_000), ("chr2", 50_000), ("chrX", 50_000)]


def cfg1_genome():
    """synthetic chr1 >= 11.2 Mb (seed 1) so example/example.vcf's positions exist."""
    return synth.make_genome(CFG1_CONTIGS, seed=1)


def _gref(genome, chrom, pos, n=1):
    s = genome[chrom][pos - 1:pos - 1 + n].upper()
    return "".join(c if c in "ACGT" else "A" for c in s)


def edge_variants(genome):
    """Hand-made rows covering what fetchSeqs/encodeSeqs do with indels, mismatches, case and non-ACGT
    alleles (chromatin.py:164,203-209).  [chrom, pos(1-based), id, ref, alt]."""
    rng = np.random.default_rng(7)
    rows = []

    def pos_on(chrom):
        return int(rng.integers(4000, len(genome[chrom]) - 4000))

    for chrom in ("chr1", "chr2", "chrX"):
        p = pos_on(chrom)
        r = _gref(genome, chrom, p)
        rows.append([chrom, p, "snv", r, "ACGT"[("ACGT".index(r) + 1) % 4]])
    p = pos_on("chr1"); r = _gref(genome, "chr1", p)
    rows.append(["chr1", p, "mismatch", "ACGT"[("ACGT".index(r) + 2) % 4], "ACGT"[("ACGT".index(r) + 3) % 4]])
    for d in (1, 2, 3, 10, 51, 150):  # insertions: odd/even length deltas, one longer than the 100-bp margin
        p = pos_on("chr2"); r = _gref(genome, "chr2", p)
        rows.append(["chr2", p, f"ins{d}", r, r + "".join("ACGT"[i] for i in rng.integers(0, 4, d))])
    for d in (1, 2, 3, 10, 51, 99, 100, 101, 120):  # deletions; >= 101 drives the crop start negative
        p = pos_on("chr1"); r = _gref(genome, "chr1", p, d + 1)
        rows.append(["chr1", p, f"del{d}", r, r[0]])
    p = pos_on("chrX"); r = _gref(genome, "chrX", p, 4)
    rows.append(["chrX", p, "mnv", r, r[::-1]])
    p = pos_on("chrX"); r = _gref(genome, "chrX", p)
    rows.append(["chrX", p, "lower", r.lower(), "ACGT"[("ACGT".index(r) + 1) % 4].lower()])
    p = pos_on("chr1"); r = _gref(genome, "chr1", p)
    rows.append(["chr1", p, "altN", r, "N"])
    p = pos_on("chr1"); r = _gref(genome, "chr1", p, 3)
    rows.append(["chr1", p, "altdash", r, "-"])
    p = pos_on("chr2"); r = _gref(genome, "chr2", p)
    rows.append(["chr2", p, "mixed", r, "aNgH-T"])
    return rows


def cfg1_edge_variants(genome):
    """A short VCF for the whole-script run: exercises chromatin.py:240-241 (prefixing + CHRS filter) too."""
    rows = []
    r = _gref(genome, "chr1", 2_000_000)
    rows.append(["1", 2_000_000, "unprefixed", r, "ACGT"[("ACGT".index(r) + 1) % 4]])
    r = _gref(genome, "chr2", 20_000)
    rows.append(["chr2", 20_000, "ins3", r, r + "GAT"])
    r = _gref(genome, "chrX", 25_000, 6)
    rows.append(["chrX", 25_000, "del5", r, r[0]])
    rows.append(["chrUn_gl000220", 1234, "filtered", "A", "C"])
    r = _gref(genome, "chr1", 9_000_001)
    rows.append(["chr1", 9_000_001, "mismatch", "ACGT"[("ACGT".index(r) + 2) % 4], "ACGT"[("ACGT".index(r) + 1) % 4]])
    r = _gref(genome, "chr1", 5_555_555)
    rows.append(["chr1", 5_555_555, "lower", r.lower(), "ACGT"[("ACGT".index(r) + 3) % 4].lower()])
    r = _gref(genome, "chr2", 30_000, 121)
    rows.append(["chr2", 30_000, "del120", r, r[0]])
    return rows


def cfg4_genes():
    """[gene id, chrom, strand, CAGE_representative_TSS] for the compute_expecto_features run (>= 21 kb from ends)."""
    return [["ENSG_T1", "chr1", "+", 5_000_000], ["ENSG_T2", "chr1", "-", 7_654_321]]
