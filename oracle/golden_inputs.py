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


# --- consensus front end (geuvadis_predict_consensus.py) -------------------------------------------------------
CONSENSUS_LEN = 393216
CONSENSUS_KEEP_SHIFTS = [0, 37, 66, 99, 100, 101, 133, 170, 199]   # shift indices stored in the fixture
CONSENSUS_FEATURE_STEP = 7                                          # every 7th chromatin feature stored
CONSENSUS_MODEL_SEED = 9
# genes file rows (ens_id, chrom, bp, gene_symbol, strand); the second gene has no symbol -> its directory is
# the lower-cased ens_id (geuvadis_predict_consensus.py:62)
CONSENSUS_GENES = [["ENSG00000000011", "chr1", 1_200_000, "Abc1", "+"], ["ENSG00000000002", "chr2", 900_000, "", "-"]]


def consensus_records(gene_index: int):
    """{file stem: [(record id, sequence as stored in the FASTA)]} for one gene: a full-length record, one missing its
    end (N-padded at the end by the reference), one missing its beginning (interval starting with '-', N-padded in
    front) — the truncations reach into the +-21 kb the windows read.  Individuals differ from the gene's base
    sequence by 60 substitutions near the TSS."""
    rng = np.random.default_rng(100 + gene_index)
    base = np.frombuffer(synth.random_sequence(rng, CONSENSUS_LEN).encode(), dtype=np.uint8)
    chrom = CONSENSUS_GENES[gene_index][1]

    def individual(k):
        r = np.random.default_rng(1000 * gene_index + k)
        s = base.copy()
        at = CONSENSUS_LEN // 2 + r.integers(-21000, 21000, size=60)
        s[at] = np.frombuffer(b"ACGT", dtype=np.uint8)[r.integers(0, 4, size=60)]
        return s.tobytes().decode()

    start = 500_000 + 1000 * gene_index
    full_id = f"{chrom}:{start}-{start + CONSENSUS_LEN - 1}"
    out = {"HG001": [(full_id, individual(0))]}
    if gene_index == 0:
        out["HG002"] = [(full_id, individual(1)[:210_000])]
        out["NA003"] = [(f"{chrom}:-5000-{CONSENSUS_LEN - 5001}", individual(2)[190_000:]), (full_id, individual(3))]
    else:
        out["HG010"] = [(full_id, individual(1)[:CONSENSUS_LEN - 7])]
    return out


def write_consensus_dir(root: str):
    """{root}/consensus/{gene}/samples/{stem}.fa + {root}/genes.csv; returns (consensus_dir, genes_file, gene names)."""
    import os
    names = []
    for gi, row in enumerate(CONSENSUS_GENES):
        gene = (row[3] or row[0]).lower()
        names.append(gene)
        d = os.path.join(root, "consensus", gene, "samples")
        os.makedirs(d, exist_ok=True)
        for stem, recs in consensus_records(gi).items():
            with open(os.path.join(d, stem + ".fa"), "w") as f:
                for rid, seq in recs:
                    f.write(f">{rid} sample={stem}\n")
                    for i in range(0, len(seq), 80):
                        f.write(seq[i:i + 80] + "\n")
    genes_file = os.path.join(root, "genes.csv")
    with open(genes_file, "w") as f:
        for row in CONSENSUS_GENES:
            f.write(",".join(str(x) for x in row) + "\n")
    return os.path.join(root, "consensus"), genes_file, names


def consensus_ref_record(gene_index: int):
    """(record id, sequence as stored in ``{gene}/ref.fa``) — the one reference-genome record per gene that
    geuvadis_predict_ref_all_genes.py reads; the second gene's record misses its last 12,345 bases (N-padded by the
    reference, the padding reaching no window) and is lower case in part (the reference upper-cases)."""
    rng = np.random.default_rng(100 + gene_index)
    base = synth.random_sequence(rng, CONSENSUS_LEN)      # same base sequence as consensus_records()
    chrom = CONSENSUS_GENES[gene_index][1]
    start = 500_000 + 1000 * gene_index
    rid = f"{chrom}:{start}-{start + CONSENSUS_LEN - 1}"
    return (rid, base) if gene_index == 0 else (rid, base[:CONSENSUS_LEN - 12_345])


def write_ref_fastas(consensus_dir: str):
    """{consensus_dir}/{gene}/ref.fa for every gene of CONSENSUS_GENES (the directory may already hold samples/)."""
    import os
    for gi, row in enumerate(CONSENSUS_GENES):
        gene = (row[3] or row[0]).lower()
        os.makedirs(os.path.join(consensus_dir, gene), exist_ok=True)
        rid, seq = consensus_ref_record(gi)
        with open(os.path.join(consensus_dir, gene, "ref.fa"), "w") as f:
            f.write(f">{rid} reference\n")
            for i in range(0, len(seq), 70):
                f.write(seq[i:i + 70] + "\n")


def consensus_model():
    """Seeded gblinear model over the 20030-wide legacy layout: (weights float32 [20030], bias, base_score)."""
    rng = np.random.default_rng(CONSENSUS_MODEL_SEED)
    return (rng.standard_normal(20030) * 0.01).astype(np.float32), np.float32(-0.125), np.float32(0.5)
