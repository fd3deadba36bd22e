"""Degenerate inputs through the drop-in CLIs on a B200: one variant with one / three shifts, zero and one gene."""
import os

import numpy as np
import pytest

from expecto_b200 import chromatin, hdf5
from expecto_b200 import compute_expecto_features as cef
from expecto_b200.genome import PackedGenome

pytestmark = pytest.mark.gpu

ANNO_HEADER = "id,symbol,seqnames,strand,TSS,CAGE_representative_TSS,type\n"
TSS_FILE = "\tens_id\tchrom\ttss\tstrand\tx\tis_default_tss\n0\tNOPE\tchr1\t1\t+\t0\tTrue\n"


@pytest.mark.parametrize("maxshift", [0, 200])
def test_single_variant_single_shift(engine, small_genome, tmp_path, maxshift):
    """--maxshift 0 is one shift (chromatin.py:243 gives [0]); segment mode and the per-window path write the same
    files with the same bits."""
    pg = PackedGenome.from_sequences(small_genome)
    chrom = next(iter(small_genome))
    ref = small_genome[chrom][9999].upper().replace("N", "A")
    vcf = tmp_path / "one.vcf"
    vcf.write_text(f"{chrom}\t10000\tv\t{ref}\t{'C' if ref != 'C' else 'G'}\n")
    got = {}
    for mode in ("segments", "windows"):
        out = tmp_path / mode
        args = chromatin.build_parser().parse_args([str(vcf), "--output_dir", str(out), "--maxshift", str(maxshift),
                                                    "--mode", mode])
        assert chromatin.run(args, engine=engine, genome=pg)["n_variants"] == 1
        names = sorted(os.listdir(out))
        assert names == sorted([f"snps.shift_{s}.diff.h5" for s in chromatin.shift_list(maxshift)] + ["snps_hg19.vcf"])
        got[mode] = {}
        for s in chromatin.shift_list(maxshift):
            with hdf5.File(str(out / f"snps.shift_{s}.diff.h5")) as f:
                got[mode][s] = {k: np.array(f[k][...], copy=True) for k in f}
                assert got[mode][s]["diff"].shape == (2, 2002)
                assert np.array_equal(got[mode][s]["diff"], got[mode][s]["alt"] - got[mode][s]["ref"])
    for s in got["segments"]:
        for k in ("ref", "alt", "diff"):
            assert np.array_equal(got["segments"][s][k], got["windows"][s][k])
    assert np.abs(got["windows"][0]["diff"]).max() > 0


def test_zero_and_one_gene(engine, small_genome, tmp_path):
    pg = PackedGenome.from_sequences(small_genome)
    chrom = next(iter(small_genome))
    (tmp_path / "tss.tsv").write_text(TSS_FILE)
    (tmp_path / "anno0.csv").write_text(ANNO_HEADER)
    tss = len(small_genome[chrom]) // 2
    (tmp_path / "anno1.csv").write_text(ANNO_HEADER + f"ENSG1,G1,{chrom},-,{tss},{tss},protein_coding\n")
    a0 = cef.build_parser().parse_args([str(tmp_path / "anno0.csv"), str(tmp_path / "tss.tsv"), "-o", str(tmp_path / "f0")])
    x0 = np.load(cef.run(a0, engine=engine, genome=pg))
    assert x0.shape == (0, 20020) and x0.dtype == np.float64
    a1 = cef.build_parser().parse_args([str(tmp_path / "anno1.csv"), str(tmp_path / "tss.tsv"), "-o", str(tmp_path / "f1")])
    x1 = np.load(cef.run(a1, engine=engine, genome=pg))
    assert x1.shape == (1, 20020) and x1.dtype == np.float64 and np.isfinite(x1).all() and x1.max() > 0
