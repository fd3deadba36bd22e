"""make_closest_genes_file.py drop-in vs the UNMODIFIED reference script's outputs (oracle/make_golden.py ran it
on a 944-gene slice of the reference's geneanno.csv and 94 variants incl. ties, exact-TSS hits and shared TSSs)."""
import numpy as np
import pytest

from expecto_b200 import make_closest_genes_file as mcg


@pytest.mark.parametrize("mode,flags", [("closest", []), ("rf", ["--all_in_receptive_field"])])
def test_outputs_identical_to_reference_script(golden, tmp_path, mode, flags):
    g = golden("closest_genes_golden.npz")
    (tmp_path / "geneanno.csv").write_text(str(g["geneanno_csv"]))
    (tmp_path / "in.vcf").write_text(str(g["vcf"]))
    args = mcg.build_parser().parse_args([str(tmp_path / "in.vcf"), "--geneanno_file", str(tmp_path / "geneanno.csv"),
                                          "-o", str(tmp_path / "out")] + flags)
    info = mcg.run(args)
    assert (tmp_path / "out" / "closest_genes.tsv").read_text() == str(g[f"{mode}_tsv"])
    assert (tmp_path / "out" / "snps_hg19.vcf").read_text() == str(g[f"{mode}_vcf"])
    assert info["rows"] == str(g[f"{mode}_tsv"]).count("\n")


def test_tie_breaks_like_argmin():
    idx = mcg.GeneIndex(["chr1"] * 5, [100, 300, 300, 500, 100], ["+", "-", "+", "+", "-"])
    # equidistant between 100 and 300 -> annotation row 0 (first in file order); on a shared TSS -> lower row
    assert list(idx.closest("chr1", np.array([200, 300, 99, 1000, 400]))) == [0, 1, 0, 3, 1]
    assert list(idx.in_receptive_field("chr1", 20400)) == [0, 1, 2, 3, 4]  # all within their strand-specific bounds
    assert list(idx.in_receptive_field("chr1", 100 + 20301)) == [1, 2, 3, 4]  # '+' row 0 out (d = 20301 > 20300)


def test_unknown_contig_raises_like_reference():
    idx = mcg.GeneIndex(["chr1"], [100], ["+"])
    with pytest.raises(ValueError):
        mcg.match_genes(["chr9"], [5], idx, False)
