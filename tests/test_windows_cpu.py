"""Host logic of kernel 1 against the oracle and the reference-generated fixture (bit-exact)."""
import os

import numpy as np
import pytest

from expecto_b200 import synth
from expecto_b200.genome import PackedGenome, encode_bases
from expecto_b200.windows import (AlleleTable, allele_match_counts, chromatin_window_descs, crop_bounds, tss_descs,
                                  variant_descs)
from helpers import decode_descs
from oracle import beluga_oracle as O
from oracle import golden_inputs as GI


def test_crop_bounds_matches_python_slice():
    import math
    for total in list(range(0, 40)) + list(range(1960, 2300)):
        lo = int(math.floor((total - 2000) / 2.0))
        hi = int(math.floor(total - (total - 2000) / 2.0))
        s, e, _ = slice(lo, hi).indices(total)
        a, b = crop_bounds(np.array([total]))
        assert (int(a[0]), int(max(b[0], a[0]))) == (s, max(e, s)), total


def test_descriptors_match_reference_fixture(small_genome, golden):
    g = golden("encode_golden.npz")
    rows = [[r[0], int(r[1]), r[2], r[3], r[4]] for r in g["rows"]]
    assert rows == GI.edge_variants(small_genome), "fixture inputs drifted"
    pg = PackedGenome.from_sequences(small_genome)
    chroms = [r[0] for r in rows]
    pos = np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    tab = AlleleTable(refs + alts)
    n = len(rows)
    for k, (shift, which) in enumerate(zip(g["shifts"], g["which"])):
        la, lo = (tab.lengths[:n], tab.offsets[:n]) if which == "ref" else (tab.lengths[n:], tab.offsets[n:])
        fwd = variant_descs(pg, chroms, pos, tab.lengths[:n], la, lo, int(shift), rc=False)
        rc = variant_descs(pg, chroms, pos, tab.lengths[:n], la, lo, int(shift), rc=True)
        ours = decode_descs(pg, np.concatenate([fwd, rc]), tab.codes)
        want = np.unpackbits(g["packed"][k], axis=1)[:, :8000].reshape(2 * n, 4, 2000)
        assert np.array_equal(ours, want), (shift, which)


def test_descriptors_match_oracle_all_shifts(small_genome):
    rows = GI.edge_variants(small_genome)
    pg = PackedGenome.from_sequences(small_genome)
    dg = O.DictGenome(small_genome)
    chroms = [r[0] for r in rows]
    pos = np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    # the default range, then shifts around and beyond +-1049, where the variant leaves the 2100-bp fetch and the
    # reference's splice follows Python's negative / past-the-end slice semantics (chromatin.py:203-209)
    # (oracle/make_golden.py asserts the oracle equals the reference's own fetchSeqs at these shifts)
    for shift in list(O.SHIFTS_CHROMATIN(800)) + GI.EXTREME_SHIFTS:
        descs, codes = chromatin_window_descs(pg, chroms, pos, refs, alts, shift)
        ours = decode_descs(pg, descs, codes).reshape(len(rows), 4, 4, 2000)
        seqs = [O.fetch_seqs(dg, c, int(p), r, a, shift=shift) for c, p, r, a in zip(chroms, pos, refs, alts)]
        ref_enc = O.encode_seqs([s[0] for s in seqs])
        alt_enc = O.encode_seqs([s[1] for s in seqs])
        n = len(rows)
        assert np.array_equal(ours[:, 0], ref_enc[:n]) and np.array_equal(ours[:, 1], ref_enc[n:])
        assert np.array_equal(ours[:, 2], alt_enc[:n]) and np.array_equal(ours[:, 3], alt_enc[n:])
    want = (sum(s[2] for s in [O.fetch_seqs(dg, c, int(p), r, a) for c, p, r, a in zip(chroms, pos, refs, alts)]),
            sum(s[3] for s in [O.fetch_seqs(dg, c, int(p), r, a) for c, p, r, a in zip(chroms, pos, refs, alts)]),
            len(rows))
    assert allele_match_counts(pg, chroms, pos, refs, alts) == want


def test_tss_windows_match_oracle(small_genome):
    pg = PackedGenome.from_sequences(small_genome)
    dg = O.DictGenome(small_genome)
    shifts = list(range(-2000, 2000, 200))
    descs = tss_descs(pg, ["chr1", "chr2"], [30000, 20000], [1, -1], shifts)
    ours = decode_descs(pg, descs, np.zeros(1, np.uint8)).reshape(2, len(shifts), 2, 4, 2000)
    for gi, (chrom, tss, strand) in enumerate([("chr1", 30000, 1), ("chr2", 20000, -1)]):
        enc = O.encode_seqs(O.tss_windows(dg, chrom, tss, strand, shifts), concat_rc=False)
        assert np.array_equal(ours[gi, :, 0], enc)
        assert np.array_equal(ours[gi, :, 1], enc[:, ::-1, ::-1])


def test_bad_allele_raises_keyerror():
    with pytest.raises(KeyError):
        encode_bases("ACGR")
    assert list(encode_bases("AGCTNnH-acgt")) == [0, 1, 2, 3, 4, 4, 4, 4, 0, 2, 1, 3]


def test_invalid_genome_char_detected():
    from expecto_b200.windows import check_valid_bases
    seq = "ACGT" * 2000
    seq = seq[:5000] + "R" + seq[5001:]
    pg = PackedGenome.from_sequences({"chr1": seq})
    d_ok = tss_descs(pg, ["chr1"], [2000], [1], [0])
    check_valid_bases(pg, d_ok)
    d_bad = tss_descs(pg, ["chr1"], [5200], [1], [0])
    with pytest.raises(KeyError):
        check_valid_bases(pg, d_bad)


@pytest.mark.parametrize("maxshift", [800, 1000, 400])
def test_segment_descriptors_cover_every_shifted_window(small_genome, maxshift):
    """Segment mode: each window of every shift is the 2000-bp slice at offset 200*k of its segment."""
    from expecto_b200.windows import chromatin_segment_descs, chromatin_shift_list, segment_eligible
    rows = GI.edge_variants(small_genome)
    ok = segment_eligible([len(r[3]) for r in rows], [len(r[4]) for r in rows], maxshift)
    assert 0 < ok.sum() < len(rows)            # the 120-bp deletion etc. are not eligible
    assert not ok[[r[2] for r in rows].index("del120")] and ok[[r[2] for r in rows].index("ins150")]
    assert not segment_eligible([1, 60], [1, 1], 1200).any()          # variant outside the fetch at +-1200
    assert list(segment_eligible([1, 50, 51], [1, 1, 1], 1000)) == [True, True, False]   # right flank at shift -1000
    rows = [r for r, e in zip(rows, ok) if e]
    pg = PackedGenome.from_sequences(small_genome)
    chroms, pos = [r[0] for r in rows], np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    S = 2 * maxshift // 200 + 1
    L = 2000 + 2 * maxshift
    sdescs, scodes, k_of = chromatin_segment_descs(pg, chroms, pos, refs, alts, maxshift)
    seg = decode_descs(pg, sdescs, scodes, length=L).reshape(len(rows), 4, 4, L)
    for si, shift in enumerate(chromatin_shift_list(maxshift)):
        wdescs, wcodes = chromatin_window_descs(pg, chroms, pos, refs, alts, shift)
        win = decode_descs(pg, wdescs, wcodes).reshape(len(rows), 4, 4, 2000)
        k = int(k_of[si])
        assert k == (shift + maxshift) // 200
        for q in range(4):
            off = 200 * (S - 1 - k) if q % 2 else 200 * k
            assert np.array_equal(win[:, q], seg[:, q, :, off:off + 2000]), (shift, q)


def test_delta_patches_are_slices_of_the_alt_segments(small_genome):
    """Substitution patches: the alt patch descriptors decode to positions [16*row0, 16*row0 + patch_len) of the
    alt segments on both strands, the ref descriptors are the plain ref segments, and outside the patch's valid
    receptive-field rows ref and alt segments agree (what the gather relies on)."""
    from expecto_b200.windows import (RECEPTIVE_FIELD, chromatin_delta_descs, chromatin_segment_descs, delta_eligible,
                                      delta_patch_geometry, segment_geometry, sub_descs)
    rows = GI.edge_variants(small_genome)
    ok = delta_eligible([len(r[3]) for r in rows], [len(r[4]) for r in rows])
    assert 0 < ok.sum() < len(rows)
    rows = [r for r, e in zip(rows, ok) if e]
    # an MNV of the longest handled length on top of the fixture's substitutions
    c, p = rows[0][0], rows[0][1]
    rows.append((c, p, "mnv16", small_genome[c][p - 1:p + 15].upper(), "ACGTTGCAACGTTGCA"))
    pg = PackedGenome.from_sequences(small_genome)
    chroms, pos = [r[0] for r in rows], np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    n = len(rows)
    rd, pd, codes, k_of, row0, plen = chromatin_delta_descs(pg, chroms, pos, refs, alts, 800)
    assert (row0, plen) == delta_patch_geometry(800) and plen % 16 == 0
    sd, scodes, k2 = chromatin_segment_descs(pg, chroms, pos, refs, alts, 800)
    assert np.array_equal(k_of, k2) and np.array_equal(codes, scodes)
    seg = decode_descs(pg, sd, scodes, length=3600).reshape(n, 4, 4, 3600)
    assert np.array_equal(decode_descs(pg, rd, codes, length=3600).reshape(n, 2, 4, 3600), seg[:, 0:2])
    patch = decode_descs(pg, pd, codes, length=plen).reshape(n, 2, 4, plen)
    assert np.array_equal(patch, seg[:, 2:4, :, 16 * row0:16 * row0 + plen])
    # every position where alt differs from ref lies inside the receptive field of valid patch rows only
    g = segment_geometry(plen)
    for strand in range(2):
        diff = np.flatnonzero((seg[:, strand] != seg[:, 2 + strand]).any(axis=(0, 1)))
        if diff.size == 0:
            continue
        for phase_off, valid in ((0, g["valid_a"]), (8, g["valid_b"])):
            first = (diff.min() - phase_off - (RECEPTIVE_FIELD - 1) + 15) // 16      # first row that sees a change
            last = (diff.max() - phase_off) // 16
            assert row0 <= first and last < row0 + valid, (strand, phase_off, first, last)
    # general sub-ranges, including ones that start inside the allele or the right flank
    for start, length in ((0, 3600), (1790, 64), (1800, 32), (1805, 200), (3000, 600)):
        sub = sub_descs(sd, start, length, 3600)
        got = decode_descs(pg, sub, scodes, length=length).reshape(n, 4, 4, length)
        assert np.array_equal(got, seg[..., start:start + length]), (start, length)


def test_tss_segment_descriptors(small_genome):
    from expecto_b200.windows import tss_segment_descs
    pg = PackedGenome.from_sequences(small_genome)
    shifts = list(range(-2000, 2000, 200))
    genes = [("chr1", 30000, 1), ("chr2", 20000, -1)]
    wd = tss_descs(pg, [g[0] for g in genes], [g[1] for g in genes], [g[2] for g in genes], shifts)
    win = decode_descs(pg, wd, np.zeros(1, np.uint8)).reshape(2, len(shifts), 2, 4, 2000)
    sd, k_of, seg_len = tss_segment_descs(pg, [g[0] for g in genes], [g[1] for g in genes], [g[2] for g in genes], shifts)
    assert seg_len == 2000 + 200 * (len(shifts) - 1)
    seg = decode_descs(pg, sd, np.zeros(1, np.uint8), length=seg_len).reshape(2, 2, 4, seg_len)
    S = len(shifts)
    for gi in range(2):
        for i in range(S):
            for q in range(2):
                k = int(k_of[i])
                if int(sd.reshape(2, 2)[gi, q]["flags"]) & 2:
                    k = S - 1 - k
                assert np.array_equal(win[gi, i, q], seg[gi, q, :, 200 * k:200 * k + 2000]), (gi, i, q)


def test_allele_match_counts_vectorised_vs_oracle(small_genome):
    from expecto_b200 import synth
    rows = synth.make_vcf_rows(small_genome, 400, seed=9, odd_names=False, margin=3000)
    rows += [["chr1", 5000, "n_ref", "N", "A"], ["chr1", 5001, "lower", small_genome["chr1"][5000].lower(), "n"]]
    # put the first two on genome N / lower-case positions if there are any
    npos = [i for i, c in enumerate(small_genome["chr1"][3000:50000]) if c == "N"][:3]
    rows += [["chr1", 3001 + i, "on_N", "N", "N"] for i in npos]
    pg = PackedGenome.from_sequences(small_genome)
    dg = O.DictGenome(small_genome)
    chroms, pos = [r[0] for r in rows], np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    flags = [O.fetch_seqs(dg, c, int(p), r, a)[2:] for c, p, r, a in zip(chroms, pos, refs, alts)]
    want = (sum(f[0] for f in flags), sum(f[1] for f in flags), len(rows))
    assert allele_match_counts(pg, chroms, pos, refs, alts) == want
    assert want[0] > 300 and len(npos) > 0


def test_packed_genome_from_fasta_roundtrip(tmp_path, small_genome):
    from expecto_b200 import synth
    fa = tmp_path / "g.fa"
    synth.write_fasta(str(fa), small_genome, width=60)
    pg = PackedGenome.from_fasta(str(fa))
    assert (tmp_path / "g.fa.bx2.npz").exists()
    again = PackedGenome.from_fasta(str(fa))  # served from the cache
    ref = PackedGenome.from_sequences(small_genome)
    for g in (pg, again):
        assert g.names == ref.names and g.offsets == ref.offsets and g.lengths == ref.lengths
        assert np.array_equal(g.bases2, ref.bases2) and np.array_equal(g.nmask, ref.nmask)
    s = small_genome["chr2"]
    assert pg.upper_string(pg.offsets["chr2"] + 100, 50) == s[100:150].upper()


def test_packed_layout_against_a_per_base_loop(tmp_path, monkeypatch):
    """The word layout of genome.py's docstring, checked base by base with plain Python (independent of the packer's
    byte tricks): CRLF line ends, ragged line widths, lower case, every zero-column character, characters outside
    ``mydict`` (recorded as invalid, chromatin.py:155-160 would raise KeyError), contigs that end inside a word, and
    contigs longer than one packing pass."""
    from expecto_b200 import genome as G
    monkeypatch.setattr(G.PackedGenome, "_CHUNK", 64)   # several passes per contig
    rng = np.random.default_rng(11)
    alphabet = np.frombuffer(b"ACGTacgtNnH-RY", dtype=np.uint8)
    seqs = {"chrA": alphabet[rng.integers(0, len(alphabet), size=333)].tobytes().decode(),
            "chrB": "ACGT" * 16 + "N",
            "chrC": alphabet[rng.integers(0, 8, size=64 * 3)].tobytes().decode()}
    fa = tmp_path / "ragged.fa"
    with open(fa, "wb") as f:
        for name, s in seqs.items():
            f.write(f">{name} some description\r\n".encode())
            i, widths = 0, [7, 60, 1, 33]
            while i < len(s):
                w = widths[(i + len(name)) % len(widths)]
                f.write(s[i:i + w].encode() + (b"\r\n" if name != "chrB" else b"\n"))
                i += w
    g = G.PackedGenome.from_fasta(str(fa), cache=False)
    assert g.names == list(seqs) and g.lengths == {k: len(v) for k, v in seqs.items()}
    row = {"A": 0, "G": 1, "C": 2, "T": 3}
    want_invalid = []
    for name, s in seqs.items():
        off = g.offsets[name]
        assert off % 32 == 0
        for i, ch in enumerate(s):
            j = off + i
            two = (int(g.bases2[j >> 4]) >> (2 * (j & 15))) & 3
            msk = (int(g.nmask[j >> 5]) >> (j & 31)) & 1
            if ch.upper() in row:
                assert (two, msk) == (row[ch.upper()], 0), (name, i, ch)
            else:
                assert msk == 1, (name, i, ch)
                if ch not in "NnH-":
                    want_invalid.append(j)
        # padding after the contig reads as zero columns
        assert all((int(g.nmask[j >> 5]) >> (j & 31)) & 1 for j in range(off + len(s), off + len(s) + 64))
    assert g.invalid.tolist() == want_invalid
    same = G.PackedGenome.from_sequences(seqs)
    assert np.array_equal(same.bases2, g.bases2) and np.array_equal(same.nmask, g.nmask)
    assert np.array_equal(same.invalid, g.invalid)


def test_fasta_cache_is_written_atomically(tmp_path, small_genome):
    from expecto_b200 import synth
    fa = tmp_path / "g.fa"
    synth.write_fasta(str(fa), small_genome, width=50)
    PackedGenome.from_fasta(str(fa))
    left = sorted(p.name for p in tmp_path.iterdir())
    assert left == ["g.fa", "g.fa.bx2.npz"]   # no temporary file left behind


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_random_variants_descriptors_match_oracle(small_genome, seed):
    """Randomised host-side version of tools/fuzz_segments.py: SNVs, MNVs, insertions up to 300 bp, deletions up to
    120 bp, long ref alleles, ref mismatches, N / lower-case alleles at random shifts (incl. beyond +-1049): the window
    descriptors decode to exactly what fetchSeqs + encodeSeqs (chromatin.py:138-209, restated in the oracle) produce."""
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "fuzz_segments", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "fuzz_segments.py"))
    fuzz = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fuzz)
    rng = np.random.default_rng(100 + seed)
    rows = fuzz.random_variants(rng, small_genome, 40, margin=3200)
    pg = PackedGenome.from_sequences(small_genome)
    dg = O.DictGenome(small_genome)
    chroms = [r[0] for r in rows]
    pos = np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    shifts = [0] + [int(s) for s in rng.choice(np.arange(-2200, 2201, 50), size=4, replace=False)]
    n = len(rows)
    for shift in shifts:
        descs, codes = chromatin_window_descs(pg, chroms, pos, refs, alts, shift)
        ours = decode_descs(pg, descs, codes).reshape(n, 4, 4, 2000)
        seqs = [O.fetch_seqs(dg, c, int(p), r, a, shift=shift) for c, p, r, a in zip(chroms, pos, refs, alts)]
        ref_enc = O.encode_seqs([s[0] for s in seqs])
        alt_enc = O.encode_seqs([s[1] for s in seqs])
        assert np.array_equal(ours[:, 0], ref_enc[:n]) and np.array_equal(ours[:, 1], ref_enc[n:]), shift
        assert np.array_equal(ours[:, 2], alt_enc[:n]) and np.array_equal(ours[:, 3], alt_enc[n:]), shift


def test_windows_running_off_a_contig_read_zero_columns_not_the_neighbour():
    """with 4096 bases of padding between contigs, a compute_expecto_features window near a contig edge
    (TSS +-20,000 shift +-1,000 bp, compute_expecto_features.py:105-111) landed inside the neighbouring contig.  The
    padding now exceeds that reach: every off-contig position must decode as code 4 (an all-zero column), for plain TSS
    windows, TSS segments and variant windows at the largest chromatin shift."""
    from expecto_b200.genome import PAD
    from expecto_b200.windows import tss_descs, tss_segment_descs, variant_descs
    assert PAD >= 21000 + 32 and PAD % 32 == 0
    genome = {"chrA": "ACGT" * 10000, "chrB": "ACGT" * 15000, "chrC": "TTGA" * 2000}
    pg = PackedGenome.from_sequences(genome)
    shifts = np.arange(-20000, 20000, 200)
    for chrom, tss, strand in (("chrB", 5000, 1), ("chrB", 5000, -1), ("chrB", 59000, 1), ("chrA", 1, -1),
                               ("chrC", 7999, 1)):
        d = tss_descs(pg, [chrom], [tss], [strand], shifts).reshape(len(shifts), 2)
        off, ln = pg.offsets[chrom], pg.lengths[chrom]
        for si, s in enumerate(shifts):
            codes = pg.codes(int(d[si, 0]["g1"]), 2000)
            start0 = tss + int(s) * strand - 999 - 1           # 0-based contig coordinate of the window's first base
            inside = (np.arange(start0, start0 + 2000) >= 0) & (np.arange(start0, start0 + 2000) < ln)
            assert np.all(codes[~inside] == 4), (chrom, tss, strand, int(s))
            want = pg.codes(off + max(start0, 0), 1)[0] if inside.any() else None
            assert np.all(codes[inside] < 4) and (want is None or want < 4)
        sd, _, seg_len = tss_segment_descs(pg, [chrom], [tss], [strand], shifts)
        codes = pg.codes(int(sd[0]["g1"]), seg_len)
        g0 = int(sd[0]["g1"]) - off
        inside = (np.arange(g0, g0 + seg_len) >= 0) & (np.arange(g0, g0 + seg_len) < ln)
        assert np.all(codes[~inside] == 4) and np.all(codes[inside] < 4)
    for shift in (800, -800):     # a variant two bases from a contig start / end
        for chrom, pos in (("chrB", 2), ("chrA", 39999)):
            d = variant_descs(pg, [chrom], [pos], np.array([1]), np.array([1]), np.array([0]), shift, rc=False)[0]
            off, ln = pg.offsets[chrom], pg.lengths[chrom]
            for g, n in ((int(d["g1"]), int(d["len1"])), (int(d["g2"]), int(d["len2"]))):
                idx = np.arange(g, g + n) - off
                codes = pg.codes(g, n)
                assert np.all(codes[(idx < 0) | (idx >= ln)] == 4)


def test_cache_with_narrow_padding_is_rebuilt(tmp_path, small_genome):
    """A .bx2.npz written before the padding was widened (no ``pad`` field, 4096 bases) must not be reused."""
    import expecto_b200.genome as G
    fa = str(tmp_path / "g.fa")
    synth.write_fasta(fa, {k: v[:5000] for k, v in small_genome.items()})
    g = PackedGenome.from_fasta(fa)
    z = dict(np.load(fa + ".bx2.npz"))
    assert int(z["pad"]) == G.PAD
    z.pop("pad")
    z["offsets"] = z["offsets"] - 1                       # a stale layout: would be wrong if it were loaded
    np.savez(fa + ".bx2.npz", **z)
    g2 = PackedGenome.from_fasta(fa)
    assert g2.offsets == g.offsets and g2.pad == G.PAD
