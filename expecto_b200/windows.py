"""Window descriptors: the host half of kernel 1.

``fetchSeqs`` (chromatin.py:175-209) fetches 2100 bp around ``pos+shift``, splices the VCF allele in at
``mutpos = 1049 - shift`` and ``encodeSeqs`` (chromatin.py:164) centre-crops the result to 2000 bp.
Both are pure index arithmetic, so each model input is described to the GPU as
``genome[g1:g1+len1] + allele[aoff:aoff+alen] + genome[g2:g2+len2]`` (bx_window_desc) and the bases
themselves never leave HBM.  Everything here is vectorised numpy over the variant axis.
"""
from __future__ import annotations

import numpy as np

from ._lib import DESC_DTYPE
from .genome import PackedGenome, encode_bases

INPUTSIZE = 2000
WINDOWSIZE = INPUTSIZE + 100  # chromatin.py:117,202


def crop_bounds(total: np.ndarray, inputsize: int = INPUTSIZE):
    """``line[int(floor((len-inputsize)/2.0)) : int(floor(len-(len-inputsize)/2.0))]`` with Python slice
    semantics (negative start counts from the end), chromatin.py:164."""
    total = np.asarray(total, dtype=np.int64)
    a = np.floor_divide(total - inputsize, 2)
    b = np.floor_divide(total + inputsize, 2)  # floor(len - (len - inputsize)/2) == floor((len + inputsize)/2)
    a = np.where(a < 0, np.maximum(a + total, 0), a)
    a = np.minimum(a, total)
    b = np.clip(b, 0, total)
    b = np.maximum(b, a)
    return a, b


class AlleleTable:
    """Concatenated allele codes (0..3 = A,G,C,T; 4 = zero column) with per-allele offsets."""

    def __init__(self, alleles: list[str]):
        lens = np.fromiter((len(a) for a in alleles), dtype=np.int64, count=len(alleles))
        self.lengths = lens
        self.offsets = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64) if len(alleles) else lens
        joined = "".join(alleles)
        self.codes = encode_bases(joined) if joined else np.zeros(0, dtype=np.uint8)  # KeyError on bad chars


def variant_descs(genome: PackedGenome, chroms, pos, ref_len: np.ndarray, allele_len: np.ndarray,
                  allele_off: np.ndarray, shift: int, rc: bool, inputsize: int = INPUTSIZE) -> np.ndarray:
    """Descriptors of the windows fetchSeqs+encodeSeqs build for one allele (ref or alt) of each variant.

    chroms: sequence of contig names (already normalised to ``chr*``); pos: 1-based positions;
    ref_len: len(ref) per variant (the spliced-out span is always len(ref), chromatin.py:209);
    allele_len/off: the allele that is spliced in (ref itself for the ref window)."""
    n = len(pos)
    windowsize = inputsize + 100
    pos = np.asarray(pos, dtype=np.int64)
    ref_len = np.asarray(ref_len, dtype=np.int64)
    la = np.asarray(allele_len, dtype=np.int64)
    coff = np.fromiter((genome.offsets[c] for c in chroms), dtype=np.int64, count=n)
    half = int(windowsize / 2 - 1)                       # 1049
    mutpos = int(windowsize / 2 - 1 - shift)             # chromatin.py:203
    s0 = coff + pos + shift - half - 1                   # global index of seq[0] (1-based start -> 0-based)
    # Python slice semantics of seq[:mutpos] and seq[mutpos+len(ref):] on the 2100-character fetch, chromatin.py:209:
    # beyond |shift| = 1049 the variant lies outside the fetch, mutpos goes negative (counts from the end) or past the
    # end, and the reference still splices the alleles in at that accidental place
    def norm(x):
        x = np.asarray(x, dtype=np.int64)
        return np.where(x < 0, np.maximum(0, windowsize + x), np.minimum(x, windowsize))
    len_p = np.full(n, int(norm(mutpos)), dtype=np.int64)  # seq[:mutpos]
    suffix0 = norm(mutpos + ref_len)                         # first character of seq[mutpos+len(ref):]
    len_s = windowsize - suffix0
    total = len_p + la + len_s
    a, b = crop_bounds(total, inputsize)
    d = np.zeros(n, dtype=DESC_DTYPE)
    p_lo, p_hi = np.minimum(a, len_p), np.minimum(b, len_p)
    d["len1"] = np.maximum(0, p_hi - p_lo)
    d["g1"] = s0 + p_lo
    al_lo, al_hi = np.clip(a - len_p, 0, la), np.clip(b - len_p, 0, la)
    d["alen"] = np.maximum(0, al_hi - al_lo)
    d["aoff"] = np.asarray(allele_off, dtype=np.int64) + al_lo
    s_lo, s_hi = np.clip(a - len_p - la, 0, len_s), np.clip(b - len_p - la, 0, len_s)
    d["len2"] = np.maximum(0, s_hi - s_lo)
    d["g2"] = s0 + suffix0 + s_lo
    d["flags"] = 1 if rc else 0
    return d


def tss_descs(genome: PackedGenome, chroms, tss, strand, shifts, windowsize: int = INPUTSIZE) -> np.ndarray:
    """compute_expecto_features.py:105-111: for gene g and shift s the window is
    ``start = tss + s*strand - 999 .. stop = tss + s*strand + 1000`` (1-based inclusive), NOT reverse
    complemented for '-' genes.  Returns descriptors ordered (gene, shift, {fwd, rc})."""
    n = len(tss)
    tss = np.asarray(tss, dtype=np.int64)
    strand = np.asarray(strand, dtype=np.int64)
    shifts = np.asarray(shifts, dtype=np.int64)
    coff = np.fromiter((genome.offsets[c] for c in chroms), dtype=np.int64, count=n)
    start0 = (coff + tss)[:, None] + shifts[None, :] * strand[:, None] - int(windowsize / 2 - 1) - 1
    d = np.zeros((n, len(shifts), 2), dtype=DESC_DTYPE)
    d["g1"] = start0[:, :, None]
    d["len1"] = windowsize
    d["flags"][:, :, 1] = 1
    return d.reshape(-1)


def variant_columns(genome: PackedGenome, chroms, pos, refs: list[str], alts: list[str]) -> dict:
    """Whole-VCF column arrays for the in-library variant loop (include/beluga_b200.h: bx_chromatin_open): the global
    packed-genome index of every variant's first base, allele lengths, and the allele codes with their offsets.
    One vectorised pass; the descriptors themselves are built on the device (csrc/variants.cuh)."""
    n = len(pos)
    tab = AlleleTable(list(refs) + list(alts))
    if tab.codes.size >= 2 ** 31:
        raise ValueError("allele table exceeds 2 GiB")
    coff = np.fromiter((genome.offsets[c] for c in chroms), dtype=np.int64, count=n)
    return {"n": n,
            "gpos": np.ascontiguousarray(coff + np.asarray(pos, dtype=np.int64) - 1),
            "ref_len": np.ascontiguousarray(tab.lengths[:n], dtype=np.int32),
            "alt_len": np.ascontiguousarray(tab.lengths[n:], dtype=np.int32),
            "ref_aoff": np.ascontiguousarray(tab.offsets[:n], dtype=np.int32),
            "alt_aoff": np.ascontiguousarray(tab.offsets[n:], dtype=np.int32),
            "alleles": np.ascontiguousarray(tab.codes, dtype=np.uint8)}


def variant_extent_invalid(genome: PackedGenome, cols: dict, maxshift: int) -> int:
    """Index of the first variant whose fetches (chromatin.py:202-206: 2100 bp around pos + shift for every shift)
    touch a genome character outside encodeSeqs' dictionary, or -1.  Conservative by at most the 50-bp crop margin:
    callers fall back to the per-chunk host check, which follows the cropped windows exactly."""
    if genome.invalid.size == 0 or cols["n"] == 0:
        return -1
    reach = maxshift + 1050
    return genome.any_invalid(cols["gpos"] - reach, np.full(cols["n"], 2 * reach + 1, dtype=np.int64) + cols["ref_len"])


def chromatin_window_descs(genome: PackedGenome, chroms, pos, refs: list[str], alts: list[str], shift: int):
    """All four windows per variant in the order the fused output kernels expect:
    (variant, {ref fwd, ref rc, alt fwd, alt rc}).  Returns (descs [4N], allele codes)."""
    n = len(pos)
    tab = AlleleTable(list(refs) + list(alts))
    rl, al = tab.lengths[:n], tab.lengths[n:]
    ro, ao = tab.offsets[:n], tab.offsets[n:]
    out = np.zeros((n, 4), dtype=DESC_DTYPE)
    out[:, 0] = variant_descs(genome, chroms, pos, rl, rl, ro, shift, rc=False)
    out[:, 1] = variant_descs(genome, chroms, pos, rl, rl, ro, shift, rc=True)
    out[:, 2] = variant_descs(genome, chroms, pos, rl, al, ao, shift, rc=False)
    out[:, 3] = variant_descs(genome, chroms, pos, rl, al, ao, shift, rc=True)
    return out.reshape(-1), tab.codes


def check_valid_bases(genome: PackedGenome, descs: np.ndarray) -> None:
    """Raise KeyError, as ``mydict[c]`` would (chromatin.py:166), if any window touches a genome
    character outside encodeSeqs' dictionary."""
    for gs, ln in ((descs["g1"], descs["len1"]), (descs["g2"], descs["len2"])):
        i = genome.any_invalid(gs.astype(np.int64), ln.astype(np.int64))
        if i >= 0:
            raise KeyError(f"window {i} covers a base that encodeSeqs cannot encode")


def allele_match_counts(genome: PackedGenome, chroms, pos, refs: list[str], alts: list[str]):
    """The three diagnostics chromatin.py:256-260 prints at shift 0: how many variants have
    ``seq[mutpos:mutpos+len(ref)].upper() == ref.upper()`` / ``== alt.upper()``, and the count.
    Single-base ref alleles (the bulk of any VCF) are compared vectorised on the packed codes; longer ones
    go through the string comparison."""
    n = len(refs)
    if n == 0:
        return 0, 0, 0
    pos = np.asarray(pos, dtype=np.int64)
    coff = np.fromiter((genome.offsets[c] for c in chroms), dtype=np.int64, count=n)
    rl = np.fromiter((len(r) for r in refs), dtype=np.int64, count=n)
    al = np.fromiter((len(a) for a in alts), dtype=np.int64, count=n)
    ref_ok = alt_ok = 0
    snv = rl == 1
    if snv.any():
        g = coff[snv] + pos[snv] - 1
        inside = (g >= 0) & (g < genome.n_bases)
        j = np.where(inside, g, 0)
        two = (genome.bases2[j >> 4] >> ((j & 15) * 2).astype(np.uint32)) & 3
        msk = (genome.nmask[j >> 5] >> (j & 31).astype(np.uint32)) & 1
        gcode = np.where((msk == 1) | ~inside, 4, two).astype(np.int64)   # 4 = 'N' after .upper()
        lut = np.full(256, -1, dtype=np.int64)                              # allele letter -> code, 'N' -> 4
        for ch, code in (("A", 0), ("G", 1), ("C", 2), ("T", 3), ("N", 4)):
            lut[ord(ch)] = lut[ord(ch.lower())] = code
        idx = np.flatnonzero(snv)
        rcode = lut[np.frombuffer("".join(refs[i] for i in idx).encode("latin-1", "replace"), dtype=np.uint8)]
        ref_ok += int(np.sum(rcode == gcode))
        a1 = al[snv] == 1   # a longer alt can never equal the one genome base
        if a1.any():
            acode = lut[np.frombuffer("".join(alts[i] for i in idx[a1]).encode("latin-1", "replace"), dtype=np.uint8)]
            alt_ok += int(np.sum(acode == gcode[a1]))
    for i in np.flatnonzero(~snv):
        g = genome.upper_string(int(coff[i] + pos[i] - 1), int(rl[i]))
        ref_ok += g == refs[i].upper()
        alt_ok += g == alts[i].upper()
    return int(ref_ok), int(alt_ok), n


# ------------------------------------------------------------------ segment mode (shared activations)
SHIFT_STEP = 200


def chromatin_shift_list(maxshift: int) -> list[int]:
    return [0] + list(range(-200, -maxshift - 1, -200)) + list(range(200, maxshift + 1, 200))  # chromatin.py:243


def segment_eligible(ref_len: np.ndarray, alt_len: np.ndarray, maxshift: int = 800) -> np.ndarray:
    """Variants whose 2*maxshift/200+1 shifted windows are all plain 2000-bp substrings of one spliced sequence:
    the centre crop must start inside the left flank for both alleles (deletions of at most 100 bp, chromatin.py:164),
    the variant must lie inside the 2100-bp fetch at every shift (maxshift <= 1000: mutpos = 1049 - shift >= 0,
    chromatin.py:203) and the right flank must exist at every shift (len(ref) <= 1050 - maxshift, i.e. 250 for the
    default range, chromatin.py:209).  The rest (rare long indels, exotic shift ranges) go through the per-window
    path, which follows the reference's Python slice semantics literally."""
    ref_len = np.asarray(ref_len, dtype=np.int64)
    alt_len = np.asarray(alt_len, dtype=np.int64)
    if maxshift > 1000:
        return np.zeros(ref_len.shape, dtype=bool)
    return (ref_len <= 1050 - maxshift) & (alt_len - ref_len >= -100) & (alt_len <= 1000) & (ref_len >= 1) & \
        (alt_len >= 0)


def _allele_segment(genome, coff, pos, ref_len, la, aoff, maxshift, rc):
    n = len(pos)
    seg_len = INPUTSIZE + 2 * maxshift
    a = np.floor_divide(100 + (la - ref_len), 2)             # crop start, chromatin.py:164
    v = coff + pos - 1                                        # global index of the variant's first base
    len1 = maxshift + 1049 - a                                # left flank inside the segment
    d = np.zeros(n, dtype=DESC_DTYPE)
    d["g1"] = v - len1
    d["len1"] = len1
    d["alen"] = la
    d["aoff"] = aoff
    d["g2"] = v + ref_len
    d["len2"] = seg_len - len1 - la
    d["flags"] = 3 if rc else 0                               # rc segments: reversed strand AND reversed offsets
    return d


def chromatin_segment_descs(genome: PackedGenome, chroms, pos, refs: list[str], alts: list[str], maxshift: int = 800):
    """Four segments per variant, order {ref fwd, ref rc, alt fwd, alt rc}: the sequences every shifted window of
    fetchSeqs + encodeSeqs (chromatin.py:175-209,164) is a substring of.  Window of shift s starts 200*k positions
    into a forward segment, k = (s + maxshift)/200, and 200*(n_shifts-1-k) into a reverse-complement segment.
    Returns (descs [4N], allele codes, k_of_shift int32 [n_shifts]); callers must pre-filter with segment_eligible."""
    n = len(pos)
    tab = AlleleTable(list(refs) + list(alts))
    rl, al = tab.lengths[:n], tab.lengths[n:]
    ro, ao = tab.offsets[:n], tab.offsets[n:]
    if not np.all(segment_eligible(rl, al, maxshift)):
        raise ValueError("chromatin_segment_descs needs variants pre-filtered with segment_eligible")
    pos = np.asarray(pos, dtype=np.int64)
    coff = np.fromiter((genome.offsets[c] for c in chroms), dtype=np.int64, count=n)
    out = np.zeros((n, 4), dtype=DESC_DTYPE)
    out[:, 0] = _allele_segment(genome, coff, pos, rl, rl, ro, maxshift, False)
    out[:, 1] = _allele_segment(genome, coff, pos, rl, rl, ro, maxshift, True)
    out[:, 2] = _allele_segment(genome, coff, pos, rl, al, ao, maxshift, False)
    out[:, 3] = _allele_segment(genome, coff, pos, rl, al, ao, maxshift, True)
    k = np.array([(s + maxshift) // SHIFT_STEP for s in chromatin_shift_list(maxshift)], dtype=np.int32)
    return out.reshape(-1), tab.codes, k


RECEPTIVE_FIELD = 310  # input positions one conv6 row depends on (8,8,pool4,8,8,pool4,8,8 with strides 1,1,4,4,4,16,16,16)
DELTA_MAX_ALLELE = 16  # longest substitution handled by the patch path


def segment_geometry(L: int) -> dict:
    """Row counts of every activation for a segment of L positions (mirror of capi.cu: make_geom)."""
    p2 = (L - 14) // 4
    p4 = (p2 - 14) // 4
    l4b = (p2 - 16) // 4
    return {"p1": (L - 7 + 3) // 4 * 4, "p2": p2, "p3": (p2 - 7 + 3) // 4 * 4, "p4": p4, "l4b": l4b, "p5": p4 - 7,
            "p6": p4 - 14, "valid_a": p4 - 14, "valid_b": l4b - 14}


def delta_eligible(ref_len: np.ndarray, alt_len: np.ndarray) -> np.ndarray:
    """Substitutions (SNVs / MNVs): the alt segment equals the ref segment outside the allele, position for position."""
    ref_len = np.asarray(ref_len, dtype=np.int64)
    alt_len = np.asarray(alt_len, dtype=np.int64)
    return (ref_len == alt_len) & (ref_len >= 1) & (ref_len <= DELTA_MAX_ALLELE)


def sub_descs(d: np.ndarray, start: int, length: int, seg_len: int) -> np.ndarray:
    """Descriptors of the strand-oriented positions [start, start+length) of segments described by ``d``
    (``seg_len`` positions each).  Reverse-complement segments are stored forward, so their range is mirrored."""
    rc = (d["flags"] & 1).astype(bool)
    f0 = np.where(rc, seg_len - start - length, start).astype(np.int64)   # first forward-strand position kept
    a0, a1 = d["len1"].astype(np.int64), d["len1"].astype(np.int64) + d["alen"]
    out = d.copy()
    len1 = np.clip(a0 - f0, 0, length)
    alen = np.clip(np.minimum(a1, f0 + length) - np.maximum(a0, f0), 0, None)
    out["g1"] = d["g1"] + f0
    out["len1"] = len1
    out["alen"] = alen
    out["aoff"] = d["aoff"] + np.maximum(0, f0 - a0)
    out["g2"] = d["g2"] + np.maximum(0, f0 - a1)
    out["len2"] = length - len1 - alen
    return out


def delta_patch_geometry(maxshift: int, max_allele: int = DELTA_MAX_ALLELE) -> tuple[int, int]:
    """(patch_row0, patch_len) for substitutions of up to ``max_allele`` bases: the patch starts 16*patch_row0
    positions into the alt segment and its valid conv6 rows cover every row whose receptive field reaches the allele,
    in both pooling phases (phase B rows sit 8 positions later), on both strands."""
    seg_len = INPUTSIZE + 2 * maxshift
    v0_f = maxshift + 999                       # allele start in the forward segment (_allele_segment, la == ref_len)
    starts = [v0_f, seg_len - v0_f - max_allele, seg_len - v0_f - 1]
    ends = [v0_f + max_allele, seg_len - v0_f]
    row0 = (min(starts) - (RECEPTIVE_FIELD + 7)) // 16           # first row of either phase that can see the allele
    last = (max(ends) - 1) // 16
    rows = last - row0 + 1
    L = 16 * ((16 * (rows - 1) + RECEPTIVE_FIELD + 8 + 15) // 16)
    while min(segment_geometry(L)["valid_a"], segment_geometry(L)["valid_b"]) < rows:
        L += 16
    if row0 < 0 or 16 * row0 + L > seg_len:
        raise ValueError("segment too short for a delta patch")
    return int(row0), int(L)


def chromatin_delta_descs(genome: PackedGenome, chroms, pos, refs: list[str], alts: list[str], maxshift: int = 800):
    """Segment mode for substitutions: per variant the two full ref segments {ref fwd, ref rc} and two short alt
    patches {alt fwd, alt rc}.  Returns (ref_descs [2N], patch_descs [2N], allele codes, k_of_shift, patch_row0,
    patch_len); callers must pre-filter with delta_eligible."""
    n = len(pos)
    rl = np.fromiter((len(r) for r in refs), dtype=np.int64, count=n)
    al = np.fromiter((len(a) for a in alts), dtype=np.int64, count=n)
    if not np.all(delta_eligible(rl, al)):
        raise ValueError("chromatin_delta_descs needs variants pre-filtered with delta_eligible")
    descs, codes, k = chromatin_segment_descs(genome, chroms, pos, refs, alts, maxshift)
    d = descs.reshape(n, 4)
    row0, plen = delta_patch_geometry(maxshift)
    seg_len = INPUTSIZE + 2 * maxshift
    patches = sub_descs(d[:, 2:4].reshape(-1), 16 * row0, plen, seg_len)
    return np.ascontiguousarray(d[:, 0:2]).reshape(-1), patches, codes, k, row0, plen


def tss_segment_descs(genome: PackedGenome, chroms, tss, strand, shifts, windowsize: int = INPUTSIZE):
    """Two segments per gene, order {fwd, rc}, covering the windows of compute_expecto_features.py:105-111
    (``tss + shift*strand - 999 .. + 1000``).  ``shifts`` must be an arithmetic progression with step 200.
    For '-' genes the windows run backwards along the forward segment (flags bit 1)."""
    shifts = np.asarray(shifts, dtype=np.int64)
    S = len(shifts)
    if S > 1 and not np.all(np.diff(shifts) == SHIFT_STEP):
        raise ValueError("segment mode needs shifts in steps of 200")
    n = len(tss)
    tss = np.asarray(tss, dtype=np.int64)
    strand = np.asarray(strand, dtype=np.int64)
    coff = np.fromiter((genome.offsets[c] for c in chroms), dtype=np.int64, count=n)
    seg_len = windowsize + SHIFT_STEP * (S - 1)
    first = np.where(strand > 0, shifts[0], -shifts[-1])       # smallest shift*strand
    start0 = coff + tss + first - int(windowsize / 2 - 1) - 1
    d = np.zeros((n, 2), dtype=DESC_DTYPE)
    d["g1"] = start0[:, None]
    d["len1"] = seg_len
    d["flags"][:, 0] = np.where(strand > 0, 0, 2)
    d["flags"][:, 1] = np.where(strand > 0, 3, 1)
    return d.reshape(-1), np.arange(S, dtype=np.int32), seg_len
