"""Deterministic synthetic resources (no network: hg19.fa / deepsea.beluga.pth are not in the reference
repo — download_resources.sh:2, .gitignore:5).  Used by tests, the oracle harness and bench.py.

* calibrated random Beluga weights in the reference's state_dict layout (SURVEY §3.2, §7 hard part 1):
  variance-preserving init, last layer scaled/biased so probabilities look like Beluga's
  (mostly << 0.5, SNV diffs ~1e-3..1e-1), optionally rounded ONCE to the kernel's operand type so the
  fp32 oracle and the 16-bit kernels see identical weight values.
* hg19-shaped synthetic genome / FASTA and VCF generators.
"""
from __future__ import annotations

import numpy as np

STATE_KEYS = [
    ("model.0.0", (320, 4, 1, 8)), ("model.0.2", (320, 320, 1, 8)), ("model.0.6", (480, 320, 1, 8)),
    ("model.0.8", (480, 480, 1, 8)), ("model.0.12", (640, 480, 1, 8)), ("model.0.14", (640, 640, 1, 8)),
    ("model.1.2.1", (2003, 67840)), ("model.1.4.1", (2002, 2003)),
]

HG19_CONTIGS = [
    ("chr1", 249250621), ("chr2", 243199373), ("chr3", 198022430), ("chr4", 191154276), ("chr5", 180915260),
    ("chr6", 171115067), ("chr7", 159138663), ("chr8", 146364022), ("chr9", 141213431), ("chr10", 135534747),
    ("chr11", 135006516), ("chr12", 133851895), ("chr13", 115169878), ("chr14", 107349540), ("chr15", 102531392),
    ("chr16", 90354753), ("chr17", 81195210), ("chr18", 78077248), ("chr19", 59128983), ("chr20", 63025520),
    ("chr21", 48129895), ("chr22", 51304566), ("chrX", 155270560), ("chrY", 59373566),
]


def round_to(a: np.ndarray, dtype: str | None) -> np.ndarray:
    """Round fp32 values to the nearest fp16 / bf16 value (kept as fp32)."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    if dtype in (None, "f32"):
        return a
    if dtype == "f16":
        return a.astype(np.float16).astype(np.float32)
    if dtype == "bf16":
        u = a.view(np.uint32).astype(np.uint64)
        u = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16  # round to nearest even
        return u.astype(np.uint32).view(np.float32)
    raise ValueError(dtype)


def make_beluga_weights(seed: int = 0, round_dtype: str | None = "f16") -> dict[str, np.ndarray]:
    """Calibrated random weights keyed like ``Beluga().state_dict()`` (fp32 numpy arrays)."""
    rng = np.random.default_rng(seed)
    sd: dict[str, np.ndarray] = {}
    for key, shape in STATE_KEYS:
        if len(shape) == 4:
            fan_in = shape[1] * shape[3]
            w = rng.standard_normal(shape, dtype=np.float32) * np.float32(np.sqrt(2.0 / fan_in))
            b = rng.standard_normal(shape[0], dtype=np.float32) * np.float32(0.05)
        elif shape[0] == 2003:
            w = rng.standard_normal(shape, dtype=np.float32) * np.float32(np.sqrt(2.0 / shape[1]))
            b = rng.standard_normal(shape[0], dtype=np.float32) * np.float32(0.05)
        else:
            # last layer: logits std ~2 around -3 => probabilities mostly well below 0.5
            w = rng.standard_normal(shape, dtype=np.float32) * np.float32(2.0 / np.sqrt(shape[1] * 0.6))
            b = rng.standard_normal(shape[0], dtype=np.float32) * np.float32(0.5) - np.float32(3.0)
        sd[key + ".weight"] = round_to(w, round_dtype)
        sd[key + ".bias"] = np.ascontiguousarray(b, dtype=np.float32)
    return sd


def random_sequence(rng: np.random.Generator, n: int, p_lower: float = 0.038, p_n: float = 0.002) -> str:
    """Random DNA with mixed case and sparse N, as in the survey probe."""
    alphabet = np.frombuffer(b"ACGTacgtN", dtype=np.uint8)
    pu = (1.0 - p_lower - p_n) / 4
    pl = p_lower / 4
    probs = [pu] * 4 + [pl] * 4 + [p_n]
    idx = rng.choice(9, size=n, p=probs)
    return alphabet[idx].tobytes().decode("ascii")


def make_genome(contigs: list[tuple[str, int]], seed: int = 1) -> dict[str, str]:
    rng = np.random.default_rng(seed)
    return {name: random_sequence(rng, length) for name, length in contigs}


def write_fasta(path: str, genome: dict[str, str], width: int = 60) -> None:
    with open(path, "w") as f:
        for name, seq in genome.items():
            f.write(f">{name}\n")
            for i in range(0, len(seq), width):
                f.write(seq[i:i + width] + "\n")


def make_vcf_rows(genome: dict[str, str], n: int, seed: int = 3, margin: int = 3000, frac_mismatch: float = 0.05,
                  frac_indel: float = 0.05, max_indel: int = 10, odd_names: bool = True) -> list[list]:
    """Synthetic VCF rows: 90 % SNVs with ref from the genome, some ref-mismatch SNVs,
    some indels; optionally un-prefixed chromosome names and a non-CHRS contig to exercise the filter."""
    rng = np.random.default_rng(seed)
    names = list(genome.keys())
    lens = np.array([len(genome[k]) for k in names], dtype=np.float64)
    usable = [k for k in names if len(genome[k]) > 2 * margin + 10]
    w = np.array([len(genome[k]) for k in usable], dtype=np.float64)
    rows = []
    bases = "ACGT"
    for i in range(n):
        chrom = usable[rng.choice(len(usable), p=w / w.sum())]
        seq = genome[chrom]
        pos = int(rng.integers(margin, len(seq) - margin))  # 1-based
        u = rng.random()
        gref = seq[pos - 1].upper()
        if gref not in bases:
            gref = "A"
        if u < frac_indel:
            d = int(rng.integers(1, max_indel + 1))
            if rng.random() < 0.5:  # deletion
                ref = "".join(c.upper() if c.upper() in bases else "A" for c in seq[pos - 1:pos + d])
                alt = ref[0]
            else:  # insertion
                ref = gref
                alt = gref + "".join(bases[j] for j in rng.integers(0, 4, size=d))
        elif u < frac_indel + frac_mismatch:
            ref = bases[(bases.index(gref) + 1) % 4]  # deliberately disagrees with the genome
            alt = bases[(bases.index(gref) + 2) % 4]
        else:
            ref = gref
            alt = bases[(bases.index(gref) + int(rng.integers(1, 4))) % 4]
        name = chrom
        if odd_names and i % 7 == 3:
            name = chrom.replace("chr", "")  # chromatin.py:240 re-adds the prefix
        rows.append([name, pos, f"rs{i}", ref, alt])
    if odd_names and n >= 4:
        rows[n // 2][0] = "chrUn_gl000220"  # filtered out by chromatin.py:241
    return rows


def write_vcf(path: str, rows: list[list], header: bool = True) -> None:
    with open(path, "w") as f:
        if header:
            f.write("##fileformat=VCFv4.1\n#CHROM\tPOS\tID\tREF\tALT\n")
        for r in rows:
            f.write("\t".join(str(x) for x in r) + "\n")


def random_onehot(n: int, seed: int = 2) -> np.ndarray:
    """uint8 [n,4,2000] one-hot of uniform random bases."""
    rng = np.random.default_rng(seed)
    codes = rng.integers(0, 4, size=(n, 2000))
    x = np.zeros((n, 4, 2000), dtype=np.uint8)
    np.put_along_axis(x, codes[:, None, :], 1, axis=1)
    return x


HG19_CONTIGS = [("chr1", 249250621), ("chr2", 243199373), ("chr3", 198022430), ("chr4", 191154276),
                ("chr5", 180915260), ("chr6", 171115067), ("chr7", 159138663), ("chr8", 146364022),
                ("chr9", 141213431), ("chr10", 135534747), ("chr11", 135006516), ("chr12", 133851895),
                ("chr13", 115169878), ("chr14", 107349540), ("chr15", 102531392), ("chr16", 90354753),
                ("chr17", 81195210), ("chr18", 78077248), ("chr19", 59128983), ("chr20", 63025520),
                ("chr21", 48129895), ("chr22", 51304566), ("chrX", 155270560), ("chrY", 59373566)]


def make_packed_genome(contigs=HG19_CONTIGS, seed: int = 3, n_frac: float = 0.002):
    """hg19-shaped synthetic genome built directly in the packed form (3.1 Gb would be a 3 GB
    Python string otherwise): uniformly random bases, N in runs of 64 bases covering ``n_frac`` of every contig."""
    from .genome import PAD, PackedGenome

    rng = np.random.default_rng(seed)
    g = PackedGenome()
    total = PAD
    for name, ln in contigs:
        total = (total + 31) // 32 * 32
        g.names.append(name)
        g.offsets[name] = total
        g.lengths[name] = int(ln)
        total += int(ln) + PAD
    total = (total + 31) // 32 * 32
    g.n_bases = total
    g.bases2 = rng.integers(0, 1 << 32, size=total // 16, dtype=np.uint32)
    g.nmask = np.full(total // 32, 0xFFFFFFFF, dtype=np.uint32)   # padding between contigs is N
    for name, ln in contigs:
        off = g.offsets[name]
        w0, w1 = off // 32, (off + ln) // 32
        g.nmask[w0:w1] = 0
        tail = (off + ln) % 32
        if tail:
            g.nmask[w1] = np.uint32((0xFFFFFFFF << tail) & 0xFFFFFFFF)
        n_runs = int(ln * n_frac / 64)
        if n_runs:
            at = rng.integers(w0, max(w0 + 1, w1 - 2), size=n_runs)
            g.nmask[at] = 0xFFFFFFFF
            g.nmask[at + 1] = 0xFFFFFFFF
    return g


def make_vcf_rows_packed(pg, n: int, seed: int = 3, margin: int = 3000, frac_mismatch: float = 0.05,
                         frac_indel: float = 0.05, max_indel: int = 10):
    """make_vcf_rows for a packed genome, vectorised over variants (1 M rows in seconds): returns
    (chroms list, pos int64 array, refs list, alts list) sorted by contig order and position."""
    rng = np.random.default_rng(seed)
    names = [k for k in pg.names if pg.lengths[k] > 2 * margin + 10]
    w = np.array([pg.lengths[k] for k in names], dtype=np.float64)
    ci = np.sort(rng.choice(len(names), size=n, p=w / w.sum()))
    lens = np.array([pg.lengths[k] for k in names], dtype=np.int64)[ci]
    pos = (margin + rng.random(n) * (lens - 2 * margin)).astype(np.int64)
    order = np.lexsort((pos, ci))
    ci, pos = ci[order], pos[order]
    offs = np.array([pg.offsets[k] for k in names], dtype=np.int64)[ci]
    gidx = offs + pos - 1
    span = max_indel + 1
    j = gidx[:, None] + np.arange(span, dtype=np.int64)[None, :]
    two = (pg.bases2[j >> 4] >> ((j & 15) * 2).astype(np.uint32)) & 3
    msk = (pg.nmask[j >> 5] >> (j & 31).astype(np.uint32)) & 1
    codes = np.where(msk == 1, 0, two).astype(np.int64)        # N -> "A", as make_vcf_rows does
    letters = np.frombuffer(b"AGCT", dtype=np.uint8)             # channel order of the packed codes
    u = rng.random(n)
    kind = np.where(u < frac_indel, np.where(rng.random(n) < 0.5, 1, 2), np.where(u < frac_indel + frac_mismatch, 3, 0))
    d = rng.integers(1, max_indel + 1, size=n)
    step = rng.integers(1, 4, size=n)
    ins = rng.integers(0, 4, size=(n, max_indel))
    chroms, refs, alts = [names[i] for i in ci], [], []
    txt = letters[codes]                                          # [n, span] ASCII
    for i in range(n):
        k = kind[i]
        g0 = int(codes[i, 0])
        if k == 1:      # deletion
            r = txt[i, :d[i] + 1].tobytes().decode()
            refs.append(r); alts.append(r[0])
        elif k == 2:    # insertion
            r = chr(letters[g0])
            refs.append(r); alts.append(r + letters[ins[i, :d[i]]].tobytes().decode())
        elif k == 3:    # ref allele disagrees with the genome
            refs.append(chr(letters[(g0 + 1) % 4])); alts.append(chr(letters[(g0 + 2) % 4]))
        else:
            refs.append(chr(letters[g0])); alts.append(chr(letters[(g0 + int(step[i])) % 4]))
    return chroms, pos, refs, alts
