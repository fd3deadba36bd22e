"""Test-only helpers."""
import numpy as np


def decode_descs(genome, descs, alleles, length=2000):
    """Host emulation of kernel 1 (kernels_misc.cuh: window_code / encode_onehot_kernel): uint8 [n,4,length]."""
    n = len(descs)
    out = np.zeros((n, 4, length), dtype=np.uint8)
    for i, d in enumerate(descs):
        parts = [genome.codes(int(d["g1"]), int(d["len1"])),
                 np.asarray(alleles[int(d["aoff"]):int(d["aoff"]) + int(d["alen"])], dtype=np.uint8),
                 genome.codes(int(d["g2"]), int(d["len2"]))]
        codes = np.concatenate(parts)[:length]
        codes = np.concatenate([codes, np.full(length - codes.size, 4, dtype=np.uint8)])
        if int(d["flags"]) & 1:
            codes = codes[::-1]
            codes = np.where(codes < 4, 3 - codes, 4)
        ok = codes < 4
        out[i, codes[ok], np.flatnonzero(ok)] = 1
    return out


def pearson(a, b):
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    a = a - a.mean()
    b = b - b.mean()
    return float((a @ b) / np.sqrt((a @ a) * (b @ b)))
