"""Test-only stand-in for natsort (geuvadis_predict_consensus.py:11,58), used to run the UNMODIFIED reference script:
``natsorted`` with the default algorithm — digit runs compare as unsigned integers, the rest as text."""
import re


def natsorted(seq, key=None, reverse=False):
    def k(x):
        s = key(x) if key else x
        return tuple(int(p) if i % 2 else p for i, p in enumerate(re.split(r"(\d+)", s)))
    return sorted(seq, key=k, reverse=reverse)
