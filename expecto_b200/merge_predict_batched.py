"""Last step of the chunked workflow (``chromatin --chunk_i k`` -> ``predict`` per chunk -> one table): drop-in for the
reference's ``merge_predict_batched.py`` with the same three options and the same ``sed.tsv``.

    python -m expecto_b200.merge_predict_batched --batch_dir batches --n_chunks 1000 -o out

What the reference does and this reproduces (file:line into /root/reference/merge_predict_batched.py):
every ``sed.tsv`` one directory below ``--batch_dir``, taken in natural order (``chunk_2`` before ``chunk_10``, :22);
their number must equal ``--n_chunks`` (AssertionError otherwise, :23); the tables are stacked and written tab-separated
WITH pandas' index column, i.e. the per-chunk row numbers, which restart at 0 in every chunk (:25-26).
"""
from __future__ import annotations

import argparse
import os
import sys

from .geuvadis_predict_consensus import natural_key

OPTIONS = (
    ("--batch_dir", dict(dest="batch_dir", type=str)),
    ("--n_chunks", dict(dest="n_chunks", type=int, default=1000, help="Expected number of chunks")),
    ("-o", dict(dest="out_dir", type=str, default="temp_cluster_by_pwm", help="Output directory")),
)


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser(description="Merge output batches of predict.py")
    for flag, kwargs in OPTIONS:
        parser.add_argument(flag, **kwargs)
    return parser


def chunk_tables(batch_dir: str) -> list[str]:
    """``{batch_dir}/<chunk>/sed.tsv`` for every chunk directory that has one, in natural order of the full path."""
    found = []
    with os.scandir(batch_dir) as it:
        for entry in it:
            table = os.path.join(batch_dir, entry.name, "sed.tsv")
            if not entry.name.startswith(".") and entry.is_dir() and os.path.isfile(table):  # glob's `*` skips dot names
                found.append(table)
    return sorted(found, key=natural_key)


def merge(tables: list[str], destination: str) -> int:
    """Stack the per-chunk tables (each keeps its own 0-based index) and write them; returns the row count."""
    import pandas as pd

    stacked = pd.concat([pd.read_csv(path, sep="\t") for path in tables])
    stacked.to_csv(destination, sep="\t")
    return len(stacked)


def run(args) -> str:
    os.makedirs(args.out_dir, exist_ok=True)
    tables = chunk_tables(args.batch_dir)
    assert len(tables) == args.n_chunks, f"Expected {args.n_chunks} chunks but got {len(tables)} sed files"
    destination = os.path.join(args.out_dir, "sed.tsv")
    merge(tables, destination)
    return destination


def main(argv=None) -> int:
    run(build_parser().parse_args(argv))
    return 0


if __name__ == "__main__":
    sys.exit(main())
