"""Drop-in for the reference's ``merge_predict_batched.py``: the last step of the chunked workflow (``chromatin
--chunk_i k`` -> ``predict`` per chunk -> one table), same CLI and the same ``sed.tsv``.

    python -m expecto_b200.merge_predict_batched --batch_dir batches --n_chunks 1000 -o out

Reference behaviour reproduced (file:line into /root/reference/merge_predict_batched.py): ``{batch_dir}/*/sed.tsv`` in
natural order (:22), the count checked against ``--n_chunks`` (:23), the tables concatenated and written with pandas'
default index column — the per-chunk row numbers, restarting at 0 in every chunk (:25-26).
"""
from __future__ import annotations

import argparse
import glob
import os
import sys

from .geuvadis_predict_consensus import natural_key


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Merge output batches of predict.py")
    p.add_argument("--batch_dir", dest="batch_dir", type=str)
    p.add_argument("--n_chunks", dest="n_chunks", type=int, default=1000, help="Expected number of chunks")
    p.add_argument("-o", dest="out_dir", type=str, default="temp_cluster_by_pwm", help="Output directory")
    return p


def run(args) -> str:
    import pandas as pd

    os.makedirs(args.out_dir, exist_ok=True)
    sed_files = sorted(glob.glob(f"{args.batch_dir}/*/sed.tsv"), key=natural_key)
    assert len(sed_files) == args.n_chunks, f"Expected {args.n_chunks} chunks but got {len(sed_files)} sed files"
    sed_df = pd.concat([pd.read_csv(f, sep="\t") for f in sed_files])
    out = f"{args.out_dir}/sed.tsv"
    sed_df.to_csv(out, sep="\t")
    return out


def main(argv=None) -> int:
    run(build_parser().parse_args(argv))
    return 0


if __name__ == "__main__":
    sys.exit(main())
