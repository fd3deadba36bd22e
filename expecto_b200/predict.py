"""Drop-in for the reference's ``predict.py`` (same CLI, same ``sed*.tsv`` out), transform + scoring on a B200.

    python -m expecto_b200.predict --model_save_file M --belugaFeatures resources/deepsea_beluga_2002_features.tsv \\
        --coorFile_chromatin out/snps_hg19.vcf --geneFile snps.closestgene \\
        --snpEffectFilePattern out/snps.shift_SHIFT.diff.h5 -o out_dir

Reference behaviour reproduced (file:line into /root/reference/predict.py):
  * arguments and defaults                                                               :15-58
  * per-shift HDF5 read of ``diff``/``ref``/``alt`` and the (fwd + rc) / 2.0 strand average :173-194
  * duplicate handling, variant <-> gene expansion with ``get_num_repeats`` / ``np.repeat`` :197-235
  * ``dist = -gene[:, -1]`` (or ``--fixeddist``), gene name, strand                          :238-246
  * ``compute_effects``: decay weights, 10 x 2002 features, keep-mask column subset, gblinear
    scores for ref and alt features; ``SED = ALT - REF``                                     :70-162,264
  * ``sed.tsv``, ``sed_sorted_by_magnitude.tsv``, ``sed_sorted_by_proportion.tsv``            :255-280
By design the [n,20020] float64 feature matrices are never built: kernel ``decay_score_variant`` forms each
variant's features in registers and contracts them with the model weights; columns dropped by
the keep mask are zero-weighted, which is the same sum.  ``--modelList`` (upstream ExPecto's interface, README.md:25)
scores every model of a list in the same launch and appends one column per model.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

from .chromatin import shift_list


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Process some integers.")
    p.add_argument("--model_save_file", action="store", dest="model_save_file")
    p.add_argument("--belugaFeatures", action="store", dest="belugaFeatures")
    p.add_argument("--coorFile_chromatin", action="store", dest="coorFile_chromatin")
    p.add_argument("--geneFile", action="store", dest="geneFile")
    p.add_argument("--snpEffectFilePattern", action="store", dest="snpEffectFilePattern")
    p.add_argument("--rsat_clusters_tab", action="store", dest="rsat_clusters_tab")
    p.add_argument("--nfeatures", action="store", dest="nfeatures", type=int, default=2002)
    p.add_argument("--fixeddist", action="store", dest="fixeddist", default=0, type=int)
    p.add_argument("--maxshift", action="store", dest="maxshift", type=int, default=800)
    p.add_argument("--batchSize", action="store", dest="batchSize", type=int, default=500)
    p.add_argument("--splitIndex", action="store", dest="splitIndex", type=int, default=0)
    p.add_argument("--splitFold", action="store", dest="splitFold", type=int, default=10)
    p.add_argument("--threads", action="store", dest="threads", type=int, default=16)
    p.add_argument("--no_tf_features", action="store_true", dest="no_tf_features", default=False)
    p.add_argument("--no_dnase_features", action="store_true", dest="no_dnase_features", default=False)
    p.add_argument("--no_histone_features", action="store_true", dest="no_histone_features", default=False)
    p.add_argument("--intersect_with_lambert", action="store_true", dest="intersect_with_lambert", default=False)
    p.add_argument("--no_pol2", action="store_true", dest="no_pol2", default=False)
    p.add_argument("-o", action="store", dest="out_dir")
    # extensions
    p.add_argument("--modelList", default=None, help="file with one model path per line (upstream interface)")
    p.add_argument("--lambert_csv", default="./resources/Lambert-hgnc-symbol-check.csv",
                   help="--intersect_with_lambert: table with an 'Approved symbol' column (cluster_utils.py:5)")
    p.add_argument("--hgnc_mapping_csv", default="./resources/beluga_hgnc_mapping.csv",
                   help="--intersect_with_lambert: Beluga assay -> HGNC symbol table (cluster_utils.py:6)")
    p.add_argument("--device", type=int, default=0)
    p.add_argument("--vcf", default=None,
                   help="compute the chromatin effects of this VCF in-process (segment + delta mode) instead of reading "
                        "--snpEffectFilePattern files: chromatin.py and predict.py in one pass, no .diff.h5 in between; "
                        "snps_hg19.vcf is written to -o and used as --coorFile_chromatin unless that is given")
    p.add_argument("--genome", default="./resources/hg19.fa", help="--vcf: FASTA (chromatin.py:44)")
    p.add_argument("--weights", default="./resources/deepsea.beluga.pth", help="--vcf: Beluga state_dict (chromatin.py:103)")
    p.add_argument("--dtype", default="f16", choices=["f16", "bf16", "f16x3"], help="--vcf: tensor-core operand type")
    p.add_argument("--vcf_block", type=int, default=131072, help="--vcf: variants resident in HBM at a time")
    return p


def lambert_mask(assays, lambert_csv: str, mapping_csv: str) -> np.ndarray:
    """The --intersect_with_lambert clause of cluster_utils.get_keep_mask (cluster_utils.py:26-42): every Beluga assay
    name found in the index of the HGNC mapping table (rows without an approved symbol dropped) is replaced by that
    row's approved symbol in upper case — when the name has several rows, by the first one whose ``Match type`` is
    "Approved symbol" — and a feature is kept when the resulting name is one of the Lambert table's approved symbols."""
    import pandas as pd

    lambert = pd.read_csv(lambert_csv, index_col=0)
    mapping = pd.read_csv(mapping_csv, index_col=0).dropna(subset=["Approved symbol"])
    symbol_of, n_rows = {}, {}
    for name, mtype, sym in zip(mapping.index, mapping["Match type"], mapping["Approved symbol"]):
        n_rows[name] = n_rows.get(name, 0) + 1
        if name not in symbol_of or (symbol_of[name][0] != "Approved symbol" and mtype == "Approved symbol"):
            symbol_of[name] = (mtype, str(sym).upper())
    for name, k in n_rows.items():
        if k > 1 and symbol_of[name][0] != "Approved symbol":
            # the reference indexes .iloc[0] of an empty selection here
            raise IndexError(f"{mapping_csv}: {name!r} has {k} rows, none with Match type 'Approved symbol'")
    approved = set(lambert["Approved symbol"].values)
    out = np.zeros(len(assays), dtype=bool)
    for i, a in enumerate(assays):
        a = symbol_of[a][1] if a in symbol_of else a
        out[i] = (a in approved) and not (isinstance(a, float) and np.isnan(a))
    return out


def keep_mask(features_tsv: str, args) -> np.ndarray:
    """cluster_utils.get_keep_mask (cluster_utils.py:8-49)."""
    import pandas as pd

    df = pd.read_csv(features_tsv, sep="\t", index_col=0)
    mask = np.ones(df.shape[0], dtype=bool)
    if args.no_tf_features:
        print("not including TF features")
        mask &= (df["Assay type"] != "TF").values
    if args.no_dnase_features:
        print("not including DNase features")
        mask &= (df["Assay type"] != "DNase").values
    if args.no_histone_features:
        print("not including histone features")
        mask &= (df["Assay type"] != "Histone").values
    if args.intersect_with_lambert:
        print("intersecting with Lambert data")
        mask &= lambert_mask(list(df["Assay"].values),
                             getattr(args, "lambert_csv", "./resources/Lambert-hgnc-symbol-check.csv"),
                             getattr(args, "hgnc_mapping_csv", "./resources/beluga_hgnc_mapping.csv"))
    if args.no_pol2:
        print("taking out Pol2*")
        mask &= ~df["Assay"].astype(str).str.startswith("Pol").values
    print(f"Number of features included in model: {int(mask.sum())}")
    return mask


def num_repeats(gene_df) -> list[int]:
    """get_num_repeats, predict.py:202-213: run lengths of consecutive rows sharing columns 0-4."""
    if len(gene_df) == 0:
        return [0]
    cols = [gene_df.iloc[:, i].astype(str) for i in range(5)]
    keys = cols[0]
    for c in cols[1:]:
        keys = keys + ":" + c
    keys = keys.values
    change = np.flatnonzero(keys[1:] != keys[:-1]) + 1
    edges = np.concatenate([[0], change, [len(keys)]])
    return [int(x) for x in np.diff(edges)]


def load_strand_averaged(pattern: str, maxshift: int, keys=("diff", "ref", "alt"), device=None, n_threads: int = 4,
                         slab_rows: int = 8192):
    """predict.py:173-194 -> dict of float32 CUDA tensors [S, N, 2002]: every per-shift dataset is read by
    ``n_threads`` positional readers straight into pinned slabs, copied to HBM asynchronously and its forward /
    reverse-complement halves averaged there (``strand_average``)."""
    from concurrent.futures import ThreadPoolExecutor

    import torch

    from . import hdf5
    from .transform import padded_rows, strand_average

    shifts = shift_list(maxshift)
    out = {}
    ring = [[torch.empty((slab_rows, 2002), dtype=torch.float32, pin_memory=True), None] for _ in range(2 * n_threads)]
    slot = 0
    with ThreadPoolExecutor(max_workers=n_threads) as pool:
        for si, s in enumerate(shifts):
            with hdf5.File(pattern.replace("SHIFT", str(s))) as f:
                for k in keys:
                    d = f[k]
                    n = int(d.shape[0] / 2)
                    if k not in out:
                        out[k] = padded_rows((len(shifts), n), device)   # rows on 128-byte boundaries for the scoring kernel
                    x = torch.empty((2 * n, 2002), dtype=torch.float32, device=device)
                    jobs = []
                    for r0 in range(0, 2 * n, slab_rows):
                        r1 = min(2 * n, r0 + slab_rows)
                        buf, ev = ring[slot]
                        if ev is not None:
                            ev.synchronize()     # the H2D copy that last used this slab has finished
                        jobs.append((pool.submit(d.read_rows_into, buf.numpy(), r0, r1), slot, r0, r1))
                        slot = (slot + 1) % len(ring)
                        if len(jobs) >= n_threads:           # keep n_threads reads in flight, copy in order
                            fut, sl, a0, a1 = jobs.pop(0)
                            fut.result()
                            x[a0:a1].copy_(ring[sl][0][:a1 - a0], non_blocking=True)
                            ring[sl][1] = torch.cuda.Event()
                            ring[sl][1].record(torch.cuda.current_stream(device))
                    for fut, sl, a0, a1 in jobs:
                        fut.result()
                        x[a0:a1].copy_(ring[sl][0][:a1 - a0], non_blocking=True)
                        ring[sl][1] = torch.cuda.Event()
                        ring[sl][1].record(torch.cuda.current_stream(device))
                    out[k][si].copy_(strand_average(x))
                    del x
    torch.cuda.synchronize(device)
    return out


def expand_weights(models, mask: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    """Model weights over the kept columns (10 x n_marks, predict.py:139-147) -> zero-filled [m, 10*2002]."""
    n_marks = int(mask.sum())
    W = np.zeros((len(models), 10, mask.size), dtype=np.float32)
    for i, m in enumerate(models):
        if m.weights.size != 10 * n_marks:
            raise ValueError(f"model {i} has {m.weights.size} weights, expected 10 x {n_marks}")
        W[i][:, mask] = m.weights.reshape(10, n_marks)
    return W.reshape(len(models), -1), np.array([m.offset for m in models], dtype=np.float32)


def _table_text(df) -> str:
    """``df.to_csv(sep='\\t', index=False)``.  The reference's columns (everything up to ``SED``) are formatted by
    pandas itself, so they are byte-identical to predict.py's files.  The extra per-model ``SED_<model>`` columns of
    ``--modelList`` (an extension: 217 float columns per row for the 218-model list) go through Arrow's CSV writer —
    shortest round-trip float32 text as well, ten times faster, different exponent style (``1e-5``/``0.00001``)."""
    cols = list(df.columns)
    n_ref = cols.index("SED") + 1 if "SED" in cols else len(cols)
    extra = cols[n_ref:]
    if not extra or len(df) == 0 or not all(str(df[c].dtype).startswith("float") for c in extra):
        return df.to_csv(None, header=True, sep="\t", index=False)
    try:
        import io

        import pyarrow as pa
        import pyarrow.csv as pc
    except ImportError:
        return df.to_csv(None, header=True, sep="\t", index=False)
    base = df[cols[:n_ref]].to_csv(None, header=True, sep="\t", index=False).split("\n")
    values = [df[c].to_numpy() for c in extra]

    def block(ab):  # Arrow releases the GIL: row blocks are formatted on several threads
        a, b = ab
        tb = pa.Table.from_arrays([pa.array(v[a:b], from_pandas=True) for v in values], names=[str(c) for c in extra])
        buf = io.BytesIO()
        pc.write_csv(tb, buf, pc.WriteOptions(include_header=False, delimiter="\t"))
        return buf.getvalue()

    n_threads = max(1, min(8, os.cpu_count() or 1))
    edges = np.linspace(0, len(df), 2 * n_threads + 1).astype(np.int64)
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(n_threads) as ex:
        parts = list(ex.map(block, [(int(a), int(b)) for a, b in zip(edges[:-1], edges[1:]) if b > a]))
    more = b"".join(parts).decode("ascii").split("\n")
    if len(base) != len(df) + 2 or len(more) != len(df) + 1:
        return df.to_csv(None, header=True, sep="\t", index=False)
    head = base[0] + "\t" + "\t".join(str(c) for c in extra)
    return "\n".join([head] + [b + "\t" + m for b, m in zip(base[1:-1], more[:-1])]) + "\n"


def write_sed_tables(df, out_dir: str) -> None:
    """sed.tsv, sed_sorted_by_magnitude.tsv, sed_sorted_by_proportion.tsv (predict.py:255-280).  The three files hold
    the same rows, so every row is formatted once (with a --modelList of 218 models the float formatting is the
    longest phase of the whole script) and the sorted files re-emit those lines in the order pandas' own
    sort_values gives, with the extra column appended; any surprise in the text falls back to three to_csv calls."""
    text = _table_text(df)
    with open(f"{out_dir}/sed.tsv", "w", newline="") as f:
        f.write(text)
    lines = text.split("\n")
    header, body = lines[0], lines[1:-1]
    extras = (("SED_MAGNITUDES", np.abs(df["SED"]), "magnitude"),
              ("SED_PROPORTION", np.abs(df["SED"] / ((df["REF"] + df["ALT"]) / 2)), "proportion"))
    for col, values, name in extras:
        path = f"{out_dir}/sed_sorted_by_{name}.tsv"
        if len(body) != len(df) or lines[-1] != "":
            s = df.copy()
            s[col] = values
            s.sort_values(by=col, axis=0, ascending=False).to_csv(path, header=True, sep="\t", index=False)
            continue
        values = values.reset_index(drop=True)
        order = values.sort_values(ascending=False).index.to_numpy()
        # pandas' own cell text for this dtype (a lone empty cell comes back quoted)
        cells = ["" if c == '""' else c for c in values.to_csv(None, header=False, index=False).split("\n")]
        if len(cells) != len(df) + 1:
            raise RuntimeError("unexpected text for the sort column")
        with open(path, "w", newline="") as f:
            f.write(header + "\t" + col + "\n")
            f.write("".join(body[i] + "\t" + cells[i] + "\n" for i in order))


def _chromatin_blocks(args, dev, engine=None, genome=None):
    """--vcf: yield (v0, v1, {"ref","alt"} -> [S, v1-v0, 2002] strand-averaged device tensors) for blocks of the
    filtered variants, straight from the CUDA forward (chromatin.py:211-281 + predict.py:173-194 without the files)."""
    from . import chromatin
    from .genome import PackedGenome

    if genome is None:
        genome = PackedGenome.from_fasta(args.genome, device=engine.device if engine is not None else dev)
    if engine is None:
        from .engine import BelugaEngine
        engine = BelugaEngine(dev.index or 0, dtype=args.dtype, capacity=4096)
        engine.load_weights_file(args.weights)
    if engine._genome is not genome:
        engine.set_genome(genome)
    vcf = chromatin.read_vcf(args.vcf)
    chromatin.write_snps_vcf(vcf, args.out_dir)
    chroms, pos, refs, alts = chromatin.variant_columns(chromatin.normalise_and_filter(vcf))
    missing = sorted({c for c in chroms if c not in genome.offsets})
    if missing:
        raise KeyError(f"contig(s) {missing} are not in the genome")
    n = len(pos)
    # The consumer sends, with each request for a block, the order it wants that block's variants in (a permutation of
    # range(v1 - v0), or None): row i of the block's tensors is then variant v0 + perm[i].  run() asks for them grouped
    # by side of the TSS (transform.side_order) when every variant has exactly one gene row, so that the scoring
    # kernel's tiles are one-sided without any gather.
    perm = yield n
    for v0 in range(0, n, args.vcf_block):
        v1 = min(n, v0 + args.vcf_block)
        ii = np.arange(v0, v1) if perm is None else v0 + np.asarray(perm, dtype=np.int64)
        if len(ii) != v1 - v0 or not np.array_equal(np.sort(ii), np.arange(v0, v1)):
            raise ValueError("block order must be a permutation of the block's variants")
        data = chromatin.predict_all_shifts_averaged_device(engine, genome, [chroms[i] for i in ii], pos[ii],
                                                            [refs[i] for i in ii], [alts[i] for i in ii],
                                                            int(args.maxshift), padded=True)
        data.pop("diff")
        perm = yield v0, v1, data
        del data


def _file_blocks(n_pred: int, data):
    """The one block of the file path, behind the same protocol as _chromatin_blocks (the requested order is ignored:
    the rows are already in HBM in file order)."""
    yield n_pred
    yield 0, n_pred, data


def run(args, engine=None, genome=None) -> dict:
    import pandas as pd
    import torch

    from .gblinear import load_model
    from .transform import ScoreModels, decay_score_variant, select_variants, side_order

    if not torch.cuda.is_available():
        raise RuntimeError("predict needs a CUDA device (B200); there is no CPU fallback")
    import time
    timing, t_last = {}, time.perf_counter()

    def lap(name):
        nonlocal t_last
        now = time.perf_counter()
        timing[name] = timing.get(name, 0.0) + now - t_last
        t_last = now

    os.makedirs(args.out_dir, exist_ok=True)
    mask = keep_mask(args.belugaFeatures, args)
    paths = [args.model_save_file.strip()] if args.model_save_file else []
    if args.modelList:
        paths += [l.strip() for l in open(args.modelList) if l.strip() and not l.startswith("#")]
    if not paths:
        raise ValueError("give --model_save_file or --modelList")
    W, offsets = expand_weights([load_model(p) for p in paths], mask)
    lap("models_s")

    dev = torch.device("cuda", args.device)
    if args.vcf:
        blocks = _chromatin_blocks(args, dev, engine, genome)
        n_pred = next(blocks)
        if not args.coorFile_chromatin:
            args.coorFile_chromatin = f"{args.out_dir}/snps_hg19.vcf"
    else:
        data = load_strand_averaged(args.snpEffectFilePattern, int(args.maxshift), keys=("ref", "alt"), device=dev)
        torch.cuda.synchronize(dev)
        n_pred = int(data["ref"].shape[1])
        blocks = _file_blocks(n_pred, data)
        next(blocks)
        del data
    lap("load_h5_s")
    coor = pd.read_csv(args.coorFile_chromatin, sep="\t", header=None, comment="#")
    gene = pd.read_csv(args.geneFile, sep="\t", header=None, comment="#")
    gene = gene.drop_duplicates(keep="first")
    coor_mask = ~coor.duplicated(keep="first").values
    coor = coor[coor_mask]
    repeats = num_repeats(gene)
    idx = np.repeat(np.flatnonzero(coor_mask), repeats)  # row of the prediction arrays for every gene row
    coor = pd.DataFrame(np.repeat(coor.values, repeats, axis=0), columns=coor.columns)
    if len(idx) != gene.shape[0]:
        raise ValueError("Gene association file does not match the vcf file.")
    if args.fixeddist == 0:
        dist = -np.asarray(gene.iloc[:, -1], dtype=np.float64)
    else:
        dist = np.full(gene.shape[0], float(args.fixeddist))
    genename, strand = np.asarray(gene.iloc[:, -2]), np.asarray(gene.iloc[:, -3])

    lap("tables_s")
    if len(idx) and (int(idx.max()) >= n_pred or int(idx.min()) < 0):
        raise IndexError(f"gene rows point at variant {int(idx.max())} but only {n_pred} were predicted "
                         "(rows filtered by chromatin.py:241 shift the two tables apart, predict.py:217)")
    shifts = shift_list(int(args.maxshift))
    models = ScoreModels(W, offsets, dev)   # staged (and laid out for the tensor-core kernel) once
    scores = {k: np.empty((len(idx), len(paths)), dtype=np.float32) for k in ("ref", "alt")}
    block_rows = int(args.vcf_block) if args.vcf else max(1, n_pred)
    for b0 in range(0, n_pred, block_rows):  # one block when reading files; --vcf computes --vcf_block variants at a time
        b1 = min(n_pred, b0 + block_rows)
        rows = np.flatnonzero((idx >= b0) & (idx < b1))   # idx is non-decreasing: a contiguous run of gene rows
        one_to_one = len(rows) == b1 - b0 and bool(np.array_equal(idx[rows] - b0, np.arange(b1 - b0)))
        want = None
        if args.vcf and one_to_one and rows.size:
            # every variant of the block has exactly one gene row: have the block computed in side order (no gather)
            want = side_order(dist[rows], strand[rows], shifts)
            rows = rows[want]
        v0, v1, data = blocks.send(want)
        if (v0, v1) != (b0, b1):
            raise RuntimeError("block protocol out of step")
        if rows.size:
            local = idx[rows] - v0 if want is None else np.arange(v1 - v0)
            identity = len(local) == v1 - v0 and bool(np.array_equal(local, np.arange(v1 - v0)))
            if not identity:
                # the rows are gathered anyway: gather them grouped by side of the TSS, which lets the scoring kernel
                # skip the operand columns that are zero for whole tiles (transform.side_order; same scores bit for bit)
                order = side_order(dist[rows], strand[rows], shifts)
                rows, local = rows[order], local[order]
            sel = None if identity else torch.from_numpy(local).to(dev)
            for key in ("ref", "alt"):
                x = data[key] if identity else select_variants(data[key], sel)
                scores[key][rows] = decay_score_variant(x, dist[rows], strand[rows], models, shifts=shifts).cpu().numpy()
        del data
        lap("chromatin_and_score_s" if args.vcf else "score_s")
    ref, alt = scores["ref"], scores["alt"]
    lap("score_s")

    df = coor
    # predict.py:240: with --fixeddist the column holds the int scalar itself ("500", not "500.0")
    df["dist"] = int(args.fixeddist) if args.fixeddist else -np.asarray(gene.iloc[:, -1])
    df["gene"] = genename
    df["strand"] = strand
    # the reference keeps xgboost's float32 predictions in float64 arrays (compute_effects' np.zeros), so REF / ALT
    # print as widened float32 values and SED = alt - ref is a float64 difference (predict.py:158-161,264)
    ref64, alt64 = ref[:, 0].astype(np.float64), alt[:, 0].astype(np.float64)
    cols = {"REF": ref64, "ALT": alt64, "SED": alt64 - ref64}
    for i, p in enumerate(paths[1:], start=1):  # --modelList: one SED column per extra model
        cols[f"SED_{os.path.basename(p)}"] = alt[:, i] - ref[:, i]
    df = pd.concat([df.reset_index()] + [pd.DataFrame(v, columns=[k]) for k, v in cols.items()], axis=1,
                   ignore_index=False)
    write_sed_tables(df, args.out_dir)
    lap("write_tsv_s")
    return {"n_rows": len(df), "ref": ref, "alt": alt, "models": paths, "timing": timing}


def main(argv=None) -> int:
    run(build_parser().parse_args(argv))
    return 0


if __name__ == "__main__":
    sys.exit(main())
