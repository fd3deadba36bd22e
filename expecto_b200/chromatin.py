"""Drop-in for the reference's ``chromatin.py`` (same CLI, same files out), computed on a B200.

    python -m expecto_b200.chromatin ./example/example.vcf [--maxshift 800 --output_dir chromatin_out ...]

Reference behaviour reproduced (file:line into /root/reference/chromatin.py):
  * arguments and defaults                                                     :26-42
  * VCF read, optional ``--chunk_i/--chunk_size`` row slice                    :211-214
  * ``{out}/snps_hg19.vcf`` = two header lines + ALL rows (before filtering)   :232-237
  * chromosome normalisation ``'chr'+str(c).replace('chr','')`` and CHRS filter :240-241
  * shift order ``[0,-200..-maxshift,200..maxshift]``                          :243
  * the three allele-match counters printed at shift 0                         :256-260
  * ``{out}/snps.shift_{s}.diff.h5`` with float32 ``diff``/``ref``/``alt`` [2N,2002], forward rows then
    reverse-complement rows, diff = alt - ref                                   :281-286
What differs by design: the windows are never built as strings — descriptors index a 2-bit genome in HBM
(expecto_b200/windows.py) and kernel 1 encodes them inside the forward; ``--cuda`` is accepted and ignored
(there is no CPU path); ``--hg38`` needs the third-party ``liftover`` package + chain files and is refused.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

CHRS = ["chr%d" % i for i in range(1, 23)] + ["chrX", "chrY"]  # chromatin.py:108-110


def shift_list(maxshift: int) -> list[int]:
    return [0] + list(range(-200, -maxshift - 1, -200)) + list(range(200, maxshift + 1, 200))  # chromatin.py:243


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Predict variant chromatin effects")
    p.add_argument("inputfile", type=str, help="Input file in vcf format")
    p.add_argument("--hg38", action="store_true", help="(unsupported here: needs the liftover package)")
    p.add_argument("--chunk_size", action="store", dest="chunk_size", type=int, default=int(1e5))
    p.add_argument("--chunk_i", action="store", dest="chunk_i", type=int, default=None)
    p.add_argument("--maxshift", action="store", dest="maxshift", type=int, default=800)
    p.add_argument("--inputsize", action="store", dest="inputsize", type=int, default=2000)
    p.add_argument("--batchsize", action="store", dest="batchsize", type=int, default=32,
                   help="accepted for compatibility; the GPU batch is --gpu_batch variants")
    p.add_argument("--output_dir", action="store", dest="output_dir", type=str, default="chromatin_out")
    p.add_argument("--cuda", action="store_true", help="accepted for compatibility (always on)")
    # extensions
    p.add_argument("--genome", default="./resources/hg19.fa", help="FASTA (chromatin.py:44 hard-codes this path)")
    p.add_argument("--weights", default="./resources/deepsea.beluga.pth", help="state_dict (chromatin.py:103)")
    p.add_argument("--dtype", default="f16", choices=["f16", "bf16", "f16x3"],
                   help="tensor-core operand type; f16x3 = fp16 (hi, lo) pairs, ~1e-4 of the fp32 reference at a third "
                        "of the throughput")
    p.add_argument("--device", type=int, default=0)
    p.add_argument("--gpu_batch", type=int, default=4096, help="variants per GPU sub-batch (x4 windows)")
    p.add_argument("--gather", default="nccl", choices=["nccl", "file"],
                   help="multi-GPU (torchrun) only: gather rows to rank 0 over NCCL, or every rank writes its slabs")
    p.add_argument("--mode", default="segments", choices=["segments", "windows"],
                   help="segments: conv stack once per variant segment (shared activations across shifts, "
                        "bit-identical); windows: every window separately, as the reference does")
    p.add_argument("--no_delta", action="store_true",
                   help="segment mode: run the alt segments of substitutions in full instead of as patches around the "
                        "variant (same results bit for bit; for comparisons)")
    p.add_argument("--host_descs", action="store_true",
                   help="build the window / segment descriptors on the host (windows.py) and loop over chunks in Python "
                        "instead of inside the library (same results bit for bit; for comparisons)")
    p.add_argument("--upstream_layout", action="store_true",
                   help="also write upstream ExPecto's {inputfile}.shift_{s}.diff.h5 with one 'pred' dataset")
    return p


def read_vcf(path: str, chunk_i=None, chunk_size=int(1e5)):
    """chromatin.py:211-214."""
    import pandas as pd

    vcf = pd.read_csv(path, sep="\t", header=None, comment="#")
    if chunk_i is not None:
        vcf = vcf.iloc[chunk_i * chunk_size:(chunk_i + 1) * chunk_size]
    return vcf


def write_snps_vcf(vcf, output_dir: str) -> str:
    """chromatin.py:232-237 — the same pandas call, so the bytes are identical."""
    path = f"{output_dir}/snps_hg19.vcf"
    with open(path, "w") as f:
        print("##fileformat=VCFv4.3", file=f)
        print("#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO", file=f)
    vcf.to_csv(path, sep="\t", header=False, index=False, mode="a")
    return path


def normalise_and_filter(vcf):
    """chromatin.py:240-241."""
    vcf = vcf.copy()
    vcf[vcf.columns[0]] = "chr" + vcf.iloc[:, 0].map(str).str.replace("chr", "")
    return vcf[vcf.iloc[:, 0].isin(CHRS)]


def variant_columns(vcf):
    chroms = [str(c) for c in vcf.iloc[:, 0]]
    pos = np.asarray([int(p) for p in vcf.iloc[:, 1]], dtype=np.int64)
    refs = [str(x) for x in vcf.iloc[:, 3]]
    alts = [str(x) for x in vcf.iloc[:, 4]]
    return chroms, pos, refs, alts


class SlabWriter:
    """Background file writer: the HDF5 row-slab writes of slab k run on threads while the GPU pipeline
    computes slab k+1 into the other set of pinned buffers (file output is as large as the GPU produces it —
    48 KB per variant-shift — so the disk, not the GPU, bounds chromatin.py-style output; SURVEY §7 hard part 4).
    Submissions are handled in order by one dispatcher; the slab writes inside one submission are independent
    positional writes and go out on ``n_threads`` workers (one thread's page-cache copy tops out near 3.6 GB/s on
    the B200 host, four reach ~5 GB/s: tools/disk_bench.py)."""

    def __init__(self, n_sets: int, n_threads: int = 4):
        import queue
        import threading
        from concurrent.futures import ThreadPoolExecutor
        self._q = queue.Queue()
        self._free = [threading.Event() for _ in range(n_sets)]
        for e in self._free:
            e.set()
        self._err = None
        self._next = 0
        self.busy_s = 0.0     # dispatcher time inside slab writes
        self.blocked_s = 0.0  # producer time spent waiting for a free buffer set
        self.bytes = 0
        self._pool = ThreadPoolExecutor(max_workers=max(1, n_threads)) if n_threads > 1 else None
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def _write(self, sink, tasks):
        import time
        t0 = time.perf_counter()
        try:
            self._write_tasks(sink, tasks)
        finally:
            self.busy_s += time.perf_counter() - t0
            self.bytes += sum(int(getattr(t[-1], "nbytes", 0)) for t in tasks)

    def _write_tasks(self, sink, tasks):
        if self._pool is None or len(tasks) < 2:
            for t in tasks:
                sink(*t)
        else:
            for f in [self._pool.submit(sink, *t) for t in tasks]:
                f.result()

    def _run(self):
        while True:
            item = self._q.get()
            if item is None:
                return
            set_idx, sink, tasks = item
            try:
                if tasks is None:
                    sink()  # ordered call (closing a finished shift's files)
                elif self._err is None:
                    self._write(sink, tasks)
            except BaseException as e:  # surfaced by wait_free / drain
                self._err = e
            finally:
                if set_idx is not None:
                    self._free[set_idx].set()

    def next_set(self) -> int:
        """Round-robin buffer set, continuing across shifts; blocks until its previous contents are on disk."""
        i = self._next
        self._next = (i + 1) % len(self._free)
        self.wait_free(i)
        return i

    def wait_free(self, set_idx: int) -> None:
        import time
        t0 = time.perf_counter()
        self._free[set_idx].wait()
        self.blocked_s += time.perf_counter() - t0
        if self._err is not None:
            raise self._err

    def submit(self, set_idx: int, sink, tasks) -> None:
        self._free[set_idx].clear()
        self._q.put((set_idx, sink, tasks))

    def call(self, fn) -> None:
        """Run ``fn()`` on the writer thread after everything submitted so far."""
        self._q.put((None, fn, None))

    def write(self, sink, tasks) -> None:
        """Ordered slab writes that do not release a buffer set (the set's last ``submit`` does)."""
        self._q.put((None, sink, tasks))

    def drain(self) -> None:
        for i in range(len(self._free)):
            self.wait_free(i)

    def close(self) -> None:
        self._q.put(None)
        self._t.join()
        if self._pool is not None:
            self._pool.shutdown()
        if self._err is not None:
            raise self._err


def predict_shift(engine, genome, chroms, pos, refs, alts, shift: int, gpu_batch: int, sink, pinned=None,
                  writer: SlabWriter | None = None) -> None:
    """One iteration of chromatin.py:243-286 for all variants.  ``sink(name, row0, array)`` receives
    float32 row slabs of the [2N,2002] datasets (forward rows at v, reverse-complement rows at N+v).
    Each slab of ``gpu_batch`` variants goes through the library's three-stream host pipeline
    (bx_chromatin_shift_host): descriptors H2D, kernel 1 + forward + diff, ref/alt/diff D2H.
    ``pinned``: one set of three pinned [2*gpu_batch,2002] tensors, or a list of sets used round-robin with
    ``writer`` so the file writes of one slab overlap the GPU work of the next."""
    import torch

    from .windows import check_valid_bases, chromatin_window_descs

    n = len(pos)
    sets = None
    if pinned is not None:
        sets = pinned if isinstance(pinned, list) else [pinned]
    for k, v0 in enumerate(range(0, n, gpu_batch)):
        v1 = min(n, v0 + gpu_batch)
        m = v1 - v0
        descs, alleles = chromatin_window_descs(genome, chroms[v0:v1], pos[v0:v1], refs[v0:v1], alts[v0:v1], shift)
        check_valid_bases(genome, descs)
        out, si = None, 0
        if sets is not None:
            si = writer.next_set() if writer is not None else k % len(sets)
            out = tuple(t[:2 * m] for t in sets[si])
        ref, alt, diff = engine.chromatin_shift_host(descs, alleles, m, average=False, out=out)
        tasks = []
        for name, t in (("diff", diff), ("ref", ref), ("alt", alt)):
            h = t.numpy()
            tasks.append((name, v0, h[:m]))
            tasks.append((name, n + v0, h[m:2 * m]))
        if writer is not None and sets is not None:
            writer.submit(si, sink, tasks)
        else:
            for name, row0, arr in tasks:
                sink(name, row0, arr)
    if writer is None:
        torch.cuda.synchronize(engine.device)


def _runs(idx: np.ndarray):
    """(offset in idx, first value, length) of each run of consecutive integers in ``idx``."""
    if len(idx) == 0:
        return
    cut = np.flatnonzero(np.diff(idx) != 1) + 1
    starts = np.concatenate([[0], cut])
    ends = np.concatenate([cut, [len(idx)]])
    for a, b in zip(starts, ends):
        yield int(a), int(idx[a]), int(b - a)


def _segment_chunks(engine, genome, chroms, pos, refs, alts, maxshift: int, ids, cap: int, delta: bool, stats=None):
    """Enqueue segment-mode forwards for the variants ``ids`` in chunks of at most ``cap`` and yield
    (variant indices, probs [S, m, 4, 2002] on the device).  Substitutions (SNVs / MNVs, ``windows.delta_eligible``)
    go through ``forward_segments_delta`` — ref segments in full, alt alleles as short patches — when ``delta``."""
    from .windows import (check_valid_bases, chromatin_delta_descs, chromatin_segment_descs, delta_eligible)

    seg_len = 2000 + 2 * maxshift
    ids = np.asarray(ids, dtype=np.int64)
    sub = delta_eligible([len(refs[i]) for i in ids], [len(alts[i]) for i in ids]) if (delta and len(ids)) else \
        np.zeros(len(ids), bool)
    if stats is not None:
        stats["delta_variants"] = int(sub.sum())
    import torch

    for c0 in range(0, len(ids), cap):
        # chunks stay contiguous in variant order (one row run per dataset and strand when they are written): the
        # chunk's substitutions and its other variants are run separately and merged on the device
        idx, is_sub = ids[c0:c0 + cap], sub[c0:c0 + cap]
        parts = []
        for kind, sel in (("delta", np.flatnonzero(is_sub)), ("full", np.flatnonzero(~is_sub))):
            if len(sel) == 0:
                continue
            part = idx[sel]
            m = len(part)
            args = (genome, [chroms[i] for i in part], pos[part], [refs[i] for i in part], [alts[i] for i in part],
                    maxshift)
            if kind == "delta":
                rd, pd, codes, k_of, row0, plen = chromatin_delta_descs(*args)
                check_valid_bases(genome, rd)
                probs = engine.forward_segments_delta(rd, pd, codes, m, seg_len, plen, row0, k_of, shift_major=True)
            else:
                descs, codes, k_of = chromatin_segment_descs(*args)
                check_valid_bases(genome, descs)
                probs = engine.forward_segments(descs, codes, m, 4, seg_len, k_of, shift_major=True)
            parts.append((sel, probs))
        if len(parts) == 1:
            yield idx, parts[0][1]
        else:
            merged = torch.empty((parts[0][1].shape[0], len(idx), 4, 2002), dtype=torch.float32, device=engine.device)
            for sel, probs in parts:
                # indices go up through the engine's pinned staging ring: a pageable source would make the copy wait
                # for the stream's earlier work and serialise this thread with the GPU
                merged.index_copy_(1, engine._stage(sel.astype(np.int64)).view(torch.int64), probs)
            yield idx, merged


def predict_all_shifts_segments(engine, genome, chroms, pos, refs, alts, maxshift: int, sinks: dict,
                                writer: SlabWriter | None = None, delta: bool = True, host_descs: bool = False) -> dict:
    """The whole shift loop (chromatin.py:243-286) for all variants: ``sinks[shift](name, row0, array)`` receives the
    float32 row slabs of that shift's ``diff`` / ``ref`` / ``alt`` datasets (forward rows at v, reverse-complement rows
    at n + v).

    The loop itself runs inside the library (include/beluga_b200.h: bx_chromatin_open / _submit / _wait): this function
    hands over the VCF columns once, then only moves finished chunks from pinned buffers to the writer.  Descriptors
    are built on the device, substitutions go through delta mode, other variants through four full segments or — long
    indels — window by window, and FC2's epilogue writes ref / alt / diff = alt - ref in their final row order; chunk
    k + 1 computes while chunk k travels to the host and chunk k - 1 is written.

    ``host_descs`` (and genomes holding characters outside encodeSeqs' dictionary, whose windows must raise KeyError
    as chromatin.py:166 does) use the host-side descriptor builder of windows.py instead — same bits."""
    import time

    import torch

    from .windows import variant_columns, variant_extent_invalid

    cols = variant_columns(genome, chroms, pos, refs, alts)
    if host_descs or variant_extent_invalid(genome, cols, maxshift) >= 0:
        return _predict_all_shifts_segments_host(engine, genome, chroms, pos, refs, alts, maxshift, sinks, writer, delta)
    n = cols["n"]
    shifts = shift_list(maxshift)
    S = len(shifts)
    n_chunks, chunk = engine.chromatin_open(cols, maxshift, delta=delta)
    n_sets = 3  # one being filled by the GPU, one waiting to be handed over, one being written
    own_writer = writer is None
    if own_writer:
        writer = SlabWriter(n_sets)
    rows_cap = max(1, min(chunk, n))
    pinned = [tuple(torch.empty(S * 2 * rows_cap * 2002, dtype=torch.float32, pin_memory=True) for _ in range(3))
              for _ in range(n_sets)]
    t_host = {"gpu_wait_s": 0.0}
    counts = {"delta": 0, "full": 0, "window": 0}

    def flush(p):
        k, si_set, v0, m, bufs = p
        t0 = time.perf_counter()
        engine.chromatin_wait(k)
        t_host["gpu_wait_s"] += time.perf_counter() - t0
        hs = [b.numpy() for b in bufs]           # (ref, alt, diff), each [S, 2, m, 2002]
        for si, shift in enumerate(shifts):
            tasks = []
            for name, h in (("diff", hs[2]), ("ref", hs[0]), ("alt", hs[1])):
                tasks.append((name, v0, h[si, 0]))
                tasks.append((name, n + v0, h[si, 1]))
            # every shift's tasks read the same buffer set; the last submit of the chunk (FIFO) frees it
            if si == S - 1:
                writer.submit(si_set, sinks[shift], tasks)
            else:
                writer.write(sinks[shift], tasks)

    pending = None
    try:
        for k in range(n_chunks):
            info = engine.chromatin_chunk_info(k)
            m = info["m"]
            for key in counts:
                counts[key] += info[key]
            si_set = writer.next_set()  # blocks until the set's previous contents are on disk
            bufs = tuple(t[:S * 2 * m * 2002].view(S, 2, m, 2002) for t in pinned[si_set])
            engine.chromatin_submit(k, *bufs)
            if pending is not None:
                flush(pending)
            pending = (k, si_set, info["v0"], m, bufs)
        if pending is not None:
            flush(pending)
        if own_writer:
            writer.close()
        else:
            writer.drain()
    finally:
        engine.chromatin_close()
    return {"segment_variants": counts["delta"] + counts["full"], "window_variants": counts["window"],
            "chunk_variants": int(chunk), "delta_variants": counts["delta"], "writer_busy_s": writer.busy_s,
            "writer_bytes": writer.bytes, "producer_blocked_s": writer.blocked_s, "descriptors": "device",
            **{"host_" + k: v for k, v in t_host.items()}}


def _predict_all_shifts_segments_host(engine, genome, chroms, pos, refs, alts, maxshift: int, sinks: dict,
                                      writer: SlabWriter | None = None, delta: bool = True) -> dict:
    """predict_all_shifts_segments with the descriptors built on the host (windows.py) and the chunk loop in Python: the ref/alt x fwd/rc sequences of a variant are
    encoded once as 2000+2*maxshift-bp segments, the conv stack runs once per segment and all shifts' ref/alt/diff
    rows come out together (bit-identical to the per-shift path).  ``sinks[shift](name, row0, array)`` receives the
    row slabs of that shift's datasets.  Variants whose windows are not plain substrings of one segment (long
    indels, ``windows.segment_eligible``) go through the per-window path."""
    import torch

    from .engine import chromatin_outputs
    from .windows import check_valid_bases, chromatin_window_descs, segment_eligible

    n = len(pos)
    shifts = shift_list(maxshift)
    S = len(shifts)
    seg_len = 2000 + 2 * maxshift
    ok = segment_eligible([len(r) for r in refs], [len(a) for a in alts], maxshift) if n else np.zeros(0, bool)
    elig, rest = np.flatnonzero(ok), np.flatnonzero(~ok)
    cap = max(1, min(engine.segment_capacity(seg_len) // 4, max(1, len(elig))))
    n_sets = 3  # one being filled by the GPU, one waiting to be handed over, one being written
    pinned = [torch.empty((S, 3, 2 * cap, 2002), dtype=torch.float32, pin_memory=True) for _ in range(n_sets)]
    own_writer = writer is None
    if own_writer:
        writer = SlabWriter(n_sets)
    # Software pipeline over chunks: while the GPU runs chunk k (compute stream) and its results stream to pinned
    # memory (copy stream), the host builds chunk k+1's descriptors and hands chunk k-1's slabs to the writer.
    import time
    copy_stream = torch.cuda.Stream(device=engine.device)
    pending = None
    t_host = {"enqueue_s": 0.0, "gpu_wait_s": 0.0}
    stats = {}

    def flush(p):
        done, si_set, idx, m = p
        t0 = time.perf_counter()
        done.synchronize()
        t_host["gpu_wait_s"] += time.perf_counter() - t0
        buf = pinned[si_set]
        for si, shift in enumerate(shifts):
            tasks = []
            for j, name in enumerate(("diff", "ref", "alt")):
                h = buf[si, j].numpy()
                for off, g0, ln in _runs(idx):
                    tasks.append((name, g0, h[off:off + ln]))
                    tasks.append((name, n + g0, h[m + off:m + off + ln]))
            # every shift's tasks read the same buffer set; the last submit of the chunk (FIFO) frees it
            if si == S - 1:
                writer.submit(si_set, sinks[shift], tasks)
            else:
                writer.write(sinks[shift], tasks)

    with torch.cuda.device(engine.device):
        t1 = time.perf_counter()
        for idx, probs in _segment_chunks(engine, genome, chroms, pos, refs, alts, maxshift, elig, cap, delta, stats):
            m = len(idx)
            outs = [chromatin_outputs(probs[si].reshape(4 * m, 2002), m) for si in range(S)]
            computed = torch.cuda.Event()
            computed.record()
            t_host["enqueue_s"] += time.perf_counter() - t1   # descriptors + launches of this chunk
            si_set = writer.next_set()  # blocks until the set's previous contents are on disk
            buf = pinned[si_set]
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(computed)
                for si, (ref, alt, diff) in enumerate(outs):
                    for j, t in enumerate((diff, ref, alt)):
                        buf[si, j, :2 * m].copy_(t, non_blocking=True)
                        t.record_stream(copy_stream)
                done = torch.cuda.Event()
                done.record(copy_stream)
            del probs, outs
            if pending is not None:
                flush(pending)
            pending = (done, si_set, idx, m)
            t1 = time.perf_counter()
        if pending is not None:
            flush(pending)
    # long indels etc.: per-window path, one variant-run at a time
    for shift in shifts if len(rest) else []:
        sub = [(chroms[i], pos[i], refs[i], alts[i]) for i in rest]
        descs, codes = chromatin_window_descs(genome, [x[0] for x in sub], np.array([x[1] for x in sub]),
                                              [x[2] for x in sub], [x[3] for x in sub], shift)
        check_valid_bases(genome, descs)
        m = len(rest)
        ref, alt, diff = engine.chromatin_shift_host(descs, codes, m, average=False)
        tasks = []
        for name, t in (("diff", diff), ("ref", ref), ("alt", alt)):
            h = t.numpy()
            for off, g0, ln in _runs(rest):
                tasks.append((name, g0, h[off:off + ln].copy()))
                tasks.append((name, n + g0, h[m + off:m + off + ln].copy()))
        writer.call(lambda sink=sinks[shift], tasks=tasks: [sink(*t) for t in tasks])
    if own_writer:
        writer.close()
    return {"segment_variants": int(len(elig)), "window_variants": int(len(rest)), "chunk_variants": int(cap),
            "delta_variants": int(stats.get("delta_variants", 0)),
            "writer_busy_s": writer.busy_s, "writer_bytes": writer.bytes, "producer_blocked_s": writer.blocked_s,
            **{"host_" + k: v for k, v in t_host.items()}}


def predict_all_shifts_segments_device(engine, genome, chroms, pos, refs, alts, maxshift: int, delta: bool = True,
                                       host_descs: bool = False):
    """The shift loop with the results left in HBM: {shift: (ref, alt, diff)} device tensors [2n,2002] in the strand-
    stacked row order (bx_chromatin_submit_device; ``host_descs``: the windows.py builder + Python chunk loop)."""
    import torch

    from .windows import variant_columns, variant_extent_invalid

    cols = variant_columns(genome, chroms, pos, refs, alts)
    if not host_descs and variant_extent_invalid(genome, cols, maxshift) < 0:
        n = cols["n"]
        shifts = shift_list(maxshift)
        S = len(shifts)
        full = [torch.empty((S, 2 * n, 2002), dtype=torch.float32, device=engine.device) for _ in range(3)]
        n_chunks, _ = engine.chromatin_open(cols, maxshift, delta=delta)
        try:
            with torch.cuda.device(engine.device):
                for k in range(n_chunks):
                    engine.chromatin_submit_device(k, *[t.view(-1, 2002) for t in full], 2 * n, n,
                                                   engine.chromatin_chunk_info(k)["v0"])
        finally:
            engine.chromatin_close()
        return {s: tuple(t[si] for t in full) for si, s in enumerate(shifts)}

    from .engine import chromatin_outputs
    from .windows import check_valid_bases, chromatin_window_descs, segment_eligible

    n = len(pos)
    shifts = shift_list(maxshift)
    seg_len = 2000 + 2 * maxshift
    outs = {s: tuple(torch.empty((2 * n, 2002), dtype=torch.float32, device=engine.device) for _ in range(3))
            for s in shifts}
    ok = segment_eligible([len(r) for r in refs], [len(a) for a in alts], maxshift) if n else np.zeros(0, bool)
    elig, rest = np.flatnonzero(ok), np.flatnonzero(~ok)
    cap = max(1, engine.segment_capacity(seg_len) // 4)

    def scatter(shift, idx, parts):
        m = len(idx)
        ii = torch.from_numpy(np.asarray(idx, dtype=np.int64)).to(engine.device)
        for full, part in zip(outs[shift], parts):
            full.index_copy_(0, ii, part[:m])
            full.index_copy_(0, ii + n, part[m:2 * m])

    for idx, probs in _segment_chunks(engine, genome, chroms, pos, refs, alts, maxshift, elig, cap, delta):
        m = len(idx)
        for si, shift in enumerate(shifts):
            scatter(shift, idx, chromatin_outputs(probs[si].reshape(4 * m, 2002), m))
        del probs
    if len(rest):
        for shift in shifts:
            descs, codes = chromatin_window_descs(genome, [chroms[i] for i in rest], pos[rest], [refs[i] for i in rest],
                                                  [alts[i] for i in rest], shift)
            check_valid_bases(genome, descs)
            probs = engine.forward_windows(descs, codes)
            scatter(shift, rest, chromatin_outputs(probs, len(rest)))
    return outs


def predict_all_shifts_averaged_device(engine, genome, chroms, pos, refs, alts, maxshift: int, delta: bool = True,
                                       host_descs: bool = False, padded: bool = False):
    """The chromatin step as ``predict.py`` consumes it, left in HBM: {"ref", "alt", "diff"} -> float32 [S, n, 2002],
    already averaged over the forward / reverse-complement rows (predict.py:186-190).  This is what reading the nine
    ``.diff.h5`` files back and averaging gives, without the 432 KB per variant of file traffic in between.  The
    averaging happens in FC2's epilogue (BX_CHROM_AVERAGE).  ``padded``: the tensors are views of buffers whose rows
    are ``transform.X_PITCH`` floats apart (128-byte aligned rows: what the scoring kernel reads fastest)."""
    import torch

    from .transform import padded_rows

    def alloc(S, n):
        if padded:
            return {k: padded_rows((S, n), engine.device) for k in ("ref", "alt", "diff")}
        return {k: torch.empty((S, n, 2002), dtype=torch.float32, device=engine.device) for k in ("ref", "alt", "diff")}

    from .windows import variant_columns, variant_extent_invalid

    cols = variant_columns(genome, chroms, pos, refs, alts)
    if not host_descs and variant_extent_invalid(genome, cols, maxshift) < 0:
        n = cols["n"]
        S = len(shift_list(maxshift))
        res = alloc(S, n)
        n_chunks, _ = engine.chromatin_open(cols, maxshift, delta=delta, average=True, windows_only=maxshift % 200 != 0)
        try:
            with torch.cuda.device(engine.device):
                for k in range(n_chunks):
                    engine.chromatin_submit_device(k, res["ref"], res["alt"], res["diff"], n, 0,
                                                   engine.chromatin_chunk_info(k)["v0"])
        finally:
            engine.chromatin_close()
        return res

    from .engine import chromatin_outputs
    from .windows import check_valid_bases, chromatin_window_descs, segment_eligible

    n = len(pos)
    shifts = shift_list(maxshift)
    S = len(shifts)
    seg_len = 2000 + 2 * maxshift
    out = alloc(S, n)
    use_segments = maxshift % 200 == 0
    ok = segment_eligible([len(r) for r in refs], [len(a) for a in alts], maxshift) if (n and use_segments) else np.zeros(n, bool)
    elig, rest = np.flatnonzero(ok), np.flatnonzero(~ok)
    if len(elig):
        cap = max(1, engine.segment_capacity(seg_len) // 4)
        for idx, probs in _segment_chunks(engine, genome, chroms, pos, refs, alts, maxshift, elig, cap, delta):
            m = len(idx)
            contiguous = m == int(idx[-1] - idx[0] + 1)
            ii = None if contiguous else engine._stage(idx.astype(np.int64)).view(torch.int64)
            for si in range(S):
                parts = chromatin_outputs(probs[si].reshape(4 * m, 2002), m, average=True)
                for k, t in zip(("ref", "alt", "diff"), parts):
                    if contiguous:
                        out[k][si, int(idx[0]):int(idx[0]) + m] = t
                    else:
                        out[k][si].index_copy_(0, ii, t)
            del probs
    gb = 1024
    for v0 in range(0, len(rest), gb):
        part = rest[v0:v0 + gb]
        m = len(part)
        ii = engine._stage(part.astype(np.int64)).view(torch.int64)
        for si, shift in enumerate(shifts):
            descs, codes = chromatin_window_descs(genome, [chroms[i] for i in part], pos[part], [refs[i] for i in part],
                                                  [alts[i] for i in part], shift)
            check_valid_bases(genome, descs)
            parts = chromatin_outputs(engine.forward_windows(descs, codes), m, average=True)
            for k, t in zip(("ref", "alt", "diff"), parts):
                out[k][si].index_copy_(0, ii, t)
    return out


def predict_all_shifts_gathered(engine, cols: dict, maxshift: int, n_total: int, sinks: dict | None,
                                writer: SlabWriter | None = None, delta: bool = True) -> dict:
    """The shift loop for variants sharded over ranks with the results gathered on rank 0 (north_star:
    "a final NCCL gather of the diff tensors over NVLink"), streamed chunk by chunk instead of holding every shift's
    [2 n_local, 2002] tensors until the end:

      round k, every rank:  chunk k of its shard -> device buffer [3, S, 2, chunk, 2002] (bx_chromatin_submit_device)
                            -> ``dist.gather`` to rank 0, in order on the compute stream;
      rank 0:               the gathered buffers -> pinned host memory on a side stream (copy engine), overlapping
                            chunk k + 1's compute, then row slabs to
                            ``sinks[shift](name, global_row0, array)`` (rank order == row order inside each strand
                            half of chromatin.py:281-286's layout) through ``writer``.

    ``cols``: this rank's columns (windows.variant_columns of its contiguous shard, shard.shard_range order).  Rank 0
    receives 3 x S x 2 x 2002 x 4 bytes = 432 KB per variant of the WHOLE job and must move them over its one PCIe
    link, which is what bounds this mode at 8 GPUs (the timings returned say so); ``--gather file`` has every rank
    copy and write its own rows instead.

    Why the gather is not overlapped with compute: NCCL's send / receive kernels need SMs, and the forward's GEMM kernels
    are persistent grids of one CTA per SM with statically strided tiles.  Issued on a side stream, the gather kernels
    got in only at kernel boundaries and the CTA pairs they displaced then delayed whole GEMM launches: measured at N = 2,
    65,536 variants: 2.87 s (gather 11.7 GB/s, 42 % of the wall time) against 1.30 s for the no-collective mode
    (profiles/r02_bench_n2_gather_on_side_stream.json).  In order, the gather has the GPU to itself for the
    millisecond it needs; only rank 0's device -> host copies, which run on a copy engine, overlap the next chunk."""
    import time

    import torch
    import torch.distributed as dist

    from . import shard

    world, rank = (dist.get_world_size(), dist.get_rank()) if dist.is_initialized() else (1, 0)
    dev = engine.device
    cuda = dev.type == "cuda"
    shifts = shift_list(maxshift)
    S = len(shifts)
    bounds = shard.shard_bounds(n_total, world)
    if cols["n"] != int(bounds[rank + 1] - bounds[rank]):
        raise ValueError(f"rank {rank} holds {cols['n']} variants, expected {int(bounds[rank + 1] - bounds[rank])}")
    # rank 0 stages world x chunk variants per round in pinned memory (two sets): chunks shrink as the world grows so
    # that a set stays under 768 MB (pinning gigabytes takes seconds)
    cap = max(64, int(768e6 // (world * 3 * S * 2 * 2002 * 4)))
    n_chunks, chunk = engine.chromatin_open(cols, maxshift, delta=delta, chunk_variants=cap)
    # every rank takes part in every round, so all agree on the chunk size and the number of rounds
    if world > 1:
        t = torch.tensor([n_chunks, -chunk], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        chunk_all = int(-t[1].item())
        if chunk_all != chunk:
            engine.chromatin_close()
            n_chunks, chunk = engine.chromatin_open(cols, maxshift, delta=delta, chunk_variants=chunk_all)
        t = torch.tensor([n_chunks], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        rounds = int(t.item())
    else:
        rounds = n_chunks
    chunk = max(chunk, 1)
    shape = (3, S, 2, chunk, 2002)
    send = [torch.zeros(shape, dtype=torch.float32, device=dev) for _ in range(2)]
    recv = [[torch.empty(shape, dtype=torch.float32, device=dev) for _ in range(world)] for _ in range(2)] \
        if (rank == 0 and world > 1) else None
    n_sets = 2
    host = [torch.empty((world,) + shape, dtype=torch.float32, pin_memory=cuda) for _ in range(n_sets)] if rank == 0 else None
    own_writer = writer is None and rank == 0
    if own_writer:
        writer = SlabWriter(n_sets)
    side = torch.cuda.Stream(device=dev) if cuda else None
    timing = {"gather_ms": 0.0, "d2h_ms": 0.0, "wait_s": 0.0, "rounds": rounds, "chunk_variants": chunk}
    ev_pairs = []
    # chunk sizes of every rank in every round (shards differ by at most one variant)
    sizes = [[max(0, min(chunk, int(bounds[r + 1] - bounds[r]) - k * chunk)) for r in range(world)] for k in range(rounds)]

    def flush(p):
        k, b, done = p
        t0 = time.perf_counter()
        if done is not None:
            done.synchronize()
        timing["wait_s"] += time.perf_counter() - t0
        h = host[b].numpy()
        for si, shift in enumerate(shifts):
            tasks = []
            for r in range(world):
                m = sizes[k][r]
                if m == 0:
                    continue
                g0 = int(bounds[r]) + k * chunk
                for j, name in ((2, "diff"), (0, "ref"), (1, "alt")):
                    tasks.append((name, g0, h[r, j, si, 0, :m]))
                    tasks.append((name, n_total + g0, h[r, j, si, 1, :m]))
            if si == S - 1:
                writer.submit(b, sinks[shift], tasks)
            else:
                writer.write(sinks[shift], tasks)

    pending = None
    d2h_done = [None, None]
    try:
        for k in range(rounds):
            b = k & 1
            if cuda and d2h_done[b] is not None:   # rank 0: recv[b] (send[b] when alone) has left for the host
                torch.cuda.current_stream(dev).wait_event(d2h_done[b])
            if k < n_chunks:
                flat = send[b].view(3, -1, 2002)
                engine.chromatin_submit_device(k, flat[0], flat[1], flat[2], 2 * chunk, chunk, 0)
            if rank == 0:
                hb = writer.next_set()     # == b: the pinned set's previous contents are on disk
                assert hb == b
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)] if (cuda and rank == 0) else None
            if e:
                e[0].record()
            if world > 1:
                # in order on the compute stream: with every SM free the 113 MB of a chunk cross NVLink in about a
                # millisecond, where a gather overlapped with the persistent GEMM kernels crawled (docstring)
                dist.gather(send[b], recv[b] if rank == 0 else None, dst=0)
            done = None
            if rank == 0:
                src = recv[b] if world > 1 else [send[b]]
                if cuda:
                    e[1].record()
                    side.wait_event(e[1])
                with (torch.cuda.stream(side) if cuda else _nullcontext()):
                    for r in range(world):
                        if sizes[k][r]:
                            host[b][r].copy_(src[r], non_blocking=True)
                    if cuda:
                        e[2].record()
                        done = d2h_done[b] = e[2]
                        ev_pairs.append(e)
                        for t_ in src:
                            t_.record_stream(side)
                if pending is not None:
                    flush(pending)
                pending = (k, b, done)
        if rank == 0 and pending is not None:
            flush(pending)
        if cuda:
            torch.cuda.synchronize(dev)
        if rank == 0:
            if own_writer:
                writer.close()
            else:
                writer.drain()
            for i, e in enumerate(ev_pairs):
                timing["gather_ms"] += e[0].elapsed_time(e[1])
                timing["d2h_ms"] += e[1].elapsed_time(e[2])
                if i:   # compute stream of rank 0 between the end of one gather and the start of the next
                    timing["compute_ms"] = timing.get("compute_ms", 0.0) + ev_pairs[i - 1][1].elapsed_time(e[0])
    finally:
        engine.chromatin_close()
    timing["gathered_bytes"] = int(3 * S * 2 * 2002 * 4) * (n_total - int(bounds[1] - bounds[0])) if world > 1 else 0
    return timing


class _nullcontext:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


def predict_shift_device(engine, genome, chroms, pos, refs, alts, shift: int, gpu_batch: int):
    """Same computation with the results left in HBM: (ref, alt, diff) device tensors [2n,2002] in the strand-
    stacked row order — the operands of the end-of-shift NCCL gather when variants are sharded over GPUs."""
    import torch

    from .engine import chromatin_outputs
    from .windows import check_valid_bases, chromatin_window_descs

    n = len(pos)
    outs = [torch.empty((2 * n, 2002), dtype=torch.float32, device=engine.device) for _ in range(3)]
    for v0 in range(0, n, gpu_batch):
        v1 = min(n, v0 + gpu_batch)
        m = v1 - v0
        descs, alleles = chromatin_window_descs(genome, chroms[v0:v1], pos[v0:v1], refs[v0:v1], alts[v0:v1], shift)
        check_valid_bases(genome, descs)
        probs = engine.forward_windows(descs, alleles)
        for full, part in zip(outs, chromatin_outputs(probs, m)):
            full[v0:v1] = part[:m]
            full[n + v0:n + v1] = part[m:]
    return tuple(outs)


def run(args, engine=None, genome=None) -> dict:
    """The whole script; returns {'n_variants', 'counters', 'files'} for callers (tests, bench).

    Under torchrun (one process per GPU) the filtered variants are split into contiguous per-rank shards
    (expecto_b200/shard.py); ``--gather nccl`` collects each shift's ref/alt/diff rows on rank 0 over NVLink and
    rank 0 writes the file, ``--gather file`` has every rank pwrite() its own row slabs (no collective)."""
    import torch

    from . import shard
    from .hdf5 import H5Writer
    from .windows import allele_match_counts

    if args.hg38:
        raise NotImplementedError("--hg38 needs the third-party `liftover` package and UCSC chain files, neither of "
                                  "which is available offline; lift the VCF to hg19 first (out of scope, SURVEY §8b)")
    if args.inputsize != 2000:
        raise ValueError("Beluga's architecture fixes --inputsize at 2000 (Linear(67840,...) = 640*106, Beluga.py:43)")
    rank, world, local = shard.init_distributed()
    os.makedirs(args.output_dir, exist_ok=True)
    if genome is None:
        genome = shard.load_genome(args.genome, device=engine.device if engine is not None else
                                   (local if world > 1 else args.device))
    if engine is None:
        from .engine import BelugaEngine
        engine = BelugaEngine(local if world > 1 else args.device, dtype=args.dtype,
                              capacity=min(4096, 4 * args.gpu_batch))
        engine.load_weights_file(args.weights)
    if engine._genome is not genome:
        engine.set_genome(genome)

    vcf = read_vcf(args.inputfile, args.chunk_i, args.chunk_size)
    if rank == 0:
        write_snps_vcf(vcf, args.output_dir)
    vcf = normalise_and_filter(vcf)
    chroms, pos, refs, alts = variant_columns(vcf)
    missing = sorted({c for c in chroms if c not in genome.offsets})
    if missing:
        raise KeyError(f"contig(s) {missing} are not in the genome")  # pyfasta raises KeyError too
    n = len(pos)
    lo, hi = shard.shard_range(n, rank, world)
    chroms, pos, refs, alts = chroms[lo:hi], pos[lo:hi], refs[lo:hi], alts[lo:hi]
    n_local = hi - lo
    files, counters = [], None
    if args.mode == "segments" and args.maxshift % 200 == 0 and (world == 1 or args.gather == "file"):
        return _run_segments(args, engine, genome, chroms, pos, refs, alts, n, lo, n_local, rank, world)
    gb = max(1, min(args.gpu_batch, max(n_local, 1)))
    seg_outs = None
    if args.mode == "segments" and args.maxshift % 200 == 0 and world > 1 and args.gather == "nccl":
        from .windows import variant_columns as vcols, variant_extent_invalid
        cols = vcols(genome, chroms, pos, refs, alts)
        bad = shard.allreduce_sum_ints([int(variant_extent_invalid(genome, cols, args.maxshift) >= 0)], device=engine.device)[0]
        if not bad and not getattr(args, "host_descs", False):
            return _run_gathered(args, engine, genome, cols, chroms, pos, refs, alts, n, rank, world)
        seg_outs = predict_all_shifts_segments_device(engine, genome, chroms, pos, refs, alts, args.maxshift,
                                                      delta=not args.no_delta, host_descs=True)
    pinned, writer = None, None
    if world == 1 or args.gather == "file":
        n_sets = 2  # slab k's file writes overlap slab k+1's GPU work (also across shifts)
        pinned = [tuple(torch.empty((2 * gb, 2002), dtype=torch.float32, pin_memory=True) for _ in range(3))
                  for _ in range(n_sets)]
        writer = SlabWriter(n_sets)
    shapes = {k: (2 * n, 2002) for k in ("diff", "ref", "alt")}
    for shift in shift_list(args.maxshift):
        if shift == 0:
            mine = allele_match_counts(genome, chroms, pos, refs, alts)
            counters = tuple(shard.allreduce_sum_ints(mine, device=engine.device if world > 1 else None))
            if rank == 0:
                print(f"Number of variants with reference allele matched with reference genome: {counters[0]}")
                print(f"Number of variants with alternate allele matched with reference genome: {counters[1]}")
                print(f"Number of input variants: {counters[2]}")
        path = f"{args.output_dir}/snps.shift_{str(shift)}.diff.h5"
        up_path = f"{args.inputfile}.shift_{str(shift)}.diff.h5"
        if world > 1 and args.gather == "nccl":
            local_out = seg_outs.pop(shift) if seg_outs is not None else \
                predict_shift_device(engine, genome, chroms, pos, refs, alts, shift, gb)
            full = [shard.gather_strand_rows(t, n) for t in local_out]  # ref, alt, diff
            if rank == 0:
                with H5Writer(path, shapes) as w:
                    for name, t in zip(("ref", "alt", "diff"), full):
                        w.write(name, t.cpu().numpy())
                if args.upstream_layout:
                    with H5Writer(up_path, {"pred": (2 * n, 2002)}) as up:
                        up.write("pred", full[2].cpu().numpy())
            del local_out, full
        else:
            # single process, or every rank writing its own slabs: rank 0 creates the files, the others attach
            if rank == 0:
                w = H5Writer(path, shapes)
                up = H5Writer(up_path, {"pred": (2 * n, 2002)}) if args.upstream_layout else None
            shard.barrier()
            if rank != 0:
                w = H5Writer(path, shapes, create=False)
                up = H5Writer(up_path, {"pred": (2 * n, 2002)}, create=False) if args.upstream_layout else None

            def sink(name, row0, arr, w=w, up=up):
                # local strand-stacked row -> global row: forward rows at lo+v, reverse-complement rows at n+lo+v
                g0 = lo + row0 if row0 < n_local else n + lo + (row0 - n_local)
                w.write_rows(name, g0, arr)
                if up is not None and name == "diff":
                    up.write_rows("pred", g0, arr)

            predict_shift(engine, genome, chroms, pos, refs, alts, shift, gb, sink, pinned, writer)

            def finish(w=w, up=up):
                w.close()
                if up is not None:
                    up.close()

            writer.call(finish)  # files of this shift close once their slabs are on disk; the next shift starts now
        files.append(path)
    if writer is not None:
        writer.close()
        shard.barrier()
    return {"n_variants": n, "counters": counters, "files": files, "rank": rank, "world": world}


def _run_gathered(args, engine, genome, cols, chroms, pos, refs, alts, n, rank, world) -> dict:
    """run() under torchrun with ``--gather nccl``: per-chunk NCCL gather to rank 0, which writes every file."""
    from . import shard
    from .hdf5 import H5Writer
    from .windows import allele_match_counts

    mine = allele_match_counts(genome, chroms, pos, refs, alts)
    counters = tuple(shard.allreduce_sum_ints(mine, device=engine.device))
    shifts = shift_list(args.maxshift)
    paths = {s: f"{args.output_dir}/snps.shift_{str(s)}.diff.h5" for s in shifts}
    ws, us, sinks = {}, {}, None
    if rank == 0:
        print(f"Number of variants with reference allele matched with reference genome: {counters[0]}")
        print(f"Number of variants with alternate allele matched with reference genome: {counters[1]}")
        print(f"Number of input variants: {counters[2]}")
        shapes = {k: (2 * n, 2002) for k in ("diff", "ref", "alt")}
        for s in shifts:
            ws[s] = H5Writer(paths[s], shapes)
            if args.upstream_layout:
                us[s] = H5Writer(f"{args.inputfile}.shift_{str(s)}.diff.h5", {"pred": (2 * n, 2002)})

        def make_sink(s):
            w, up = ws[s], us.get(s)

            def sink(name, row0, arr):
                w.write_rows(name, row0, arr)
                if up is not None and name == "diff":
                    up.write_rows("pred", row0, arr)
            return sink
        sinks = {s: make_sink(s) for s in shifts}
    info = predict_all_shifts_gathered(engine, cols, args.maxshift, n, sinks, delta=not args.no_delta)
    for s in ws:
        ws[s].close()
        if s in us:
            us[s].close()
    shard.barrier()
    return {"n_variants": n, "counters": counters, "files": [paths[s] for s in shifts], "rank": rank, "world": world,
            "mode": "segments", "gather": "nccl", **info}


def _run_segments(args, engine, genome, chroms, pos, refs, alts, n, lo, n_local, rank, world) -> dict:
    """run() in segment mode: all shifts of a variant come out of one pass, so every shift's file is open at once."""
    from . import shard
    from .hdf5 import H5Writer
    from .windows import allele_match_counts

    mine = allele_match_counts(genome, chroms, pos, refs, alts)
    counters = tuple(shard.allreduce_sum_ints(mine, device=engine.device if world > 1 else None))
    if rank == 0:
        print(f"Number of variants with reference allele matched with reference genome: {counters[0]}")
        print(f"Number of variants with alternate allele matched with reference genome: {counters[1]}")
        print(f"Number of input variants: {counters[2]}")
    shapes = {k: (2 * n, 2002) for k in ("diff", "ref", "alt")}
    shifts = shift_list(args.maxshift)
    paths = {s: f"{args.output_dir}/snps.shift_{str(s)}.diff.h5" for s in shifts}
    ups = {s: f"{args.inputfile}.shift_{str(s)}.diff.h5" for s in shifts}
    ws, us = {}, {}
    if rank == 0:
        for s in shifts:
            ws[s] = H5Writer(paths[s], shapes)
            if args.upstream_layout:
                us[s] = H5Writer(ups[s], {"pred": (2 * n, 2002)})
    shard.barrier()
    if rank != 0:
        for s in shifts:
            ws[s] = H5Writer(paths[s], shapes, create=False)
            if args.upstream_layout:
                us[s] = H5Writer(ups[s], {"pred": (2 * n, 2002)}, create=False)

    def make_sink(s):
        w, up = ws[s], us.get(s)

        def sink(name, row0, arr):
            # local strand-stacked row -> global row: forward rows at lo+v, reverse-complement rows at n+lo+v
            g0 = lo + row0 if row0 < n_local else n + lo + (row0 - n_local)
            w.write_rows(name, g0, arr)
            if up is not None and name == "diff":
                up.write_rows("pred", g0, arr)
        return sink

    info = predict_all_shifts_segments(engine, genome, chroms, pos, refs, alts, args.maxshift,
                                       {s: make_sink(s) for s in shifts}, delta=not args.no_delta,
                                       host_descs=getattr(args, "host_descs", False))
    for s in shifts:
        ws[s].close()
        if s in us:
            us[s].close()
    shard.barrier()
    return {"n_variants": n, "counters": counters, "files": [paths[s] for s in shifts], "rank": rank, "world": world,
            "mode": "segments", **info}


def main(argv=None) -> int:
    from . import shard
    args = build_parser().parse_args(argv)
    run(args)
    shard.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
