#!/usr/bin/env python
"""Benchmark of the Beluga chromatin hot path (BASELINE.json metric: Beluga 2-kb windows/s; tensor-pipe
utilisation vs the measured bf16 peak).

    python bench.py [--gpus N --steps K --warmup W --batch B --dtype f16|bf16]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      # the reference's CPU path (oracle port) on the host cores

A step = one Beluga forward over a batch of B synthetic 2000-bp one-hot windows per GPU (BASELINE configs[1],
B inside its 64..8192 sweep), random-init calibrated weights.  ``value`` is timed on the device with the batch
resident in HBM; ``e2e`` goes through the host-buffer C-ABI call (bx_forward_onehot_host: pinned host one-hot
in, pinned host probabilities out, copies inside the timed region).  One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_PER_WINDOW = 7_803_176_492          # SURVEY §3.2 (2*MAC, valid conv, every position the reference computes)
STAGE_FLOP = {"conv1": 40_816_640, "conv2": 3_253_862_400, "conv3": 1_201_766_400, "conv4": 1_776_844_800,
              "conv5": 555_417_600, "conv6": 694_681_600, "fc1": 271_767_040, "fc2": 8_020_012}
METRIC = "beluga_2kb_windows_per_sec"
UNIT = "windows/s"


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"bf16_tflops": float(p["bf16_tflops"]), "bf16_tflops_sustained": float(p.get("bf16_tflops_sustained",
                p["bf16_tflops"])), "hbm_gbs": float(p["hbm_gbs"]), "source": "measured (MEASURED_PEAKS.json)"}
    return {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "hbm_gbs": 6650.0,
            "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """Samples SM clock / throttle reasons with NVML while the timed region runs (B200_PROFILING.md clocks line)."""

    def __init__(self, index: int, period: float = 0.02):
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:  # pragma: no cover
            self.nv, self.err = None, str(e)

    _NAMES = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
              0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x10: "sync_boost",
              0x100: "display_clock_setting"}

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for bit, name in self._NAMES.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(self.period)

    def start(self):
        if self.nv is not None:
            self._stop.clear()
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        if self._thr is not None:
            self._stop.set()
            self._thr.join()
            self._thr = None

    def summary(self):
        if self.nv is None or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "NVML unavailable"}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def cpu_forward_rate(batch: int, min_seconds: float, max_windows: int, seed: int = 2):
    """The oracle's fp32 Beluga forward (restating Beluga.py:21-51 with torch CPU ops) on all host threads."""
    import numpy as np
    import torch

    from expecto_b200 import synth
    from oracle import beluga_oracle as O

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = {k: torch.from_numpy(v) for k, v in synth.make_beluga_weights(0, round_dtype="f16").items()}
    x = synth.random_onehot(batch, seed=seed).astype(np.float32)
    O.beluga_forward(x, sd)  # warm-up (oneDNN primitive creation)
    done, t0 = 0, time.perf_counter()
    while True:
        O.beluga_forward(x, sd)
        done += batch
        dt = time.perf_counter() - t0
        if dt >= min_seconds or done >= max_windows:
            break
    return done / dt, cores, torch.get_num_threads(), done, dt


def variant_workflow(eng, n_variants: int, rank: int = 0, world: int = 1, sync=None, cpu: bool = False):
    """BASELINE configs[2] shape at a stated reduced size, outside the timed region: a synthetic VCF (90 % SNV, 5 %
    ref-mismatch, 5 % indel) on a synthetic genome -> ref/alt x fwd/rc x 9 shifts -> pinned host ref/alt/diff, through
    the drop-in's segment + delta mode (host descriptors in, host arrays out, no file).  Reported in the headline
    metric's unit: windows/s = 36 windows per variant, every one bit-identical to the per-window forward."""
    import numpy as np
    import torch

    from expecto_b200 import chromatin, synth
    from expecto_b200.genome import PackedGenome

    genome = synth.make_genome([("chr1", 6_000_000), ("chr2", 4_000_000)], seed=3)
    pg = PackedGenome.from_sequences(genome)
    eng.set_genome(pg)
    rows = synth.make_vcf_rows(genome, n_variants, seed=3 + rank, odd_names=False)  # every rank: its own shard
    chroms, pos = [r[0] for r in rows], np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    sinks = {s: (lambda *a: None) for s in chromatin.shift_list(800)}
    k = min(1200, n_variants)
    chromatin.predict_all_shifts_segments(eng, pg, chroms[:k], pos[:k], refs[:k], alts[:k], 800, sinks)  # warm-up
    torch.cuda.synchronize()
    if sync is not None:
        sync(0.0)                      # barrier: all ranks start together
    t0 = time.perf_counter()
    info = chromatin.predict_all_shifts_segments(eng, pg, chroms, pos, refs, alts, 800, sinks)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if sync is not None:
        dt = sync(dt)                  # slowest rank
    n_all = n_variants * world
    out = {"workload": f"{n_variants} synthetic variants per GPU x (ref, alt) x (fwd, rc) x 9 shifts, segment + delta "
                       "mode, host descriptors -> pinned host ref/alt/diff, no file; variants sharded over the GPUs, "
                       "no collective",
           "variants": n_all, "windows": 36 * n_all, "seconds": dt, "windows_per_s": 36 * n_all / dt,
           "variants_per_s": n_all / dt, "d2h_bytes": 9 * 3 * 2 * n_all * 2002 * 4,
           "delta_variants": info["delta_variants"], "window_path_variants": info["window_variants"]}
    if cpu and rank == 0:
        out["cpu_baseline"] = chromatin_cpu_baseline(genome, rows)
    return out


def sharded_chromatin(eng, total_variants: int, rank: int, world: int, barrier, max_over_ranks):
    """The multi-GPU design of the variant path in front of the driver (strong scaling: ONE synthetic VCF of
    ``total_variants`` rows split into contiguous per-rank shards), both ways the drop-in offers:

    * ``gather_nccl``  — every chunk's ref / alt / diff rows are gathered to rank 0 over NCCL on a side stream while the
      next chunk computes, and rank 0 copies them to pinned host memory (``chromatin --gather nccl`` minus the file);
    * ``local_copy``   — every rank copies its own rows to its own pinned host memory, no collective
      (``chromatin --gather file`` minus the file).

    Timed barrier to barrier, slowest rank; windows/s = 36 windows per variant of the whole job."""
    import numpy as np
    import torch

    from expecto_b200 import chromatin, shard, synth
    from expecto_b200.genome import PackedGenome
    from expecto_b200.windows import variant_columns

    genome = synth.make_genome([("chr1", 6_000_000), ("chr2", 4_000_000)], seed=3)
    pg = PackedGenome.from_sequences(genome)
    eng.set_genome(pg)
    rows = synth.make_vcf_rows(genome, total_variants, seed=17, odd_names=False)     # the same table on every rank
    lo, hi = shard.shard_range(total_variants, rank, world)
    part = rows[lo:hi]
    chroms, pos = [r[0] for r in part], np.array([r[1] for r in part])
    refs, alts = [r[3] for r in part], [r[4] for r in part]
    cols = variant_columns(pg, chroms, pos, refs, alts)
    shifts = chromatin.shift_list(800)
    noop = {s: (lambda *a: None) for s in shifts}
    out = {"workload": f"one synthetic VCF of {total_variants} variants (90 % SNV, 5 % ref-mismatch, 5 % indel) split "
                       f"over {world} GPU(s) in contiguous shards x (ref, alt) x (fwd, rc) x 9 shifts, segment + delta "
                       "mode with device-built descriptors; results delivered to pinned host memory",
           "scaling": "strong", "total_variants": total_variants, "windows": 36 * total_variants}
    k = min(1200, len(pos))
    chromatin.predict_all_shifts_segments(eng, pg, chroms[:k], pos[:k], refs[:k], alts[:k], 800, noop)  # warm-up
    # warm-up of the gather mode on a small table sharded the same way (same chunk size, so the pinned and device
    # staging buffers of the timed call come from the allocators' caches)
    n_warm = min(total_variants, 700 * world)
    wlo, whi = shard.shard_range(n_warm, rank, world)
    wpart = rows[wlo:whi]
    wcols = variant_columns(pg, [r[0] for r in wpart], np.array([r[1] for r in wpart]), [r[3] for r in wpart],
                            [r[4] for r in wpart])
    chromatin.predict_all_shifts_gathered(eng, wcols, 800, n_warm, noop if rank == 0 else None)
    for mode in ("gather_nccl", "local_copy"):
        barrier()
        t0 = time.perf_counter()
        if mode == "gather_nccl":
            info = chromatin.predict_all_shifts_gathered(eng, cols, 800, total_variants, noop if rank == 0 else None)
        else:
            info = chromatin.predict_all_shifts_segments(eng, pg, chroms, pos, refs, alts, 800, noop)
        torch.cuda.synchronize()
        dt = max_over_ranks(time.perf_counter() - t0)
        barrier()
        res = {"seconds": dt, "windows_per_s": 36 * total_variants / dt, "variants_per_s": total_variants / dt}
        if mode == "gather_nccl":
            gb = info["gathered_bytes"] / 1e9
            res.update({"gathered_GB_to_rank0": gb, "rank0_gather_ms": info["gather_ms"], "rank0_d2h_ms": info["d2h_ms"],
                        "rank0_host_wait_s": info["wait_s"], "rank0_compute_ms": info.get("compute_ms"),
                        "gather_GBps": (gb / (info["gather_ms"] * 1e-3)) if info["gather_ms"] > 0 and gb > 0 else None,
                        "frac_of_wall_in_gather": info["gather_ms"] * 1e-3 / dt,
                        "frac_of_wall_in_rank0_d2h": info["d2h_ms"] * 1e-3 / dt,
                        "rank0_d2h_GBps": (36 * total_variants / 4 * 3 * 2002 * 4 / 1e9) / (info["d2h_ms"] * 1e-3)
                        if info["d2h_ms"] > 0 else None,
                        "note": "gather / d2h times are CUDA-event sums on rank 0's side stream; they overlap compute"})
        out[mode] = res
    return out


def transform_rooflines(dev, peaks, iters: int = 10, cpu: bool = True):
    """Kernel 4 (decay-weighted shift reductions) as HBM GB/s over the algorithmic bytes of SURVEY §8(d); inputs
    (1.6 GB / 1.2 GB) are larger than the 126 MB L2."""
    import numpy as np
    import torch

    from expecto_b200.transform import (chromatin_shifts, decay_transform_fixed, decay_transform_variant,
                                        fixed_pos_weights)

    out = {}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(dev)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1) / iters * 1e-3

    G = 512
    probs = torch.rand((G, 200, 2, 2002), dtype=torch.float32, device=dev)
    w = torch.from_numpy(fixed_pos_weights().astype(np.float32)).to(dev)  # staged once: the timed region is the kernel
    res = torch.empty((G, 20020), dtype=torch.float32, device=dev)
    t = timed(lambda: decay_transform_fixed(probs, w, out=res))
    b = G * (200 * 2 * 2002 * 4 + 20020 * 4)
    out["fixed"] = {"genes": G, "bytes_per_gene": b // G, "ms": t * 1e3, "achieved": b / t / 1e9, "unit": "GB/s",
                    "peak": peaks["hbm_gbs"], "frac": b / t / 1e9 / peaks["hbm_gbs"], "genes_per_s": G / t}
    del probs, res
    n = 16384
    x = torch.rand((9, n, 2002), dtype=torch.float32, device=dev)
    dist = np.random.default_rng(0).integers(-40000, 40000, size=n).astype(np.float64)
    strand = np.where(np.arange(n) % 2 == 0, 1, -1)
    res = torch.empty((n, 20020), dtype=torch.float32, device=dev)
    dist_d = torch.from_numpy(dist).to(dev)
    strand_d = torch.from_numpy(strand.astype(np.int8)).to(dev)
    shifts_d = torch.tensor(chromatin_shifts(800), dtype=torch.int32, device=dev)
    t = timed(lambda: decay_transform_variant(x, dist_d, strand_d, shifts=shifts_d, out=res))
    b = n * (9 * 2002 * 4 + 20020 * 4 + 9)
    out["variant"] = {"variants": n, "bytes_per_variant": b // n, "ms": t * 1e3, "achieved": b / t / 1e9,
                      "unit": "GB/s", "peak": peaks["hbm_gbs"], "frac": b / t / 1e9 / peaks["hbm_gbs"],
                      "variants_per_s": n / t, "note": "dist/strand staged on the device before the timed region"}
    try:  # DRAM bytes per launch from the ncu capture of this command (profiles/r01_ncu_transform_launches.csv)
        tr = json.load(open(os.path.join(ROOT, "profiles", "dram_traffic.json")))["transform"]
        out["fixed"]["traffic"] = int(tr["fixed_per_gene_bytes"] * G)
        out["variant"]["traffic"] = int(tr["variant_per_variant_bytes"] * n)
        out["fixed"]["traffic_source"] = out["variant"]["traffic_source"] = \
            "ncu capture r01 (profiles/r01_ncu_transform_launches.csv), static: not measured in this run"
    except Exception:
        out["fixed"]["traffic"] = out["variant"]["traffic"] = None
    del res
    # ---- transform fused with gblinear scoring (predict.py:120-160) as one tcgen05 GEMM: HBM-bound on the x read
    from expecto_b200.transform import X_PITCH, ScoreModels, decay_score_variant, padded_rows
    m = 218                                    # resources/modellist
    rng = np.random.default_rng(5)
    models_w = (rng.standard_normal((m, 20020)) * 0.01).astype(np.float32)
    models_b = rng.standard_normal(m).astype(np.float32)
    models = ScoreModels(models_w, models_b, dev)
    models.prepared()
    # four full waves of 74 CTA pairs x 256 variants
    del x
    n = 74 * 256 * 4
    x = padded_rows((9, n), dev)               # rows X_PITCH floats apart, as FC2's epilogue leaves them for predict
    x.uniform_()
    dist_d = torch.from_numpy(np.random.default_rng(1).integers(-40000, 40000, size=n).astype(np.float64)).to(dev)
    strand_d = torch.from_numpy(np.where(np.arange(n) % 2 == 0, 1, -1).astype(np.int8)).to(dev)
    t = timed(lambda: decay_score_variant(x, dist_d, strand_d, models, shifts=shifts_d))
    b = n * (9 * 2002 * 4 + m * 4 + 9)
    flop = 2.0 * n * 20020 * m
    out["score"] = {"variants": n, "models": m, "bytes_per_variant": b // n, "ms": t * 1e3, "achieved": b / t / 1e9,
                    "unit": "GB/s", "peak": peaks["hbm_gbs"], "frac": b / t / 1e9 / peaks["hbm_gbs"],
                    "variants_per_s": n / t, "kernel": "score_tma_kernel",
                    "algorithmic_tflops": flop / t / 1e12, "executed_tflops": 3 * flop * (20160 / 20020) * (224 / m) / t / 1e12,
                    "x_row_pitch_floats": X_PITCH,
                    "note": "features never leave shared memory; every product is three fp16 MMAs (hi/lo split operands); "
                            "bytes are the algorithmic 2002 floats per row, rows are padded to x_row_pitch_floats"}
    # rows grouped by side of the TSS (transform.side_order; predict hands its rows over like that wherever it gathers
    # them anyway): whole K steps of the operand are zero for every row of most tiles and the kernel skips them
    from expecto_b200.transform import side_order
    so = torch.from_numpy(side_order(dist_d.cpu().numpy(), strand_d.cpu().numpy())).to(dev)
    dist_s, strand_s = dist_d[so].contiguous(), strand_d[so].contiguous()
    t_so = timed(lambda: decay_score_variant(x, dist_s, strand_s, models, shifts=shifts_d))
    out["score"]["rows_ordered_by_side"] = {"ms": t_so * 1e3, "achieved": b / t_so / 1e9, "frac": b / t_so / 1e9 / peaks["hbm_gbs"]}
    # one model, as a single predict.py run has it (--model_save_file): the x stream with next to no MMA work
    one = ScoreModels(np.ascontiguousarray(models_w[:1]), models_b[:1].copy(), dev)
    one.prepared()
    t_one = timed(lambda: decay_score_variant(x, dist_d, strand_d, one, shifts=shifts_d))
    out["score"]["one_model"] = {"ms": t_one * 1e3, "achieved": n * (9 * 2002 * 4 + 4 + 9) / t_one / 1e9,
                                 "frac": n * (9 * 2002 * 4 + 4 + 9) / t_one / 1e9 / peaks["hbm_gbs"]}
    del one
    # the same launch back to back for ~2 s: the board reaches its power cap and the SM clock drops (kernel 5)
    e0.record()
    for _ in range(800):
        decay_score_variant(x, dist_d, strand_d, models, shifts=shifts_d)
    e1.record()
    torch.cuda.synchronize(dev)
    t_sus = e0.elapsed_time(e1) / 800 * 1e-3
    out["score"]["ms_sustained"] = t_sus * 1e3
    out["score"]["frac_sustained"] = b / t_sus / 1e9 / peaks["hbm_gbs"]
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "dram_traffic.json")))["transform"]
        out["score"]["traffic"] = int(tr["score_per_variant_bytes"] * n)
        out["score"]["traffic_source"] = "ncu launch list r02 (profiles/r02_ncu_score_launches.csv), static: not measured in this run"
    except Exception:
        out["score"]["traffic"] = None
    xd = x.contiguous()                        # dense 8008-byte rows (a caller's own [9, n, 2002] array): score_tc_kernel
    out["score"]["ms_dense_rows"] = timed(lambda: decay_score_variant(xd, dist_d, strand_d, models, shifts=shifts_d)) * 1e3
    del xd
    t_simt = timed(lambda: decay_score_variant(x[:, :2048].contiguous(), dist_d[:2048], strand_d[:2048], models,
                                               shifts=shifts_d, tensor_cores=False))
    out["score"]["simt_kernel_ms_per_variant"] = t_simt * 1e3 / 2048
    out["score"]["speedup_vs_simt_kernel"] = (t_simt / 2048) / (t / n)
    del x
    if cpu:
        out["fixed"]["cpu_baseline"], out["variant"]["cpu_baseline"] = transform_cpu_baselines()
    return out


def transform_cpu_baselines(genes: int = 64, variants: int = 5000):
    """The reference's numpy expressions for kernel 4 on this box's host (oracle restatements, bounded samples):
    compute_expecto_features.py:123-124 on ``genes`` genes and predict.py:120-136 (``compute_effects``' reduce over
    np.tile / np.repeat temporaries, 500-row batches) on ``variants`` variants, as GB/s over the same algorithmic
    bytes the GPU figures use."""
    import numpy as np

    from oracle import beluga_oracle as O

    rng = np.random.default_rng(0)
    cores = os.cpu_count() or 1
    pw = O.fixed_pos_weights()
    fwd = rng.random((200, 2002), dtype=np.float32)
    rc = rng.random((200, 2002), dtype=np.float32)
    O.fixed_transform(fwd, rc, pw)
    t0 = time.perf_counter()
    for _ in range(genes):
        O.fixed_transform(fwd, rc, pw)
    dt = time.perf_counter() - t0
    b = genes * (200 * 2 * 2002 * 4 + 20020 * 4)
    fixed = {"value": b / dt / 1e9, "unit": "GB/s", "genes_per_s": genes / dt, "cores": cores, "kind": "port",
             "sample": f"{genes} genes, numpy float64 expression of compute_expecto_features.py:123-124 "
                       "(numpy's own threading; the expression is a broadcast multiply + sum, single-threaded)"}
    x = [rng.random((500, 2002), dtype=np.float32) for _ in range(9)]
    done, t0 = 0, time.perf_counter()
    while done < variants:
        dist = rng.integers(-40000, 40000, size=500).astype(np.float64)
        strand = np.where(np.arange(500) % 2 == 0, "+", "-")
        O.variant_transform(x, O.variant_weights(dist, strand, 800))
        done += 500
    dt = time.perf_counter() - t0
    b = done * (9 * 2002 * 4 + 20020 * 4 + 9)
    variant = {"value": b / dt / 1e9, "unit": "GB/s", "variants_per_s": done / dt, "cores": cores, "kind": "port",
               "sample": f"{done} variants in 500-row batches (predict.py's --batchSize), one of diff/ref/alt: numpy "
                         "float64 np.tile / np.repeat reduce of predict.py:87-124"}
    return fixed, variant


def chromatin_cpu_baseline(genome: dict, rows, n_variants: int = 8):
    """The reference's chromatin.py shift loop on the host (oracle.chromatin_shift restating chromatin.py:243-281:
    fetchSeqs + encodeSeqs + fp32 forward for ref and alt, all 9 shifts) on the first ``n_variants`` rows of the same
    VCF the GPU workflow runs; windows/s with 4 windows per variant-shift."""
    import torch

    from expecto_b200 import synth
    from oracle import beluga_oracle as O

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = {k: torch.from_numpy(v) for k, v in synth.make_beluga_weights(0, round_dtype="f16").items()}
    dg = O.DictGenome(genome)
    part = [list(r) for r in rows[:n_variants]]
    O.chromatin_shift(dg, part[:1], sd, 0)   # warm-up
    t0 = time.perf_counter()
    for shift in O.SHIFTS_CHROMATIN(800):
        O.chromatin_shift(dg, part, sd, shift)
    dt = time.perf_counter() - t0
    w = 36 * len(part)
    return {"value": w / dt, "unit": UNIT, "variants_per_s": len(part) / dt, "cores": cores, "kind": "port",
            "sample": f"{len(part)} variants x 9 shifts x (ref, alt) x (fwd, rc) = {w} windows in {dt:.1f} s: fetchSeqs + "
                      f"encodeSeqs + fp32 torch CPU forward at the reference's batch of 32, {torch.get_num_threads()} "
                      "threads"}


def run_reference(args) -> int:
    """--impl reference: the reference's CPU implementation of the path (oracle port of Beluga.forward, torch
    fp32 on every host core), each step a bounded sample (--ref_batch windows, the reference's default batch of
    32) of the same workload.  Rank 0 only; other ranks exit 0."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    import numpy as np
    import torch

    from expecto_b200 import synth
    from oracle import beluga_oracle as O

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = {k: torch.from_numpy(v) for k, v in synth.make_beluga_weights(0, round_dtype="f16").items()}
    b = args.ref_batch
    x = synth.random_onehot(b, seed=2).astype(np.float32)
    for _ in range(max(1, args.warmup)):
        O.beluga_forward(x, sd)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        O.beluga_forward(x, sd)
    dt = time.perf_counter() - t0
    value = args.steps * b / dt
    sample = f"{args.steps} steps x {b} windows (fp32 torch CPU forward, {torch.get_num_threads()} threads)"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"Beluga forward over {args.batch} synthetic 2000-bp one-hot windows per GPU per step "
                               f"(BASELINE configs[1] batch sweep point B={args.batch}), calibrated random-init weights",
                   "sample": f"CPU arm times a bounded sample of {b} windows per step (the reference's default batch)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def run_gpu(args) -> int:
    import numpy as np
    import torch
    import torch.distributed as dist

    from expecto_b200 import _lib, synth
    from expecto_b200.engine import BelugaEngine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if args.gpus > 1 and world == 1:
        raise SystemExit("for --gpus N>1 launch with: python -m torch.distributed.run --nnodes=1 --nproc-per-node N "
                         "--master-addr 127.0.0.1 --master-port P bench.py --gpus N ...")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # keep stdout to the one JSON line
        # NCCL_DEBUG=VERSION makes NCCL printf its version to stdout at its first initialisation: file descriptor 1
        # points at stderr until the first collective is through, so that stdout carries the one JSON line only
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize(dev)
        finally:
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(v: float) -> float:
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    _lib.build()
    B, K, W = args.batch, args.steps, max(3, args.warmup)
    eng = BelugaEngine(local, dtype=args.dtype, cta_group=args.cta_group, capacity=min(B, args.capacity))
    eng.load_state_dict(synth.make_beluga_weights(0, round_dtype=None if args.dtype == "f16x3" else args.dtype))
    x_host = torch.from_numpy(synth.random_onehot(B, seed=2 + rank)).pin_memory()
    x = x_host.to(dev)
    out = torch.empty((B, 2002), dtype=torch.float32, device=dev)
    out_host = torch.empty((B, 2002), dtype=torch.float32, pin_memory=True)
    sampler = ClockSampler(local)
    L = _lib.lib()

    # ---------------- device-resident throughput (value) + per-stage event timing (roofline)
    for _ in range(W):
        eng.forward_onehot(x, out)
    barrier()
    eng.profile(True)
    eng.profile_read(reset=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = L.bx_launch_count()
    sampler.start()
    barrier()
    e0.record()
    for _ in range(K):
        eng.forward_onehot(x, out)
    e1.record()
    barrier()
    sampler.stop()
    launches = L.bx_launch_count() - launches0
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    stages = eng.profile_read(reset=True)
    eng.profile(False)
    checksum = float(out.sum().item())
    value = world * B * K / (ms_total * 1e-3)

    # ---------------- end to end through the host-buffer C-ABI call
    for _ in range(2):
        eng.forward_onehot_host(x_host, out_host)
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        eng.forward_onehot_host(x_host, out_host)   # returns when out_host is complete
    torch.cuda.synchronize(dev)
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    e2e_value = world * B * K / e2e_s
    e2e_ok = bool(torch.equal(out_host, out.cpu()))

    wf = sc = None
    if world > 1 and not args.no_workflow:   # all ranks take part (N = 1 runs it after the transform measurement)
        wf = variant_workflow(eng, args.workflow_variants, rank, world, max_over_ranks)
    if world > 1 and not args.no_sharded:
        sc = sharded_chromatin(eng, args.sharded_variants, rank, world, barrier, max_over_ranks)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peaks = load_peaks()
    per = {k: (ms / max(cnt, 1)) for k, (ms, cnt) in stages.items() if cnt}
    launches_per_step = {k: cnt / K for k, (ms, cnt) in stages.items() if cnt}
    chunk = min(B, args.capacity)
    top = max((k for k in per if k in STAGE_FLOP), key=lambda k: per[k] * launches_per_step[k])
    top_tf = chunk * STAGE_FLOP[top] / (per[top] * 1e-3) / 1e12 if B % chunk == 0 else \
        B * STAGE_FLOP[top] / (per[top] * launches_per_step[top] * 1e-3) / 1e12
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "dram_traffic.json")
    if os.path.exists(tpath):
        try:
            per_window = json.load(open(tpath))["per_window_bytes"].get(top)
            traffic = None if per_window is None else int(per_window * chunk)  # conv traffic is linear in windows
        except Exception:
            traffic = None
    fwd_tf = value / world * FLOP_PER_WINDOW / 1e12
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": f"Beluga forward over {B} synthetic 2000-bp one-hot windows per GPU per step "
                               f"(BASELINE configs[1] batch sweep point B={B}), calibrated random-init weights",
                   "batch_per_gpu": B, "chunk_windows": chunk, "cta_group": args.cta_group,
                   "operands": ("fp16 (hi, lo) pairs, Ah*Wh + Ah*Wl + Al*Wh -> fp32 accumulate: three tcgen05.mma "
                                "kind::f16 per product, FLOP figures count the algorithmic product once"
                                if args.dtype == "f16x3" else
                                f"{args.dtype} x {args.dtype} -> fp32 accumulate (tcgen05.mma kind::f16)"),
                   "l2": "per-step activation traffic (2.7 MB/window) far exceeds the 126 MB L2; no explicit flush",
                   "parallelism": f"dp{world} (windows sharded, no collective on the path)"},
        "tflops_per_gpu": fwd_tf,
        "roofline": {"bound": "tensor", "kernel": f"gemm_tc_kernel ({top})", "achieved": top_tf,
                     "peak": peaks["bf16_tflops"], "unit": "TFLOP/s", "frac": top_tf / peaks["bf16_tflops"],
                     "traffic": traffic,
                     "traffic_source": "ncu --set full captures (profiles/dram_traffic.json: conv2 from round 2 at this chunk size, "
                                       "other stages round 1), static: not measured in this run",
                     "peak_source": peaks["source"] + ", burst",
                     "peak_sustained": peaks["bf16_tflops_sustained"],
                     "frac_sustained": top_tf / peaks["bf16_tflops_sustained"],
                     "flop_per_launch": chunk * STAGE_FLOP[top], "ms_per_launch": per[top],
                     "share_of_step": per[top] * launches_per_step[top] / (ms_total / K)},
        "forward_roofline": {"achieved": fwd_tf, "frac": fwd_tf / peaks["bf16_tflops"],
                             "frac_sustained": fwd_tf / peaks["bf16_tflops_sustained"], "unit": "TFLOP/s"},
        "stage_ms_per_launch": {k: round(v, 5) for k, v in per.items()},
        "stage_tflops": {k: chunk * STAGE_FLOP[k] / (per[k] * 1e-3) / 1e12 for k in per if k in STAGE_FLOP},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(x_host.numel()),
                "d2h_bytes_per_step": int(out_host.numel() * 4), "api": "bx_forward_onehot_host",
                "matches_device_path": e2e_ok},
        "gpu_launches": int(launches),
        "clocks": sampler.summary(),
        "checksum": checksum,
    }
    if world == 1 and not args.no_transform:
        del x, out
        torch.cuda.empty_cache()
        line["transform_roofline"] = transform_rooflines(dev, peaks, cpu=not args.no_cpu_baseline)
    if not args.no_workflow:
        line["variant_workflow"] = wf if world > 1 else variant_workflow(eng, args.workflow_variants,
                                                                         cpu=not args.no_cpu_baseline)
    if not args.no_sharded:
        line["sharded_chromatin"] = sc if world > 1 else sharded_chromatin(eng, args.sharded_variants, 0, 1, barrier,
                                                                           max_over_ranks)
    if world == 1 and not args.no_cpu_baseline:
        rate, cores, threads, done, dt = cpu_forward_rate(32, args.cpu_seconds, 4096)
        line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{done} windows in {dt:.1f} s, batch 32 (reference default), fp32 torch CPU "
                                          f"forward restating Beluga.py:21-51, {threads} threads"}
        # the batch the reference's authors run with (scripts/sed_immune_atlas/chromatin_immune_atlas.sh:12)
        rate, cores, threads, done, dt = cpu_forward_rate(512, 4.0, 1024)
        line["cpu_baseline"]["batch_512"] = {"value": rate, "unit": UNIT, "cores": cores,
                                             "sample": f"{done} windows in {dt:.1f} s, batch 512, {threads} threads"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=8192, help="windows per GPU per step (BASELINE configs[1]: 64..8192)")
    ap.add_argument("--capacity", type=int, default=4096, help="windows per forward chunk (activation scratch)")
    ap.add_argument("--dtype", default="f16", choices=["f16", "bf16", "f16x3"],
                    help="tensor-core operand mode; f16x3 = split fp16 operands (3 MMAs per product, ~1e-4 of fp32)")
    ap.add_argument("--cta_group", type=int, default=2, choices=[1, 2])
    ap.add_argument("--ref_batch", type=int, default=32)
    ap.add_argument("--cpu_seconds", type=float, default=12.0)
    ap.add_argument("--no_cpu_baseline", action="store_true")
    ap.add_argument("--no_transform", action="store_true", help="skip the kernel-4 GB/s measurement")
    ap.add_argument("--no_workflow", action="store_true", help="skip the variant-workflow (segment mode) measurement")
    ap.add_argument("--workflow_variants", type=int, default=8192)
    ap.add_argument("--no_sharded", action="store_true", help="skip the sharded-pipeline (strong scaling) measurement")
    ap.add_argument("--sharded_variants", type=int, default=65536,
                    help="variants of the one VCF that the sharded-pipeline measurement splits over the GPUs")
    args = ap.parse_args()
    return run_reference(args) if args.impl == "reference" else run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
