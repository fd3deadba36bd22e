"""The variant loop inside the library (bx_chromatin_open / _submit / _wait / _submit_device): device-side descriptor
builders against their host oracle (expecto_b200/windows.py, itself pinned to the reference's fetchSeqs + encodeSeqs),
and the fused FC2 epilogue (ref / alt / diff = alt - ref layout, strand average) against the per-window forward +
chromatin_outputs kernels — bit for bit."""
import numpy as np
import pytest
import torch

from expecto_b200 import chromatin
from expecto_b200.engine import chromatin_outputs, variant_descs_host
from expecto_b200.genome import PackedGenome
from expecto_b200.windows import (chromatin_delta_descs, chromatin_segment_descs, chromatin_shift_list,
                                  chromatin_window_descs, delta_eligible, segment_eligible, variant_columns)
from oracle import golden_inputs as GI

pytestmark = pytest.mark.gpu
FIELDS = ("g1", "g2", "len1", "alen", "len2", "aoff", "flags")


def _variants(small_genome, n_random=60, seed=5):
    """Edge variants (indels up to 150 bp, 120-bp deletion, mismatches, N / lower-case alleles) + random SNVs, MNVs
    and indels, in one table."""
    rows = [list(r) for r in GI.edge_variants(small_genome)]
    rng = np.random.default_rng(seed)
    names = sorted(small_genome)
    for i in range(n_random):
        c = names[i % len(names)]
        p = int(rng.integers(3000, len(small_genome[c]) - 3000))
        kind = i % 5
        g = "".join(b if b in "ACGT" else "A" for b in small_genome[c][p - 1:p + 30].upper())
        if kind < 2:
            ref, alt = g[0], "ACGT"[("ACGT".index(g[0]) + 1 + i % 3) % 4]
        elif kind == 2:
            ln = int(rng.integers(2, 17))
            ref, alt = g[:ln], g[:ln][::-1]
        elif kind == 3:
            ref, alt = g[:int(rng.integers(2, 25))], g[0]
        else:
            ref, alt = g[0], g[0] + "".join("ACGT"[j] for j in rng.integers(0, 4, int(rng.integers(1, 40))))
        rows.append([c, p, f"r{i}", ref, alt])
    return rows


def _cols(pg, rows):
    return variant_columns(pg, [r[0] for r in rows], np.array([r[1] for r in rows]), [r[3] for r in rows],
                           [r[4] for r in rows])


def _same(a, b, what):
    for f in FIELDS:
        assert np.array_equal(a[f], b[f]), (what, f, np.flatnonzero(a[f] != b[f])[:5])


def test_device_descriptor_builders_match_windows_py(small_genome):
    pg = PackedGenome.from_sequences(small_genome)
    rows = _variants(small_genome)
    chroms, pos = [r[0] for r in rows], np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    cols = _cols(pg, rows)
    # windows: every shift of the default range plus the ones where the variant leaves the 2100-bp fetch
    for shift in chromatin_shift_list(800) + GI.EXTREME_SHIFTS:
        want, codes = chromatin_window_descs(pg, chroms, pos, refs, alts, shift)
        assert np.array_equal(codes, cols["alleles"])
        _same(variant_descs_host(2, cols, shift=shift).reshape(-1), want, f"windows shift {shift}")
    for maxshift in (800, 400, 1000, 0):
        ok = segment_eligible(cols["ref_len"], cols["alt_len"], maxshift)
        part = [r for r, e in zip(rows, ok) if e]
        want, _, _ = chromatin_segment_descs(pg, [r[0] for r in part], np.array([r[1] for r in part]),
                                             [r[3] for r in part], [r[4] for r in part], maxshift)
        _same(variant_descs_host(0, _cols(pg, part), maxshift=maxshift).reshape(-1), want, f"segments {maxshift}")
        if maxshift == 0:
            continue
        sub = [r for r in part if delta_eligible([len(r[3])], [len(r[4])])[0]]
        rd, pd, _, _, row0, plen = chromatin_delta_descs(pg, [r[0] for r in sub], np.array([r[1] for r in sub]),
                                                         [r[3] for r in sub], [r[4] for r in sub], maxshift)
        g_rd, g_pd, g_row0, g_plen = variant_descs_host(1, _cols(pg, sub), maxshift=maxshift)
        assert (g_row0, g_plen) == (row0, plen)
        _same(g_rd.reshape(-1), rd, f"delta ref {maxshift}")
        _same(g_pd.reshape(-1), pd, f"delta patches {maxshift}")


def _reference_outputs(engine, pg, rows, maxshift, average):
    """Per-window forward + chromatin_outputs kernels, shift by shift: [S, 2n (or n), 2002] x (ref, alt, diff)."""
    chroms, pos = [r[0] for r in rows], np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    outs = []
    for shift in chromatin.shift_list(maxshift):
        wd, codes = chromatin_window_descs(pg, chroms, pos, refs, alts, shift)
        outs.append(chromatin_outputs(engine.forward_windows(wd, codes), len(rows), average=average))
    return [torch.stack([o[j] for o in outs]) for j in range(3)]


@pytest.mark.parametrize("average", [False, True])
def test_pipeline_device_outputs_equal_per_window_path(engine, small_genome, average):
    """All three paths in one table (substitutions -> delta, short indels -> full segments, long indels -> windows),
    chunks of 7 variants so groups straddle chunks, results left in HBM."""
    pg = PackedGenome.from_sequences(small_genome)
    engine.set_genome(pg)
    rows = _variants(small_genome, n_random=25)
    n = len(rows)
    want = _reference_outputs(engine, pg, rows, 800, average)
    cols = _cols(pg, rows)
    S = 9
    rows_per_shift = n if average else 2 * n
    got = [torch.full((S, rows_per_shift, 2002), float("nan"), device=engine.device) for _ in range(3)]
    n_chunks, chunk = engine.chromatin_open(cols, 800, average=average, chunk_variants=7)
    assert chunk == 7 and n_chunks == (n + 6) // 7
    seen = {"delta": 0, "full": 0, "window": 0}
    for k in range(n_chunks):
        info = engine.chromatin_chunk_info(k)
        for key in seen:
            seen[key] += info[key]
        engine.chromatin_submit_device(k, *[t.view(-1, 2002) for t in got], rows_per_shift, n, info["v0"])
    engine.chromatin_close()
    torch.cuda.synchronize()
    assert min(seen.values()) > 0 and sum(seen.values()) == n, seen
    for name, g, w in zip(("ref", "alt", "diff"), got, want):
        assert torch.equal(g, w), (name, average)
    if not average:
        assert torch.equal(got[2], got[1] - got[0])           # diff == alt - ref exactly (chromatin.py:281)


def test_pipeline_host_chunks_and_flags(engine, small_genome):
    """Host-buffer mode with two chunks in flight, BX_CHROM_NO_DELTA and BX_CHROM_WINDOWS: same bits."""
    pg = PackedGenome.from_sequences(small_genome)
    engine.set_genome(pg)
    rows = _variants(small_genome, n_random=20)
    n = len(rows)
    want = [t.cpu() for t in _reference_outputs(engine, pg, rows, 400, False)]
    cols = _cols(pg, rows)
    S = 5
    for kw in ({}, {"delta": False}, {"windows_only": True}):
        n_chunks, chunk = engine.chromatin_open(cols, 400, chunk_variants=9, **kw)
        bufs = [[torch.empty((S, 2, chunk, 2002), dtype=torch.float32, pin_memory=True) for _ in range(3)] for _ in range(2)]
        got = [torch.zeros((S, 2 * n, 2002)) for _ in range(3)]

        def collect(k):
            engine.chromatin_wait(k)
            info = engine.chromatin_chunk_info(k)
            v0, m = info["v0"], info["m"]
            for j in range(3):
                dense = bufs[k & 1][j].view(-1)[:S * 2 * m * 2002].view(S, 2, m, 2002)
                got[j][:, v0:v0 + m] = dense[:, 0]
                got[j][:, n + v0:n + v0 + m] = dense[:, 1]

        for k in range(n_chunks):
            m = engine.chromatin_chunk_info(k)["m"]
            engine.chromatin_submit(k, *[b.view(-1)[:S * 2 * m * 2002].view(S, 2, m, 2002) for b in bufs[k & 1]])
            if k >= 1:
                collect(k - 1)
        collect(n_chunks - 1)
        engine.chromatin_close()
        for name, g, w in zip(("ref", "alt", "diff"), got, want):
            assert torch.equal(g, w), (name, kw)


def test_pipeline_odd_maxshift_and_empty_input(engine, small_genome):
    """maxshift 300 (chromatin.py:243's range() then yields shifts 0, -200, 200: per-window path) and zero variants."""
    pg = PackedGenome.from_sequences(small_genome)
    engine.set_genome(pg)
    rows = _variants(small_genome, n_random=5)[-8:]
    want = _reference_outputs(engine, pg, rows, 300, False)
    n = len(rows)
    got = [torch.empty((3, 2 * n, 2002), device=engine.device) for _ in range(3)]
    n_chunks, _ = engine.chromatin_open(_cols(pg, rows), 300)
    for k in range(n_chunks):
        info = engine.chromatin_chunk_info(k)
        assert info["window"] == info["m"]
        engine.chromatin_submit_device(k, *[t.view(-1, 2002) for t in got], 2 * n, n, info["v0"])
    engine.chromatin_close()
    for g, w in zip(got, want):
        assert torch.equal(g, w)
    assert engine.chromatin_open(_cols(pg, []), 800) [0] == 0
    engine.chromatin_close()


def test_cli_sinks_device_descriptors_equal_host_descriptors(engine, small_genome):
    """predict_all_shifts_segments (the drop-in's default path) with the loop inside the library vs the windows.py
    builder + Python chunk loop: every slab identical."""
    pg = PackedGenome.from_sequences(small_genome)
    engine.set_genome(pg)
    rows = _variants(small_genome, n_random=30)
    chroms, pos = [r[0] for r in rows], np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    n = len(rows)
    res = {}
    for host in (False, True):
        store = {s: {k: np.full((2 * n, 2002), np.nan, np.float32) for k in ("diff", "ref", "alt")}
                 for s in chromatin.shift_list(800)}

        def make(s):
            def sink(name, row0, arr):
                store[s][name][row0:row0 + len(arr)] = arr
            return sink
        info = chromatin.predict_all_shifts_segments(engine, pg, chroms, pos, refs, alts, 800,
                                                     {s: make(s) for s in store}, host_descs=host)
        res[host] = (store, info)
    assert res[False][1]["descriptors"] == "device" and res[False][1]["window_variants"] == res[True][1]["window_variants"]
    assert res[False][1]["delta_variants"] == res[True][1]["delta_variants"] > 0
    for s in res[False][0]:
        for k in ("diff", "ref", "alt"):
            assert np.array_equal(res[False][0][s][k], res[True][0][s][k]), (s, k)
            assert not np.isnan(res[False][0][s][k]).any()


@pytest.mark.parametrize("average", [False, True])
def test_fused_epilogue_writes_only_its_own_rows(engine, small_genome, average):
    """compute-sanitizer is closed on this pool (profiles/r02_sanitizer_memcheck.log), so the FC2 chromatin epilogue is
    checked with guard rows: the chunk's rows sit inside larger tensors (row0 > 0, strides wider than the chunk,
    spare rows after every strand block) pre-filled with a sentinel; every row the call does not own must come back
    intact and every owned row must be written — for all three paths and a partial last chunk."""
    pg = PackedGenome.from_sequences(small_genome)
    engine.set_genome(pg)
    rows = _variants(small_genome, n_random=12)
    n = len(rows)
    SENT = -5.5
    S, pad0, gap = 9, 3, 2
    strand_stride = n + gap
    shift_stride = (1 if average else 2) * strand_stride + gap
    total = pad0 + S * shift_stride + 4
    outs = [torch.full((total, 2002), SENT, device=engine.device) for _ in range(3)]
    n_chunks, _ = engine.chromatin_open(_cols(pg, rows), 800, average=average, chunk_variants=5)
    for k in range(n_chunks):
        engine.chromatin_submit_device(k, *outs, shift_stride, strand_stride, pad0 + engine.chromatin_chunk_info(k)["v0"])
    engine.chromatin_close()
    torch.cuda.synchronize()
    owned = torch.zeros(total, dtype=torch.bool, device=engine.device)
    for si in range(S):
        for t in range(1 if average else 2):
            a = pad0 + si * shift_stride + t * strand_stride
            owned[a:a + n] = True
    for name, o in zip(("ref", "alt", "diff"), outs):
        assert bool((o[~owned] == SENT).all()), name
        assert not bool((o[owned] == SENT).any()), name
        if name != "diff":
            assert bool(((o[owned] > 0) & (o[owned] < 1)).all()), name


@pytest.mark.parametrize("average", [False, True])
def test_device_outputs_with_padded_rows_equal_dense_outputs(engine, small_genome, average):
    """bx_chromatin_set_row_pitch: the FC2 epilogue writing rows 2016 floats apart (what predict keeps for the scoring
    kernel) gives the dense call's values bit for bit; the padding columns keep their sentinel, except that averaged
    plans zero whole rows before delta mode's alt rows add their halves (alt / diff padding reads 0 or the sentinel)."""
    from expecto_b200.transform import X_PITCH
    pg = PackedGenome.from_sequences(small_genome)
    engine.set_genome(pg)
    rows = _variants(small_genome, n_random=12)
    n = len(rows)
    S, SENT = 9, -5.5
    rows_per_shift = n if average else 2 * n
    res = {}
    for pitch in (2002, X_PITCH):
        bufs = [torch.full((S * rows_per_shift, pitch), SENT, device=engine.device) for _ in range(3)]
        outs = [b[:, :2002].view(S, rows_per_shift, 2002) if pitch == 2002 else b[:, :2002].unflatten(0, (S, rows_per_shift))
                for b in bufs]
        n_chunks, _ = engine.chromatin_open(_cols(pg, rows), 800, average=average, chunk_variants=7)
        for k in range(n_chunks):
            engine.chromatin_submit_device(k, *outs, rows_per_shift, n, engine.chromatin_chunk_info(k)["v0"])
        engine.chromatin_close()
        torch.cuda.synchronize()
        res[pitch] = (bufs, outs)
    for name, d, p_, pb in zip(("ref", "alt", "diff"), res[2002][1], res[X_PITCH][1], res[X_PITCH][0]):
        assert not bool((d == SENT).any()) and torch.equal(d, p_), name
        tail = pb[:, 2002:]
        assert bool(((tail == SENT) | ((tail == 0) if (average and name != "ref") else False)).all()), name
    # a later dense call on the same context is dense again (the pitch belongs to the plan)
    dense2 = [torch.empty((S, rows_per_shift, 2002), device=engine.device) for _ in range(3)]
    n_chunks, _ = engine.chromatin_open(_cols(pg, rows), 800, average=average)
    for k in range(n_chunks):
        engine.chromatin_submit_device(k, *dense2, rows_per_shift, n, engine.chromatin_chunk_info(k)["v0"])
    engine.chromatin_close()
    for a, b in zip(dense2, res[2002][1]):
        assert torch.equal(a, b)
    with pytest.raises(ValueError):
        engine.chromatin_submit_device(0, res[2002][1][0], res[X_PITCH][1][1], res[X_PITCH][1][2], rows_per_shift, n, 0)


def test_set_row_pitch_argument_and_state_errors(engine, small_genome):
    """bx_chromatin_set_row_pitch: refused without an open plan, for odd pitches and for pitches below 2002; a refused
    call leaves the plan's pitch as it was (dense)."""
    from expecto_b200 import _lib
    L = _lib.lib()
    assert L.bx_chromatin_set_row_pitch(engine._ctx, 2016) == _lib.BX_E_STATE
    pg = PackedGenome.from_sequences(small_genome)
    engine.set_genome(pg)
    rows = _variants(small_genome, n_random=3)
    n = len(rows)
    n_chunks, _ = engine.chromatin_open(_cols(pg, rows), 800)
    for bad in (2001, 2003, 2000, 0, -8):
        assert L.bx_chromatin_set_row_pitch(engine._ctx, bad) == _lib.BX_E_ARG
        assert b"row_pitch" in L.bx_last_error()
    got = [torch.full((9, 2 * n, 2002), -5.5, device=engine.device) for _ in range(3)]
    for k in range(n_chunks):
        engine.chromatin_submit_device(k, *got, 2 * n, n, engine.chromatin_chunk_info(k)["v0"])
    engine.chromatin_close()
    torch.cuda.synchronize()
    for g, w in zip(got, _reference_outputs(engine, pg, rows, 800, False)):
        assert torch.equal(g, w)
