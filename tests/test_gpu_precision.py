"""Precision of the tcgen05 forward against the fp32 oracle at sizes beyond the golden fixtures (VERDICT r01 item 1,
ADVICE r01: weight rounding), and the split-operand mode (BX_DTYPE_F16X3) that closes the gap.

Bars (BASELINE.json north_star): probabilities within 2e-3 absolute of the reference fp32 forward (Beluga.py:21-51);
Pearson >= 0.999 on the ref-alt diffs (chromatin.py:281).  tools/parity_at_scale.py runs the same comparisons at
BASELINE sizes (2048 windows, 1000 variants x 9 shifts, 16 genes) and its numbers are under profiles/.
"""
import numpy as np
import pytest
import torch

from expecto_b200 import synth
from expecto_b200.genome import PackedGenome
from oracle import beluga_oracle as O
from oracle import golden_inputs as GI
from helpers import pearson

pytestmark = pytest.mark.gpu


def _windows(n, seed):
    """Random one-hot windows; every 4th has runs of all-zero (N) columns, window 1 is all N."""
    x = synth.random_onehot(n, seed=seed)
    rng = np.random.default_rng(seed + 1)
    for i in range(0, n, 4):
        for _ in range(int(rng.integers(1, 8))):
            at, run = int(rng.integers(0, 1900)), int(rng.integers(1, 100))
            x[i, :, at:at + run] = 0
    x[1] = 0
    return x


def _oracle(x, sd, batch=64):
    sdt = {k: torch.from_numpy(v) for k, v in sd.items()}
    return np.vstack([O.beluga_forward(x[i:i + batch].astype(np.float32), sdt) for i in range(0, len(x), batch)])


@pytest.fixture(scope="module")
def weights_fp32():
    """The calibrated weights NOT rounded to the operand type: what a user's fp32 deepsea.beluga.pth looks like."""
    return synth.make_beluga_weights(GI.WEIGHT_SEED, round_dtype=None)


@pytest.fixture(scope="module")
def engine_x3(weights_fp32):
    from expecto_b200.engine import BelugaEngine
    eng = BelugaEngine(0, dtype="f16x3", cta_group=2, capacity=256)
    eng.load_state_dict(weights_fp32)
    yield eng
    eng.close()


def test_f16_forward_512_windows_incl_n_rich_within_bar(engine, weights_f16):
    """fp16 operands, same (pre-rounded) weights on both sides: 512 windows x 2002 outputs under the 2e-3 bar, with the
    error distribution the design quotes (p99.99 around 1.2e-3)."""
    x = _windows(512, seed=21)
    got = engine.forward_onehot(torch.from_numpy(x).cuda()).cpu().numpy()
    want = _oracle(x, weights_f16)
    d = np.abs(got - want)
    assert d.max() <= 2e-3, d.max()
    assert np.quantile(d, 0.9999) <= 1.5e-3 and d.mean() <= 2e-4
    assert np.abs(got[1] - want[1]).max() <= 2e-3          # all-N window: biases only


def test_f16x3_matches_fp32_oracle_with_unrounded_weights(engine_x3, weights_fp32):
    """Split operands (Ah*Wh + Ah*Wl + Al*Wh): the fp32 reference with its fp32 weights to ~1.5e-4 at worst (first
    B200 run: max 1.45e-4 on 96 windows; what remains is the tensor core's truncating fp32 accumulation over K up to
    5120 x 3, not operand rounding) — an order of magnitude inside the bar, where plain fp16 operands with engine-side
    weight rounding reach 2.3e-3 on 2048 windows (profiles/r02_parity_at_scale_f16.json)."""
    x = _windows(96, seed=22)
    got = engine_x3.forward_onehot(torch.from_numpy(x).cuda()).cpu().numpy()
    want = _oracle(x, weights_fp32)
    d = np.abs(got - want)
    assert d.max() <= 4e-4, d.max()
    assert d.mean() <= 3e-5, d.mean()


def test_f16x3_is_deterministic_across_cta_groups_and_batching(engine_x3, weights_fp32):
    from expecto_b200.engine import BelugaEngine
    x = torch.from_numpy(_windows(300, seed=23)).cuda()   # 300 > capacity 256: two chunks
    a = engine_x3.forward_onehot(x)
    b = engine_x3.forward_onehot(x[37:113])
    assert torch.equal(a[37:113], b)
    eng1 = BelugaEngine(0, dtype="f16x3", cta_group=1, capacity=128)
    eng1.load_state_dict(weights_fp32)
    c = eng1.forward_onehot(x[:130])
    eng1.close()
    assert torch.equal(a[:130], c)


def test_f16x3_segment_and_delta_modes_equal_windows_bitwise(engine_x3, small_genome):
    """The shared-activation paths (segments, delta patches, FC1 slice reuse, pooled-pair phases) keep their
    bit-identity to the per-window path when every activation is an fp16 (hi, lo) pair."""
    from expecto_b200.windows import (chromatin_delta_descs, chromatin_segment_descs, chromatin_shift_list,
                                      chromatin_window_descs, delta_eligible, segment_eligible, tss_descs,
                                      tss_segment_descs)
    pg = PackedGenome.from_sequences(small_genome)
    engine_x3.set_genome(pg)
    rows = GI.edge_variants(small_genome)
    ok = segment_eligible([len(r[3]) for r in rows], [len(r[4]) for r in rows])
    rows = [r for r, e in zip(rows, ok) if e]
    sub = delta_eligible([len(r[3]) for r in rows], [len(r[4]) for r in rows])
    for kind, part in (("full", [r for r, s in zip(rows, sub) if not s][:6]), ("delta", [r for r, s in zip(rows, sub) if s])):
        chroms, pos = [r[0] for r in part], np.array([r[1] for r in part])
        refs, alts = [r[3] for r in part], [r[4] for r in part]
        n = len(part)
        assert n > 0
        if kind == "delta":
            rd, pd, codes, k_of, row0, plen = chromatin_delta_descs(pg, chroms, pos, refs, alts, 800)
            seg = engine_x3.forward_segments_delta(rd, pd, codes, n, 3600, plen, row0, k_of, shift_major=True)
        else:
            sd, codes, k_of = chromatin_segment_descs(pg, chroms, pos, refs, alts, 800)
            seg = engine_x3.forward_segments(sd, codes, n, 4, 3600, k_of, shift_major=True)
        for si, shift in enumerate(chromatin_shift_list(800)):
            wd, wcodes = chromatin_window_descs(pg, chroms, pos, refs, alts, shift)
            win = engine_x3.forward_windows(wd, wcodes).reshape(n, 4, 2002)
            assert torch.equal(seg[si], win), (kind, shift)
    shifts = np.arange(-1000, 1000, 200)
    chroms, tss, strand = ["chr1", "chr2"], [30000, 20000], [1, -1]
    sd, k_of, seg_len = tss_segment_descs(pg, chroms, tss, strand, shifts)
    seg = engine_x3.forward_segments(sd, np.zeros(1, np.uint8), 2, 2, seg_len, k_of, shift_major=False)
    wd = tss_descs(pg, chroms, tss, strand, shifts)
    win = engine_x3.forward_windows(wd, np.zeros(1, np.uint8)).reshape(2, len(shifts), 2, 2002)
    assert torch.equal(seg, win)


def test_chromatin_200_variants_three_shifts_vs_oracle(engine, weights_f16, small_genome):
    """A cfg3-shaped sample (SNVs, 5 % ref mismatches, 5 % indels) through the drop-in's segment + delta mode against
    oracle.chromatin_shift for shifts 0 and +-800: probabilities within the bar, diff = alt - ref exactly, Pearson on
    the diffs >= 0.999 (measured 0.99996)."""
    from expecto_b200 import chromatin
    pg = PackedGenome.from_sequences(small_genome)
    engine.set_genome(pg)
    rows = synth.make_vcf_rows(small_genome, 200, seed=31, odd_names=False)
    chroms, pos = [r[0] for r in rows], np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    outs = chromatin.predict_all_shifts_segments_device(engine, pg, chroms, pos, refs, alts, 800)
    dg = O.DictGenome(small_genome)
    sdt = {k: torch.from_numpy(v) for k, v in weights_f16.items()}
    for shift in (0, -800, 800):
        ref, alt, diff = (t.cpu().numpy() for t in outs[shift])
        d_o, r_o, a_o, _ = O.chromatin_shift(dg, rows, sdt, shift, batchsize=64)
        assert np.abs(ref - r_o).max() <= 2e-3 and np.abs(alt - a_o).max() <= 2e-3
        assert np.array_equal(diff, alt - ref)
        assert pearson(diff, d_o) >= 0.999
