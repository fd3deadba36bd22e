#!/usr/bin/env python
"""Smallest run that touches every kernel (for compute-sanitizer): 2 variants through the per-window path, the
segment path (3 shifts), the output / transform / scoring kernels."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from expecto_b200 import synth
    from expecto_b200.engine import BelugaEngine, chromatin_outputs
    from expecto_b200.genome import PackedGenome
    from expecto_b200.transform import (decay_score_variant, decay_transform_fixed, decay_transform_variant,
                                        strand_average)
    from expecto_b200.windows import chromatin_segment_descs, chromatin_shift_list, chromatin_window_descs

    genome = synth.make_genome([("chr1", 30000)], seed=11)
    pg = PackedGenome.from_sequences(genome)
    eng = BelugaEngine(0, dtype="f16", cta_group=2, capacity=32)
    eng.load_state_dict(synth.make_beluga_weights(0, round_dtype="f16"))
    eng.set_genome(pg)
    chroms, pos, refs, alts = ["chr1", "chr1"], np.array([15000, 9000]), ["A", "AC"], ["C", "A"]
    sd, codes, k_of = chromatin_segment_descs(pg, chroms, pos, refs, alts, 200)
    seg = eng.forward_segments(sd, codes, 2, 4, 2400, k_of, shift_major=True)
    ok = True
    stack = []
    for si, s in enumerate(chromatin_shift_list(200)):
        wd, wc = chromatin_window_descs(pg, chroms, pos, refs, alts, s)
        win = eng.forward_windows(wd, wc)
        ok &= bool(torch.equal(seg[si].reshape(8, 2002), win))
        ref, alt, diff = chromatin_outputs(win, 2)
        stack.append(strand_average(diff))
        eng.encode_windows(wd, wc)
    x = torch.stack(stack)
    decay_transform_variant(x, np.array([100.0, -300.0]), np.array(["+", "-"]), shifts=chromatin_shift_list(200))
    decay_score_variant(x, np.array([100.0, -300.0]), np.array(["+", "-"]), np.zeros((3, 20020), np.float32),
                        np.zeros(3, np.float32), shifts=chromatin_shift_list(200))
    decay_transform_fixed(torch.rand((2, 5, 2, 2002), device="cuda"), np.ones((10, 5)))
    eng.forward_onehot_host(synth.random_onehot(3, seed=1))
    # round 2: the in-library variant loop (device descriptor builders, delta + full-segment + per-window paths, the
    # fused FC2 epilogue in both output forms) and the tensor-core scoring kernel
    from expecto_b200.transform import ScoreModels
    from expecto_b200.windows import variant_columns
    g = genome["chr1"]
    rows = [("chr1", 15000, "A", "C"), ("chr1", 9000, "AC", "A"), ("chr1", 12000, g[11999:12130].upper().replace("N", "A"), "A")]
    cols = variant_columns(pg, [r[0] for r in rows], np.array([r[1] for r in rows]), [r[2] for r in rows], [r[3] for r in rows])
    for average in (False, True):
        rows_per_shift = 3 if average else 6
        outs = [torch.zeros((9, rows_per_shift, 2002), device="cuda") for _ in range(3)]
        n_chunks, _ = eng.chromatin_open(cols, 800, average=average, chunk_variants=2)
        for k in range(n_chunks):
            eng.chromatin_submit_device(k, *[t.view(-1, 2002) for t in outs], rows_per_shift, 3,
                                        eng.chromatin_chunk_info(k)["v0"])
        eng.chromatin_close()
    x9 = outs[0]
    models = ScoreModels(np.full((3, 20020), 1e-3, np.float32), np.zeros(3, np.float32), "cuda:0")
    decay_score_variant(x9, np.array([100.0, -300.0, 0.0]), np.array(["+", "-", "+"]), models)
    torch.cuda.synchronize()
    print("sanitize_small ok" if ok else "sanitize_small MISMATCH")
    eng.close()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
