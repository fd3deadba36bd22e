#!/usr/bin/env python
"""cProfile of the host side of the segment-mode chromatin loop (no files): where the Python thread spends its time."""
import cProfile, os, pstats, sys, io
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from expecto_b200 import chromatin, synth
from expecto_b200.engine import BelugaEngine
from expecto_b200.genome import PackedGenome
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
contigs = [("chr1", 24_000_000), ("chr2", 20_000_000)]
genome = synth.make_genome(contigs, seed=3)
pg = PackedGenome.from_sequences(genome)
eng = BelugaEngine(0, dtype="f16", capacity=4096)
eng.load_state_dict(synth.make_beluga_weights(0, round_dtype="f16"))
eng.set_genome(pg)
rows = synth.make_vcf_rows(genome, n, seed=3, odd_names=False)
chroms = [r[0] for r in rows]; pos = np.array([r[1] for r in rows]); refs = [r[3] for r in rows]; alts = [r[4] for r in rows]
sinks = {s: (lambda *a: None) for s in chromatin.shift_list(800)}
chromatin.predict_all_shifts_segments(eng, pg, chroms[:1200], pos[:1200], refs[:1200], alts[:1200], 800, sinks)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
info = chromatin.predict_all_shifts_segments(eng, pg, chroms, pos, refs, alts, 800, sinks)
torch.cuda.synchronize()
pr.disable()
print(info)
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(22)
print(s.getvalue()[:5000])
