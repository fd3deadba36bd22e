"""expecto_b200 — B200-native (sm_100a) implementation of ExPecto's Beluga chromatin-effect hot path.

Scope: window extraction + one-hot/reverse-complement encoding from a 2-bit packed
genome, the Beluga CNN forward as tcgen05/TMEM implicit GEMMs, ref/alt diff outputs, and the
exponential-decay spatial transforms, behind the reference's chromatin.py /
compute_expecto_features.py / predict.py interfaces.  No CPU fallback, no Triton, no torch.compile.
"""
__version__ = "0.1.0"
