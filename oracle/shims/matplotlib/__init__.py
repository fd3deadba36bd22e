"""Test-only stub: predict.py:6,10 imports matplotlib but the scoring path never plots."""
