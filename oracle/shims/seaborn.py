"""Test-only stand-in for seaborn (imported by geuvadis_predict_consensus.py:19, used only by its unused plot helper)."""
