"""Test-only stand-in for Biopython (geuvadis_predict_consensus.py:14): only ``SeqIO.parse(path, 'fasta')``."""
