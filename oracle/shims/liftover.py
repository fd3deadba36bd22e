"""Test-only stand-in for liftover==1.1.11 (requirements.txt:11): identity mapping.  chromatin.py only calls it
under --hg38; compute_expecto_features.py:45 builds a converter unconditionally."""


class _Identity:
    def convert_coordinate(self, chrom, pos):
        return [(chrom, pos, "+")]


def get_lifter(a, b):
    return _Identity()
