"""Test-only stand-in for pyfasta==0.5.2 (requirements.txt:20), used to run the UNMODIFIED reference scripts.
``Fasta(path).sequence({'chr','start','stop'})`` is 1-based inclusive and case preserving; the reference never
passes ``strand`` (call sites chromatin.py:205, compute_expecto_features.py:108)."""


class Fasta:
    def __init__(self, path):
        self.seqs = {}
        name, chunks = None, []
        with open(path) as f:
            for line in f:
                if line.startswith(">"):
                    if name is not None:
                        self.seqs[name] = "".join(chunks)
                    name, chunks = line[1:].split()[0], []
                else:
                    chunks.append(line.strip())
        if name is not None:
            self.seqs[name] = "".join(chunks)

    def sequence(self, f):
        return self.seqs[f["chr"]][f["start"] - 1:f["stop"]]
