"""``SeqIO.parse(handle_or_path, 'fasta')`` -> records with ``.id`` (header up to the first whitespace) and ``.seq``
(the record's lines joined), as geuvadis_predict_consensus.py:143-145,175 uses them."""


class _Record:
    def __init__(self, rid, seq):
        self.id, self.seq = rid, seq


def parse(handle, fmt):
    assert fmt == "fasta"
    f = open(handle) if isinstance(handle, str) else handle
    rid, chunks = None, []
    try:
        for line in f:
            if line.startswith(">"):
                if rid is not None:
                    yield _Record(rid, "".join(chunks))
                rid, chunks = line[1:].split()[0], []
            elif rid is not None:
                chunks.append(line.strip())
        if rid is not None:
            yield _Record(rid, "".join(chunks))
    finally:
        if isinstance(handle, str):
            f.close()
