"""CPU restatement of the reference's algorithm for the Beluga chromatin-effect hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``expecto_b200/`` imports this module; it is used by
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs
as the checker and the timed CPU baseline, never as the product path.

Pinned against the reference itself: ``oracle/make_golden.py`` runs the UNMODIFIED reference code
(``/root/reference/expecto_utils.py``, ``Beluga.py``, and ``chromatin.py`` /
``compute_expecto_features.py`` as subprocesses behind import shims) on seeded synthetic inputs,
asserts these restatements agree, and commits the outputs under ``tests/golden/``.  The reference has
no tests and its own example fixtures need hg19 + trained weights that are not in the repository
(SURVEY §4), so parity with the *trained* model is unpinned; parity with the reference *code* is pinned.

Each function cites the reference lines it follows.
"""
from __future__ import annotations

import math

import numpy as np

INPUTSIZE = 2000
SHIFTS_CHROMATIN = lambda maxshift=800: [0] + list(range(-200, -maxshift - 1, -200)) + list(  # noqa: E731
    range(200, maxshift + 1, 200))  # chromatin.py:243 / predict.py:173
CHRS = ["chr%d" % i for i in range(1, 23)] + ["chrX", "chrY"]  # chromatin.py:108-110

_ROW = {"A": 0, "G": 1, "C": 2, "T": 3, "a": 0, "g": 1, "c": 2, "t": 3}  # chromatin.py:155-160
_ZERO = set("NHn-")


def encode_seqs(seqs, inputsize=INPUTSIZE, concat_rc=True):
    """encodeSeqs, chromatin.py:138-172 == expecto_utils.py:5-39 (``concat_rc=False`` is the variant at
    compute_expecto_features.py:184-218).  bool [2n (or n), 4, inputsize], channel order A,G,C,T."""
    out = np.zeros((len(seqs), 4, inputsize), np.bool_)
    for n, line in enumerate(seqs):
        lo = int(math.floor((len(line) - inputsize) / 2.0))
        hi = int(math.floor(len(line) - (len(line) - inputsize) / 2.0))
        cline = line[lo:hi]
        for i, c in enumerate(cline):
            if c in _ROW:
                out[n, _ROW[c], i] = True
            elif c not in _ZERO:
                raise KeyError(c)  # mydict[c], chromatin.py:166
    if concat_rc:
        out = np.concatenate([out, out[:, ::-1, ::-1]], axis=0)  # chromatin.py:170-171
    return out


class DictGenome:
    """pyfasta.Fasta stand-in over in-memory strings: ``sequence({'chr','start','stop'})`` is 1-based
    inclusive and case preserving (pyfasta 0.5.2, call sites chromatin.py:205, compute_expecto_features.py:108)."""

    def __init__(self, seqs):
        self.seqs = seqs

    def sequence(self, f):
        return self.seqs[f["chr"]][f["start"] - 1:f["stop"]]


def fetch_seqs(genome, chr, pos, ref, alt, shift=0, inputsize=INPUTSIZE):
    """fetchSeqs, chromatin.py:175-209."""
    windowsize = inputsize + 100
    mutpos = int(windowsize / 2 - 1 - shift)
    seq = genome.sequence({"chr": chr, "start": pos + shift - int(windowsize / 2 - 1),
                           "stop": pos + shift + int(windowsize / 2)})
    ref_matched = seq[mutpos:(mutpos + len(ref))].upper() == ref.upper()
    alt_matched = seq[mutpos:(mutpos + len(ref))].upper() == alt.upper()
    return (seq[:mutpos] + ref + seq[(mutpos + len(ref)):], seq[:mutpos] + alt + seq[(mutpos + len(ref)):],
            ref_matched, alt_matched)


def normalise_vcf_rows(rows):
    """chromatin.py:240-241: chrom <- 'chr' + str(chrom).replace('chr',''), keep chr1-22,X,Y."""
    out = []
    for r in rows:
        c = "chr" + str(r[0]).replace("chr", "")
        if c in CHRS:
            out.append([c] + list(r[1:]))
    return out


def beluga_forward(x, sd):
    """Beluga.forward, Beluga.py:21-51, as fp32 torch functional ops on CPU.  ``x``: float32 [B,4,2000]
    (the reference's [B,4,1,2000] without the dummy axis); ``sd``: state_dict-keyed numpy/torch arrays."""
    import torch
    import torch.nn.functional as F

    t = lambda k: sd[k] if isinstance(sd[k], torch.Tensor) else torch.from_numpy(np.asarray(sd[k]))  # noqa: E731
    y = torch.as_tensor(x, dtype=torch.float32)
    convs = ["model.0.0", "model.0.2", "model.0.6", "model.0.8", "model.0.12", "model.0.14"]
    pool_after = {1, 3}  # MaxPool2d((1,4),(1,4)) after conv2 and conv4, Beluga.py:28,34 (Dropout = identity in eval)
    with torch.no_grad():
        for i, k in enumerate(convs):
            y = F.relu(F.conv1d(y, t(k + ".weight")[:, :, 0, :], t(k + ".bias")))
            if i in pool_after:
                y = F.max_pool1d(y, 4, 4)
        y = y.reshape(y.shape[0], -1)  # x.view(x.size(0), -1): channel-major flatten, Beluga.py:42
        y = F.relu(F.linear(y, t("model.1.2.1.weight"), t("model.1.2.1.bias")))
        y = torch.sigmoid(F.linear(y, t("model.1.4.1.weight"), t("model.1.4.1.bias")))
    return y.numpy()


def chromatin_shift(genome, rows, sd, shift, inputsize=INPUTSIZE, batchsize=32):
    """One iteration of the shift loop, chromatin.py:244-281: returns (diff, ref, alt, counters)."""
    refseqs, altseqs, rm, am = [], [], [], []
    for r in rows:
        a, b, c, d = fetch_seqs(genome, r[0], int(r[1]), r[3], r[4], shift=shift, inputsize=inputsize)
        refseqs.append(a), altseqs.append(b), rm.append(c), am.append(d)
    ref_enc = encode_seqs(refseqs, inputsize).astype(np.float32)
    alt_enc = encode_seqs(altseqs, inputsize).astype(np.float32)

    def run(enc):
        outs = [beluga_forward(enc[i:i + batchsize], sd) for i in range(0, enc.shape[0], batchsize)]
        return np.vstack(outs)

    ref_p, alt_p = run(ref_enc), run(alt_enc)
    return alt_p - ref_p, ref_p, alt_p, (int(np.sum(rm)), int(np.sum(am)), len(rm))


def fixed_pos_weights(shifts=None):
    """compute_expecto_features.py:88-101: float64 [10, S]."""
    s = np.array(list(range(-20000, 20000, 200))) if shifts is None else np.asarray(shifts)
    rates = [0.01, 0.02, 0.05, 0.1, 0.2]
    return np.vstack([np.exp(-k * np.abs(s) / 200) * (s <= 0) for k in rates] +
                     [np.exp(-k * np.abs(s) / 200) * (s >= 0) for k in rates])


def fixed_transform(pred_fwd, pred_rc, pos_weights):
    """compute_expecto_features.py:123-124: float64 [10*2002] from fp32 [S,2002] forward / rc predictions."""
    pred = 0.5 * (pred_fwd + pred_rc)
    return np.sum(pos_weights[:, :, None] * pred[None, :, :], axis=1).flatten()


def tss_windows(genome, chrom, tss, strand, shifts=None, windowsize=INPUTSIZE):
    """compute_expecto_features.py:106-111."""
    s = list(range(-20000, 20000, 200)) if shifts is None else shifts
    return [genome.sequence({"chr": chrom, "start": tss + (sh * strand) - int(windowsize / 2 - 1),
                             "stop": tss + (sh * strand) + int(windowsize / 2)}) for sh in s]


def variant_weights(snpdists, snpstrands, maxshift=800):
    """predict.py:87-109: list over shifts of float64 [n,10]."""
    sgn = (np.asarray(snpstrands) == "+") * 2 - 1
    d0 = np.asarray(snpdists) * sgn
    rates = [0.01, 0.02, 0.05, 0.1, 0.2]
    out = []
    for dist in SHIFTS_CHROMATIN(maxshift):
        p = d0 + dist * sgn
        w = [np.exp(-k * np.floor(np.abs(p / 200.0))) * (p <= 0) for k in rates] + \
            [np.exp(-k * np.floor(np.abs(p / 200.0))) * (p >= 0) for k in rates]
        out.append(np.vstack(w).T)
    return out


def variant_transform(x_by_shift, weights, nfeatures=2002):
    """predict.py:123-124: reduce(+, tile(X_j,10) * repeat(W_j, nfeatures)) -> float64 [n, 10*nfeatures]."""
    acc = None
    for xj, wj in zip(x_by_shift, weights):
        term = np.tile(np.asarray(xj), 10) * np.repeat(wj, nfeatures, axis=1)
        acc = term if acc is None else acc + term
    return acc


def strand_average(d):
    """predict.py:183-190: (top half + bottom half) / 2.0."""
    n = int(d.shape[0] / 2)
    return (np.asarray(d[0:n, :]) + np.asarray(d[n:2 * n, :])) / 2.0


def gblinear_predict(features, weights, bias, base_score=0.0):
    """xgboost 0.7 gblinear Booster.predict (predict.py:150-160; dump layout predict_by_cluster.py:73-75):
    base_score + bias + w . x  (xgboost==0.7.post4 is a pinned third-party dependency, requirements.txt:38)."""
    return base_score + bias + np.asarray(features) @ np.asarray(weights)


# ---------------------------------------------------------------------------------------------------------------
# consensus-sequence front end (geuvadis_predict_consensus.py) and per-gene prediction dumps
# (replicate_expecto_features.py).  Pinned by oracle/make_golden.py stage_consensus / stage_replicate.
ENFORMER_SEQ_LENGTH = 393216  # geuvadis_predict_consensus.py:23


def consensus_pad(record_id, seq):
    """gen_sample_seqs_and_id_for_gene, geuvadis_predict_consensus.py:143-167: upper-case; a record whose interval
    starts with ``-`` is missing its beginning (N-padded in front), otherwise a short record is N-padded at the end."""
    seq = seq.upper()
    interval = record_id.split(":")[1]
    if interval.startswith("-"):
        bp_start = -int(interval.split("-")[-2])
        bp_end = int(interval.split("-")[-1])
        assert bp_end - bp_start + 1 == ENFORMER_SEQ_LENGTH
        seq = "N" * (ENFORMER_SEQ_LENGTH - len(seq)) + seq
    else:
        bp_start, bp_end = map(int, interval.split("-"))
        assert bp_end - bp_start + 1 == ENFORMER_SEQ_LENGTH
        if len(seq) < ENFORMER_SEQ_LENGTH:
            seq = seq + "N" * (ENFORMER_SEQ_LENGTH - len(seq))
    assert len(seq) == ENFORMER_SEQ_LENGTH
    return seq


def consensus_seq_shifts(sample_seq, strand, shifts, windowsize=INPUTSIZE):
    """get_seq_shifts_for_sample_seq, geuvadis_predict_consensus.py:210-243: TSS at ``len // 2``; ``strand`` '+'/'-'."""
    tss_i = len(sample_seq) // 2
    sgn = {"+": 1, "-": -1}[strand]
    out = []
    for shift in shifts:
        seq = sample_seq[tss_i + (shift * sgn) - int(windowsize / 2 - 1): tss_i + (shift * sgn) + int(windowsize / 2) + 1]
        assert len(seq) == windowsize
        out.append(seq)
    return out


def consensus_chromatin(sample_seq, strand, sd, shifts, chunk=50):
    """geuvadis_predict_consensus.py:91-103 for one record: encodeSeqs (forward windows then their reverse
    complements), Beluga, the two halves averaged in float64 -> [len(shifts), 2002]."""
    enc = encode_seqs(consensus_seq_shifts(sample_seq, strand, shifts)).astype(np.float32)
    p = np.zeros((enc.shape[0], 2002))
    for i in range(0, enc.shape[0], chunk):
        p[i:i + chunk] = beluga_forward(enc[i:i + chunk], sd)
    return (p[:p.shape[0] // 2] + p[p.shape[0] // 2:]) / 2


def consensus_features(preds, shifts=None):
    """geuvadis_predict_consensus.py:108-124: float64 [n, 20030] (zero column in front of each basis function)."""
    pw = fixed_pos_weights(shifts)
    features = np.sum(pw[None, :, :, None] * np.asarray(preds)[:, None, :, :], axis=2)
    return np.concatenate([np.zeros((features.shape[0], 10, 1)), features], axis=2).reshape((-1, 20030))


def ref_all_genes_scores(records, strands, sd, weights, bias, base_score, shifts=None):
    """geuvadis_predict_ref_all_genes.py:61-102: one reference record per gene -> padded, 200 shifted windows, Beluga
    on the forward windows and their reverse complements, the halves averaged (:77), the legacy 20030-wide features
    (:80-96) and the gblinear score (:98); float32 [n_genes] as ``np.array(list).squeeze()`` leaves them (:100)."""
    shifts = list(range(-20000, 20000, 200)) if shifts is None else shifts
    out = []
    for (rid, seq), strand in zip(records, strands):
        preds = consensus_chromatin(consensus_pad(rid, seq), strand, sd, shifts)
        feats = consensus_features(preds[None], shifts)
        out.append(np.asarray(gblinear_predict(feats, weights, bias, base_score), dtype=np.float32))
    return np.array(out).squeeze()


def replicate_prediction(genome, chrom, tss, strand, sd, shifts=None, chunk=50):
    """replicate_expecto_features.py:65-83: float32 [S,2002] = 0.5 * (prediction + rc prediction)."""
    enc = encode_seqs(tss_windows(genome, chrom, tss, strand, shifts), concat_rc=False)
    fwd_in = enc.astype(np.float32)
    rc_in = np.ascontiguousarray(enc[:, ::-1, ::-1]).astype(np.float32)
    fwd = np.vstack([beluga_forward(fwd_in[i:i + chunk], sd) for i in range(0, len(fwd_in), chunk)])
    rc = np.vstack([beluga_forward(rc_in[i:i + chunk], sd) for i in range(0, len(rc_in), chunk)])
    return 0.5 * (fwd + rc)
