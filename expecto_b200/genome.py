"""2-bit packed genome (replaces ``pyfasta.Fasta('./resources/hg19.fa')``, chromatin.py:44).

Layout handed to the kernels (include/beluga_b200.h, bx_ctx_set_genome):

* ``bases2``  uint32, 16 bases per word, 2 bits each, code A=0 G=1 C=2 T=3 — the channel order of
  ``encodeSeqs``' ``mydict`` (chromatin.py:155-160), so reverse complement is ``3 - code``;
* ``nmask``   uint32, 1 bit per base; 1 = all-zero column (``N n H -`` in ``mydict``).

Contigs are concatenated in file order, each starting on a 32-base boundary after ``PAD`` bases of
N, so a window that runs off a contig reads zero columns.  ``PAD`` exceeds the farthest any window of the
path reaches from a coordinate inside its contig: compute_expecto_features.py:105-111 goes 20,000 + 1,000 bp
either side of a TSS, chromatin.py:202-206 at most maxshift + 1,050 bp plus the allele.  (The reference's
behaviour off a contig end is accidental — negative Python slice indices / short strings from pyfasta — and is not
reproduced; SURVEY §7 hard part 2.)  Characters outside ``mydict`` make the reference raise ``KeyError`` when a
window touches them; they are recorded in ``invalid`` and checked on the host before any launch.
"""
from __future__ import annotations

import os

import numpy as np

PAD = 24576  # >= 21,000 (TSS windows: +-20,000 shift +-1,000 half window) + slack, a multiple of 32
_CODE = np.full(256, 255, dtype=np.uint8)
for _i, _c in enumerate("AGCT"):
    _CODE[ord(_c)] = _i
    _CODE[ord(_c.lower())] = _i
for _c in "NnH-":
    _CODE[ord(_c)] = 4
_DECODE = np.frombuffer(b"AGCTN", dtype=np.uint8)


def encode_bases(seq: str | bytes) -> np.ndarray:
    """Allele / sequence string -> codes 0..3 (A,G,C,T), 4 (zero column); KeyError like ``mydict[c]``."""
    b = seq.encode("ascii", "replace") if isinstance(seq, str) else seq
    codes = _CODE[np.frombuffer(b, dtype=np.uint8)]
    bad = np.flatnonzero(codes == 255)
    if bad.size:
        raise KeyError(chr(b[int(bad[0])]))
    return codes


class PackedGenome:
    def __init__(self):
        self.names: list[str] = []
        self.offsets: dict[str, int] = {}   # global index of base 0 of each contig
        self.lengths: dict[str, int] = {}
        self.bases2 = np.zeros(0, dtype=np.uint32)
        self.nmask = np.zeros(0, dtype=np.uint32)
        self.n_bases = 0
        self.pad = PAD                               # N bases before / after every contig in this layout
        self.invalid = np.zeros(0, dtype=np.int64)  # sorted global indices of characters outside mydict

    # ------------------------------------------------------------------ construction
    @classmethod
    def from_sequences(cls, seqs: dict[str, str | bytes]) -> "PackedGenome":
        raws = {name: np.frombuffer(s.encode("ascii", "replace") if isinstance(s, str) else s, dtype=np.uint8)
                for name, s in seqs.items()}
        g = cls._allocate({name: r.size for name, r in raws.items()})
        invalid = []
        for name, raw in raws.items():
            g._pack_contig(g.offsets[name], raw, invalid)
        g.invalid = np.sort(np.concatenate(invalid)) if invalid else np.zeros(0, dtype=np.int64)
        return g

    @classmethod
    def _allocate(cls, lengths: dict[str, int], host_arrays: bool = True) -> "PackedGenome":
        """Offsets (PAD bases of N before/after every contig, contigs on 32-base boundaries) and all-N arrays."""
        g = cls()
        total = PAD
        for name, n in lengths.items():
            total = (total + 31) // 32 * 32
            g.names.append(name)
            g.offsets[name] = total
            g.lengths[name] = int(n)
            total += int(n) + PAD
        total = (total + 31) // 32 * 32
        g.n_bases = total
        if host_arrays:
            g.bases2 = np.zeros(total // 16, dtype=np.uint32)
            g.nmask = np.full(total // 32, 0xFFFFFFFF, dtype=np.uint32)  # padding is N
        return g

    _CHUNK = 1 << 26  # bases packed per pass (multiple of 32): bounds the temporaries at a few hundred MB

    def _pack_contig(self, off: int, raw: np.ndarray, invalid: list) -> None:
        """Pack one contig's characters at global index ``off`` (a multiple of 32).  Byte-wise bit packing: four
        2-bit codes per byte and eight mask bits per byte, both little-endian, so the byte stream viewed as
        little-endian uint32 is the word layout of the module docstring (base i of a word at bits 2*(i%16))."""
        for c0 in range(0, raw.size, self._CHUNK):
            piece = raw[c0:c0 + self._CHUNK]
            n = piece.size
            codes = _CODE[piece]
            bad = np.flatnonzero(codes == 255)
            if bad.size:
                invalid.append(bad.astype(np.int64) + (off + c0))
                codes[bad] = 4
            n32 = (n + 31) // 32 * 32
            if n32 != n:  # only the last piece of a contig: tail padded with N
                full = np.full(n32, 4, dtype=np.uint8)
                full[:n] = codes
            else:
                full = codes
            q = (full & 3).reshape(-1, 4)          # code 4 (zero column) packs as 0
            packed = q[:, 0] | (q[:, 1] << 2) | (q[:, 2] << 4) | (q[:, 3] << 6)
            w0 = (off + c0) // 16
            self.bases2[w0: w0 + n32 // 16] = np.ascontiguousarray(packed).view("<u4")
            m0 = (off + c0) // 32
            self.nmask[m0: m0 + n32 // 32] = np.packbits(full == 4, bitorder="little").view("<u4")

    @classmethod
    def from_fasta(cls, path: str, cache: bool = True, device=None) -> "PackedGenome":
        """Parse a (multi-)FASTA; a ``<path>.bx2.npz`` cache is written next to it when possible (atomically: a
        temporary file renamed into place, so concurrent readers never see a partial cache).

        ``device`` (a CUDA device): the file is copied to HBM and packed there by the ``bx_fasta_*`` kernels
        (csrc/genome_pack.cuh) — what the drop-in CLIs do; the packed planes stay on that device for
        ``BelugaEngine.set_genome``.  ``device=None`` packs with numpy on the host (tests, tools without a GPU); the
        two produce identical arrays (tests/test_gpu_parity.py)."""
        cpath = path + ".bx2.npz"
        if cache and os.path.exists(cpath) and os.path.getmtime(cpath) >= os.path.getmtime(path):
            g = cls.load(cpath)
            if g.pad >= PAD:   # caches written with a narrower inter-contig padding are rebuilt
                return g
        if device is not None:
            g = cls._from_fasta_device(path, device)
            if cache:
                g._save_cache(cpath)
            return g
        buf = np.fromfile(path, dtype=np.uint8)
        starts = np.flatnonzero(buf == ord(">"))
        # keep only '>' at line starts
        starts = starts[(starts == 0) | (buf[np.maximum(starts - 1, 0)] == 10)]
        bodies: dict[str, tuple[int, int]] = {}
        lengths: dict[str, int] = {}
        for i, s in enumerate(starts):
            e = int(starts[i + 1]) if i + 1 < len(starts) else buf.size
            nl = int(s) + int(np.argmax(buf[s:e] == 10)) if np.any(buf[s:min(e, s + 4096)] == 10) else e
            header = buf[s + 1:nl].tobytes().decode("ascii", "replace").strip()
            name = header.split()[0] if header else f"contig{i}"
            b0 = min(nl + 1, e)
            body = buf[b0:e]
            bodies[name] = (b0, e)
            lengths[name] = int(body.size - np.count_nonzero(body == 10) - np.count_nonzero(body == 13))
        g = cls._allocate(lengths)
        invalid = []
        for name, (b0, e) in bodies.items():  # one contig at a time: the stripped copy is the only large temporary
            body = buf[b0:e]
            g._pack_contig(g.offsets[name], body[(body != 10) & (body != 13)], invalid)
        g.invalid = np.sort(np.concatenate(invalid)) if invalid else np.zeros(0, dtype=np.int64)
        del buf
        if cache:
            g._save_cache(cpath)
        return g

    def _save_cache(self, cpath: str) -> None:
        tmp = f"{cpath}.{os.getpid()}.tmp.npz"
        try:
            self.save(tmp)
            os.replace(tmp, cpath)
        except OSError:
            try:
                os.unlink(tmp)
            except OSError:
                pass

    # ------------------------------------------------------------------ device-side packing
    @classmethod
    def _from_fasta_device(cls, path: str, device, chunk_bytes: int = 1 << 26, header_cap: int = 1 << 20,
                           invalid_cap: int = 1 << 22) -> "PackedGenome":
        """FASTA -> packed planes on ``device`` (include/beluga_b200.h: bx_fasta_find_headers / _count / _pack).
        The host only streams the file into HBM through two pinned buffers and reads the header lines."""
        import ctypes as C

        import torch

        from ._lib import check, lib

        L = lib()
        dev = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        if dev.type != "cuda":
            raise ValueError("device must be a CUDA device")
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        size = os.path.getsize(path)
        st = torch.cuda.current_stream(dev)
        with torch.cuda.device(dev):
            text = torch.empty((size + 15) // 16 * 16 + 16, dtype=torch.uint8, device=dev)
            text[size:] = 10  # the kernels load whole 16-byte vectors
            pinned = [torch.empty(chunk_bytes, dtype=torch.uint8, pin_memory=True) for _ in range(2)]
            done = [torch.cuda.Event(), torch.cuda.Event()]
            with open(path, "rb", buffering=0) as f:
                off, i = 0, 0
                while off < size:
                    done[i].synchronize()  # the copy that last used this buffer
                    n = f.readinto(memoryview(pinned[i].numpy())[:min(chunk_bytes, size - off)])
                    if not n:
                        raise IOError(f"{path}: short read at byte {off}")
                    text[off:off + n].copy_(pinned[i][:n], non_blocking=True)
                    done[i].record(st)
                    off, i = off + n, i ^ 1
            pos = torch.empty(header_cap, dtype=torch.int64, device=dev)
            cnt = torch.zeros(1, dtype=torch.int64, device=dev)
            check(L.bx_fasta_find_headers(text.data_ptr(), size, pos.data_ptr(), header_cap, cnt.data_ptr(), st.cuda_stream))
            n_hdr = int(cnt.item())
            if n_hdr > header_cap:
                raise ValueError(f"{path}: {n_hdr} contigs exceed the header buffer ({header_cap})")
            starts = np.sort(pos[:n_hdr].cpu().numpy())
            bodies, lengths = [], {}
            with open(path, "rb") as f:
                for k, s0 in enumerate(starts):
                    e = int(starts[k + 1]) if k + 1 < n_hdr else size
                    f.seek(int(s0))
                    line = f.readline()
                    header = line[1:].decode("ascii", "replace").strip()
                    name = header.split()[0] if header else f"contig{k}"
                    bodies.append((name, min(int(s0) + len(line), e), e))
            bodies = list({name: (name, b0, e) for name, b0, e in bodies}.values())  # a repeated name: the last record wins
            offs = [torch.empty(int(L.bx_fasta_blocks(b0, e)) + 1, dtype=torch.int64, device=dev) for _, b0, e in bodies]
            n_out = C.c_int64(0)
            for (name, b0, e), o in zip(bodies, offs):
                check(L.bx_fasta_count(text.data_ptr(), b0, e, o.data_ptr(), C.byref(n_out), st.cuda_stream))
                lengths[name] = int(n_out.value)
            g = cls._allocate(lengths, host_arrays=False)
            bases2 = torch.zeros(g.n_bases // 16, dtype=torch.int32, device=dev)
            nmask = torch.full((g.n_bases // 32,), -1, dtype=torch.int32, device=dev)  # padding is N
            invalid = torch.empty(invalid_cap, dtype=torch.int64, device=dev)
            n_invalid = torch.zeros(1, dtype=torch.int64, device=dev)
            for (name, b0, e), o in zip(bodies, offs):
                check(L.bx_fasta_pack(text.data_ptr(), b0, e, o.data_ptr(), g.offsets[name], bases2.data_ptr(),
                                      nmask.data_ptr(), invalid.data_ptr(), invalid_cap, n_invalid.data_ptr(),
                                      st.cuda_stream))
            n_bad = int(n_invalid.item())
            if n_bad > invalid_cap:
                raise ValueError(f"{path}: {n_bad} characters outside ACGTacgtNnH- exceed the list of {invalid_cap}")
            g.invalid = np.sort(invalid[:n_bad].cpu().numpy()).astype(np.int64)
            g.bases2 = bases2.cpu().numpy().view(np.uint32)
            g.nmask = nmask.cpu().numpy().view(np.uint32)
            g._device_planes = (dev, bases2, nmask)  # BelugaEngine.set_genome reuses them on the same device
        return g

    def save(self, path: str) -> None:
        np.savez(path, bases2=self.bases2, nmask=self.nmask, invalid=self.invalid, pad=np.int64(self.pad),
                 names=np.array(self.names), offsets=np.array([self.offsets[n] for n in self.names], dtype=np.int64),
                 lengths=np.array([self.lengths[n] for n in self.names], dtype=np.int64))

    @classmethod
    def load(cls, path: str) -> "PackedGenome":
        z = np.load(path, allow_pickle=False)
        g = cls()
        g.names = [str(x) for x in z["names"]]
        g.offsets = {n: int(o) for n, o in zip(g.names, z["offsets"])}
        g.lengths = {n: int(o) for n, o in zip(g.names, z["lengths"])}
        g.bases2, g.nmask, g.invalid = z["bases2"], z["nmask"], z["invalid"]
        g.pad = int(z["pad"]) if "pad" in z.files else 4096  # the padding of caches that predate the field
        g.n_bases = int(g.nmask.size) * 32
        return g

    # ------------------------------------------------------------------ host-side reads (diagnostics, tests)
    def codes(self, gstart: int, n: int) -> np.ndarray:
        """Codes 0..4 of global positions [gstart, gstart+n); outside the array -> 4."""
        idx = np.arange(gstart, gstart + n, dtype=np.int64)
        ok = (idx >= 0) & (idx < self.n_bases)
        j = np.where(ok, idx, 0)
        two = (self.bases2[j >> 4] >> ((j & 15) * 2).astype(np.uint32)) & 3
        msk = (self.nmask[j >> 5] >> (j & 31).astype(np.uint32)) & 1
        out = np.where((msk == 1) | ~ok, 4, two).astype(np.uint8)
        return out

    def upper_string(self, gstart: int, n: int) -> str:
        """Upper-case bases (N for any zero column) — what ``seq[...].upper()`` compares against
        (chromatin.py:207-208) for genomes made of ACGTN."""
        return _DECODE[self.codes(gstart, n)].tobytes().decode("ascii")

    def global_index(self, chrom: str, pos0: np.ndarray | int):
        """0-based contig coordinate -> global index (may fall in the N padding)."""
        return self.offsets[chrom] + pos0

    def any_invalid(self, gstart: np.ndarray, length: np.ndarray) -> int:
        """Index of the first window [gstart, gstart+length) touching an invalid base, or -1."""
        if self.invalid.size == 0:
            return -1
        lo = np.searchsorted(self.invalid, gstart, side="left")
        hi = np.searchsorted(self.invalid, gstart + length, side="left")
        hit = np.flatnonzero(hi > lo)
        return int(hit[0]) if hit.size else -1
