"""Host-side handle on the CUDA Beluga engine (one per GPU).

Mirrors what the reference scripts do with ``model = Beluga(); model.load_state_dict(...); model.cuda()``
(chromatin.py:102-106) and ``model.forward(input)`` (chromatin.py:270), but every FLOP runs in the
hand-written sm_100a kernels behind the C ABI.  PyTorch is used only for device buffers and streams.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import check, lib

N_FEATURES = 2002
WINDOW = 2000
_ACT_SHAPES = {0: (1996, 64), 1: (1996, 320), 2: (496, 320), 3: (492, 480), 4: (120, 480), 5: (113, 640),
               6: (106, 640), 7: (1, 2048)}
# rows of each activation that hold valid reference values (the rest of the pitch is junk by design)
_ACT_VALID = {0: 1993, 1: 1993, 2: 496, 3: 489, 4: 120, 5: 113, 6: 106, 7: 1}


def _stream_ptr(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


class BelugaEngine:
    """Beluga forward on one B200.  ``dtype`` is the tensor-core operand type: "f16" (default), "bf16", or "f16x3" —
    fp16 operands split into (hi, lo) pairs, three MMAs per product, ~1e-5 from the fp32 reference at a third of
    the throughput; accumulation, bias, ReLU/MaxPool inputs, sigmoid and all outputs are fp32."""

    def __init__(self, device: int | str | torch.device = 0, dtype: str = "f16", cta_group: int = 2,
                 capacity: int = 1024):
        if not torch.cuda.is_available():
            raise RuntimeError("BelugaEngine needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        self.device = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        self.dtype = dtype
        self.capacity = int(capacity)
        self._ctx = C.c_void_p()
        code = {"f16": _lib.DTYPE_F16, "bf16": _lib.DTYPE_BF16, "f16x3": _lib.DTYPE_F16X3}[dtype]
        check(lib().bx_ctx_create(self.device.index or 0, code, cta_group, self.capacity, C.byref(self._ctx)))
        self._genome = None
        self._keep = []

    def close(self) -> None:
        if getattr(self, "_ctx", None) is not None and self._ctx.value:
            lib().bx_ctx_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ weights
    def load_state_dict(self, sd) -> None:
        """``sd``: mapping with the reference's keys (``model.0.0.weight`` ...), torch tensors or numpy."""
        from .synth import STATE_KEYS

        ws, bs = [], []
        for key, shape in STATE_KEYS:
            w = sd[key + ".weight"]
            b = sd[key + ".bias"]
            w = w.detach().cpu().numpy() if isinstance(w, torch.Tensor) else np.asarray(w)
            b = b.detach().cpu().numpy() if isinstance(b, torch.Tensor) else np.asarray(b)
            w = np.ascontiguousarray(w, dtype=np.float32)
            b = np.ascontiguousarray(b, dtype=np.float32)
            if tuple(w.shape) != tuple(shape) or b.shape != (shape[0],):
                raise ValueError(f"{key}: expected weight {shape}, got {w.shape} / bias {b.shape}")
            ws.append(w)
            bs.append(b)
        wp = (C.c_void_p * 8)(*[w.ctypes.data for w in ws])
        bp = (C.c_void_p * 8)(*[b.ctypes.data for b in bs])
        with torch.cuda.device(self.device):
            check(lib().bx_ctx_load_weights(self._ctx, wp, bp))

    def load_weights_file(self, path: str) -> None:
        """``torch.load(path)`` of a Beluga state_dict (chromatin.py:103)."""
        self.load_state_dict(torch.load(path, map_location="cpu"))

    # ------------------------------------------------------------------ genome
    def set_genome(self, packed) -> None:
        """``packed``: expecto_b200.genome.PackedGenome; uploads the 2-bit planes once."""
        self._genome = packed
        planes = getattr(packed, "_device_planes", None)
        if planes is not None and planes[0] == self.device:  # packed on this device (PackedGenome.from_fasta)
            self._bases2, self._nmask = planes[1], planes[2]
        else:
            self._bases2 = torch.from_numpy(packed.bases2.view(np.int32)).to(self.device)
            self._nmask = torch.from_numpy(packed.nmask.view(np.int32)).to(self.device)
        check(lib().bx_ctx_set_genome(self._ctx, self._bases2.data_ptr(), self._nmask.data_ptr(), packed.n_bases))

    # ------------------------------------------------------------------ compute
    def _stage(self, host: np.ndarray) -> torch.Tensor:
        """Bytes of ``host`` -> device through a ring of reusable pinned buffers, so the H2D copy is truly
        asynchronous (a pageable source makes cudaMemcpyAsync wait for the stream's earlier work once it exceeds the
        driver's 64 KB staging chunk, which serialised the host with the GPU in the chunked segment loops)."""
        raw = np.ascontiguousarray(host).view(np.uint8).reshape(-1)
        ring = self.__dict__.setdefault("_pin_ring", [])
        slot = self.__dict__.get("_pin_next", 0)
        self._pin_next = (slot + 1) % 32
        if len(ring) <= slot:
            ring.append([None, None])
        buf, ev = ring[slot]
        if ev is not None:
            ev.synchronize()           # the copy that last used this buffer has finished
        if buf is None or buf.numel() < raw.size:
            buf = torch.empty(max(raw.size, 1 << 16), dtype=torch.uint8, pin_memory=True)
        buf[:raw.size].numpy()[:] = raw
        dev = buf[:raw.size].to(self.device, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        ring[slot] = [buf, ev]
        return dev

    def _upload_descs(self, descs: np.ndarray, alleles: np.ndarray):
        if alleles.size == 0:
            alleles = np.zeros(1, dtype=np.uint8)
        return self._stage(descs), self._stage(np.ascontiguousarray(alleles, dtype=np.uint8))

    def encode_windows(self, descs: np.ndarray, alleles: np.ndarray) -> torch.Tensor:
        """encodeSeqs materialised on the GPU: uint8 [n,4,2000] (bit-exact test hook)."""
        n = len(descs)
        d, a = self._upload_descs(descs, alleles)
        out = torch.empty((n, 4, WINDOW), dtype=torch.uint8, device=self.device)
        check(lib().bx_encode_windows(self._ctx, d.data_ptr(), a.data_ptr(), n, out.data_ptr(),
                                      _stream_ptr(self.device)))
        return out

    def forward_windows(self, descs, alleles, out: torch.Tensor | None = None) -> torch.Tensor:
        """Beluga probabilities fp32 [n,2002] for windows described against the packed genome.
        ``descs``/``alleles`` may be numpy (uploaded here) or device tensors already resident in HBM."""
        if isinstance(descs, np.ndarray):
            n = len(descs)
            d, a = self._upload_descs(descs, alleles)
        else:
            d, a = descs, alleles
            n = d.numel() // 40
        if out is None:
            out = torch.empty((n, N_FEATURES), dtype=torch.float32, device=self.device)
        check(lib().bx_forward_windows(self._ctx, d.data_ptr(), a.data_ptr(), n, out.data_ptr(),
                                       _stream_ptr(self.device)))
        return out

    def segment_capacity(self, seg_len: int) -> int:
        """Segments of ``seg_len`` positions per chunk of this engine's activation scratch."""
        r = lib().bx_segment_capacity(self._ctx, int(seg_len))
        if r < 0:
            check(r)
        return int(r)

    def forward_segments(self, descs: np.ndarray, alleles: np.ndarray, n_groups: int, segs_per_group: int,
                         seg_len: int, k_of_shift: np.ndarray, shift_major: bool,
                         out: torch.Tensor | None = None) -> torch.Tensor:
        """Segment mode (bx_forward_segments): all shifted windows of ``n_groups`` variants / genes with the conv
        stack evaluated once per segment.  Returns fp32 [n_shifts, n_groups, segs, 2002] (``shift_major``) or
        [n_groups, n_shifts, segs, 2002]; bit-identical to ``forward_windows`` on the same windows."""
        k = np.ascontiguousarray(k_of_shift, dtype=np.int32)
        S = int(k.size)
        assert len(descs) == n_groups * segs_per_group
        d, a = self._upload_descs(descs, alleles)
        shape = (S, n_groups, segs_per_group, N_FEATURES) if shift_major else (n_groups, S, segs_per_group, N_FEATURES)
        if out is None:
            out = torch.empty(shape, dtype=torch.float32, device=self.device)
        check(lib().bx_forward_segments(self._ctx, d.data_ptr(), a.data_ptr(), n_groups, segs_per_group, int(seg_len), S,
                                        k.ctypes.data_as(C.POINTER(C.c_int32)), 1 if shift_major else 0,
                                        out.data_ptr(), _stream_ptr(self.device)))
        return out

    def forward_segments_delta(self, ref_descs: np.ndarray, patch_descs: np.ndarray, alleles: np.ndarray,
                               n_groups: int, seg_len: int, patch_len: int, patch_row0: int, k_of_shift: np.ndarray,
                               shift_major: bool, out: torch.Tensor | None = None) -> torch.Tensor:
        """Segment mode for substitutions (bx_forward_segments_delta): the two ref segments of each variant run in
        full, the alt alleles only as ``patch_len``-position patches around the variant.  Returns fp32
        [n_shifts, n_groups, 4, 2002] (``shift_major``) or [n_groups, n_shifts, 4, 2002] in the order ref fwd, ref rc,
        alt fwd, alt rc — bit-identical to ``forward_segments`` on the four full segments."""
        k = np.ascontiguousarray(k_of_shift, dtype=np.int32)
        S = int(k.size)
        assert len(ref_descs) == 2 * n_groups and len(patch_descs) == 2 * n_groups
        both = np.concatenate([np.ascontiguousarray(ref_descs), np.ascontiguousarray(patch_descs)])
        d, a = self._upload_descs(both, alleles)
        shape = (S, n_groups, 4, N_FEATURES) if shift_major else (n_groups, S, 4, N_FEATURES)
        if out is None:
            out = torch.empty(shape, dtype=torch.float32, device=self.device)
        check(lib().bx_forward_segments_delta(self._ctx, d.data_ptr(), d.data_ptr() + 2 * n_groups * both.dtype.itemsize,
                                              a.data_ptr(), n_groups, int(seg_len), int(patch_len), int(patch_row0), S,
                                              k.ctypes.data_as(C.POINTER(C.c_int32)), 1 if shift_major else 0,
                                              out.data_ptr(), _stream_ptr(self.device)))
        return out

    # ------------------------------------------------------------------ the variant loop inside the library
    def chromatin_open(self, cols: dict, maxshift: int = 800, delta: bool = True, average: bool = False,
                       windows_only: bool = False, chunk_variants: int = 0) -> tuple[int, int]:
        """bx_chromatin_open: hand the whole-VCF columns (``windows.variant_columns``) to the library.  Returns
        (number of chunks, variants per chunk)."""
        flags = (0 if delta else _lib.CHROM_NO_DELTA) | (_lib.CHROM_AVERAGE if average else 0) | \
            (_lib.CHROM_WINDOWS if windows_only else 0)
        nch, chunk = C.c_int(0), C.c_int(0)
        p = lambda a: a.ctypes.data if a.size else None  # noqa: E731
        check(lib().bx_chromatin_open(self._ctx, int(cols["n"]), p(cols["gpos"]), p(cols["ref_len"]), p(cols["alt_len"]),
                                      p(cols["ref_aoff"]), p(cols["alt_aoff"]), p(cols["alleles"]),
                                      int(cols["alleles"].size), int(maxshift), flags, int(chunk_variants),
                                      C.byref(nch), C.byref(chunk)))
        self._plan = {"S": 1 + 2 * (int(maxshift) // 200), "average": average, "n": int(cols["n"]), "chunk": chunk.value}
        return nch.value, chunk.value

    def chromatin_chunk_info(self, k: int) -> dict:
        v0, m, nd, nf, nw = C.c_int64(0), C.c_int(0), C.c_int(0), C.c_int(0), C.c_int(0)
        check(lib().bx_chromatin_chunk_info(self._ctx, k, C.byref(v0), C.byref(m), C.byref(nd), C.byref(nf), C.byref(nw)))
        return {"v0": v0.value, "m": m.value, "delta": nd.value, "full": nf.value, "window": nw.value}

    def chromatin_submit(self, k: int, ref: torch.Tensor, alt: torch.Tensor, diff: torch.Tensor) -> None:
        """Enqueue chunk ``k``; its ref / alt / diff arrive in the given pinned host tensors as dense
        [S, 2, m, 2002] (``average``: [S, m, 2002]).  ``chromatin_wait(k)`` before touching them."""
        for t in (ref, alt, diff):
            if t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous():
                raise ValueError("chromatin_submit takes contiguous float32 host tensors")
        check(lib().bx_chromatin_submit(self._ctx, k, ref.data_ptr(), alt.data_ptr(), diff.data_ptr()))

    def chromatin_wait(self, k: int) -> None:
        check(lib().bx_chromatin_wait(self._ctx, k))

    def chromatin_submit_device(self, k: int, ref: torch.Tensor, alt: torch.Tensor, diff: torch.Tensor,
                                shift_stride: int, strand_stride: int, row0: int) -> None:
        """Chunk ``k`` with the results left in HBM on the current stream: row (shift si, strand t, variant v of the
        chunk) goes to row ``si * shift_stride + t * strand_stride + row0 + v`` of each [..., 2002] device tensor.
        The tensors are dense, or dense apart from a common row padding (``transform.padded_rows``: the row pitch is
        handed to the library with bx_chromatin_set_row_pitch)."""
        pitches = set()
        for t in (ref, alt, diff):
            if not t.is_cuda or t.dtype != torch.float32 or t.dim() < 2 or t.shape[-1] != N_FEATURES or t.stride(-1) != 1:
                raise ValueError("chromatin_submit_device takes float32 CUDA tensors [..., 2002]")
            pitch = t.stride(-2) if t.shape[-2] > 1 or t.dim() == 2 else N_FEATURES
            if any(t.shape[i] > 1 and t.stride(i) != t.stride(i + 1) * t.shape[i + 1] for i in range(t.dim() - 2)):
                raise ValueError("chromatin_submit_device: only the rows may be padded")
            pitches.add(int(pitch))
        if len(pitches) != 1:
            raise ValueError("ref / alt / diff must share one row pitch")
        check(lib().bx_chromatin_set_row_pitch(self._ctx, pitches.pop()))
        check(lib().bx_chromatin_submit_device(self._ctx, k, ref.data_ptr(), alt.data_ptr(), diff.data_ptr(),
                                               int(shift_stride), int(strand_stride), int(row0), _stream_ptr(self.device)))

    def chromatin_close(self) -> None:
        check(lib().bx_chromatin_close(self._ctx))

    def forward_onehot(self, onehot: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
        """``onehot``: uint8/bool [n,4,2000] (or [n,4,1,2000]) on this device -> fp32 [n,2002]."""
        x = onehot.reshape(-1, 4, WINDOW)
        if x.dtype == torch.bool:
            x = x.view(torch.uint8)
        if x.dtype != torch.uint8 or x.device != self.device:
            raise ValueError("onehot must be a uint8/bool tensor on the engine's device")
        x = x.contiguous()
        n = x.shape[0]
        if out is None:
            out = torch.empty((n, N_FEATURES), dtype=torch.float32, device=self.device)
        check(lib().bx_forward_onehot(self._ctx, x.data_ptr(), n, out.data_ptr(), _stream_ptr(self.device)))
        return out

    __call__ = forward_onehot

    # ------------------------------------------------------------------ host-buffer entry points
    def forward_onehot_host(self, onehot: np.ndarray | torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
        """``model.forward`` as chromatin.py:266-270 uses it, host to host: ``onehot`` uint8/bool [n,4,2000] in
        HOST memory (pinned for asynchronous copies) -> fp32 [n,2002] in pinned host memory.  H2D, compute and
        D2H of consecutive chunks overlap inside the library."""
        x = torch.from_numpy(onehot) if isinstance(onehot, np.ndarray) else onehot
        if x.is_cuda:
            raise ValueError("forward_onehot_host takes host memory; use forward_onehot for device tensors")
        x = x.reshape(-1, 4, WINDOW)
        if x.dtype == torch.bool:
            x = x.view(torch.uint8)
        if x.dtype != torch.uint8:
            raise ValueError("onehot must be uint8/bool")
        x = x.contiguous()
        n = x.shape[0]
        if out is None:
            out = torch.empty((n, N_FEATURES), dtype=torch.float32, pin_memory=True)
        check(lib().bx_forward_onehot_host(self._ctx, x.data_ptr(), n, out.data_ptr()))
        return out

    def chromatin_shift_host(self, descs: np.ndarray, alleles: np.ndarray, n_var: int, average: bool = False,
                             out=None):
        """One shift of chromatin.py:243-286 from host descriptors to host (pinned) ref/alt/diff arrays."""
        descs = np.ascontiguousarray(descs)
        assert descs.size == 4 * n_var and descs.dtype.itemsize == 40
        alleles = np.ascontiguousarray(alleles, dtype=np.uint8)
        rows = n_var if average else 2 * n_var
        if out is None:
            out = tuple(torch.empty((rows, N_FEATURES), dtype=torch.float32, pin_memory=True) for _ in range(3))
        ref, alt, diff = out
        check(lib().bx_chromatin_shift_host(self._ctx, descs.ctypes.data, alleles.ctypes.data if alleles.size else None,
                                            alleles.size, n_var, 1 if average else 0, ref.data_ptr(), alt.data_ptr(),
                                            diff.data_ptr()))
        return ref, alt, diff

    # ------------------------------------------------------------------ per-stage device timing
    def profile(self, enable: bool = True) -> None:
        check(lib().bx_ctx_profile(self._ctx, 1 if enable else 0))

    def profile_read(self, reset: bool = True) -> dict[str, tuple[float, int]]:
        """{stage: (summed ms, count)} measured with CUDA events around each stage's launches."""
        ms = (C.c_double * _lib.NSTAGES)()
        cnt = (C.c_int64 * _lib.NSTAGES)()
        check(lib().bx_ctx_profile_read(self._ctx, ms, cnt, 1 if reset else 0))
        return {name: (ms[i], cnt[i]) for i, name in enumerate(_lib.STAGE_NAMES)}

    def debug_activation(self, layer: int, n: int) -> torch.Tensor:
        """Internal activation after the last forward: [n, pitch, C] in the operand dtype (test hook)."""
        rows, cols = _ACT_SHAPES[layer]
        tdt = torch.bfloat16 if self.dtype == "bf16" else torch.float16
        dst = torch.empty((n, rows, cols), dtype=tdt, device=self.device)
        cnt = C.c_int64()
        check(lib().bx_debug_activation(self._ctx, layer, n, dst.data_ptr(), C.byref(cnt), _stream_ptr(self.device)))
        assert cnt.value == dst.numel()
        return dst


def variant_descs_host(kind: int, cols: dict, maxshift: int = 800, shift: int = 0):
    """The device descriptor builders (csrc/variants.cuh) on host columns — test hook (bx_variant_descs_host).
    kind 0: [n, 4] full segments; 1: ([n, 2] ref segments, [n, 2] alt patches, patch_row0, patch_len);
    2: [n, 4] windows at ``shift``."""
    n = int(cols["n"])
    out = np.zeros(4 * n, dtype=_lib.DESC_DTYPE)
    row0, plen = C.c_int(0), C.c_int(0)
    p = lambda a: a.ctypes.data if a.size else None  # noqa: E731
    check(lib().bx_variant_descs_host(kind, n, p(cols["gpos"]), p(cols["ref_len"]), p(cols["alt_len"]),
                                      p(cols["ref_aoff"]), p(cols["alt_aoff"]), int(maxshift), int(shift),
                                      C.byref(row0), C.byref(plen), out.ctypes.data if n else None))
    if kind == 1:
        return out[:2 * n].reshape(n, 2), out[2 * n:].reshape(n, 2), row0.value, plen.value
    return out.reshape(n, 4)


def chromatin_outputs(probs: torch.Tensor, n_var: int, average: bool = False):
    """ref/alt/diff in chromatin.py:281-286 layout (or strand-averaged as predict.py:186-190 reads them)
    from probabilities ordered (variant, {ref fwd, ref rc, alt fwd, alt rc})."""
    rows = n_var if average else 2 * n_var
    ref = torch.empty((rows, N_FEATURES), dtype=torch.float32, device=probs.device)
    alt = torch.empty_like(ref)
    diff = torch.empty_like(ref)
    fn = lib().bx_chromatin_outputs_avg if average else lib().bx_chromatin_outputs
    with torch.cuda.device(probs.device):  # context-free entry point: launches on the current device
        check(fn(probs.data_ptr(), n_var, ref.data_ptr(), alt.data_ptr(), diff.data_ptr(), _stream_ptr(probs.device)))
    return ref, alt, diff
