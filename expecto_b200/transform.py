"""Host wrappers of kernel 4, the decay-weighted shift reductions (include/beluga_b200.h:
bx_decay_transform_fixed / bx_decay_transform_variant).

* fixed weights   — compute_expecto_features.py:88-101,123-124: ``pos_weights float64[10,S]`` applied to the
  strand-averaged predictions of the S shifted windows of one gene;
* per-variant weights — predict.py:87-136 (``compute_effects``): weights derived from each variant's signed
  distance to the TSS and the gene strand, applied to the 9 per-shift diff/ref/alt matrices.

Outputs are fp32 [n, 10*2002] with column index ``basis*2002 + feature`` as in the reference; callers widen to
float64 where the reference stores float64 (the ``.npy`` of compute_expecto_features.py:128).
"""
from __future__ import annotations

import numpy as np
import torch

from ._lib import check, lib

N_FEATURES = 2002
N_BASIS = 10
RATES = (0.01, 0.02, 0.05, 0.1, 0.2)
TSS_SHIFTS = np.arange(-20000, 20000, 200)  # compute_expecto_features.py:88


def fixed_pos_weights(shifts=TSS_SHIFTS) -> np.ndarray:
    """float64 [10,S]: exp(-k|s|/200) masked to s<=0 (first five) / s>=0 (last five),
    compute_expecto_features.py:90-101."""
    s = np.asarray(shifts, dtype=np.float64)
    decay = np.stack([np.exp(-k * np.abs(s) / 200) for k in RATES])
    return np.concatenate([decay * (s <= 0), decay * (s >= 0)], axis=0)


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def decay_transform_fixed(probs: torch.Tensor, weights, out: torch.Tensor | None = None) -> torch.Tensor:
    """``probs`` fp32 CUDA [G,S,2,2002] (shift-major; forward then reverse-complement prediction);
    ``weights`` [10,S] -> fp32 [G,20020] = sum_s w[k,s] * 0.5*(fwd+rc)[s,f]."""
    if not probs.is_cuda or probs.dtype != torch.float32 or probs.dim() != 4 or probs.shape[2:] != (2, N_FEATURES):
        raise ValueError("probs must be a CUDA float32 tensor [G,S,2,2002]")
    G, S = probs.shape[:2]
    if isinstance(weights, torch.Tensor) and weights.is_cuda:  # already staged by the caller (no copy per call)
        w = weights.to(dtype=torch.float32).contiguous()
    else:
        w = torch.as_tensor(np.ascontiguousarray(np.asarray(weights, dtype=np.float32))).to(probs.device)
    if tuple(w.shape) != (N_BASIS, S):
        raise ValueError(f"weights must be [10,{S}]")
    probs = probs.contiguous()
    if out is None:
        out = torch.empty((G, N_BASIS * N_FEATURES), dtype=torch.float32, device=probs.device)
    with torch.cuda.device(probs.device):  # the context-free entry points launch on the current device
        check(lib().bx_decay_transform_fixed(probs.data_ptr(), w.data_ptr(), G, S, out.data_ptr(), _stream(probs)))
    return out


def strand_average(x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """predict.py:183-190 on the device: strand-stacked fp32 [2n,2002] -> [n,2002] = (top half + bottom half) / 2."""
    if not x.is_cuda or x.dtype != torch.float32 or x.dim() != 2 or x.shape[1] != N_FEATURES or x.shape[0] % 2:
        raise ValueError("x must be a CUDA float32 tensor [2n,2002]")
    n = x.shape[0] // 2
    x = x.contiguous()
    if out is None:
        out = torch.empty((n, N_FEATURES), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        check(lib().bx_strand_average(x.data_ptr(), n, out.data_ptr(), _stream(x)))
    return out


def chromatin_shifts(maxshift: int = 800) -> list[int]:
    return [0] + list(range(-200, -maxshift - 1, -200)) + list(range(200, maxshift + 1, 200))  # predict.py:173


X_PITCH = 2016   # row pitch (floats) of averaged chromatin effects kept for scoring: rows on 128-byte boundaries


def padded_rows(shape_prefix, device) -> torch.Tensor:
    """float32 [..., 2002] view of a buffer whose rows are X_PITCH floats apart (what ``decay_score_variant``'s
    tensor-core kernel reads fastest: 8008-byte rows make most of its 32-byte reads straddle two DRAM sectors)."""
    buf = torch.empty(tuple(shape_prefix) + (X_PITCH,), dtype=torch.float32, device=device)
    return buf[..., :N_FEATURES]


def select_variants(x: torch.Tensor, sel: torch.Tensor) -> torch.Tensor:
    """``x.index_select(1, sel)`` of a [S, n, 2002] tensor that keeps the row padding of ``padded_rows`` tensors (the
    result of a plain index_select is dense, i.e. back on 8008-byte rows)."""
    S, n, f = x.shape
    pitch = _row_pitch(x)
    if pitch <= f or x.stride(0) != n * pitch:
        return x.index_select(1, sel)
    return x.as_strided((S, n, pitch), (n * pitch, pitch, 1)).index_select(1, sel)[..., :f]


def side_order(dist, strand, shifts=None) -> np.ndarray:
    """Row order that groups variants by the side of the TSS their shifted positions fall on (predict.py:87-109:
    ``pos = (dist + shift) * sgn``): all nine upstream, straddling, all nine downstream — a stable sort, so rows keep
    their order within a group.  The scoring kernel skips the operand columns that are zero for every row of a
    256-variant tile; rows handed over in this order make all but a few tiles one-sided (about a fifth less time).
    Any order gives the same scores bit for bit."""
    d = np.asarray(dist, dtype=np.float64).reshape(-1)
    st = np.asarray(strand).reshape(-1)
    if st.dtype.kind in "USO":
        sgn = np.where(np.asarray([str(v) for v in st]) == "+", 1.0, -1.0)
    else:
        sgn = np.where(st.astype(np.float64) > 0, 1.0, -1.0)
    sh = np.asarray(chromatin_shifts(800) if shifts is None else shifts, dtype=np.float64)
    pos = d[:, None] * sgn[:, None] + sh[None, :] * sgn[:, None]
    key = np.where((pos < 0).all(axis=1), 0, np.where((pos > 0).all(axis=1), 2, 1))
    return np.argsort(key, kind="stable")


def _row_pitch(x: torch.Tensor) -> int:
    """Row pitch of a [S,n,2002] tensor that is dense apart from row padding, or 0."""
    S, n, f = x.shape
    st = x.stride()
    if st[2] != 1 or st[1] < f or (n > 1 and S > 1 and st[0] != st[1] * n):
        return 0
    return int(st[1]) if n > 1 else (int(st[0]) if S > 1 and st[0] >= f else f)


def _variant_args(x: torch.Tensor, dist, strand, shifts, allow_pitch: bool = False):
    if not x.is_cuda or x.dtype != torch.float32 or x.dim() != 3 or x.shape[2] != N_FEATURES:
        raise ValueError("x must be a CUDA float32 tensor [S,n,2002]")
    if not (allow_pitch and _row_pitch(x) >= N_FEATURES and _row_pitch(x) % 2 == 0):
        x = x.contiguous()
    S, n = x.shape[:2]
    shifts_dev = shifts if isinstance(shifts, torch.Tensor) and shifts.is_cuda else None
    if shifts_dev is None:
        shifts = chromatin_shifts((S - 1) // 2 * 200) if shifts is None else list(shifts)
    elif shifts_dev.dtype != torch.int32:
        raise ValueError("device-resident shifts must be int32")
    if len(shifts) != S:
        raise ValueError("len(shifts) must equal x.shape[0]")
    if isinstance(dist, torch.Tensor) and dist.is_cuda and isinstance(strand, torch.Tensor) and strand.is_cuda:
        # already staged by the caller: float64 distances and int8 signs on the device (no copy per call)
        if dist.dtype != torch.float64 or strand.dtype != torch.int8 or dist.numel() != n or strand.numel() != n:
            raise ValueError("device-resident dist/strand must be float64 / int8 (+1, -1) with one entry per variant")
        sh = shifts_dev if shifts_dev is not None else torch.tensor(shifts, dtype=torch.int32, device=x.device)
        return x, dist.contiguous(), strand.contiguous(), sh.contiguous(), n, S
    st = np.asarray(strand)
    if st.dtype.kind in "USO":  # '+' / '-' as predict.py:90 compares them (pandas 3 hands back object arrays)
        sgn = np.where(np.asarray([str(v) for v in st.ravel()]) == "+", 1, -1).astype(np.int8)
    else:
        sgn = np.where(st > 0, 1, -1).astype(np.int8)
    d = torch.from_numpy(np.ascontiguousarray(np.asarray(dist, dtype=np.float64))).to(x.device)
    sg = torch.from_numpy(np.ascontiguousarray(sgn)).to(x.device)
    sh = shifts_dev if shifts_dev is not None else torch.tensor(shifts, dtype=torch.int32, device=x.device)
    if d.numel() != n or sg.numel() != n:
        raise ValueError("dist/strand must have one entry per variant")
    return x, d, sg, sh, n, S


def decay_transform_variant(x: torch.Tensor, dist, strand, shifts=None, out: torch.Tensor | None = None) -> torch.Tensor:
    """``x`` fp32 CUDA [S,n,2002] (one matrix per shift, in ``shifts`` order); ``dist`` = variant - TSS
    (predict.py:241 ``-gene[:, -1]``); ``strand`` array of '+'/'-' (or +1/-1) -> fp32 [n,20020]."""
    x, d, sg, sh, n, S = _variant_args(x, dist, strand, shifts)
    if out is None:
        out = torch.empty((n, N_BASIS * N_FEATURES), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        check(lib().bx_decay_transform_variant(x.data_ptr(), d.data_ptr(), sg.data_ptr(), sh.data_ptr(), n, S,
                                               out.data_ptr(), _stream(x)))
    return out


class ScoreModels:
    """gblinear models staged once on the device for ``decay_score_variant``: fp32 weights [m, 20020] (columns dropped
    by the keep mask = 0) for the SIMT kernel, and — prepared lazily — the tensor-core kernel's operand layout (fp16
    (hi, lo) pairs, K axis permuted to the order the feature tiles are generated in; bx_score_prepare_weights)."""

    def __init__(self, weights, offsets, device):
        w = np.ascontiguousarray(np.asarray(weights, dtype=np.float32)).reshape(-1, N_BASIS * N_FEATURES)
        off = np.ascontiguousarray(np.asarray(offsets, dtype=np.float32)).reshape(-1)
        if off.size != w.shape[0]:
            raise ValueError("one offset per model")
        self.m = int(w.shape[0])
        self.device = torch.device(device)
        self.w = torch.from_numpy(w).to(self.device)
        self.off = torch.from_numpy(off).to(self.device)
        self._wp = None

    def prepared(self) -> torch.Tensor:
        if self._wp is None:
            nbytes = int(lib().bx_score_weights_bytes(self.m))
            self._wp = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
            with torch.cuda.device(self.device):
                check(lib().bx_score_prepare_weights(self.w.data_ptr(), self.m, self._wp.data_ptr(),
                                                     torch.cuda.current_stream(self.device).cuda_stream))
        return self._wp


def decay_score_variant(x: torch.Tensor, dist, strand, weights, offsets=None, shifts=None,
                        tensor_cores: bool | None = None) -> torch.Tensor:
    """predict.py:123-160 fused (transform + gblinear scoring): ``weights`` [m,20020] (dropped columns = 0) with
    ``offsets`` [m] = base_score + bias, or a ``ScoreModels`` staged once -> fp32 [n,m] scores; the feature matrix is
    never materialised.  With the nine shifts of --maxshift 800 the contraction runs as one tcgen05 GEMM whose feature
    tiles are generated in shared memory (bx_decay_score_variant_tc); other shift counts, or ``tensor_cores=False``,
    use the SIMT kernel (bx_decay_score_variant)."""
    use_tc = (x.shape[0] == 9) if tensor_cores is None else bool(tensor_cores)
    x, d, sg, sh, n, S = _variant_args(x, dist, strand, shifts, allow_pitch=use_tc)
    models = weights if isinstance(weights, ScoreModels) else ScoreModels(weights, offsets, x.device)
    if models.device != x.device:
        raise ValueError("models are staged on another device")
    m = models.m
    out = torch.empty((n, m), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        if use_tc:
            pitch = _row_pitch(x) if x.numel() else N_FEATURES
            check(lib().bx_decay_score_variant_tc(x.data_ptr(), pitch, d.data_ptr(), sg.data_ptr(), sh.data_ptr(),
                                                  models.prepared().data_ptr(), models.off.data_ptr(), n, S, m,
                                                  out.data_ptr(), _stream(x)))
        else:
            check(lib().bx_decay_score_variant(x.data_ptr(), d.data_ptr(), sg.data_ptr(), sh.data_ptr(),
                                               models.w.data_ptr(), models.off.data_ptr(), n, S, m, out.data_ptr(),
                                               _stream(x)))
    return out
