"""ctypes binding of the C ABI declared in include/beluga_b200.h.

The shared library is built in-tree (``expecto_b200/lib/libbeluga_b200.so``) by :func:`build` with
``nvcc -gencode arch=compute_100a,code=sm_100a``.  There is no CPU fallback: if the library is
missing or no B200 is present, the compute entry points raise.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
LIB_PATH = os.path.join(_HERE, "lib", "libbeluga_b200.so")
SRC = os.path.join(_HERE, "csrc", "capi.cu")
SRC_DEPS = [os.path.join(_HERE, "csrc", f) for f in ("capi.cu", "gemm_tc.cuh", "ptx.cuh", "kernels_misc.cuh", "encode.cuh", "genome_pack.cuh", "variants.cuh", "score_tc.cuh")] + [
    os.path.join(ROOT, "include", "beluga_b200.h")
]

BX_OK, BX_E_ARG, BX_E_CUDA, BX_E_STATE, BX_E_BASE = 0, -1, -2, -3, -4
DTYPE_F16, DTYPE_BF16, DTYPE_F16X3 = 0, 1, 2


class WindowDesc(C.Structure):
    """bx_window_desc (include/beluga_b200.h)."""

    _fields_ = [
        ("g1", C.c_int64),
        ("g2", C.c_int64),
        ("len1", C.c_int32),
        ("alen", C.c_int32),
        ("len2", C.c_int32),
        ("aoff", C.c_int32),
        ("flags", C.c_int32),
        ("reserved", C.c_int32),
    ]


# numpy structured dtype with the same memory layout as WindowDesc
DESC_DTYPE = [
    ("g1", "<i8"), ("g2", "<i8"), ("len1", "<i4"), ("alen", "<i4"), ("len2", "<i4"), ("aoff", "<i4"),
    ("flags", "<i4"), ("reserved", "<i4"),
]

EXPORTS = [
    "bx_version", "bx_last_error", "bx_launch_count", "bx_ctx_create", "bx_ctx_destroy", "bx_ctx_load_weights",
    "bx_ctx_set_genome", "bx_encode_windows", "bx_forward_windows", "bx_forward_onehot", "bx_debug_activation", "bx_debug_run_layer",
    "bx_chromatin_outputs", "bx_chromatin_outputs_avg", "bx_decay_transform_fixed", "bx_decay_transform_variant",
    "bx_forward_onehot_host", "bx_chromatin_shift_host", "bx_ctx_profile", "bx_ctx_profile_read",
    "bx_decay_score_variant", "bx_forward_segments", "bx_segment_capacity", "bx_strand_average",
    "bx_forward_segments_delta", "bx_fasta_blocks", "bx_fasta_find_headers", "bx_fasta_count", "bx_fasta_pack",
    "bx_chromatin_open", "bx_chromatin_chunk_info", "bx_chromatin_submit", "bx_chromatin_wait",
    "bx_chromatin_submit_device", "bx_chromatin_set_row_pitch", "bx_chromatin_close", "bx_variant_descs_host",
    "bx_score_weights_bytes", "bx_score_prepare_weights", "bx_decay_score_variant_tc", "bx_fasta_count_noalloc",
]
CHROM_NO_DELTA, CHROM_AVERAGE, CHROM_WINDOWS = 1, 2, 4
NSTAGES = 10
STAGE_NAMES = ["encode", "conv1", "conv2", "conv3", "conv4", "conv5", "conv6", "fc1", "fc2", "outputs"]


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(s) > t for s in SRC_DEPS if os.path.exists(s))


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile the CUDA extension for sm_100a (nvcc cross-compiles without a GPU)."""
    if not force and not needs_build():
        return LIB_PATH
    os.makedirs(os.path.dirname(LIB_PATH), exist_ok=True)
    cmd = [
        "nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--shared",
        "-Xcompiler", "-fPIC", "-o", LIB_PATH, SRC,
    ]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    cmd[1:1] = os.environ.get("BX_NVCC_FLAGS", "").split()   # experiments (e.g. -DBX_SCORE_SPAN=4)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        sys.stderr.write(res.stderr)
    return LIB_PATH


_lib = None


def lib() -> C.CDLL:
    """Load the C-ABI library (loading needs no GPU; calling compute entry points does)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback for the Beluga hot path)")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64 = C.c_void_p, C.c_int, C.c_int64
    L.bx_version.restype = i32
    L.bx_last_error.restype = C.c_char_p
    L.bx_launch_count.restype = i64
    L.bx_ctx_create.argtypes = [i32, i32, i32, i32, C.POINTER(vp)]
    L.bx_ctx_destroy.argtypes = [vp]
    L.bx_ctx_load_weights.argtypes = [vp, C.POINTER(vp), C.POINTER(vp)]
    L.bx_ctx_set_genome.argtypes = [vp, vp, vp, i64]
    L.bx_encode_windows.argtypes = [vp, vp, vp, i32, vp, vp]
    L.bx_forward_windows.argtypes = [vp, vp, vp, i32, vp, vp]
    L.bx_forward_onehot.argtypes = [vp, vp, i32, vp, vp]
    L.bx_debug_activation.argtypes = [vp, i32, i32, vp, C.POINTER(i64), vp]
    L.bx_debug_run_layer.argtypes = [vp, i32, i32, vp, vp]
    L.bx_chromatin_outputs.argtypes = [vp, i32, vp, vp, vp, vp]
    L.bx_chromatin_outputs_avg.argtypes = [vp, i32, vp, vp, vp, vp]
    L.bx_decay_transform_fixed.argtypes = [vp, vp, i32, i32, vp, vp]
    L.bx_decay_transform_variant.argtypes = [vp, vp, vp, vp, i32, i32, vp, vp]
    L.bx_forward_onehot_host.argtypes = [vp, vp, i32, vp]
    L.bx_chromatin_shift_host.argtypes = [vp, vp, vp, i64, i32, i32, vp, vp, vp]
    L.bx_decay_score_variant.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, i32, vp, vp]
    L.bx_forward_segments.argtypes = [vp, vp, vp, i32, i32, i32, i32, C.POINTER(C.c_int32), i32, vp, vp]
    L.bx_forward_segments_delta.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, i32, C.POINTER(C.c_int32), i32, vp, vp]
    L.bx_segment_capacity.argtypes = [vp, i32]
    L.bx_strand_average.argtypes = [vp, i32, vp, vp]
    L.bx_fasta_blocks.argtypes = [i64, i64]
    L.bx_fasta_blocks.restype = i64
    L.bx_fasta_find_headers.argtypes = [vp, i64, vp, i64, vp, vp]
    L.bx_fasta_count.argtypes = [vp, i64, i64, vp, C.POINTER(i64), vp]
    L.bx_fasta_pack.argtypes = [vp, i64, i64, vp, i64, vp, vp, vp, i64, vp, vp]
    ip = C.POINTER(i32)
    L.bx_chromatin_open.argtypes = [vp, i64, vp, vp, vp, vp, vp, vp, i64, i32, i32, i32, ip, ip]
    L.bx_chromatin_chunk_info.argtypes = [vp, i32, C.POINTER(i64), ip, ip, ip, ip]
    L.bx_chromatin_submit.argtypes = [vp, i32, vp, vp, vp]
    L.bx_chromatin_wait.argtypes = [vp, i32]
    L.bx_chromatin_submit_device.argtypes = [vp, i32, vp, vp, vp, i64, i64, i64, vp]
    L.bx_chromatin_set_row_pitch.argtypes = [vp, i64]
    L.bx_chromatin_close.argtypes = [vp]
    L.bx_variant_descs_host.argtypes = [i32, i64, vp, vp, vp, vp, vp, i32, i32, ip, ip, vp]
    L.bx_score_weights_bytes.argtypes = [i32]
    L.bx_score_weights_bytes.restype = i64
    L.bx_score_prepare_weights.argtypes = [vp, i32, vp, vp]
    L.bx_decay_score_variant_tc.argtypes = [vp, i64, vp, vp, vp, vp, vp, i32, i32, i32, vp, vp]
    L.bx_fasta_count_noalloc.argtypes = [vp, i64, i64, vp, vp, vp]
    L.bx_ctx_profile.argtypes = [vp, i32]
    L.bx_ctx_profile_read.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(i64), i32]
    for name in EXPORTS:
        if name not in ("bx_last_error", "bx_launch_count", "bx_fasta_blocks", "bx_score_weights_bytes"):
            getattr(L, name).restype = i32
    _lib = L
    return L


class BelugaError(RuntimeError):
    pass


def check(rc: int) -> None:
    if rc != 0:
        msg = lib().bx_last_error().decode("utf-8", "replace")
        if rc == BX_E_BASE:
            raise KeyError(msg)
        raise BelugaError(f"beluga_b200 error {rc}: {msg}")
