"""The C-ABI library: builds for sm_100a without a GPU, exports every symbol include/beluga_b200.h declares,
and refuses to compute without a B200 (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, has_gpu
from expecto_b200 import _lib


@pytest.fixture(scope="module")
def library():
    _lib.build()
    return _lib.lib()


def header_symbols():
    text = open(os.path.join(ROOT, "include", "beluga_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bx_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree(library):
    syms = header_symbols()
    assert len(syms) >= 20
    assert sorted(_lib.EXPORTS) == syms
    for s in syms:
        assert hasattr(library, s), f"{s} declared in include/beluga_b200.h but not exported"


def test_desc_struct_layout():
    assert C.sizeof(_lib.WindowDesc) == 40 == np.dtype(_lib.DESC_DTYPE).itemsize
    for (name, _), f in zip(_lib.DESC_DTYPE, _lib.WindowDesc._fields_):
        assert name == f[0]
        assert np.dtype(_lib.DESC_DTYPE).fields[name][1] == getattr(_lib.WindowDesc, name).offset


def test_sass_is_blackwell_native(library):
    """tcgen05.mma / tcgen05.ld / TMA show up as UTCHMMA / LDTM / UTMALDG in the SASS (B200_PROFILING.md)."""
    out = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump not available")
    sass = out.stdout
    assert "UTCHMMA" in sass and "LDTM" in sass and "UTMALDG" in sass
    assert "HMMA.16816" not in sass, "legacy mma.sync path must not exist"
    assert "sm_100a" in sass


@pytest.mark.skipif(has_gpu(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(library):
    ctx = C.c_void_p()
    rc = library.bx_ctx_create(0, _lib.DTYPE_F16, 2, 64, C.byref(ctx))
    assert rc == _lib.BX_E_CUDA
    assert b"no CPU fallback" in library.bx_last_error()
    with pytest.raises(RuntimeError):
        from expecto_b200.engine import BelugaEngine
        BelugaEngine(0)


def test_argument_errors(library):
    ctx = C.c_void_p()
    assert library.bx_ctx_create(0, 7, 2, 64, C.byref(ctx)) == _lib.BX_E_ARG
    assert library.bx_ctx_create(0, 0, 3, 64, C.byref(ctx)) == _lib.BX_E_ARG
    assert library.bx_ctx_create(0, 0, 2, 0, C.byref(ctx)) == _lib.BX_E_ARG
    assert library.bx_chromatin_outputs(None, 1, None, None, None, None) == _lib.BX_E_ARG
    assert library.bx_chromatin_outputs(None, 0, None, None, None, None) == 0   # empty input is a no-op
    assert library.bx_chromatin_outputs(None, -1, None, None, None, None) == _lib.BX_E_ARG
    assert library.bx_decay_transform_fixed(None, None, 1, 200, None, None) == _lib.BX_E_ARG
    assert library.bx_version() >= 100
