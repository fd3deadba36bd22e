"""bx_fasta_count_noalloc (include/beluga_b200.h): the allocation-free variant of bx_fasta_count, whose scan scratch
lives in a file-scope device array instead of coming from cudaMallocAsync.  It has no counterpart in the reference.
Collected last on purpose; it
depends on nothing the other tests set up and passes alone in a fresh process as well as after them.

Integer work: the block offsets and the contig length must EQUAL those of the existing entry point, on two calls in a
row that share the scratch array (the second on a different, smaller contig, so stale scratch contents would show)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _fasta_text(n_bases: int, width: int, seed: int, crlf: bool = False) -> bytes:
    rng = np.random.default_rng(seed)
    body = np.frombuffer(b"ACGTacgtN", dtype=np.uint8)[rng.integers(0, 9, size=n_bases)]
    eol = b"\r\n" if crlf else b"\n"
    return b">c1 test\n" + eol.join(body[i:i + width].tobytes() for i in range(0, n_bases, width)) + eol


def test_noalloc_count_equals_bx_fasta_count_on_two_calls_in_a_row():
    import torch

    from expecto_b200 import _lib
    from expecto_b200._lib import check, lib

    _lib.build()
    L = lib()
    dev = torch.device("cuda", 0)
    st = torch.cuda.current_stream(dev).cuda_stream
    SENT = -7
    # second call: a smaller contig with CRLF line ends and a ragged width
    for n_bases, width, seed, crlf in ((1_500_007, 61, 1, False), (70_001, 50, 2, True)):
        raw = _fasta_text(n_bases, width, seed, crlf)
        body0, body1 = raw.index(b"\n") + 1, len(raw)
        text = torch.full(((len(raw) + 15) // 16 * 16 + 16,), 10, dtype=torch.uint8, device=dev)
        text[:len(raw)] = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(dev)
        nb = int(L.bx_fasta_blocks(body0, body1))
        assert nb > 1
        want = torch.full((nb + 1,), SENT, dtype=torch.int64, device=dev)
        n_host = C.c_int64(0)
        check(L.bx_fasta_count(text.data_ptr(), body0, body1, want.data_ptr(), C.byref(n_host), st))
        assert n_host.value == n_bases
        # the caller owns every buffer; guard words around the outputs must come back intact
        got = torch.full((nb + 1 + 8,), SENT, dtype=torch.int64, device=dev)
        n_dev = torch.full((3,), SENT, dtype=torch.int64, device=dev)
        check(L.bx_fasta_count_noalloc(text.data_ptr(), body0, body1, got.data_ptr() + 4 * 8, n_dev.data_ptr() + 8, st))
        torch.cuda.synchronize(dev)
        assert torch.equal(got[4:4 + nb + 1], want)
        assert n_dev.tolist() == [SENT, n_bases, SENT]
        assert bool((got[:4] == SENT).all()) and bool((got[4 + nb + 1:] == SENT).all())


def test_noalloc_count_rejects_sizes_beyond_its_static_scratch_and_bad_arguments():
    import torch

    from expecto_b200 import _lib
    from expecto_b200._lib import BX_E_ARG, lib

    _lib.build()
    L = lib()
    dev = torch.device("cuda", 0)
    text = torch.full((4096,), 10, dtype=torch.uint8, device=dev)
    off = torch.zeros(16, dtype=torch.int64, device=dev)
    n_dev = torch.zeros(1, dtype=torch.int64, device=dev)
    # a body of 2^20 + 1 blocks: refused before anything is launched or read
    too_long = 16 + ((1 << 20) + 1) * 4096
    assert L.bx_fasta_count_noalloc(text.data_ptr(), 16, too_long, off.data_ptr(), n_dev.data_ptr(), None) == BX_E_ARG
    assert b"exceeds" in L.bx_last_error() and b"bx_fasta_count" in L.bx_last_error()
    assert L.bx_fasta_count_noalloc(text.data_ptr(), 16, 64, None, n_dev.data_ptr(), None) == BX_E_ARG
    assert L.bx_fasta_count_noalloc(text.data_ptr() + 1, 16, 64, off.data_ptr(), n_dev.data_ptr(), None) == BX_E_ARG
    # an empty body is fine: total 0, nothing launched
    assert L.bx_fasta_count_noalloc(text.data_ptr(), 32, 32, off.data_ptr(), n_dev.data_ptr(), None) == 0
    torch.cuda.synchronize(dev)
    assert int(n_dev.item()) == 0
