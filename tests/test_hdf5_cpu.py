"""Dependency-free HDF5 subset (chromatin.py:282-286 writes it, predict.py:173-190 reads it)."""
import os

import numpy as np
import pytest

from expecto_b200 import hdf5

EXAMPLE = "/root/reference/example/example.vcf.shift_0.diff.h5"


def test_roundtrip_three_datasets(tmp_path):
    rng = np.random.default_rng(0)
    arrs = {k: rng.standard_normal((20, 2002)).astype(np.float32) for k in ("diff", "ref", "alt")}
    p = str(tmp_path / "snps.shift_0.diff.h5")
    hdf5.write_datasets(p, arrs)
    with hdf5.File(p) as f:
        assert sorted(f.keys()) == ["alt", "diff", "ref"]
        for k, v in arrs.items():
            assert f[k].shape == (20, 2002)
            assert np.array_equal(np.asarray(f[k]), v)
            assert np.array_equal(f[k][3:7, :], v[3:7])


def test_row_slab_writes(tmp_path):
    """Ranks / pipeline stages fill one contiguous dataset by row slabs."""
    p = str(tmp_path / "slabs.h5")
    a = np.arange(12 * 2002, dtype=np.float32).reshape(12, 2002)
    with hdf5.H5Writer(p, {"diff": (12, 2002)}) as w:
        w.write_rows("diff", 6, a[6:])
        w.write_rows("diff", 0, a[:6])
    with hdf5.File(p) as f:
        assert np.array_equal(np.asarray(f["diff"]), a)


def test_empty_dataset(tmp_path):
    p = str(tmp_path / "empty.h5")
    hdf5.write_datasets(p, {"diff": np.zeros((0, 2002), np.float32)})
    with hdf5.File(p) as f:
        assert f["diff"].shape == (0, 2002)


@pytest.mark.skipif(not os.path.exists(EXAMPLE), reason="reference example files only exist in the build container")
def test_reads_reference_example_and_writer_is_byte_compatible(tmp_path):
    with hdf5.File(EXAMPLE) as f:
        assert list(f.keys()) == ["pred"]
        pred = np.asarray(f["pred"])
    assert pred.shape == (20, 2002) and pred.dtype == np.float32
    assert abs(float(pred.min()) + 0.111) < 2e-3 and abs(float(pred.max()) - 0.191) < 2e-3
    p = str(tmp_path / "pred.h5")
    hdf5.write_datasets(p, {"pred": pred})
    ours, theirs = open(p, "rb").read(), open(EXAMPLE, "rb").read()
    assert len(ours) == len(theirs)
    assert ours[0x800:] == theirs[0x800:]      # raw data region
    assert ours[:8] == theirs[:8]               # signature


def test_slab_writer_threads_and_order(tmp_path):
    """chromatin's background writer: slabs of one submission go out on several threads (positional writes),
    submissions stay ordered, an ordered call closes the file after its slabs, errors surface to the producer."""
    from expecto_b200.chromatin import SlabWriter

    p = str(tmp_path / "bg.h5")
    n = 64
    a = np.random.default_rng(1).standard_normal((2 * n, 2002)).astype(np.float32)
    w = hdf5.H5Writer(p, {"diff": (2 * n, 2002), "ref": (2 * n, 2002)})
    sw = SlabWriter(2, n_threads=4)
    sink = lambda name, row0, arr: w.write_rows(name, row0, arr)
    for k in range(0, n, 16):  # forward rows at v, reverse-complement rows at n + v (chromatin.py:170-171 order)
        s = sw.next_set()
        tasks = [(name, r0, a[r0:r0 + 16]) for name in ("diff", "ref") for r0 in (k, n + k)]
        sw.write(sink, tasks[:2])
        sw.submit(s, sink, tasks[2:])
    sw.call(w.close)
    sw.close()
    assert sw.bytes == 2 * a.nbytes and sw.busy_s > 0
    with hdf5.File(p) as f:
        assert np.array_equal(np.asarray(f["diff"]), a) and np.array_equal(np.asarray(f["ref"]), a)

    sw = SlabWriter(1)
    s = sw.next_set()
    sw.submit(s, lambda *t: (_ for _ in ()).throw(OSError("disk full")), [("diff", 0, a[:1])])
    with pytest.raises(OSError):
        sw.wait_free(s)
    with pytest.raises(OSError):
        sw.close()


def test_positional_row_reads(tmp_path):
    """predict's loader reads row ranges of a dataset straight into (pinned) buffers from several threads."""
    p = str(tmp_path / "r.h5")
    a = np.random.default_rng(2).standard_normal((50, 2002)).astype(np.float32)
    hdf5.write_datasets(p, {"ref": a, "alt": a * 2})
    with hdf5.File(p) as f:
        buf = np.zeros((20, 2002), np.float32)
        f["alt"].read_rows_into(buf, 7, 19)
        assert np.array_equal(buf[:12], a[7:19] * 2) and not buf[12:].any()
        f["ref"].read_rows_into(buf, 30, 50)
        assert np.array_equal(buf, a[30:50])
