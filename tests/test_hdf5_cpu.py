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
def test_reads_reference_example_and_writer_is_byte_identical_to_h5py(tmp_path):
    """example/example.vcf.shift_0.diff.h5 was written by h5py (upstream chromatin.py: one ``pred`` dataset).  Writing
    the same array here must reproduce the file byte for byte — metadata skeleton (superblock, root header, B-tree,
    heap, object header, symbol node, in libhdf5's allocation order) and data — except the 4-byte modification time
    of the dataset's mtime message at 0x3bc."""
    with hdf5.File(EXAMPLE) as f:
        assert list(f.keys()) == ["pred"]
        pred = np.asarray(f["pred"])
    assert pred.shape == (20, 2002) and pred.dtype == np.float32
    assert abs(float(pred.min()) + 0.111) < 2e-3 and abs(float(pred.max()) - 0.191) < 2e-3
    p = str(tmp_path / "pred.h5")
    hdf5.write_datasets(p, {"pred": pred})
    ours, theirs = open(p, "rb").read(), open(EXAMPLE, "rb").read()
    assert len(ours) == len(theirs)
    differing = [i for i in range(len(ours)) if ours[i] != theirs[i]]
    assert set(differing) <= {0x3BC, 0x3BD, 0x3BE, 0x3BF}, [hex(i) for i in differing[:20]]


GOLDEN_SKELETON = os.path.join(os.path.dirname(__file__), "golden", "h5_pred_skeleton.bin")


def test_pred_skeleton_matches_committed_h5py_bytes(tmp_path):
    """The same byte-equality against the first 0x800 bytes of the h5py-written example, committed as a fixture
    (tests/golden/h5_pred_skeleton.bin) so it also runs where /root/reference does not exist."""
    theirs = open(GOLDEN_SKELETON, "rb").read()
    assert len(theirs) == 0x800
    p = str(tmp_path / "pred.h5")
    hdf5.write_datasets(p, {"pred": np.zeros((20, 2002), np.float32)})
    ours = open(p, "rb").read()[:0x800]
    differing = [i for i in range(0x800) if ours[i] != theirs[i]]
    assert set(differing) <= {0x3BC, 0x3BD, 0x3BE, 0x3BF}, [hex(i) for i in differing[:20]]


@pytest.mark.skipif(not os.path.exists(EXAMPLE), reason="reference example files only exist in the build container")
def test_spec_validator_accepts_the_h5py_written_examples():
    """The independent validator (tests/h5spec.py, written from the format specification) is itself pinned on files
    libhdf5 wrote."""
    import glob

    import h5spec

    for p in sorted(glob.glob(os.path.join(os.path.dirname(EXAMPLE), "*.diff.h5"))):
        info = h5spec.validate(p)
        assert info["pred"]["shape"] == (20, 2002) and info["pred"]["address"] == 0x800


def test_three_dataset_file_passes_the_spec_validator(tmp_path):
    """chromatin.py:282-286's diff/ref/alt form: superblock v0 -> root symbol table -> TREE keys -> SNOD entries sorted
    by heap name -> v1 object headers (message sizes, alignment, NIL fill) -> layout v3 address/size inside EOF, no
    overlaps; and the metadata sits where libhdf5's allocation order puts it (header, symbol node, header, header)."""
    import h5spec

    rng = np.random.default_rng(3)
    arrs = {k: rng.standard_normal((14, 2002)).astype(np.float32) for k in ("diff", "ref", "alt")}
    p = str(tmp_path / "snps.shift_-200.diff.h5")
    hdf5.write_datasets(p, arrs)
    info = h5spec.validate(p)
    assert list(info) == ["alt", "diff", "ref"]                      # symbol node order
    nb = 14 * 2002 * 4
    assert [info[k]["header"] for k in ("diff", "ref", "alt")] == [0x320, 0x578, 0x688]
    assert [info[k]["address"] for k in ("diff", "ref", "alt")] == [0x800, 0x800 + nb, 0x800 + 2 * nb]
    for k, v in arrs.items():
        assert np.array_equal(h5spec.read(p, k), v)
    # row-slab writers attached to an existing file see the same layout
    with hdf5.H5Writer(p, {k: (14, 2002) for k in ("diff", "ref", "alt")}, create=False) as w:
        w.write_rows("ref", 3, arrs["alt"][3:5])
    assert np.array_equal(h5spec.read(p, "ref")[3:5], arrs["alt"][3:5])


def test_other_forms_pass_the_spec_validator(tmp_path):
    """float64 + fixed-length strings (geuvadis_predict_consensus.py:127-133), an empty dataset (address UNDEF, as
    libhdf5 leaves unallocated storage), eight datasets with long names (larger heap, second metadata block)."""
    import h5spec

    p = str(tmp_path / "c.h5")
    ids = np.array([b"ENSG1:1-5", b"ENSG22:7-9", b"x"])
    hdf5.write_datasets(p, {"chromatin_preds": np.arange(2 * 5 * 2002, dtype=np.float64).reshape(2, 5, 2002),
                            "record_ids": ids})
    info = h5spec.validate(p)
    assert info["chromatin_preds"]["dtype"] == "<f8" and info["record_ids"]["dtype"] == "S10"
    assert np.array_equal(h5spec.read(p, "record_ids"), ids)
    p = str(tmp_path / "e.h5")
    hdf5.write_datasets(p, {"diff": np.zeros((0, 2002), np.float32), "ref": np.zeros((0, 2002), np.float32)})
    info = h5spec.validate(p)
    assert info["diff"]["shape"] == (0, 2002) and info["diff"]["address"] == h5spec.UNDEF
    with hdf5.File(p) as f:
        assert np.asarray(f["ref"]).shape == (0, 2002)
    p = str(tmp_path / "l.h5")
    arrs = {f"a_rather_long_dataset_name_{i}": np.full((3, 4), i, np.float32) for i in range(8)}
    hdf5.write_datasets(p, arrs)
    assert len(h5spec.validate(p)) == 8
    with hdf5.File(p) as f:
        for k, v in arrs.items():
            assert np.array_equal(np.asarray(f[k]), v)


def test_h5py_reads_our_files_when_available(tmp_path):
    """If a real h5py / libhdf5 is importable (it is not in the build image), it must read the writer's output."""
    h5py = pytest.importorskip("h5py")
    if not hasattr(h5py, "version"):
        pytest.skip("h5py here is the oracle's shim, not libhdf5")
    rng = np.random.default_rng(4)
    arrs = {k: rng.standard_normal((6, 2002)).astype(np.float32) for k in ("diff", "ref", "alt")}
    p = str(tmp_path / "x.h5")
    hdf5.write_datasets(p, arrs)
    with h5py.File(p, "r") as f:
        assert sorted(f.keys()) == ["alt", "diff", "ref"]
        for k, v in arrs.items():
            assert np.array_equal(f[k][...], v)


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
