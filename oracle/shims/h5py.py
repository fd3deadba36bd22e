"""Test-only stand-in for h5py (requirements.txt:7) so the UNMODIFIED reference scripts run without libhdf5.
'w' files capture ``create_dataset(name, data=...)`` into ``<path>.npz``; 'r' files expose ``f[name]`` with
``.shape`` and slicing (predict.py:174-190) from ``<path>.npz`` when present, else from a real HDF5 file through
the repo's dependency-free reader (that path also exercises our writer against the reference's reader code)."""
import os
import sys

import numpy as np


class File:
    def __init__(self, path, mode="r"):
        self.path, self.mode, self._w = path, mode, {}
        if mode == "r":
            if os.path.exists(path + ".npz"):
                self._r = dict(np.load(path + ".npz"))
            else:
                sys.path.insert(0, os.environ["BX_REPO_ROOT"])
                from expecto_b200 import hdf5
                f = hdf5.File(path)
                self._r = {k: np.asarray(f[k]) for k in f}

    def create_dataset(self, name, data=None, **kw):
        self._w[name] = np.asarray(data)

    def __getitem__(self, k):
        return self._r[k]

    def close(self):
        if self.mode != "r":
            np.savez(self.path + ".npz", **self._w)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
