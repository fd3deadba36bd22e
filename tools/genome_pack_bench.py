#!/usr/bin/env python
"""Cold start from a FASTA: host packer (numpy) vs the device packer (bx_fasta_* kernels), same file, arrays compared
word for word.  Also times the kernels alone (text already in HBM) for their HBM GB/s:
algorithmic bytes per base = 2 x 1 (text read by the count and the pack pass) + 0.375 (packed planes written) = 2.375 B.   Usage: python tools/genome_pack_bench.py [Mbases=400]"""
import ctypes as C
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch

    from expecto_b200._lib import check, lib
    from expecto_b200.genome import PackedGenome

    mb = int(sys.argv[1]) if len(sys.argv) > 1 else 400
    n, W = mb * 1_000_000, 50
    rng = np.random.default_rng(0)
    tmp = tempfile.mkdtemp()
    fa = os.path.join(tmp, "g.fa")
    with open(fa, "wb") as f:
        for c in range(4):
            seq = np.frombuffer(b"ACGTACGTACGTACGTacgtN", dtype=np.uint8)[rng.integers(0, 21, size=n // 4)]
            lines = np.empty((seq.size // W, W + 1), dtype=np.uint8)
            lines[:, :W] = seq.reshape(-1, W)
            lines[:, W] = 10
            f.write(f">chr{c + 1}\n".encode())
            f.write(lines.tobytes())
    size = os.path.getsize(fa)
    torch.cuda.init()
    torch.zeros(1, device="cuda")
    t = time.perf_counter(); host = PackedGenome.from_fasta(fa, cache=False); t_host = time.perf_counter() - t
    t = time.perf_counter(); dev = PackedGenome.from_fasta(fa, cache=False, device=0); t_dev = time.perf_counter() - t
    t = time.perf_counter(); dev = PackedGenome.from_fasta(fa, cache=False, device=0); t_dev2 = time.perf_counter() - t
    same = (np.array_equal(host.bases2, dev.bases2) and np.array_equal(host.nmask, dev.nmask)
            and host.offsets == dev.offsets and host.lengths == dev.lengths and np.array_equal(host.invalid, dev.invalid))
    # kernels alone: one contig-sized body already in HBM
    L = lib()
    text = torch.from_numpy(np.fromfile(fa, dtype=np.uint8)).cuda()
    pad = torch.full((32,), 10, dtype=torch.uint8, device="cuda")
    text = torch.cat([text, pad])
    b0, b1 = 6, size
    nb = int(L.bx_fasta_blocks(b0, b1))
    off = torch.empty(nb + 1, dtype=torch.int64, device="cuda")
    nout = C.c_int64(0)
    st = torch.cuda.current_stream().cuda_stream
    check(L.bx_fasta_count(text.data_ptr(), b0, b1, off.data_ptr(), C.byref(nout), st))
    nbases = int(nout.value)
    bases2 = torch.zeros((nbases + 64) // 16 + 8, dtype=torch.int32, device="cuda")
    nmask = torch.zeros((nbases + 64) // 32 + 8, dtype=torch.int32, device="cuda")
    inv = torch.empty(1 << 20, dtype=torch.int64, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    reps = 5
    for it in range(reps + 1):
        if it == 1:
            torch.cuda.synchronize(); e[0].record()
        check(L.bx_fasta_count(text.data_ptr(), b0, b1, off.data_ptr(), C.byref(nout), st))
    e[1].record()
    for it in range(reps):
        check(L.bx_fasta_pack(text.data_ptr(), b0, b1, off.data_ptr(), 0, bases2.data_ptr(), nmask.data_ptr(),
                              inv.data_ptr(), 1 << 20, cnt.data_ptr(), st))
    e[2].record(); torch.cuda.synchronize()
    ms_count, ms_pack = e[0].elapsed_time(e[1]) / reps, e[1].elapsed_time(e[2]) / reps
    bytes_alg = 2.375 * nbases
    print(json.dumps({"fasta_bytes": size, "bases": int(sum(host.lengths.values())), "identical": bool(same),
                      "host_packer_s": round(t_host, 3), "device_packer_s_first": round(t_dev, 3),
                      "device_packer_s": round(t_dev2, 3), "speedup": round(t_host / t_dev2, 1),
                      "kernels": {"bases": nbases, "count_scan_ms": round(ms_count, 3), "pack_ms": round(ms_pack, 3),
                                  "algorithmic_bytes": int(bytes_alg),
                                  "GBps": round(bytes_alg / ((ms_count + ms_pack) * 1e-3) / 1e9, 1)},
                      "hg19_estimate_s": {"host": round(t_host * 3.1e9 / size, 1), "device": round(t_dev2 * 3.1e9 / size, 1)}}))
    os.unlink(fa)


if __name__ == "__main__":
    main()
