#!/usr/bin/env python
"""Kernel 4 alone (bench.py's transform_rooflines): HBM GB/s of the decay-weighted shift reductions with every
argument resident on the device.  Used under ncu for the per-kernel duration / DRAM throughput capture."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench

if __name__ == "__main__":
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    peaks = bench.load_peaks() if hasattr(bench, "load_peaks") else {"hbm_gbs": 6552.0}
    print(json.dumps(bench.transform_rooflines(dev, peaks, iters=int(sys.argv[1]) if len(sys.argv) > 1 else 10)))
