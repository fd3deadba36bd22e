#!/usr/bin/env python
"""Write-throughput probe for the HDF5 output path: pwrite of 64 MiB slabs from pinned-like buffers with
1..8 threads, buffered vs O_DIRECT, into one file (disjoint offsets).  Prints one JSON line per setting."""
import json, mmap, os, sys, tempfile, threading, time

def run(path, n_threads, direct, total_mb=6144, blk_mb=64):
    flags = os.O_WRONLY | os.O_CREAT | os.O_TRUNC | (os.O_DIRECT if direct else 0)
    fd = os.open(path, flags, 0o644)
    blk = blk_mb << 20
    bufs = [mmap.mmap(-1, blk) for _ in range(n_threads)]  # page-aligned
    for b in bufs:
        b.write(os.urandom(1 << 20) * blk_mb)
    n_blk = total_mb // blk_mb
    def work(t):
        for i in range(t, n_blk, n_threads):
            os.pwrite(fd, bufs[t], i * blk)
    t0 = time.perf_counter()
    ths = [threading.Thread(target=work, args=(t,)) for t in range(n_threads)]
    [t.start() for t in ths]; [t.join() for t in ths]
    t1 = time.perf_counter()
    os.fsync(fd) if os.environ.get("FSYNC") else None
    t2 = time.perf_counter()
    os.close(fd)
    os.unlink(path)
    return {"threads": n_threads, "direct": direct, "GBps": round(total_mb / 1024 / (t1 - t0), 2), "fsync_s": round(t2 - t1, 2)}

if __name__ == "__main__":
    d = sys.argv[1] if len(sys.argv) > 1 else tempfile.gettempdir()
    os.system(f"df -h {d} | tail -1; grep -E 'MemTotal|MemAvailable|Dirty:' /proc/meminfo | tr '\\n' ' '; echo")
    for direct in (False, True):
        for nt in (1, 2, 4, 8):
            try:
                print(json.dumps(run(os.path.join(d, "bx_disk_probe.bin"), nt, direct)), flush=True)
            except OSError as e:
                print(json.dumps({"threads": nt, "direct": direct, "error": str(e)}), flush=True)
