"""Cycle stamps of one CTA's pipeline events in score_tma_kernel (library built with BX_NVCC_FLAGS=-DBX_SC_TRACE):
per chunk the two generator warps traced (first / last), per K-block the MMA issuer, per chunk the x producer.
Prints, for a window of chunks in the steady state, where each role's time goes."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from expecto_b200 import _lib
from expecto_b200.transform import ScoreModels, decay_score_variant, chromatin_shifts, padded_rows

n, m = 74 * 256, int(os.environ.get("PROBE_MODELS", 218))
rng = np.random.default_rng(0)
dev = torch.device("cuda:0")
x = padded_rows((9, n), dev); x.uniform_()
dist = torch.from_numpy(rng.integers(-40000, 40000, size=n).astype(np.float64)).to(dev)
strand = torch.from_numpy(np.where(np.arange(n) % 2 == 0, 1, -1).astype(np.int8)).to(dev)
shifts = torch.tensor(chromatin_shifts(800), dtype=torch.int32, device=dev)
models = ScoreModels((rng.standard_normal((m, 20020)) * 0.01).astype(np.float32), rng.standard_normal(m).astype(np.float32), dev)
for _ in range(3):
    decay_score_variant(x, dist, strand, models, shifts=shifts)
torch.cuda.synchronize()
tr = np.zeros((4, 512, 8), dtype=np.int64)
L = _lib.lib()
L.bx_debug_score_trace.argtypes = [C.c_void_p]
assert L.bx_debug_score_trace(tr.ctypes.data) == 0
t0 = tr[0, 0, 0]
g0, g1, mm, xp = (tr[i] - t0 for i in range(4))
lo, hi = 100, 228   # chunks 100..227 = super-blocks 25..56
per = (g0[hi, 0] - g0[lo, 0]) / (hi - lo)
print(f"models {m}: chunk period {per:.0f} cycles")
for name, g in (("generator warp 0", g0), ("generator warp 15", g1)):
    d = g[lo:hi]
    print(f"{name}: wait x {np.mean(d[:,1]-d[:,0]):.0f}  LDS+arithmetic {np.mean(d[:,2]-d[:,1]):.0f}  wait emptyA {np.mean(d[:,3]-d[:,2]):.0f}  "
          f"late K stores {np.mean(d[:,4]-d[:,3]):.0f}  selects+stores {np.mean(d[:,6]-d[:,4]):.0f}  fence {np.mean(d[:,7]-d[:,6]):.0f}  "
          f"syncwarp+arrive {np.mean(d[:,5]-d[:,7]):.0f}")
k0, k1 = lo * 5 // 4, hi * 5 // 4
d = mm[k0:k1]
print(f"MMA issuer per K-block: wait A {np.mean(d[:,1]-d[:,0]):.0f}  wait B {np.mean(d[:,2]-d[:,1]):.0f}  issue+commit {np.mean(d[:,3]-d[:,2]):.0f}  "
      f"period {(mm[k1,0]-mm[k0,0])/(k1-k0):.0f}")
d = xp[lo:hi]
print(f"x producer per chunk: wait emptyX {np.mean(d[:,1]-d[:,0]):.0f}")
# latency from x issue to generator's fullX pass, and lead of x issue over its use
print(f"x box: issue -> first generator read {np.mean(g0[lo:hi,1]-xp[lo:hi,1]):.0f} cycles")
np.save(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out", "score_trace.npy"), tr)
if os.environ.get("TRACE_ROWS"):
    for ch in range(lo, lo + 6):
        print(ch, "gen0", (g0[ch] - g0[lo, 0]).tolist(), "x", (xp[ch] - g0[lo, 0])[:2].tolist())
    for kb in range(k0, k0 + 8):
        print(kb, "mma", (mm[kb] - g0[lo, 0])[:4].tolist())
