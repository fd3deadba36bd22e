"""One launch of the tensor-core scoring kernel (for ncu captures): 74 x 256 variants, 218 models."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from expecto_b200.transform import ScoreModels, decay_score_variant, chromatin_shifts

n, m = 74 * 256 * int(sys.argv[1] if len(sys.argv) > 1 else 1), int(os.environ.get('PROBE_MODELS', 218))
n = int(os.environ.get('PROBE_VARIANTS', n))
rng = np.random.default_rng(0)
dev = torch.device("cuda:0")
pitch = int(sys.argv[2]) if len(sys.argv) > 2 else 2002
x = torch.rand((9, n, pitch), dtype=torch.float32, device=dev)[..., :2002]
dist = torch.from_numpy(rng.integers(-40000, 40000, size=n).astype(np.float64)).to(dev)
st_h = np.where(np.arange(n) % 2 == 0, 1, -1).astype(np.int8)
if os.environ.get("PROBE_SORTED"):       # rows grouped by side of the TSS (transform.side_order)
    from expecto_b200.transform import side_order
    o = side_order(dist.cpu().numpy(), st_h)
    dist = dist[torch.from_numpy(o).to(dev)]
    st_h = st_h[o]
strand = torch.from_numpy(st_h).to(dev)
shifts = torch.tensor(chromatin_shifts(800), dtype=torch.int32, device=dev)
models = ScoreModels((rng.standard_normal((m, 20020)) * 0.01).astype(np.float32), rng.standard_normal(m).astype(np.float32), dev)
models.prepared()
for _ in range(3):
    out = decay_score_variant(x, dist, strand, models, shifts=shifts)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
iters = int(os.environ.get('PROBE_ITERS', 5))
for _ in range(iters):
    out = decay_score_variant(x, dist, strand, models, shifts=shifts)
e1.record()
torch.cuda.synchronize()
print("pitch", pitch, "ms per launch", e0.elapsed_time(e1) / iters, "variants", n)
