#!/usr/bin/env python
"""FC1 / FC2 / conv1 stage times (CUDA events per stage, bx_ctx_profile) at several batch sizes: python tools/dense_sweep.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from expecto_b200 import synth
from expecto_b200.engine import BelugaEngine
from expecto_b200._lib import STAGE_NAMES

def main():
    eng = BelugaEngine(0, dtype="f16", cta_group=2, capacity=4096)
    eng.load_state_dict(synth.make_beluga_weights(0, "f16"))
    res = {}
    for n in (64, 256, 1024, 4096):
        x = torch.from_numpy(synth.random_onehot(64, seed=2)).cuda().repeat(n // 64, 1, 1).contiguous()
        out = torch.empty((n, 2002), dtype=torch.float32, device="cuda")
        for _ in range(3):
            eng.forward_onehot(x, out)
        torch.cuda.synchronize()
        eng.profile(True)
        for _ in range(10):
            eng.forward_onehot(x, out)
        torch.cuda.synchronize()
        prof = eng.profile_read(reset=True)
        eng.profile(False)
        res[n] = {k: round(v[0] / max(v[1], 1), 4) for k, v in prof.items() if v[1]}
        res[n]["checksum"] = float(out.double().sum().item())
    print(json.dumps(res))

if __name__ == "__main__":
    main()
