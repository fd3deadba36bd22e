#!/usr/bin/env python
"""Does conv1 (bound by the SM->L2 store rate, i.e. by the SM clock) speed up when it runs twice in a row after the
power-capped tensor layers?  Times each of three back-to-back conv1 launches that follow a hot conv2..fc2 sequence."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from expecto_b200 import synth
from expecto_b200.engine import BelugaEngine
from expecto_b200._lib import lib, check

n = 4096
eng = BelugaEngine(0, dtype="f16", cta_group=2, capacity=n)
eng.load_state_dict(synth.make_beluga_weights(0, "f16"))
x = torch.from_numpy(synth.random_onehot(64, seed=2)).cuda().repeat(n // 64, 1, 1).contiguous()
out = torch.empty((n, 2002), dtype=torch.float32, device="cuda")
for _ in range(3):
    eng.forward_onehot(x, out)
torch.cuda.synchronize()
st = torch.cuda.current_stream().cuda_stream
L = lib()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
res = []
for rep in range(6):
    for _ in range(2):                      # two hot forwards' worth of tensor layers
        for layer in range(1, 8):
            check(L.bx_debug_run_layer(eng._ctx, layer, n, out.data_ptr(), st))
    ev[0].record()
    for k in range(4):
        check(L.bx_debug_run_layer(eng._ctx, 0, n, out.data_ptr(), st))
        ev[k + 1].record()
    torch.cuda.synchronize()
    res.append([round(ev[k].elapsed_time(ev[k + 1]), 3) for k in range(4)])
print(json.dumps({"conv1_ms_back_to_back_after_hot_layers": res}))
