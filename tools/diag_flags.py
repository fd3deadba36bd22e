#!/usr/bin/env python
"""Timing diagnostics: per-layer times with parts of the GEMM pipeline switched off (BX_DEBUG_FLAGS:
1 = no B TMA loads, 2 = no A TMA loads, 4 = no epilogue work).  Results are WRONG by construction; only the
times matter — and only as CYCLES: with flag 4 no layer writes its output, so the next layer multiplies zeros, the
tensor pipe draws far less power and the clock rises by ~20 % (tools/diag_epilogue_cycles.sh: conv2 5.78 M vs
5.69 M cycles, 1.44 vs 1.76 GHz).  Operand data, not the epilogue, moves the clock.  Usage: python tools/diag_flags.py [n]"""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

def stage(n):
    import torch
    from expecto_b200 import synth
    from expecto_b200.engine import BelugaEngine
    from expecto_b200._lib import lib, check
    eng = BelugaEngine(0, dtype="f16", cta_group=int(os.environ.get("BX_CG", "2")), capacity=n)
    eng.load_state_dict({k: (v * 0 + 0.01).astype("float32") for k, v in synth.make_beluga_weights(0, None).items()} if False else synth.make_beluga_weights(0, "f16"))
    x = torch.from_numpy(synth.random_onehot(64, seed=2)).cuda().repeat(n // 64, 1, 1).contiguous()
    out = torch.empty((n, 2002), dtype=torch.float32, device="cuda")
    eng.forward_onehot(x, out)
    torch.cuda.synchronize()
    st = torch.cuda.current_stream().cuda_stream
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    res = {}
    for layer in range(8):
        for _ in range(2):
            check(lib().bx_debug_run_layer(eng._ctx, layer, n, out.data_ptr(), st))
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5):
            check(lib().bx_debug_run_layer(eng._ctx, layer, n, out.data_ptr(), st))
        e1.record()
        torch.cuda.synchronize()
        res[layer] = round(e0.elapsed_time(e1) / 5, 4)
    print(json.dumps({"flags": int(os.environ.get("BX_DEBUG_FLAGS", "0")), "n": n, "ms": res}))

if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "stage":
        stage(int(sys.argv[2]))
    else:
        n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
        for flags in [int(f) for f in os.environ.get("BX_DIAG_LIST", "0,1,2,3,4,7").split(",")]:
            env = dict(os.environ, BX_DEBUG_FLAGS=str(flags))
            r = subprocess.run([sys.executable, __file__, "stage", str(n)], env=env, capture_output=True, text=True, timeout=300)
            print(r.stdout.strip() or r.stderr[-800:], flush=True)
