#!/usr/bin/env python
"""Per-layer numerical check and timing of the CUDA engine on a real B200 (development tool).

Each stage runs in its own subprocess with a timeout so that a trap in one configuration does not
take the others down.  Usage on the GPU box:  python tools/gpu_check.py all
"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def ref_layers(x_onehot, sd, acts, hid, probs, dev):
    """Layer-isolated fp32 torch references: layer i is computed from OUR layer i-1 output."""
    import torch
    import torch.nn.functional as F

    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    W = {k: torch.from_numpy(v).to(dev) for k, v in sd.items()}
    convs = ["model.0.0", "model.0.2", "model.0.6", "model.0.8", "model.0.12", "model.0.14"]
    valid_in = [2000, 1993, 496, 489, 120, 113]
    pool = [False, True, False, True, False, False]
    out = {}
    n = x_onehot.shape[0]
    for i in range(6):
        if i == 0:
            xin = x_onehot.float()
        else:
            xin = acts[i][:, :valid_in[i], :].float().transpose(1, 2).contiguous()
        y = F.relu(F.conv1d(xin, W[convs[i] + ".weight"][:, :, 0, :], W[convs[i] + ".bias"]))
        if pool[i]:
            y = F.max_pool1d(y, 4, 4)
        out[i + 1] = y.transpose(1, 2).contiguous()  # [n, L, C]
    flat = acts[6].float().transpose(1, 2).reshape(n, -1)  # channel-major flatten (Beluga.py:42)
    h = F.relu(flat @ W["model.1.2.1.weight"].t() + W["model.1.2.1.bias"])
    out[7] = h
    z = hid[:, 0, :2003].float() @ W["model.1.4.1.weight"].t() + W["model.1.4.1.bias"]
    out[8] = torch.sigmoid(z)
    # end-to-end fp32 chain
    y = x_onehot.float()
    for i in range(6):
        y = F.relu(F.conv1d(y, W[convs[i] + ".weight"][:, :, 0, :], W[convs[i] + ".bias"]))
        if pool[i]:
            y = F.max_pool1d(y, 4, 4)
    y = F.relu(y.reshape(n, -1) @ W["model.1.2.1.weight"].t() + W["model.1.2.1.bias"])
    out["e2e"] = torch.sigmoid(y @ W["model.1.4.1.weight"].t() + W["model.1.4.1.bias"])
    return out


def stage_layers(dtype, cg, n):
    import numpy as np
    import torch
    from expecto_b200.engine import BelugaEngine, _ACT_VALID
    from expecto_b200 import synth

    dev = torch.device("cuda", 0)
    sd = synth.make_beluga_weights(0, round_dtype=dtype)
    eng = BelugaEngine(0, dtype=dtype, cta_group=cg, capacity=max(n, 8))
    eng.load_state_dict(sd)
    x = torch.from_numpy(synth.random_onehot(n, seed=2)).to(dev)
    x[0, :, 100:110] = 0  # some N columns
    probs = eng.forward_onehot(x)
    torch.cuda.synchronize()
    acts = {i: eng.debug_activation(i, n) for i in range(0, 8)}
    torch.cuda.synchronize()
    ref = ref_layers(x, sd, acts, acts[7], probs, dev)
    res = {"stage": "layers", "dtype": dtype, "cg": cg, "n": n}
    ok = True
    tol = 3e-3 if dtype == "f16" else 2e-2
    for i in range(1, 7):
        v = _ACT_VALID[i]
        ours = acts[i][:, :v, :].float()
        r = ref[i]
        err = (ours - r).abs().max().item()
        scale = r.abs().max().item()
        res[f"L{i}"] = {"max_abs_err": err, "ref_max": scale, "rel": err / max(scale, 1e-9),
                        "ours_nonzero": float((ours != 0).float().mean().item())}
        ok &= err <= tol * max(scale, 1.0)
    ours = acts[7][:, 0, :2003].float()
    err = (ours - ref[7]).abs().max().item()
    res["FC1"] = {"max_abs_err": err, "ref_max": ref[7].abs().max().item(),
                  "pad_max": acts[7][:, 0, 2003:].float().abs().max().item()}
    ok &= err <= tol * max(ref[7].abs().max().item(), 1.0)
    err = (probs - ref[8]).abs().max().item()
    res["FC2"] = {"max_abs_err": err}
    ok &= err <= 1e-3
    e2e = (probs - ref["e2e"]).abs()
    res["e2e"] = {"max_abs_err": e2e.max().item(), "mean_abs_err": e2e.mean().item(),
                  "p_mean": ref["e2e"].mean().item(), "p_max": ref["e2e"].max().item(),
                  "frac_gt_half": (ref["e2e"] > 0.5).float().mean().item()}
    res["ok"] = bool(ok)
    print(json.dumps(res))
    return 0 if ok else 1


def stage_time(dtype, cg, n, iters=5):
    import ctypes as C
    import torch
    from expecto_b200.engine import BelugaEngine
    from expecto_b200 import synth
    from expecto_b200._lib import lib, check

    dev = torch.device("cuda", 0)
    sd = synth.make_beluga_weights(0, round_dtype=dtype)
    eng = BelugaEngine(0, dtype=dtype, cta_group=cg, capacity=n)
    eng.load_state_dict(sd)
    x = torch.from_numpy(synth.random_onehot(min(n, 256), seed=2)).to(dev)
    x = x.repeat((n + x.shape[0] - 1) // x.shape[0], 1, 1)[:n].contiguous()
    out = torch.empty((n, 2002), dtype=torch.float32, device=dev)
    for _ in range(2):
        eng.forward_onehot(x, out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        eng.forward_onehot(x, out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    res = {"stage": "time", "dtype": dtype, "cg": cg, "n": n, "ms": ms, "windows_per_s": n / ms * 1e3,
           "tflops": n * 7.803176492e9 / ms / 1e9}
    flops = [40816640, 3253862400, 1201766400, 1776844800, 555417600, 694681600, 271767040, 8020012]
    st = torch.cuda.current_stream(dev).cuda_stream
    per = {}
    for layer in range(8):
        check(lib().bx_debug_run_layer(eng._ctx, layer, n, out.data_ptr(), st))
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            check(lib().bx_debug_run_layer(eng._ctx, layer, n, out.data_ptr(), st))
        e1.record()
        torch.cuda.synchronize()
        lms = e0.elapsed_time(e1) / iters
        per[f"layer{layer}"] = {"ms": lms, "tflops": n * flops[layer] / lms / 1e9}
    res["layers"] = per
    print(json.dumps(res))
    return 0


def main():
    if len(sys.argv) >= 2 and sys.argv[1] == "stage":
        kind, dtype, cg, n = sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5])
        sys.exit(stage_layers(dtype, cg, n) if kind == "layers" else stage_time(dtype, cg, n))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    plan = [
        ("layers", "f16", 1, 3), ("layers", "f16", 2, 3), ("layers", "f16", 1, 70), ("layers", "f16", 2, 70),
        ("layers", "bf16", 2, 5), ("time", "f16", 1, 512), ("time", "f16", 2, 512), ("time", "f16", 2, 2048),
    ]
    if len(sys.argv) >= 2 and sys.argv[1] != "all":
        plan = [p for p in plan if p[0] == sys.argv[1]]
    rc_all = 0
    with open(os.path.join(ROOT, "gpurun_out", "gpu_check.log"), "a") as log:
        for kind, dtype, cg, n in plan:
            t0 = time.time()
            try:
                r = subprocess.run([sys.executable, __file__, "stage", kind, dtype, str(cg), str(n)],
                                   capture_output=True, text=True, timeout=420)
                out, err, rc = r.stdout, r.stderr, r.returncode
            except subprocess.TimeoutExpired as e:
                out, err, rc = (e.stdout or b"").decode() if isinstance(e.stdout, bytes) else (e.stdout or ""), "TIMEOUT", 124
            line = f"== {kind} {dtype} cg={cg} n={n} rc={rc} {time.time()-t0:.1f}s\n{out}\n{err[-3000:]}\n"
            print(line)
            log.write(line)
            log.flush()
            rc_all |= rc
    sys.exit(1 if rc_all else 0)


if __name__ == "__main__":
    main()
