"""Measured errors of the CUDA path against the reference's golden vectors, per GEMM mode: the numbers the bounds in
tests/test_gpu_parity.py (TOL) are derived from (bound <= 3 x measured, rounded up).  Run on a B200:
    python scripts/measure_errors.py > profiles/r2_measured_errors.txt"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from multidisttorch_b200 import engine as E
from multidisttorch_b200.synth import fingerprint_error
from oracle import vae_numpy as O
from tests.golden_cases import PARAM_SEED, batch_seed, load

dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
print(f"{'mode':8s} {'case':18s} {'loss':>9s} {'grad0':>9s} {'paramN':>9s} {'mN':>9s} {'vN':>9s}")
for mode in ("fp32", "tf32x3", "tf32", "bf16"):
    for tag in ("step_b128", "step_b48", "step_b1", "step_b512", "step_h4096_b128"):
        g = load(tag)
        B, n = int(g["B"]), int(g["n_steps"])
        H = int(g["H"]) if "H" in g else 400
        if H != 400 and mode == "fp32":
            continue
        eng = E.VAEEngine(hidden=H, max_batch=B, gemm_mode=mode)
        eng.load_segments(O.init_params(PARAM_SEED, O.Dims(H=H)))
        e_loss = e_grad = 0.0
        for t in range(n):
            x, eps = O.make_batch(batch_seed(t), B, O.Dims(H=H))
            loss = float(eng.train_step(dev(x), dev(eps)).item())
            e_loss = max(e_loss, abs(loss - g["losses"][t]) / abs(g["losses"][t]))
            if t == 0:
                gr = eng.segments_to_numpy(eng.grads)
                e_grad = max(max(fingerprint_error(g, "grad0", k, gr[k])[:1]) for k in O.SEGMENTS)
        res = []
        for pre, buf in (("paramN", eng.params), ("mN", eng.exp_avg), ("vN", eng.exp_avg_sq)):
            s = eng.segments_to_numpy(buf)
            res.append(max(fingerprint_error(g, pre, k, s[k])[0] for k in O.SEGMENTS))
        print(f"{mode:8s} {tag:18s} {e_loss:9.2e} {e_grad:9.2e} {res[0]:9.2e} {res[1]:9.2e} {res[2]:9.2e}")
        eng.close()
