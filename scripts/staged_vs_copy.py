"""How much of a step is the D2D staging copy?  Times train_step(x) against train_step_staged (x already in the
engine's input buffer).  Wall clock around many steps with a synchronize on both sides."""
import sys, time
import torch
sys.path.insert(0, ".")
from multidisttorch_b200.engine import VAEEngine

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
eng = VAEEngine(max_batch=B, gemm_mode="tf32x3")
eng.set_rng(1234)
NP = int(sys.argv[2]) if len(sys.argv) > 2 else 64
pool = torch.rand(NP, B, 784, device="cuda")
for i in range(50):
    eng.train_step(pool[i % NP], want_loss=False)
eng.synchronize()
for name in ("copy", "staged", "copy", "staged"):
    n = 4000
    eng.synchronize(); t0 = time.perf_counter()
    if name == "copy":
        for i in range(n):
            eng.train_step(pool[i % NP], want_loss=False)
    else:
        for i in range(n):
            eng.train_step_staged(B, want_loss=False)
    eng.synchronize(); dt = time.perf_counter() - t0
    print(f"{name}: {dt / n * 1e6:.2f} us/step")
