"""GPU-side debugging of the tcgen05 GEMM: compares mdt_linear in tf32/tf32x3 against fp64 numpy per shape and
prints error statistics instead of asserting (run under gpurun)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from multidisttorch_b200 import engine as E

def dev(a): return torch.from_numpy(np.ascontiguousarray(a)).cuda()

shapes = [(128, 32, 32, "none"), (128, 64, 64, "none"), (128, 400, 784, "relu"), (128, 40, 400, "none"), (128, 400, 20, "relu"),
          (128, 784, 400, "sigmoid"), (48, 400, 784, "relu"), (1, 784, 400, "none"), (300, 400, 784, "none"), (128, 4096, 784, "none")]
for mode in sys.argv[1:] or ["tf32", "tf32x3"]:
    for (M, N, K, act) in shapes:
        rs = np.random.RandomState(M + N + K)
        x = rs.random_sample((M, K)).astype(np.float32)
        w = (rs.standard_normal((N, K)) / np.sqrt(K)).astype(np.float32)
        b = rs.standard_normal(N).astype(np.float32)
        try:
            y = E.linear(dev(x), dev(w), dev(b), act=act, gemm_mode=mode)
            torch.cuda.synchronize()
            y = y.cpu().numpy()
        except Exception as ex:
            print(mode, (M, N, K, act), "EXC", repr(ex)[:300]); break
        ref = x.astype(np.float64) @ w.astype(np.float64).T + b
        ref = {"none": ref, "relu": np.maximum(ref, 0), "sigmoid": 1 / (1 + np.exp(-ref))}[act]
        err = np.linalg.norm(y - ref) / np.linalg.norm(ref)
        bad = np.argwhere(np.abs(y - ref) > 1e-2 * (np.abs(ref) + 1e-3))
        print(mode, (M, N, K, act), f"relL2={err:.3e} nbad={len(bad)}", "first_bad=" + str(bad[:4].tolist()) if len(bad) else "")
        if len(bad) and M * N <= 128 * 64:
            print(" y[0,:8]", y[0, :8], "\n r[0,:8]", ref[0, :8])
