"""Phase timing of CTA (0,0) of every tcgen05 GEMM in one un-graphed training step (MDT_TC05_DBG=1)."""
import os, sys, ctypes as C
os.environ["MDT_TC05_DBG"] = "1"
if os.environ.get("SOURCED", "0") == "1":
    os.environ.setdefault("MDT_STEPS_PER_GRAPH", "2")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from multidisttorch_b200 import engine as E, _lib
from oracle import vae_numpy as O
lib = _lib.load()
names = ["entry", "setup", "tma_issued", "first_data", "mma_done", "accum_seen", "drained", "epi_done", "exit"]
labels = ["fc1_fwd", "fc2x_reparam", "fc3_fwd", "fc4_bce", "fc4_dX", "fc4_dW", "dz_reparam_bwd", "fc3_dW", "fc2x_dX", "fc1_dW", "fc2x_dW"]
if os.environ.get("MDT_TC_LATENT", "0") == "0":   # the latent-side layers are CUDA-core kernels (no stamps): tcgen05 launches left
    # latent_*: stamps are entry / after wait / rows in smem / dots done / element phase done / (fwd: fc3 done) / exit
    labels = ["fc1_fwd", "latent_fwd", "fc4_bce", "fc4_dX", "fc4_dW", "fc3_dW", "latent_bwd", "fc2x_dW", "fc1_dW"]
    if not os.environ.get("MDT_DW3_EARLY"):
        labels = ["fc1_fwd", "latent_fwd", "fc4_bce", "fc4_dX", "fc4_dW", "latent_bwd", "fc3_dW", "fc2x_dW", "fc1_dW"]
    if os.environ.get("MDT_DW4_LATE"):
        labels = ["fc1_fwd", "latent_fwd", "fc4_bce", "fc4_dX", "latent_bwd", "fc3_dW", "fc2x_dW", "fc1_dW", "fc4_dW"]
out = (C.c_longlong * 512)()
for mode in sys.argv[1:] or ["tf32", "tf32x3"]:
    H = int(os.environ.get("H", "400")); B = int(os.environ.get("B", "128"))
    eng = E.VAEEngine(hidden=H, max_batch=B, gemm_mode=mode, use_graph=os.environ.get("GRAPH", "0") == "1")
    eng.load_segments(O.init_params(1, O.Dims(H=H)))
    x = torch.rand(B, 784, device="cuda")
    pool = torch.rand(8, B, 784, device="cuda")
    sourced = os.environ.get("SOURCED", "0") == "1"      # the pipelined steps of mdt_train_steps (no prologue kernel)
    step = (lambda: eng.train_steps(pool, 0, 2)) if sourced else (lambda: eng.train_step(x, None))
    for _ in range(4):        # an even count: single-step graphs come in two parities (Tc05Plan::w3par)
        step()
    eng.synchronize(); lib.mdt_debug_tc05_stamps(out, 1)
    if sourced:
        eng.train_steps(pool, 3, 2)        # two steps: the stamps below are those of the second one
        eng.synchronize()
        tmp = list(out)
        lib.mdt_debug_tc05_stamps(out, 1)
        full = list(out)
        n = len(labels)
        for i in range(n):
            for j in range(16):
                out[16 * i + j] = full[16 * (n + i) + j]
        t_prev_end = max(full[16 * i + 15] for i in range(n))
        print(f"   (previous step's last kernel left {(full[16 * n + 13] - t_prev_end) / 1e3:.1f} us before this step's first GEMM passed its wait)")
    else:
        step()
        eng.synchronize(); lib.mdt_debug_tc05_stamps(out, 1)
    print(f"== {mode} H={H} B={B}: cycles from CTA(0,0) entry (1 cycle = 0.51 ns)")
    for i, lab in enumerate(labels):
        st = list(out)[16 * i: 16 * i + 16]
        d = {n: st[j] - st[0] for j, n in enumerate(names)}
        if i == 0:
            t_first = st[13]
        print(f"  {lab:16s} go +{(st[13] - t_first) / 1e3:6.1f} us  end +{(st[15] - t_first) / 1e3:6.1f} us  dur {(st[15] - st[13]) / 1e3:5.1f} us |",
              " ".join(f"{n}={v}" for n, v in d.items() if n != "entry"), "| tile(0,0) last split: ticket +%.1f reduced +%.1f epilogue +%.1f us after go" % tuple((st[j] - st[13]) / 1e3 for j in (9, 10, 11)) if abs(st[9] - st[0]) > 10**8 else ("split done kb0..3: %s" % [st[j] - st[0] for j in (9, 10, 11, 12)]))
    eng.close()
