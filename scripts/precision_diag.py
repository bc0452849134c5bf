"""Per-tensor error of one training step in every gemm mode against a float64 torch-autograd reference
(run under gpurun).  Prints relative L2 errors; asserts nothing."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import torch.nn.functional as F
from multidisttorch_b200.engine import VAEEngine
from oracle import vae_numpy as O

def rel(a, b):
    a = np.asarray(a, np.float64).ravel(); b = np.asarray(b, np.float64).ravel()
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)

H = int(os.environ.get("H", "400")); B = int(os.environ.get("B", "128"))
dims = O.Dims(H=H); Z = 20
p0 = O.init_params(1234, dims)
x, eps = O.make_batch(4321, B, dims)
# float64 reference
P = {k: torch.tensor(v, dtype=torch.float64, requires_grad=True) for k, v in p0.items()}
X = torch.tensor(x, dtype=torch.float64); EPS = torch.tensor(eps, dtype=torch.float64)
h1 = F.relu(X @ P["W1"].T + P["b1"]); ml = h1 @ P["W2x"].T + P["b2x"]; mu, lv = ml[:, :Z], ml[:, Z:]
z = mu + EPS * torch.exp(0.5 * lv); h3 = F.relu(z @ P["W3"].T + P["b3"]); r = torch.sigmoid(h3 @ P["W4"].T + P["b4"])
L = F.binary_cross_entropy(r, X, reduction="sum") - 0.5 * torch.sum(1 + lv - mu ** 2 - lv.exp())
for t in (h1, ml, z, h3, r): t.retain_grad()
L.backward()
ref_act = {"h1": h1, "ml": ml, "z": z, "h3": h3, "recon": r}
print(f"H={H} B={B} loss64={L.item():.6f}")
_, g32, _ = O.loss_and_grads(p0, x, eps, dims)
print("numpy-fp32 oracle grads:", {k: f"{rel(g32[k], P[k].grad.numpy()):.1e}" for k in O.SEGMENTS})
for mode in sys.argv[1:] or ["fp32", "tf32x3", "tf32"]:
    eng = VAEEngine(hidden=H, max_batch=B, gemm_mode=mode, use_graph=False)
    eng.load_segments(p0)
    loss = eng.train_step(torch.from_numpy(x).cuda(), torch.from_numpy(eps).cuda()).item()
    acts = {k: f"{rel(eng.workspace(k, B).cpu().numpy(), v.detach().numpy()):.1e}" for k, v in ref_act.items()}
    da4 = eng.workspace("da4", B).cpu().numpy(); a4g = (r.grad * r * (1 - r)).detach().numpy()
    g = eng.segments_to_numpy(eng.grads)
    print(f"[{mode}] loss rel {abs(loss - L.item()) / L.item():.1e} acts {acts} da4 {rel(da4, a4g):.1e}")
    print(f"[{mode}] grads", {k: f"{rel(g[k], P[k].grad.numpy()):.1e}" for k in O.SEGMENTS})
    eng.close()
