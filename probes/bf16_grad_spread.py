"""Probe: per-parameter agreement of the bf16 training step (tensor-core pixel + token path) with the fp32 parity path: relative
max error, cosine, norm ratio.  python probes/bf16_grad_spread.py [case]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from test_gpu_train_bf16 import _run

name = sys.argv[1] if len(sys.argv) > 1 else "train_learned_bbox_wh_small"
cuda = torch.device("cuda:0")
h32, o32, l32 = _run(name, "fp32", cuda)
h16, o16, l16 = _run(name, "bf16", cuda)
print("loss", l32, l16, "depth rel", float((o16.depth_pred - o32.depth_pred).abs().max() / o32.depth_pred.abs().max()))
g32 = {k: p.grad for k, p in h32.named_parameters()}
gmax = max(float(g.abs().max()) for g in g32.values() if g is not None)
rows = []
for k, p in h16.named_parameters():
    if g32[k] is None or p.grad is None or float(g32[k].abs().max()) == 0:
        continue
    a, r = p.grad.double().flatten(), g32[k].double().flatten()
    err = float((a - r).abs().max()) / max(float(r.abs().max()), 1e-2 * gmax)
    cos = float((a @ r) / (a.norm() * r.norm()).clamp_min(1e-300))
    rows.append((err, cos, float(a.norm() / r.norm()), k))
for err, cos, nr, k in sorted(rows, reverse=True)[:25]:
    print(f"{err:9.3e} cos {cos:7.4f} norm {nr:6.3f}  {k}")
print("min cosine", min(r[1] for r in rows), "params", len(rows))
