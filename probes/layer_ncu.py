"""One image-encoder layer forward + backward of the training step (B=8, S=221) for an ncu capture of its kernels."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from objcavit_b200 import ops
import objcavit_b200.engine_train as et

dev = torch.device("cuda:0")
B, S, E, F = 8, 221, 128, 1024
g = torch.Generator().manual_seed(0)
P = {}
pre = "l."
shapes = {"self_attn.in_proj_weight": (3 * E, E), "self_attn.in_proj_bias": (3 * E,), "self_attn.out_proj.weight": (E, E),
          "self_attn.out_proj.bias": (E,), "linear1.weight": (F, E), "linear1.bias": (F,), "linear2.weight": (E, F),
          "linear2.bias": (E,), "norm1.weight": (E,), "norm1.bias": (E,), "norm2.weight": (E,), "norm2.bias": (E,)}
for k, sh in shapes.items():
    P[pre + k] = (torch.randn(*sh, generator=g) * 0.05).to(dev)
G = {k: torch.zeros_like(v) for k, v in P.items()}
x = torch.randn(B, S, E, generator=g).to(dev)
drop = et.DropCtx(0.1, 1234)
for _ in range(3):
    out, sv = et.encoder_layer_fwd(P, pre, x, None, drop)
    dx = et.encoder_layer_bwd(P, G, pre, sv, torch.ones_like(out))
    et.join_side(dev)
torch.cuda.synchronize()
print("ok")
