"""Debug probe: AdaBins training case under dropout, engine vs fp64 autograd through the oracle on the GPU with shared masks."""
import json, os, sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from helpers import GOLD
from objcavit_b200 import engine_train, ops
from objcavit_b200.config import builtin_args
from oracle import head as oh, synth
from oracle.cases import TRAIN_CASES, case_cfg, case_gt, case_inputs

cuda = torch.device("cuda:0")
name = sys.argv[1] if len(sys.argv) > 1 else "train_adabins_small"
which = set(int(c) for c in (sys.argv[2] if len(sys.argv) > 2 else "0123"))
case = TRAIN_CASES[name]
args = builtin_args(case["args"])
sd = synth.make_state_dict(json.load(open(os.path.join(GOLD, f"manifest_{case['args']}.json"))), seed=case["seed"])
feats, objs, boxes = case_inputs(case)
gt, mask = case_gt(case, args)
p, seed = 0.1, 4242

# selective sites: patch ops so that site kinds not in `which` are identity
real_drop, real_core, real_core_bwd = ops.dropout, ops.mha_core, ops.mha_core_bwd
kind_of = {}
class Ctx(engine_train.DropCtx):
    def site(self):
        s = super().site()
        kind_of[s] = (self.count - 1) % 4
        return s
def drop_sel(x, p_, seed_, residual=None, out=None):
    if kind_of.get(seed_, 0) in which:
        return real_drop(x, p_, seed_, residual=residual, out=out)
    r = x if residual is None else x + residual
    if out is not None:
        out.copy_(r); return out
    return r.clone()
ops.dropout = drop_sel
ops.mha_core = lambda q, k, v, m=None, h=4, drop=None: real_core(q, k, v, m, h, drop=drop if 0 in which else None)
ops.mha_core_bwd = lambda *a, drop=None, **kw: real_core_bwd(*a, drop=drop if 0 in which else None, **kw)

P = {k: v.to(cuda) for k, v in sd.items()}
G = engine_train.Touched({k: torch.zeros_like(v) for k, v in P.items()})
cfg = dict(case_cfg(case, args), train=True)
lo, hi = float(args[args.basic.dataset].min_depth), float(args[args.basic.dataset].max_depth)
ecfg = dict(kind="adabins", min_depth=lo, max_depth=hi) if objs is None else dict(cfg, kind="graphbins", min_depth=lo, max_depth=hi)
dev = lambda l: None if l is None else [None if t is None else t.to(cuda) for t in l]
depth, edges, tape = engine_train.forward(P, ecfg, feats.to(cuda), dev(objs), dev(boxes), drop=Ctx(p, seed))
l1, dd = ops.silog(depth, gt.to(cuda), mask.to(cuda), want_grad=True, grad_scale=1.0)
l2, de = ops.bins_chamfer(edges, gt.to(cuda), mask.to(cuda), want_grad=True, grad_scale=0.1)
engine_train.backward(P, G, tape, dd, de)

ctx2 = Ctx(p, seed)
def hook(t):
    s = ctx2.site()
    if kind_of[s] not in which:
        return t
    return t * real_drop(torch.ones(t.shape, device=cuda), p, s).to(t.dtype)
sdg = {k: v.to(cuda).double().requires_grad_(True) for k, v in sd.items()}
oh.DROPOUT_HOOK = hook
d64 = lambda l: None if l is None else [None if t is None else t.to(cuda).double() for t in l]
od, oe, _, _ = oh.head_forward(sdg, cfg, feats.to(cuda).double(), d64(objs), d64(boxes))
oh.DROPOUT_HOOK = None
(oh.silog_loss(od, gt.to(cuda).double(), mask.to(cuda)) + 0.1 * oh.bins_chamfer_loss(oe, gt.to(cuda).double(), mask.to(cuda))).backward()
print("forward depth err", float((depth - od.float()).abs().max() / od.abs().max()))
gmax = max(float(v.grad.abs().max()) for v in sdg.values() if v.grad is not None)
rows = []
for k, v in sdg.items():
    if v.grad is None: continue
    err = float((G[k].double() - v.grad).abs().max()) / max(float(v.grad.abs().max()), 1e-3 * gmax)
    rows.append((err, k))
rows.sort(reverse=True)
for err, k in rows[:8]:
    print(f"{err:.2e} {k}")
k = rows[0][1]
e = (G[k].double() - sdg[k].grad).abs().flatten()
top = torch.topk(e, 6)
print(k, "top errors", top.values.tolist(), "idx", top.indices.tolist(), "n>1e-3*max", int((e > 1e-3 * sdg[k].grad.abs().max()).sum()), "of", e.numel())
print("mine", G[k].flatten()[top.indices].tolist()); print("ref ", sdg[k].grad.flatten()[top.indices].tolist())
