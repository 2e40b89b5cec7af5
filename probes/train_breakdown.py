"""Per-shape breakdown of one training step (CUDA events around every C-ABI call): which GEMM shapes the time goes to."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from objcavit_b200 import ops
from objcavit_b200.config import builtin_args
from objcavit_b200.modules import build_head
from objcavit_b200.trainer import HeadTrainer
from oracle import synth  # inputs only

prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
dev = torch.device("cuda:0")
B, H, W, NOBJ = 8, 208, 272, 32
head = build_head(builtin_args("nyu_ocv_learned_bbox_wh"), precision=prec).to(dev).train()
tr = HeadTrainer(head)
x = synth.make_features(B, H, W, seed=1).to(dev)
objs, boxes = synth.make_objects(B, H, W, NOBJ, seed=1)
objs, boxes = [o.to(dev) for o in objs], [b.to(dev) for b in boxes]
gt, mask = synth.make_gt(B, H, W, 10.0, seed=1)
gt, mask = gt.to(dev), mask.to(dev)
orig = ops.gemm_ex
def tagged(a, b, M, N, K, *args, **kw):
    ops.profile_tag = f"[{M}x{N}x{K} b{kw.get('batch', 1)} opb{kw.get('op_b', 0)}]"
    try:
        return orig(a, b, M, N, K, *args, **kw)
    finally:
        ops.profile_tag = ""
ops.gemm_ex = tagged
orig_lin = ops.linear
def tagged_lin(a, w, *args, **kw):
    ops.profile_tag = f"[{a.shape[0]}x{w.shape[0]}x{a.shape[1]}]"
    try:
        return orig_lin(a, w, *args, **kw)
    finally:
        ops.profile_tag = ""
ops.linear = tagged_lin
import objcavit_b200.engine_train as et
for _ in range(3):
    tr.step(x, list(objs), list(boxes), gt, mask)
torch.cuda.synchronize()
ops.profile_events = []
tr.step(x, list(objs), list(boxes), gt, mask)
torch.cuda.synchronize()
agg, cnt = {}, {}
for name, e0, e1 in ops.profile_events:
    agg[name] = agg.get(name, 0.0) + e0.elapsed_time(e1); cnt[name] = cnt.get(name, 0) + 1
tot = sum(agg.values())
print(f"precision {prec}: sum of calls {tot:.2f} ms, {len(ops.profile_events)} calls")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:40]:
    print(f"{v:8.3f} ms  x{cnt[k]:<4d} {k}")
