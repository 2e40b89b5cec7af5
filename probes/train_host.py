"""Host issue time vs GPU time of the training step (is the step launch-bound on the host?)."""
import sys, os, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from objcavit_b200 import ops
import objcavit_b200.engine_train as et
from objcavit_b200.config import builtin_args
from objcavit_b200.modules import build_head
from objcavit_b200.trainer import HeadTrainer
from oracle import synth  # inputs only

prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
et.ASYNC_WGRAD = (sys.argv[2] != "sync") if len(sys.argv) > 2 else True
dev = torch.device("cuda:0")
B, H, W, NOBJ = 8, 208, 272, 32
head = build_head(builtin_args("nyu_ocv_learned_bbox_wh"), precision=prec).to(dev).train()
tr = HeadTrainer(head)
x = synth.make_features(B, H, W, seed=1).to(dev)
objs, boxes = synth.make_objects(B, H, W, NOBJ, seed=1)
objs, boxes = [o.to(dev) for o in objs], [b.to(dev) for b in boxes]
gt, mask = synth.make_gt(B, H, W, 10.0, seed=1)
gt, mask = gt.to(dev), mask.to(dev)
for _ in range(5):
    tr.step(x, list(objs), list(boxes), gt, mask)
torch.cuda.synchronize()
K = 20
t0 = time.perf_counter()
for _ in range(K):
    tr.step(x, list(objs), list(boxes), gt, mask)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"{prec} async_wgrad={et.ASYNC_WGRAD}: host issue {1e3 * (t1 - t0) / K:.2f} ms/step, total {1e3 * (t2 - t0) / K:.2f} ms/step")
# host issue time of ONE step with the GPU held busy by a spin kernel (the host never waits for a full launch queue: one step is
# ~560 launches), i.e. the time below which the step cannot go however fast the kernels are
ts = []
for _ in range(10):
    torch.cuda._sleep(int(40e-3 * 1.9e9))
    t0 = time.perf_counter()
    tr.step(x, list(objs), list(boxes), gt, mask)
    ts.append(time.perf_counter() - t0)
    torch.cuda.synchronize()
ts.sort()
print(f"host issue of one step behind a busy GPU: median {1e3 * ts[len(ts) // 2]:.2f} ms, min {1e3 * ts[0]:.2f} ms")
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(5):
    tr.step(x, list(objs), list(boxes), gt, mask)
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
