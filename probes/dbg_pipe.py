import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from objcavit_b200.config import builtin_args
from objcavit_b200.modules import build_head
dev = torch.device("cuda:0")
head = build_head(builtin_args(bench.CFG_NAME), precision="bf16").to(dev).eval()
head.load_state_dict(bench.load_weights(bench.CFG_NAME), strict=True)
feats, objs, boxes = bench.make_inputs(16, 0)
x = feats.permute(0, 2, 3, 1).contiguous().bfloat16().to(dev).permute(0, 3, 1, 2)
objs = [o.to(dev) for o in objs]; boxes = [b.to(dev) for b in boxes]
head.enable_pipeline(2, 0)
with torch.no_grad():
    for i in range(8):
        t0 = time.perf_counter()
        head(x, list(objs), boxes)
        print(f"call {i}: host {1e3*(time.perf_counter()-t0):.3f} ms", file=sys.stderr)
head.join(); torch.cuda.synchronize()
import cProfile, pstats
pr = cProfile.Profile()
with torch.no_grad():
    pr.enable()
    for i in range(200):
        head(x, list(objs), boxes)
    pr.disable()
head.join(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
