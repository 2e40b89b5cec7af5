"""Host-side cost of one head call: python marshalling vs the native executor's launch sequence."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from objcavit_b200 import engine, ops  # noqa: E402
from objcavit_b200.config import builtin_args  # noqa: E402
from objcavit_b200.modules import build_head  # noqa: E402

dev = torch.device("cuda:0")
B = 16
head = build_head(builtin_args(bench.CFG_NAME), precision="bf16").to(dev).eval()
head.load_state_dict(bench.load_weights(bench.CFG_NAME), strict=True)
feats, objs, boxes = bench.make_inputs(B, 0)
x = feats.permute(0, 2, 3, 1).contiguous().bfloat16().to(dev).permute(0, 3, 1, 2)
objs = [o.to(dev) for o in objs]
boxes = [b.to(dev) for b in boxes]
with torch.no_grad():
    for _ in range(5):
        head(x, list(objs), boxes)
torch.cuda.synchronize()
N = 50
t0 = time.perf_counter()
for _ in range(N):
    f, bx, counts = engine.marshal_objects(list(objs), boxes, dev)
    batch = engine.ObjectBatch(f, counts)
torch.cuda.synchronize()
print(f"python marshalling: {(time.perf_counter() - t0) / N * 1e3:.3f} ms/call")
nh = head.__dict__["_ocv_native"]
xn = x.permute(0, 2, 3, 1)
tok = torch.empty(f.shape[0], 128, device=dev)
torch.cuda.synchronize()
for lanes in (0, 3):
    nh.set_pipeline(lanes, 0)
    nh.forward(xn, f, bx, batch.offsets, f.shape[0], batch.n_max, tok)
    torch.cuda.synchronize()
    # host issue time: small batches of calls so the launch queue never fills
    ts = []
    for _ in range(10):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        nh.forward(xn, f, bx, batch.offsets, f.shape[0], batch.n_max, tok)
        ts.append(time.perf_counter() - t0)
    nh.join()
    torch.cuda.synchronize()
    print(f"lanes {lanes}: native forward host issue time: median {sorted(ts)[5] * 1e3:.3f} ms/call")
    t0 = time.perf_counter()
    with torch.no_grad():
        for _ in range(N):
            head(x, list(objs), boxes)
    t1 = time.perf_counter()
    head.join()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"lanes {lanes}: module call loop: host {(t1 - t0) / N * 1e3:.3f} ms/call, incl. drain {(t2 - t0) / N * 1e3:.3f} ms/call")
