"""Sweep the pipelined serving mode (lanes x pixel_ctas) on the headline workload: ms/step over 40 steps (CUDA events)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from objcavit_b200.config import builtin_args  # noqa: E402
from objcavit_b200.modules import build_head  # noqa: E402

dev = torch.device("cuda:0")
B = int(os.environ.get("B", "16"))
head = build_head(builtin_args(bench.CFG_NAME), precision="bf16").to(dev).eval()
head.load_state_dict(bench.load_weights(bench.CFG_NAME), strict=True)
feats, objs, boxes = bench.make_inputs(B, 0)
x = feats.permute(0, 2, 3, 1).contiguous().bfloat16().to(dev).permute(0, 3, 1, 2)
objs = [o.to(dev) for o in objs]
boxes = [b.to(dev) for b in boxes]


def run(steps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    with torch.no_grad():
        for _ in range(steps):
            out = head(x, list(objs), boxes)
    head.join()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, out


with torch.no_grad():
    ref = head(x, list(objs), boxes).depth_pred.clone()
ms, _ = run(10)
ms, _ = run(40)
print(f"classic: {ms:.4f} ms/step  {B / ms * 1e3:.0f} maps/s")
for lanes in (1, 2, 3):
    for ctas in (0, 140, 132, 124, 116, 100):
        head.enable_pipeline(lanes, ctas)
        run(10)
        ms, out = run(40)
        same = torch.equal(out.depth_pred, ref)
        print(f"lanes {lanes} pixel_ctas {ctas:3d}: {ms:.4f} ms/step  {B / ms * 1e3:.0f} maps/s  bit-equal {same}")
head.enable_pipeline(0)
