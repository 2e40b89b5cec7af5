"""Device-resident throughput of the other BASELINE.json inference configurations (they are parity-test cases, not bench lines):
config 2 (KITTI grid_random_roi_align, 64 objects, batch 8, features 176x608) and config 3 (NYU 2-SACA, batch 32), bf16 head,
classic calls and the pipelined serving mode.  CUDA events, 20 steps after 5 warm-up steps; inputs > L2 (no flush needed)."""
import json, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from objcavit_b200.config import builtin_args
from objcavit_b200.modules import build_head
from oracle import synth  # inputs / seeded weights only

dev = torch.device("cuda:0")
CASES = [("nyu_ocv_learned_bbox_wh", 16, 240, 320, 32, "config 1 (bench line)"),
         ("kitti_ocv_grid_random_roi_align", 8, 176, 608, 64, "config 2"),
         ("nyu_ocv_learned_bbox_wh_2saca", 32, 240, 320, 32, "config 3"),
         ("nyu_adabins", 16, 240, 320, 0, "AdaBins head (config 0 shape, batch 16)")]
for name, B, h, w, nobj, label in CASES:
    args = builtin_args(name)
    with open(os.path.join(ROOT, "tests", "golden", f"manifest_{name}.json")) as f:
        sd = synth.make_state_dict(json.load(f), seed=0)
    head = build_head(args, precision="bf16").to(dev).eval()
    head.load_state_dict(sd, strict=True)
    x = synth.make_features(B, h, w, seed=1).to(dev)
    from objcavit_b200 import ops
    xh = ops.nchw_to_nhwc_bf16(x).permute(0, 3, 1, 2)          # channels-last bf16 view, as bench.py feeds it
    if nobj:
        objs, boxes = synth.make_objects(B, h, w, nobj, seed=1)
        objs, boxes = [o.to(dev) for o in objs], [b.to(dev) for b in boxes]
        call = lambda: head(xh, list(objs), boxes)
    else:
        call = lambda: head(xh)
    out = {}
    for lanes in (0, 2):
        head.enable_pipeline(lanes)
        with torch.no_grad():
            for _ in range(5):
                call()
            head.join(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                call()
            head.join()
            e1.record(); torch.cuda.synchronize()
        out[lanes] = e0.elapsed_time(e1) / 20
    head.enable_pipeline(0)
    print(f"{label:42s} {name:34s} B={B:<3d} {h}x{w}: classic {out[0]:.3f} ms/step = {B / out[0] * 1e3:8.0f} maps/s; "
          f"pipelined(2) {out[2]:.3f} ms/step = {B / out[2] * 1e3:8.0f} maps/s")
