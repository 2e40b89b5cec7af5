"""Probe: where does the transposed pixel kernel's time go?  Runs ocv_pixel_fused_bf16 at the bench shape with the
measurement knobs of OCV_PIXEL_DEBUG (results are wrong with any knob set; only the time matters).

The shipped library ignores OCV_PIXEL_DEBUG: build a measurement library first with
``OCV_NVCC_EXTRA=-DOCV_PROBES python -m objcavit_b200.build -f`` (and rebuild without it afterwards)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from objcavit_b200 import ops

dev = torch.device("cuda")
B, h, w = 16, 240, 320
g = torch.Generator().manual_seed(0)
x = (torch.randn(B, h, w, 128, generator=g) * 0.5).bfloat16().to(dev)
w3p = ops.pack_conv3x3_bf16((torch.randn(128, 128, 3, 3, generator=g) * 0.03).to(dev))
wout = (torch.randn(256, 128, 1, 1, generator=g) * 0.1).to(dev); bout = torch.zeros(256, device=dev)
tok = (torch.randn(B, 128, 128, generator=g) * 0.2).to(dev)
Mb, blog = ops.fold_queries_bf16(wout, bout, tok, torch.zeros(128, device=dev))
centres = torch.rand(B, 256, device=dev).sort(1).values.contiguous()


def timed(label):
    for _ in range(3):
        ops.pixel_fused_bf16(x, w3p, Mb, blog, centres)
    torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10):
            ops.pixel_fused_bf16(x, w3p, Mb, blog, centres)
        b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) / 10)
    print(f"{label:60s} {min(ts) * 1e3:7.1f} us")


for t, dbg, label in [("0", "0", "pixels-on-M kernel (N128 MMAs)"), ("1", "0", "transposed kernel"), ("1", "1", "transposed, weight/Mb slabs loaded once"),
                      ("1", "4", "transposed, planes loaded once"), ("1", "5", "transposed, no TMA traffic after the first block"),
                      ("1", "2", "transposed, no softmax math"), ("1", "7", "transposed, no TMA traffic, no softmax math")]:
    os.environ["OCV_PIXEL_T"] = t
    os.environ["OCV_PIXEL_DEBUG"] = dbg
    timed(label)

os.environ["OCV_PIXEL_T"] = "1"
for dbg, reps, pause in (("8", 3, 0.0), ("8", 3, 0.2), ("8", 40, 0.0)):
    os.environ["OCV_PIXEL_DEBUG"] = dbg
    print("OCV_PIXEL_DEBUG =", dbg, "reps", reps, "pause", pause, flush=True)
    import time
    for i in range(reps):
        if pause:
            time.sleep(pause)
        if reps > 10 and i < reps - 2:
            os.environ["OCV_PIXEL_DEBUG"] = "0"
            for _ in range(5):
                ops.pixel_fused_bf16(x, w3p, Mb, blog, centres)
            os.environ["OCV_PIXEL_DEBUG"] = dbg
            continue
        ops.pixel_fused_bf16(x, w3p, Mb, blog, centres)
    torch.cuda.synchronize()
