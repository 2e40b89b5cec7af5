"""Where does a head step go?  (a) host time of one module call (no sync), (b) GPU time of the same call replayed from a
CUDA graph (no host in the loop), (c) per-kernel GPU times from a graph-free run with events between native stages."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from objcavit_b200 import engine, engine_bf16, ops
from objcavit_b200.config import builtin_args
from objcavit_b200.modules import build_head
from oracle import synth

name = sys.argv[1] if len(sys.argv) > 1 else "nyu_ocv_learned_bbox_wh"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 16
H, W, N = 240, 320, 32
dev = torch.device("cuda:0")
with open(os.path.join(ROOT, "tests", "golden", f"manifest_{name}.json")) as f:
    sd = synth.make_state_dict(json.load(f), seed=0)
head = build_head(builtin_args(name), precision="bf16").to(dev).eval()
head.load_state_dict(sd)
feats = synth.make_features(B, H, W, seed=1)
objs, boxes = synth.make_objects(B, H, W, N, seed=1)
x = feats.permute(0, 2, 3, 1).contiguous().bfloat16().to(dev).permute(0, 3, 1, 2)
objs = [o.to(dev) for o in objs]; boxes = [b.to(dev) for b in boxes]

def step():
    with torch.no_grad():
        return head(x, list(objs), boxes)

for _ in range(5): step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(50): step()
host = (time.perf_counter() - t0) / 50
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(50): step()
torch.cuda.synchronize()
tot = (time.perf_counter() - t0) / 50
print(f"host issue time per step {host*1e3:.3f} ms; synced step {tot*1e3:.3f} ms")

# marshal only
t0 = time.perf_counter()
for _ in range(50):
    f, b, c = engine.marshal_objects(list(objs), boxes, dev); ob = engine.ObjectBatch(f, c)
print(f"marshal host time {(time.perf_counter()-t0)/50*1e3:.3f} ms")

# graph replay of the native call
nh = head.__dict__["_ocv_native"]
f, b, c = engine.marshal_objects(list(objs), boxes, dev); ob = engine.ObjectBatch(f, c)
x_nhwc = engine_bf16.features_nhwc_bf16(x)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    nh.forward(x_nhwc, f, b, ob.offsets, f.shape[0], ob.n_max)
    torch.cuda.synchronize()
    with torch.cuda.graph(g, stream=s):
        d, e = nh.forward(x_nhwc, f, b, ob.offsets, f.shape[0], ob.n_max)
for _ in range(3): g.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50): g.replay()
e1.record(); torch.cuda.synchronize()
print(f"graph replay per step {e0.elapsed_time(e1)/50:.3f} ms")
# direct native call, host-issued
e0.record()
for _ in range(50): nh.forward(x_nhwc, f, b, ob.offsets, f.shape[0], ob.n_max)
e1.record(); torch.cuda.synchronize()
print(f"native call per step {e0.elapsed_time(e1)/50:.3f} ms")
