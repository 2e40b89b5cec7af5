"""Concurrent GPU timeline of one graph-replayed training step (CUPTI through torch.profiler): per-stream busy time, the span of
the step, idle gaps of the whole GPU, and the kernels ranked by summed duration.  Not a bench number (profiler attached)."""
import os, sys, json, collections, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from objcavit_b200.config import builtin_args
from objcavit_b200.modules import build_head
from objcavit_b200.trainer import HeadTrainer
from oracle import synth  # inputs only
from torch.profiler import profile, ProfilerActivity

dev = torch.device("cuda:0")
B, H, W, NOBJ = 8, 208, 272, 32
head = build_head(builtin_args("nyu_ocv_learned_bbox_wh"), precision="bf16").to(dev).train()
tr = HeadTrainer(head, graph=("eager" not in sys.argv))
x = synth.make_features(B, H, W, seed=1).to(dev)
objs, boxes = synth.make_objects(B, H, W, NOBJ, seed=1)
objs, boxes = [o.to(dev) for o in objs], [b.to(dev) for b in boxes]
gt, mask = synth.make_gt(B, H, W, 10.0, seed=1)
gt, mask = gt.to(dev), mask.to(dev)
for _ in range(6):
    tr.step(x, list(objs), list(boxes), gt, mask)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        tr.step(x, list(objs), list(boxes), gt, mask)
    torch.cuda.synchronize()
path = "/tmp/trace.json"
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
ev.sort(key=lambda e: e["ts"])


def short(n: str, k: int = 44) -> str:
    n = n.replace("(anonymous namespace)::", "").replace("void ", "")
    return n.split("(")[0][:k]

# split into steps at the largest gaps: take the middle step = events between the 1st and 2nd adamw groups
names = [e["name"] for e in ev]
adam = [i for i, n in enumerate(names) if "adamw" in n]
groups = []
for i in adam:
    if not groups or i - groups[-1][-1] > 20:
        groups.append([i])
    else:
        groups[-1].append(i)
lo, hi = groups[0][-1] + 1, groups[1][-1] + 1
step = ev[lo:hi]
t0 = step[0]["ts"]; t1 = max(e["ts"] + e["dur"] for e in step)
print(f"step span {1e-3 * (t1 - t0):.3f} ms, {len(step)} GPU activities")
os.makedirs("gpurun_out", exist_ok=True)
with open("gpurun_out/train_timeline_events.csv", "w") as f:
    f.write("start_us,dur_us,stream,grid,block,name\n")
    for e in step:
        a = e["args"]
        f.write(f'{e["ts"] - t0:.1f},{e["dur"]:.1f},{a.get("stream")},{"x".join(map(str, a.get("grid", [])))},{"x".join(map(str, a.get("block", [])))},{short(e["name"], 60)}\n')
per_stream = collections.defaultdict(float)
for e in step:
    per_stream[e["args"].get("stream")] += e["dur"]
for s, d in sorted(per_stream.items(), key=lambda kv: -kv[1]):
    print(f"  stream {s}: busy {1e-3 * d:.3f} ms")
# GPU-wide idle: union of intervals
iv = sorted((e["ts"], e["ts"] + e["dur"]) for e in step)
busy, cur_s, cur_e = 0.0, iv[0][0], iv[0][1]
gaps = []
for s, e in iv[1:]:
    if s > cur_e:
        busy += cur_e - cur_s; gaps.append((s - cur_e, cur_e - t0)); cur_s, cur_e = s, e
    else:
        cur_e = max(cur_e, e)
busy += cur_e - cur_s
print(f"GPU busy (any stream) {1e-3 * busy:.3f} ms, idle {1e-3 * (t1 - t0 - busy):.3f} ms in {len(gaps)} gaps; largest gaps (us @ ms):",
      [(round(g, 1), round(1e-3 * at, 2)) for g, at in sorted(gaps, reverse=True)[:8]])
agg = collections.defaultdict(lambda: [0.0, 0])
for e in step:
    n = short(e["name"], 60)
    agg[n][0] += e["dur"]; agg[n][1] += 1
print("kernels by summed duration:")
for n, (d, c) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:32]:
    print(f"  {1e-3 * d:7.3f} ms  x{c:<4d} {n}")
# coarse phase timeline: 20 slices, number of concurrently running kernels on average and top kernel
S = 50
for k in range(S):
    a, b = t0 + (t1 - t0) * k / S, t0 + (t1 - t0) * (k + 1) / S
    occ = collections.defaultdict(float)
    for e in step:
        o = min(b, e["ts"] + e["dur"]) - max(a, e["ts"])
        if o > 0:
            occ[f'{short(e["name"], 30)}@{e["args"].get("stream")}'] += o
    tot = sum(occ.values())
    top = sorted(occ.items(), key=lambda kv: -kv[1])[:3]
    print(f"  {1e-3 * (a - t0):5.2f} ms  conc {tot / (b - a):4.2f}  " + ", ".join(f"{n} {100 * v / (b - a):.0f}%" for n, v in top))
