#!/usr/bin/env python
"""Headline benchmark: ObjCAViT head depth-maps/s (BASELINE.json metric) on N GPUs of one node.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference ...                      (the reference head's own modules on the host cores)
    python bench.py --workload train ...                      (BASELINE.json configs[4] as the headline of the line)

A step = one pass of the head (ObjCAViT + conv_out + bin-centre expectation) over one batch of synthetic decoder features +
synthetic YOLO boxes / CLIP embeddings (BASELINE.json configs[1]: nyu learned_bbox_wh, 32 boxes, batch 16 per GPU, 240x320
features of a 480x640 image).  Prints ONE JSON line on rank 0.  Besides the headline (`value`, `e2e`, `roofline`,
`cpu_baseline`) the line carries, measured in the same run:

  train               BASELINE.json configs[4]: the training step (fwd + SILog/chamfer + bwd + NCCL all-reduce + clip + AdamW),
                      batch 8 / GPU, with the exposed all-reduce time of the step (`allreduce_ms`)
  configs             device-resident depth-maps/s of BASELINE.json configs[0], [2], [3] (config 4 both as the strong-scaling
                      split of a global batch of 32 and with 32 images per GPU)
  e2e_fp32_nchw       the end-to-end figure with the reference caller's input format (pinned fp32 NCHW, modules/GraphBins.py:107)
  h2d_ceiling         a plain pinned cudaMemcpyAsync host->device probe (what bounds both e2e figures)
  gpu_eager_baseline  the reference's own modules in stock eager PyTorch on this GPU (fp32 and autocast bf16)
  parity              the GPU outputs of this run against the CPU baseline leg's outputs (same images)

See DESIGN.md "Measurement" for how each field is obtained.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

_REAL_STDOUT = None


def capture_stdout() -> None:
    """From here on everything libraries write to fd 1 (NCCL prints its version banner there under torchrun) goes to stderr; the
    ONE JSON line of the contract is written to the real stdout by ``emit``."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(obj) -> None:
    line = (json.dumps(obj) + "\n").encode()
    sys.stdout.flush()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, line)


METRIC = "ObjCAViT head depth-maps/sec @480x640"
UNIT = "depth-maps/s"
CFG_NAME = "nyu_ocv_learned_bbox_wh"
H, W, NOBJ = 240, 320, 32
# algorithmic FLOPs per image (general computation)
GF_HEAD = 34.0
GF_PIXEL = 22.649 + 2.517 + 5.033      # conv3x3 + dot-product + conv_out: the fused pixel kernel (algorithmic)
GF_PIXEL_EXECUTED = 27.7               # what the kernel executes after folding the range-attention channels
GF_CONV3 = 22.649
# ncu --set full capture of one pixel_fused launch (reduced by probes/ncu_summary.py): where `roofline.traffic` comes from
NCU_SUMMARY = os.path.join("profiles", "pixel_fused_ncu_summary.json")

# BASELINE.json configs[0], [2], [3]: name, builtin args, batch, h, w, objects, algorithmic GF / image
OTHER_CONFIGS = [
    ("config1_adabins_b1", "nyu_adabins", 1, 240, 320, 0, 33.69),
    ("config3_kitti_roi_align_b8", "kitti_ocv_grid_random_roi_align", 8, 176, 608, 64, 47.51),
    ("config4_nyu_2saca_b32_per_gpu", "nyu_ocv_learned_bbox_wh_2saca", 32, 240, 320, 32, 36.11),
]


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


def ncu_traffic(batch):
    """dram__bytes_read.sum + dram__bytes_write.sum of one pixel_fused launch from the committed ncu summary, scaled to ``batch``
    images (the kernel streams every image once: traffic is proportional to the batch).  None when the file is missing."""
    try:
        with open(os.path.join(ROOT, NCU_SUMMARY)) as f:
            row = json.load(f)[0]
        return row["dram_bytes_per_launch"] * batch / row["batch"], row
    except (OSError, KeyError, IndexError, ValueError, TypeError):
        return None, None


def load_weights(name):
    from oracle import synth  # seeded synthetic weights only (no computation)
    with open(os.path.join(ROOT, "tests", "golden", f"manifest_{name}.json")) as f:
        return synth.make_state_dict(json.load(f), seed=0)


def make_inputs(B, rank, h=H, w=W, nobj=NOBJ):
    from oracle import synth
    feats = synth.make_features(B, h, w, seed=1234 + rank)
    if not nobj:
        return feats, None, None
    objs, boxes = synth.make_objects(B, h, w, nobj, seed=1234 + rank)
    return feats, objs, boxes


class ClockSampler:
    """nvidia-smi clocks / throttle reasons (B200_PROFILING.md recipe), one sample every 500 ms from process start to the end of the
    last timed region.  The query is kept to the clock fields (power / utilisation queries take a driver lock for milliseconds and a
    sample that lands inside a 12 ms timed window doubles it); samples are attributed to the timed windows by their timestamps."""

    Q = ("timestamp,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.windows = index, [], None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "500", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def window(self, t0, t1):
        self.windows.append((t0, t1))

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.6)
        self.proc.terminate()
        rows = [(t, r[1:]) for t, r in self.rows if len(r) >= 7 and r[1].isdigit()]
        lo = min((w[0] for w in self.windows), default=0.0) - 0.6
        hi = max((w[1] for w in self.windows), default=float("inf")) + 0.6
        busy = [r for t, r in rows if lo <= t <= hi] or [r for _, r in rows]      # samples between the first and the last timed region
        sm = sorted(int(r[0]) for r in busy)
        mx = [int(r[1]) for _, r in rows if r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in busy for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(rows), "samples_in_timed_span": len(busy)}


_SAMPLER = None


# ---- the reference's own CPU implementation of the path -----------------------------------------------------------------

def reference_runner(name=CFG_NAME, train=False):
    """-> (kind, module-like callable): the UNMODIFIED reference head modules (``oracle/_ref`` on the GPU box, ``/root/reference``
    in the build container: ``kind = "reference"``) around the 8-line tail of ``modules/GraphBins.py:107-119``; when neither is
    present, the restatement ``oracle/head.py`` (``kind = "port"``)."""
    from objcavit_b200.config import builtin_args
    from oracle import ref_shim
    sd = load_weights(name)
    args = builtin_args(name)
    if ref_shim.reference_available():
        ref = ref_shim.ReferenceHead(args)
        ref.load_state_dict(sd, strict=True)
        ref = ref.train() if train else ref.eval()
        return "reference", ref, sd
    return "port", None, sd


def _port_cfg(train=False):
    cfg = dict(kind="graphbins", strategy="learned_bbox_wh", min_depth=0.001, max_depth=10.0)
    if train:
        cfg["train"] = True
    return cfg


def cpu_reference_rate(feats, objs, boxes, threads, repeats=3):
    """The reference head (fp32 torch CPU ops) on the host cores over the given images.  -> (maps/s, seconds, kind, depth)."""
    import copy
    from oracle import head as oh
    torch.set_num_threads(threads)
    kind, ref, sd = reference_runner()

    def run(f, o, b):
        if ref is not None:
            return ref(f, [t.clone() for t in o], copy.deepcopy(b))
        return oh.head_forward(sd, _port_cfg(), f, o, b)

    best, out = float("inf"), None
    with torch.no_grad():
        run(feats[:1], objs[:1], boxes[:1])          # warm-up
        for _ in range(repeats):
            t0 = time.perf_counter()
            out = run(feats, objs, boxes)
            best = min(best, time.perf_counter() - t0)
    return feats.shape[0] / best, best, kind, out


def cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def run_reference(a):
    """--impl reference: the reference's own CPU implementation of the path, all host threads, each step a bounded sample of
    the workload (the full batch of 16 when one pass stays under ~6 s on this host, else 4 images)."""
    import copy
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    from oracle import head as oh
    kind, ref, sd = reference_runner()
    feats16, objs16, boxes16 = make_inputs(16, 0)

    def run(n):
        f, o, b = feats16[:n], objs16[:n], boxes16[:n]
        if ref is not None:
            return ref(f, [t.clone() for t in o], copy.deepcopy(b))
        return oh.head_forward(sd, _port_cfg(), f, o, b)

    with torch.no_grad():
        run(1)
        t0 = time.perf_counter()
        run(2)
        per_image = (time.perf_counter() - t0) / 2
        sample = 16 if per_image * 16 < 6.0 else 4
        warm = min(a.warmup, 2)
        for _ in range(warm):
            run(sample)
        steps = max(1, min(a.steps, 5))
        t0 = time.perf_counter()
        for _ in range(steps):
            run(sample)
        dt = (time.perf_counter() - t0) / steps
    v = sample / dt
    what = ("the UNMODIFIED reference modules (modules/ObjCAViT.py + tail of modules/GraphBins.py:107-119) in eager PyTorch fp32"
            if kind == "reference" else "torch fp32 CPU ops of oracle/head.py (reference modules not shipped to this box)")
    emit(({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{CFG_NAME} head, {NOBJ} boxes/img, features 128x{H}x{W}, "
                               + ("the full batch of 16 images per step" if sample == 16 else f"bounded sample of {sample} images per step"),
                   "same_images_as_gpu_arm": "the first images of rank 0's batch (seed 1234)"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": kind, "cpu": cpu_model(),
                         "sample": f"{sample} images per step, {steps} steps, {what}"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ---- config 5: the training step ----------------------------------------------------------------------------------
TRAIN_H, TRAIN_W = 208, 272          # NYU-train features of a 416x544 crop
GF_TRAIN = 56.5                      # algorithmic GFLOP / image / step (fwd 25.02 + bwd 31.5)
TRAIN_METRIC = "ObjCAViT head training images/sec (config 5: fwd+bwd, SILog+BinsChamfer, grad all-reduce, AdamW)"


def cpu_train_rate(sample_images, threads, repeats=2):
    """The reference's training step of the head on the host cores (the unmodified reference modules in train() mode under torch
    autograd when available, else the oracle port: forward with dropout 0.1, SILog + 0.1 chamfer, backward; no optimizer) on a
    bounded sample of the config-5 workload.  The chamfer term is the restatement in both cases (pytorch3d is not installable)."""
    import copy
    from oracle import head as oh
    from oracle import synth
    torch.set_num_threads(threads)
    kind, ref, sd0 = reference_runner(train=True)
    feats = synth.make_features(sample_images, TRAIN_H, TRAIN_W, seed=1234)
    objs, boxes = synth.make_objects(sample_images, TRAIN_H, TRAIN_W, NOBJ, seed=1234)
    gt, mask = synth.make_gt(sample_images, TRAIN_H, TRAIN_W, 10.0, seed=1234)
    best = float("inf")
    if ref is not None:
        for i in range(repeats + 1):
            t0 = time.perf_counter()
            depth, edges, _, _ = ref(feats, [o.clone() for o in objs], copy.deepcopy(boxes))
            (oh.silog_loss(depth, gt, mask) + 0.1 * oh.bins_chamfer_loss(edges, gt, mask)).backward()
            if i > 0:
                best = min(best, time.perf_counter() - t0)
            ref.zero_grad(set_to_none=True)
        return sample_images / best, best, kind
    sd = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    oh.DROPOUT_HOOK = lambda t: torch.nn.functional.dropout(t, 0.1, True)      # train(): dropout 0.1 at every encoder-layer site
    try:
        for i in range(repeats + 1):
            t0 = time.perf_counter()
            depth, edges, _, _ = oh.head_forward(sd, _port_cfg(train=True), feats, objs, boxes)
            (oh.silog_loss(depth, gt, mask) + 0.1 * oh.bins_chamfer_loss(edges, gt, mask)).backward()
            if i > 0:
                best = min(best, time.perf_counter() - t0)
            for v in sd.values():
                v.grad = None
    finally:
        oh.DROPOUT_HOOK = None
    return sample_images / best, best, kind


def run_reference_train(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = 2
    reps = max(1, min(a.steps, 3))
    v, secs, kind = cpu_train_rate(sample, threads, repeats=reps)
    emit(({
        "impl": "reference", "metric": TRAIN_METRIC, "value": v, "unit": "images/s", "n_gpus": a.gpus, "steps": reps,
        "warmup": 1, "ms_per_step": secs * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": f"{CFG_NAME} head training step (fwd + SILog/chamfer + bwd), {NOBJ} boxes/img, features "
                               f"128x{TRAIN_H}x{TRAIN_W}, bounded sample of {sample} images per step"},
        "cpu_baseline": {"value": v, "unit": "images/s", "cores": threads, "kind": kind, "cpu": cpu_model(),
                         "sample": f"{sample} images per step, best of {reps}, torch autograd over "
                                   + ("the unmodified reference modules in train() mode" if kind == "reference" else "oracle/head.py")},
        "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ---- shared measurement plumbing -------------------------------------------------------------------------------------

class Ctx:
    def __init__(self):
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback)"
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        torch.cuda.synchronize()

    def timed(self, fn, steps):
        """CUDA events around ``steps`` calls of ``fn``, barrier + synchronize on both sides -> (MAX over ranks ms, per-rank ms)."""
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.time()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        self.barrier()
        if _SAMPLER is not None:
            _SAMPLER.window(w0, time.time())
        ms = e0.elapsed_time(e1)
        if self.dist is None:
            return ms, [ms]
        t = torch.tensor([ms], device=self.dev)
        allms = [torch.empty_like(t) for _ in range(self.world)]
        self.dist.all_gather(allms, t)
        per = [float(x.item()) for x in allms]
        return max(per), per

    def timed_median(self, fn, steps, repeats=3):
        """``timed`` over ``repeats`` back-to-back windows of exactly ``steps`` steps each; the window with the MEDIAN job time
        (max over ranks) is reported.  A 12 ms window is at the mercy of one host hiccup (observed on this pool: single windows
        2-3x slower about once in eight processes, the neighbours of the same process unaffected)."""
        runs = sorted((self.timed(fn, steps) for _ in range(repeats)), key=lambda r: r[0])
        return runs[len(runs) // 2]

    def timed_windows(self, fn, steps, min_total_ms=1000.0, min_windows=3, max_windows=400):
        """``timed`` over back-to-back windows of exactly ``steps`` steps each until the windows add up to ``min_total_ms`` of
        device time (max over ranks, so every rank stops at the same window).  -> (median window ms, its per-rank ms, stats):
        a headline taken from 3 windows of 12 ms depended on a quiet host for those 36 ms (round 2: the same code gave 13.2k and
        27.3k maps/s on two boxes); one second of windows shows the spread instead of sampling it."""
        runs, total = [], 0.0
        while len(runs) < min_windows or (total < min_total_ms and len(runs) < max_windows):
            r = self.timed(fn, steps)
            runs.append(r)
            total += r[0]
        runs.sort(key=lambda r: r[0])
        q = lambda f: runs[min(len(runs) - 1, int(f * len(runs)))][0] / steps
        stats = {"windows": len(runs), "steps_per_window": steps, "device_time_s": total / 1e3,
                 "ms_per_step": {"min": q(0.0), "p10": q(0.1), "median": q(0.5), "p90": q(0.9), "max": runs[-1][0] / steps},
                 "spread_p90_over_p10": q(0.9) / q(0.1)}
        med = runs[len(runs) // 2]
        return med[0], med[1], stats

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


def rank_stats(per, steps):
    s = sorted(x / steps for x in per)
    return {"min": s[0], "median": s[len(s) // 2], "max": s[-1]}


def bf16_inputs(ctx, feats, objs, boxes):
    """Device-resident inputs in the dtype / layout the bf16 path consumes (NHWC bf16 memory behind a logical NCHW view)."""
    x_host = feats.permute(0, 2, 3, 1).contiguous().bfloat16().pin_memory()
    x_dev = x_host.to(ctx.dev).permute(0, 3, 1, 2)
    objs_dev = None if objs is None else [o.to(ctx.dev) for o in objs]
    boxes_dev = None if boxes is None else [b.to(ctx.dev) for b in boxes]
    return x_host, x_dev, objs_dev, boxes_dev


def head_call(head, x, ob, bx):
    with torch.no_grad():
        return head(x) if ob is None else head(x, list(ob), bx)


def pipelined_rate(ctx, head, x, ob, bx, B, K, W_, lanes, pixel_ctas=0, want_classic=True, min_total_ms=0.0):
    """Device-resident throughput of a bf16 head: classic calls (single-call latency), then the pipelined serving mode.
    ``x``: one input tensor, or a list of equal tensors at different addresses that the calls rotate over (inputs smaller than
    L2: the rotation makes every call read its features from HBM, like a flush would)."""
    from objcavit_b200 import ops
    xs = x if isinstance(x, (list, tuple)) else [x]
    it = [0]

    def call():
        it[0] += 1
        return head_call(head, xs[it[0] % len(xs)], ob, bx)

    for _ in range(W_):
        call()
    ms_classic = ctx.timed_median(call, K)[0] if want_classic else None
    if lanes:
        head.enable_pipeline(lanes, pixel_ctas)
        for _ in range(max(W_, 3 * lanes)):          # incl. the calls that capture each lane's token-path graph
            call()
        head.join()
    l0 = ops.launch_count

    issue = []                                        # host time spent issuing the K calls of a window (before the join)

    def timed_steps():
        t0 = time.perf_counter()
        for _ in range(K):
            call()
        issue.append((time.perf_counter() - t0) * 1e3)
        head.join()                                   # every one of the K steps has finished when the stop event fires

    stats = None
    if min_total_ms > 0:                              # the headline: windows of exactly K steps until >= min_total_ms of device time
        ms, per, stats = ctx.timed_windows(timed_steps, 1, min_total_ms)
        stats["steps_per_window"] = K
        for k in stats["ms_per_step"]:
            stats["ms_per_step"][k] /= K
        n_windows = stats["windows"]
    else:
        ms, per = ctx.timed_median(timed_steps, 1)
        n_windows = 3
    launches = (ops.launch_count - l0) // n_windows
    issue.sort()
    host_ms = issue[len(issue) // 2] / K
    if lanes:
        head.enable_pipeline(0)
    return dict(ms=ms, per=per, ms_classic=ms_classic, launches=launches, value=ctx.world * B * K / (ms * 1e-3), windows=stats,
                host_issue_ms_per_call=host_ms, host_bound=bool(host_ms > 0.8 * ms / K))


def measure_other_configs(ctx, K, W_, lanes, pk):
    """BASELINE.json configs[0], [2], [3] on the bf16 head, device-resident inputs (larger than L2, or rotated over enough copies
    to be)."""
    from objcavit_b200.config import builtin_args
    from objcavit_b200.modules import build_head
    out = {}
    todo = list(OTHER_CONFIGS)
    # config 4 as BASELINE states it: a global batch of 32 split over the GPUs (strong scaling); per-GPU batch 32/N
    if 32 % ctx.world == 0:
        todo.append((f"config4_nyu_2saca_global32_split{ctx.world}", "nyu_ocv_learned_bbox_wh_2saca", 32 // ctx.world, 240, 320, 32, 36.11))
    for label, name, B, h, w, nobj, gf in todo:
        head = build_head(builtin_args(name), precision="bf16").to(ctx.dev).eval()
        head.load_state_dict(load_weights(name), strict=True)
        feats, objs, boxes = make_inputs(B, ctx.rank, h, w, nobj)
        _, x, ob, bx = bf16_inputs(ctx, feats, objs, boxes)
        nbytes = x.numel() * 2
        copies = 1 if nbytes > 200e6 else int(-(-320e6 // nbytes))     # inputs below L2 size: rotate over > 2.5 x L2 of copies
        xs = [x] + [x.clone() for _ in range(copies - 1)]
        r = pipelined_rate(ctx, head, xs if copies > 1 else x, ob, bx, B, K, W_, lanes)
        r["l2"] = (f"inputs {nbytes / 1e6:.0f} MB/step > 126 MB L2, no explicit flush" if copies == 1 else
                   f"inputs {nbytes / 1e6:.0f} MB/step < L2: calls rotate over {copies} copies at different addresses ({copies * nbytes / 1e6:.0f} MB), "
                   "so every call reads its features from HBM")
        del xs
        strong = "global32" in label
        out[label] = {"value": r["value"], "unit": UNIT, "ms_per_step": r["ms"] / K,
                      "ms_per_step_single_call": (r["ms_classic"] / K) if r.get("ms_classic") else None,
                      "batch_per_gpu": B, "features": f"128x{h}x{w} bf16 NHWC", "objects_per_image": nobj,
                      "scaling": "strong" if strong else "weak", "l2": r["l2"],
                      "head_roofline": {"gf_per_image": gf, "achieved_tflops": gf * r["value"] / 1e3 / ctx.world,
                                        "frac_of_burst_peak": gf * r["value"] / 1e3 / ctx.world / pk["tf_burst"],
                                        "frac_of_sustained_peak": gf * r["value"] / 1e3 / ctx.world / pk["tf_sust"]}}
        del head, x, ob, bx
        torch.cuda.empty_cache()
    return out


def measure_train(ctx, a, K, W_, with_breakdown=True, with_e2e=True):
    """BASELINE.json configs[4] — frozen features + ObjCAViT head fwd/bwd, SILog + 0.1 chamfer, one all-reduce of the flat
    head-gradient buffer, clip 0.1, AdamW + OneCycle; batch 8 per GPU (weak scaling).  Returns the fields of a bench line."""
    from objcavit_b200 import ops
    from objcavit_b200.config import builtin_args
    from objcavit_b200.modules import build_head
    from objcavit_b200.trainer import HeadTrainer, OneCycle
    from oracle import synth

    dev, rank, world = ctx.dev, ctx.rank, ctx.world
    B = a.train_batch
    tprec = a.train_precision   # "bf16" (default): tensor-core pixel + token path; "fp32": the all-SIMT parity path
    head = build_head(builtin_args(CFG_NAME), precision=tprec).to(dev).train()
    head.load_state_dict(load_weights(CFG_NAME), strict=True)
    trainer = HeadTrainer(head, lr=0.000357, weight_decay=0.1, clip_grad_norm=0.1,
                          schedule=OneCycle(0.000357, max(W_ + 3 * K + 16, 100), 25, 100), graph=bool(a.graph))
    feats = synth.make_features(B, TRAIN_H, TRAIN_W, seed=1234 + rank)
    objs, boxes = synth.make_objects(B, TRAIN_H, TRAIN_W, NOBJ, seed=1234 + rank)
    gt, mask = synth.make_gt(B, TRAIN_H, TRAIN_W, 10.0, seed=1234 + rank)
    host = dict(x=feats.pin_memory(), gt=gt.pin_memory(), mask=mask.pin_memory(), objs=[o.pin_memory() for o in objs],
                boxes=[b.pin_memory() for b in boxes])
    x_dev, gt_dev, mask_dev = host["x"].to(dev), host["gt"].to(dev), host["mask"].to(dev)
    objs_dev, boxes_dev = [o.to(dev) for o in host["objs"]], [b.to(dev) for b in host["boxes"]]

    def step():
        return trainer.step(x_dev, list(objs_dev), boxes_dev, gt_dev, mask_dev)

    for _ in range(W_):
        step()
    trainer.allreduce_events = [] if world > 1 else None
    l0 = ops.launch_count
    ms, per = ctx.timed_median(step, K)
    launches = (ops.launch_count - l0) // 3
    ar_ms = None
    if trainer.allreduce_events:
        torch.cuda.synchronize()
        ar = [e0.elapsed_time(e1) for e0, e1 in trainer.allreduce_events]
        ar_ms = sum(ar) / len(ar)
        if ctx.dist is not None:
            t = torch.tensor([ar_ms], device=dev)
            ctx.dist.all_reduce(t, op=ctx.dist.ReduceOp.MAX)
            ar_ms = float(t.item())
    trainer.allreduce_events = None
    value = world * B * K / (ms * 1e-3)
    pk = peaks()
    step_tf = GF_TRAIN * B * K / ms                  # GFLOP/ms == TFLOP/s per GPU, whole step on the algorithmic FLOPs
    roofline = {"bound": "tensor", "achieved": step_tf, "peak": pk["tf_burst"], "unit": "TFLOP/s", "frac": step_tf / pk["tf_burst"],
                "frac_of_sustained_peak": step_tf / pk["tf_sust"], "traffic": None,
                "peak_source": pk["src"] + " (bf16 burst); achieved = 56.5 GF/image x images/s of the WHOLE step per GPU"}
    if with_breakdown:
        # per-entry-point breakdown of one step (CUDA events around every C-ABI call on the launching stream)
        ops.profile_events = []
        trainer.graph = False                            # the breakdown needs the individual calls: one eager step
        step()
        trainer.graph = bool(a.graph)
        torch.cuda.synchronize()
        agg = {}
        for name, e0, e1 in ops.profile_events:
            agg[name] = agg.get(name, 0.0) + e0.elapsed_time(e1)
        ops.profile_events = None
        top = sorted(agg.items(), key=lambda kv: -kv[1])
        if top:
            roofline.update(kernel=f"{top[0][0]} (largest entry point of one eager step)", kernel_ms=top[0][1],
                            kernel_share_of_step=top[0][1] / max(sum(agg.values()), 1e-9),
                            breakdown_ms={k: round(v, 3) for k, v in top[:10]})
    out = {
        "metric": TRAIN_METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": K, "warmup": W_,
        "ms_per_step": ms / K, "ms_per_step_ranks": rank_stats(per, K), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16" if tprec == "bf16" else "f32", "data": "synthetic",
        "config": {"workload": f"{CFG_NAME} head training step, {NOBJ} boxes/img, frozen features 128x{TRAIN_H}x{TRAIN_W} fp32, "
                               f"GT {2 * TRAIN_H}x{2 * TRAIN_W}, batch {B}/GPU, encoder dropout {head.dropout} (train() mode, as the reference)",
                   "batch_per_gpu": B,
                   "launch": ("zero-grads + forward + losses + backward replayed from ONE CUDA graph per step (same kernels, "
                              "streams and order as the eager step; inputs copied into static buffers inside the timed region), "
                              "all-reduce + clip + AdamW eager") if a.graph else "eager launches",
                   "parallelism": f"dp{world} (batch sharded, one NCCL all-reduce of "
                                  f"{trainer.grads.flat.numel() * 4 / 1e6:.1f} MB head gradients per step)",
                   "l2": f"a step streams {B * 128 * TRAIN_H * TRAIN_W * 4 / 1e6:.0f} MB of features and {3 * B * 128 * TRAIN_H * TRAIN_W * 2 / 1e6:.0f} MB of bf16 "
                         "conv activations / gradients > 126 MB L2, no explicit flush",
                   "timing": "median of 3 back-to-back windows of exactly `steps` steps"},
        "gpu_launches": launches, "roofline": roofline,
        "allreduce_ms": ar_ms, "allreduce_bytes": trainer.grads.flat.numel() * 4 if world > 1 else 0,
        "allreduce_share_of_step": (ar_ms / (ms / K)) if ar_ms is not None else 0.0,
    }
    if with_e2e:
        # end to end: every step's inputs come from pinned host memory and every step's loss goes back to the host; the copy of
        # step i+1 runs on the prefetcher's stream while step i computes, the loss is read one step late (objcavit_b200/prefetch.py)
        from objcavit_b200.prefetch import DevicePrefetcher, LaggedScalars

        def host_batches():
            while True:
                yield (host["x"], host["objs"], host["boxes"], host["gt"], host["mask"])

        feed = DevicePrefetcher(host_batches(), dev)
        losses = LaggedScalars(2, lag=1)

        def e2e_step():
            x, ob, bx, g, m = next(feed)
            l1, l2 = trainer.step(x, list(ob), list(bx), g, m)
            return losses.push(l1, l2)                    # D2H of this step's loss; returns the previous step's values

        for _ in range(3):
            e2e_step()
        ms_e2e, _ = ctx.timed_median(e2e_step, K)
        h2d = sum(t.numel() * t.element_size() for t in (host["x"], host["gt"], host["mask"])) + \
            sum(t.numel() * 4 for t in host["objs"] + host["boxes"])
        out["e2e"] = {"value": world * B * K / (ms_e2e * 1e-3), "unit": "images/s", "h2d_bytes_per_step": h2d,
                      "d2h_bytes_per_step": 8, "ms_per_step": ms_e2e / K}
    del trainer, head
    torch.cuda.empty_cache()
    return out


def run_train(a):
    """--workload train: the training step as the headline of the line."""
    global _SAMPLER
    ctx = Ctx()
    W_, K = max(a.warmup, 3), a.steps
    sampler = _SAMPLER = ClockSampler(ctx.local).start() if ctx.rank == 0 else None      # NVML start-up is over before anything is timed
    out = measure_train(ctx, a, K, W_)
    out["clocks"] = sampler.stop() if sampler else None
    if ctx.rank == 0 and ctx.world == 1 and not a.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, secs, kind = cpu_train_rate(2, threads)
        out["cpu_baseline"] = {"value": v, "unit": "images/s", "cores": threads, "kind": kind, "cpu": cpu_model(),
                               "sample": f"2 images of the same workload, best of 2 ({secs:.2f} s per pass: forward with dropout 0.1 + losses + "
                                         "backward, no optimizer), torch autograd over "
                                         + ("the unmodified reference modules" if kind == "reference" else "oracle/head.py")}
    if ctx.rank == 0:
        emit(out)
    ctx.close()


def h2d_probe(ctx, nbytes, reps=6):
    """Plain pinned host -> device cudaMemcpyAsync (torch ``copy_(non_blocking=True)`` on a stream: ONE copy per chunk), every
    rank at the same time.  -> GB/s of this rank (best of ``reps``)."""
    src = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    dst = torch.empty(nbytes, dtype=torch.uint8, device=ctx.dev)
    dst.copy_(src, non_blocking=True)
    best = float("inf")
    for _ in range(reps):
        ctx.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dst.copy_(src, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    gbs = nbytes / best / 1e6
    if ctx.dist is not None:
        t = torch.tensor([gbs], device=ctx.dev)
        lo = t.clone()
        ctx.dist.all_reduce(lo, op=ctx.dist.ReduceOp.MIN)
        ctx.dist.all_reduce(t, op=ctx.dist.ReduceOp.SUM)
        return float(lo.item()), float(t.item())
    return gbs, gbs


def gpu_eager_baseline(ctx, feats, objs, boxes, K):
    """The reference's OWN modules (unmodified, ``oracle/_ref``) in stock eager PyTorch on this GPU, fp32 and under
    ``torch.autocast(bf16)``: the number to beat, since the reference has no Blackwell kernel of its own.
    Library kernels only (cuDNN / cuBLAS / ATen); none of this repo's kernels run here."""
    import copy
    try:
        kind, ref, _ = reference_runner()
        if ref is None:
            return {"unavailable": "reference modules not present (oracle/_ref missing): run python -m oracle.vendor_ref in the build container"}
        ref = ref.to(ctx.dev).eval()
        f = feats.to(ctx.dev)
        o = [t.to(ctx.dev) for t in objs]
        b = [t.to(ctx.dev) for t in boxes]
        B = f.shape[0]
        res = {"kind": kind, "what": "unmodified reference modules + tail (GraphBins.py:107-119), eager PyTorch on cuda, same images and weights",
               "batch": B, "torch": torch.__version__}
        for tag, ctxmgr in (("fp32", None), ("autocast_bf16", torch.autocast("cuda", dtype=torch.bfloat16))):
            def run():
                with torch.no_grad():
                    if ctxmgr is None:
                        return ref(f, [t.clone() for t in o], copy.copy(b))
                    with ctxmgr:
                        return ref(f, [t.clone() for t in o], copy.copy(b))
            for _ in range(3):
                run()
            n = max(3, min(K, 10))
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n):
                run()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / n
            res[tag] = {"value": B / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms}
        del ref
        torch.cuda.empty_cache()
        return res
    except Exception as e:      # a baseline leg must never take the headline down
        return {"unavailable": f"{type(e).__name__}: {e}"[:300]}


def parity_against_cpu_leg(ctx, head16, feats, objs, boxes, ref_out):
    """GPU outputs of THIS run against the CPU baseline leg's outputs on the same images: relative depth error of the bf16 and
    fp32 paths, and the bf16 bin argmax against the argmax of the reference's fp32 logits."""
    from objcavit_b200 import engine_bf16
    from objcavit_b200.config import builtin_args
    from objcavit_b200.modules import build_head
    depth_ref, edges_ref, _, ram_ref = ref_out
    n = feats.shape[0]
    dev = ctx.dev

    def rel(a, b):
        return float((a.double().cpu() - b.double()).abs().max() / b.double().abs().max())

    res = {"images": n, "reference": "outputs of the cpu_baseline leg"}
    f = feats.to(dev)
    with torch.no_grad():
        out16 = head16(f, [o.to(dev) for o in objs], [b.to(dev) for b in boxes])
        res["bf16_depth_rel_err"] = rel(out16.depth_pred, depth_ref)
        res["bf16_edges_rel_err"] = rel(out16.bin_edges, edges_ref)
        h32 = build_head(builtin_args(CFG_NAME), precision="fp32").to(dev).eval()
        h32.load_state_dict(load_weights(CFG_NAME), strict=True)
        out32 = h32(f, [o.to(dev) for o in objs], [b.to(dev) for b in boxes])
        res["fp32_depth_rel_err"] = rel(out32.depth_pred, depth_ref)
        del h32
        sd = load_weights(CFG_NAME)
        logits = torch.einsum("kq,bqp->bkp", sd["conv_out.0.weight"].view(256, -1), ram_ref.flatten(2)) + sd["conv_out.0.bias"].view(1, -1, 1)
        top2 = logits.topk(2, dim=1)
        _, _, amax = engine_bf16.head_forward_staged(head16, head16._params(), "objcavit.", f, [o.to(dev) for o in objs],
                                                     [b.to(dev) for b in boxes], 0.001, 10.0, want_argmax=True)
        agree = amax.cpu().long().view(n, -1) == top2.indices[:, 0]
        margin = top2.values[:, 0] - top2.values[:, 1]
        decisive = margin > 2e-2 * float(logits.abs().max())
        res["argmax_overall_agreement"] = float(agree.float().mean())
        res["argmax_decisive_fraction"] = float(decisive.float().mean())
        res["argmax_decisive_agreement"] = float(agree[decisive].float().mean()) if bool(decisive.any()) else 1.0
    res["tolerances"] = {"fp32": 1e-4, "bf16": 2e-2, "argmax_on_decisive_pixels": 0.999}
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=None, choices=[None, "bf16", "fp32"])
    ap.add_argument("--batch", type=int, default=16, help="images per GPU per step (headline workload)")
    ap.add_argument("--train-batch", type=int, default=8, help="images per GPU per training step (config 5)")
    ap.add_argument("--train-precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--lanes", type=int, default=2, help="pipelined serving mode of the bf16 head (0 = classic calls)")
    ap.add_argument("--pixel-ctas", type=int, default=0, help="grid cap of the fused pixel kernel in pipelined mode (0 = 148)")
    ap.add_argument("--graph", type=int, default=1, help="training step: replay zero-grads + forward + losses + backward from a CUDA graph (0 = eager launches)")
    ap.add_argument("--workload", default="infer", choices=["infer", "train"],
                    help="infer = BASELINE.json configs[1] as the headline (the line also carries train / configs sub-records); "
                         "train = configs[4] as the headline")
    ap.add_argument("--min-timed-ms", type=float, default=1000.0,
                    help="headline: keep timing windows of exactly --steps steps until they add up to this much device time "
                         "(0 = three windows, for runs under a profiler)")
    ap.add_argument("--only-headline", action="store_true", help="skip the train / configs / baseline sub-records")
    a = ap.parse_args()
    capture_stdout()
    if a.precision is not None and a.workload == "train":
        a.train_precision = a.precision
    if a.impl != "reference" and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        from objcavit_b200.sharding import bind_host_to_gpu
        bind_host_to_gpu(int(os.environ.get("LOCAL_RANK", "0")))     # NUMA-local pinned buffers for the e2e (host-buffer) legs
    if a.impl == "reference":
        return run_reference_train(a) if a.workload == "train" else run_reference(a)
    if a.workload == "train":
        return run_train(a)

    from objcavit_b200 import ops
    from objcavit_b200.config import builtin_args
    from objcavit_b200.modules import build_head

    global _SAMPLER
    ctx = Ctx()
    world, rank, dev = ctx.world, ctx.rank, ctx.dev
    sampler = _SAMPLER = ClockSampler(ctx.local).start() if rank == 0 else None      # NVML start-up is over before anything is timed
    precision = a.precision
    if precision is None:
        try:
            from objcavit_b200 import engine_bf16  # noqa: F401
            precision = "bf16"
        except ImportError:
            precision = "fp32"
    W_ = max(a.warmup, 3)
    K = a.steps
    B = a.batch
    pk = peaks()

    head = build_head(builtin_args(CFG_NAME), precision=precision).to(dev).eval()
    head.load_state_dict(load_weights(CFG_NAME), strict=True)
    feats, objs, boxes = make_inputs(B, rank)
    # device-resident inputs for `value`: the dtype/layout the path consumes
    if precision == "bf16":
        x_host, x_dev, objs_dev, boxes_dev = bf16_inputs(ctx, feats, objs, boxes)
    else:
        x_host = feats.pin_memory()
        x_dev = x_host.to(dev)
        objs_dev, boxes_dev = [o.to(dev) for o in objs], [b.to(dev) for b in boxes]
    objs_host = [o.pin_memory() for o in objs]
    boxes_host = [b.pin_memory() for b in boxes]

    # ---- device-resident throughput (`value`) ------------------------------------------------------
    # classic mode first (one call = one stream-ordered pass: the single-call latency), then the pipelined serving mode
    # (consecutive calls overlap: token path of call i+1 next to the pixel kernel of call i) which is what `value` reports.
    lanes = a.lanes if precision == "bf16" else 0
    r = pipelined_rate(ctx, head, x_dev, objs_dev, boxes_dev, B, K, W_, lanes, a.pixel_ctas, min_total_ms=a.min_timed_ms)
    ms, ms_classic, launches, value = r["ms"], r["ms_classic"], r["launches"], r["value"]
    # the same pipelined loop with a FRESHLY ALLOCATED input tensor every call (a serving loop does not reuse pointers)
    fresh = None
    if lanes:
        pool = [x_dev.clone() for _ in range(3)]
        head.enable_pipeline(lanes, a.pixel_ctas)
        it = [0]

        def fresh_call():
            it[0] += 1
            head_call(head, pool[it[0] % 3].clone() if it[0] % 2 else pool[it[0] % 3], objs_dev, boxes_dev)

        for _ in range(max(W_, 3 * lanes)):
            fresh_call()
        head.join()

        def fresh_steps():
            for _ in range(K):
                fresh_call()
            head.join()

        ms_fresh, _ = ctx.timed_median(fresh_steps, 1)
        head.enable_pipeline(0)
        fresh = {"value": world * B * K / (ms_fresh * 1e-3), "ms_per_step": ms_fresh / K,
                 "what": "input tensor at a different address every call (rotating + freshly cloned buffers; the clone is inside the timed region)"}
        del pool

    # ---- dominant kernel, timed live with CUDA events on the launching stream ------------------------
    durs = []
    if precision == "bf16":
        from objcavit_b200 import _lib
        dom = "pixel_fused_t_kernel<0> (inside ocv_head_forward)"
        nh = head.__dict__["_ocv_native"]
        _lib.check(_lib.lib().ocv_head_enable_timing(nh._h, 1), "ocv_head_enable_timing")
        for _ in range(min(K, 10)):
            head_call(head, x_dev, objs_dev, boxes_dev)
            durs.append(float(_lib.lib().ocv_head_pixel_ms(nh._h)))
        _lib.lib().ocv_head_enable_timing(nh._h, 0)
    else:
        dom = "ocv_conv3x3_f32"
        ops.timed_kernel = dom
        ops.timed_events = []
        for _ in range(min(K, 10)):
            head_call(head, x_dev, objs_dev, boxes_dev)
        torch.cuda.synchronize()
        durs = [s.elapsed_time(e) for s, e in ops.timed_events]
        ops.timed_kernel = None
    dom_ms = sum(durs) / max(len(durs), 1)
    gf = (GF_PIXEL if precision == "bf16" else GF_CONV3) * B
    achieved = gf / dom_ms if dom_ms > 0 else 0.0     # GFLOP/ms == TFLOP/s
    traffic, ncu_row = ncu_traffic(B) if precision == "bf16" else (None, None)
    # peak = the BURST cuBLAS figure: the timed region is milliseconds at full clock (see `clocks`), nowhere near the power-capped
    # state the sustained figure was measured in; frac <= 1 unless the kernel beats cuBLAS
    roofline = {"bound": "tensor", "kernel": dom, "achieved": achieved, "peak": pk["tf_burst"], "unit": "TFLOP/s",
                "frac": achieved / pk["tf_burst"], "traffic": traffic,
                "traffic_source": (f"{NCU_SUMMARY} <- {ncu_row['source']} (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum "
                                   f"of one launch at batch {ncu_row['batch']}, scaled by batch)") if ncu_row else None,
                "algorithmic_bytes": B * (H * W * 128 * 2 + H * W * 4),
                "peak_source": pk["src"] + " (bf16 burst: cuBLAS 8192^3 best of 10)",
                "frac_executed_flops": achieved / pk["tf_burst"] * (GF_PIXEL_EXECUTED / GF_PIXEL) if precision == "bf16" else None,
                "frac_of_sustained_peak": achieved / pk["tf_sust"],
                "kernel_ms": dom_ms, "kernel_share_of_step": dom_ms / (ms / K),
                "head_frac_of_burst_peak": (GF_HEAD * B * K / ms) / pk["tf_burst"],
                "head_frac_of_sustained_peak": (GF_HEAD * B * K / ms) / pk["tf_sust"]}

    # ---- end to end through the public module API with HOST buffers ----------------------------------
    pipe = None
    e2e_nchw = None
    h2d_ceiling = None
    if precision == "bf16":
        from objcavit_b200.modules import HostPipeline
        pipe = HostPipeline(head, chunks=4)   # chunked H2D / compute / D2H overlap (the public host-buffer entry point)
        depth_host = torch.empty(B, 1, H, W, dtype=torch.float32).pin_memory()
        edges_host = torch.empty(B, 257, dtype=torch.float32).pin_memory()

    def e2e_step():
        if pipe is not None:
            return pipe(x_host, objs_host, boxes_host, depth_host, edges_host)
        x = x_host.to(dev, non_blocking=True)
        ob = [o.to(dev, non_blocking=True) for o in objs_host]
        bx = [b.to(dev, non_blocking=True) for b in boxes_host]
        out = head_call(head, x, ob, bx)
        return out.depth_pred.to("cpu", non_blocking=False)

    for _ in range(3):
        e2e_step()
    ms_e2e, _ = ctx.timed_median(e2e_step, K)
    e2e_value = world * B * K / (ms_e2e * 1e-3)
    small = sum(o.numel() * 4 for o in objs_host) + sum(b.numel() * 4 for b in boxes_host)
    h2d = x_host.numel() * x_host.element_size() + small
    d2h = B * H * W * 4 + B * 257 * 4
    if pipe is not None:
        # plain-copy ceiling: one chunk of the bf16 features per copy, all ranks at once
        lo_gbs, sum_gbs = h2d_probe(ctx, x_host.numel() * 2 // 4)
        h2d_ceiling = {"gb_per_s_per_rank_min": lo_gbs, "gb_per_s_all_ranks": sum_gbs,
                       "what": "pinned cudaMemcpyAsync host->device, one copy of a chunk-sized buffer (78.6 MB), best of 6, all ranks at once"}
        # the reference caller's format: fp32 NCHW pinned host features (modules/GraphBins.py:107), converted on the device
        x32_host = feats.contiguous().pin_memory()
        pipe32 = HostPipeline(head, chunks=4)

        def e2e32():
            return pipe32(x32_host, objs_host, boxes_host, depth_host, edges_host)

        for _ in range(3):
            e2e32()
        ms32, _ = ctx.timed_median(e2e32, K)
        h2d32 = x32_host.numel() * 4 + small
        e2e_nchw = {"value": world * B * K / (ms32 * 1e-3), "unit": UNIT, "ms_per_step": ms32 / K, "h2d_bytes_per_step": h2d32,
                    "d2h_bytes_per_step": d2h, "achieved_h2d_gb_per_s_per_rank": h2d32 / (ms32 / K) / 1e6,
                    "frac_of_h2d_ceiling": (h2d32 / (ms32 / K) / 1e6) / lo_gbs,
                    "what": "pinned fp32 NCHW host features (the reference caller's format) -> H2D -> ocv_nchw_to_nhwc_bf16 on the device "
                            "-> head -> D2H, all inside the timed region"}
        del pipe32, x32_host

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W_,
        "ms_per_step": ms / K, "ms_per_step_ranks": rank_stats(r["per"], K), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16" if precision == "bf16" else "f32", "data": "synthetic",
        "config": {"workload": f"{CFG_NAME} head (ObjCAViT + conv_out + bin expectation), {NOBJ} boxes/img, "
                               f"features 128x{H}x{W} ({'bf16 NHWC' if precision == 'bf16' else 'fp32 NCHW'}), "
                               f"batch {B}/GPU, inference",
                   "batch_per_gpu": B, "parallelism": f"dp{world} (batch sharded, no collective)",
                   "l2": f"inputs {x_host.numel() * x_host.element_size() / 1e6:.0f} MB/step > 126 MB L2, no explicit flush",
                   "pipeline": (f"{lanes} lanes (calls overlap on the head's own streams, token path replayed from a CUDA graph), "
                                f"pixel kernel grid cap {a.pixel_ctas or 148}") if lanes else "off",
                   "ms_per_step_single_call": ms_classic / K,
                   "timing": "`value` / `ms_per_step` = the MEDIAN window of back-to-back windows of exactly `steps` steps each, run "
                             "until they add up to >= --min-timed-ms of device time (`windows`); every other figure = the median of 3 "
                             "such windows (CUDA events, barrier + synchronize on both sides, max over ranks per window)",
                   "windows": r["windows"], "host_issue_ms_per_call": r["host_issue_ms_per_call"], "host_bound": r["host_bound"]},
        "gpu_launches": launches, "roofline": roofline,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / K, "input_format": "pinned bf16 NHWC host features (the decoder hand-off format, DESIGN.md §5c)"
                if precision == "bf16" else "pinned fp32 NCHW host features",
                "achieved_h2d_gb_per_s_per_rank": h2d / (ms_e2e / K) / 1e6,
                "frac_of_h2d_ceiling": (h2d / (ms_e2e / K) / 1e6) / h2d_ceiling["gb_per_s_per_rank_min"] if h2d_ceiling else None},
    }
    if fresh:
        out["value_fresh_inputs"] = fresh
    if e2e_nchw:
        out["e2e_fp32_nchw"] = e2e_nchw
        out["h2d_ceiling"] = h2d_ceiling

    # ---- the other BASELINE configurations and the training step, same run -----------------------------------------------
    if precision == "bf16" and not a.only_headline:
        del pipe
        torch.cuda.empty_cache()
        out["configs"] = measure_other_configs(ctx, K, W_, lanes, pk)
        tr = measure_train(ctx, a, K, W_, with_breakdown=False)
        out["train"] = {k: tr[k] for k in ("metric", "value", "unit", "ms_per_step", "ms_per_step_ranks", "steps", "warmup", "scaling",
                                           "dtype", "config", "gpu_launches", "roofline", "allreduce_ms", "allreduce_bytes",
                                           "allreduce_share_of_step", "e2e")}
    out["clocks"] = sampler.stop() if sampler else None

    # ---- baselines beside it (rank 0 of a 1-GPU run only) ------------------------------------------------------------------
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        threads = os.cpu_count() or 1
        n_cpu = 2
        v, secs, kind, ref_out = cpu_reference_rate(feats[:n_cpu], objs[:n_cpu], boxes[:n_cpu], threads)
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": kind, "cpu": cpu_model(),
                               "sample": f"the first {n_cpu} images of the same batch, best of 3 ({secs:.2f} s per pass), "
                                         + ("the unmodified reference modules (oracle/_ref) in eager PyTorch fp32 on the host cores"
                                            if kind == "reference" else "oracle/head.py fp32 on torch CPU")}
        if precision == "bf16" and not a.only_headline:
            try:
                out["parity"] = parity_against_cpu_leg(ctx, head, feats[:n_cpu], objs[:n_cpu], boxes[:n_cpu], ref_out)
            except Exception as e:
                out["parity"] = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
            out["gpu_eager_baseline"] = gpu_eager_baseline(ctx, feats, objs, boxes, K)
    if rank == 0:
        emit(out)
    ctx.close()


if __name__ == "__main__":
    main()
