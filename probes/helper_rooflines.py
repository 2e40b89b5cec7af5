"""Achieved bandwidth / throughput of the kernels outside the inference head's fused pipeline (rows N1-N4 of SURVEY.md §8f and the
helpers of the training step), timed alone with CUDA events (best of 20 after warm-up; a 512 MB L2 flush write between runs),
against MEASURED_PEAKS.json (HBM copy bandwidth, burst bf16 TFLOP/s: kernels timed alone)."""
import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from objcavit_b200 import ops

dev = torch.device("cuda:0")
pk = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
HBM, TF = pk["hbm_gbs"], pk["bf16_tflops"]
flush = torch.empty(512 << 20, device=dev, dtype=torch.uint8)

def best(fn, n=20):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(n):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts)

rows = []
def hbm(name, bytes_, fn, note=""):
    ms = best(fn); gbs = bytes_ / ms / 1e6
    rows.append(f"{name:34s} {ms*1e3:9.1f} us  {bytes_/1e6:9.1f} MB algorithmic  {gbs:8.0f} GB/s  {gbs/HBM:5.2f} of {HBM:.0f} GB/s  {note}")
def tens(name, flops, fn, note=""):
    ms = best(fn); tf = flops / ms / 1e9
    rows.append(f"{name:34s} {ms*1e3:9.1f} us  {flops/1e9:9.1f} GFLOP algorithmic  {tf:8.0f} TFLOP/s  {tf/TF:5.2f} of {TF:.0f} TFLOP/s (burst)  {note}")

torch.manual_seed(0)
# ---- N1: decoder hand-off conv (config 2 shape: B=16, 240x320) and the training shape
for B, h, w in ((16, 240, 320), (8, 208, 272)):
    x = torch.randn(B, h, w, 128, device=dev).bfloat16()
    w3 = ops.pack_conv3x3_bf16(torch.randn(128, 128, 3, 3, device=dev) * 0.03)
    bias = torch.zeros(128, device=dev)
    tens(f"conv3x3_nhwc_bf16 B{B} {h}x{w}", 2.0 * B * h * w * 128 * 1152, lambda: ops.conv3x3_nhwc_bf16(x, w3, bias))
# ---- N2/N3: evaluation rows at NYU size (prediction 240x320, GT 480x640), B=16
B, h, w, H, W = 16, 240, 320, 480, 640
a = torch.rand(B, 1, h, w, device=dev) * 10; b = torch.rand(B, 1, h, w, device=dev) * 10
gt = torch.rand(B, 1, H, W, device=dev) * 10
hbm("tta_merge B16 240x320", 12.0 * B * h * w, lambda: ops.tta_merge(a, b, 0.001, 10.0))
acc = torch.zeros(9, device=dev, dtype=torch.float64)
hbm("depth_metrics B16 480x640", 4.0 * B * H * W + 4.0 * B * h * w, lambda: ops.depth_metrics(a, gt, 0.001, 10.0, (45, 471, 41, 601), acc))
# ---- training-step helpers at the config-5 shape (B=8, 208x272, K=256)
B, h, w, K = 8, 208, 272, 256
Pn = h * w
logits = torch.randn(B, K, Pn, device=dev)
bias = torch.randn(K, device=dev); cen = torch.rand(B, K, device=dev).cumsum(1); g = torch.randn(B, Pn, device=dev)
hbm("tail_logits_f32", 4.0 * B * K * Pn, lambda: ops.tail_logits(logits, bias, cen))
hbm("tail_logits_bwd_bf16", 6.0 * B * K * Pn, lambda: ops.tail_logits_bwd(logits, bias, cen, g), "(4K read + 2K written per pixel; the second read pass is extra traffic)")
dlT = torch.randn(B, K, Pn, device=dev).bfloat16()
hbm("transpose_bf16 [8,256,P]", 4.0 * B * K * Pn, lambda: ops.transpose_bf16(dlT))
xp = torch.randn(B, 128, h, w, device=dev).bfloat16()
hbm("im2col3x3_bf16", 2.0 * B * Pn * (128 * 9 + 8) + 2.0 * B * Pn * 128, lambda: ops.im2col3x3_bf16(xp))
x32 = torch.randn(B, 128, h, w, device=dev)
hbm("nchw_to_nhwc_bf16 (fp32 in)", 6.0 * B * Pn * 128, lambda: ops.nchw_to_nhwc_bf16(x32))
col = ops.im2col3x3_bf16(xp)
dx3T = torch.randn(128, B * Pn, device=dev).bfloat16()
gw = torch.zeros(128, 1152, device=dev)
tens("conv3x3 wgrad GEMM 128x1152xBP", 2.0 * 128 * 1152 * B * Pn, lambda: ops.gemm_bf16(dx3T, col[:1152], residual=gw, out=gw, split_k=32),
     "(K-major split-K GEMM; 1.16 GB of operands: HBM-bound at this shape)")
Mb = torch.randn(K, 128, device=dev).bfloat16(); x3 = torch.randn(Pn, 128, device=dev).bfloat16(); lg = torch.empty(K, Pn, device=dev)
hbm("logits GEMM 256xPx128 (1 image)", 4.0 * K * Pn + 2.0 * Pn * 128, lambda: ops.gemm_bf16(Mb, x3, out=lg), "(output-write bound)")
print(f"peaks: MEASURED_PEAKS.json hbm {HBM} GB/s, bf16 burst {TF} TFLOP/s (kernels timed alone, L2 flushed before each run)")
print("\n".join(rows))
