"""Per-kernel latency of the token-path kernels in a dependent chain replayed from a CUDA graph (L2 warm), PDL on/off via
OCV_NO_PDL.  Probe only."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from objcavit_b200 import ops
dev = torch.device("cuda:0")
B, S, E = 16, 300, 128
M = B * S
g = torch.Generator(device="cpu").manual_seed(0)
def rnd(*s, sc=1.0): return (torch.randn(*s, generator=g) * sc).to(dev)
x32 = rnd(M, E); x16 = x32.bfloat16()
w_qkv = rnd(3 * E, E, sc=0.05).bfloat16(); b_qkv = rnd(3 * E, sc=0.1)
w_o = rnd(E, E, sc=0.05).bfloat16(); b_o = rnd(E, sc=0.1)
w1 = rnd(1024, E, sc=0.05).bfloat16(); b1 = rnd(1024, sc=0.1)
w2 = rnd(E, 1024, sc=0.03).bfloat16(); b2 = rnd(E, sc=0.1)
g1, e1 = torch.ones(E, device=dev), torch.zeros(E, device=dev)

def layer(x32, x16):
    _, qkv = ops.gemm_bf16_ex(x16, w_qkv, b_qkv)
    q = qkv.view(B, S, 3 * E)
    att = ops.mha_core_bf16(q[:, :, :E], q[:, :, E:2 * E], q[:, :, 2 * E:], None, 4)
    y32, y16 = ops.gemm_bf16_ex(att.view(M, E), w_o, b_o, residual=x32, ln=(g1, e1), want32=True)
    _, h16 = ops.gemm_bf16_ex(y16, w1, b1, act=ops.ACT_RELU)
    z32, z16 = ops.gemm_bf16_ex(h16, w2, b2, residual=y32, ln=(g1, e1), want32=True)
    return z32, z16

def chain(fn, n):
    """graph of n dependent calls; returns us per call"""
    s = torch.cuda.Stream()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s):
        fn(2); torch.cuda.synchronize()
        with torch.cuda.graph(gr, stream=s):
            fn(n)
    for _ in range(3): gr.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): gr.replay()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / 10 / n * 1e3

def f_layer(n):
    a, b = x32, x16
    for _ in range(n): a, b = layer(a, b)
def f_qkv(n):
    t = x16
    for _ in range(n):
        _, o = ops.gemm_bf16_ex(t, w_qkv, b_qkv); t = o[:, :E]
def f_oproj(n):
    a, b = x32, x16
    for _ in range(n): a, b = ops.gemm_bf16_ex(b, w_o, b_o, residual=a, ln=(g1, e1), want32=True)
def f_ffn1(n):
    t = x16
    for _ in range(n):
        _, o = ops.gemm_bf16_ex(t, w1, b1, act=ops.ACT_RELU); t = o[:, :E]
def f_ffn2(n):
    _, h = ops.gemm_bf16_ex(x16, w1, b1, act=ops.ACT_RELU)
    a = x32
    for _ in range(n): a, _ = ops.gemm_bf16_ex(h, w2, b2, residual=a, ln=(g1, e1), want32=True, want16=False)
def f_mha(n):
    _, qkv = ops.gemm_bf16_ex(x16, w_qkv, b_qkv)
    q = qkv.view(B, S, 3 * E).clone()
    for _ in range(n):
        o = ops.mha_core_bf16(q[:, :, :E], q[:, :, E:2 * E], q[:, :, 2 * E:], None, 4)
        q[:, :, :E].copy_(o) if False else None
def f_cast(n):
    t = x32
    for _ in range(n): ops.cast_bf16(t)

print("PDL", "off" if os.environ.get("OCV_NO_PDL") == "1" else "on")
for name, fn, n in (("layer (5 kernels)", f_layer, 8), ("qkv gemm", f_qkv, 20), ("out-proj+LN gemm", f_oproj, 20),
                    ("ffn1 gemm", f_ffn1, 20), ("ffn2+LN gemm", f_ffn2, 20), ("mha", f_mha, 20), ("cast", f_cast, 20)):
    print(f"{name:22s} {chain(fn, n):8.2f} us per call")
