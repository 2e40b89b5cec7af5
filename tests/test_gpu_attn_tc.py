"""GPU: the tcgen05 attention core (csrc/attn_tc.cu) against fp64 torch math on the same bf16 operands — packed qkv and separate
q / k / v views, key-padding masks incl. an all-masked image (NaN rows like torch), every Sk regime (one 64-key tile, an odd tile
count, two MMAs of N <= 256, the 448-key limit), query tiles with a ragged tail, the fp32 copy and the softmax statistics, and the
probability dropout against the fp32 kernel's mask convention.  reference sites: nn.MultiheadAttention inside
modules/ObjCAViT.py:155-169, 188, 195-207 and modules/layers.py:8-9.
Tolerance: 1.5e-2 of the largest |output| (P is rounded to bf16 before P V, the output is stored in bf16)."""
import math

import pytest
import torch

from objcavit_b200 import ops

pytestmark = pytest.mark.gpu


def _ref(q, k, v, mask, H=4):
    B, Sq, E = q.shape
    Sk, d = k.shape[1], E // H
    qh = q.double().view(B, Sq, H, d).transpose(1, 2)
    kh = k.double().view(B, Sk, H, d).transpose(1, 2)
    vh = v.double().view(B, Sk, H, d).transpose(1, 2)
    s = qh @ kh.transpose(-1, -2) / math.sqrt(d)
    if mask is not None:
        s = s.masked_fill(mask.bool()[:, None, None, :], float("-inf"))
    p = torch.softmax(s, dim=-1)
    return (p @ vh).transpose(1, 2).reshape(B, Sq, E), s


@pytest.mark.parametrize("B,Sq,Sk,packed,masked", [
    (16, 300, 300, True, False),     # config 2 image encoder
    (8, 418, 418, True, False),      # KITTI: two score MMAs, four query tiles
    (8, 221, 221, True, False),      # training shape
    (16, 32, 32, True, True),        # object encoder with key padding
    (4, 300, 300, False, True),      # cross attention: separate q / k / v tensors, right-padded mask (F5)
    (3, 64, 448, False, False),      # the Sk limit
    (2, 130, 50, False, True),       # one key tile: the second softmax half has no keys
])
def test_attention_core_on_tcgen05(cuda, B, Sq, Sk, packed, masked):
    g = torch.Generator().manual_seed(Sq * 1000 + Sk)
    E = 128
    if packed:
        qkv = torch.randn(B, Sq, 3 * E, generator=g).to(cuda).bfloat16()
        q, k, v = qkv[:, :, 0:E], qkv[:, :, E:2 * E], qkv[:, :, 2 * E:]
    else:
        q = torch.randn(B, Sq, E, generator=g).to(cuda).bfloat16()
        k = torch.randn(B, Sk, E, generator=g).to(cuda).bfloat16()
        v = torch.randn(B, Sk, E, generator=g).to(cuda).bfloat16()
    mask = None
    if masked:
        mask = torch.zeros(B, Sk, dtype=torch.uint8, device=cuda)
        for b in range(B):
            mask[b, max(1, (b * 7 + 3) % Sk):] = 1
        mask[B - 1, :] = 1                       # an image whose keys are all padding: NaN rows, as torch gives
    want, scores = _ref(q, k, v, mask)
    out16, out32, stats = ops.mha_core_tc_bf16(q, k, v, mask, want32=True, want_stats=True)
    torch.cuda.synchronize()
    assert torch.equal(torch.isnan(out32), torch.isnan(want)), "NaN pattern differs"
    ok = ~torch.isnan(want)
    scale = float(want[ok].abs().max())
    assert float((out32.double()[ok] - want[ok]).abs().max()) < 1.5e-2 * scale
    assert float((out16.double()[ok] - want[ok]).abs().max()) < 1.5e-2 * scale
    # statistics: log2-domain row maximum and the sum of 2^(s - max)
    s2 = scores * math.log2(math.e)
    rows = ok.view(B, Sq, 4, 32)[..., 0].permute(0, 2, 1)                     # [B,H,Sq] rows with at least one key
    mx = s2.max(dim=-1).values
    assert float((stats[..., 0].double()[rows] - mx[rows]).abs().max()) < 2e-2
    sm = torch.exp2(s2 - mx[..., None]).sum(-1)
    assert float(((stats[..., 1].double()[rows] - sm[rows]).abs() / sm[rows]).max()) < 1e-2
    # the dispatching entry point (mma.sync kernel by default, tcgen05 with OCV_ATTN_TC=1) agrees to bf16 rounding
    again = ops.mha_core_bf16(q, k, v, mask)
    torch.cuda.synchronize()
    assert float((again.double()[ok] - out16.double()[ok]).abs().max()) < 1.5e-2 * scale


def test_probability_dropout_uses_the_fp32_kernels_masks(cuda):
    """Same (seed, element) -> same keep decision as ocv_mha_core_dropout_f32, so the fp32 backward kernel regenerates the masks
    of a tensor-core forward."""
    g = torch.Generator().manual_seed(4)
    B, S, E, p, seed = 3, 221, 128, 0.25, 987654321
    qkv = torch.randn(B, S, 3 * E, generator=g).to(cuda)
    q, k, v = qkv[:, :, 0:E], qkv[:, :, E:2 * E], qkv[:, :, 2 * E:]
    want = ops.mha_core(q, k, v, None, 4, drop=(p, seed))
    qb = qkv.bfloat16()
    _, got, _ = ops.mha_core_tc_bf16(qb[:, :, 0:E], qb[:, :, E:2 * E], qb[:, :, 2 * E:], None, want32=True, drop=(p, seed))
    torch.cuda.synchronize()
    assert float((got - want).abs().max()) < 3e-2 * float(want.abs().max())


@pytest.mark.parametrize("B,Sq,Sk,masked,p", [
    (8, 221, 221, False, 0.0),       # config 5 image encoder
    (8, 32, 32, True, 0.0),          # object encoder with key padding
    (3, 242, 242, False, 0.0),       # KITTI-train tokens
    (2, 130, 200, True, 0.0),        # Sq != Sk, ragged tiles
    (4, 221, 221, False, 0.2),       # probability dropout: masks regenerated from (seed, element)
])
def test_attention_backward_on_tcgen05(cuda, B, Sq, Sk, masked, p):
    """dq / dk / dv of the tcgen05 backward against fp64 autograd through softmax attention on the same bf16 q / k / v and the
    same dropout masks (those of the forward kernel, recovered from its output).  Tolerance 2e-2 of the largest |gradient|: P and
    dS are rounded to bf16 before the three gradient GEMMs."""
    g = torch.Generator().manual_seed(Sq * 7 + Sk)
    E, H, d = 128, 4, 32
    q = torch.randn(B, Sq, E, generator=g).to(cuda).bfloat16()
    k = torch.randn(B, Sk, E, generator=g).to(cuda).bfloat16()
    v = torch.randn(B, Sk, E, generator=g).to(cuda).bfloat16()
    dout = torch.randn(B, Sq, E, generator=g).to(cuda).bfloat16()
    mask = None
    if masked:
        mask = torch.zeros(B, Sk, dtype=torch.uint8, device=cuda)
        for b in range(B):
            mask[b, max(1, (b * 5 + 9) % Sk):] = 1
    seed = 424242
    drop = (p, seed) if p > 0 else None
    out16, out32, stats = ops.mha_core_tc_bf16(q, k, v, mask, want32=True, want_stats=True, drop=drop)
    (dq, dk, dv), (dq16, dk16, dv16) = ops.mha_core_bwd_tc_bf16(q, k, v, out16, dout, stats, mask, drop=drop)
    torch.cuda.synchronize()
    # reference: fp64 autograd; the dropout keep mask of element (b,h,q,k) is regenerated with the elementwise kernel's convention
    q64, k64, v64 = (t.double().requires_grad_(True) for t in (q, k, v))
    qh = q64.view(B, Sq, H, d).transpose(1, 2)
    kh = k64.view(B, Sk, H, d).transpose(1, 2)
    vh = v64.view(B, Sk, H, d).transpose(1, 2)
    s = qh @ kh.transpose(-1, -2) / math.sqrt(d)
    if mask is not None:
        s = s.masked_fill(mask.bool()[:, None, None, :], float("-inf"))
    P = torch.softmax(s, dim=-1)
    if p > 0:
        ones = torch.ones(B * H * Sq * Sk, device=cuda)
        keep = ops.dropout(ones, p, seed).view(B, H, Sq, Sk).double()         # 0 or 1 / (1 - p)
        P = P * keep
    o = (P @ vh).transpose(1, 2).reshape(B, Sq, E)
    (o * dout.double()).sum().backward()
    for name, got, got16, want in (("dq", dq, dq16, q64.grad), ("dk", dk, dk16, k64.grad), ("dv", dv, dv16, v64.grad)):
        scale = float(want.abs().max())
        err = float((got.double() - want).abs().max()) / scale
        err16 = float((got16.double() - want).abs().max()) / scale
        assert err < 2e-2 and err16 < 2.5e-2, (name, err, err16)
