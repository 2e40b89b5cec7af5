"""GPU, row N1: the decoder's final conv3x3 (128 -> 128, bias, no activation) on the tcgen05 pipeline, bf16 NHWC out, against
``F.conv2d`` on the same bf16-rounded operands (fp32 math).  Tolerance 1e-2 relative to the largest output (bf16 output rounding
2^-8 on values up to ~4 sigma, fp32 accumulation order)."""
import pytest
import torch
import torch.nn.functional as F

from helpers import rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,h,w", [(2, 240, 320), (1, 176, 608), (3, 37, 53)])
def test_final_conv_nhwc_matches_conv2d(cuda, B, h, w):
    from objcavit_b200.modules import FinalConvNHWC
    g = torch.Generator().manual_seed(h)
    x = torch.randn(B, 128, h, w, generator=g)
    m = FinalConvNHWC().to(cuda)
    with torch.no_grad():
        m.weight.copy_(torch.randn(128, 128, 3, 3, generator=g) * 0.03)
        m.bias.copy_(torch.randn(128, generator=g))
    want = F.conv2d(x.bfloat16().float(), m.weight.detach().cpu().bfloat16().float(), m.bias.detach().cpu(), padding=1)
    y = m(x.to(cuda))
    assert y.shape == (B, 128, h, w) and y.dtype == torch.bfloat16 and y.permute(0, 2, 3, 1).is_contiguous()
    assert rel_err(y.float(), want) < 1e-2
    # channels_last bf16 input is consumed in place and gives the same bits
    y2 = m(x.to(cuda).bfloat16().contiguous(memory_format=torch.channels_last))
    assert torch.equal(y, y2)


def test_handoff_feeds_the_head_in_place(cuda):
    """decoder conv -> head with no layout / dtype conversion in between: same result as handing the head the same map."""
    import json, os
    from helpers import GOLD
    from objcavit_b200.config import builtin_args
    from objcavit_b200.modules import FinalConvNHWC, build_head
    from oracle import synth
    name = "nyu_adabins"
    with open(os.path.join(GOLD, f"manifest_{name}.json")) as f:
        sd = synth.make_state_dict(json.load(f), seed=0)
    head = build_head(builtin_args(name), precision="bf16").to(cuda).eval()
    head.load_state_dict(sd, strict=True)
    m = FinalConvNHWC().to(cuda)
    x = synth.make_features(2, 240, 320, seed=3).to(cuda)
    feats = m(x)
    with torch.no_grad():
        a = head(feats)
        b = head(feats.float().contiguous())          # the reference-style fp32 NCHW hand-off of the same values
    assert torch.equal(a.depth_pred, b.depth_pred)
