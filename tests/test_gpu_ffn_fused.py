"""GPU: the fused feed-forward kernel (ocv_ffn_fused_bf16) against the two-GEMM path it replaces (same bf16 rounding of the hidden
activations, same K order: expected equal to fp32 rounding noise) and against torch fp32 on the bf16-rounded operands (2e-2)."""
import pytest
import torch
import torch.nn.functional as F

from helpers import rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,Fh", [(4800, 1024), (333, 1024), (128, 128), (512, 256)])
def test_ffn_fused_matches_two_gemms_and_torch(cuda, M, Fh):
    from objcavit_b200 import ops
    g = torch.Generator().manual_seed(M + Fh)
    x32 = torch.randn(M, 128, generator=g)
    w1 = torch.randn(Fh, 128, generator=g) * 0.08
    b1 = torch.randn(Fh, generator=g) * 0.1
    w2 = torch.randn(128, Fh, generator=g) * 0.03
    b2 = torch.randn(128, generator=g) * 0.1
    gamma = 1 + 0.1 * torch.randn(128, generator=g)
    beta = 0.1 * torch.randn(128, generator=g)
    zero = (torch.rand(M, generator=g) < 0.1).to(torch.uint8)
    d = lambda t: t.to(cuda)
    x16, w1h, w2h = d(x32).bfloat16(), d(w1).bfloat16(), d(w2).bfloat16()
    for zr in (None, d(zero)):
        o32, o16 = ops.ffn_fused_bf16(x16, d(x32), w1h, d(b1), w2h, d(b2), d(gamma), d(beta), zr)
        _, h16 = ops.gemm_bf16_ex(x16, w1h, d(b1), act=ops.ACT_RELU)
        r32, r16 = ops.gemm_bf16_ex(h16, w2h, d(b2), residual=d(x32), ln=(d(gamma), d(beta)), zero_rows=zr, want32=True)
        assert rel_err(o32, r32) < 1e-5
        assert rel_err(o16.float(), r16.float()) < 1e-2
        h = F.relu(x16.float().cpu() @ w1h.float().cpu().T + b1).bfloat16().float()
        want = F.layer_norm(x32 + h @ w2h.float().cpu().T + b2, (128,), gamma, beta, 1e-5)
        if zr is not None:
            want = want.masked_fill(zero.bool()[:, None], 0.0)
        assert rel_err(o32, want) < 2e-3
    torch.cuda.synchronize()
