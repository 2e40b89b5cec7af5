"""GPU: backward primitives of the fp32 parity path against torch autograd of the same op on the CPU (fp32/fp64)."""
import math

import pytest
import torch
import torch.nn.functional as F

from helpers import rel_err

pytestmark = pytest.mark.gpu


def g(seed):
    return torch.Generator().manual_seed(seed)


@pytest.mark.parametrize("M,N,K,split", [(70, 50, 33, 1), (130, 128, 4800, 8), (256, 128, 300, 3)])
def test_gemm_ex_dense_modes(cuda, M, N, K, split):
    from objcavit_b200 import ops
    A = torch.randn(M, K, generator=g(1)); At = A.t().contiguous()
    Bn = torch.randn(K, N, generator=g(2)); Bt = Bn.t().contiguous()
    C0 = torch.randn(M, N, generator=g(3))
    want = A.double() @ Bn.double()
    for ta, a, lda in ((False, A, K), (True, At, M)):
        for ob, b, ldb in ((0, Bn, N), (1, Bt, K)):
            out = C0.clone().to(cuda)
            ops.gemm_ex(a.to(cuda), b.to(cuda), M, N, K, lda, ldb, out, N, trans_a=ta, op_b=ob, beta=1.0, split_k=split)
            assert rel_err(out, C0.double() + want) < 1e-5
            out2 = torch.empty(M, N, device=cuda)
            ops.gemm_ex(a.to(cuda), b.to(cuda), M, N, K, lda, ldb, out2, N, trans_a=ta, op_b=ob, beta=0.0, split_k=split)
            assert rel_err(out2, want) < 1e-5


def test_gemm_ex_batched_reduce(cuda):
    """dWout = sum_b dl[b] @ ram[b]^T (reduce over the batch) and a per-batch product."""
    from objcavit_b200 import ops
    Bn, M, N, K = 3, 40, 24, 500
    A = torch.randn(Bn, M, K, generator=g(4)); Bm = torch.randn(Bn, N, K, generator=g(5))
    out = torch.zeros(M, N, device=cuda)
    ops.gemm_ex(A.to(cuda), Bm.to(cuda), M, N, K, K, K, out, N, op_b=1, batch=Bn, stride_a=M * K, stride_b=N * K, reduce_batch=True,
                split_k=4)
    assert rel_err(out, torch.einsum("bmk,bnk->mn", A.double(), Bm.double())) < 1e-5
    out2 = torch.empty(Bn, M, N, device=cuda)
    ops.gemm_ex(A.to(cuda), Bm.to(cuda), M, N, K, K, K, out2, N, op_b=1, batch=Bn, stride_a=M * K, stride_b=N * K, stride_c=M * N)
    assert rel_err(out2, torch.einsum("bmk,bnk->bmn", A.double(), Bm.double())) < 1e-5


def test_conv_weight_gradients_by_gathered_gemm(cuda):
    """conv3x3 (pad 1) and the 16x16/stride-16 patch conv: weight gradients without an im2col buffer."""
    from objcavit_b200 import ops
    B, C, h, w, O = 2, 6, 11, 13, 5
    x = torch.randn(B, C, h, w, generator=g(6))
    wt = torch.randn(O, C, 3, 3, generator=g(7), requires_grad=True)
    dy = torch.randn(B, O, h, w, generator=g(8))
    F.conv2d(x, wt, padding=1).backward(dy)
    dw = torch.zeros(O, C * 9, device=cuda)
    ops.gemm_ex(dy.to(cuda), x.to(cuda), O, C * 9, h * w, h * w, 0, dw, C * 9, op_b=3, batch=B, stride_a=O * h * w,
                stride_b=C * h * w, reduce_batch=True, split_k=2, geom=(C, h, w, 0))
    assert rel_err(dw.view(O, C, 3, 3), wt.grad) < 1e-5
    P, E = 4, 7
    h2, w2 = 9, 14          # ragged: floor(h/P) x floor(w/P) patches
    x2 = torch.randn(B, C, h2, w2, generator=g(9))
    wp = torch.randn(E, C, P, P, generator=g(10), requires_grad=True)
    tok = F.conv2d(x2, wp, stride=P)                      # [B,E,gh,gw]
    dtok = torch.randn(B, tok.shape[2] * tok.shape[3], E, generator=g(11))      # [B,S,E] like the head's token layout
    tok.flatten(2).permute(0, 2, 1).backward(dtok)
    S = dtok.shape[1]
    dwp = torch.zeros(E, C * P * P, device=cuda)
    ops.gemm_ex(dtok.view(B * S, E).to(cuda), x2.to(cuda), E, C * P * P, B * S, E, 0, dwp, C * P * P, trans_a=True, op_b=2,
                geom=(C, h2, w2, P))
    assert rel_err(dwp.view(E, C, P, P), wp.grad) < 1e-5


def test_linear_grads_and_reductions(cuda):
    from objcavit_b200 import ops
    M, N, K = 333, 96, 128
    x = torch.randn(M, K, generator=g(12)); w = torch.randn(N, K, generator=g(13), requires_grad=True)
    b = torch.randn(N, generator=g(14), requires_grad=True)
    xr = x.clone().requires_grad_(True)
    dy = torch.randn(M, N, generator=g(15))
    F.linear(xr, w, b).backward(dy)
    assert rel_err(ops.linear_dgrad(dy.to(cuda), w.detach().to(cuda)), xr.grad) < 1e-5
    dw = torch.ones(N, K, device=cuda); db = torch.ones(N, device=cuda)
    ops.linear_wgrad(dy.to(cuda), x.to(cuda), dw, db, beta=1.0)          # accumulates into existing gradients
    assert rel_err(dw, w.grad + 1) < 1e-5 and rel_err(db, b.grad + 1) < 1e-5
    t = torch.randn(2, 3, 5, 77, generator=g(16))
    out = torch.empty(2, 5, device=cuda)
    ops.reduce_channels(t.to(cuda), out, 2, 3, 5, 77)
    assert rel_err(out, t.sum(dim=(1, 3))) < 1e-5
    o = F.leaky_relu(torch.randn(50, 20, generator=g(17)), 0.01)
    d = torch.randn(50, 20, generator=g(18))
    assert torch.equal(ops.act_bwd(d.to(cuda), o.to(cuda), ops.ACT_LEAKY_RELU).cpu(), torch.where(o > 0, d, 0.01 * d))
    assert torch.equal(ops.act_bwd(d.to(cuda), o.relu().to(cuda), ops.ACT_RELU).cpu(), torch.where(o > 0, d, torch.zeros(())))


def test_layernorm_backward(cuda):
    from objcavit_b200 import ops
    rows, E = 301, 128
    z = (torch.randn(rows, E, generator=g(19)) * 2 + 0.5).requires_grad_(True)
    gamma = (1 + 0.1 * torch.randn(E, generator=g(20))).requires_grad_(True)
    beta = torch.randn(E, generator=g(21)).requires_grad_(True)
    dy = torch.randn(rows, E, generator=g(22))
    F.layer_norm(z, (E,), gamma, beta, 1e-5).backward(dy)
    dg = torch.zeros(E, device=cuda); db = torch.zeros(E, device=cuda)
    dz = ops.layernorm_bwd(dy.to(cuda), z.detach().to(cuda), gamma.detach().to(cuda), dg, db, beta=1.0)
    assert rel_err(dz, z.grad) < 1e-5 and rel_err(dg, gamma.grad) < 1e-5 and rel_err(db, beta.grad) < 1e-5


@pytest.mark.parametrize("Sq,Sk,masked", [(37, 37, False), (20, 45, True), (221, 221, False)])
def test_mha_core_backward(cuda, Sq, Sk, masked):
    from objcavit_b200 import ops
    B, H, d = 2, 4, 32
    q = torch.randn(B, Sq, H * d, generator=g(23)).requires_grad_(True)
    k = torch.randn(B, Sk, H * d, generator=g(24)).requires_grad_(True)
    v = torch.randn(B, Sk, H * d, generator=g(25)).requires_grad_(True)
    mask = None
    if masked:
        mask = torch.zeros(B, Sk, dtype=torch.bool)
        mask[0, 30:] = True
        mask[1, 5:9] = True
    qh, kh, vh = (t.view(B, -1, H, d).transpose(1, 2) for t in (q, k, v))
    s = qh @ kh.transpose(-1, -2) / math.sqrt(d)
    if mask is not None:
        s = s.masked_fill(mask[:, None, None, :], float("-inf"))
    out = (s.softmax(-1) @ vh).transpose(1, 2).reshape(B, Sq, H * d)
    dout = torch.randn(B, Sq, H * d, generator=g(26))
    out.backward(dout)
    m8 = None if mask is None else mask.to(torch.uint8).to(cuda)
    got_out = ops.mha_core(q.detach().to(cuda), k.detach().to(cuda), v.detach().to(cuda), m8, H)
    dq, dk, dv = ops.mha_core_bwd(q.detach().to(cuda), k.detach().to(cuda), v.detach().to(cuda), got_out, dout.to(cuda), m8, H)
    assert rel_err(dq, q.grad) < 2e-5 and rel_err(dk, k.grad) < 2e-5 and rel_err(dv, v.grad) < 2e-5


def test_tail_and_bins_backward(cuda):
    from objcavit_b200 import ops
    B, K, P = 2, 256, 700
    logits = (torch.randn(B, K, P, generator=g(27)) * 2).requires_grad_(True)
    y = torch.randn(B, K, generator=g(28)).requires_grad_(True)
    lo, hi = 0.001, 10.0
    u = F.relu(y) + 0.1
    wn = u / u.sum(1, keepdim=True)
    edges = torch.cumsum(F.pad((hi - lo) * wn, (1, 0), value=lo), 1)
    centres = 0.5 * (edges[:, :-1] + edges[:, 1:])
    depth = (logits.softmax(1) * centres[:, :, None]).sum(1)                      # [B,P]
    gd = torch.randn(B, P, generator=g(29))
    ge = torch.randn(B, K + 1, generator=g(30))
    (depth * gd).sum().backward(retain_graph=True)
    dl_want, dy_from_depth = logits.grad.clone(), y.grad.clone()
    dl, pg = ops.tail_bwd(logits.detach().clone().to(cuda), centres.detach().to(cuda), gd.to(cuda))
    assert rel_err(dl, dl_want) < 1e-5
    dcent = torch.empty(B, K, device=cuda)
    ops.reduce_channels(pg, dcent, B, 1, K, P)
    y.grad = None
    ((depth * gd).sum() + (edges * ge).sum()).backward()
    dy = ops.bins_bwd(y.detach().to(cuda), ge.to(cuda), dcent, lo, hi)
    assert rel_err(dy, y.grad) < 1e-4
    assert rel_err(ops.bins_bwd(y.detach().to(cuda), None, dcent, lo, hi), dy_from_depth) < 1e-4
