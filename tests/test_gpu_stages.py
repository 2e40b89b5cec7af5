"""GPU: every C-ABI stage against the CPU oracle / plain torch fp32 on the same seeded inputs.
Index work is compared bit-exactly; fp32 stages to 1e-5..1e-4 relative (tolerances from north_star)."""
import math
import os

import pytest
import torch
import torch.nn.functional as F

from helpers import GOLD, rel_err
from oracle import head as oh
from oracle import synth

pytestmark = pytest.mark.gpu


def g(seed):
    return torch.Generator().manual_seed(seed)


def test_linear_and_batched_matmul(cuda):
    from objcavit_b200 import ops
    a = torch.randn(301, 130, generator=g(1)); w = torch.randn(77, 130, generator=g(2)); b = torch.randn(77, generator=g(3))
    r = torch.randn(301, 77, generator=g(4))
    want = F.leaky_relu(a @ w.T + b, 0.01) + r
    got = ops.linear(a.to(cuda), w.to(cuda), b.to(cuda), ops.ACT_LEAKY_RELU, r.to(cuda))
    assert rel_err(got, want) < 1e-5
    # strided A (token-0 view) and relu
    tok = torch.randn(3, 20, 128, generator=g(5))
    w2 = torch.randn(256, 128, generator=g(6))
    got = ops.linear(tok.to(cuda)[:, 0, :], w2.to(cuda), None, ops.ACT_RELU)
    assert rel_err(got, torch.relu(tok[:, 0, :] @ w2.T)) < 1e-5
    # PixelWiseDotProduct form
    q = torch.randn(2, 131, 128, generator=g(7)); x = torch.randn(2, 128, 333, generator=g(8))
    got = ops.batched_matmul_nn(q.to(cuda)[:, 1:129, :], x.to(cuda))
    assert rel_err(got, q[:, 1:129, :] @ x) < 1e-5


def test_patch_embed(cuda):
    from objcavit_b200 import ops
    x = torch.randn(2, 128, 48, 64, generator=g(1)); w = torch.randn(128, 128, 16, 16, generator=g(2)) * 0.01
    b = torch.randn(128, generator=g(3)); pos = torch.rand(50, 128, generator=g(4))
    want = F.conv2d(x, w, b, stride=16).flatten(2).permute(0, 2, 1) + pos[:12]
    got = ops.patch_embed(x.to(cuda), w.to(cuda), b.to(cuda), pos.to(cuda))
    assert rel_err(got, want) < 1e-5


def test_add_layernorm_and_zero_rows(cuda):
    from objcavit_b200 import ops
    x = torch.randn(37, 128, generator=g(1)); r = torch.randn(37, 128, generator=g(2))
    gm = torch.randn(128, generator=g(3)); bt = torch.randn(128, generator=g(4))
    z = (torch.arange(37) % 5 == 0).to(torch.uint8)
    want = F.layer_norm(x + r, (128,), gm, bt, 1e-5).masked_fill(z.bool()[:, None], 0.0)
    got = ops.add_layernorm(x.to(cuda), r.to(cuda), gm.to(cuda), bt.to(cuda), z.to(cuda))
    assert rel_err(got, want) < 1e-5


@pytest.mark.parametrize("Sq,Sk", [(300, 300), (45, 77), (9, 9)])
def test_mha_core_with_key_padding(cuda, Sq, Sk):
    from objcavit_b200 import ops
    B = 3
    qkv = torch.randn(B, max(Sq, Sk), 384, generator=g(Sq))
    q, k, v = qkv[:, :Sq, 0:128], qkv[:, :Sk, 128:256], qkv[:, :Sk, 256:]
    mask = torch.zeros(B, Sk, dtype=torch.bool)
    mask[0, Sk // 2:] = True
    mask[2, 1:] = True
    sc = (q.view(B, Sq, 4, 32).transpose(1, 2) / math.sqrt(32)) @ k.reshape(B, Sk, 4, 32).transpose(1, 2).transpose(-1, -2)
    sc = sc.masked_fill(mask[:, None, None, :], float("-inf"))
    want = (torch.softmax(sc, -1) @ v.reshape(B, Sk, 4, 32).transpose(1, 2)).transpose(1, 2).reshape(B, Sq, 128)
    d = qkv.to(cuda)
    got = ops.mha_core(d[:, :Sq, 0:128], d[:, :Sk, 128:256], d[:, :Sk, 256:], mask.to(torch.uint8).to(cuda))
    assert rel_err(got, want) < 1e-5


def test_pad_objects_bit_exact(cuda):
    from objcavit_b200 import engine, ops
    counts = [3, 1, 7, 4]
    rows = torch.randn(sum(counts), 128, generator=g(1))
    seqs = list(torch.split(rows, counts))
    want, wmask = oh.pad_objects(seqs)
    ob = engine.ObjectBatch(rows.to(cuda), counts)
    got, mask = ops.pad_objects(ob.rows, ob.offsets, 4, 7, 7, 0.0001, False)
    assert torch.equal(got.cpu(), want) and torch.equal(mask.cpu().bool(), wmask)
    want_s, wmask_s = oh.pad_objects(seqs, S=20)          # left-padded objects, right-padded mask (F5)
    uni = torch.arange(0, 5 * 7, 7, dtype=torch.int32).to(cuda)
    got_s, mask_s = ops.pad_objects(got.view(28, 128), uni, 4, 20, 7, 0.0001, True, mask_offsets=ob.offsets)
    assert torch.equal(got_s.cpu(), want_s) and torch.equal(mask_s.cpu().bool(), wmask_s)


def test_grid_sample_index_mapping_bit_exact(cuda):
    from objcavit_b200 import ops
    gold = torch.load(os.path.join(GOLD, "grid_ops.pt"))
    table = gold["table"]
    for (h, w) in ((176, 608), (240, 320)):
        grid, gh, gw = oh._grid_for(table, h, w)
        pts = gold["cases"][f"gs_{h}x{w}"]["pts"]
        want, widx, _ = oh.grid_sample_points(grid, pts[:, 0], pts[:, 1])
        got, idx = ops.grid_sample_pos(table.to(cuda), pts.to(cuda).contiguous(), gh, gw, 0, want_idx=True)
        assert torch.equal(idx.cpu().long(), widx)
        assert torch.equal(got.cpu(), want)                       # same op order as the restatement
        assert float((got.cpu() - gold["cases"][f"gs_{h}x{w}"]["want"]).abs().max()) < 5e-6   # vs F.grid_sample


def test_ps_roi_align_bit_exact_incl_nan(cuda):
    from objcavit_b200 import ops
    gold = torch.load(os.path.join(GOLD, "grid_ops.pt"))
    table = gold["table"]
    for (h, w) in ((176, 608), (240, 320)):
        grid, gh, gw = oh._grid_for(table, h, w)
        for tag, scale in (("obj", 1 / 32), ("img", 1 / 16)):
            c = gold["cases"][f"roi_{tag}_{h}x{w}"]          # want = torchvision.ops.ps_roi_align itself
            got, plan = ops.ps_roi_align_pos(table.to(cuda), c["xyxy"].to(cuda).contiguous(), gh, gw, scale, False,
                                             want_plan=True)
            _, plans = oh.ps_roi_align_1x1(grid, c["xyxy"], scale)
            assert plan.cpu().tolist() == [[p[0], p[1]] for p in plans]
            want = c["want"]
            assert torch.equal(torch.isnan(got.cpu()), torch.isnan(want))
            m = ~torch.isnan(want)
            assert torch.equal(got.cpu()[m], want[m])


def test_grid_random_module_quirks(cuda):
    """GridRandomPositionalEmbeddings.forward through the module surface, 'img' and 'obj', both modes."""
    from objcavit_b200.config import builtin_args
    from objcavit_b200.modules import GridRandomPositionalEmbeddings
    args = builtin_args("kitti_ocv_grid_random_roi_align")
    feats = torch.zeros(2, 128, 176, 208)
    _, boxes = synth.make_objects(1, 176, 208, 12, seed=3)
    boxes = torch.cat([boxes[0], torch.tensor([[-1., -1., -1., -1.]])])
    pc = oh.patch_coords(2, 11, 13)
    for mode in ("centre", "roi_align"):
        mod = GridRandomPositionalEmbeddings(args, 128, 16, mode).to(cuda)
        table = mod.positional_encodings.data.cpu()
        n = 2 if mode == "centre" else 4
        want_obj = oh.grid_random_pos(table, mode, boxes[:, :n], 176, 208, "obj")
        want_img = oh.grid_random_pos(table, mode, pc[..., :n], 176, 208, "img")
        with torch.no_grad():
            got_obj = mod(boxes[:, :n].to(cuda), feats.to(cuda), "obj").cpu()
            got_img = mod(pc[..., :n].to(cuda), feats.to(cuda), "img").cpu()
        assert torch.equal(torch.isnan(got_obj), torch.isnan(want_obj))
        m = ~torch.isnan(want_obj)
        assert torch.equal(got_obj[m], want_obj[m]) and torch.equal(got_img, want_img)


@pytest.mark.parametrize("norm", ["linear", "softmax", "sigmoid"])
def test_bins(cuda, norm):
    from objcavit_b200 import ops
    y = torch.randn(5, 256, generator=g(2)) * 3
    if norm == "linear":
        wn = torch.relu(y) + 0.1
    elif norm == "softmax":
        wn = torch.exp(y - y.max(1, keepdim=True).values)
    else:
        wn = torch.sigmoid(y)
    wn = wn / wn.sum(1, keepdim=True)
    edges, centres = oh.bin_edges_and_centres(wn, 0.001, 80.0)
    w, e, c = ops.bins(y.to(cuda), norm, 0.001, 80.0)
    assert rel_err(w, wn) < 1e-5 and rel_err(e, edges) < 1e-5 and rel_err(c, centres) < 1e-5
    e2, c2 = ops.edges_from_widths(wn.to(cuda), 0.001, 80.0)
    assert rel_err(e2, edges) < 5e-6 and rel_err(c2, centres) < 5e-6   # 256-term fp32 running sum


def test_conv3x3(cuda):
    from objcavit_b200 import ops
    x = torch.randn(2, 128, 37, 70, generator=g(1)); w = torch.randn(128, 128, 3, 3, generator=g(2)) * 0.03
    b = torch.randn(128, generator=g(3))
    got = ops.conv3x3(x.to(cuda), w.to(cuda), b.to(cuda))
    assert rel_err(got, F.conv2d(x, w, b, padding=1)) < 1e-5


def test_tail(cuda):
    from objcavit_b200 import ops
    ram = torch.randn(2, 128, 23, 41, generator=g(1)) * 2
    w = torch.randn(256, 128, 1, 1, generator=g(2)) * 0.2; b = torch.randn(256, generator=g(3))
    wn = torch.rand(2, 256, generator=g(4)) + 0.1
    wn = wn / wn.sum(1, keepdim=True)
    edges, centres = oh.bin_edges_and_centres(wn, 0.001, 10.0)
    want = (torch.softmax(F.conv2d(ram, w, b), 1) * centres[:, :, None, None]).sum(1, keepdim=True)
    got = ops.tail(ram.to(cuda), w.to(cuda), b.to(cuda), centres.to(cuda))
    assert rel_err(got, want) < 1e-5


@pytest.mark.parametrize("M,N,K,split", [(300, 128, 128, 1), (4800, 384, 128, 1), (520, 1024, 128, 1), (777, 128, 1024, 4),
                                          (600, 128, 32768, 8)])
def test_gemm_bf16_tcgen05(cuda, M, N, K, split):
    """tcgen05 GEMM vs an fp32 matmul of the same bf16-rounded operands."""
    from objcavit_b200 import ops
    a = (torch.randn(M, K, generator=g(M)) * 0.5).bfloat16(); w = (torch.randn(N, K, generator=g(N + K)) * 0.5).bfloat16()
    b = torch.randn(N, generator=g(3)); r = torch.randn(M, N, generator=g(4))
    want = torch.relu(a.float() @ w.float().T + b) + r
    got = ops.gemm_bf16(a.to(cuda), w.to(cuda), b.to(cuda), ops.ACT_RELU, r.to(cuda), split_k=split)
    assert rel_err(got, want) < 2e-5 * max(1.0, math.sqrt(K / 128))
    got16 = ops.gemm_bf16(a.to(cuda), w.to(cuda), b.to(cuda), ops.ACT_RELU, None, out_dtype=torch.bfloat16, split_k=split)
    assert rel_err(got16.float(), torch.relu(a.float() @ w.float().T + b)) < 1e-2


def test_layout_and_weight_packing(cuda):
    from objcavit_b200 import ops
    x = torch.randn(2, 128, 19, 45, generator=g(1))
    got = ops.nchw_to_nhwc_bf16(x.to(cuda))
    assert torch.equal(got.cpu(), x.permute(0, 2, 3, 1).contiguous().bfloat16())
    got = ops.nchw_to_nhwc_bf16(x.bfloat16().to(cuda))
    assert torch.equal(got.cpu(), x.bfloat16().permute(0, 2, 3, 1).contiguous())
    xb = torch.randn(3, 128, 24, 40, generator=g(5))               # h*w % 64 == 0, fp32, 128 channels: the tiled fast path
    got = ops.nchw_to_nhwc_bf16(xb.to(cuda))
    assert torch.equal(got.cpu(), xb.permute(0, 2, 3, 1).contiguous().bfloat16())
    w = torch.randn(128, 128, 3, 3, generator=g(2))
    assert torch.equal(ops.pack_conv3x3_bf16(w.to(cuda)).cpu(), w.permute(2, 3, 0, 1).reshape(9, 128, 128).bfloat16())
    wout = torch.randn(256, 128, 1, 1, generator=g(3)); bout = torch.randn(256, generator=g(4))
    tok = torch.randn(2, 140, 128, generator=g(5)); b3 = torch.randn(128, generator=g(6))
    Mb, blog = ops.fold_queries_bf16(wout.to(cuda), bout.to(cuda), tok.to(cuda)[:, 1:129, :], b3.to(cuda), 1.0)
    want = wout.view(256, 128) @ tok[:, 1:129, :]
    assert rel_err(Mb.float(), want) < 1e-2
    assert rel_err(blog, bout + want @ b3) < 1e-5


def _pixel_reference(x, w3, b3, wout, bout, queries, centres):
    """conv3x3 -> dot product -> conv_out -> softmax -> sum p*c, fp32 torch on the bf16-rounded operands the
    kernel consumes (x, conv weights, folded Mb, conv tile rounded to bf16)."""
    xb = x.bfloat16().float()
    conv = F.conv2d(xb, w3.bfloat16().float(), None, padding=1).bfloat16().float()          # [B,128,h,w]
    Mb = (wout.view(256, 128) @ queries) * 1.4426950408889634                                 # [B,256,128]
    blog = (bout + (wout.view(256, 128) @ queries) @ b3) * 1.4426950408889634                 # [B,256]
    logits = torch.einsum("bkc,bchw->bkhw", Mb.bfloat16().float(), conv) + blog[:, :, None, None]
    p = torch.softmax(logits * math.log(2.0), 1)
    return (p * centres[:, :, None, None]).sum(1, keepdim=True)


@pytest.mark.parametrize("transposed", [1, 0])
@pytest.mark.parametrize("B,h,w,max_ctas", [(1, 16, 8, 0), (2, 48, 40, 0), (3, 37, 29, 2), (2, 64, 96, 5), (1, 240, 24, 3), (2, 176, 40, 0),
                                            (1, 30, 8, 0)])
def test_pixel_fused_bf16(cuda, monkeypatch, B, h, w, max_ctas, transposed):
    """The fused tcgen05 pixel kernel (conv3x3 halo tile + logits GEMM + softmax + expectation), incl. image
    borders (zero padding through TMA OOB fill), ragged tiles and CTAs that loop over many tiles.  Both formulations: the
    transposed one (channels on M, N = 8 x th pixels per MMA: block heights 16 / 24 / 20 / 32 / 30 / 30 / 30 for these shapes) that
    ocv_pixel_fused_bf16 launches by default and the pixels-on-M one (OCV_PIXEL_T=0)."""
    from objcavit_b200 import ops
    monkeypatch.setenv("OCV_PIXEL_T", str(transposed))
    x = torch.randn(B, 128, h, w, generator=g(1))
    w3 = torch.randn(128, 128, 3, 3, generator=g(2)) * 0.03; b3 = torch.randn(128, generator=g(3)) * 0.1
    wout = torch.randn(256, 128, 1, 1, generator=g(4)) * 0.1; bout = torch.randn(256, generator=g(5)) * 0.1
    tok = torch.randn(B, 140, 128, generator=g(6)) * 0.2     # logits of std ~2.5: already sharper than the head's
    wn = torch.rand(B, 256, generator=g(7)) + 0.1
    wn = wn / wn.sum(1, keepdim=True)
    _, centres = oh.bin_edges_and_centres(wn, 0.001, 10.0)
    queries = tok[:, 1:129, :]
    want16 = _pixel_reference(x, w3, b3, wout, bout, queries, centres)
    want32 = (torch.softmax(F.conv2d(torch.einsum("bqc,bchw->bqhw", queries, F.conv2d(x, w3, b3, padding=1)), wout, bout), 1)
              * centres[:, :, None, None]).sum(1, keepdim=True)
    xd = ops.nchw_to_nhwc_bf16(x.to(cuda))
    w3p = ops.pack_conv3x3_bf16(w3.to(cuda))
    Mb, blog = ops.fold_queries_bf16(wout.to(cuda), bout.to(cuda), tok.to(cuda)[:, 1:129, :], b3.to(cuda))
    got = ops.pixel_fused_bf16(xd, w3p, Mb, blog, centres.to(cuda), max_ctas=max_ctas)
    assert got.shape == (B, 1, h, w)
    # same rounded operands: accumulation order / ex2.approx differ, and a 1-ulp flip of a bf16-rounded conv value
    # moves an isolated pixel by a few 1e-3 -> tight on the mean, looser on the max
    err = (got.cpu() - want16).abs()
    assert float(err.mean() / want16.abs().mean()) < 1e-4 and rel_err(got, want16) < 1e-2
    assert rel_err(got, want32) < 2e-2          # north_star bf16 tolerance vs the fp32 chain


def test_gemm_bf16_layernorm_epilogue(cuda):
    """out_proj/linear2 + residual + LayerNorm fused in the tcgen05 GEMM epilogue; fp32 and bf16 outputs; zero rows."""
    from objcavit_b200 import ops
    for (M, K) in ((300, 128), (1111, 1024)):
        a = (torch.randn(M, K, generator=g(M)) * 0.5).bfloat16(); w = (torch.randn(128, K, generator=g(K)) * 0.1).bfloat16()
        b = torch.randn(128, generator=g(3)); r = torch.randn(M, 128, generator=g(4))
        gm = torch.randn(128, generator=g(5)); bt = torch.randn(128, generator=g(6))
        z = (torch.arange(M) % 7 == 0).to(torch.uint8)
        want = F.layer_norm(a.float() @ w.float().T + b + r, (128,), gm, bt, 1e-5).masked_fill(z.bool()[:, None], 0.0)
        o32, o16 = ops.gemm_bf16_ex(a.to(cuda), w.to(cuda), b.to(cuda), residual=r.to(cuda), ln=(gm.to(cuda), bt.to(cuda)),
                                    zero_rows=z.to(cuda), want32=True, want16=True)
        assert rel_err(o32, want) < 2e-5 * max(1.0, math.sqrt(K / 128))
        assert torch.equal(o16.cpu(), o32.cpu().bfloat16())
    # plain epilogue with both outputs + relu
    a = (torch.randn(500, 128, generator=g(7)) * 0.5).bfloat16(); w = (torch.randn(1024, 128, generator=g(8)) * 0.1).bfloat16()
    b = torch.randn(1024, generator=g(9))
    o32, o16 = ops.gemm_bf16_ex(a.to(cuda), w.to(cuda), b.to(cuda), act=ops.ACT_RELU, want32=True, want16=True)
    want = torch.relu(a.float() @ w.float().T + b)
    assert rel_err(o32, want) < 2e-5 and torch.equal(o16.cpu(), o32.cpu().bfloat16())
    _, o16b = ops.gemm_bf16_ex(a.to(cuda), w.to(cuda), b.to(cuda), act=ops.ACT_RELU)
    assert torch.equal(o16b.cpu(), o16.cpu())


@pytest.mark.parametrize("Sq,Sk", [(300, 300), (45, 77), (9, 9), (418, 418), (300, 32)])
def test_mha_core_bf16(cuda, Sq, Sk):
    from objcavit_b200 import ops
    B = 3
    qkv = torch.randn(B, max(Sq, Sk), 384, generator=g(Sq)).bfloat16()
    qf = qkv.float()
    q, k, v = qf[:, :Sq, 0:128], qf[:, :Sk, 128:256], qf[:, :Sk, 256:]
    mask = torch.zeros(B, Sk, dtype=torch.bool)
    mask[0, Sk // 2:] = True
    mask[2, 1:] = True
    sc = (q.reshape(B, Sq, 4, 32).transpose(1, 2) / math.sqrt(32)) @ k.reshape(B, Sk, 4, 32).transpose(1, 2).transpose(-1, -2)
    sc = sc.masked_fill(mask[:, None, None, :], float("-inf"))
    want = (torch.softmax(sc, -1) @ v.reshape(B, Sk, 4, 32).transpose(1, 2)).transpose(1, 2).reshape(B, Sq, 128)
    d = qkv.to(cuda)
    got = ops.mha_core_bf16(d[:, :Sq, 0:128], d[:, :Sk, 128:256], d[:, :Sk, 256:], mask.to(torch.uint8).to(cuda))
    assert rel_err(got.float(), want) < 1.5e-2          # P and the output are rounded to bf16
    assert float((got.float().cpu() - want).abs().mean() / want.abs().mean()) < 4e-3


@pytest.mark.parametrize("B,h,w", [(2, 48, 64), (3, 240, 320), (2, 176, 608), (1, 37, 50)])
def test_patch_embed_bf16(cuda, B, h, w):
    from objcavit_b200 import ops
    x = torch.randn(B, 128, h, w, generator=g(1)).bfloat16()
    wt = (torch.randn(128, 128, 16, 16, generator=g(2)) * 0.01)
    b = torch.randn(128, generator=g(3)); pos = torch.rand(500, 128, generator=g(4))
    S = (h // 16) * (w // 16)
    want = F.conv2d(x.float(), wt.bfloat16().float(), b, stride=16).flatten(2).permute(0, 2, 1) + pos[:S]
    wp = ops.pack_patch_weights_bf16(wt.to(cuda))
    assert torch.equal(wp.cpu(), wt.permute(0, 2, 3, 1).reshape(128, -1).bfloat16())
    xn = x.permute(0, 2, 3, 1).contiguous().to(cuda)
    t32, t16 = ops.patch_embed_bf16(xn, wp, b.to(cuda), pos.to(cuda))
    assert rel_err(t32, want) < 1e-4 and torch.equal(t16.cpu(), t32.cpu().bfloat16())


def test_mlp_chain(cuda):
    from objcavit_b200 import ops
    dims = [4, 32, 64, 128, 256, 128]
    Ws = [torch.randn(dims[i + 1], dims[i], generator=g(i)) / math.sqrt(dims[i]) for i in range(5)]
    bs = [torch.randn(dims[i + 1], generator=g(10 + i)) * 0.1 for i in range(5)]
    x = torch.randn(77, 6, generator=g(20)) * 50
    res = torch.randn(77, 128, generator=g(21))
    want = x[:, :4]
    for i in range(5):
        want = want @ Ws[i].T + bs[i]
        if i < 4:
            want = F.leaky_relu(want, 0.01)
    want = want + res
    o32, o16 = ops.mlp(x.to(cuda)[:, :4], [w.to(cuda) for w in Ws], [b.to(cuda) for b in bs], residual=res.to(cuda),
                       want32=True, want16=True)
    assert rel_err(o32, want) < 1e-5 and torch.equal(o16.cpu(), o32.cpu().bfloat16())


@pytest.mark.parametrize("R", [1, 16, 37])
def test_mlp_rows_regressor(cuda, R):
    """The per-layer few-rows MLP (bin-width regressor, miniViT.py:16-20) on a strided token-0 view."""
    from objcavit_b200 import ops
    dims = [128, 256, 256, 256]
    Ws = [torch.randn(dims[i + 1], dims[i], generator=g(i)) / math.sqrt(dims[i]) for i in range(3)]
    bs = [torch.randn(dims[i + 1], generator=g(10 + i)) * 0.1 for i in range(3)]
    tok = torch.randn(R, 5, 128, generator=g(30))
    want = tok[:, 0, :]
    for i in range(3):
        want = want @ Ws[i].T + bs[i]
        if i < 2:
            want = F.leaky_relu(want, 0.01)
    out = ops.mlp_rows(tok.to(cuda)[:, 0, :], [w.to(cuda) for w in Ws], [b.to(cuda) for b in bs])
    assert rel_err(out, want) < 1e-5
