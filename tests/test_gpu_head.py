"""GPU: the whole head through the nn.Module boundary against (a) the fixtures produced by the real
reference and (b) the CPU oracle on the same seeded inputs.  fp32 path: depth within 1e-4 relative."""
import copy

import pytest
import torch

from helpers import load_case, rel_err
from objcavit_b200.modules import ObjCAViT, build_head, mViT
from oracle import head as oh
from oracle.cases import CASES

pytestmark = pytest.mark.gpu

FP32_TOL = 1e-4


@pytest.mark.parametrize("name", list(CASES))
def test_head_fp32_matches_reference_fixture(cuda, name):
    case, args, sd, feats, objs, boxes, gold, cfg = load_case(name)
    head = build_head(args, precision="fp32").to(cuda).eval()
    head.load_state_dict(sd, strict=True)
    with torch.no_grad():
        if objs is None:
            out = head(feats.to(cuda))
        else:
            out = head(feats.to(cuda), [o.to(cuda) for o in objs], [None if b is None else b.to(cuda) for b in boxes])
    assert rel_err(out.depth_pred, gold["depth"]) < FP32_TOL
    assert rel_err(out.bin_edges, gold["bin_edges"]) < 1e-5
    assert out.depth_pred.shape == gold["depth"].shape and out[1].shape == gold["bin_edges"].shape


def test_mvit_module_surface(cuda):
    """mViT.forward -> (bin_widths_normed, range_attention_maps), loaded from a reference-format state_dict."""
    case, args, sd, feats, _, _, gold, cfg = load_case("adabins_small")
    m = mViT(128, n_query_channels=128, patch_size=16, dim_out=256, embedding_dim=128, norm="linear").to(cuda).eval()
    m.load_state_dict({k[len("adaptive_bins_layer."):]: v for k, v in sd.items() if k.startswith("adaptive_bins_layer.")})
    with torch.no_grad():
        y, ram = m(feats.to(cuda))
    assert rel_err(y, gold["bin_widths_normed"]) < 1e-5
    assert float((ram[:, ::8, ::8, ::8].cpu() - gold["ram_sample"]).abs().max() / gold["ram_absmax"]) < 1e-5
    with torch.no_grad():
        tgt = m.patch_transformer(feats.to(cuda))                 # [S,N,E] like the reference
    want = oh.patch_transformer_encoder(sd, "adaptive_bins_layer.patch_transformer.", feats)
    assert tgt.shape == (132, 1, 128) and rel_err(tgt.permute(1, 0, 2), want) < 1e-5


def test_objcavit_module_surface_and_list_mutation(cuda):
    case, args, sd, feats, objs, boxes, gold, cfg = load_case("ocv_2saca_small")
    m = ObjCAViT(args).to(cuda).eval()
    m.load_state_dict({k[len("objcavit."):]: v for k, v in sd.items() if k.startswith("objcavit.")})
    obj_list = [o.to(cuda) for o in objs]
    with torch.no_grad():
        y, ram = m(feats.to(cuda), obj_list, [b.to(cuda) for b in boxes])
    assert rel_err(y, gold["bin_widths_normed"]) < 1e-5
    assert float((ram[:, ::8, ::8, ::8].cpu() - gold["ram_sample"]).abs().max() / gold["ram_absmax"]) < 2e-5
    # F11: the caller's list now holds the embedded [N_i,128] object tokens
    assert all(o.shape == (objs[i].shape[0], 128) for i, o in enumerate(obj_list))


def test_larger_batch_against_oracle(cuda):
    """Ragged objects incl. an image without detections, at a size the oracle still finishes in seconds."""
    from objcavit_b200.config import builtin_args
    from oracle import synth
    import json, os
    from helpers import GOLD
    args = builtin_args("nyu_ocv_learned_bbox_wh")
    with open(os.path.join(GOLD, "manifest_nyu_ocv_learned_bbox_wh.json")) as f:
        sd = synth.make_state_dict(json.load(f), seed=11)
    feats = synth.make_features(3, 208, 272, seed=5)
    objs, boxes = synth.make_objects(3, 208, 272, 32, seed=5, ragged=True, none_at=2)
    cfg = dict(kind="graphbins", strategy="learned_bbox_wh", min_depth=0.001, max_depth=10)
    want_depth, want_edges, _, _ = oh.head_forward(sd, cfg, feats, objs, boxes)
    head = build_head(args).to(cuda).eval()
    head.load_state_dict(sd)
    with torch.no_grad():
        out = head(feats.to(cuda), [o.to(cuda) for o in objs], [None if b is None else b.to(cuda) for b in boxes])
    assert rel_err(out.depth_pred, want_depth) < FP32_TOL and rel_err(out.bin_edges, want_edges) < 1e-5


BF16_TOL = 2e-2   # north_star: bf16 depth within 2e-2 relative of the reference fp32 head


@pytest.mark.parametrize("name", list(CASES))
def test_head_bf16_matches_reference_fixture(cuda, name):
    """Throughput path (tcgen05 pixel kernel + bf16 token path) against the fixtures of the real reference."""
    case, args, sd, feats, objs, boxes, gold, cfg = load_case(name)
    head = build_head(args, precision="bf16").to(cuda).eval()
    head.load_state_dict(sd, strict=True)

    def run(x):
        with torch.no_grad():
            if objs is None:
                return head(x)
            return head(x, [o.to(cuda) for o in objs], [None if b is None else b.to(cuda) for b in boxes])
    out = run(feats.to(cuda))
    assert out.depth_pred.shape == gold["depth"].shape and out[1].shape == gold["bin_edges"].shape
    assert rel_err(out.depth_pred, gold["depth"]) < BF16_TOL
    assert rel_err(out.bin_edges, gold["bin_edges"]) < BF16_TOL
    # channels_last bf16 features are consumed in place and give the same bits as the layout kernel's copy
    x_cl = feats.to(cuda).bfloat16().contiguous(memory_format=torch.channels_last)
    out2 = run(x_cl)
    assert torch.equal(out2.depth_pred, out.depth_pred)
