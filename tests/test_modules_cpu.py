"""CPU: the nn.Module boundary — state_dict surface, args handling, loud failure without CUDA."""
import glob
import json
import os

import pytest
import torch

from helpers import GOLD
from objcavit_b200 import _lib
from objcavit_b200.config import AttrDict, builtin_args
from objcavit_b200.modules import (AdaBinsHead, GraphBinsHead, ObjCAViT, PatchTransformerEncoder, build_head, mViT)
from oracle import synth

MANIFESTS = sorted(glob.glob(os.path.join(GOLD, "manifest_*.json")))


@pytest.mark.parametrize("mf", MANIFESTS, ids=[os.path.basename(m)[9:-5] for m in MANIFESTS])
def test_state_dict_surface_matches_reference(mf):
    name = os.path.basename(mf)[len("manifest_"):-5]
    with open(mf) as f:
        manifest = json.load(f)          # dumped from the real reference modules
    head = build_head(builtin_args(name))
    assert {k: list(v.shape) for k, v in head.state_dict().items()} == manifest
    head.load_state_dict(synth.make_state_dict(manifest, 0), strict=True)


def test_key_counts_of_survey_f9():
    assert len(mViT(128, dim_out=256).state_dict()) == 59
    assert len(ObjCAViT(builtin_args("nyu_ocv_learned_bbox_wh")).state_dict()) == 150
    assert len(ObjCAViT(builtin_args("nyu_ocv_learned_bbox_wh_2saca")).state_dict()) == 278
    assert len(ObjCAViT(builtin_args("nyu_ocv_learned_bbox_wh_no_obj_sa")).state_dict()) == 90
    oc = ObjCAViT(builtin_args("kitti_ocv_grid_random_roi_align"))
    assert len(oc.state_dict()) == 141
    assert tuple(oc.positional_encoder.positional_encodings.shape) == (24 * 78, 128)


def test_unknown_strategy_exits_like_reference():
    args = builtin_args("nyu_ocv_learned")
    args.graphbins.objcavit.positional_embedding_strategy = "pixelwise"
    with pytest.raises(SystemExit):
        ObjCAViT(args)


def test_no_cpu_fallback():
    head = AdaBinsHead(builtin_args("nyu_adabins")).eval()
    with torch.no_grad(), pytest.raises(_lib.OcvError):
        head(torch.randn(1, 128, 176, 192))


def test_training_mode_has_no_cpu_path_either():
    """Grad mode routes the head through the training kernels (engine_train): on CPU tensors that must raise, never fall
    back to torch ops; gradients w.r.t. the (frozen-extractor) features are refused explicitly."""
    head = AdaBinsHead(builtin_args("nyu_adabins"))
    with pytest.raises(_lib.OcvError):
        head(torch.randn(1, 128, 176, 192))
    with pytest.raises(NotImplementedError):
        head(torch.randn(1, 128, 176, 192, requires_grad=True))


def test_attrdict_access_styles():
    a = AttrDict({"graphbins": {"objcavit": {"x": 1}}, "model": {"name": "graphbins"}})
    assert a[a.model.name].objcavit.x == 1 and a.graphbins.objcavit.get("no_obj_sa") is None


# ---- every params file of the reference ("all must construct through the new modules") -----------------
with open(os.path.join(GOLD, "params_sweep.json")) as _f:
    SWEEP = json.load(_f)          # oracle/make_golden.py:dump_params_sweep — state_dict surfaces of the UNMODIFIED reference heads


@pytest.mark.parametrize("fname", sorted(SWEEP["files"]), ids=[n[:-5] for n in sorted(SWEEP["files"])])
def test_params_sweep(fname):
    """The product head built from each reference params file has exactly the reference head's ``state_dict`` keys and shapes
    (checkpoint drop-in, F9).  The committed fixture carries the head-relevant keys of each file; in the build container the
    real file is loaded through ``load_params`` as well."""
    entry = SWEEP["files"][fname]
    want = SWEEP["manifests"][entry["manifest"]]
    head = build_head(AttrDict(entry["args"]))
    got = {k: list(v.shape) for k, v in head.state_dict().items()}
    assert got == want
    assert len(got) == entry["n_keys"] and sum(p.numel() for p in head.parameters()) == entry["n_params"]
    real = os.path.join(os.environ.get("OBJCAVIT_REFERENCE_ROOT", "/root/reference"), "params", fname)
    if os.path.isfile(real):
        from objcavit_b200.config import load_params
        head2 = build_head(load_params(real))
        assert {k: list(v.shape) for k, v in head2.state_dict().items()} == want


def test_params_sweep_covers_the_reference_tree():
    assert len(SWEEP["files"]) == 56 and len(SWEEP["skipped"]) == 1      # 57 shipped files, one is not valid YAML (F14)
