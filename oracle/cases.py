"""TEST INFRASTRUCTURE ONLY — the seeded parity cases shared by the golden generator and the tests.

Small enough for the CPU oracle to finish in seconds, but valid for the head: S >= 129 tokens so that
128 range-attention queries exist (``modules/miniViT.py:27``).
"""
from __future__ import annotations

from oracle import synth

# name -> args (builtin config), B, h, w (feature-map size), objects, seed
CASES = {
    "adabins_small": dict(args="nyu_adabins", B=1, h=176, w=192, seed=1),
    "ocv_learned_bbox_wh_small": dict(args="nyu_ocv_learned_bbox_wh", B=2, h=176, w=192, n_obj=6, ragged=True,
                                      none_at=1, seed=2),
    "ocv_learned_small": dict(args="nyu_ocv_learned", B=2, h=176, w=192, n_obj=5, ragged=False, seed=3),
    "ocv_2saca_small": dict(args="nyu_ocv_learned_bbox_wh_2saca", B=2, h=176, w=192, n_obj=7, ragged=True, seed=4),
    "ocv_no_obj_sa_small": dict(args="nyu_ocv_learned_bbox_wh_no_obj_sa", B=2, h=176, w=192, n_obj=4, ragged=True,
                                seed=5),
    "ocv_grid_random_small": dict(args="nyu_ocv_grid_random", B=2, h=176, w=192, n_obj=6, ragged=True, seed=6),
    "kitti_roi_align_small": dict(args="kitti_ocv_grid_random_roi_align", B=2, h=176, w=208, n_obj=9, ragged=True,
                                  seed=7),
}

# training-step cases (config 5 semantics: grad mode, dropout disabled, SILog + 0.1 chamfer): gradients w.r.t. every parameter
TRAIN_CASES = {
    "train_learned_bbox_wh_small": dict(args="nyu_ocv_learned_bbox_wh", B=2, h=176, w=192, n_obj=6, ragged=True, none_at=1,
                                        seed=8),
    "train_2saca_small": dict(args="nyu_ocv_learned_bbox_wh_2saca", B=2, h=176, w=192, n_obj=7, ragged=True, seed=9),
    "train_no_obj_sa_small": dict(args="nyu_ocv_learned_bbox_wh_no_obj_sa", B=2, h=176, w=192, n_obj=4, ragged=True, seed=10),
    "train_adabins_small": dict(args="nyu_adabins", B=2, h=176, w=192, seed=11),
}


def grad_probe_indices(numel: int, n: int = 8):
    """Fixed flat positions at which a parameter's gradient is stored in the training fixtures."""
    return sorted({(i * 2654435761 + 12345) % numel for i in range(n)})


def case_gt(case, args):
    ds = args[args.basic.dataset]
    return synth.make_gt(case["B"], case["h"], case["w"], ds.max_depth, ds.min_depth, seed=1234 + case["seed"])


def case_inputs(case):
    feats = synth.make_features(case["B"], case["h"], case["w"], seed=1234 + case["seed"])
    if "n_obj" not in case:
        return feats, None, None
    objs, boxes = synth.make_objects(case["B"], case["h"], case["w"], case["n_obj"], seed=1234 + case["seed"],
                                     ragged=case.get("ragged", False), none_at=case.get("none_at"))
    return feats, objs, boxes


def case_cfg(case, args):
    ds = args[args.basic.dataset]
    cfg = dict(kind=args.model.name, min_depth=ds.min_depth, max_depth=ds.max_depth)
    if args.model.name != "adabins":
        oc = args[args.model.name].objcavit
        cfg.update(strategy=oc.positional_embedding_strategy,
                   no_obj_sa=args.graphbins.objcavit.get("no_obj_sa") is True,
                   use_2_saca=args.graphbins.objcavit.get("use_2_saca") is True)
    return cfg
