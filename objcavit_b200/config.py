"""``args`` objects for the head.

The reference builds its modules from an OmegaConf ``DictConfig`` loaded from a params yaml
(``main.py:157-165``) and reads it with attribute access, item access and ``.get``
(``modules/ObjCAViT.py:35-36,159,244,293``; ``modules/GraphBins.py:45,111-112``).  OmegaConf is not a
dependency here: any mapping with those three access styles works — an OmegaConf ``DictConfig`` passed
in by an existing caller is used untouched, and :class:`AttrDict` gives the same behaviour over
``yaml.safe_load``.
"""
from __future__ import annotations

import copy
import os
from typing import Any, Mapping

import yaml


class AttrDict(dict):
    """dict with attribute access, recursive over nested mappings (OmegaConf stand-in)."""

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        for k, v in list(self.items()):
            self[k] = self._wrap(v)

    @classmethod
    def _wrap(cls, v):
        if isinstance(v, Mapping) and not isinstance(v, AttrDict):
            return cls(v)
        if isinstance(v, list):
            return [cls._wrap(x) for x in v]
        return v

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:  # match OmegaConf: missing key -> AttributeError-ish
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = self._wrap(v)

    def __deepcopy__(self, memo):
        return AttrDict(copy.deepcopy(dict(self), memo))


def load_params(path: str) -> AttrDict:
    """Load a reference-format params yaml (``params/*.yaml``) into an ``args`` object."""
    with open(path, "r") as f:
        return AttrDict(yaml.safe_load(f))


_CONFIG_DIR = os.path.join(os.path.dirname(__file__), "configs")

# name -> (file, description); the head-relevant subset of the reference params files named in
# SURVEY.md Appendix B / BASELINE.json configs (keys re-stated, values as in the reference files).
BUILTIN = {
    "nyu_adabins": "nyu_adabins.yaml",
    "nyu_ocv_learned_bbox_wh": "nyu_ocv_learned_bbox_wh.yaml",
    "nyu_ocv_learned": "nyu_ocv_learned.yaml",
    "nyu_ocv_learned_bbox_wh_2saca": "nyu_ocv_learned_bbox_wh_2saca.yaml",
    "nyu_ocv_learned_bbox_wh_no_obj_sa": "nyu_ocv_learned_bbox_wh_no_obj_sa.yaml",
    "nyu_ocv_grid_random": "nyu_ocv_grid_random.yaml",
    "kitti_ocv_grid_random_roi_align": "kitti_ocv_grid_random_roi_align.yaml",
}


def builtin_args(name: str) -> AttrDict:
    return load_params(os.path.join(_CONFIG_DIR, BUILTIN[name]))


def dataset_cfg(args: Any):
    return args[args.basic.dataset]


def model_cfg(args: Any):
    return args[args.model.name]
