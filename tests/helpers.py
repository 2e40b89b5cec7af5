"""Shared helpers for the parity tests (oracle = checker only)."""
import json
import os

import torch

from objcavit_b200.config import builtin_args
from oracle import synth
from oracle.cases import CASES, case_cfg, case_inputs

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_case(name):
    case = CASES[name]
    args = builtin_args(case["args"])
    with open(os.path.join(GOLD, f"manifest_{case['args']}.json")) as f:
        manifest = json.load(f)
    sd = synth.make_state_dict(manifest, seed=case["seed"])
    feats, objs, boxes = case_inputs(case)
    gold = torch.load(os.path.join(GOLD, f"{name}.pt"))
    return case, args, sd, feats, objs, boxes, gold, case_cfg(case, args)


def rel_err(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    m = torch.isfinite(b)
    assert torch.equal(m, torch.isfinite(a)), "NaN/inf pattern differs"
    if not m.any():
        return 0.0
    return float((a[m] - b[m]).abs().max() / b[m].abs().max().clamp_min(1e-30))
