"""TEST INFRASTRUCTURE ONLY — pin the oracle against the REAL reference and write ``tests/golden``.

Run in the build container (needs ``/root/reference``):  ``python -m oracle.make_golden``

For every case it (1) constructs the unmodified reference module(s) through ``oracle.ref_shim``,
(2) dumps the ``state_dict`` manifest (pins SURVEY.md F9), (3) loads seeded weights with
``load_state_dict(strict=True)``, (4) runs the reference under ``eval()+no_grad`` on the seeded inputs
of ``oracle.synth``, (5) checks the restatement in ``oracle.head`` against it (hard assert) and
(6) stores the reference outputs as fixtures.  The GPU box regenerates weights/inputs from the
manifest + seeds and compares against these fixtures without ever touching the reference tree.
"""
from __future__ import annotations

import copy
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from objcavit_b200.config import builtin_args  # noqa: E402  (args loader only; no kernels involved)
from oracle import head as oh  # noqa: E402
from oracle import synth  # noqa: E402
from oracle.cases import CASES, case_cfg, case_inputs  # noqa: E402
from oracle.ref_shim import ReferenceHead, import_reference  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def rel_err(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def run_case(name: str):
    case = CASES[name]
    args = builtin_args(case["args"])
    torch.manual_seed(0)
    ref = ReferenceHead(args).eval()
    manifest = {k: list(v.shape) for k, v in ref.state_dict().items()}
    with open(os.path.join(GOLD, f"manifest_{case['args']}.json"), "w") as f:
        json.dump(manifest, f, indent=0, sort_keys=True)
    sd = synth.make_state_dict(manifest, seed=case["seed"])
    ref.load_state_dict(sd, strict=True)
    feats, objs, boxes = case_inputs(case)
    with torch.no_grad():
        if objs is None:
            depth, edges, widths, ram = ref(feats)
        else:
            depth, edges, widths, ram = ref(feats, [o.clone() for o in objs], copy.deepcopy(boxes))
    # pin the restatement
    o_depth, o_edges, o_widths, o_ram = oh.head_forward(sd, case_cfg(case, args), feats, objs, boxes)
    finite = torch.isfinite(depth)
    assert torch.equal(finite, torch.isfinite(o_depth)), f"{name}: NaN pattern differs"
    e = dict(depth=rel_err(o_depth[finite], depth[finite]) if finite.any() else 0.0,
             edges=rel_err(o_edges[torch.isfinite(edges)], edges[torch.isfinite(edges)]) if torch.isfinite(edges).any() else 0.0,
             ram=rel_err(o_ram[torch.isfinite(ram)], ram[torch.isfinite(ram)]) if torch.isfinite(ram).any() else 0.0)
    print(f"[pin] {name}: oracle-vs-reference rel err {e}")
    assert e["depth"] < 2e-5 and e["edges"] < 2e-5 and e["ram"] < 5e-5, f"{name}: restatement does not match reference"
    torch.save(dict(depth=depth.contiguous(), bin_edges=edges.contiguous(), bin_widths_normed=widths.contiguous(),
                    ram_sample=ram[:, ::8, ::8, ::8].contiguous(),
                    ram_absmax=ram[torch.isfinite(ram)].abs().max() if torch.isfinite(ram).any() else torch.tensor(0.)),
               os.path.join(GOLD, f"{name}.pt"))


def pin_grid_ops():
    """ps_roi_align / grid_sample restatements vs the installed torchvision / torch ops."""
    import torch.nn.functional as F
    import torchvision

    g = torch.Generator().manual_seed(5)
    table = torch.rand(24 * 78, 128, generator=g)
    out = {}
    for (h, w) in ((176, 608), (240, 320)):
        grid, gh, gw = oh._grid_for(table, h, w)
        gridNCHW = grid.permute(2, 0, 1).unsqueeze(0).contiguous()
        _, boxes = synth.make_objects(1, h, w, 48, seed=11 + h)
        xywh = boxes[0]
        xywh = torch.cat([xywh, torch.tensor([[-1., -1., -1., -1.], [3., 2., 0., 0.], [2. * w - 1, 2. * h - 1, 9., 9.]])])
        xyxy = oh._xywh_to_clamped_xyxy(xywh)
        for scale, tag in ((1 / 32, "obj"), (1 / 16, "img")):
            want = torchvision.ops.ps_roi_align(gridNCHW, [xyxy], output_size=[1, 1], spatial_scale=scale).squeeze(-1).squeeze(-1)
            got, _ = oh.ps_roi_align_1x1(grid, xyxy, scale)
            assert torch.equal(torch.isnan(want), torch.isnan(got))
            m = ~torch.isnan(want)
            err = float((want[m] - got[m]).abs().max())
            print(f"[pin] ps_roi_align {tag} {h}x{w}: max abs err {err:.3e}")
            assert err < 1e-6
            out[f"roi_{tag}_{h}x{w}"] = dict(xyxy=xyxy, want=want)
        # grid_sample
        pts = (torch.rand(64, 2, generator=g) * 2.6 - 1.3)
        want = F.grid_sample(gridNCHW, pts.view(1, 1, -1, 2), align_corners=False).squeeze(2).squeeze(0).T
        got, _, _ = oh.grid_sample_points(grid, pts[:, 0], pts[:, 1])
        err = float((want - got).abs().max())
        print(f"[pin] grid_sample {h}x{w}: max abs err {err:.3e}")
        assert err < 5e-6  # a few ulp: ATen's vectorised kernel orders the 4-corner sum differently
        out[f"gs_{h}x{w}"] = dict(pts=pts, want=want)
    torch.save(dict(table=table, cases=out), os.path.join(GOLD, "grid_ops.pt"))


def pin_losses():
    ref = import_reference()
    g = torch.Generator().manual_seed(3)
    depth = 0.5 + 9 * torch.rand(2, 1, 24, 32, generator=g)
    gt, mask = synth.make_gt(2, 24, 32, 10.0)
    want = ref.SILogLoss(None)(depth, gt, mask)
    got = oh.silog_loss(depth, gt, mask)
    assert abs(float(want) - float(got)) < 1e-6 * abs(float(want))
    widths = torch.rand(2, 256, generator=g) + 0.1
    widths = widths / widths.sum(1, keepdim=True)
    edges, _ = oh.bin_edges_and_centres(widths, 0.001, 10.0)
    a, b = oh.bins_chamfer_loss(edges, gt, mask), oh.bins_chamfer_loss_bruteforce(edges, gt, mask)
    assert abs(float(a) - float(b)) < 1e-5 * abs(float(b)), (a, b)
    print(f"[pin] silog {float(want):.6f} == {float(got):.6f}; chamfer(search) {float(a):.6f} == brute {float(b):.6f} (pytorch3d unpinned)")
    torch.save(dict(depth=depth, silog=want, edges=edges, chamfer=b), os.path.join(GOLD, "losses.pt"))


def main():
    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    pin_grid_ops()
    pin_losses()
    for name in CASES:
        run_case(name)
    print("golden fixtures written to", GOLD)


if __name__ == "__main__":
    main()
