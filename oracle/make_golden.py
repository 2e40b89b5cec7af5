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
from oracle.cases import CASES, TRAIN_CASES, case_cfg, case_gt, case_inputs, grad_probe_indices  # noqa: E402
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


def pin_eval():
    """Rows N2-N4: the restatement in ``oracle/eval.py`` against the reference's own code, imported unmodified:
    ``metrics/MetricsPreprocess.py`` (torch only), the six metric classes of ``metrics/*.py`` through a minimal ``torchmetrics``
    stub (``Metric`` = nn.Module whose ``add_state`` sets an attribute, call = ``update``), and
    ``ObjectLanguageStrategy.get_single_relative_size_clause`` through an ``nltk`` stub with a duck-typed ``self``."""
    import types

    import torch.nn as nn

    from oracle import eval as oe
    from oracle.ref_shim import REFERENCE_ROOT

    class Metric(nn.Module):
        def add_state(self, name, default, dist_reduce_fx=None):
            setattr(self, name, default.clone())

        def forward(self, *a, **kw):
            return self.update(*a, **kw)

    tm = types.ModuleType("torchmetrics")
    tm.Metric = Metric
    sys.modules.setdefault("torchmetrics", tm)
    nl, nc = types.ModuleType("nltk"), types.ModuleType("nltk.corpus")
    nc.wordnet = None
    nl.corpus = nc
    sys.modules.setdefault("nltk", nl)
    sys.modules.setdefault("nltk.corpus", nc)
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import_reference()
    from metrics.AbsRel import AbsRel  # type: ignore
    from metrics.AccThresh import AccThresh  # type: ignore
    from metrics.Log10 import Log10  # type: ignore
    from metrics.MetricsPreprocess import MetricsPreprocess  # type: ignore
    from metrics.RMSE import RMSE  # type: ignore
    from metrics.RMSELog import RMSELog  # type: ignore
    from metrics.SqRel import SqRel  # type: ignore
    from modules.ObjectLanguageStrategy import ObjectLanguageStrategy  # type: ignore
    from objcavit_b200.config import AttrDict

    out = {}
    for tag, c in oe.EVAL_CASES.items():
        dataset, lo, hi, garg, eigen = c["dataset"], c["lo"], c["hi"], c["garg"], c["eigen"]
        args = AttrDict({"basic": {"dataset": dataset},
                         dataset: {"min_depth": lo, "max_depth": hi, "garg_crop": garg, "eigen_crop": eigen}})
        a, b, gt = oe.eval_case_inputs(tag)
        H, W = gt.shape[-2:]
        # reference: validation_step lines 159-183 verbatim on the two passes
        p1 = torch.clamp(a, min=lo, max=hi)
        p2 = torch.clamp(b.flip(dims=[3]), min=lo, max=hi)
        final = 0.5 * (p1 + p2)
        pm, mask = MetricsPreprocess(args)(depth_pred=final.clone(), depth_gt=gt.clone())
        pm_m, gt_m = pm[mask], gt[mask]
        ms = [AbsRel(args), SqRel(args), RMSE(args), RMSELog(args), Log10(args), AccThresh(args, 1.25), AccThresh(args, 1.25 ** 2),
              AccThresh(args, 1.25 ** 3)]
        for m in ms:
            m(depth_pred=pm_m.clone(), depth_gt=gt_m.clone())
        ref_vals = torch.tensor([float(m.compute()) for m in ms])
        # restatement
        crop = oe.crop_rect(dataset, garg, eigen, H, W)
        o_final = oe.tta_merge(a, b, lo, hi)
        assert torch.equal(torch.isnan(o_final), torch.isnan(final)) and torch.equal(o_final.nan_to_num(), final.nan_to_num())
        o_pm, o_mask = oe.metrics_preprocess(o_final, gt, lo, hi, crop)
        assert torch.equal(o_mask, mask) and torch.equal(o_pm, pm)
        sums = oe.metric_sums(o_pm[o_mask], gt[o_mask])
        mv = oe.metrics_from_sums(sums)
        got = torch.tensor([mv[k] for k in ("abs_rel", "sq_rel", "rmse", "rmse_log", "log10", "acc_1", "acc_2", "acc_3")])
        err = float(((got - ref_vals).abs() / ref_vals.abs().clamp_min(1e-12)).max())
        print(f"[pin] eval {tag}: metrics vs reference classes rel err {err:.2e}; valid {int(sums[8])} px")
        assert err < 2e-5, (got, ref_vals)
        out[tag] = dict(crop=crop, final=final[:, :, ::3, ::3].contiguous(), upsampled=pm[:, :, ::5, ::5].contiguous(), sums=sums,
                        metrics=ref_vals)
    # object glue
    xyxy = oe.glue_case_inputs()
    o_xywh, o_bucket = oe.object_glue(xyxy)
    fake = types.SimpleNamespace(rel_size_scale=list(oe.REL_SIZE_SCALE), synset_to_name=lambda s: s)
    names = [[f"obj{j}" for j in range(len(x))] for x in o_xywh]
    clauses = ObjectLanguageStrategy.get_single_relative_size_clause(fake, o_xywh, None, None, None, names)
    for cl, bl in zip(clauses, o_bucket):
        for c, bi in zip(cl, bl):
            if bi < 0:
                assert c == ""
            else:
                assert f" appears {oe.REL_SIZE_SCALE[bi]} the " in c, (c, bi)
    print(f"[pin] object glue: {sum(len(b) for b in o_bucket)} relative-size buckets match the reference clauses")
    out["glue"] = dict(xywh=o_xywh, bucket=o_bucket)
    torch.save(out, os.path.join(GOLD, "eval_rows.pt"))


GRAD_FLOOR = 1e-3   # per-parameter gradient error is relative to max(|g|_max, GRAD_FLOOR * largest |g| of the head)


def training_loss(depth, edges, gt, mask):
    """config 5: 1 * SILog + 0.1 * BinsChamfer (``params/*.yaml`` loss.names / loss.coeffs)."""
    return oh.silog_loss(depth, gt, mask) + 0.1 * oh.bins_chamfer_loss(edges, gt, mask)


def run_train_case(name: str):
    """Gradients of the training loss w.r.t. every head parameter from the UNMODIFIED reference modules in train() mode with
    dropout disabled ("parity of grads checked with dropout disabled"), pinned against autograd through
    the restatement (``train=True``: slow-path semantics), stored as a compact fixture: per-parameter L2 norm, sum and 8 probes."""
    case = TRAIN_CASES[name]
    args = builtin_args(case["args"])
    torch.manual_seed(0)
    ref = ReferenceHead(args).train()
    for m in ref.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
        if isinstance(m, torch.nn.MultiheadAttention):
            m.dropout = 0.0
    with open(os.path.join(GOLD, f"manifest_{case['args']}.json")) as f:
        manifest = json.load(f)
    sd = synth.make_state_dict(manifest, seed=case["seed"])
    ref.load_state_dict(sd, strict=True)
    feats, objs, boxes = case_inputs(case)
    gt, mask = case_gt(case, args)
    if objs is None:
        depth, edges, _, _ = ref(feats)
    else:
        depth, edges, _, _ = ref(feats, [o.clone() for o in objs], copy.deepcopy(boxes))
    silog_ref = import_reference().SILogLoss(None)(depth, gt, mask)
    loss = silog_ref + 0.1 * oh.bins_chamfer_loss(edges, gt, mask)          # chamfer: restatement (pytorch3d unavailable)
    loss.backward()
    ref_grads = {k: p.grad for k, p in ref.named_parameters()}
    # restatement under autograd
    sd_g = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    cfg = dict(case_cfg(case, args), train=True)
    o_depth, o_edges, _, _ = oh.head_forward(sd_g, cfg, feats, objs, boxes)
    o_loss = training_loss(o_depth, o_edges, gt, mask)
    o_loss.backward()
    worst = 0.0
    gmax = max(float(g.abs().max()) for g in ref_grads.values() if g is not None)
    fixture = {}
    for k, g in ref_grads.items():
        og = sd_g[k].grad
        if g is None:
            assert og is None or float(og.abs().max()) == 0.0, f"{name}: {k} has no reference gradient but the oracle gives one"
            fixture[k] = None
            continue
        assert og is not None, f"{name}: oracle gives no gradient for {k}"
        # parameters whose true gradient is a rounding-noise cancellation (e.g. the block-2 object path: |g| ~ 1e-9 of the
        # largest gradient) are compared on the global scale
        err = float((og - g).abs().max()) / max(float(g.abs().max()), GRAD_FLOOR * gmax)
        worst = max(worst, err)
        idx = torch.tensor(grad_probe_indices(g.numel()))
        fixture[k] = dict(norm=g.double().norm().float(), sum=g.double().sum().float(), probes=g.flatten()[idx].clone())
    print(f"[pin] {name}: largest |grad| {gmax:.3e}; loss ref {float(loss):.6f} oracle {float(o_loss):.6f}; worst per-parameter grad rel err {worst:.2e}; "
          f"{sum(v is None for v in fixture.values())} parameters without gradient")
    assert abs(float(loss) - float(o_loss)) < 1e-5 * abs(float(loss)) and worst < 2e-3, f"{name}: restatement gradients differ"
    torch.save(dict(loss=loss.detach(), gmax=torch.tensor(gmax), depth=depth.detach()[:, :, ::4, ::4].contiguous(), bin_edges=edges.detach(), grads=fixture),
               os.path.join(GOLD, f"{name}.pt"))


def head_args_subset(args):
    """The keys of a params file the head reads (SURVEY.md §8b "args keys consumed by the head")."""
    model = args["model"]["name"]
    dataset = args["basic"]["dataset"]
    out = {"model": {"name": model}, "basic": {"dataset": dataset}}
    m = {k: args[model][k] for k in ("n_bins", "do_final_upscale") if k in args[model]}
    if "objcavit" in args[model]:
        m["objcavit"] = dict(args[model]["objcavit"])
    out[model] = m
    if model != "graphbins" and "graphbins" in args and "objcavit" in args["graphbins"]:
        out["graphbins"] = {"objcavit": dict(args["graphbins"]["objcavit"])}
    out[dataset] = {k: args[dataset][k] for k in ("min_depth", "max_depth", "dimensions_train", "dimensions_test") if k in args[dataset]}
    return out


def dump_params_sweep():
    """Every ``params/*.yaml`` of the reference ("all must construct through the new modules"): build the
    UNMODIFIED reference head for it, record its ``state_dict`` {key: shape} (deduplicated by content hash) and the head-relevant
    subset of the file.  ``tests/test_modules_cpu.py::test_params_sweep`` builds the product head from that subset (and, in the
    build container, from the real file) and compares the surfaces."""
    import glob
    import hashlib

    import yaml

    from objcavit_b200.config import AttrDict
    from oracle.ref_shim import REFERENCE_ROOT

    files, manifests, skipped = {}, {}, {}
    for path in sorted(glob.glob(os.path.join(REFERENCE_ROOT, "params", "*.yaml"))):
        name = os.path.basename(path)
        try:
            with open(path) as f:
                raw = yaml.safe_load(f)
        except yaml.YAMLError as e:           # one shipped file is not valid YAML
            skipped[name] = "not valid YAML: " + str(e).splitlines()[0]
            continue
        if not isinstance(raw, dict) or "model" not in raw:
            skipped[name] = "no model section (basicParams.yaml: validation-time overrides only)"
            continue
        args = AttrDict(raw)
        torch.manual_seed(0)
        ref = ReferenceHead(args)
        man = {k: list(v.shape) for k, v in ref.state_dict().items()}
        blob = json.dumps(man, sort_keys=True)
        h = hashlib.sha1(blob.encode()).hexdigest()[:12]
        manifests[h] = man
        files[name] = {"args": head_args_subset(raw), "manifest": h, "n_keys": len(man),
                       "n_params": int(sum(p.numel() for p in ref.parameters()))}
    with open(os.path.join(GOLD, "params_sweep.json"), "w") as f:
        json.dump({"files": files, "manifests": manifests, "skipped": skipped}, f, indent=0, sort_keys=True)
    print(f"[pin] params sweep: {len(files)} files -> {len(manifests)} distinct state_dict surfaces; skipped {skipped}")


def main():
    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    pin_grid_ops()
    pin_losses()
    pin_eval()
    for name in CASES:
        run_case(name)
    for name in TRAIN_CASES:
        run_train_case(name)
    dump_params_sweep()
    print("golden fixtures written to", GOLD)


if __name__ == "__main__":
    if "--params-sweep" in sys.argv:       # only the params sweep (leaves the other fixtures untouched)
        dump_params_sweep()
    else:
        main()
