"""Autograd boundary of the kernel modules in grad mode (config 5 semantics).

One ``torch.autograd.Function`` serves both surfaces the reference trains through:

* the head classes (``GraphBins.forward`` / ``AdaBins.forward``, ``modules/GraphBins.py:107-119``): outputs ``(depth_pred, bin_edges)``;
* the inner modules (``self.objcavit(...)`` / ``self.adaptive_bins_layer(...)``, ``modules/GraphBins.py:107``, ``modules/AdaBins.py:76``):
  outputs ``(bin_widths_normed, range_attention_maps)`` — ``cfg["surface"] == "module"``.

forward = ``engine_train.forward`` (activations kept on a tape), backward = ``engine_train.backward`` / ``backward_module`` into ONE
zero-filled flat buffer whose slices are returned as the parameter gradients (autograd accumulates them into ``.grad``, e.g. the
views of a ``FlatGradBuffer``).  Parameters the backward never writes get ``None`` (no gradient in the reference either, F10).
"""
from __future__ import annotations

import torch


def draw_dropout_seed() -> int:
    """Base seed of one training forward's dropout masks, drawn from torch's CPU generator: reproducible under
    ``torch.manual_seed`` like the reference's masks, without a device synchronisation."""
    return int(torch.randint(0, 2 ** 47, (1,)).item())


class TrainFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cfg, x, object_features, object_xywh_list, keys, *params):
        from .. import engine_train
        P = {k: p.detach() for k, p in zip(keys, params)}
        drop = engine_train.DropCtx(cfg.get("dropout", 0.0), cfg.get("dropout_seed", 0))
        a, b, tape = engine_train.forward(P, cfg, x.detach().contiguous().float(), object_features, object_xywh_list, drop=drop)
        ctx.tape, ctx.P, ctx.keys, ctx.surface = tape, P, keys, cfg.get("surface", "head")
        ctx.needs = [p.requires_grad for p in params]
        return a, b

    @staticmethod
    def backward(ctx, d_a, d_b):
        from .. import engine_train
        P, keys = ctx.P, ctx.keys
        sizes = [(P[k].numel() + 63) // 64 * 64 for k in keys]
        flat = torch.zeros(sum(sizes), device=P[keys[0]].device, dtype=torch.float32)
        G, off = engine_train.Touched(), 0
        for k, n in zip(keys, sizes):
            G[k] = flat[off:off + P[k].numel()].view_as(P[k])
            off += n
        if ctx.surface == "module":
            engine_train.backward_module(P, G, ctx.tape, d_a, d_b)
        else:
            if d_a is None:
                d_a = torch.zeros_like(ctx.tape.x[:, :1])
            engine_train.backward(P, G, ctx.tape, d_a.contiguous().float(), d_b)
        ctx.tape = None
        touched = set(G.touched)     # parameters the backward never wrote have no gradient at all (None, as in the reference: F10)
        return (None,) * 5 + tuple(dict.__getitem__(G, k) if (need and k in touched) else None for k, need in zip(keys, ctx.needs))


def train_forward(module, cfg, x, object_features, object_xywh_list):
    """Grad-mode forward of ``module`` (a head class or an inner mViT / ObjCAViT) through :class:`TrainFn`."""
    if torch.is_tensor(x) and x.requires_grad:
        raise NotImplementedError("objcavit_b200: gradients w.r.t. the decoder features are not produced (the extractor is "
                                  "frozen in config 5); pass features.detach()")
    named = list(module.named_parameters())
    keys = tuple(k for k, _ in named)
    if module.training and module.dropout > 0:
        cfg = dict(cfg, dropout=float(module.dropout), dropout_seed=draw_dropout_seed())
    return TrainFn.apply(cfg, x, object_features, object_xywh_list, keys, *[p for _, p in named])
