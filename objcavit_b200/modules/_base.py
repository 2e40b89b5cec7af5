"""Shared plumbing for the boundary modules: parameter lookup and the inference-only guard."""
from __future__ import annotations

import torch
import torch.nn as nn


class KernelModule(nn.Module):
    """nn.Module whose forward runs on the objcavit_b200 kernels.

    Parameters stay ordinary ``nn.Parameter``s (optimizers, ``state_dict`` and checkpoints see exactly the
    reference layout); ``_params()`` hands the kernels a ``{state_dict key: tensor}`` view of them.
    """

    def _params(self):
        cache = self.__dict__.get("_ocv_param_cache")
        if cache is None:
            cache = {k: v for k, v in self.named_parameters()}
            self.__dict__["_ocv_param_cache"] = cache
        return {k: v.data for k, v in cache.items()}

    def _apply(self, fn, *a, **kw):  # .to() / .cuda() / .half(): drop derived caches
        self.__dict__.pop("_ocv_param_cache", None)
        for m in self.modules():
            m.__dict__.pop("_ocv_param_cache", None)
            m.__dict__.pop("_ocv_packed", None)
            m.__dict__.pop("_ocv_native", None)
        return super()._apply(fn, *a, **kw)

    def _guard_inference(self, *tensors):
        if torch.is_grad_enabled() and (any(torch.is_tensor(t) and t.requires_grad for t in tensors)
                                        or any(p.requires_grad for p in self.parameters())):
            raise NotImplementedError(
                "objcavit_b200: the kernel path implements inference (eval + no_grad); the training step "
                "(backward kernels, SURVEY.md config 5) is not built yet — wrap the call in torch.no_grad().")

    @property
    def device(self):  # the reference modules are LightningModules and use self.device
        try:
            return next(self.parameters()).device
        except StopIteration:
            return torch.device("cpu")
