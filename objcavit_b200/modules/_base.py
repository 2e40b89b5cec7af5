"""Shared plumbing for the boundary modules: parameter lookup and the inference-only guard."""
from __future__ import annotations

import torch
import torch.nn as nn


class KernelModule(nn.Module):
    """nn.Module whose forward runs on the objcavit_b200 kernels.

    Parameters stay ordinary ``nn.Parameter``s (optimizers, ``state_dict`` and checkpoints see exactly the
    reference layout); ``_params()`` hands the kernels a ``{state_dict key: tensor}`` view of them.
    """

    # Dropout probability of the encoder layers while ``self.training`` (nn.TransformerEncoderLayer's default, which the
    # reference never overrides: modules/layers.py:8, modules/ObjCAViT.py:155-161).  The reference's modules are switched off
    # with ``m.p = 0`` / ``mha.dropout = 0``; here it is this one attribute.  ``eval()`` or 0.0 = no dropout kernels at all.
    dropout = 0.1

    def _wants_grad(self, *tensors):
        return torch.is_grad_enabled() and (any(p.requires_grad for p in self.parameters()) or
                                            any(torch.is_tensor(t) and t.requires_grad for t in tensors))

    def _params(self):
        cache = self.__dict__.get("_ocv_param_cache")
        if cache is None:
            cache = {k: v for k, v in self.named_parameters()}
            self.__dict__["_ocv_param_cache"] = cache
        # detach() (not .data): the view shares the parameter's version counter, so load_state_dict, an optimizer step or any
        # other in-place update shows up in ``_version`` and invalidates the derived bf16 / packed copies (ops.NativeHead.bind,
        # engine_bf16.Packed.signature).  Writes through ``p.data`` bypass torch's version counters by design: callers doing
        # that must call :meth:`invalidate_caches`.
        return {k: v.detach() for k, v in cache.items()}

    def invalidate_caches(self):
        """Drop every derived copy of the parameters (bf16 GEMM weights, packed conv weights, native plan).  Needed only after
        writes torch cannot see (``p.data.copy_``, a kernel writing into parameter storage); everything else is detected."""
        for m in self.modules():
            m.__dict__.pop("_ocv_param_cache", None)
            m.__dict__.pop("_ocv_packed", None)
            m.__dict__.pop("_ocv_native", None)

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
                f"objcavit_b200: {type(self).__name__} on its own runs inference only (wrap the call in torch.no_grad()); training "
                "goes through mViT / ObjCAViT or the head classes, whose forward is an autograd node over the training kernels.")

    @property
    def device(self):  # the reference modules are LightningModules and use self.device
        try:
            return next(self.parameters()).device
        except StopIteration:
            return torch.device("cpu")
