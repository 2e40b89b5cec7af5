"""Head classes: the part of ``AdaBins`` / ``GraphBins`` between the decoder's feature map and
``depth_pred`` (``modules/AdaBins.py:42-54,76-89``; ``modules/GraphBins.py:45-62,105-121``).

The encoder/decoder, YOLOv7 and CLIP stay outside (SURVEY.md §2 "out of scope"): these classes take
the dense features (and the object embeddings / boxes) the reference models compute just before the
head, expose the same attribute names (``adaptive_bins_layer`` / ``objcavit``, ``conv_out``) so
checkpoints map with a prefix strip, and return the reference's ``ReturnType`` namedtuples.
"""
from __future__ import annotations

from collections import namedtuple

import torch
import torch.nn as nn

from .. import engine
from ._base import KernelModule
from .miniViT import mViT
from .ObjCAViT import ObjCAViT


from ._autograd import draw_dropout_seed, train_forward  # noqa: E402,F401  (draw_dropout_seed is re-exported for trainer.py)


class _HeadBase(KernelModule):
    precision = "fp32"   # "fp32": parity path (SIMT fp32).  "bf16": tensor-core path (tcgen05).

    def enable_pipeline(self, lanes: int = 2, pixel_ctas: int = 0):
        """Pipelined serving mode of the bf16 path (``ocv_head_set_pipeline``): consecutive ``forward`` calls overlap — the
        latency-bound token path of call i+1 runs next to the tensor-bound fused pixel kernel of call i.  Calls return
        immediately; outputs are valid on the current stream only after :meth:`join`, and the inputs of a call must stay
        untouched until then.  ``lanes=0`` switches back."""
        assert self.precision == "bf16", "the pipelined mode drives the bf16 throughput path"
        self.__dict__["_ocv_pipeline"] = (int(lanes), int(pixel_ctas)) if lanes else None
        nh = self.__dict__.get("_ocv_native")
        if nh is not None:
            nh.set_pipeline(int(lanes), int(pixel_ctas))

    def join(self):
        nh = self.__dict__.get("_ocv_native")
        if nh is not None:
            nh.join()

    def _train_forward(self, cfg, x, object_features, object_xywh_list):
        """Grad-mode forward (the reference's training step): explicit forward + backward kernels.  ``precision="fp32"``: every
        kernel fp32 SIMT (the parity path); ``"bf16"``: the pixel path (conv3x3, logits and their gradients: 90 % of the FLOPs)
        and the encoder GEMMs on tcgen05 with bf16 operands."""
        if self.precision == "bf16":
            cfg = dict(cfg, pixel="bf16")
        return train_forward(self, cfg, x, object_features, object_xywh_list)

    def _depth_range(self):
        ds = self.args[self.args.basic.dataset]
        return float(ds.min_depth), float(ds.max_depth)

    def get_non_encoder_params(self):  # AdaBins.py:63-65 — the head is all "non encoder"
        yield from self.parameters()

    def get_encoder_params(self):
        return iter(())

    def get_frozen_params(self):
        return iter(())


class AdaBinsHead(_HeadBase):
    def __init__(self, args, precision: str = "fp32"):
        super().__init__()
        self.args = args
        self.precision = precision
        self.n_bins = self.args.adabins.n_bins
        self.ReturnType = namedtuple('ReturnType', ['depth_pred', 'bin_edges'])
        max_seq_len = 1200 if self.args[self.args.model.name].get('do_final_upscale') else 500
        self.adaptive_bins_layer = mViT(128, n_query_channels=128, patch_size=16, dim_out=self.n_bins,
                                        embedding_dim=128, norm='linear', max_seq_len=max_seq_len)
        self.conv_out = nn.Sequential(nn.Conv2d(128, self.n_bins, kernel_size=1, stride=1, padding=0),
                                      nn.Softmax(dim=1))

    def forward(self, unet_out):
        """``unet_out``: decoder features ``[B,128,h,w]`` -> ``ReturnType(depth_pred [B,1,h,w], bin_edges [B,K+1])``."""
        lo, hi = self._depth_range()
        if self._wants_grad(unet_out):
            depth, edges = self._train_forward(dict(kind="adabins", min_depth=lo, max_depth=hi), unet_out, None, None)
            return self.ReturnType(depth_pred=depth, bin_edges=edges)
        P, fp = self._param_state()
        if self.precision == "bf16":
            from .. import engine_bf16
            depth, edges = engine_bf16.head_forward(self, P, "adaptive_bins_layer.", unet_out, None, None, lo, hi, fingerprint=fp)
        else:
            x = unet_out.contiguous().float()
            widths, ram = engine.mvit_forward(P, "adaptive_bins_layer.", x, 128, 'linear')
            depth, edges = engine.head_tail(P, "conv_out.", ram, widths, lo, hi)
        return self.ReturnType(depth_pred=depth, bin_edges=edges)


class GraphBinsHead(_HeadBase):
    def __init__(self, args, precision: str = "fp32", obj_feature_dim: int = 512):
        super().__init__()
        self.args = args
        self.precision = precision
        self.ReturnType = namedtuple('ReturnType', ['depth_pred', 'bin_edges', 'detections'])
        emb = self.args[self.args.model.name].objcavit.embedding_dim
        max_seq_len = 1200 if self.args[self.args.model.name].get('do_final_upscale') else 500
        self.objcavit = ObjCAViT(self.args, n_query_channels=128, patch_size=16, im_feature_dim=128,
                                 obj_feature_dim=obj_feature_dim, embedding_dim=emb,
                                 dim_out=self.args.graphbins.n_bins, norm='linear', max_seq_len=max_seq_len)
        self.conv_out = nn.Sequential(nn.Conv2d(emb, self.args.graphbins.n_bins, kernel_size=1, stride=1, padding=0),
                                      nn.Softmax(dim=1))

    def forward(self, dense_features, object_features, object_xywh_list, detections=None):
        """-> ``ReturnType(depth_pred, bin_edges, detections)`` (``GraphBins.py:105-121``)."""
        lo, hi = self._depth_range()
        oc = self.objcavit
        if self._wants_grad(dense_features):
            cfg = dict(kind="graphbins", strategy=oc.strategy, no_obj_sa=oc.no_obj_sa, use_2_saca=oc.use_2_saca, min_depth=lo,
                       max_depth=hi)
            depth, edges = self._train_forward(cfg, dense_features, object_features, object_xywh_list)
            return self.ReturnType(depth_pred=depth, bin_edges=edges, detections=detections)
        P, fp = self._param_state()
        if self.precision == "bf16":
            from .. import engine_bf16
            depth, edges = engine_bf16.head_forward(self, P, "objcavit.", dense_features, object_features,
                                                    object_xywh_list, lo, hi, fingerprint=fp)
        else:
            x = dense_features.contiguous().float()
            widths, ram = engine.objcavit_forward(P, "objcavit.", x, object_features, object_xywh_list, oc.strategy,
                                                  oc.no_obj_sa, oc.use_2_saca, 128, 'linear')
            depth, edges = engine.head_tail(P, "conv_out.", ram, widths, lo, hi)
        return self.ReturnType(depth_pred=depth, bin_edges=edges, detections=detections)


class HostPipeline:
    """Chunked host -> device -> host streaming of one batch through a bf16 head (the e2e path): the batch is cut into
    chunks; chunk c+1's features cross PCIe on a copy stream while chunk c runs, and chunk c's depth map goes back on a
    second copy stream.  Images are independent, so chunking changes nothing numerically (up to the patch-embedding
    split-K plan, see tests/test_gpu_fullsize.py).  Host tensors must be pinned for the copies to be asynchronous."""

    def __init__(self, head, chunks: int = 4):
        assert head.precision == "bf16", "the streaming pipeline drives the bf16 throughput path"
        self.head, self.chunks = head, chunks
        self.h2d, self.d2h = torch.cuda.Stream(), torch.cuda.Stream()
        self.xbuf = [None, None]
        self.done = [None, None]

    @torch.no_grad()
    def __call__(self, feats_host, object_features=None, object_xywh_list=None, depth_out=None, edges_out=None,
                 sync: bool = True):
        """``feats_host`` (pinned host memory), either layout:

        * ``[B,h,w,128]`` bf16 NHWC — the decoder hand-off format (``modules/decoder_handoff.py``): copied and consumed in place;
        * ``[B,128,h,w]`` fp32 NCHW — what the reference's caller holds (``modules/GraphBins.py:107``): each chunk crosses PCIe
          as fp32 (twice the bytes) and is converted on the device by ``ocv_nchw_to_nhwc_bf16`` before the head runs.

        Returns (depth [B,1,h,w] fp32, bin_edges [B,K+1]) in pinned host memory.  The results live in HOST memory, so the call
        waits (``sync=True``, the default) until the last device->host copy has landed before it returns; with ``sync=False``
        it returns at once and the caller must ``pipe.done_event.synchronize()`` before reading the outputs on the CPU."""
        from .. import ops
        head = self.head
        dev = head.device
        nchw32 = feats_host.dtype == torch.float32
        if nchw32:
            B, C, h, w = feats_host.shape
            chunk_shape = lambda n: (n, C, h, w)
        else:
            assert feats_host.dtype == torch.bfloat16, "host features: bf16 NHWC or fp32 NCHW"
            B, h, w, C = feats_host.shape
            chunk_shape = lambda n: (n, h, w, C)
        n = min(self.chunks, B)
        bounds = [(B * i // n, B * (i + 1) // n) for i in range(n)]
        if depth_out is None:
            depth_out = torch.empty(B, 1, h, w, dtype=torch.float32).pin_memory()
        if edges_out is None:
            edges_out = torch.empty(B, head.conv_out[0].out_channels + 1, dtype=torch.float32).pin_memory()
        cur = torch.cuda.current_stream()
        self.h2d.wait_stream(cur)
        results = []
        for i, (lo, hi) in enumerate(bounds):
            slot = i & 1
            with torch.cuda.stream(self.h2d):
                if self.done[slot] is not None:
                    self.h2d.wait_event(self.done[slot])      # the chunk that used this buffer two steps ago has been consumed
                buf = self.xbuf[slot]
                if buf is None or buf.shape[0] < hi - lo or buf.shape[1:] != chunk_shape(1)[1:] or buf.dtype != feats_host.dtype:
                    if buf is not None:
                        buf.record_stream(cur)                # the compute stream may still read the buffer being replaced
                    self.xbuf[slot] = torch.empty(chunk_shape(max(hi - lo, (B + n - 1) // n)), device=dev, dtype=feats_host.dtype)
                x = self.xbuf[slot][:hi - lo]
                x.copy_(feats_host[lo:hi], non_blocking=True)
                objs = boxes = None
                if object_features is not None:
                    objs = [o.to(dev, non_blocking=True) for o in object_features[lo:hi]]
                    boxes = [None if b is None else b.to(dev, non_blocking=True) for b in object_xywh_list[lo:hi]]
                    for t in objs + [b for b in boxes if b is not None]:
                        t.record_stream(cur)                  # allocated on the copy stream, consumed on the compute stream
                ready = torch.cuda.Event()
                ready.record(self.h2d)
            cur.wait_event(ready)
            if nchw32:
                xv = ops.nchw_to_nhwc_bf16(x).permute(0, 3, 1, 2)   # layout + precision change on the device, then as below
            else:
                xv = x.permute(0, 3, 1, 2)                     # logical NCHW view of NHWC memory: consumed in place
            out = head(xv) if objs is None else head(xv, objs, boxes)
            ev = torch.cuda.Event()
            ev.record(cur)
            self.done[slot] = ev
            with torch.cuda.stream(self.d2h):
                self.d2h.wait_event(ev)
                depth_out[lo:hi].copy_(out.depth_pred, non_blocking=True)
                edges_out[lo:hi].copy_(out.bin_edges, non_blocking=True)
            results.append(out)                                # keep device tensors alive until the copies are done
        cur.wait_stream(self.d2h)
        self._keep = results
        self.done_event = torch.cuda.Event()
        self.done_event.record(self.d2h)                       # after the last device->host copy
        if sync:
            self.done_event.synchronize()                      # stream order is not host order: the outputs are host memory
        return depth_out, edges_out


def build_head(args, precision: str = "fp32"):
    """Head for a params file: ``model.name == 'adabins'`` -> :class:`AdaBinsHead`, else :class:`GraphBinsHead`."""
    return AdaBinsHead(args, precision) if args.model.name == "adabins" else GraphBinsHead(args, precision)
