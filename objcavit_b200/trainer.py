"""The data-parallel training step of the head (BASELINE.json config 5) without torch autograd or torch
optimizers on the path:

    zero flat grads -> engine_train.forward -> SILog / chamfer kernels (loss + d depth, d edges) -> engine_train.backward
    (accumulates straight into the flat gradient buffer) -> ONE all-reduce of that buffer (NCCL over NVLink; skipped at world 1)
    -> deterministic sum-of-squares -> fused clip + AdamW over the flat parameter buffer (2 launches for the whole head).

replaces: ``GraphBinsLM.training_step`` (``modules/GraphBinsLM.py:121-139``: model forward + LossWrapper), Lightning's backward,
DDP's bucketed gradient all-reduce with unused-parameter search (``main.py:66``), ``gradient_clip_val`` (``main.py:35-38``),
``torch.optim.AdamW`` + ``OneCycleLR(cycle_momentum)`` (``modules/GraphBinsLM.py:455-480``).
"""
from __future__ import annotations

import math
from typing import List, Optional

import torch
import torch.distributed as dist

from . import engine_train, ops
from .grad_sync import FlatGradBuffer
from .modules.heads import draw_dropout_seed


class OneCycle:
    """``torch.optim.lr_scheduler.OneCycleLR`` (cos annealing, two phases, ``cycle_momentum=True``) as a pure host function of
    the step index: returns (lr, beta1) for optimizer step ``i`` (0-based)."""

    def __init__(self, max_lr: float, total_steps: int, div_factor: float = 25.0, final_div_factor: float = 100.0,
                 pct_start: float = 0.3, base_momentum: float = 0.85, max_momentum: float = 0.95):
        self.total = total_steps
        self.initial_lr = max_lr / div_factor
        self.max_lr = max_lr
        self.min_lr = self.initial_lr / final_div_factor
        self.base_m, self.max_m = base_momentum, max_momentum
        self.end1 = float(pct_start * total_steps) - 1.0
        self.end2 = float(total_steps) - 1.0

    @staticmethod
    def _cos(start, end, pct):
        return end + (start - end) / 2.0 * (math.cos(math.pi * pct) + 1.0)

    def at(self, i: int):
        i = min(i, self.total - 1)
        if i <= self.end1 or self.end1 <= 0:
            pct = i / self.end1 if self.end1 > 0 else 1.0
            return self._cos(self.initial_lr, self.max_lr, pct), self._cos(self.max_m, self.base_m, pct)
        pct = (i - self.end1) / (self.end2 - self.end1)
        return self._cos(self.max_lr, self.min_lr, pct), self._cos(self.base_m, self.max_m, pct)


class HeadTrainer:
    """Owns the flat parameter / gradient / AdamW-moment buffers of one head replica and runs whole training steps.

    ``MAX_GRAPHS``: captured step graphs kept alive (LRU over input signatures); the seen-signature counter is bounded too.

    The head's ``nn.Parameter``s are re-pointed at slices of ONE flat fp32 buffer (their values are preserved; ``state_dict``
    and checkpoints are unchanged), so the optimizer step is a single kernel over [param | grad | exp_avg | exp_avg_sq].
    """

    MAX_GRAPHS = 8

    def __init__(self, head, lr: float = 0.000357, weight_decay: float = 0.1, betas=(0.9, 0.999), eps: float = 1e-8,
                 clip_grad_norm: float = 0.1, loss_coeffs=(1.0, 0.1), schedule: Optional[OneCycle] = None, group=None,
                 graph: bool = False):
        self.head = head
        self.graph = False
        self.allreduce_events = None     # a list while the caller wants the exposed all-reduce time of every step (bench.py)
        self.flat16, self._P16 = None, None
        self._graphs, self._graph_seen, self._graph_pool, self._salt = {}, {}, None, None
        self._capture_stream = None
        self.group = group
        self.lr, self.wd, self.betas, self.eps, self.clip = lr, weight_decay, betas, eps, clip_grad_norm
        self.coeffs = loss_coeffs
        self.schedule = schedule
        self.grads = FlatGradBuffer(head.parameters())
        flat = torch.empty_like(self.grads.flat)
        flat.zero_()
        for p, off in zip(self.grads.params, self.grads.offsets):      # parameters become views of one buffer
            view = flat[off:off + p.numel()].view_as(p)
            view.copy_(p.data)
            p.data = view
        self.flat_params = flat
        self.exp_avg = torch.zeros_like(flat)
        self.exp_avg_sq = torch.zeros_like(flat)
        self.sumsq = torch.zeros(1, device=flat.device, dtype=torch.float32)
        self.step_index = 0
        self.ranges = None
        self.keys = [k for k, p in head.named_parameters() if p.requires_grad]
        head.invalidate_caches()
        oc = getattr(head, "objcavit", None)
        lo, hi = head._depth_range()
        if oc is None:
            self.cfg = dict(kind="adabins", min_depth=lo, max_depth=hi)
        else:
            self.cfg = dict(kind="graphbins", strategy=oc.strategy, no_obj_sa=oc.no_obj_sa, use_2_saca=oc.use_2_saca,
                            min_depth=lo, max_depth=hi)
        if getattr(head, "precision", "fp32") == "bf16":     # pixel path of the step on tensor cores (engine_train._pixel_*_bf16)
            self.cfg["pixel"] = "bf16"
        if graph:
            self.enable_graph(True)

    def world(self) -> int:
        return dist.get_world_size(self.group) if dist.is_available() and dist.is_initialized() else 1

    def _maps(self):
        named = self.__dict__.get("_named")
        if named is None:                                   # the walk over the module tree is host time on every step otherwise
            named = self._named = dict(self.head.named_parameters())
            self._modules = list(self.head.modules())
        P = {k: named[k].data for k in self.keys}
        G = engine_train.Touched({k: named[k].grad for k in self.keys})
        if self.cfg.get("pixel") == "bf16" and engine_train.TOKEN_BF16:
            # bf16 operands of the encoder GEMMs: views (same offsets) of ONE bf16 mirror of the flat parameter buffer, refreshed by a
            # single cast kernel at the start of every step (forward_backward)
            if self.flat16 is None:
                self.flat16 = torch.empty(self.flat_params.numel(), device=self.flat_params.device, dtype=torch.bfloat16)
                offs = {id(p): off for p, off in zip(self.grads.params, self.grads.offsets)}
                self._P16 = {k: self.flat16[offs[id(named[k])]:offs[id(named[k])] + named[k].numel()].view_as(named[k])
                             for k in self.keys if named[k].dim() == 2 and id(named[k]) in offs}
            return P, G, self._P16
        return P, G, None

    def forward_backward(self, features, object_features, object_xywh_list, depth_gt, depth_mask, xh=None):
        """Local forward + loss + backward; gradients accumulate into the flat buffer.  Returns the loss (device scalar).
        ``xh``: the features already in NHWC bf16 (graph mode converts the caller's tensor straight into its static buffer);
        ``features`` is then a shape placeholder."""
        P, G, P16 = self._maps()
        if P16 is not None:
            ops.cast_bf16(self.flat_params, out=self.flat16)
        x = features if xh is not None else features.contiguous().float()
        p_drop = float(self.head.dropout) if self.head.training else 0.0      # encoder-layer dropout, as the reference in train()
        drop = engine_train.DropCtx(p_drop, draw_dropout_seed() if p_drop > 0 else 0)
        gt = depth_gt.contiguous().float()                # before the forward: the chamfer branch below forks inside it
        depth, edges, tape = engine_train.forward(P, self.cfg, x, object_features, object_xywh_list, drop=drop, P16=P16, xh=xh)
        # the chamfer loss needs only the bin edges (ready before the pixel path of the forward) and the ground truth: it runs on a
        # branch stream from that point, next to the pixel forward; SILog (needs the depth map) stays on the launching stream
        dev, fork = depth.device, getattr(tape, "edges_ready", None)
        with engine_train.on_branch(dev, fork, edges, gt, depth_mask, name="loss"):
            l2, d_edges = ops.bins_chamfer(edges, gt, depth_mask, want_grad=True, grad_scale=self.coeffs[1])
        l1, d_depth = ops.silog(depth, gt, depth_mask, want_grad=True, grad_scale=self.coeffs[0])
        engine_train.join_branch(dev, fork, l2, d_edges, name="loss")
        engine_train.backward(P, G, tape, d_depth, d_edges)
        if self.ranges is None:                      # static per configuration: merge the touched parameters into flat ranges
            named = self._named
            spans = sorted((off, off + (p.numel() + 63) // 64 * 64) for (p, off) in zip(self.grads.params, self.grads.offsets)
                           if any(named[k] is p for k in G.touched))
            merged = []
            for lo, hi in spans:
                if merged and merged[-1][1] == lo:
                    merged[-1][1] = hi
                else:
                    merged.append([lo, hi])
            self.ranges = [(lo, min(hi, self.flat_params.numel())) for lo, hi in merged]
        self.last_outputs = (depth, edges)
        return l1, l2

    # ---- CUDA-graph replay of zero-grads + forward + losses + backward ------------------------------------------------------
    # The step is ~560 short launches sequenced from the host (4.7 ms of host time against ~5.4 ms on the GPU): with
    # ``graph=True`` the trainer captures that part once per input signature (shapes, object counts, train/eval, dropout p) after
    # two eager steps and replays it with ONE launch; the all-reduce, the clip and AdamW stay eager (their arguments — learning
    # rate, beta1, step index — change every step).  Inputs are copied into static buffers; the dropout site seeds are frozen
    # kernel arguments, so every replay first advances a 64-bit salt in device memory that all dropout kernels fold into their
    # seed (``ocv_set_dropout_salt``) — masks differ from step to step and stay reproducible under ``torch.manual_seed``.
    # Same kernels, same order, same streams as the eager step: with dropout off the results are bit-identical.
    def enable_graph(self, on: bool = True) -> None:
        from . import _lib
        self.graph = bool(on)
        if on and self._salt is None:
            self._salt = torch.zeros(1, device=self.flat_params.device, dtype=torch.int64) + int(torch.randint(0, 2 ** 62, (1,)).item())
            self._graph_pool = torch.cuda.graph_pool_handle()
        if on:
            self._arm_salt()
        else:
            if self._salt is not None:      # only if the process-wide pointer is still ours (another live trainer may own it)
                _lib.check(_lib.lib().ocv_release_dropout_salt(self._salt.data_ptr()), "ocv_release_dropout_salt")
            self._graphs.clear()

    def _arm_salt(self) -> None:
        """Install THIS trainer's salt word as the process-wide dropout salt (a host pointer write).  Done before every step of
        a graph-mode trainer, so a second trainer in the process (or one that was deleted) cannot leave a capture with a NULL or
        foreign salt — which would freeze the dropout masks of every replay."""
        from . import _lib
        _lib.check(_lib.lib().ocv_set_dropout_salt(self._salt.data_ptr()), "ocv_set_dropout_salt")

    def __del__(self):
        try:
            if self._salt is not None:
                from . import _lib
                _lib.lib().ocv_release_dropout_salt(self._salt.data_ptr())
        except Exception:
            pass

    def _signature(self, features, object_features, object_xywh_list, depth_gt, depth_mask):
        if object_features is None:
            objs = None
        else:
            if any(b is None for b in object_xywh_list):
                return None                                   # placeholder boxes are built on the host: eager only
            objs = tuple(int(o.shape[0]) for o in object_features) + (int(object_features[0].shape[1]),)
        p_drop = float(self.head.dropout) if self.head.training else 0.0
        return (tuple(features.shape), features.dtype, objs, tuple(depth_gt.shape), depth_gt.dtype, tuple(depth_mask.shape),
                depth_mask.dtype, p_drop)

    def _graph_step(self, sig, features, object_features, object_xywh_list, depth_gt, depth_mask):
        entry = self._graphs.get(sig)
        dev = self.flat_params.device
        if entry is None:
            # tensor-core mode: the layout conversion NCHW fp32 -> NHWC bf16 is the only reader of the caller's features, so it runs
            # eagerly from the caller's tensor into a static bf16 buffer right before the replay — no 232 MB device-to-device copy
            # into a static fp32 buffer, and the captured step starts from the converted map ("x" is a storage-less shape carrier)
            nhwc = engine_train.features_nhwc_ok(self.cfg, features.shape) and features.dtype in (torch.float32, torch.bfloat16)
            if nhwc:
                Bf, Cf, hf, wf = features.shape
                st = {"x": torch.empty(1, device=dev, dtype=torch.float32).as_strided(tuple(features.shape), (0, 0, 0, 0)),
                      "xh": torch.empty(Bf, hf, wf, Cf, device=dev, dtype=torch.bfloat16)}
            else:
                st = {"x": torch.empty(features.shape, device=dev, dtype=torch.float32)}
            st.update({
                  "gt": torch.empty(depth_gt.shape, device=dev, dtype=torch.float32),
                  "mask": torch.empty(depth_mask.shape, device=dev, dtype=depth_mask.dtype)})
            if object_features is not None:
                counts = [int(o.shape[0]) for o in object_features]
                st["feats"] = torch.empty(sum(counts), object_features[0].shape[1], device=dev, dtype=torch.float32)
                st["boxes"] = torch.empty(sum(counts), 4, device=dev, dtype=torch.float32)
                offs = [0]
                for c in counts:
                    offs.append(offs[-1] + c)
                st["obj_views"] = [st["feats"][offs[i]:offs[i + 1]] for i in range(len(counts))]
                st["box_views"] = [st["boxes"][offs[i]:offs[i + 1]] for i in range(len(counts))]
            self._load_static(st, features, object_features, object_xywh_list, depth_gt, depth_mask)
            torch.cuda.synchronize(dev)
            g = torch.cuda.CUDAGraph()
            l0 = ops.launch_count
            if self._capture_stream is None:      # the launching stream of the captured step: high priority (engine_train._branch)
                self._capture_stream = torch.cuda.Stream(dev, priority=-1 if engine_train.CHAIN_PRIORITY else 0)
            # thread_local: other threads of the process (the NCCL watchdog polling its events, the input prefetcher) may make
            # CUDA calls while this thread captures
            from . import engine
            refs = engine.offsets_capture_refs = []      # the graph bakes in the addresses of the CSR offset tensors: own them
            try:
                with torch.cuda.graph(g, pool=self._graph_pool, stream=self._capture_stream, capture_error_mode="thread_local"):
                    self._salt.add_(-7046029254386353131)       # 0x9E3779B97F4A7C15 as int64: a new mask stream per replay
                    self.grads.zero()
                    l1, l2 = self.forward_backward(st["x"], None if object_features is None else list(st["obj_views"]),
                                                   None if object_features is None else list(st["box_views"]), st["gt"], st["mask"],
                                                   xh=st.get("xh"))
            finally:
                engine.offsets_capture_refs = None
            st["_offset_refs"] = refs
            while len(self._graphs) >= self.MAX_GRAPHS:       # each entry pins its static buffers and pool memory: evict the
                self._graphs.pop(next(iter(self._graphs)))    # least recently used signature
            entry = self._graphs[sig] = (g, st, l1, l2, self.last_outputs, ops.launch_count - l0)
        else:
            self._graphs[sig] = self._graphs.pop(sig)         # most recently used last
            self._load_static(entry[1], features, object_features, object_xywh_list, depth_gt, depth_mask)
        g, st, l1, l2, outs, launches = entry
        g.replay()
        ops.launch_count += launches
        self.last_outputs = outs
        return l1, l2                                         # static device scalars: valid until the next step

    @staticmethod
    def _load_static(st, features, object_features, object_xywh_list, depth_gt, depth_mask) -> None:
        if "xh" in st:
            ops.nchw_to_nhwc_bf16(features.contiguous(), out=st["xh"])
        elif features.data_ptr() != st["x"].data_ptr():
            st["x"].copy_(features, non_blocking=True)
        st["gt"].copy_(depth_gt, non_blocking=True)
        st["mask"].copy_(depth_mask, non_blocking=True)
        if object_features is not None:
            torch.cat([o.to(st["feats"].device, torch.float32, non_blocking=True) for o in object_features], 0, out=st["feats"])
            torch.cat([b.to(st["boxes"].device, torch.float32, non_blocking=True) for b in object_xywh_list], 0, out=st["boxes"])

    def _local_step(self, features, object_features, object_xywh_list, depth_gt, depth_mask):
        if self.graph:
            self._arm_salt()
            sig = self._signature(features, object_features, object_xywh_list, depth_gt, depth_mask)
            if sig is not None:
                seen = self._graph_seen.get(sig, 0)
                if seen >= 2:                                 # two eager steps first: allocations, caches, parameter ranges
                    return self._graph_step(sig, features, object_features, object_xywh_list, depth_gt, depth_mask)
                if len(self._graph_seen) >= 4096:             # signatures that never repeat (ragged object counts of real data)
                    self._graph_seen.clear()
                self._graph_seen[sig] = seen + 1
        self.grads.zero()
        return self.forward_backward(features, object_features, object_xywh_list, depth_gt, depth_mask)

    def step(self, features, object_features, object_xywh_list, depth_gt, depth_mask):
        """One optimizer step on this rank's shard of the batch.  No host synchronisation anywhere in the step."""
        l1, l2 = self._local_step(features, object_features, object_xywh_list, depth_gt, depth_mask)
        world = self.world()
        if world > 1:
            if self.allreduce_events is not None:      # bench: CUDA events on the launching stream around the collective
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            dist.all_reduce(self.grads.flat, op=dist.ReduceOp.SUM, group=self.group)   # the ONE collective of the step
            if self.allreduce_events is not None:
                e1.record()
                self.allreduce_events.append((e0, e1))
        scale = 1.0 / world
        lr, beta1 = (self.lr, self.betas[0]) if self.schedule is None else self.schedule.at(self.step_index)
        self.step_index += 1
        clip = self.clip and self.clip > 0
        if clip:
            ops.sumsq(self.grads.flat, scale, out=self.sumsq)
        for lo, hi in self.ranges:                   # 1-3 ranges: everything except the parameters that never get a gradient
            ops.adamw_step(self.flat_params[lo:hi], self.grads.flat[lo:hi], self.exp_avg[lo:hi], self.exp_avg_sq[lo:hi], lr, beta1,
                           self.betas[1], self.eps, self.wd, self.step_index, grad_scale=scale,
                           grad_sumsq=self.sumsq if clip else None, max_norm=float(self.clip or 0.0))
        for m in self._modules:                    # the kernel wrote the parameters behind torch's version counters:
            m.__dict__.pop("_ocv_packed", None)    # drop the derived bf16 / packed inference operands
            m.__dict__.pop("_ocv_native", None)
        return l1, l2

    def loss_value(self, l1, l2) -> float:
        return float(self.coeffs[0] * l1.item() + self.coeffs[1] * l2.item())
