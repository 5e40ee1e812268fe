"""The training step as one object: forward -> VAELoss -> backward -> (data-parallel gradient
all-reduce) -> clip + Adam/AdamW, restating what the reference's Lightning module does around the
hot path (src/lightning_module.py:98-218 training_step, :390-477 optimizer / clipping / NaN guard)
and what Lightning's DDP strategy adds when ``devices > 1`` (main.py:86-99): batch sharding, one
process per GPU, gradients averaged over ranks by a bucketed all-reduce that overlaps backward.

Pieces:
  * ``GradBuckets``  — one flat fp32 gradient buffer laid out in reverse execution order and cut
    into buckets; every parameter's ``.grad`` is a view into it.  A bucket is all-reduced (NCCL on
    the GPU box, gloo in the CPU tests) the moment its last gradient has been written, while the
    rest of the backward schedule keeps running.
  * ``TrainStep``    — owns model + criterion + FusedAdamW + GradBuckets; ``step(x, cond)`` accepts
    host (pinned) or device tensors, optionally splits the per-GPU batch into micro-batches
    (gradients accumulate in the flat buffer; GroupNorm / attention are per sample, so the result
    is the same as one big batch) and returns the loss dict as 0-d device tensors.
  * ``shard_batch``  — the batch split DDP's DistributedSampler would do.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence

import torch
import torch.distributed as dist
import torch.nn as nn


def shard_batch(n: int, rank: int, world: int):
    """contiguous, equal shards (n must divide by world): -> slice of the global batch for `rank`."""
    if n % world != 0:
        raise ValueError(f"global batch {n} is not divisible by world size {world}")
    per = n // world
    return slice(rank * per, (rank + 1) * per)


class GradBuckets:
    """Flat gradient storage + bucketed all-reduce.

    params must be given in the order their gradients become ready during backward (last layer
    first).  ``mark_ready(p)`` is called once per parameter per reduction round; when every
    parameter of a bucket is ready the bucket's slice of the flat buffer is all-reduced
    asynchronously (op AVG where the backend has it, else SUM followed by a scale).
    ``finish()`` waits for all outstanding reductions.
    """

    ALIGN = 64   # elements

    def __init__(self, params: Sequence[nn.Parameter], bucket_bytes: int = 32 << 20, group=None):
        self.params = list(params)
        self.group = group
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.rank = dist.get_rank(group) if (dist.is_available() and dist.is_initialized()) else 0
        dev = self.params[0].device
        pad = lambda n: (n + self.ALIGN - 1) // self.ALIGN * self.ALIGN   # 256-byte aligned views: 16-byte accesses
        total = sum(pad(p.numel()) for p in self.params)
        self.flat = torch.zeros(total, dtype=torch.float32, device=dev)
        self.views: Dict[int, torch.Tensor] = {}
        self.bucket_of: Dict[int, int] = {}
        self.bounds: List[List[int]] = []       # [start, end) element offsets per bucket
        self.members: List[int] = []            # parameters per bucket
        off, start, count = 0, 0, 0
        for p in self.params:
            n = p.numel()
            self.views[id(p)] = self.flat[off:off + n].view(p.shape)
            self.bucket_of[id(p)] = len(self.bounds)
            off += pad(n)
            count += 1
            if (off - start) * 4 >= bucket_bytes:
                self.bounds.append([start, off]); self.members.append(count)
                start, count = off, 0
        if count:
            self.bounds.append([start, off]); self.members.append(count)
        self._pending = list(self.members)
        self._touched = set()
        self._works: list = []
        self._add_dst, self._add_src, self._add_par = [], [], []
        self.reduce_enabled = True
        self.launched = 0
        self.trace = None            # set to a list to collect (start, end) CUDA events around finish()'s waits
        for p in self.params:
            p.grad = self.views[id(p)]

    # ---- per step
    def zero(self):
        self.flat.zero_()
        self._pending = list(self.members)
        self._touched = set()

    def untouched(self):
        """parameters no gradient was produced for since zero() (unused parameters: modality_embedding, FiLM / inject
        layers, projectors of absent modalities): their .grad is a zero view, where autograd would leave None"""
        return [p for p in self.params if id(p) not in self._touched]

    def begin_round(self):
        if self._add_par:
            raise RuntimeError("GradBuckets.begin_round(): the previous round still has batched adds pending; "
                               "call flush() when a backward pass ends")
        self._pending = list(self.members)

    def view_of(self, p: nn.Parameter):
        """the slice of the flat buffer that is p.grad (kernels may accumulate into it directly and
        then call mark_ready), or None for a parameter this object does not own"""
        return self.views.get(id(p))

    FLUSH_AT = 64

    def accumulate(self, p: nn.Parameter, g: torch.Tensor):
        """grad sink used by the encoder/decoder backward schedules: add into the flat view.  The small
        gradients (GroupNorm affine, biases: ~270 tensors per backward pass at C4) are batched and added
        by one multi-tensor launch per FLUSH_AT tensors instead of one launch each."""
        self._add_dst.append(self.views[id(p)])
        self._add_src.append(g.view(p.shape))
        self._add_par.append(p)
        if len(self._add_par) >= self.FLUSH_AT:
            self.flush()

    def flush(self):
        """apply the batched adds and report their parameters ready (call before the round ends)"""
        if self._add_par:
            torch._foreach_add_(self._add_dst, self._add_src)
            pars = self._add_par
            self._add_dst, self._add_src, self._add_par = [], [], []
            for p in pars:
                self.mark_ready(p)

    def mark_ready(self, p: nn.Parameter):
        b = self.bucket_of.get(id(p))
        if b is None:
            return
        self._touched.add(id(p))
        self._pending[b] -= 1
        if self._pending[b] == 0 and self.reduce_enabled and self.world > 1:
            self._launch(b)

    def _launch(self, b: int):
        s, e = self.bounds[b]
        t = self.flat[s:e]
        backend = dist.get_backend(self.group)
        if backend == "nccl":
            w = dist.all_reduce(t, op=dist.ReduceOp.AVG, group=self.group, async_op=True)
            self._works.append((w, None))
        else:
            w = dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
            self._works.append((w, t))
        self.launched += 1

    def finish(self):
        """wait for every outstanding bucket; buckets that never filled (unused parameters) are
        reduced now so all ranks stay in step."""
        self.flush()
        ev = None
        if self.trace is not None and torch.cuda.is_available():
            # exposed part of the gradient all-reduce: time the compute stream spends between the end of backward and
            # the last bucket's completion (the waits below are stream dependencies, not host blocks)
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            ev[0].record()
        if self.world > 1 and self.reduce_enabled:
            for b, left in enumerate(self._pending):
                if left > 0:
                    self._pending[b] = 0
                    self._launch(b)
        for w, t in self._works:
            w.wait()
            if t is not None:
                t.div_(self.world)
        self._works = []
        if ev is not None:
            ev[1].record()
            self.trace.append(ev)


def backward_order(model: nn.Module) -> List[nn.Parameter]:
    """parameters in the order backward produces their gradients: decoder (last stage first), then
    encoder, then whatever is left (condition_proj, projector heads ...)."""
    seen, out = set(), []

    def push(ps):
        for p in ps:
            if id(p) not in seen and p.requires_grad:
                seen.add(id(p)); out.append(p)

    for name in ("decoder", "encoder"):
        stack = getattr(model, name, None)
        if stack is not None and hasattr(stack, "stages"):
            for st in reversed(stack.stages(True)):
                push(st.params_backward_order())
            push(stack.parameters())
    push(model.parameters())
    return out


class TrainStep:
    """One optimisation step of a BaseVAE / BetaVAE / ConditionalVAE / DisentangledConditionalVAE on this
    rank's shard."""

    def __init__(self, model: nn.Module, criterion: nn.Module, *, lr: float = 1e-4, weight_decay: float = 1e-4,
                 betas=(0.9, 0.999), eps: float = 1e-8, decoupled: bool = True, max_norm: float = 1.0,
                 micro_batch: Optional[int] = None, bucket_bytes: int = 32 << 20, group=None, use_graph: bool = False):
        """use_graph: capture the whole step (forward, loss, backward, all-reduce, optimizer) in a CUDA graph
        after two eager steps and replay it afterwards — for shapes whose step is bound by the ~3000 host
        launches per pass (the 28x28 configs).  The step counter, learning rate and Philox stream position
        then live on the device (``step_state``) so the captured kernels do not bake them in."""
        from .optim import FusedAdamW
        self.model, self.criterion = model, criterion
        self.use_graph = bool(use_graph)
        if self.use_graph and hasattr(criterion, "separation_weight"):
            # the disentangled model takes its modality indices from the host: the input projector, the decoder heads,
            # the reconstruction width (all-gray batch -> 1 channel) and the grouping of the two latent losses would be
            # frozen at capture and replayed for later batches with other indices
            raise NotImplementedError("TrainStep(use_graph=True) does not support DisentangledConditionalVAE / "
                                      "DisentangledVAELoss: its host-side modality indices cannot be replayed from a "
                                      "captured graph; use the eager step (the 128x128 workload gains nothing from a graph)")
        self._graph = None
        self._graph_key = None
        self._graph_eager_steps = 0
        self.graph_launches = 0            # C-ABI kernel launches captured in the graph (= per replay)
        self.micro_batch = micro_batch
        self.order = backward_order(model)
        self.buckets = GradBuckets(self.order, bucket_bytes, group)
        self.optimizer = FusedAdamW(self.order, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                                    decoupled=decoupled, max_norm=max_norm)
        self.device = self.order[0].device
        if self.buckets.world > 1 and self.device.type == "cuda":
            # persistent one-CTA-per-SM kernels and NCCL's CTAs otherwise wait for each other: leave a few SMs to the
            # overlapped gradient all-reduce (MVAE_RESERVE_SMS, default 0 = off until measured)
            import os
            from . import _lib as L
            L.lib().mvae_set_debug(12, int(os.environ.get("MVAE_RESERVE_SMS", "0")))
        # device-resident {optimizer step, learning rate, Philox position}, advanced by one kernel per step
        self.step_state = torch.tensor([0.0, float(lr), 0.0], dtype=torch.float64, device=self.device)
        if self.use_graph:
            self.optimizer.step_state = self.step_state
            model._step_state = self.step_state
            for name in ("decoder", "encoder"):
                stack = getattr(model, name, None)
                if stack is not None and hasattr(stack, "_step_state"):
                    stack._step_state = self.step_state
        self._hooks = []
        sunk = set()
        for name in ("decoder", "encoder"):
            stack = getattr(model, name, None)
            if stack is not None and hasattr(stack, "stages"):
                stack._grad_sink = self.buckets.accumulate
                stack._grad_views = self.buckets
                sunk.update(id(p) for p in stack.parameters())
        for p in self.order:
            if id(p) not in sunk:
                self._hooks.append(p.register_post_accumulate_grad_hook(self.buckets.mark_ready))

    # ------------------------------------------------------------------ checkpoint / resume
    def state_dict(self) -> Dict[str, object]:
        """everything a resumed run needs to continue bit for bit: weights, Adam moments, step count, learning rate
        and the Philox stream positions (what Lightning's checkpoint holds for the reference, lightning_module.py +
        configs/callbacks: model state_dict + optimizer state)"""
        rng = {"model": getattr(self.model, "_rng_step", 0)}
        for name in ("decoder", "encoder"):
            stack = getattr(self.model, name, None)
            if stack is not None and hasattr(stack, "_dropout_step"):
                rng[name] = stack._dropout_step
        return {"model": {k: v.detach().clone() for k, v in self.model.state_dict().items()},
                "optimizer": {k: ([t.detach().clone() for t in v] if isinstance(v, list) else v)
                              for k, v in self.optimizer.state_dict().items()},
                "rng": rng, "step_state": self.step_state.detach().clone()}

    def load_state_dict(self, sd: Dict[str, object]):
        with torch.no_grad():
            own = self.model.state_dict()
            for k, v in sd["model"].items():
                own[k].copy_(v)                 # in place: gradient views, pointer tables and graphs stay valid
        self.optimizer.load_state_dict(sd["optimizer"])
        self.step_state.copy_(sd["step_state"])
        # a checkpoint written by an eager TrainStep never advanced step_state[0] (eager AdamW takes t from the host
        # counter); a graph-mode step reads t from the device: keep the two in step whichever mode wrote the checkpoint
        self.step_state[0:1].fill_(float(self.optimizer.step_count))
        self.model._rng_step = int(sd["rng"].get("model", 0))
        for name in ("decoder", "encoder"):
            stack = getattr(self.model, name, None)
            if stack is not None and name in sd["rng"]:
                stack._dropout_step = int(sd["rng"][name])
        torch._C._increment_version(self.optimizer.params)
        # the packed operands follow NOW: a captured step assumes they are current when it starts (its own re-pack
        # launch sits at its end)
        from .blocks import refresh_packed
        refresh_packed(self._stacks())

    def _stacks(self):
        return [s for s in (getattr(self.model, "encoder", None), getattr(self.model, "decoder", None))
                if s is not None and hasattr(s, "_pack_cache")]

    def close(self):
        for h in self._hooks:
            h.remove()
        for name in ("decoder", "encoder"):
            stack = getattr(self.model, name, None)
            if stack is not None and hasattr(stack, "_grad_sink"):
                stack._grad_sink = None
                stack._grad_views = None

    def _to_device(self, t):
        if t is None or t.device == self.device:
            return t
        return t.to(self.device, non_blocking=True)

    # ------------------------------------------------------------------ CUDA-graph mode
    def set_lr(self, lr: float):
        """learning rate for the next steps (what a scheduler calls in graph mode: the captured AdamW kernel
        reads it from the device)"""
        self.optimizer.lr = float(lr)          # the setter mirrors it into step_state when the optimizer has one
        if self.optimizer.step_state is None:
            self.step_state[1:2].fill_(float(lr))

    def _graph_step(self, x, cond):
        from . import _lib as L
        from . import ops
        disentangled = hasattr(self.criterion, "separation_weight")
        key = (tuple(x.shape), None if cond is None else tuple(cond.shape), disentangled)
        if self._graph is None or self._graph_key != key:
            if self._graph_key != key:
                self._graph, self._graph_key, self._graph_eager_steps = None, key, 0
            if self._graph_eager_steps < 2:
                # eager steps first: packed-operand caches, scratch buffers, optimizer tables, kernel attributes
                self._graph_eager_steps += 1
                ops.step_state_advance(self.step_state, 1.0)
                return self._eager_step(x, cond)
            if disentangled and cond is not None and cond.device.type != "cpu":
                raise RuntimeError("graph mode needs the modality indices on the host (the model's output-width "
                                   "decision would otherwise synchronise inside the capture)")
            self._gx = torch.empty_like(x, device=self.device)
            self._gc = None if cond is None else (cond.clone() if disentangled else torch.empty_like(cond, device=self.device))
            self._gx.copy_(x, non_blocking=True)
            if self._gc is not None and not disentangled:
                self._gc.copy_(cond, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            n0 = L.lib().mvae_launch_count()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                ops.step_state_advance(self.step_state, 1.0)
                self._gout = self._eager_step(self._gx, self._gc)
            self.graph_launches = int(L.lib().mvae_launch_count() - n0)
            # the scratch buffers the captured kernels point at must outlive later re-allocations
            self._graph_keepalive = list(ops._scratch.values())
            self._graph = g
            # the capture itself did not run anything: undo its host-side bookkeeping, then fall through to replay
            self.optimizer.step_count -= 1
        self._gx.copy_(x, non_blocking=True)
        if self._gc is not None and not disentangled:
            self._gc.copy_(cond, non_blocking=True)
        self._graph.replay()
        self.optimizer.step_count += 1
        # parameters changed behind autograd's back (version bump for any cache keyed on them); the graph's last node
        # re-packed the conv operands, so those are current
        torch._C._increment_version(self.optimizer.params)
        from .blocks import mark_packed_current
        mark_packed_current(self._stacks())
        return dict(self._gout)

    def release_graph(self):
        """drop the captured graph (before destroying a process group whose collectives it contains)"""
        import gc
        self._graph = None
        self._gout = None
        self._graph_keepalive = None
        self._graph_key = None
        gc.collect()
        if torch.cuda.is_available():
            torch.cuda.synchronize()

    def step(self, x: torch.Tensor, cond: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
        if self.use_graph:
            return self._graph_step(x, cond)
        return self._eager_step(x, cond)

    def step_eager(self, x: torch.Tensor, cond: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
        """one step launched kernel by kernel even in graph mode (per-launch profiling)"""
        if self.use_graph:
            from . import ops
            ops.step_state_advance(self.step_state, 1.0)
        return self._eager_step(x, cond)

    def _eager_step(self, x: torch.Tensor, cond: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
        """x: (B, C, H, W) fp32 in [-1, 1]; cond: (B, n_modalities) one-hot, None, or — for a
        DisentangledConditionalVAE with DisentangledVAELoss — the (B,) int64 modality indices
        (lightning_module.py:116-118).  Host tensors are copied to the device here (use pinned
        memory for an asynchronous copy); a host index vector also stays on the host, which saves
        the model's one device->host read (the reconstruction-width decision).

        Micro-batching a disentangled step evaluates the two batch-coupled latent losses per
        micro-batch (what the reference does under ``accumulate_grad_batches``); leave
        ``micro_batch`` unset to get the whole-shard losses of the reference's plain step."""
        disentangled = hasattr(self.criterion, "separation_weight")   # the reference's own test (:191)
        x = self._to_device(x)
        if not disentangled:
            cond = self._to_device(cond)
        n = x.shape[0]
        mb = self.micro_batch or n
        chunks = [(i, min(i + mb, n)) for i in range(0, n, mb)]
        self.buckets.zero()
        totals: Dict[str, torch.Tensor] = {}
        # verification hook: externally supplied reparameterisation noise follows the micro-batch split
        eps_full = getattr(self.model, "eps_override", None)
        # own Philox noise: one stream position per optimizer step for every micro-batch and rank, the element counter
        # addressed by the sample's index in the global batch (equal shards) — ranks draw different noise and the
        # noise of a sample does not depend on the number of ranks or the micro-batch split
        rng0 = getattr(self.model, "_rng_step", None)
        drop0 = {name: getattr(self.model, name)._dropout_step for name in ("decoder", "encoder")
                 if hasattr(getattr(self.model, name, None), "_dropout_step")}
        for ci, (a, b) in enumerate(chunks):
            if rng0 is not None:
                self.model._rng_step = rng0
                self.model._rng_sample_base = self.buckets.rank * n + a
                for name in ("decoder", "encoder"):
                    stack = getattr(self.model, name, None)
                    if stack is not None and hasattr(stack, "_rng_sample_base"):
                        stack._rng_sample_base = self.buckets.rank * n + a
                        stack._dropout_step = drop0[name]     # one dropout stream position per optimizer step
            last = ci == len(chunks) - 1
            self.buckets.reduce_enabled = last
            self.buckets.begin_round()
            xs = x[a:b]
            if eps_full is not None and len(chunks) > 1:
                self.model.eps_override = eps_full[a:b]
            if cond is not None:
                out = self.model(xs, cond[a:b])
            else:
                out = self.model(xs)
            if disentangled:
                ld = self.criterion(out, xs)
            else:
                ld = self.criterion(inputs=xs, reconstructions=out["reconstruction"], posteriors=out["posterior"],
                                    priors=out["prior"])
            w = (b - a) / n
            (ld["loss"] * w if len(chunks) > 1 else ld["loss"]).backward()
            self.buckets.flush()
            for k, v in ld.items():
                v = v.detach() * w if len(chunks) > 1 else v.detach()
                totals[k] = v if k not in totals else totals[k] + v
            del out, ld
        if eps_full is not None:
            self.model.eps_override = eps_full
        if rng0 is not None:
            self.model._rng_sample_base = 0
            for name in ("decoder", "encoder"):
                stack = getattr(self.model, name, None)
                if stack is not None and hasattr(stack, "_rng_sample_base"):
                    stack._rng_sample_base = 0
        self.buckets.finish()
        # torch.optim leaves a parameter without a gradient alone (no weight decay, no moment update)
        self.optimizer.step(skip=self.buckets.untouched())
        # the bf16 operands of every conv follow their fp32 masters with ONE launch (instead of ~110 per-conv packs
        # at the head of the next forward)
        from .blocks import refresh_packed
        refresh_packed(self._stacks())
        return totals

    @torch.no_grad()
    def validate(self, x: torch.Tensor, cond: Optional[torch.Tensor] = None, with_latent: bool = False) -> Dict[str, float]:
        """validation_step of the reference (src/lightning_module.py:220-310): eval-mode forward, the
        reconstruction / KL metrics of src/utils/metrics.py and the criterion's loss with its NaN/Inf ->
        1e6 guard.  The metric kernels leave everything on the device; ONE device->host read returns
        the whole dict (the reference pays one ``.item()`` per metric)."""
        from .metrics import KL_KEYS, LATENT_KEYS, RECON_KEYS, validation_metrics_device
        disentangled = hasattr(self.criterion, "separation_weight")
        x = self._to_device(x)
        if not disentangled:
            cond = self._to_device(cond)
        was_training = self.model.training
        self.model.eval()
        try:
            out = self.model(x, cond) if cond is not None else self.model(x)
            if disentangled:
                ld = self.criterion(out, x)
            else:
                ld = self.criterion(inputs=x, reconstructions=out["reconstruction"], posteriors=out["posterior"],
                                    priors=out["prior"])
            vec = torch.cat([validation_metrics_device(x, out, with_latent), ld["loss"].detach().float().reshape(1)])
        finally:
            self.model.train(was_training)
        vals = vec.tolist()
        keys = RECON_KEYS + KL_KEYS + (LATENT_KEYS if with_latent else ())
        res = {f"val/{k}": v for k, v in zip(keys, vals[:-1])}
        loss = vals[-1]
        res["val/loss"] = 1e6 if (loss != loss or loss in (float("inf"), float("-inf"))) else loss
        return res
