"""Data-parallel training across the GPUs of one box (north_star item 4; the reference has no distributed
code at all -- SURVEY section 5 / 8e).  One process per GPU, torch.distributed for the plumbing, NCCL all-reduce
of gradient buckets on a side stream, launched as soon as each backward stage has finished writing its slice
of the engine's flat gradient buffer, so that the transfer overlaps the remaining backward kernels.

The unit of sharding is the clip (frames folded into the batch): no data-path collective, one gradient
exchange per step; sampling needs no exchange at all (replicas only).
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int):
    """Contiguous [lo, hi) slice of n_items owned by `rank` (remainder spread over the first ranks)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def plan_buckets(stage_end, bucket_elems):
    """Group consecutive backward stages into buckets of >= bucket_elems fp32 elements.
    stage_end[i] = end offset of stage i in the flat buffer.  Returns [(last_stage, lo, hi)]."""
    buckets, lo = [], 0
    for i, end in enumerate(stage_end):
        last = i == len(stage_end) - 1
        if end - lo >= bucket_elems or (last and end > lo):
            buckets.append((i, lo, end))
            lo = end
    if buckets and buckets[-1][2] != stage_end[-1]:
        buckets.append((len(stage_end) - 1, lo, stage_end[-1]))
    return buckets


class GradientAllReducer:
    """Attach to a UNetModel / UNetAudio: averages every gradient across the ranks while backward is still running.

    Torso: the engine writes all torso gradients into one flat fp32 buffer in backward-completion order; its stage hook
    all-reduces contiguous >= `bucket_mb` slices on a side stream as soon as the stages that fill them are done.

    Every other parameter (the wav2vec2 encoder, 94 M of the 110 M parameters) lives in ONE persistent flat fp32 buffer
    with a FIXED layout over all of them (reverse registration order = the order autograd finishes them, cut into
    >= `bucket_mb` buckets).  A post-accumulate-grad hook per parameter counts its bucket down; when the last gradient
    of a bucket has arrived the bucket's gradients are copied into the flat buffer by one multi-tensor copy, the slice
    is all-reduced on the side stream -- while the encoder's backward continues -- and `p.grad` becomes a view of the
    buffer (nothing is copied back).  A parameter that received no gradient on this rank (HF LayerDrop skips whole
    encoder layers per rank and step) contributes zeros: the collective's length never depends on which layers ran,
    which is what made the round-1 `torch.cat` of the present gradients unsafe.  Buckets are issued STRICTLY in index
    order on every rank (NCCL matches collectives by issue order).  Which parameters will receive a gradient in this
    backward is read off the autograd graph when the torso's backward starts (the AccumulateGrad nodes reachable from
    the audio features' grad_fn -- exact whatever the encoder's Python code does), so a bucket whose layer was skipped
    does not wait for gradients that will never come; whatever is still pending when backward ends is flushed by
    `finish()`, which then makes the compute stream wait for the side stream; it must run before `optimizer.step()`."""

    def __init__(self, model, group=None, bucket_mb: float = 25.0, force: bool = False, overlap=None):
        self.model, self.group = model, group
        # overlap=False (B2D_DP_OVERLAP=0): no collective is issued while backward runs; finish() all-reduces the two
        # flat buffers with one call each.  NCCL's CTAs hold SMs while a collective is in flight, and a one-wave
        # persistent kernel launched meanwhile runs its displaced CTAs as a second wave (measured in
        # profiles/r2_parity_and_dp.md), so "overlapped" is not automatically the faster schedule.
        if overlap is None:
            overlap = os.environ.get("B2D_DP_OVERLAP", "1") != "0"
        self.overlap = bool(overlap)
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.active = self.world > 1 or force          # force: run the whole path on one rank (tests)
        self.bucket_elems = int(bucket_mb * (1 << 20) / 4)
        self._buckets = None
        self._next = 0
        self._comm = None
        self._torso_ids = {id(p) for p in model.engine().torso_params()}
        self.other_params = [p for p in reversed(list(model.parameters()))
                             if id(p) not in self._torso_ids and p.requires_grad]
        model._grad_stage_hook = self._on_stage
        model._grad_final_hook = self._on_final
        model._grad_begin_hook = self._on_begin
        model._grad_backward_end_hook = self._on_backward_end
        self.bytes_reduced = 0
        # ---- fixed layout of the non-torso gradients
        self._flat = None
        self._views = None
        self._bucket_of = {}
        self._obuckets = []                            # [lo, hi, [param indices]]
        lo = off = 0
        cur = []
        for i, p in enumerate(self.other_params):
            self._bucket_of[id(p)] = len(self._obuckets)
            cur.append(i)
            off += (p.numel() + 3) // 4 * 4            # 16-byte aligned views
            if off - lo >= self.bucket_elems:
                self._obuckets.append([lo, off, cur])
                lo, cur = off, []
        if cur:
            self._obuckets.append([lo, off, cur])
        self._other_total = off
        self._remaining = [0] * len(self._obuckets)    # expected gradients of the bucket that have not arrived yet
        self._expected = set()
        self._next_other = 0                           # buckets [0, _next_other) have been issued this step
        self._index = {id(p): i for i, p in enumerate(self.other_params)}
        self._handles = []
        if self.active:
            for p in self.other_params:
                self._handles.append(p.register_post_accumulate_grad_hook(self._on_param))
        self._armed = False          # the expected set of this backward is known (set by _on_begin)
        self._flushed = False        # this backward's remaining buckets have been issued (by the end-of-backward hook)

    def detach(self):
        for h in self._handles:
            h.remove()
        self._handles = []
        self.model._grad_stage_hook = None
        self.model._grad_final_hook = None
        self.model._grad_begin_hook = None
        self.model._grad_backward_end_hook = None

    def _stream(self, device):
        if self._comm is None and device.type == "cuda":
            self._comm = torch.cuda.Stream(device=device)
        return self._comm

    def _all_reduce(self, t):
        if self.world > 1 or (self.active and dist.is_initialized()):
            if t.is_cuda:
                dist.all_reduce(t, op=dist.ReduceOp.AVG, group=self.group)
            else:                                  # gloo (CPU tests) has no AVG
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
                t.div_(self.world)
        self.bytes_reduced += t.numel() * 4

    def _reduce_async(self, view):
        """All-reduce `view` on the side stream, ordered after everything enqueued on the current stream so far."""
        comm = self._stream(view.device)
        if comm is not None:
            comm.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(comm):
                self._all_reduce(view)
            view.record_stream(comm)
        else:
            self._all_reduce(view)

    # ---- torso: stage hooks of the engine's backward
    def _on_stage(self, G, stage):
        if not self.active:
            return
        if self._buckets is None:
            self._buckets = plan_buckets(G.stage_end, self.bucket_elems if self.overlap else 1 << 62)
        if stage == 0:
            self._next = 0
        while self._next < len(self._buckets) and self._buckets[self._next][0] <= stage:
            _, lo, hi = self._buckets[self._next]
            self._next += 1
            self._reduce_async(G.flat[lo:hi])

    def _on_final(self, G):
        if self._comm is not None:
            torch.cuda.current_stream().wait_stream(self._comm)

    # ---- everything else: per-parameter hooks into the persistent flat buffer
    def _ensure_flat(self, device):
        if self._flat is None or self._flat.device != device:
            self._flat = torch.zeros(max(self._other_total, 1), device=device, dtype=torch.float32)
            self._views, off = [], 0
            for p in self.other_params:
                self._views.append(self._flat[off:off + p.numel()].view(p.shape))
                off += (p.numel() + 3) // 4 * 4

    def _on_begin(self, roots):
        """Called when the torso's backward starts with the grad_fn(s) through which gradients will flow on into the
        rest of the model: collects the parameters whose AccumulateGrad node is reachable, i.e. exactly those that
        will get a gradient in this backward."""
        if not self.active or not self.other_params:
            return
        self._expected.clear()
        self._remaining = [0] * len(self._obuckets)
        self._flushed = False
        seen, stack = set(), [r for r in roots if r is not None]
        while stack:
            node = stack.pop()
            if node in seen:
                continue
            seen.add(node)
            var = getattr(node, "variable", None)           # AccumulateGrad
            if var is not None and id(var) in self._index and id(var) not in self._expected:
                self._expected.add(id(var))
                self._remaining[self._bucket_of[id(var)]] += 1
            for nxt, _ in node.next_functions:
                if nxt is not None:
                    stack.append(nxt)
        self._armed = True

    def _on_param(self, p):
        self._flushed = False
        if not self._armed or not self.overlap:   # no torso hook (unknown expected set) / no overlap: all in finish()
            return
        b = self._bucket_of[id(p)]
        if b < self._next_other:
            raise RuntimeError("GradientAllReducer: a gradient arrived after its bucket had been all-reduced")
        if id(p) in self._expected:
            self._expected.discard(id(p))
            self._remaining[b] -= 1
        while self._next_other < len(self._obuckets) and self._remaining[self._next_other] == 0 \
                and self._next_other <= b:
            self._launch_bucket(self._next_other)

    def _launch_bucket(self, b):
        lo, hi, idxs = self._obuckets[b]
        ps = [self.other_params[i] for i in idxs]
        self._ensure_flat(ps[0].device)
        have = [(self._views[i], p.grad) for i, p in zip(idxs, ps)
                if p.grad is not None and p.grad.data_ptr() != self._views[i].data_ptr()]
        for i, p in zip(idxs, ps):
            if p.grad is None:
                self._views[i].zero_()
        if have:
            with torch.no_grad():
                torch._foreach_copy_([v for v, _ in have], [g for _, g in have])
        self._reduce_async(self._flat[lo:hi])
        for i, p in zip(idxs, ps):
            p.grad = self._views[i]
        self._next_other = b + 1

    def _launch_all(self):
        """Every non-torso gradient in ONE collective (no-overlap mode)."""
        ps = self.other_params
        self._ensure_flat(ps[0].device)
        have = [(v, p.grad) for v, p in zip(self._views, ps) if p.grad is not None and p.grad.data_ptr() != v.data_ptr()]
        for v, p in zip(self._views, ps):
            if p.grad is None:
                v.zero_()
        if have:
            with torch.no_grad():
                torch._foreach_copy_([v for v, _ in have], [g for _, g in have])
        self._reduce_async(self._flat[:self._other_total])
        for v, p in zip(self._views, ps):
            p.grad = v

    def _flush_other(self):
        if self.other_params and self.active:
            if not self.overlap and self._next_other == 0:
                self._launch_all()
            else:
                for b in range(self._next_other, len(self._obuckets)):
                    self._launch_bucket(b)
            self._remaining = [0] * len(self._obuckets)
            self._expected.clear()
            self._next_other = 0
            self._armed = False
            self._flushed = True

    def _on_backward_end(self):
        """The engine deferred its stage hooks to the end of the backward pass (deferred weight gradients): issue the
        remaining buckets of the other parameters FIRST, so that every rank issues [other buckets in index order, torso
        buckets in stage order] whatever the timing of its hooks was."""
        self._flush_other()

    def finish(self):
        """Flush the buckets whose parameters did not all receive a gradient this step, then make the compute stream
        wait for every in-flight all-reduce.  Call between `loss.backward()` and `optimizer.step()`."""
        if not self.active:
            return
        if not self._flushed:        # the end-of-backward hook already issued them when the engine deferred its stages
            self._flush_other()
        ov = getattr(self.model.engine(), "_overflow", None) if hasattr(self.model, "engine") else None
        if ov is not None and ov.is_cuda and self.world > 1:
            # a step is skipped on every rank or on none: MAX of the saturation counters (one 4-byte collective)
            comm = self._stream(ov.device)
            comm.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(comm):
                dist.all_reduce(ov[:1], op=dist.ReduceOp.MAX, group=self.group)
        if self._comm is not None:
            torch.cuda.current_stream().wait_stream(self._comm)


def shard_rows(L: int, rank: int, world: int, align: int = 256):
    """Query rows [q_row0, q_row0 + q_rows) of an attention over L positions owned by `rank`: equal `align`-aligned
    shares, or None when L does not split that way (the layer then runs replicated)."""
    if world <= 1 or L % (world * align) != 0:
        return None
    per = L // world
    return rank * per, per


class SequenceParallelAttention:
    """ONE clip (UNetModel(dims=3): attention over L = T*H*W positions, 1,048,576 at cfg5) split
    across the GPUs of a box.  Attention is >= 99 % of that step's FLOPs, so everything else (convolutions, GroupNorm,
    weight gradients) runs replicated on every rank with identical inputs and only the attention core is sharded, by
    QUERY rows: each rank already holds the full qkv tensor, so there is no K/V exchange --
      forward : rank r computes out[q rows of r] against all keys; one all-gather of `out` (L x C fp16);
      backward: rank r visits its query rows only (b2d_attention_bwd_rows): dQ for its rows, partial dK / dV; one
                all-reduce(sum) of dqkv (L x 3C fp16) yields the full gradient on every rank.
    Layers whose L does not split into 256-aligned shares, or with head dim 256 (0.3 % of the FLOPs), run replicated.
    Parameter gradients are identical on all ranks up to atomic-order rounding; `GradientAllReducer` (AVG) keeps the
    replicas bit-identical."""

    def __init__(self, model, group=None, min_rows: int = 4096):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.min_rows = min_rows
        self.bytes_moved = 0
        model.engine().seq_parallel = self if self.world > 1 else None

    def rows(self, n_samples: int, L: int, head_dim: int, heads: int = 1):
        """(q_row0, q_rows) of this rank, or None when the layer runs replicated: one sample (the row shares must be
        contiguous for the in-place all-gather), a head dim whose forward has a row-range kernel (64, 128), enough
        rows to be worth two collectives."""
        if n_samples != 1 or head_dim not in (64, 128) or L < self.min_rows * self.world:
            return None
        return shard_rows(L, self.rank, self.world)

    def _gather(self, flat, rows):
        mine = flat[rows[0]:rows[0] + rows[1]]
        if flat.is_cuda:
            dist.all_gather_into_tensor(flat, mine, group=self.group)      # in place: the input is this rank's slice
        else:                                                               # gloo (CPU tests)
            parts = [torch.empty_like(mine) for _ in range(self.world)]
            dist.all_gather(parts, mine.clone(), group=self.group)
            for r, part in enumerate(parts):
                flat[r * rows[1]:(r + 1) * rows[1]].copy_(part)
        self.bytes_moved += flat.numel() * flat.element_size()

    def all_gather_rows(self, out, rows):
        """out (1, H', W, C) contiguous with only this rank's rows valid -> all rows valid (in place)."""
        self._gather(out.view(-1, out.shape[-1]), rows)

    def all_gather_lse(self, lse, rows):
        if lse.shape[0] * lse.shape[1] != 1:
            raise RuntimeError("sequence-parallel lse gather expects one sample and one head")
        self._gather(lse.view(-1), rows)

    def all_reduce_max(self, t):
        """MAX over the ranks of a small fp32 tensor (the loss-scale maximum: every rank must pick the same scale)."""
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)

    def all_reduce_sum(self, t):
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        if t.dtype == torch.float16:
            # the kernels store gradients with saturating conversions (+-65504); the fp16 sum of the shards' partial
            # dK / dV can pass that and become inf where the un-sharded kernel would have saturated: do the same here
            t.clamp_(-65504.0, 65504.0)
        self.bytes_moved += t.numel() * t.element_size()


def broadcast_parameters(model, src: int = 0, group=None):
    """Identical weights on every rank before the first step."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    with torch.no_grad():
        for t in list(model.parameters()) + list(model.buffers()):
            dist.broadcast(t, src=src, group=group)
    eng = model.engine() if hasattr(model, "engine") else None
    if eng is not None:
        eng._weights_maybe_stale = True     # collectives write parameters without bumping `_version`: re-pack on next use
