"""Graph-capturable data-parallel gradient averaging (replaces the DistributedDataParallel wrapper of
runs/_model.py:76-82 inside the captured training step).

torch's DDP reducer is host-driven (bucket bookkeeping in autograd hooks, work handles, a rebuilt-bucket pass after the
first iteration) and cannot live inside one CUDA graph.  This reducer keeps the same semantics — every trainable
parameter's gradient is replaced by its mean over ranks before the optimizer runs — with static structure:

* trainable parameters are laid into flat fp32 buckets in REVERSE registration order (heads and decoder first, then
  encoder, backbone last: the order in which backward produces them);
* `p.grad` is a persistent view into its bucket, so autograd accumulates straight into the bucket (no gather copy);
* a post-accumulate hook per parameter counts the bucket down; when the last gradient of a bucket has landed, the
  bucket is all-reduced (NCCL AVG) on a side stream that forks from the backward stream at that point, so the
  transfer over NVLink overlaps the rest of backward;
* `finish()` joins the side stream back before clipping / AdamW.

Under CUDA-graph capture the fork/join events and the NCCL kernels are captured as parallel branches of the step's
graph.  The same code runs with gloo on CPU (no streams), which is how tests/test_dist_cpu.py covers it.
"""
import torch
import torch.distributed as dist


class BucketedGradReducer:
    def __init__(self, params, bucket_mb=32.0, group=None, overlap=True, order=None, wgrad_stream=True, tail_mb=0.0):
        """params: iterable of parameters (trainable ones are taken).  order: optional explicit layout order of the
        trainable parameters (default: reverse registration order).  tail_mb: the LAST gradients of backward get a
        small bucket of their own — its all-reduce is the only one nothing can hide (the optimizer waits for it), so
        its latency should be that of a few MB, not of a full bucket.  Off by default: at 8 GPUs the exposed
        tail is 126 us of a 28.4 ms step and a 4 MB tail bucket changed nothing measurable."""
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.params = [p for p in params if p.requires_grad]
        order = list(order) if order is not None else list(reversed(self.params))
        assert len(order) == len(self.params)
        cap = int(bucket_mb * (1 << 20)) // 4
        # ONE arena holds every bucket back to back (the flat optimizer walks it in a single launch); a bucket is a
        # slice of it.  16-byte slots keep every gradient view vector-load friendly.
        self.offset, self.buckets, bounds = {}, [], []
        total = sum((p.numel() + 7) // 8 * 8 for p in order)
        tail_from = total - int(tail_mb * (1 << 20)) // 4 if 0 < tail_mb < bucket_mb else total + 1
        n, b0, cur = 0, 0, []
        for p in order:
            assert p.dtype == torch.float32, "gradient buckets are fp32 (checkpoint contract: fp32 parameters)"
            if cur and (n - b0 + p.numel() > cap or (b0 < tail_from <= n)):
                self.buckets.append(cur)
                bounds.append((b0, n))
                cur, b0 = [], n
            self.offset[p] = n
            cur.append(p)
            n += (p.numel() + 7) // 8 * 8  # 8-element slots: 16-byte aligned in the bf16 shadow arena too (TMA operands)
        if cur:
            self.buckets.append(cur)
            bounds.append((b0, n))
        dev = order[0].device if order else "cpu"
        self.order, self.numel = order, n
        self.arena = torch.zeros(n, dtype=torch.float32, device=dev)
        self.flat = [self.arena[a:b] for a, b in bounds]
        self._bucket_of, self._pending, self._launched, self._handles = {}, [], [], []
        for bi, plist in enumerate(self.buckets):
            for p in plist:
                o = self.offset[p]
                p.grad = self.arena[o:o + p.numel()].view_as(p)
                self._bucket_of[p] = bi
                self._handles.append(p.register_post_accumulate_grad_hook(self._hook))
            self._pending.append(len(plist))
            self._launched.append(False)
        self.cuda = self.params[0].is_cuda if self.params else False
        self.overlap = overlap and self.cuda
        self.side = torch.cuda.Stream() if self.overlap else None
        self.enabled = True
        self._ready = set()
        self.fused = {}  # autograd.FusedParams views over adjacent slots (their use counters are reset with the step)
        # second stream for gradient kernels that are off backward's critical path (autograd._OffCriticalPath); the
        # all-reduce of a bucket and the optimizer wait for it
        self.wgrad_stream = torch.cuda.Stream() if (self.cuda and wgrad_stream) else None
        self.keepalive = []  # operands of kernels in flight on wgrad_stream (see autograd._OffCriticalPath)

    # ------------------------------------------------------------------ per step
    def zero_grad(self):
        """Replaces optimizer.zero_grad(set_to_none=True): the views must survive, so the buckets are memset."""
        self.arena.zero_()
        self.rearm()

    def rearm(self):
        """Re-arm the per-bucket countdowns (the gradients themselves are zeroed by zero_grad or the fused optimizer)."""
        for bi, plist in enumerate(self.buckets):
            self._pending[bi] = len(plist)
            self._launched[bi] = False
        self._ready.clear()
        self.keepalive.clear()
        for f in self.fused.values():
            f._fod_uses = 0

    def _reduce(self, bi):
        self._launched[bi] = True
        if self.world == 1 or not self.enabled:
            return
        flat = self.flat[bi]
        if self.overlap:
            self.side.wait_stream(torch.cuda.current_stream())
            if self.wgrad_stream is not None:
                self.side.wait_stream(self.wgrad_stream)
            with torch.cuda.stream(self.side):
                self._all_reduce_avg(flat)
        else:
            if self.wgrad_stream is not None:
                torch.cuda.current_stream().wait_stream(self.wgrad_stream)
            self._all_reduce_avg(flat)

    def _all_reduce_avg(self, flat):
        if flat.is_cuda:
            dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=self.group)
        else:  # gloo has no AVG
            dist.all_reduce(flat, group=self.group)
            flat.div_(self.world)

    def _hook(self, p):
        if p in self._ready:
            # Second signal for the same parameter in one step: a gradient kernel wrote the slot directly
            # (mark_ready) AND autograd ran the parameter's AccumulateGrad node with an undefined gradient (it does,
            # once every producer has finished).  Either signal means "all contributions are enqueued".
            return
        self._ready.add(p)
        bi = self._bucket_of[p]
        if p.grad.data_ptr() != self.arena.data_ptr() + 4 * self.offset[p]:
            raise RuntimeError("a gradient view was replaced (optimizer.zero_grad(set_to_none=True) or p.grad = ...); "
                               "use BucketedGradReducer.zero_grad()")
        self._pending[bi] -= 1
        if self._pending[bi] == 0:
            self._reduce(bi)

    def mark_ready(self, p):
        """For gradients written into the arena by a kernel instead of autograd's AccumulateGrad (fused weight-gradient
        accumulation): counts the parameter's bucket down exactly as the post-accumulate hook would.  A fused view
        (autograd.FusedParams) signals every member."""
        for m in getattr(p, "_fod_members", (p,)):
            self._hook(m)

    def finish(self):
        """Call after backward: reduces buckets whose countdown never reached zero (parameters that got no gradient
        keep their zero slot, DDP's find_unused_parameters=False would have raised) and joins the side stream."""
        for bi in range(len(self.buckets)):
            if not self._launched[bi]:
                self._reduce(bi)
        if self.wgrad_stream is not None:
            torch.cuda.current_stream().wait_stream(self.wgrad_stream)
            self.keepalive.clear()  # the backward stream is ordered after every side-stream reader now
        if self.overlap and self.world > 1 and self.enabled:
            torch.cuda.current_stream().wait_stream(self.side)

    # ------------------------------------------------------------------ fused clipping on the flat buckets
    def clip_grad_norm_(self, max_norm, eps=1e-6):
        """torch.nn.utils.clip_grad_norm_ (trainer.py:186) over the flat buckets: one norm kernel per bucket instead
        of one per parameter; padding slots are zero and do not change the norm.  No host sync."""
        total = torch.linalg.vector_norm(self.arena)
        coef = torch.clamp(max_norm / (total + eps), max=1.0)
        self.arena.mul_(coef)
        return total

    def remove(self):
        for h in self._handles:
            h.remove()
        self._handles = []
