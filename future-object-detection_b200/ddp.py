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
    def __init__(self, params, bucket_mb=32.0, group=None, overlap=True):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.params = [p for p in params if p.requires_grad]
        order = list(reversed(self.params))
        cap = int(bucket_mb * (1 << 20)) // 4
        self.buckets, cur, cur_n = [], [], 0
        for p in order:
            if cur and cur_n + p.numel() > cap:
                self.buckets.append(cur)
                cur, cur_n = [], 0
            cur.append(p)
            cur_n += p.numel()
        if cur:
            self.buckets.append(cur)
        self.flat, self._bucket_of, self._offs, self._pending, self._launched, self._handles = [], {}, {}, [], [], []
        for bi, plist in enumerate(self.buckets):
            dev = plist[0].device
            # 16-byte aligned slots keep every gradient view vector-load friendly for the fused optimizer
            offs, n = [], 0
            for p in plist:
                offs.append(n)
                n += (p.numel() + 3) // 4 * 4
            flat = torch.zeros(n, dtype=torch.float32, device=dev)
            for p, o in zip(plist, offs):
                assert p.dtype == torch.float32, "gradient buckets are fp32 (checkpoint contract: fp32 parameters)"
                p.grad = flat[o:o + p.numel()].view_as(p)
                self._bucket_of[p], self._offs[p] = bi, o
                self._handles.append(p.register_post_accumulate_grad_hook(self._hook))
            self.flat.append(flat)
            self._pending.append(len(plist))
            self._launched.append(False)
        self.cuda = self.params[0].is_cuda if self.params else False
        self.overlap = overlap and self.cuda
        self.side = torch.cuda.Stream() if self.overlap else None
        self.enabled = True

    # ------------------------------------------------------------------ per step
    def zero_grad(self):
        """Replaces optimizer.zero_grad(set_to_none=True): the views must survive, so the buckets are memset."""
        for f in self.flat:
            f.zero_()
        for bi, plist in enumerate(self.buckets):
            self._pending[bi] = len(plist)
            self._launched[bi] = False

    def _reduce(self, bi):
        self._launched[bi] = True
        if self.world == 1 or not self.enabled:
            return
        flat = self.flat[bi]
        if self.overlap:
            self.side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.side):
                self._all_reduce_avg(flat)
        else:
            self._all_reduce_avg(flat)

    def _all_reduce_avg(self, flat):
        if flat.is_cuda:
            dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=self.group)
        else:  # gloo has no AVG
            dist.all_reduce(flat, group=self.group)
            flat.div_(self.world)

    def _hook(self, p):
        bi = self._bucket_of[p]
        if p.grad.data_ptr() != self.flat[bi].data_ptr() + 4 * self._offs[p]:
            raise RuntimeError("a gradient view was replaced (optimizer.zero_grad(set_to_none=True) or p.grad = ...); "
                               "use BucketedGradReducer.zero_grad()")
        self._pending[bi] -= 1
        if self._pending[bi] == 0:
            self._reduce(bi)

    def finish(self):
        """Call after backward: reduces buckets whose countdown never reached zero (parameters that got no gradient
        keep their zero slot, DDP's find_unused_parameters=False would have raised) and joins the side stream."""
        for bi in range(len(self.buckets)):
            if not self._launched[bi]:
                self._reduce(bi)
        if self.overlap and self.world > 1 and self.enabled:
            torch.cuda.current_stream().wait_stream(self.side)

    # ------------------------------------------------------------------ fused clipping on the flat buckets
    def clip_grad_norm_(self, max_norm, eps=1e-6):
        """torch.nn.utils.clip_grad_norm_ (trainer.py:186) over the flat buckets: one norm kernel per bucket instead
        of one per parameter; padding slots are zero and do not change the norm.  No host sync."""
        norms = torch._foreach_norm(self.flat)
        total = torch.linalg.vector_norm(torch.stack(norms))
        coef = torch.clamp(max_norm / (total + eps), max=1.0)
        torch._foreach_mul_(self.flat, coef)
        return total

    def remove(self):
        for h in self._handles:
            h.remove()
        self._handles = []
