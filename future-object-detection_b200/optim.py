"""Flat training state: every trainable parameter, its gradient, both AdamW moments and the bf16
operand copy the kernels read live in five contiguous arenas with one shared layout, so that

* gradient clipping + AdamW + the fp32->bf16 weight casts + zero_grad are ONE C-ABI call (fod_flat_adamw_step, three
  launches) instead of torch.nn.utils.clip_grad_norm_ + torch.optim.AdamW.step + ~290 per-weight casts
  (trainer.py:171,186-188; runs/_helper.py:84-107);
* the data-parallel gradient average is an all-reduce over slices of the gradient arena (ddp.BucketedGradReducer).

Parameters stay `nn.Parameter`s with the reference's names and shapes (the checkpoint contract): `p.data` becomes a
view into the parameter arena, `p.grad` a view into the gradient arena, and `p._fod_shadow` the bf16 view that
`model._q` hands to the GEMM kernels.
"""
import ctypes as C

import torch

from . import lib, ops
from . import profile as prof
from .ddp import BucketedGradReducer

BF16 = torch.bfloat16


class FlatTrainState:
    def __init__(self, model, lr=1e-4, lr_backbone=1e-5, weight_decay=1e-4, betas=(0.9, 0.999), eps=1e-8,
                 bucket_mb=32.0, group=None, overlap=True, wgrad_stream=True, tail_mb=0.0):
        """Parameter groups as runs/_helper.py:84-107: names containing "backbone" train at lr_backbone."""
        self.model = model
        named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
        rest = [p for n, p in named if "backbone" not in n]
        back = [p for n, p in named if "backbone" in n]
        # layout = the order in which backward produces gradients: heads/decoder, encoder, ..., backbone last
        order = list(reversed(rest)) + list(reversed(back))
        # parameters a module reads as one concatenated operand (fused_param_groups()) get adjacent slots in that order
        groups = [list(g) for m in model.modules() if hasattr(m, "fused_param_groups") for g in m.fused_param_groups()]
        groups = [g for g in groups if all(p.requires_grad for p in g)]
        member_of = {p: g for g in groups for p in g}
        assert len(member_of) == sum(len(g) for g in groups), "a parameter may belong to one fused group only"
        placed, new_order = set(), []
        for p in order:
            if p in placed:
                continue
            for q in member_of.get(p, (p,)):
                new_order.append(q)
                placed.add(q)
        order = new_order
        self.reducer = BucketedGradReducer([p for _, p in named], bucket_mb=bucket_mb, group=group, overlap=overlap,
                                           order=order, wgrad_stream=wgrad_stream, tail_mb=tail_mb)
        r = self.reducer
        dev = r.arena.device
        self.n = r.numel
        self.grads = r.arena
        self.params = torch.zeros(self.n, dtype=torch.float32, device=dev)
        self.exp_avg = torch.zeros_like(self.params)
        self.exp_avg_sq = torch.zeros_like(self.params)
        self.shadow = torch.zeros(self.n, dtype=BF16, device=dev)
        self.state = torch.zeros(8, dtype=torch.float32, device=dev)
        self.group_begin = [0] + ([r.offset[back[-1]]] if back and rest else [])
        self.group_lr = [lr if rest else lr_backbone] + ([lr_backbone] if back and rest else [])
        self.group_wd = [weight_decay] * len(self.group_begin)
        self.betas, self.eps = betas, eps
        for p in order:
            o = r.offset[p]
            view = self.params[o:o + p.numel()].view_as(p)
            view.copy_(p.data)
            p.data = view
            p._fod_shadow = self.shadow[o:o + p.numel()].view_as(p)
            p._fod_reducer = r   # autograd.py: gradient kernels accumulate straight into the arena slot
            p._fod_uses = 0
        # runs/_model.py:76-82 wraps the model in DistributedDataParallel, whose constructor broadcasts rank 0's
        # parameters and buffers (rank 0 alone loads the pretrained backbone, paper.py:96; the transformer is
        # initialised from each process's own RNG).  This object replaces that wrapper, so it does the same.
        self.group = group
        self.broadcast_from_rank0(refresh=False)
        self._args = None
        self._partial = None
        self.lr_version = 0  # bumped by set_lr(): learning rates are kernel ARGUMENTS, so a captured graph must follow
        if dev.type == "cuda":
            self._partial = torch.empty(lib.load().fod_optim_workspace_floats(), dtype=torch.float32, device=dev)
            self.refresh_shadow()
        self._hook = model.register_load_state_dict_post_hook(lambda m, keys: self.refresh_shadow())

    def broadcast_from_rank0(self, refresh=True):
        """Rank 0's trainable parameters (the arena, one collective), frozen parameters and buffers -> every rank.
        No-op without an initialised process group.  Also usable after an out-of-band load on rank 0 only."""
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(self.group) == 1:
            return
        src = dist.get_global_rank(self.group, 0) if self.group is not None else 0
        with torch.no_grad():
            dist.broadcast(self.params, src=src, group=self.group)
            in_arena = set(id(p) for p in self.reducer.order)
            rest = [p.data for p in self.model.parameters() if id(p) not in in_arena] + list(self.model.buffers())
            for t in rest:
                if t.numel():
                    dist.broadcast(t, src=src, group=self.group)
        if refresh and self.params.is_cuda:
            self.refresh_shadow()

    # ------------------------------------------------------------------ bf16 operand copies
    def refresh_shadow(self):
        """bf16(params) for the whole arena in one launch.  Runs automatically after load_state_dict; call it after
        any other out-of-band parameter edit."""
        if self.params.is_cuda:
            L = lib.load()
            lib.check(L.fod_cast_bf16(C.c_void_p(self.params.data_ptr()), C.c_void_p(self.shadow.data_ptr()), self.n,
                                      C.c_void_p(torch.cuda.current_stream().cuda_stream)), "fod_cast_bf16")
            ops.LAUNCH_COUNTS["cast_bf16_kernel"] += 1
        else:
            self.shadow.copy_(self.params)

    # ------------------------------------------------------------------ the step
    def begin_step(self):
        """Start of a training step.  Gradients were zeroed by the previous fused update (or are the initial zeros),
        so only the per-bucket countdowns of the reducer are re-armed."""
        self.reducer.rearm()
        for p in self.reducer.order:
            p._fod_uses = 0

    def finish_backward(self):
        self.reducer.finish()

    def step(self, max_norm=0.1, zero_grad=True):
        """clip_grad_norm_(max_norm) + AdamW + bf16 shadow refresh (+ zero_grad) on the arenas; no host sync.
        Statistics of the step (gradient norm before clipping, clip coefficient, step count) stay in `self.state`."""
        if not self.params.is_cuda:
            raise lib.FodError("FlatTrainState.step needs the CUDA library (no CPU fallback)")
        L = lib.load()
        if self._args is None:
            a = lib.AdamWArgs()
            a.params, a.grads = self.params.data_ptr(), self.grads.data_ptr()
            a.exp_avg, a.exp_avg_sq = self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr()
            a.shadow_bf16, a.state, a.partial = self.shadow.data_ptr(), self.state.data_ptr(), self._partial.data_ptr()
            a.n, a.num_groups = self.n, len(self.group_begin)
            for i, (b, lr, wd) in enumerate(zip(self.group_begin, self.group_lr, self.group_wd)):
                a.group_begin[i], a.group_lr[i], a.group_wd[i] = b, lr, wd
            a.beta1, a.beta2, a.eps = self.betas[0], self.betas[1], self.eps
            self._args = a
        a = self._args
        a.max_norm, a.zero_grad = float(max_norm if max_norm else 0.0), int(zero_grad)
        _t = prof.begin()
        lib.check(L.fod_flat_adamw_step(C.byref(a), C.c_void_p(torch.cuda.current_stream().cuda_stream)),
                  "fod_flat_adamw_step")
        # algorithmic bytes: sumsq reads g (4 B); update reads p,g,m,v (16 B) and writes p,m,v,g (16 B) + bf16 shadow (2 B)
        prof.end(_t, "adamw_flat_step", 0.0, 38.0 * self.n)
        ops.LAUNCH_COUNTS["sumsq_partial_kernel"] += 1
        ops.LAUNCH_COUNTS["optim_prepare_kernel"] += 1
        ops.LAUNCH_COUNTS["adamw_flat_kernel"] += 1

    def set_lr(self, lr=None, lr_backbone=None):
        """Learning-rate schedule hook (runs/_helper.py builds a StepLR over both groups).  The rates are by-value
        arguments of the update kernel, i.e. baked into a captured CUDA graph: `lr_version` tells
        train.GraphedTrainStep to re-capture before its next replay."""
        before = list(self.group_lr)
        if lr is not None:
            self.group_lr[0] = lr
        if lr_backbone is not None and len(self.group_lr) > 1:
            self.group_lr[1] = lr_backbone
        if self._args is not None:
            for i, v in enumerate(self.group_lr):
                self._args.group_lr[i] = v
        if self.group_lr != before:
            self.lr_version += 1

    @property
    def grad_norm(self):
        return self.state[0]

    # ------------------------------------------------------------------ checkpointing
    def _names(self):
        return {id(p): n for n, p in self.model.named_parameters()}

    def state_dict(self):
        """torch.optim.AdamW's own format (`state` keyed by parameter index in param_groups order with `step`,
        `exp_avg`, `exp_avg_sq`; `param_groups` with lr / betas / eps / weight_decay / params), built the way
        runs/_helper.py:84-107 builds the reference optimizer — group 0: non-backbone, group 1: backbone, each in
        named_parameters() order — so the `optimizer` entry of a reference checkpoint (trainer.py:282-299) resumes here
        and vice versa.  `param_names` (extra key, ignored by torch) pins the order by NAME, so a load never depends on
        the arena layout."""
        named = [(n, p) for n, p in self.model.named_parameters() if p.requires_grad]
        groups = [[(n, p) for n, p in named if "backbone" not in n], [(n, p) for n, p in named if "backbone" in n]]
        groups = [g for g in groups if g]
        r, step = self.reducer, self.state[2].detach().clone()
        state, pgs, names, idx = {}, [], [], 0
        for gi, g in enumerate(groups):
            ids = []
            for n, p in g:
                o = r.offset[p]
                state[idx] = {"step": step.clone(), "exp_avg": self.exp_avg[o:o + p.numel()].view_as(p).clone(),
                              "exp_avg_sq": self.exp_avg_sq[o:o + p.numel()].view_as(p).clone()}
                ids.append(idx)
                names.append(n)
                idx += 1
            pgs.append({"lr": self.group_lr[min(gi, len(self.group_lr) - 1)], "betas": tuple(self.betas), "eps": self.eps,
                        "weight_decay": self.group_wd[min(gi, len(self.group_wd) - 1)], "amsgrad": False, "params": ids})
        return {"state": state, "param_groups": pgs, "param_names": names}

    def load_state_dict(self, sd):
        """Accepts state_dict() above or a torch.optim.AdamW state_dict of the reference's optimizer.  Every entry is
        matched to its parameter (by name when `param_names` is present, else by the reference's group order) and its
        shape is checked; nothing is loaded on a mismatch."""
        named = [(n, p) for n, p in self.model.named_parameters() if p.requires_grad]
        order = [(n, p) for n, p in named if "backbone" not in n] + [(n, p) for n, p in named if "backbone" in n]
        ids = [i for g in sd["param_groups"] for i in g["params"]]
        if len(ids) != len(order):
            raise ValueError("optimizer state has %d parameters, the model has %d trainable ones" % (len(ids), len(order)))
        if "param_names" in sd:
            by_name = dict(order)
            if set(sd["param_names"]) != set(by_name):
                raise ValueError("optimizer state names do not match the model's trainable parameters")
            order = [(n, by_name[n]) for n in sd["param_names"]]
        plan = []
        for i, (n, p) in zip(ids, order):
            st = sd["state"].get(i)
            if st is None:
                continue
            if tuple(st["exp_avg"].shape) != tuple(p.shape) or tuple(st["exp_avg_sq"].shape) != tuple(p.shape):
                raise ValueError("optimizer state of %s has shape %s, the parameter %s" % (n, tuple(st["exp_avg"].shape), tuple(p.shape)))
            plan.append((p, st))
        r, steps = self.reducer, set()
        with torch.no_grad():
            for p, st in plan:
                o = r.offset[p]
                self.exp_avg[o:o + p.numel()].view_as(p).copy_(st["exp_avg"])
                self.exp_avg_sq[o:o + p.numel()].view_as(p).copy_(st["exp_avg_sq"])
                steps.add(float(st["step"]))
            if len(steps) > 1:
                raise ValueError("per-parameter step counts differ (%s): the flat AdamW keeps one counter" % sorted(steps)[:4])
            if steps:
                self.state[2] = steps.pop()
        lrs = [g["lr"] for g in sd["param_groups"]]
        self.set_lr(lrs[0], lrs[1] if len(lrs) > 1 else None)

    def close(self):
        self._hook.remove()
        self.reducer.remove()
        for p in self.reducer.order:
            for attr in ("_fod_shadow", "_fod_reducer", "_fod_uses"):
                if hasattr(p, attr):
                    delattr(p, attr)
