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
                 bucket_mb=32.0, group=None, overlap=True, wgrad_stream=True):
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
                                           order=order, wgrad_stream=wgrad_stream)
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
        self._args = None
        self._partial = None
        self.lr_version = 0  # bumped by set_lr(): learning rates are kernel ARGUMENTS, so a captured graph must follow
        if dev.type == "cuda":
            self._partial = torch.empty(lib.load().fod_optim_workspace_floats(), dtype=torch.float32, device=dev)
            self.refresh_shadow()
        self._hook = model.register_load_state_dict_post_hook(lambda m, keys: self.refresh_shadow())

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
    def state_dict(self):
        return {"step": self.state[2].clone(), "exp_avg": self.exp_avg.clone(), "exp_avg_sq": self.exp_avg_sq.clone(),
                "group_lr": list(self.group_lr), "layout_numel": self.n}

    def load_state_dict(self, sd):
        assert sd["layout_numel"] == self.n, "optimizer arena layout mismatch"
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.state[2] = sd["step"]
        self.set_lr(*sd["group_lr"]) if len(sd["group_lr"]) > 1 else self.set_lr(sd["group_lr"][0])

    def close(self):
        self._hook.remove()
        self.reducer.remove()
        for p in self.reducer.order:
            for attr in ("_fod_shadow", "_fod_reducer", "_fod_uses"):
                if hasattr(p, attr):
                    delattr(p, attr)
