"""Training-step drivers for the drop-in model: the reference Trainer's inner step (future_od/trainer.py:171-189:
zero_grad, forward, backward, clip_grad_norm_(0.1), optimizer.step) either eagerly or as ONE CUDA graph.

The whole step is capturable because nothing in it touches the host: the matcher / loss run on the device
(no `.cpu()`, no scipy, no `.item()`), shapes are static, and AdamW runs fused with capturable state.
"""
import torch


def setup_optimizer(model, detr_args, capturable=False):
    """runs/_helper.py:84-107: AdamW with the backbone parameters in their own group at lr_backbone."""
    named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
    groups = [{"params": [p for n, p in named if "backbone" not in n]},
              {"params": [p for n, p in named if "backbone" in n], "lr": detr_args.lr_backbone}]
    return torch.optim.AdamW(groups, lr=detr_args.lr, weight_decay=detr_args.weight_decay, fused=True,
                             capturable=capturable)


class TrainStep:
    """Eager step.  `model` may be DDP-wrapped; `params` are the trainable parameters that get clipped."""

    def __init__(self, model, optimizer, max_norm=0.1, distributed=False):
        self.model, self.opt, self.max_norm, self.distributed = model, optimizer, max_norm, distributed
        self.params = [p for p in model.parameters() if p.requires_grad]

    def __call__(self, data):
        self.opt.zero_grad(set_to_none=True)
        _, _, loss, stats, _ = self.model(data=data, distributed=self.distributed)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(self.params, self.max_norm)
        self.opt.step()
        return loss


class GraphedTrainStep:
    """The same step captured into one CUDA graph (single process / single GPU).  Inputs are copied into static
    device buffers, the graph is replayed, and the static loss tensor is returned (read it with .item() if needed).

    The optimizer must have been built with capturable=True.  `warmup` eager steps run first on a side stream, as
    CUDA-graph capture requires (allocator warm-up, cuDNN autotuning, lazily created kernel attributes)."""

    def __init__(self, model, optimizer, example, max_norm=0.1, warmup=3):
        self.eager = TrainStep(model, optimizer, max_norm)
        self.static = {k: (v.clone() if torch.is_tensor(v) else v) for k, v in example.items()}
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(warmup):
                self.eager(self.static)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        optimizer.zero_grad(set_to_none=True)
        with torch.cuda.graph(self.graph):
            self.loss = self.eager(self.static)
        torch.cuda.synchronize()

    def __call__(self, data):
        for k, v in data.items():
            if torch.is_tensor(v):
                self.static[k].copy_(v, non_blocking=True)
        self.graph.replay()
        return self.loss
