"""Synthetic clips and name-keyed synthetic weights (no datasets or checkpoints exist on the box).

`make_clip` emits the same dict the reference's datasets produce (future_od/datasets/nu_scenes.py:334-351,
dense 256-slot targets of future_od/datasets/utils.py:19-38).  `synthetic_state_dict` fills a name->shape
template deterministically from the parameter NAME, so the build container (reference model), the CPU
oracle and the CUDA modules all get identical weights without shipping a 216 MB checkpoint.
"""
import hashlib
import math

import torch


def make_clip(batch, frames, height, width, seed=0, max_targets=40, num_classes=8, slots=256, fixed_targets=None):
    g = torch.Generator().manual_seed(1000 + seed)
    r = lambda *s: torch.randn(*s, generator=g)
    u = lambda *s: torch.rand(*s, generator=g)
    data = {"video": r(batch, frames, 3, height, width)}
    quat = r(batch, frames, 4)
    quat = quat / quat.norm(dim=-1, keepdim=True)
    data["translation"] = r(batch, frames, 3)
    data["translation"][:, 0] = 0
    data["acceleration"] = r(batch, frames, 3)
    data["rotation"] = quat
    data["rotation"][:, 0] = torch.tensor([1.0, 0.0, 0.0, 0.0])
    data["rotation_rate"] = r(batch, frames, 3)
    data["speed"] = r(batch, frames, 1)
    boxes = torch.zeros(batch, slots, 4)
    classes = torch.zeros(batch, slots, dtype=torch.int64)
    active = torch.zeros(batch, slots, dtype=torch.int64)
    for b in range(batch):
        nt = int(torch.randint(0, max_targets, (1,), generator=g)) if fixed_targets is None else int(fixed_targets)
        x0 = u(nt) * 0.6 * width
        y0 = u(nt) * 0.6 * height
        bw = 4 + u(nt) * (0.3 * width - 4)
        bh = 4 + u(nt) * (0.3 * height - 4)
        boxes[b, :nt] = torch.stack((x0, y0, x0 + bw, y0 + bh), dim=1)
        classes[b, :nt] = torch.randint(0, num_classes, (nt,), generator=g)
        active[b, :nt] = 1
    data.update(boxes=boxes, classes=classes, active=active,
                annotated_frame_idx=torch.full((batch,), frames - 1, dtype=torch.int64),
                temporal_offsets=torch.linspace(-1.0, 0.0, frames).repeat(batch, 1))
    return data


def _gen_for(name, seed):
    h = hashlib.sha256(f"{seed}:{name}".encode()).digest()
    return torch.Generator().manual_seed(int.from_bytes(h[:7], "little"))


def synthetic_state_dict(template, seed=0, residual_gain=1.0):
    """template: dict name -> shape (or tensor).  Values depend only on (seed, name, shape).

    residual_gain: factor on the affine weight of the LAST frozen BatchNorm of every backbone bottleneck (bn3).  With
    1.0 (the fixtures under tests/golden were generated with it) the un-normalised He-initialised ResNet doubles its
    activation variance per block (layer4 std ~60, end-to-end Jacobian gain ~1e4) and Adam's coherent first steps can
    tip it into overflow within ~10-50 steps on a quarter of the runs.  0.5 keeps every stage at std 0.3-0.7, the
    regime of a trained checkpoint; training workloads (bench.py, scripts/) use that."""
    out = {}
    for name, shp in template.items():
        shape = tuple(shp.shape) if hasattr(shp, "shape") else tuple(shp)
        g = _gen_for(name, seed)
        leaf = name.rsplit(".", 1)[-1]
        if leaf == "num_batches_tracked":
            out[name] = torch.zeros(shape, dtype=torch.int64)
        elif leaf == "running_var":
            out[name] = 0.8 + 0.4 * torch.rand(shape, generator=g)
        elif leaf == "running_mean":
            out[name] = 0.05 * torch.randn(shape, generator=g)
        elif len(shape) >= 2:
            fan_out, fan_in = shape[0], int(math.prod(shape[1:]))
            if len(shape) == 4 and "input_proj" not in name:  # backbone conv: He-style
                std = math.sqrt(2.0 / (shape[0] * shape[2] * shape[3]))
                out[name] = std * torch.randn(shape, generator=g)
            elif "query_embed" in name:
                out[name] = torch.randn(shape, generator=g)
            elif "bbox_embed.layers.2" in name:
                out[name] = 0.05 * torch.randn(shape, generator=g)
            else:
                a = math.sqrt(6.0 / (fan_in + fan_out))
                if "input_proj" in name and residual_gain == 1.0:
                    # the un-normalised random ResNet emits features of std ~80; a trained input_proj hands the encoder
                    # O(1) tokens.  Without this the layer-0 attention logits have std ~2000 (one-hot softmax), a regime
                    # no trained checkpoint is in and in which bf16 gradients are meaningless.
                    a /= 64.0
                out[name] = (torch.rand(shape, generator=g) * 2 - 1) * a
        else:
            if "backbone.body" in name:  # frozen BN affine
                out[name] = (0.9 + 0.2 * torch.rand(shape, generator=g)) if leaf == "weight" else 0.02 * torch.randn(shape, generator=g)
                if leaf == "weight" and ".bn3." in name:
                    out[name] = out[name] * residual_gain
            elif "norm" in name and leaf == "weight":
                out[name] = 1.0 + 0.1 * torch.randn(shape, generator=g)
            elif "class_embed" in name:
                out[name] = -2.0 + 0.5 * torch.randn(shape, generator=g)
            else:
                out[name] = 0.05 * torch.randn(shape, generator=g)
    return out
