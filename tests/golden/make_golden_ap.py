"""Golden vector for the epoch-level AP aggregation (future_od/utils/od_map.py:288-364, driven as
future_od/trainer.py:155,205-236 does), produced by the UNMODIFIED reference functions:

  PYTHONPATH=/root/reference:/root/reference/ConditionalDETR PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_ap.py

`aggregate_mean_average_precision` times itself with CUDA events (od_map.py:336-350); this container has no GPU, so
torch.cuda.Event / synchronize are replaced by no-op stand-ins for the call — the arithmetic is untouched.
Per-iteration inputs are the reference's own `prepare_od_map_stuffs` on jittered ground-truth boxes (so that every IoU
threshold has hits and misses), 4 iterations of 3 clips, concatenated exactly as the trainer does.
"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import fod_b200  # noqa: E402,F401
from fod_b200.synthetic import make_clip  # noqa: E402

from future_od.utils import od_map  # noqa: E402


class _NoEvent:
    def __init__(self, *a, **k):
        pass

    def record(self):
        pass

    def elapsed_time(self, other):
        return 0.0


def predictions(data, g, m=128, ncls=8):
    """(B, M, 4) xyxy pixel boxes and (B, M, C+1) scores: the first queries shadow the targets with growing jitter."""
    b = data["boxes"].shape[0]
    h, w = data["video"].shape[-2:]
    boxes = torch.rand(b, m, 4, generator=g) * torch.tensor([w, h, w, h]) * 0.5
    boxes[..., 2:] += boxes[..., :2] + 2.0
    logits = torch.randn(b, m, ncls, generator=g) - 2.0
    for i in range(b):
        nt = int(data["active"][i].sum())
        k = min(nt, m // 2)
        for rep in range(2):  # two competing detections per target
            sl = slice(rep * k, (rep + 1) * k)
            wh = (data["boxes"][i, :k, 2:] - data["boxes"][i, :k, :2]).repeat(1, 2)
            jitter = torch.randn(k, 4, generator=g) * wh * (0.02 + 0.12 * torch.rand(k, 1, generator=g))
            boxes[i, sl] = data["boxes"][i, :k] + jitter
            right = torch.rand(k, generator=g) < 0.8
            cls = torch.where(right, data["classes"][i, :k], torch.randint(0, ncls, (k,), generator=g))
            logits[i, sl].scatter_(1, cls[:, None], 1.0 + 2.0 * torch.rand(k, 1, generator=g))
    sc = logits.sigmoid()
    return boxes, torch.cat((sc, sc.max(dim=2, keepdim=True)[0]), dim=2)


def main():
    torch.manual_seed(0)
    g = torch.Generator().manual_seed(77)
    h, w = 448, 800
    per_iter, lst = [], [[], [], [], []]
    for it in range(4):
        data = make_clip(3, 3, 32, 32, seed=40 + it)  # tiny video: only the annotation slots matter here
        data["video"] = torch.zeros(3, 3, 3, 1, 1).expand(3, 3, 3, h, w)
        scale = torch.tensor([w / 32.0, h / 32.0, w / 32.0, h / 32.0])
        data["boxes"] = data["boxes"] * scale
        pb, sc = predictions(data, g)
        stuffs = od_map.prepare_od_map_stuffs(pb, sc, data["boxes"], data["classes"], data["active"], (h, w))
        per_iter.append({"pred_boxes": pb, "scores": sc, "boxes": data["boxes"], "classes": data["classes"],
                         "active": data["active"], "stuffs": [t.clone() for t in stuffs]})
        for i, t in enumerate(stuffs):
            lst[i].append(t.detach().cpu())
    cat = [torch.cat(lst[0], dim=2), torch.cat(lst[1], dim=2), torch.cat(lst[2], dim=2), torch.stack(lst[3], dim=2)]
    real = (torch.cuda.Event, torch.cuda.synchronize)
    torch.cuda.Event, torch.cuda.synchronize = _NoEvent, (lambda *a, **k: None)
    try:
        ap = od_map.aggregate_mean_average_precision(*cat, device="cpu")
    finally:
        torch.cuda.Event, torch.cuda.synchronize = real
    ap = {k: torch.as_tensor(v).clone() for k, v in ap.items()}
    torch.save({"imsize": (h, w), "iterations": per_iter, "ap": ap}, os.path.join(HERE, "ap_aggregate_4x3.pt"))
    print({k: tuple(v.shape) for k, v in ap.items()})
    print("classavg threshavg", ap["classavg threshavg"], "positives", int(cat[1].sum()), "of", cat[1].numel())


if __name__ == "__main__":
    main()
