"""Whole-path parity at BASELINE.json's FULL sizes: configs[0] (1 clip x 4 frames, 448x800, S = 350 tokens per frame,
forward), the configs[1] shape (L = 3) and configs[2]'s stage-2 resolution (896x1600, S = 1400 — where the space-time
attention is tensor-bound), with the weights bench.py trains (residual_gain=0.5).  The drop-in (bf16
kernels) against goldens of the unmodified reference (tests/golden/make_golden_fullsize.py, fp32 CPU): 2e-2 relative L2
per tensor; matcher assignments, loss and AP bookkeeping exact against the oracle on the model's own fp32 outputs.
(File name: sorts last, so the kernel- and small-size model tests report first.)"""
import json
import os
from argparse import Namespace

import pytest
import torch

from conftest import GOLDEN, load_golden
from oracle import fod_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 2e-2


def rel_l2(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / max(1e-12, float(b.norm())))


@pytest.fixture(scope="module")
def model():
    from fod_b200 import SpatioTemporalDETRArgs, build_model
    from fod_b200.synthetic import synthetic_state_dict
    with open(os.path.join(GOLDEN, "state_dict_template.json")) as f:
        sd = synthetic_state_dict(json.load(f), seed=0, residual_gain=0.5)
    m = build_model(Namespace(device=DEV, distributed=False),
                    SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4),
                    dropout=0.0)
    m.load_state_dict(sd, strict=True)
    return m.eval()


@pytest.mark.parametrize("name", ["f1_b1_l4_448x800", "f2_b2_l3_448x800", "f3_b1_l3_896x1600"])
def test_full_size_forward_matches_reference_golden(model, name):
    from fod_b200.st_detr import check_match_status
    from fod_b200.synthetic import make_clip
    gold = load_golden(name)
    b, l, h, w, seed = gold["case"]
    data = make_clip(b, l, h, w, seed=seed)
    core = {}
    hook = model._model.detector.register_forward_hook(lambda m, i, o: core.update(o))
    with torch.no_grad():
        outputs, state, loss, stats, stuffs = model({k: (v.to(DEV) if torch.is_tensor(v) else v) for k, v in data.items()})
    hook.remove()
    check_match_status(model._criterion.last_status)
    errs = {"hs": rel_l2(core["hs"][-1].float(), gold["hs_last"]),
            "pred_logits": rel_l2(core["pred_logits"], gold["pred_logits"]),
            "pred_boxes": rel_l2(core["pred_boxes"], gold["pred_boxes"]),
            "aux_logits": rel_l2(core["all_logits"][:-1], gold["aux_logits"]),
            "aux_boxes": rel_l2(core["all_boxes"][:-1], gold["aux_boxes"]),
            "class_scores": rel_l2(outputs["class_scores"], gold["class_scores"]),
            "boxes": rel_l2(outputs["boxes"], gold["boxes"])}
    print(name, {k: round(v, 5) for k, v in errs.items()}, "loss", float(loss), float(gold["loss"]))
    for k, v in errs.items():
        assert v < TOL, (k, v)
    # discrete parts: exact against the oracle on the model's own fp32 outputs (identical inputs)
    targets = O.to_detr_targets(h, w, data["active"], data["boxes"], data["classes"])
    lv = model._criterion.last_match.cpu()
    for li in range(lv.shape[0]):
        ref = O.hungarian_match(core["all_logits"][li].float().cpu(), core["all_boxes"][li].float().cpu(), targets)
        for bi, (pi, ti) in enumerate(ref):
            mt = lv[li, bi]
            assert int((mt >= 0).sum()) == len(pi) and torch.equal(mt[pi].long(), ti), (li, bi)
    cpu_core = {"pred_logits": core["pred_logits"].float().cpu(), "pred_boxes": core["pred_boxes"].float().cpu(),
                "aux_outputs": [{"pred_logits": a.float().cpu(), "pred_boxes": c.float().cpu()}
                                for a, c in zip(core["all_logits"][:-1], core["all_boxes"][:-1])]}
    l_ref, s_ref, _ = O.criterion(cpu_core, targets)
    assert abs(float(loss) - float(l_ref)) < 1e-4 * max(1.0, abs(float(l_ref)))
    for k in s_ref:
        assert abs(float(stats[k]) - float(s_ref[k])) < 1e-4 * max(1.0, abs(float(s_ref[k]))), k
    st_ref, _ = O.post_process(cpu_core["pred_logits"], cpu_core["pred_boxes"], data)
    assert torch.equal(stuffs[1].cpu(), st_ref[1]) and torch.equal(stuffs[2].cpu().bool(), st_ref[2].bool())
    assert torch.equal(stuffs[3].cpu(), st_ref[3]) and torch.equal(stuffs[3].cpu(), gold["od_map"][3])
