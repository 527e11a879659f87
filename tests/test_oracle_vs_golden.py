"""Pins the CPU oracle (oracle/fod_oracle.py) to outputs of the unmodified reference
(tests/golden/*.pt, produced by tests/golden/make_golden.py).  fp32 tolerance 1e-4 (north star)."""
import pytest
import torch

from conftest import load_golden, rel_err
from fod_b200.synthetic import make_clip
from oracle import fod_oracle as O

CASES = ["g1_b2_l3_128x160", "g2_b1_l4_224x320", "g3_b3_l3_96x128_nt0"]
TOL = 1e-4


@pytest.mark.parametrize("name", CASES)
def test_forward_matches_reference(name, ref_state_dict):
    gold = load_golden(name)
    b, l, h, w, seed, nt = gold["case"]
    data = make_clip(b, l, h, w, seed=seed, fixed_targets=nt)
    torch.set_num_threads(8)
    with torch.no_grad():
        outputs, _, loss, stats, stuffs, core = O.forward(data, ref_state_dict)
    assert rel_err(core["encoder_features"], gold["encoder_features"]) < TOL
    assert rel_err(core["hs"], gold["hs"]) < TOL
    assert rel_err(core["pred_logits"], gold["pred_logits"]) < TOL
    assert rel_err(core["pred_boxes"], gold["pred_boxes"]) < TOL
    aux_l = torch.stack([a["pred_logits"] for a in core["aux_outputs"]])
    aux_b = torch.stack([a["pred_boxes"] for a in core["aux_outputs"]])
    assert rel_err(aux_l, gold["aux_logits"]) < TOL and rel_err(aux_b, gold["aux_boxes"]) < TOL
    assert rel_err(loss, gold["loss"]) < TOL
    for k, v in gold["stats"].items():
        assert rel_err(stats[k], v) < TOL, k
    # matcher assignments: bit-exact, every level and sample
    for lv_o, lv_g in zip(core["indices"], gold["indices"]):
        for (i, j), (gi, gj) in zip(lv_o, lv_g):
            assert torch.equal(i, gi) and torch.equal(j, gj)
    assert rel_err(outputs["class_scores"], gold["class_scores"]) < TOL
    assert rel_err(outputs["boxes"], gold["boxes"]) < TOL
    confs, is_pos, size_cat, num_annos = stuffs
    g_confs, g_pos, g_size, g_num = gold["od_map"]
    assert rel_err(confs, g_confs) < TOL
    assert torch.equal(is_pos, g_pos)
    assert torch.equal(size_cat.bool(), g_size.bool())
    assert torch.equal(num_annos, g_num)


def test_backward_matches_reference(ref_state_dict):
    """Autograd through the oracle reproduces the reference's parameter gradients (dropout off)."""
    gold = load_golden("g1_b2_l3_128x160")
    b, l, h, w, seed, nt = gold["case"]
    data = make_clip(b, l, h, w, seed=seed, fixed_targets=nt)
    sd = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and "backbone.body" not in k else v)
          for k, v in ref_state_dict.items()}
    _, _, loss, _, _, _ = O.forward(data, sd, with_od_map=False)
    loss.backward()
    worst = 0.0
    for k, g in gold["grads_small"].items():
        worst = max(worst, rel_err(sd[k].grad, g))
    assert worst < 2e-4, worst
    for k, n in gold["grad_norms"].items():
        if "backbone.body" in k or sd[k].grad is None:
            continue
        assert abs(float(sd[k].grad.norm()) - float(n)) <= 2e-4 * max(1.0, float(n)), k


def test_temporal_code_matches_reference(ref_state_dict):
    """PositionalEncoder(no_temporal=False) with per-sample temporal_offsets (paper.py:50-55,66-73; built off by
    runs/_model.py:70-72).  Golden: tests/golden/make_golden_temporal.py, case t1 — forward, loss, matcher indices and
    gradient norms of the unmodified reference with the switch flipped."""
    gold = load_golden("t1_temporal_b2_l3_128x160")
    b, l, h, w, seed = gold["case"]
    data = make_clip(b, l, h, w, seed=seed)
    data["temporal_offsets"] = gold["temporal_offsets"]
    temporal = {"offsets": gold["temporal_offsets"], "extra": gold["extra"]}
    sd = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and "backbone.body" not in k else v)
          for k, v in ref_state_dict.items()}
    torch.set_num_threads(8)
    _, _, loss, stats, _, core = O.forward(data, sd, with_od_map=False, temporal=temporal)
    assert rel_err(core["hs"][-1].detach(), gold["hs_last"]) < TOL
    assert rel_err(core["pred_logits"].detach(), gold["pred_logits"]) < TOL
    assert rel_err(core["pred_boxes"].detach(), gold["pred_boxes"]) < TOL
    aux_l = torch.stack([a["pred_logits"] for a in core["aux_outputs"]]).detach()
    aux_b = torch.stack([a["pred_boxes"] for a in core["aux_outputs"]]).detach()
    assert rel_err(aux_l, gold["aux_logits"]) < TOL and rel_err(aux_b, gold["aux_boxes"]) < TOL
    assert rel_err(loss.detach(), gold["loss"]) < TOL
    for k, v in gold["stats"].items():
        assert rel_err(stats[k].detach(), v) < TOL, k
    for lv_o, lv_g in zip(core["indices"], gold["indices"]):
        for (i, j), (gi, gj) in zip(lv_o, lv_g):
            assert torch.equal(i, gi) and torch.equal(j, gj)
    loss.backward()
    for k, n in gold["grad_norms"].items():
        if sd[k].grad is None:
            continue
        assert abs(float(sd[k].grad.norm()) - float(n)) <= 2e-4 * max(1.0, float(n)), k
    # ... and in this per-frame configuration the code is dead: every cross-attention sees ONE frame's encoding
    # (paper.py:381-382), so the temporal part shifts all keys of a softmax by the same vector and cancels.  The
    # reference's own output with the code ON equals the oracle's with it OFF.
    with torch.no_grad():
        off = O.forward(data, ref_state_dict, with_od_map=False)[5]
    assert rel_err(off["pred_logits"], gold["pred_logits"]) < TOL
    assert rel_err(off["pred_boxes"], gold["pred_boxes"]) < TOL


def test_temporal_code_in_joint_mode_matches_reference(ref_state_dict):
    """Case t2: frame-counter temporal code + JointEncoder (paper.py:180-203) over all T*h*w tokens — here tokens of
    different frames share a softmax and the code is live."""
    from oracle.fod_oracle import IMU_KEYS
    gold = load_golden("t2_temporal_joint_b2_l4_96x128")
    b, l, h, w, seed = gold["case"]
    data = make_clip(b, l, h, w, seed=seed)
    imu = torch.cat([data[k] for k in IMU_KEYS], dim=2)
    torch.set_num_threads(8)
    with torch.no_grad():
        on = O.core_forward(data["video"], imu, ref_state_dict, temporal={"offsets": None, "extra": gold["extra"]},
                            joint=True)
        off = O.core_forward(data["video"], imu, ref_state_dict, joint=True)
    assert rel_err(on["hs"][-1], gold["hs_last"]) < TOL
    assert rel_err(on["pred_logits"], gold["pred_logits"]) < TOL
    assert rel_err(on["pred_boxes"], gold["pred_boxes"]) < TOL
    assert rel_err(off["pred_logits"], gold["pred_logits"]) > 1e-3  # live


def test_temporal_code_frame_counter_and_normalisation():
    """paper.py:66-73: without offsets the code is the frame counter 1..T; either way it is divided by the LAST
    frame's value, so the last frame always carries the angle 2*pi*t/(t+1e-6)."""
    import math
    enc = O.temporal_encoding(None, 2, 3, 2, 2)
    assert enc.shape == (2, 3, 256, 2, 2)
    idx = torch.arange(256, dtype=torch.float32)
    dim_t = 10000.0 ** (2 * torch.div(idx, 2, rounding_mode="floor") / 256)
    for t in range(3):
        ang = torch.tensor((t + 1) / (3 + 1e-6) * 2 * math.pi, dtype=torch.float32) / dim_t
        want = torch.where(idx.long() % 2 == 0, ang.sin(), ang.cos())
        assert torch.allclose(enc[1, t, :, 0, 1], want, atol=1e-6)
    offs = torch.tensor([[-1.0, -0.5], [-0.8, -0.2]])
    e2 = O.temporal_encoding(offs, 2, 2, 1, 1)
    ang = torch.tensor(-0.8 / (-0.2 + 1e-6) * 2 * math.pi, dtype=torch.float32) / dim_t
    assert torch.allclose(e2[1, 0, :, 0, 0], torch.where(idx.long() % 2 == 0, ang.sin(), ang.cos()), atol=1e-5)


@pytest.mark.parametrize("name", ["f1_b1_l4_448x800", "f2_b2_l3_448x800", "f3_b1_l3_896x1600"])
def test_full_size_forward_matches_reference(name):
    """BASELINE.json configs[0] (1 clip x 4 frames, 448x800, CPU forward) and the configs[1] shape, at full size, with
    the weights bench.py trains (residual_gain=0.5).  Golden: tests/golden/make_golden_fullsize.py."""
    import json
    import os

    from conftest import GOLDEN
    from fod_b200.synthetic import synthetic_state_dict
    gold = load_golden(name)
    b, l, h, w, seed = gold["case"]
    with open(os.path.join(GOLDEN, "state_dict_template.json")) as f:
        sd = synthetic_state_dict(json.load(f), seed=0, residual_gain=gold["residual_gain"])
    data = make_clip(b, l, h, w, seed=seed)
    torch.set_num_threads(8)
    with torch.no_grad():
        outputs, _, loss, stats, stuffs, core = O.forward(data, sd)
    assert rel_err(core["hs"][-1], gold["hs_last"]) < TOL
    assert rel_err(core["pred_logits"], gold["pred_logits"]) < TOL
    assert rel_err(core["pred_boxes"], gold["pred_boxes"]) < TOL
    aux_l = torch.stack([a["pred_logits"] for a in core["aux_outputs"]])
    aux_b = torch.stack([a["pred_boxes"] for a in core["aux_outputs"]])
    assert rel_err(aux_l, gold["aux_logits"]) < TOL and rel_err(aux_b, gold["aux_boxes"]) < TOL
    assert rel_err(loss, gold["loss"]) < TOL
    for k, v in gold["stats"].items():
        assert rel_err(stats[k], v) < TOL, k
    for lv_o, lv_g in zip(core["indices"], gold["indices"]):
        for (i, j), (gi, gj) in zip(lv_o, lv_g):
            assert torch.equal(i, gi) and torch.equal(j, gj)
    assert rel_err(outputs["class_scores"], gold["class_scores"]) < TOL
    assert rel_err(outputs["boxes"], gold["boxes"]) < TOL
    confs, is_pos, size_cat, num_annos = stuffs
    g_confs, g_pos, g_size, g_num = gold["od_map"]
    assert rel_err(confs, g_confs) < TOL and torch.equal(is_pos, g_pos)
    assert torch.equal(size_cat.bool(), g_size.bool()) and torch.equal(num_annos, g_num)
