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
