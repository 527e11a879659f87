"""Learning-rate schedule under the captured step (the reference's StepLR, runs/_helper.py:104-106): the rates are
by-value arguments of the fused AdamW kernel, i.e. baked into the CUDA graph; FlatTrainState.set_lr() must make
GraphedTrainStep re-capture so the next replay really uses them."""
from argparse import Namespace

import pytest
import torch

pytestmark = pytest.mark.gpu


def test_set_lr_reaches_the_captured_step(ref_state_dict):
    from fod_b200 import SpatioTemporalDETRArgs, build_model
    from fod_b200 import train as T
    from fod_b200.optim import FlatTrainState
    from fod_b200.synthetic import make_clip
    dev = "cuda:0"
    dargs = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
    model = build_model(Namespace(device=dev, distributed=False), dargs, dropout=0.0)
    model.load_state_dict(ref_state_dict)
    model.train()
    batch = {k: v.to(dev) for k, v in make_clip(2, 3, 128, 160, seed=31).items()}
    flat = FlatTrainState(model, lr=dargs.lr, lr_backbone=dargs.lr_backbone, weight_decay=dargs.weight_decay)
    step = T.GraphedTrainStep(model, flat, batch, dargs.max_norm, warmup=1)
    step(batch)
    torch.cuda.synchronize()
    first_graph = step.graph
    # rate 0 in both groups: decoupled decay and the Adam update both scale with lr, so the parameters must not move
    flat.set_lr(0.0, 0.0)
    before = flat.params.clone()
    loss = step(batch)
    torch.cuda.synchronize()
    assert step.graph is not first_graph, "set_lr() did not trigger a re-capture"
    assert bool(torch.isfinite(loss)) and torch.equal(flat.params, before)
    same_graph = step.graph
    step(batch)  # unchanged rates: no further capture
    assert step.graph is same_graph
    # back to a non-zero rate: the parameters move again, by a step of the order of lr per element
    flat.set_lr(1e-3, 1e-3)
    step(batch)
    torch.cuda.synchronize()
    delta = (flat.params - before).abs()
    assert float(delta.max()) > 1e-4 and float(delta.max()) < 5e-2 and bool(torch.isfinite(flat.params).all())
    flat.close()
