"""Golden vector for the optional JointEncoder mode (future_od/models/paper.py:180-203), produced by the
UNMODIFIED reference modules in the build container:

  PYTHONPATH=/root/reference:/root/reference/ConditionalDETR PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_joint.py

The encoder weights are the `_model.separate_encoder.transformer.*` entries of the name-keyed synthetic state dict, so
only inputs and the output are stored.  Inputs are bf16-representable so the CUDA path sees identical values.
"""
import json
import os
import sys

import torch
from torch import nn

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import fod_b200  # noqa: E402,F401
from fod_b200.synthetic import synthetic_state_dict  # noqa: E402

from future_od.models import transformer  # noqa: E402
from future_od.models.paper import JointEncoder, PositionalEncoder  # noqa: E402


def main():
    torch.manual_seed(0)
    torch.set_num_threads(8)
    with open(os.path.join(HERE, "state_dict_template.json")) as f:
        sd = synthetic_state_dict(json.load(f), seed=0)
    prefix = "_model.separate_encoder.transformer."
    enc = transformer.TransformerEncoder(layers=nn.ModuleList(
        transformer.TransformerEncoderLayer(D=256, Nhead=8, Dff=2048, use_egodeep=True) for _ in range(6)))
    enc.load_state_dict({k[len(prefix):]: v for k, v in sd.items() if k.startswith(prefix)})
    joint = JointEncoder(enc).eval()
    b, l, h, w = 2, 3, 5, 7
    g = torch.Generator().manual_seed(7)
    feats = torch.randn(b, l, 256, h, w, generator=g).to(torch.bfloat16).float()
    ego = torch.randn(b, l, 256, generator=g).to(torch.bfloat16).float()
    pe = PositionalEncoder(no_temporal=True)
    pos = pe.get_spatial_encoding(b * l, 256, h, w, feats.device).view(b, l, 256, h, w)
    with torch.no_grad():
        out, _ = joint(feats, pos, ego)
    torch.save({"features": feats, "egodeep": ego, "out": out, "shape": (b, l, h, w)}, os.path.join(HERE, "joint_b2_l3_5x7.pt"))
    print("joint golden", tuple(out.shape), float(out.abs().mean()))


if __name__ == "__main__":
    main()
