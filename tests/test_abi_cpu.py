"""CPU checks of the drop-in boundary: the C-ABI library loads and exports every symbol include/fod_b200.h declares
(no compute calls without a GPU), the ctypes structs match the header's field lists, and the Python module tree
reproduces the reference's state_dict contract."""
import ctypes
import json
import os
import re
from argparse import Namespace

import pytest
import torch

from conftest import GOLDEN, ROOT

HEADER = os.path.join(ROOT, "include", "fod_b200.h")


def _header_text():
    with open(HEADER) as f:
        return re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)


def test_library_exports_every_declared_symbol():
    from fod_b200 import lib
    decl = set(re.findall(r"\b(fod_[a-z0-9_]+)\s*\(", _header_text()))
    assert {"fod_linear_fwd", "fod_attn_fwd", "fod_attn_bwd", "fod_match", "fod_set_loss", "fod_targets_compact",
            "fod_add_ln_fwd", "fod_ln_bwd", "fod_colsum", "fod_version", "fod_last_error"} <= decl
    assert decl == set(lib.SYMBOLS), decl ^ set(lib.SYMBOLS)
    handle = lib.load()
    raw = ctypes.CDLL(lib.LIB_PATH)
    for name in decl:
        assert getattr(raw, name) is not None
    assert handle.fod_version() == lib.ABI_VERSION


@pytest.mark.parametrize("struct,cls", [("fod_linear_args", "LinearArgs"), ("fod_attn_args", "AttnArgs"),
                                        ("fod_attn_bwd_args", "AttnBwdArgs"), ("fod_match_args", "MatchArgs"),
                                        ("fod_set_loss_args", "SetLossArgs"), ("fod_adamw_args", "AdamWArgs"), ("fod_ap_args", "ApArgs"),
                                        ("fod_fold_desc", "FoldDesc"), ("fod_dropout", "Dropout")])
def test_ctypes_structs_follow_the_header(struct, cls):
    from fod_b200 import lib
    body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (struct, struct), _header_text(), flags=re.S).group(1)
    names = []
    for stmt in body.split(";"):
        stmt = stmt.strip()
        if not stmt:
            continue
        stmt = re.sub(r"^(const\s+)?(void|float|int32_t|int64_t|fod_attn_args|fod_dropout)\s*\**", "", stmt)
        names += [re.sub(r"\[\d+\]$", "", n.strip().lstrip("*")) for n in stmt.split(",")]
    assert names == [f[0] for f in getattr(lib, cls)._fields_]


def test_null_arguments_are_rejected_without_a_gpu():
    """Argument validation happens before any CUDA call, so the status convention is checkable on CPU."""
    from fod_b200 import lib
    handle = lib.load()
    assert handle.fod_linear_fwd(None, None) == lib.FOD_E_ARG
    assert handle.fod_match(None, None) == lib.FOD_E_ARG
    assert b"null" in handle.fod_last_error()
    with pytest.raises(lib.FodError):
        lib.check(lib.FOD_E_ARG, "probe")


def test_state_dict_contract_matches_the_reference():
    from fod_b200 import SpatioTemporalDETRArgs, build_model
    with open(os.path.join(GOLDEN, "state_dict_template.json")) as f:
        template = json.load(f)
    model = build_model(Namespace(device="cpu", distributed=False),
                        SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4))
    sd = model.state_dict()
    assert set(sd) == set(template)
    assert all(list(sd[k].shape) == template[k] for k in template)
    assert sum(p.numel() for p in model.parameters()) == 54113614          # SURVEY.md §3.4 [probed]
    assert sum(p.numel() for p in model.parameters() if p.requires_grad) == 53891214
    assert len(list(model.parameters())) == 562
    low_lr = [n for n, p in model.named_parameters() if "backbone" in n and p.requires_grad]
    assert low_lr and all("layer1" not in n for n in low_lr)                # paper.py:102-109
    assert model.get_stat_idfs() == ["labels", "box_l1", "box_giou", "cardinality", "class_error"]
    # init conventions of paper.py:306-313
    det = model._model.detector
    assert torch.all(det.bbox_embed.layers[-1].weight == 0) and torch.allclose(det.class_embed.bias, torch.full((8,), -4.59512))


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "future-object-detection_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            with open(os.path.join(pkg, fn)) as f:
                src = f.read()
            assert not re.search(r"^\s*(from|import)\s+[\w.]*oracle", src, flags=re.M), fn
            assert "import_module(" not in src and "__import__(" not in src, fn


def test_frozen_bn_fold_cache_refreshes_in_place():
    """FrozenBatchNorm2d.fold() caches (scale, shift); load_state_dict must refresh the SAME tensors (a captured CUDA
    graph keeps reading them), and the values must equal backbone.py:53-63."""
    from fod_b200.model import FrozenBatchNorm2d
    bn = FrozenBatchNorm2d(6)
    s0, t0 = bn.fold()
    assert torch.equal(s0, torch.ones(6) * (1 + 1e-5) ** -0.5) and bn.fold()[0] is s0
    sd = {"weight": torch.rand(6) + 0.5, "bias": torch.randn(6), "running_mean": torch.randn(6), "running_var": torch.rand(6) + 0.1}
    bn.load_state_dict(sd)
    s1, t1 = bn._fold_st  # refreshed by the post hook, without another fold() call
    assert s1 is s0 and t1 is t0
    scale = sd["weight"] * (sd["running_var"] + 1e-5).rsqrt()
    assert torch.allclose(s1, scale) and torch.allclose(t1, sd["bias"] - sd["running_mean"] * scale)
    x = torch.randn(2, 6, 3, 3)
    assert torch.allclose(bn(x), x * s1.view(1, -1, 1, 1) + t1.view(1, -1, 1, 1), atol=1e-6)


def test_hot_kernels_are_tcgen05_tma_code():
    """Static guard on the built library: the GEMM, attention and stem kernels must contain tensor-core (`UTC*MMA` =
    tcgen05.mma), TMEM (`LDTM` = tcgen05.ld) and — where operands come by TMA — `UTMALDG` instructions in their SASS
    (mnemonics: /opt/skills/guides/B200_PROFILING.md).  A rebuild that silently lost them would still pass parity."""
    import re
    import shutil
    import subprocess
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe):
        pytest.skip("cuobjdump not available")
    from fod_b200 import lib
    sass = subprocess.run([exe, "-sass", lib.LIB_PATH], capture_output=True, text=True, timeout=300).stdout
    counts, fn = {}, None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = m.group(1)
            counts[fn] = {"mma": 0, "ldtm": 0, "tma": 0}
        elif fn:
            counts[fn]["mma"] += bool(re.search(r"UTC[A-Z]*MMA", line))
            counts[fn]["ldtm"] += "LDTM" in line
            counts[fn]["tma"] += "UTMALDG" in line
    families = {"gemm_bf16_kernel": True, "attn_fwd_kernel": True, "attn_bwd_dq_kernel": True,
                "attn_bwd_dkdv_kernel": True, "attn_bwd_fused_kernel": True, "stem_conv7x7s2_kernel": False}  # value: operands arrive by TMA
    for fam, wants_tma in families.items():
        kernels = {k: v for k, v in counts.items() if fam in k}
        assert kernels, fam
        for k, v in kernels.items():
            assert v["mma"] > 0 and v["ldtm"] > 0, (k, v)
            assert v["tma"] > 0 or not wants_tma, (k, v)
